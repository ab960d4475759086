"""CPU, world_size 2, gloo: the host-side multi-rank logic (row / replicate sharding and the sum of the
per-chunk sufficient statistics) without any GPU."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from cnmf_torch_b200 import dist as d


def test_shard_rows_covers_everything():
    for n, p in [(10, 3), (1000000, 8), (7, 8), (5000, 2)]:
        blocks = [d.shard_rows(n, r, p) for r in range(p)]
        assert blocks[0][0] == 0 and blocks[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(blocks, blocks[1:]))
        sizes = [hi - lo for lo, hi in blocks]
        assert max(sizes) - min(sizes) <= 1


def test_shard_replicates_is_the_reference_stride(golden_dir):
    g = np.load(os.path.join(golden_dir, "worker_filter.npz"))
    for key in g.files:
        n, w, t = [int(v) for v in key.split("_")]
        assert d.shard_replicates(list(range(n)), w, t) == g[key].tolist()


def _worker(rank, world, port, tmp):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(0)
        n, g, k, R = 101, 13, 3, 2
        x = rng.random((n, g)).astype(np.float32)
        h = rng.random((R, n, k)).astype(np.float32)
        lo, hi = d.shard_rows(n, rank, world)
        # rank-local sufficient statistics of "its" cells, as the online loop forms them per chunk
        A = torch.from_numpy(np.einsum("rnk,rnj->rkj", h[:, lo:hi], h[:, lo:hi]))
        B = torch.from_numpy(np.einsum("rnk,ng->rkg", h[:, lo:hi], x[lo:hi]))
        cnt = torch.tensor([hi - lo], dtype=torch.float64)
        d.allreduce_sum_([A, B, cnt])
        np.testing.assert_allclose(A.numpy(), np.einsum("rnk,rnj->rkj", h, h), rtol=1e-5)
        np.testing.assert_allclose(B.numpy(), np.einsum("rnk,ng->rkg", h, x), rtol=1e-5)
        assert cnt.item() == n
        assert d.allreduce_max_int(3 + rank, torch.device("cpu")) == 3 + world - 1
        tot = d.allreduce_scalars([1.0, float(rank)], torch.device("cpu"))
        assert tot[0] == world and tot[1] == sum(range(world))
        # replicate sharding: every job exactly once over the ranks
        mine = d.shard_replicates(list(range(23)), rank, world)
        gathered = [None] * world
        dist.all_gather_object(gathered, mine)
        assert sorted(sum(gathered, [])) == list(range(23))
        open(os.path.join(tmp, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_gloo_world2_statistics_allreduce(tmp_path):
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert os.path.exists(tmp_path / "ok0") and os.path.exists(tmp_path / "ok1")
