"""GPU: the 'random' init and its two streams (engine.draw_init).

nmf-torch seeds torch with ``random_state`` and draws |sqrt(mean(X)/K) * randn| for H then W on the device it
computes on; the host stream (use_gpu=False, the oracle's) and the device stream (use_gpu=True) are both offered.
PARITY UNPINNED against nmf-torch itself (absent here): what is pinned is that each stream equals what torch's
own generator of that device yields for the seed."""
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def counts():
    from cnmf_torch_b200.synth import nb_counts
    return nb_counts(500, 240, k_true=5, seed=3, libsize=500.0)


def test_device_stream_equals_torch_default_cuda_generator(counts):
    from cnmf_torch_b200.engine import DenseFrobeniusEngine
    eng = DenseFrobeniusEngine(counts, init_device="cuda")
    K, seeds = 6, [11, 59886188, 0]
    H0, W0 = eng.make_init(K, seeds)
    assert H0.is_cuda and W0.is_cuda and H0.shape == (3, 500, K) and W0.shape == (3, K, 240)
    avg = math.sqrt(eng.x_mean / K)
    for i, s in enumerate(seeds):
        torch.manual_seed(s)                                        # what run_nmf does before the draws
        h = torch.abs(avg * torch.randn((500, K), dtype=torch.float32, device="cuda"))
        w = torch.abs(avg * torch.randn((K, 240), dtype=torch.float32, device="cuda"))
        assert torch.equal(h, H0[i]) and torch.equal(w, W0[i])     # bit-identical
    H1, W1 = eng.make_init(K, seeds)
    assert torch.equal(H0, H1) and torch.equal(W0, W1)             # deterministic


def test_host_stream_equals_oracle_and_half_normal_moments(counts):
    from cnmf_torch_b200.engine import DenseFrobeniusEngine
    from oracle import nmf_oracle as no
    eng = DenseFrobeniusEngine(counts)
    H0, W0 = eng.make_init(5, [7, 8])
    assert not H0.is_cuda
    for i, s in enumerate([7, 8]):
        h, w = no.init_random(float(counts.mean(dtype=np.float64)), 500, 240, 5, s)
        assert np.array_equal(np.asarray(h), H0[i].numpy()) and np.array_equal(np.asarray(w), W0[i].numpy())
    # both streams are |avg * N(0,1)|: E = avg * sqrt(2/pi)
    Hd, Wd = eng.make_init(5, list(range(40)), init_device="cuda")
    avg = math.sqrt(eng.x_mean / 5)
    for t in (Hd, Wd):
        assert (t >= 0).all() and abs(float(t.mean()) / (avg * math.sqrt(2 / math.pi)) - 1) < 0.01


@pytest.mark.parametrize("mode", ["batch", "online", "kl"])
def test_factorize_from_device_init_equals_explicit_init(counts, mode):
    """Drawing on the device is only a different source of H0/W0: passing the same arrays explicitly (through the
    host) gives bit-identical results."""
    from cnmf_torch_b200.engine import DenseFrobeniusEngine
    from cnmf_torch_b200.kl_engine import SparseKLEngine
    eng = (SparseKLEngine if mode == "kl" else DenseFrobeniusEngine)(counts, init_device="cuda")
    seeds = [3, 4, 5]
    H0, W0 = eng.make_init(5, seeds)
    if mode == "online":
        a = eng.factorize_online(5, seeds=seeds, tol=0.0, max_pass=2, chunk_size=200, chunk_max_iter=30)
        b = eng.factorize_online(5, H0=H0.cpu(), W0=W0.cpu(), tol=0.0, max_pass=2, chunk_size=200, chunk_max_iter=30)
    else:
        a = eng.factorize(5, seeds=seeds, tol=0.0, max_iter=20)
        b = eng.factorize(5, H0=H0.cpu(), W0=W0.cpu(), tol=0.0, max_iter=20)
    assert np.array_equal(a.W, b.W) and np.array_equal(a.err, b.err)


def test_use_gpu_flag_selects_the_stream_in_cnmf_factorize(tmp_path, counts):
    """cNMF.factorize: the ledger's use_gpu flag picks the generator (device under True, host under False), and
    the constructor's init_device overrides it."""
    from cnmf_torch_b200 import cnmf as host
    from cnmf_torch_b200.engine import DenseFrobeniusEngine
    out = {}
    for tag, use_gpu, override in (("dev", True, None), ("host", False, None), ("forced", True, "cpu")):
        obj = host.cNMF(output_dir=str(tmp_path), name=tag, init_device=override)
        obj.prepare_from_matrix(counts, [4], n_iter=3, seed=14, beta_loss="frobenius", use_gpu=use_gpu)
        obj.factorize(nmf_kwargs_override=dict(mode="batch", batch_max_iter=20, tol=0.0))
        out[tag] = np.stack([host.load_df_from_npz(obj.paths["iter_spectra"] % (4, i)).values for i in range(3)])
        rp = host.load_df_from_npz(obj.paths["nmf_replicate_parameters"])
    assert np.array_equal(out["host"], out["forced"]) and not np.array_equal(out["host"], out["dev"])
    X, _, _ = obj.load_norm_counts()
    seeds = [int(s) for s in rp["nmf_seed"]]
    for tag, dev in (("host", "cpu"), ("dev", "cuda")):
        eng = DenseFrobeniusEngine(X, init_device=dev)
        ref = eng.factorize(4, seeds=seeds, tol=0.0, max_iter=20)
        assert np.array_equal(ref.W, out[tag])


@pytest.mark.parametrize("shape", [(500, 240), (5000, 2000), (37, 19)])
def test_fused_device_init_is_bit_identical_to_torch(shape):
    """cnmf_b200_init_random (all replicates of a K in one launch, written straight into the stacked K-major
    layout) restates ATen's grid-stride Philox geometry: every value equals torch's own CUDA generator's."""
    from cnmf_torch_b200.engine import DenseFrobeniusEngine, FrobBatch, SeedInit
    from cnmf_torch_b200.synth import nb_counts
    n, g = shape
    x = nb_counts(n, g, k_true=4, seed=5, libsize=300.0)
    eng = DenseFrobeniusEngine(x, init_device="cuda")
    for K, seeds in ((7, [3, 14, 59886188, 2**31 + 5]), (12, [0, 1]), (13, [9])):
        b = eng.new_batch(K, seeds=seeds)                    # SeedInit path
        H0, W0 = eng.make_init(K, seeds)                     # torch.randn per replicate
        assert torch.equal(b.Ht[:, :K, :n].permute(0, 2, 1), H0) and torch.equal(b.Wall[:, :K, :g], W0)
        assert float(b.Ht[:, K:].abs().sum()) == 0.0 and float(b.Ht[:, :, n:].abs().sum()) == 0.0
        assert float(b.Wall[:, K:].abs().sum()) == 0.0 and float(b.Wall[:, :, g:].abs().sum()) == 0.0
    # a stack of several K, and a row-sharded engine keeping its block of the global draw
    grp = [(9, SeedInit([1, 2, 3]), None), (11, SeedInit([4, 5]), None)]
    st = FrobBatch(eng, None, None, None, "mu", 0, 0, 0, 0, 1e-16, groups=grp)
    H9, W9 = eng.make_init(9, [1, 2, 3]); H11, W11 = eng.make_init(11, [4, 5])
    assert torch.equal(st.Ht[:3, :9, :n].permute(0, 2, 1), H9) and torch.equal(st.Ht[3:, :11, :n].permute(0, 2, 1), H11)
    assert torch.equal(st.Wall[:3, :9, :g], W9) and torch.equal(st.Wall[3:, :11, :g], W11)
    if n >= 100:
        lo, hi = n // 5, n // 5 + n // 2
        part = DenseFrobeniusEngine(x[lo:hi], init_device="cuda")
        part.N_total, part.row_offset, part.x_mean = n, lo, eng.x_mean      # what a row group's all-reduce sets
        pb = part.new_batch(7, seeds=[3, 14])
        H0, W0 = eng.make_init(7, [3, 14])
        assert torch.equal(pb.Ht[:, :7, :hi - lo].permute(0, 2, 1), H0[:, lo:hi]) and torch.equal(pb.Wall[:, :7, :g], W0)


def test_fused_device_init_large_tensor_multiple_grid_strides():
    """More elements than one pass of torch's capped grid (148 SMs x 8 blocks x 256 threads x 4): the grid-stride
    loop runs twice and the Philox offset of the W draw moves accordingly."""
    from cnmf_torch_b200.engine import DenseFrobeniusEngine
    x = np.ones((160000, 64), dtype=np.float32)
    eng = DenseFrobeniusEngine(x, init_device="cuda")
    b = eng.new_batch(15, seeds=[7, 8])
    H0, W0 = eng.make_init(15, [7, 8])
    assert torch.equal(b.Ht[:, :15, :160000].permute(0, 2, 1), H0) and torch.equal(b.Wall[:, :15, :64], W0)
