"""Multi-GPU check, launched as
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 scripts/multigpu_check.py
(a) cells row-sharded over the ranks, online mode, NCCL all-reduce of the per-chunk statistics, against the
    oracle restating the same execution order (row_blocks);
(b) replicates sharded over the ranks through cNMF.factorize(worker_i, total_workers): every replicate file
    written exactly once, identical to a single-worker run."""
import os, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
from cnmf_torch_b200 import dist as d
from cnmf_torch_b200.engine import DenseFrobeniusEngine
from cnmf_torch_b200.synth import nb_counts
from cnmf_torch_b200 import cnmf as host

x = nb_counts(803, 260, k_true=6, seed=14, libsize=700.0)
lo, hi = d.shard_rows(x.shape[0], rank, world)
eng = DenseFrobeniusEngine(x[lo:hi], device=f"cuda:{local}", row_group=dist.group.WORLD, row_offset=lo)
K, seeds, chunk = 6, [11, 12], 100
res = eng.factorize_online(K, seeds=seeds, tol=0.0, max_pass=3, chunk_size=chunk, chunk_max_iter=50, return_H=True)
ok = True
if rank == 0:
    from oracle import nmf_oracle as no
    blocks = [d.shard_rows(x.shape[0], r, world) for r in range(world)]
    for r, s in enumerate(seeds):
        h0, w0 = no.init_random(float(x.mean(dtype=np.float64)), x.shape[0], x.shape[1], K, s)
        ora = no.nmf_online(x, h0, w0, tol=0.0, max_pass=3, chunk_size=chunk, chunk_max_iter=50, row_blocks=blocks)
        for (p_o, obj_o), (p_g, obj_g) in zip(ora["trace"], res.trace):
            rel = abs(obj_g[r] - obj_o) / obj_o
            ok &= rel < 1e-3
            print(f"[row-sharded online, {world} ranks] replicate {r} pass {p_o}: oracle {obj_o:.4f} gpu {obj_g[r]:.4f} rel {rel:.2e}")
        c = np.corrcoef(res.W[r].ravel(), ora["W"].ravel())[0, 1]
        ok &= c >= 0.999
        print(f"  spectra Pearson {c:.6f}; local H rows Pearson {np.corrcoef(res.H[r].ravel(), ora['H'][lo:hi].ravel())[0,1]:.6f}")
# W must be identical on every rank (replicated, deterministic)
w_local = torch.from_numpy(res.W).cuda()
w0 = w_local.clone(); dist.broadcast(w0, 0)
same = bool(torch.equal(w0, w_local))
print(f"rank {rank}: W identical to rank 0: {same}")
ok &= same

# (b) replicate sharding through the file contract
tmp = [tempfile.mkdtemp() if rank == 0 else None]
dist.broadcast_object_list(tmp, 0)
obj = host.cNMF(output_dir=tmp[0], name="mg", device=f"cuda:{local}")
if rank == 0:
    obj.prepare_from_matrix(x, [4, 5], n_iter=6, seed=14, beta_loss="frobenius", var_names=[f"g{i}" for i in range(x.shape[1])])
dist.barrier()
obj.factorize(worker_i=rank, total_workers=world, nmf_kwargs_override=dict(mode="batch", batch_max_iter=60))
dist.barrier()
if rank == 0:
    obj.combine()
    ref = host.cNMF(output_dir=tmp[0], name="single", device=f"cuda:{local}")
    ref.prepare_from_matrix(x, [4, 5], n_iter=6, seed=14, beta_loss="frobenius", var_names=[f"g{i}" for i in range(x.shape[1])])
    ref.factorize(nmf_kwargs_override=dict(mode="batch", batch_max_iter=60))
    ref.combine()
    for k in (4, 5):
        a = host.load_df_from_npz(obj.paths["merged_spectra"] % k); b = host.load_df_from_npz(ref.paths["merged_spectra"] % k)
        assert list(a.index) == list(b.index) and a.shape == (6 * k, x.shape[1])
        c = np.corrcoef(a.values.ravel(), b.values.ravel())[0, 1]
        ok &= c >= 0.999
        print(f"[replicate-sharded factorize, {world} workers] k={k}: merged spectra {a.shape}, Pearson vs single worker {c:.6f}, max|diff| {np.abs(a.values-b.values).max():.2e}")
    med, labels, dens = obj.consensus(4, density_threshold=2.0)
    print("consensus spectra", med.shape, "row sums", med.sum(axis=1).round(4).tolist())
flag = torch.tensor([1 if ok else 0], device="cuda"); dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print("MULTIGPU_CHECK", "PASS" if int(flag.item()) == 1 else "FAIL")
dist.destroy_process_group()
sys.exit(0 if int(flag.item()) == 1 else 1)
