"""Scale check at BASELINE cfg-3 shape (100 000 cells x 2000 genes, ~7 % density): sparse KL MU and dense
Frobenius MU, K=9, R=100 -- time per batched iteration, monotone objective, peak device memory."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from cnmf_torch_b200.synth import nb_counts
from cnmf_torch_b200.kl_engine import SparseKLEngine
from cnmf_torch_b200.engine import DenseFrobeniusEngine
N = int(os.environ.get("N", "100000"))
t0 = time.perf_counter(); x = nb_counts(N, 2000, seed=14, libsize=170.0); print(f"synthetic {x.shape} density {(x>0).mean():.3f} in {time.perf_counter()-t0:.1f}s")
seeds = list(range(1, 101))
for name, make in (("sparse KL MU", lambda: SparseKLEngine(x)), ("dense Frobenius MU (tcgen05)", lambda: DenseFrobeniusEngine(x))):
    torch.cuda.reset_peak_memory_stats()
    t0 = time.perf_counter(); eng = make(); torch.cuda.synchronize(); t_ing = time.perf_counter() - t0
    t0 = time.perf_counter(); b = eng.new_batch(9, seeds=seeds) if name.startswith("sparse") else eng.new_batch(9, seeds=seeds, algo="mu"); torch.cuda.synchronize(); t_init = time.perf_counter() - t0
    o0 = b.objective()
    b.iterate(10); torch.cuda.synchronize()
    o1 = b.read_loss()
    t0 = time.perf_counter(); b.iterate(10); torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 10
    o2 = b.read_loss()
    assert np.all(o1 < o0) and np.all(o2 <= o1 * (1 + 1e-6)), "objective must decrease"
    print(f"{name}: ingest {t_ing:.2f}s, init+upload of 100 replicates {t_init:.2f}s, {dt*1e3:.2f} ms / batched iteration "
          f"-> {100/dt:.0f} replicate-iters/s; objective {o0.mean():.4e} -> {o1.mean():.4e} -> {o2.mean():.4e}; "
          f"peak device memory {torch.cuda.max_memory_allocated()/2**30:.2f} GiB")
    del b, eng
    torch.cuda.empty_cache()
