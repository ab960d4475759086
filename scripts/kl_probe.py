"""Dev probe: the sparse KL half-steps at cfg-3 density (20k x 2k, ~7.8 %), K=9, R=100."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cnmf_torch_b200.kl_engine import SparseKLEngine
from cnmf_torch_b200.synth import nb_counts
x = nb_counts(int(os.environ.get("N", "20000")), 2000, seed=14, libsize=170.0)
eng = SparseKLEngine(x)
b = eng.new_batch(int(os.environ.get("K", "9")), seeds=list(range(int(os.environ.get("R", "100")))))
b.iterate(2); torch.cuda.synchronize()
t0 = time.perf_counter(); b.iterate(5); torch.cuda.synchronize(); t = (time.perf_counter() - t0) / 5
print(f"KL {x.shape} d={float((x>0).mean()):.3f} nnz={eng.nnz}: {t*1e3:.2f} ms / batched iteration -> {b.R/t:.0f} replicate-iters/s")
