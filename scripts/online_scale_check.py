"""Scale check of online mode at BASELINE cfg-5 shape reduced to what one rank of an 8-GPU run holds
(125 000 cells x 2000 genes dense, chunk 5000, K=15, R=50): seconds per pass, inner-iteration counts."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from cnmf_torch_b200.synth import nb_counts
from cnmf_torch_b200.engine import DenseFrobeniusEngine
N = int(os.environ.get("N", "125000"))
x = nb_counts(N, 2000, seed=14, libsize=400.0)
print(f"synthetic {x.shape} density {(x>0).mean():.3f}")
eng = DenseFrobeniusEngine(x)
torch.cuda.reset_peak_memory_stats()
t0 = time.perf_counter()
res = eng.factorize_online(15, seeds=list(range(1, 51)), tol=0.0, max_pass=2, chunk_size=5000, chunk_max_iter=1000)
torch.cuda.synchronize(); dt = time.perf_counter() - t0
obj = [v.mean() for _, v in res.trace]
print(f"online MU K=15 R=50: {dt/2:.2f} s / pass ({N//5000} chunks), inner iterations {res.inner_iterations}, launches {res.launches}, "
      f"objective {obj[0]:.4e} -> {obj[1]:.4e} -> {obj[2]:.4e}, peak device memory {torch.cuda.max_memory_allocated()/2**30:.2f} GiB")
assert obj[2] < obj[1] < obj[0]
