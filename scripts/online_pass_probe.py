"""Dev probe: wall time per online pass (difference of a 4-pass and a 2-pass run), device-drawn init."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cnmf_torch_b200.synth import nb_counts
from cnmf_torch_b200.engine import DenseFrobeniusEngine
N = int(os.environ.get("N", "50000")); K = int(os.environ.get("K", "15")); R = int(os.environ.get("R", "50"))
x = nb_counts(N, 2000, seed=14, libsize=400.0)
eng = DenseFrobeniusEngine(x, init_device="cuda")
def run(p):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    res = eng.factorize_online(K, seeds=list(range(1, R + 1)), tol=0.0, max_pass=p, chunk_size=5000, chunk_max_iter=1000, inner_batch=int(os.environ.get("IB", "4")))
    torch.cuda.synchronize(); return time.perf_counter() - t0, res
run(1)
t2, r2 = run(2); t4, r4 = run(4)
print(f"online N={N} K={K} R={R}: 2 passes {t2*1e3:.1f} ms, 4 passes {t4*1e3:.1f} ms -> {(t4-t2)/2*1e3:.1f} ms per pass; "
      f"inner iterations (4 passes) {r4.inner_iterations}, launches {r4.launches}")
