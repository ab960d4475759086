import sys, os, time, cProfile, pstats
sys.path.insert(0, "/root/repo")
import torch
from cnmf_torch_b200.engine import DenseFrobeniusEngine
from cnmf_torch_b200.synth import nb_counts
xo = nb_counts(3000, 2000, seed=14)
e2 = DenseFrobeniusEngine(xo, device=torch.device("cuda", 0), init_device="cuda")
kw = dict(tol=0.0, max_pass=3, chunk_max_iter=1000)
for _ in range(3):
    e2.factorize_online(9, seeds=list(range(100)), **kw)
torch.cuda.synchronize()
ts = []
for _ in range(10):
    t0 = time.perf_counter(); res = e2.factorize_online(9, seeds=list(range(100)), **kw); torch.cuda.synchronize(); ts.append(time.perf_counter() - t0)
print("3 passes: min %.3f ms median %.3f ms, launches %d, inner %s" % (min(ts) * 1e3, sorted(ts)[5] * 1e3, res.launches, res.inner_iterations))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); res = e2.factorize_online(9, seeds=list(range(100)), **kw); e1.record(); torch.cuda.synchronize()
print("events: %.3f ms" % e0.elapsed_time(e1))
pr = cProfile.Profile(); pr.enable()
for _ in range(20):
    e2.factorize_online(9, seeds=list(range(100)), **kw)
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(18)
