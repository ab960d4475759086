"""Dev probe: one rank's share of cfg-5 (online mode, chunk 5000, K=15, R=50) -- ms per pass, and under
`ncu --metrics gpu__time_duration.sum` the launch list of a pass.
    python scripts/online_pass_probe.py [cells]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from cnmf_torch_b200.engine import DenseFrobeniusEngine
n = int(sys.argv[1]) if len(sys.argv) > 1 else 125000
dev = torch.device("cuda", 0)
x = bench._cfg5_block_device(n, 0, dev)
eng = DenseFrobeniusEngine(x, device=dev, init_device="cuda")
kw = dict(tol=0.0, chunk_size=5000, chunk_max_iter=200)
eng.factorize_online(15, seeds=list(range(50)), max_pass=1, **kw)
torch.cuda.synchronize()
for passes in (1, 3):
    t0 = time.perf_counter()
    res = eng.factorize_online(15, seeds=list(range(50)), max_pass=passes, **kw)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"{n} cells, {passes} pass(es): {dt * 1e3:.1f} ms total, inner iterations {res.inner_iterations}, launches {res.launches}")
