"""Dev probe: how long the HOST needs to enqueue one online pass (125 000 cells = one of 8 ranks of cfg-5) against how
long the device needs to run it.  python scripts/online_host_probe.py [cells]"""
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
for _ in range(3):
    eng.factorize_online(15, seeds=list(range(50)), max_pass=1, **kw)
torch.cuda.synchronize()
# the last thing factorize_online does is read the objective back (a sync), so host time ~ max(host enqueue, device);
# compare with the device-only time between events placed around the call
for passes in (1, 4):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record()
    res = eng.factorize_online(15, seeds=list(range(50)), max_pass=passes, **kw)
    e1.record(); torch.cuda.synchronize()
    print(f"{n} cells, {passes} pass(es): wall {1e3 * (time.perf_counter() - t0):.2f} ms, events {e0.elapsed_time(e1):.2f} ms, launches {res.launches}")
# host enqueue rate alone: the same calls with the device kept busy by a long kernel queued first
big = torch.empty(1 << 28, device=dev)
for _ in range(20):
    big.add_(1.0)            # ~20 x 1 ms of device work ahead of the pass: the host runs free
t0 = time.perf_counter()
import cnmf_torch_b200.engine as E
orig = E.FrobBatch.read_loss if hasattr(E, "FrobBatch") else None
res = eng.factorize_online(15, seeds=list(range(50)), max_pass=1, **kw)
print(f"wall with 20 ms of device work queued ahead: {1e3 * (time.perf_counter() - t0):.2f} ms (device alone ~20 + pass)")
