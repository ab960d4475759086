"""Dev probe: per-kernel event timing of one iteration's four launches for one stack (engine path, so the stream-K /
split-K-tail schedules are available): uses torch profiler-free CUDA events around single-iteration calls."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cnmf_torch_b200.engine import DenseFrobeniusEngine
from cnmf_torch_b200.synth import nb_counts
K = int(os.environ.get("K", "12")); R = int(os.environ.get("R", "400"))
eng = DenseFrobeniusEngine(nb_counts(2700, 2000, seed=14))
b = eng.new_batch(K, seeds=list(range(R)), algo="hals")
b.iterate(5); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); b.iterate(40); e1.record(); torch.cuda.synchronize()
print(f"SK={os.environ.get('CNMF_B200_TC_SK','0')} T={os.environ.get('CNMF_B200_TC_T','auto')} K={K} R={R}: {e0.elapsed_time(e1)/40*1e3:.1f} us per iteration")
