"""Dev probe: one resident batch alone on the GPU, ms per iteration (stream-K vs whole-tile scheduling via CNMF_B200_TC_SK)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cnmf_torch_b200.engine import DenseFrobeniusEngine
from cnmf_torch_b200.synth import nb_counts
eng = DenseFrobeniusEngine(nb_counts(2700, 2000, seed=14))
for K, R in ((8, 200), (12, 400), (10, 100), (7, 100), (5, 100), (13, 100)):
    b = eng.new_batch(K, seeds=list(range(R)), algo=os.environ.get("ALGO", "hals"))
    b.iterate(10); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); b.iterate(50); e1.record(); torch.cuda.synchronize()
    print(f"SK={os.environ.get('CNMF_B200_TC_SK','1')} K={K} R={R}: {e0.elapsed_time(e1)/50*1e3:.1f} us per iteration, loss[0]={b.read_loss()[0]:.6e}")
