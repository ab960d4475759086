"""Dev probe: clock64 stamps of CTA 0's first epilogue warp during one live H half-step (CNMF_B200_TC_PROF)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
prof = torch.zeros(64, dtype=torch.int64, device="cuda")
os.environ["CNMF_B200_TC_PROF"] = str(prof.data_ptr())
from cnmf_torch_b200 import _lib
from cnmf_torch_b200.engine import DenseFrobeniusEngine
from cnmf_torch_b200.synth import nb_counts
lib = _lib.lib()
K = int(os.environ.get("K", "12")); RR = int(os.environ.get("R", "120")); ALGO = os.environ.get("ALGO", "hals")
eng = DenseFrobeniusEngine(nb_counts(2700, 2000, seed=14))
b = eng.new_batch(K, seeds=list(range(RR)), algo=ALGO)
p = b.p
st = torch.cuda.current_stream().cuda_stream
part = torch.zeros(64 * b.R * 2, dtype=torch.float64, device="cuda")
Wop = b.Wall.clone(); Hop2 = b.Ht.clone()
def h(): _lib.check(lib.cnmf_b200_dense_half_step(p.X, p.ldx, p.n_cells, p.n_genes, Wop.data_ptr(), p.ldw, p.Ht, Hop2.data_ptr(), p.ldh, p.gramW, None, p.R, p.Kp, p.algo, 0., 0., p.eps, part.data_ptr(), st))
for _ in range(3): h()
torch.cuda.synchronize()
v = prof.cpu().numpy(); n = int(v[63])
names = ["sgram built", "tfull (accumulator ready)"] + [f"replicate {i}" for i in range(n - 3)] + ["tile done"]
t0 = v[0]
for i in range(n):
    print(f"{names[i]:28s} +{v[i]-t0:8d} clk   (delta {v[i]-v[i-1] if i else 0})")
