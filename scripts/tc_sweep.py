"""Dev probe: time the H and W half-steps of cfg-2 (K=10, R=100) for the tile plan chosen via CNMF_B200_TC_T."""
import os, sys, torch, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cnmf_torch_b200 import _lib
from cnmf_torch_b200.engine import DenseFrobeniusEngine
from cnmf_torch_b200.synth import nb_counts
lib = _lib.lib()
x = nb_counts(2700, 2000, seed=14)
eng = DenseFrobeniusEngine(x)
K = int(os.environ.get("K", "10"))
RR = int(os.environ.get("R", "100"))
b = eng.new_batch(K, seeds=list(range(RR)), algo="hals")
p = b.p
st = torch.cuda.current_stream().cuda_stream
frozen = torch.zeros(b.R, dtype=torch.uint8, device="cuda")
part = torch.zeros(64 * b.R * 2, dtype=torch.float64, device="cuda")
Wop = b.Wall.clone(); Hop = b.Ht.clone()
ACT = None if os.environ.get("LIVE") else frozen.data_ptr()
PART = None if os.environ.get("NOPART") else part.data_ptr()
Hop2 = b.Ht.clone(); Wop2 = b.Wall.clone()
def h(): _lib.check(lib.cnmf_b200_dense_half_step(p.X, p.ldx, p.n_cells, p.n_genes, Wop.data_ptr(), p.ldw, p.Ht, Hop2.data_ptr(), p.ldh, p.gramW, ACT, p.R, p.Kp, p.algo, 0., 0., p.eps, PART, st))
def w(): _lib.check(lib.cnmf_b200_dense_half_step(p.Xt, p.ldxt, p.n_genes, p.n_cells, Hop.data_ptr(), p.ldh, p.Wall, Wop2.data_ptr(), p.ldw, p.gramH, ACT, p.R, p.Kp, p.algo, 0., 0., p.eps, PART, st))
ws = torch.empty(lib.cnmf_b200_gram_workspace_bytes(p.R, p.Kp, p.n_cells), dtype=torch.uint8, device="cuda")
def g(): _lib.check(lib.cnmf_b200_gram(p.Ht, p.ldh, p.R, p.Kp, p.n_cells, p.gramH, ws.data_ptr(), st))
Pbuf = torch.zeros(p.R * p.Kp, p.ldh, device="cuda")
def gs(): _lib.check(lib.cnmf_b200_dense_gemm_tf32(p.X, p.ldx, p.n_cells, p.n_genes, Wop.data_ptr(), p.ldw, Pbuf.data_ptr(), p.ldh, None, p.R, p.Kp, 0, st))
for name, fn in (("H", h), ("W", w), ("gemmH", gs), ("gramH", g)):
    for _ in range(5): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50): fn()
    e1.record(); torch.cuda.synchronize()
    print(f"t={os.environ.get('CNMF_B200_TC_T','auto')} K={K} {name}: {e0.elapsed_time(e1)/50*1000:.1f} us")
