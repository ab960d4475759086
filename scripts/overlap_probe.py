"""Dev probe: does a Gram kernel of one stack run concurrently with a tcgen05 half-step of another stack?"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cnmf_torch_b200 import _lib
from cnmf_torch_b200.engine import DenseFrobeniusEngine
from cnmf_torch_b200.synth import nb_counts
lib = _lib.lib()
eng = DenseFrobeniusEngine(nb_counts(2700, 2000, seed=14))
b12 = eng.new_batch(12, seeds=list(range(400)), algo="hals"); b8 = eng.new_batch(8, seeds=list(range(200)), algo="hals")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def half(b, st):
    p = b.p
    _lib.check(lib.cnmf_b200_dense_half_step(p.X, p.ldx, p.n_cells, p.n_genes, b._wop.data_ptr(), p.ldw, p.Ht, b._hop.data_ptr(), p.ldh, p.gramW, None, p.R, p.Kp, p.algo, 0., 0., p.eps, None, st.cuda_stream))
def gram(b, st):
    p = b.p
    _lib.check(lib.cnmf_b200_gram(p.Ht, p.ldh, p.R, p.Kp, p.n_cells, p.gramH, b._gws.data_ptr(), st.cuda_stream))
for b in (b12, b8):
    b._wop = b.Wall.clone(); b._hop = b.Ht.clone()
    b._gws = torch.empty(lib.cnmf_b200_gram_workspace_bytes(b.p.R, b.p.Kp, b.p.n_cells), dtype=torch.uint8, device="cuda")
def timed(fn, n=30):
    torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    s1.synchronize(); s2.synchronize(); e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
cur = torch.cuda.current_stream()
def both(fa, fb):
    def f():
        s1.wait_stream(cur); s2.wait_stream(cur); fa(); fb(); cur.wait_stream(s1); cur.wait_stream(s2)
    return f
for _ in range(3): half(b12, s1); half(b8, s2); gram(b12, s1); gram(b8, s2)
print("H12 alone  %.1f us" % timed(both(lambda: half(b12, s1), lambda: None)))
print("H8 alone   %.1f us" % timed(both(lambda: half(b8, s1), lambda: None)))
print("G12 alone  %.1f us" % timed(both(lambda: gram(b12, s1), lambda: None)))
print("G8 alone   %.1f us" % timed(both(lambda: gram(b8, s1), lambda: None)))
print("H12 || G8  %.1f us" % timed(both(lambda: half(b12, s1), lambda: gram(b8, s2))))
print("H8 || G12  %.1f us" % timed(both(lambda: half(b8, s1), lambda: gram(b12, s2))))
print("H12 || H8  %.1f us" % timed(both(lambda: half(b12, s1), lambda: half(b8, s2))))
