"""Launch the bench's dominant kernel (H half-step of the Kp=12 stack, K=9..12 x 100 replicates, cfg-2 X) a few
times: the target of the `ncu --set full` capture kept under profiles/."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cnmf_torch_b200 import _lib
from cnmf_torch_b200.engine import DenseFrobeniusEngine, FrobBatch
from cnmf_torch_b200.synth import ledger_seeds, nb_counts
lib = _lib.lib()
x = nb_counts(2700, 2000, seed=14)
eng = DenseFrobeniusEngine(x)
seeds = ledger_seeds(600, 14).reshape(6, 100)
groups = [(k,) + tuple(eng.make_init(k, seeds[k - 7])) for k in (9, 10, 11, 12)]
b = FrobBatch(eng, None, None, None, "hals", 0.0, 0.0, 0.0, 0.0, 1e-16, groups=groups)
p = b.p
st = torch.cuda.current_stream().cuda_stream
Wop = b.Wall.clone(); _lib.check(lib.cnmf_b200_round_tf32(Wop.data_ptr(), Wop.numel(), st))
Hs, Hop = b.Ht.clone(), b.Ht.clone()
for _ in range(int(os.environ.get("N", "6"))):
    _lib.check(lib.cnmf_b200_dense_half_step(p.X, p.ldx, p.n_cells, p.n_genes, Wop.data_ptr(), p.ldw, Hs.data_ptr(),
                                             Hop.data_ptr(), p.ldh, p.gramW, None, p.R, p.Kp, p.algo, 0.0, 0.0, p.eps, None, st))
torch.cuda.synchronize()
print("ok")
