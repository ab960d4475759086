import sys; sys.path.insert(0,".")
import numpy as np
from cnmf_torch_b200.engine import DenseFrobeniusEngine
from cnmf_torch_b200.synth import ledger_seeds, nb_counts
from cnmf_torch_b200 import _lib
x = nb_counts(2700, 2000, seed=14)
seeds = [int(s) for s in ledger_seeds(100, 14)]
for impl in (_lib.IMPL_TCGEN05, _lib.IMPL_SIMT):
    eng = DenseFrobeniusEngine(x, impl=impl)
    a = eng.factorize(10, seeds=seeds, algo="hals", tol=0.0, max_iter=60)
    obj = np.stack([v for _, v in a.trace]); its=[i for i,_ in a.trace]
    d = np.diff(obj, axis=0)/obj[:-1]
    bad = np.argwhere(d > 2e-5)
    print("impl", impl, "n_iter", np.unique(a.n_iter, return_counts=True), "max rel increase", d.max(), "bad", bad[:10].tolist())
    for (c, r) in bad[:3]:
        print("  replicate", r, [f"{its[i]}:{obj[i, r]:.1f}" for i in range(len(its))])
