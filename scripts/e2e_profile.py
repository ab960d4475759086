"""Dev probe: where the end-to-end (host buffers in, spectra out) time of bench.py's e2e step goes."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from cnmf_torch_b200.engine import DenseFrobeniusEngine, FrobBatch, run_batches, pad4
from cnmf_torch_b200.synth import ledger_seeds, nb_counts
x = torch.from_numpy(nb_counts(2700, 2000, seed=14)).pin_memory()
KS = list(range(7, 13)); seeds = ledger_seeds(600, 14).reshape(6, 100)
def T(): torch.cuda.synchronize(); return time.perf_counter()
for rep in range(6):
    dev_init = 'cuda' if rep >= 3 else 'cpu'
    t0 = T(); eng = DenseFrobeniusEngine(x, init_device=dev_init); t1 = T()
    inits = {k: eng.make_init(k, seeds[i]) for i, k in enumerate(KS)}; t2 = T()
    stacks = {}
    for k in KS: stacks.setdefault(pad4(k), []).append((k,) + tuple(inits[k]))
    batches = [FrobBatch(eng, None, None, None, "hals", 0, 0, 0, 0, 1e-16, groups=g) for g in stacks.values()]; t3 = T()
    streams = [torch.cuda.Stream() for _ in batches]
    res = run_batches(batches, 0.0, 100, 10, False, streams); t4 = T()
    outs = [b.results_by_group(False) for b in batches]; t5 = T()
    print(f"rep{rep} [{dev_init} init]: engine ctor {1e3*(t1-t0):.1f}  init draws {1e3*(t2-t1):.1f}  upload+prepare {1e3*(t3-t2):.1f}  "
          f"100 iterations (lock-step, 10 checks) {1e3*(t4-t3):.1f}  spectra D2H {1e3*(t5-t4):.1f}  total {1e3*(t5-t0):.1f} ms")
