"""Dev probe: ms per batched KL iteration of the two kernel families (slab / gather) at the cfg-3 shape.
    python scripts/kl_bench.py [cells ...]      env: KS=9 (K), RS="100,13" (replicates), IMPLS="slab,gather"
"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, scipy.sparse as sp, torch
from cnmf_torch_b200.kl_engine import SparseKLEngine
from cnmf_torch_b200.synth import nb_counts

cells = [int(a) for a in sys.argv[1:]] or [20000, 100000]
K = int(os.environ.get("KS", "9"))
RS = [int(r) for r in os.environ.get("RS", "100,13").split(",")]
IMPLS = os.environ.get("IMPLS", "slab,gather").split(",")
out = []
for n in cells:
    x = sp.csr_matrix(nb_counts(n, 2000, seed=14, libsize=170.0))
    for impl in IMPLS:
        eng = SparseKLEngine(x, impl=impl, init_device="cuda")
        for R in RS:
            b = eng.new_batch(K, seeds=list(range(R)))
            b.iterate(2)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            b.iterate(10)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 10
            loss = b.read_loss()
            flops = 2.0 * 2 * eng.nnz * R * 2 * K     # 2 half-steps x (s and numerator) x K FMAs per entry and replicate
            rec = dict(cells=n, nnz=eng.nnz, impl=impl, K=K, R=R, ms_per_iter=ms, rep_iters_per_s=R / ms * 1e3,
                       tflops_fp32_algorithmic=flops / ms / 1e9, loss0=float(loss[0]), ws_gib=b.ws.numel() / 2 ** 30)
            print(json.dumps(rec), flush=True)
            out.append(rec)
            del b
        del eng
        torch.cuda.empty_cache()
