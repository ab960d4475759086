"""Scale check of BASELINE cfg-3 as ONE of 8 workers sees it: 100 000 cells x 2000 genes CSR (~7 % dense), K = 5..13 x 100
replicates in the ledger, beta_loss = kullback-leibler; this process is worker 0 of 8 (113 of the 900 jobs, 12-13 per K).
Prints the wall time of factorize (device work + spectra files) and of consensus for one K on the replicates present."""
import os, sys, time, tempfile, shutil
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, scipy.sparse as sp, torch
from cnmf_torch_b200 import cnmf as host
from cnmf_torch_b200.synth import nb_counts
N = int(os.environ.get("N", "100000"))
t0 = time.perf_counter()
x = sp.csr_matrix(nb_counts(N, 2000, seed=14, libsize=170.0))
print(f"synthetic {x.shape}, density {x.nnz / (x.shape[0] * x.shape[1]):.3f}, nnz {x.nnz} ({time.perf_counter() - t0:.1f} s to generate)")
tmp = tempfile.mkdtemp(prefix="cfg3_")
try:
    obj = host.cNMF(output_dir=tmp, name="cfg3")
    obj.prepare_from_matrix(x, list(range(5, 14)), n_iter=100, seed=14, beta_loss="kullback-leibler")
    torch.cuda.synchronize(); t0 = time.perf_counter()
    obj.factorize(worker_i=0, total_workers=8)
    torch.cuda.synchronize(); t1 = time.perf_counter()
    print(f"factorize, worker 0 of 8 (113 jobs, KL MU, tol 1e-4, <= 500 iterations): {t1 - t0:.1f} s, "
          f"peak device memory {torch.cuda.max_memory_allocated() / 2**30:.2f} GiB")
    obj.combine(components=9, skip_missing_files=True)
    t2 = time.perf_counter()
    med, labels, dens = obj.consensus(9, density_threshold=2.0, seed=1)
    t3 = time.perf_counter()
    print(f"combine k=9 {t2 - t1:.1f} s; consensus k=9 on {len(labels) // 9} replicates incl. usage refit on {N} cells: {t3 - t2:.1f} s; "
          f"consensus spectra row sums {np.round(med.sum(axis=1).values, 4)}")
finally:
    shutil.rmtree(tmp, ignore_errors=True)
