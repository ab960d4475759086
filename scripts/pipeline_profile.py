"""Dev probe: where the wall time of factorize + combine + consensus goes at cfg-2 (BASELINE metric 2)."""
import cProfile, io, os, pstats, shutil, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from cnmf_torch_b200 import cnmf as host
from cnmf_torch_b200.synth import nb_counts

x = nb_counts(bench.N_CELLS, bench.N_GENES, seed=14)
for rep in range(2):
    tmp = tempfile.mkdtemp(prefix="cnmf_prof_")
    obj = host.cNMF(output_dir=tmp, name="cfg2", device="cuda:0")
    obj.prepare_from_matrix(x, bench.KS, n_iter=bench.REPLICATES, seed=14, beta_loss="frobenius", use_gpu=True)
    pr = cProfile.Profile()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    if rep == 1: pr.enable()
    obj.factorize(nmf_kwargs_override=dict(mode="batch", algo="hals", batch_max_iter=500))
    torch.cuda.synchronize(); t1 = time.perf_counter()
    obj.combine(); t2 = time.perf_counter()
    for k in bench.KS:
        obj.consensus(k, density_threshold=0.5, local_neighborhood_size=0.30)
    torch.cuda.synchronize(); t3 = time.perf_counter()
    if rep == 1: pr.disable()
    print(f"rep {rep}: factorize {t1-t0:.3f}s combine {t2-t1:.3f}s consensus {t3-t2:.3f}s total {t3-t0:.3f}s")
    if rep == 1:
        s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(45); print(s.getvalue()[:9000])
    shutil.rmtree(tmp, ignore_errors=True)
