"""Dev probe: host profile of cNMF.consensus at the cfg-3 shape (100 000 x 2000 CSR).  Factorizes 12 replicates of
K = 9 only (short), then profiles consensus(9) twice (the second call has the caches warm)."""
import cProfile, os, pstats, shutil, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import scipy.sparse as sp
import torch
import bench
from cnmf_torch_b200 import cnmf as host
xs = sp.csr_matrix(bench._cfg3_matrix())
box = tempfile.mkdtemp(prefix="cnmf_b200_cfg3p_")
try:
    obj = host.cNMF(output_dir=box, name="p", device="cuda:0")
    obj.prepare_from_matrix(xs, [9], n_iter=12, seed=14, beta_loss="kullback-leibler", use_gpu=True)
    obj.factorize()
    obj.combine(components=[9])
    for rep in range(2):
        pr = cProfile.Profile()
        t0 = time.perf_counter()
        pr.enable()
        obj.consensus(9, density_threshold=2.0, local_neighborhood_size=0.30)
        torch.cuda.synchronize()
        pr.disable()
        print(f"consensus call {rep}: {time.perf_counter() - t0:.3f} s")
        pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
finally:
    shutil.rmtree(box, ignore_errors=True)
