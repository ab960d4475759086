"""Dev probe: stage times of the consensus front half at cfg-4 shape."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from cnmf_torch_b200 import consensus as cs
from cnmf_torch_b200.synth import replicate_spectra
s = replicate_spectra(20, 200, 2000, seed=14, outlier_frac=0.03)
def T(f, *a, **k):
    torch.cuda.synchronize(); t0 = time.perf_counter(); r = f(*a, **k); torch.cuda.synchronize(); return r, (time.perf_counter() - t0) * 1e3
for rep in range(2):
    sd, t_up = T(lambda: torch.from_numpy(s).cuda())
    l2, t_l2 = T(cs.l2_normalize_rows, sd)
    (dens, D, _), t_dens = T(cs.local_density_device, l2, 60)
    keep = dens.cpu().numpy() < 0.5
    kept = l2[torch.from_numpy(keep).cuda()].contiguous()
    Xc = kept.clone(); mean = torch.empty(2000, device="cuda")
    from cnmf_torch_b200._lib import lib, check, ptr
    _, t_center = T(lambda: check(lib().cnmf_b200_col_center(ptr(Xc), Xc.stride(0), Xc.shape[0], 2000, ptr(mean), torch.cuda.current_stream().cuda_stream)))
    t_down = 0.0
    (c0, _), t_seed = T(cs.kmeans_plusplus_device, Xc, 20, 10, 1)
    (lab, inertia, cen, nit), t_lloyd = T(cs.kmeans_lloyd_batched, Xc, c0, 1e-9, 300)
    _, t_med = T(cs.cluster_medians, kept, lab[0].astype(np.int64), 20)
    print(f"rep{rep}: l2 {t_l2:.1f} upload {t_up:.1f} dist+topk {t_dens:.1f} center {t_center:.1f} download {t_down:.1f} "
          f"seed(device) {t_seed:.1f} lloyd {t_lloyd:.1f} (iters {nit.tolist()}) median {t_med:.1f} ms")
