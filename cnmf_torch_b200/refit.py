"""``fit_H_online`` on the GPU -- the usage refit that follows the consensus front half
(/root/reference/src/cnmf/cnmf.py:260-388, called through ``refit_usage`` / ``refit_spectra`` at
cnmf.py:923-994, 1093, 1126, 1148).  Same signature and stopping rule as the reference; the arithmetic is
the online engine's own building blocks with the spectra held fixed: per chunk one fp32 GEMM x W^T, then
the inner multiplicative-update loop until the relative change of h falls below ``h_tol``.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib
from ._lib import check, lib, ptr
from .engine import pad4


def fit_H_online(X, W, H_init=None, chunk_size=5000, chunk_max_iter=200, h_tol=0.05, l1_reg_H=0.0, l2_reg_H=0.0,
                 epsilon=1e-16, device="cuda:0", seed=None, inner_batch=4):
    """Online MU for H with W fixed.  X: cells x genes (array, DataFrame or scipy sparse), W: programs x genes.
    ``H_init=None`` draws uniform(0, 1) like the reference (cnmf.py:342-343) -- from a generator seeded with
    ``seed`` when given, which the reference does not offer (its refit is not reproducible).  ``device`` must be
    a CUDA device: there is no CPU path here.  Returns H as a float32 numpy array (cells x programs)."""
    L = lib()
    if not torch.cuda.is_available() or str(device).startswith("cpu"):
        raise _lib.Cnmfb200Error("cnmf_torch_b200.fit_H_online runs on a CUDA device only")
    dev = torch.device("cuda:0" if str(device) == "cuda" else device)
    torch.cuda.set_device(dev)
    X = getattr(X, "values", X)
    W = np.asarray(getattr(W, "values", W), dtype=np.float32)
    H_init = getattr(H_init, "values", H_init)
    sparse = hasattr(X, "toarray")
    if sparse:
        X = X.tocsr()      # densified one chunk at a time below (the reference densifies the whole matrix, cnmf.py:329-330;
                           # TPM over all genes x 100 000 cells would be 8 GB of host memory) -- ON THE DEVICE: the CSR
                           # arrays are uploaded once (a twelfth of the dense bytes at 8 % density) and each chunk is
                           # scattered into the dense buffer there; numpy's toarray + a pageable H2D of 800 MB per refit
                           # were most of a consensus call at 100 000 cells
    else:
        X = np.ascontiguousarray(X, dtype=np.float32)
    n, g = X.shape
    k = W.shape[0]
    Kp = pad4(k)
    if Kp > _lib.MAX_KP:
        raise ValueError(f"{k} programs exceed the engine limit {_lib.MAX_KP}")
    st = torch.cuda.current_stream().cuda_stream
    ldg = pad4(g)
    Wd = torch.zeros((Kp, ldg), device=dev, dtype=torch.float32)
    Wd[:k, :g] = torch.from_numpy(np.array(W, dtype=np.float32, copy=True)).to(dev)   # (W may be a read-only view)
    if H_init is None:
        if isinstance(seed, torch.Generator):
            gen = seed                    # a caller's generator: consecutive refits continue one stream
        else:
            gen = torch.Generator(device="cpu")
            if seed is not None:
                gen.manual_seed(int(seed))
            else:
                gen.seed()
        h0 = torch.rand((n, k), generator=gen, dtype=torch.float32)
    else:
        h0 = torch.from_numpy(np.asarray(H_init, dtype=np.float32)).clamp(min=0.0)   # cnmf.py:345
    ldn = pad4(n)
    Ht = torch.zeros((Kp, ldn), device=dev, dtype=torch.float32)
    Ht[:k, :n] = h0.to(dev).t()
    gram = torch.zeros((1, Kp, Kp), device=dev, dtype=torch.float32)
    ws_g = torch.empty(max(256, int(L.cnmf_b200_gram_workspace_bytes(1, Kp, g))), device=dev, dtype=torch.uint8)
    check(L.cnmf_b200_gram(ptr(Wd), ldg, 1, Kp, g, ptr(gram), ptr(ws_g), st), "gram")       # W W^T, cnmf.py:346
    cmax = min(int(chunk_size), n)
    ldp = pad4(cmax)
    Pt = torch.zeros((Kp, ldp), device=dev, dtype=torch.float32)
    xbuf = torch.zeros((cmax, ldg), device=dev, dtype=torch.float32)
    iters = torch.zeros(1, device=dev, dtype=torch.int32)
    ws = torch.empty(max(256, int(L.cnmf_b200_online_loop_workspace_bytes(1))), device=dev, dtype=torch.uint8)
    Hs = torch.zeros((Kp, ldp), device=dev, dtype=torch.float32)   # chunk-local state: 16-byte aligned columns
    if sparse:
        indptr = np.asarray(X.indptr, dtype=np.int64)
        d_col = torch.from_numpy(np.asarray(X.indices, dtype=np.int64)).to(dev)
        d_val = torch.from_numpy(np.asarray(X.data, dtype=np.float32)).to(dev)
        d_row = torch.from_numpy(np.repeat(np.arange(n, dtype=np.int64), np.diff(indptr))).to(dev)
    for lo in range(0, n, int(chunk_size)):
        c = min(int(chunk_size), n - lo)
        if sparse:
            a, b_ = int(indptr[lo]), int(indptr[lo + c])
            xbuf.zero_()
            xbuf.index_put_((d_row[a:b_] - lo, d_col[a:b_]), d_val[a:b_], accumulate=True)   # (duplicates sum, as toarray)
        else:
            xbuf[:c, :g] = torch.from_numpy(X[lo:lo + c]).to(dev)                         # chunk upload (cnmf.py:351)
        check(L.cnmf_b200_gemm_nt(ptr(Wd), ldg, Kp, ptr(xbuf), ldg, c, g, ptr(Pt), ldp, st), "gemm x W^T")
        Hs[:, :c] = Ht[:, lo:lo + c]
        # the chunk's inner loop to convergence in one launch (stops on the device, cnmf.py:365-378)
        check(L.cnmf_b200_online_loop(ptr(Pt), ldp, ptr(Hs), ldp, ptr(gram), None, c, 1, Kp, float(l1_reg_H),
                                      float(l2_reg_H), float(epsilon), float(h_tol), int(chunk_max_iter), ptr(iters),
                                      None, ptr(ws), st), "online_loop")
        Ht[:, lo:lo + c] = Hs[:, :c]
    return Ht[:k, :n].t().contiguous().cpu().numpy()
