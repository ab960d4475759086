"""Rows that follow the hot path inside ``cNMF.consensus``, on the device:

* ``efficient_ols_all_cols`` -- /root/reference/src/cnmf/cnmf.py:56-126, the z-score spectra regression;
* ``silhouette_score`` / ``prediction_error`` -- the two k-selection statistics, cnmf.py:1096-1110.
"""
from __future__ import annotations

import numpy as np
import scipy.sparse as sp
import torch

from . import _lib
from ._lib import check, lib, ptr


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _require_cuda(device):
    if not torch.cuda.is_available():
        raise _lib.Cnmfb200Error("CUDA device required: cnmf_torch_b200 has no CPU path")
    dev = torch.device(device)
    torch.cuda.set_device(dev)
    return dev


def efficient_ols_all_cols(X, Y, batch_size=1024, normalize_y=False, device="cuda:0"):
    """Beta = (X^T X)^-1 X^T Y for every column of Y, same signature and return as the reference
    (``batch_size`` is accepted for compatibility: the accumulation runs column-wise over the stored entries of
    Y on the device instead of over dense row batches).  With ``normalize_y`` each column of Y is centred and
    scaled by its global mean / population variance (variance floored at 1e-12) -- applied algebraically,
    X^T((Y - m)/s) = (X^T Y - (X^T 1) m^T)/s, so Y is never densified.  fp64 accumulation as in the reference;
    the K x K normal equations are solved on the host with ``np.linalg.lstsq`` like the reference."""
    L = lib()
    dev = _require_cuda(device)
    X = np.asarray(getattr(X, "values", X))
    Y = getattr(Y, "values", Y)
    n, k = X.shape
    if Y.shape[0] != n:
        raise ValueError("X and Y must have the same number of rows.")
    if k > 32:
        raise ValueError("at most 32 predictors")
    t = Y.shape[1]
    Yc = sp.csc_matrix(Y, dtype=np.float32)
    Yc.sort_indices()
    col_ptr = torch.from_numpy(Yc.indptr.astype(np.int64)).to(dev)
    row_idx = torch.from_numpy(Yc.indices.astype(np.int32)).to(dev)
    val = torch.from_numpy(Yc.data.astype(np.float32)).to(dev)
    Xd = torch.from_numpy(np.ascontiguousarray(X, dtype=np.float32)).to(dev)
    xty = torch.zeros((k, t), device=dev, dtype=torch.float64)
    ysum = torch.zeros(t, device=dev, dtype=torch.float64)
    ysq = torch.zeros(t, device=dev, dtype=torch.float64)
    check(L.cnmf_b200_ols_xty(ptr(col_ptr), ptr(row_idx), ptr(val), t, ptr(Xd), k, ptr(xty), ptr(ysum), ptr(ysq),
                              _stream()), "ols_xty")
    XtY = xty.cpu().numpy()
    X64 = X.astype(np.float64)
    XtX = X64.T @ X64                                    # K x K: the reference accumulates the same product
    if normalize_y:
        mean = ysum.cpu().numpy() / n
        var = ysq.cpu().numpy() / n - mean ** 2          # population variance (StandardScaler.var_)
        var[var < 1e-12] = 1e-12                         # cnmf.py:93-95
        XtY = (XtY - np.outer(X64.sum(axis=0), mean)) / np.sqrt(var)
    beta, *_ = np.linalg.lstsq(XtX, XtY, rcond=None)
    return beta


def silhouette_score(l2_spectra, labels0, k, device="cuda:0"):
    """mean silhouette coefficient, metric='euclidean' (cnmf.py:1097) from the device distance matrix."""
    from .consensus import local_density_device
    L = lib()
    dev = _require_cuda(device)
    l2 = l2_spectra if torch.is_tensor(l2_spectra) else torch.from_numpy(np.ascontiguousarray(l2_spectra, dtype=np.float32)).to(dev)
    M = int(l2.shape[0])
    _, D, _ = local_density_device(l2.contiguous(), 1)
    lab = torch.from_numpy(np.asarray(labels0, dtype=np.int32)).to(dev)
    cnt = torch.from_numpy(np.bincount(np.asarray(labels0), minlength=k).astype(np.int32)).to(dev)
    sil = torch.empty(M, device=dev, dtype=torch.float32)
    check(L.cnmf_b200_silhouette(ptr(D), M, M, ptr(lab), ptr(cnt), int(k), ptr(sil), _stream()), "silhouette")
    return float(sil.double().mean().item())


def prediction_error(X, usages, spectra, device="cuda:0"):
    """||X - usages @ spectra||_F^2 (cnmf.py:1100-1104) through the engine's objective kernels (fp32 GEMM path)."""
    from .engine import DenseFrobeniusEngine, FrobBatch
    eng = DenseFrobeniusEngine(X, device=device, impl=_lib.IMPL_SIMT)
    H0 = torch.from_numpy(np.ascontiguousarray(getattr(usages, "values", usages), dtype=np.float32))[None]
    W0 = torch.from_numpy(np.ascontiguousarray(getattr(spectra, "values", spectra), dtype=np.float32))[None]
    b = FrobBatch(eng, int(W0.shape[1]), H0, W0, "mu", 0.0, 0.0, 0.0, 0.0, 1e-16)
    return float(2.0 * b.objective()[0])
