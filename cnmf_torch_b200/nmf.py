"""``run_nmf`` -- drop-in for ``from nmf import run_nmf`` at /root/reference/src/cnmf/cnmf.py:17,
called as ``(usages, spectra, err) = run_nmf(X, **nmf_kwargs)`` at cnmf.py:819.

Same keyword arguments (cnmf.py:757-771 + ``random_state`` / ``n_components`` set at cnmf.py:885-886),
same return order and types: ``H`` (N x K usages), ``W`` (K x G spectra) as host float32 arrays and
the scalar ``err = sqrt(2 * objective)``.  Everything runs on the GPU through libcnmf_b200.so; there is
no CPU path.  ``use_gpu`` (False by default in nmf-torch and in every YAML ``cnmf prepare`` writes without
``--use_gpu``, cnmf.py:461, 770, 1420) keeps the one meaning it can have here: which torch generator draws
the random init -- the host generator under False, so a replicate starts exactly where the reference's CPU
run starts, the device generator under True.

``run_nmf_batch`` is the replicate-batched form the ``cNMF.factorize`` override uses: many seeds of one
K share every pass over X.
"""
from __future__ import annotations

import numpy as np

from . import _lib

_IMPL = {"auto": _lib.IMPL_AUTO, "simt": _lib.IMPL_SIMT, "tcgen05": _lib.IMPL_TCGEN05}


ITAKURA_SAITO_REFUSAL = (
    "beta_loss='itakura-saito' is not implemented by cnmf_torch_b200: the divergence is +inf wherever X is zero, "
    "and a normalised count matrix always has zeros (scikit-learn refuses beta_loss <= 0 on such a matrix for the "
    "same reason). Use 'frobenius' or 'kullback-leibler'.")


def _reg(alpha_W, l1_ratio_W, alpha_H, l1_ratio_H):
    # L1 strength = alpha * l1_ratio, L2 strength = alpha * (1 - l1_ratio) (sklearn convention the
    # reference's kwargs follow, cnmf.py:757-761; W = spectra, H = usages in nmf-torch naming)
    return dict(l1_H=alpha_H * l1_ratio_H, l2_H=alpha_H * (1.0 - l1_ratio_H),
                l1_W=alpha_W * l1_ratio_W, l2_W=alpha_W * (1.0 - l1_ratio_W))


def nndsvd_init(engine, k, seeds, init):
    """NNDSVD family (reachable from the cnmf CLI, cnmf.py:1427): nmf-torch delegates to scikit-learn's
    ``_initialize_nmf`` on the host (sklearn/decomposition/_nmf.py:214-366); so does this.  'nndsvd' and
    'nndsvda' do not depend on the seed: the replicates of a batch then start identically."""
    import torch
    from sklearn.decomposition._nmf import _initialize_nmf
    x = engine.host_matrix()
    hs, ws = [], []
    cache = {}
    for s in seeds:
        key = int(s) if init == "nndsvdar" else 0
        if key not in cache:
            Wsk, Hsk = _initialize_nmf(x, k, init=init, random_state=int(s))   # sklearn: X ~ W_sk H_sk
            cache[key] = (torch.from_numpy(np.ascontiguousarray(Wsk, dtype=np.float32)),
                          torch.from_numpy(np.ascontiguousarray(Hsk, dtype=np.float32)))
        hs.append(cache[key][0])      # usages   (N x K)  = sklearn's W
        ws.append(cache[key][1])      # spectra  (K x G)  = sklearn's H
    return torch.stack(hs), torch.stack(ws)


def make_engine(X, beta_loss="frobenius", device="cuda:0", impl="auto", init_device="cpu"):
    """Ingest X once (cf. the per-replicate densify + upload at cnmf.py:817-819).  ``init_device``: which torch
    generator draws the 'random' init -- "cpu" (the reference with use_gpu=False, and the oracle) or "cuda" (the
    reference with use_gpu=True); see engine.draw_init."""
    if beta_loss in ("frobenius", 2, 2.0):
        from .engine import DenseFrobeniusEngine
        return DenseFrobeniusEngine(X, device=device, impl=_IMPL.get(impl, _lib.IMPL_AUTO), init_device=init_device)
    if beta_loss in ("kullback-leibler", 1, 1.0):
        from .kl_engine import SparseKLEngine
        return SparseKLEngine(X, device=device, init_device=init_device, impl=impl if impl in ("slab", "gather") else None)
    if beta_loss in ("itakura-saito", 0, 0.0):
        raise NotImplementedError(ITAKURA_SAITO_REFUSAL)
    raise ValueError(f"beta_loss={beta_loss!r}: expected 'frobenius', 'kullback-leibler' or 'itakura-saito'")


def run_nmf_batch(engine, n_components, seeds, init="random", algo="mu", mode="batch", tol=1e-4,
                  alpha_W=0.0, l1_ratio_W=0.0, alpha_H=0.0, l1_ratio_H=0.0, batch_max_iter=500,
                  online_max_pass=20, online_chunk_size=5000, online_chunk_max_iter=200, online_h_tol=0.05,
                  online_w_tol=0.05, return_H=False, H0=None, W0=None):
    """R = len(seeds) replicates of one K on a resident engine. Returns engine.BatchResult."""
    if init not in ("random", "nndsvd", "nndsvda", "nndsvdar") and (H0 is None or W0 is None):
        raise ValueError(f"init={init!r}: expected 'random' or 'nndsvd[a|ar]'")
    if init != "random" and (H0 is None or W0 is None):
        H0, W0 = nndsvd_init(engine, int(n_components), seeds, init)
    if mode == "online" and not hasattr(engine, "factorize_online"):
        # sufficient statistics (A = h^T h, B = h^T x) exist for the Frobenius loss only: nmf-torch itself
        # switches beta != 2 to batch mode, and cNMF always passes mode='online' (cnmf.py:765)
        print("Online mode only works for beta_loss='frobenius'. Switch to batch mode.")
        mode = "batch"
    if algo not in ("mu", "hals"):
        raise NotImplementedError(f"algo={algo!r}: 'mu' and 'hals' are implemented")
    reg = _reg(alpha_W, l1_ratio_W, alpha_H, l1_ratio_H)
    if mode == "batch":
        return engine.factorize(int(n_components), seeds=seeds, H0=H0, W0=W0, algo=algo, tol=tol,
                                max_iter=int(batch_max_iter), return_H=return_H, **reg)
    if mode == "online":
        return engine.factorize_online(int(n_components), seeds=seeds, H0=H0, W0=W0, tol=tol,
                                       max_pass=int(online_max_pass), chunk_size=int(online_chunk_size),
                                       chunk_max_iter=int(online_chunk_max_iter), h_tol=online_h_tol,
                                       w_tol=online_w_tol, return_H=return_H, **reg)
    raise ValueError(f"mode={mode!r}")


def run_nmf(X, n_components, init="random", beta_loss="frobenius", algo="mu", mode="batch", tol=1e-4,
            n_jobs=-1, random_state=0, use_gpu=False, alpha_W=0.0, l1_ratio_W=0.0, alpha_H=0.0,
            l1_ratio_H=0.0, fp_precision="float", batch_max_iter=500, online_max_pass=20,
            online_chunk_size=5000, online_chunk_max_iter=200, online_h_tol=0.05, online_w_tol=0.05,
            online_usage_tol=None, online_spectra_tol=None, device="cuda:0", impl="auto", init_device=None,
            **unknown):
    """One replicate, the reference's call shape.  Returns (H, W, err).

    ``use_gpu`` selects the init stream, not the processor (always the GPU): False -> torch's host generator,
    True -> the device generator, as nmf-torch draws on the device it computes on.  ``init_device`` ("cpu" /
    "cuda", not a reference kwarg) overrides that choice.  ``n_jobs`` (host threads in nmf-torch) has nothing to
    set here and is accepted."""
    if unknown:
        raise TypeError(f"run_nmf got unexpected keyword arguments {sorted(unknown)}")
    if init_device is None:
        init_device = "cuda" if use_gpu else "cpu"
    if fp_precision not in ("float", np.float32, "float32"):
        raise NotImplementedError("state precision is float32 (nmf-torch's default fp_precision='float')")
    if online_usage_tol is not None:
        online_h_tol = online_usage_tol
    if online_spectra_tol is not None:
        online_w_tol = online_spectra_tol
    if beta_loss in ("kullback-leibler", 1, 1.0) and algo != "mu":
        algo = "mu"  # only MU exists for beta != 2 (nmf-torch falls back the same way)
    eng = make_engine(X, beta_loss=beta_loss, device=device, impl=impl, init_device=init_device)
    res = run_nmf_batch(eng, n_components, [int(random_state)], init=init, algo=algo, mode=mode, tol=tol,
                        alpha_W=alpha_W, l1_ratio_W=l1_ratio_W, alpha_H=alpha_H, l1_ratio_H=l1_ratio_H,
                        batch_max_iter=batch_max_iter, online_max_pass=online_max_pass,
                        online_chunk_size=online_chunk_size, online_chunk_max_iter=online_chunk_max_iter,
                        online_h_tol=online_h_tol, online_w_tol=online_w_tol, return_H=True)
    return res.H[0], res.W[0], float(res.err[0])
