"""Deterministic synthetic inputs of the benchmark shapes.

Negative-binomial counts from K_true gene programs, then the reference's normalisation: each
gene divided by its standard deviation (ddof=1), no centring
(/root/reference/src/cnmf/cnmf.py:671-681).  Master seed 14 is the seed used throughout the
reference's docs/tests.  Host-side numpy: this produces inputs, it is not on the hot path.
"""
from __future__ import annotations

import numpy as np


def nb_counts(n_cells: int, n_genes: int, k_true: int = 10, seed: int = 14, libsize: float = 2500.0,
              dispersion: float = 2.0, chunk: int = 20000):
    """Variance-scaled NB count matrix, float32 dense [n_cells x n_genes] (zeros preserved)."""
    rng = np.random.default_rng(seed)
    programs = rng.gamma(0.3, 1.0, size=(k_true, n_genes))
    programs /= programs.sum(1, keepdims=True)
    out = np.empty((n_cells, n_genes), dtype=np.float32)
    for lo in range(0, n_cells, chunk):
        hi = min(n_cells, lo + chunk)
        usage = rng.dirichlet(np.full(k_true, 0.2), size=hi - lo)
        lib = rng.lognormal(np.log(libsize), 0.35, size=(hi - lo, 1))
        mu = lib * (usage @ programs)
        lam = rng.gamma(dispersion, mu / dispersion)
        out[lo:hi] = rng.poisson(lam)
    # guard: no all-zero gene / cell (the reference drops/raises on them, cnmf.py:688-691)
    zero_g = out.sum(0) == 0
    if zero_g.any():
        out[rng.integers(0, n_cells, zero_g.sum()), np.where(zero_g)[0]] = 1.0
    zero_c = out.sum(1) == 0
    if zero_c.any():
        out[np.where(zero_c)[0], rng.integers(0, n_genes, zero_c.sum())] = 1.0
    std = out.std(axis=0, ddof=1, dtype=np.float64)
    std[std == 0] = 1.0
    out /= std.astype(np.float32)
    return out


def replicate_spectra(k: int, n_rep: int, n_genes: int, seed: int = 14, outlier_frac: float = 0.02):
    """cfg-4 style stacked spectra: jittered copies of k base programs plus outlier rows."""
    rng = np.random.default_rng(seed)
    base = rng.gamma(0.3, 1.0, size=(k, n_genes)) + 1e-3
    rows = [base[rng.permutation(k)] * rng.lognormal(0.0, 0.1, size=(k, n_genes)) for _ in range(n_rep)]
    s = np.concatenate(rows, 0)
    n_out = max(1, int(outlier_frac * s.shape[0]))
    idx = rng.choice(s.shape[0], n_out, replace=False)
    s[idx] = rng.gamma(0.3, 1.0, size=(n_out, n_genes)) + 1e-3
    return s.astype(np.float32)


def ledger_seeds(n_runs: int, seed: int = 14):
    """Replicate seeds exactly as the reference draws them (cnmf.py:740-741)."""
    rs = np.random.RandomState(seed)
    return rs.randint(low=1, high=(2 ** 31) - 1, size=n_runs)
