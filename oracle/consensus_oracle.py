"""CPU oracle for the consensus front half -- TEST INFRASTRUCTURE ONLY.

Restates /root/reference/src/cnmf/cnmf.py:1053-1090 with the same library calls the
reference makes (sklearn ``euclidean_distances`` / ``KMeans``, ``np.argpartition``, pandas
groupby-median), on the fp32 dtype flow of this fork.  Pinned against
outputs of the reference's own ``cNMF.consensus`` run in the build container:
tests/golden/consensus_*.npz (made by tests/golden/make_golden.py).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs import this.
"""
from __future__ import annotations

import numpy as np
import pandas as pd
from sklearn.cluster import KMeans
from sklearn.metrics.pairwise import euclidean_distances


def l2_normalize(spectra: np.ndarray) -> np.ndarray:
    """cnmf.py:1056  (merged.T / sqrt((merged**2).sum(axis=1))).T, dtype preserved."""
    df = pd.DataFrame(spectra)
    return ((df.T / np.sqrt((df ** 2).sum(axis=1))).T).values


def n_neighbors_for(local_neighborhood_size: float, n_rows: int, k: int) -> int:
    return int(local_neighborhood_size * n_rows / k)  # cnmf.py:1053


def local_density(l2_spectra: np.ndarray, n_neighbors: int):
    """cnmf.py:1065-1072.  Returns (density[rows], topics_dist, neighbour index sets)."""
    topics_dist = euclidean_distances(l2_spectra)
    part = np.argpartition(topics_dist, n_neighbors + 1)[:, :n_neighbors + 1]
    d = topics_dist[np.arange(topics_dist.shape[0])[:, None], part]
    return d.sum(1) / n_neighbors, topics_dist, part


def consensus_front_half(merged_spectra: np.ndarray, k: int, density_threshold: float = 0.5,
                         local_neighborhood_size: float = 0.30):
    """cnmf.py:1053-1090.  Returns dict(local_density, keep, labels (1-based, kept rows),
    median_spectra (k x G, row-normalised), l2_spectra)."""
    n_nb = n_neighbors_for(local_neighborhood_size, merged_spectra.shape[0], k)
    l2 = l2_normalize(merged_spectra)
    dens, _, _ = local_density(l2, n_nb)
    keep = dens < density_threshold  # strict, cnmf.py:1077
    l2_kept = l2[keep]
    if l2_kept.shape[0] == 0:
        raise RuntimeError("Zero components remain after density filtering. "
                           "Consider increasing density threshold")
    km = KMeans(n_clusters=k, n_init=10, random_state=1)  # cnmf.py:1082
    km.fit(l2_kept)
    labels = km.labels_ + 1
    df = pd.DataFrame(l2_kept)
    med = df.groupby(pd.Series(labels)).median()  # cnmf.py:1087
    med = (med.T / med.sum(axis=1)).T  # cnmf.py:1090
    return dict(local_density=dens, keep=keep, labels=labels, median_spectra=med.values,
                l2_spectra=l2, inertia=float(km.inertia_), centers=km.cluster_centers_,
                n_iter=int(km.n_iter_))
