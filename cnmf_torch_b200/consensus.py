"""Consensus front half on the GPU: local-density filter, batched KMeans, cluster medians.

Replaces /root/reference/src/cnmf/cnmf.py:1056-1090.  The heavy arithmetic (pairwise distances,
top-k selection, Lloyd iterations of all 10 restarts at once, medians) runs in libcnmf_b200.so; this
file holds the control flow of ``KMeans.fit`` (restart bookkeeping, stopping rules, best-of-n
selection; sklearn/cluster/_kmeans.py:1461-1563, 630-758) so that labels come out as sklearn's.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib
from ._lib import check, lib, ptr


def _stream():
    return torch.cuda.current_stream().cuda_stream


def l2_normalize_rows(spectra) -> torch.Tensor:
    """cnmf.py:1056 on the device, bit-identical to the reference's float32 pandas expression
    (squares summed in gene order in fp32).  Accepts a host array or a CUDA tensor; returns [M, G]
    float32 on the device."""
    L = lib()
    s = spectra if torch.is_tensor(spectra) else torch.from_numpy(np.ascontiguousarray(spectra, dtype=np.float32)).cuda()
    s = s.contiguous()
    out = torch.empty_like(s)
    check(L.cnmf_b200_l2_normalize(ptr(s), s.stride(0), int(s.shape[0]), int(s.shape[1]), ptr(out), out.stride(0),
                                   _stream()), "l2_normalize")
    return out


def local_density_device(l2: torch.Tensor, n_neighbors: int, return_neighbors: bool = False):
    """cnmf.py:1065-1072 on the device.  l2: [M, G] float32 contiguous CUDA tensor."""
    L = lib()
    M, G = int(l2.shape[0]), int(l2.shape[1])
    dev = l2.device
    D = torch.empty((M, M), device=dev, dtype=torch.float32)
    sq = torch.empty(M, device=dev, dtype=torch.float64)
    check(L.cnmf_b200_pairwise_dist(ptr(l2), l2.stride(0), M, G, ptr(D), M, ptr(sq), _stream()), "pairwise_dist")
    dens = torch.empty(M, device=dev, dtype=torch.float32)
    nbr = torch.empty((M, n_neighbors + 1), device=dev, dtype=torch.int32) if return_neighbors else None
    check(L.cnmf_b200_topk_density(ptr(D), M, M, int(n_neighbors), ptr(dens), ptr(nbr), _stream()), "topk_density")
    return dens, D, nbr


def _is_same_clustering(a: np.ndarray, b: np.ndarray, k: int) -> bool:
    """Same partition up to a relabelling (sklearn/cluster/_k_means_common.pyx:_is_same_clustering)."""
    mapping = np.full(k, -1, dtype=np.int64)
    for x, y in zip(a.tolist(), b.tolist()):
        if mapping[x] == -1:
            mapping[x] = y
        elif mapping[x] != y:
            return False
    return True


def kmeans_plusplus_device(Xc: torch.Tensor, k: int, n_init: int, random_state: int):
    """k-means++ seeding of all restarts on the device.  Only the random numbers are drawn on the host,
    from the same RandomState stream and in the same order as sklearn (first centre: choice(M, p);
    then uniform(size=n_trials) per further centre; _kmeans.py:236-258); their number does not depend
    on the data, so all restarts can be seeded side by side.  Returns (centers [n_init,k,G], chosen)."""
    L = lib()
    dev = Xc.device
    M, G = int(Xc.shape[0]), int(Xc.shape[1])
    T = 2 + int(np.log(k))
    rs = np.random.RandomState(random_state)
    w = np.ones(M, dtype=np.float32)
    first = np.empty(n_init, dtype=np.int32)
    rnd = np.zeros((n_init, max(k - 1, 1), T), dtype=np.float64)
    for i in range(n_init):
        first[i] = rs.choice(M, p=w / w.sum())
        for c in range(1, k):
            rnd[i, c - 1] = rs.uniform(size=T)
    first_d = torch.from_numpy(first).to(dev)
    rnd_d = torch.from_numpy(rnd).to(dev)
    chosen = torch.empty((n_init, k), device=dev, dtype=torch.int32)
    centers = torch.empty((n_init, k, G), device=dev, dtype=torch.float32)
    ws = torch.empty(int(L.cnmf_b200_kmeanspp_workspace_bytes(M, n_init, T)), device=dev, dtype=torch.uint8)
    check(L.cnmf_b200_kmeanspp(ptr(Xc), Xc.stride(0), M, G, n_init, k, T, ptr(first_d), ptr(rnd_d), ptr(chosen),
                               ptr(centers), ptr(ws), _stream()), "kmeanspp")
    return centers, chosen


def kmeans_lloyd_batched(Xc: torch.Tensor, centers0: torch.Tensor, tol: float, max_iter: int = 300):
    """All restarts of ``_kmeans_single_lloyd`` (sklearn/cluster/_kmeans.py:630-758) side by side.
    Xc: centred data [M, G]; centers0: [n_init, k, G].  Returns labels [n_init, M] int32 (host),
    inertia [n_init] float64, centers [n_init, k, G] (device), n_iter [n_init]."""
    L = lib()
    dev = Xc.device
    M, G = int(Xc.shape[0]), int(Xc.shape[1])
    n_init, k = int(centers0.shape[0]), int(centers0.shape[1])
    ld = Xc.stride(0)
    centers = centers0.contiguous().clone()
    centers_new = torch.empty_like(centers)
    labels = torch.full((n_init, M), -1, device=dev, dtype=torch.int32)
    n_changed = torch.zeros(n_init, device=dev, dtype=torch.int32)
    shift = torch.zeros((n_init, k), device=dev, dtype=torch.float32)
    counts = torch.zeros((n_init, k), device=dev, dtype=torch.int32)
    ws = torch.empty(int(L.cnmf_b200_kmeans_workspace_bytes(n_init, k, G)), device=dev, dtype=torch.uint8)
    active_h = np.ones(n_init, dtype=bool)
    active = torch.ones(n_init, device=dev, dtype=torch.uint8)
    strict = np.zeros(n_init, dtype=bool)
    n_iter = np.full(n_init, max_iter, dtype=np.int64)
    st = _stream()
    for it in range(max_iter):
        n_changed.zero_()
        check(L.cnmf_b200_kmeans_assign(ptr(Xc), ld, M, G, ptr(centers), n_init, k, ptr(active), ptr(labels),
                                        ptr(n_changed), st), "kmeans_assign")
        check(L.cnmf_b200_kmeans_update(ptr(Xc), ld, M, G, ptr(labels), n_init, k, ptr(active), ptr(centers),
                                        ptr(centers_new), ptr(shift), ptr(counts), ptr(ws), st), "kmeans_update")
        flags = torch.cat([n_changed.to(torch.float32), (counts == 0).sum(1).to(torch.float32)]).cpu().numpy()
        shift_h = shift.cpu().numpy()
        changed_h, empty_h = flags[:n_init], flags[n_init:]
        if (empty_h[active_h] > 0).any():
            _relocate_empty_clusters(Xc, labels, centers, centers_new, counts, shift, active_h & (empty_h > 0))
            shift_h = shift.cpu().numpy()
        # adopt the new centres of the restarts that are still running
        am = torch.from_numpy(active_h).to(dev)
        centers[am] = centers_new[am]
        for i in np.where(active_h)[0]:
            if changed_h[i] == 0:                                  # labels == labels_old  -> strict convergence
                strict[i] = True
                n_iter[i] = it + 1
                active_h[i] = False
            elif np.sum(shift_h[i] ** 2, dtype=np.float32) <= tol:  # center_shift_tot <= tol
                n_iter[i] = it + 1
                active_h[i] = False
        if not active_h.any():
            break
        active.copy_(torch.from_numpy(active_h.astype(np.uint8)))
    # restarts that stopped on the shift rule (or ran out of iterations) get one more E-step so that
    # labels match the final centres (_kmeans.py:741-752)
    redo = ~strict
    if redo.any():
        n_changed.zero_()
        act = torch.from_numpy(redo.astype(np.uint8)).to(dev)
        check(L.cnmf_b200_kmeans_assign(ptr(Xc), ld, M, G, ptr(centers), n_init, k, ptr(act), ptr(labels),
                                        ptr(n_changed), st), "kmeans_assign")
    nb = int(L.cnmf_b200_kmeans_inertia_blocks(M))
    part = torch.empty((n_init, nb), device=dev, dtype=torch.float64)
    check(L.cnmf_b200_kmeans_inertia(ptr(Xc), ld, M, G, ptr(centers), n_init, k, ptr(labels), ptr(part), st),
          "kmeans_inertia")
    return labels.cpu().numpy(), part.sum(1).cpu().numpy(), centers, n_iter


def _relocate_empty_clusters(Xc, labels, centers_old, centers_new, counts, shift, which):
    """Rare branch (_k_means_common.pyx:_relocate_empty_clusters_dense): an empty cluster takes the
    sample farthest from its centre.  Expressed with torch ops on the device -- it runs only when a
    restart produces an empty cluster, which well-separated consensus spectra do not."""
    for i in np.where(which)[0]:
        lab = labels[i].long()
        cnt = counts[i].clone().float()
        empty = torch.where(cnt == 0)[0]
        d = ((Xc - centers_old[i][lab]) ** 2).sum(1)
        far = torch.argsort(d, descending=True)[: len(empty)]
        sums = centers_new[i] * torch.where(cnt > 0, cnt, torch.ones_like(cnt))[:, None]
        for e, f in zip(empty.tolist(), far.tolist()):
            old = int(lab[f])
            sums[e] = Xc[f]
            sums[old] -= Xc[f]
            cnt[e] = 1
            cnt[old] -= 1
        centers_new[i] = sums * (1.0 / cnt)[:, None]
        shift[i] = torch.sqrt(((centers_old[i] - centers_new[i]) ** 2).sum(1))


def kmeans_fit(X: torch.Tensor, k: int, n_init: int = 10, random_state: int = 1, max_iter: int = 300,
               tol: float = 1e-4, centers_init: np.ndarray | None = None):
    """``KMeans(n_clusters=k, n_init=10, random_state=1).fit(X)`` (cnmf.py:1082-1083).
    ``centers_init`` ([n_init, k, G], in the CENTRED space) replaces the k-means++ seeding (tests use it to
    feed sklearn's own seeds).  Returns dict(labels (0-based, int32 host), inertia, n_iter, centers (host, original space))."""
    L = lib()
    dev = X.device
    M, G = int(X.shape[0]), int(X.shape[1])
    if k > 32:
        raise NotImplementedError("the Lloyd kernels hold up to 32 clusters")
    if M < k:
        raise ValueError(f"n_samples={M} should be >= n_clusters={k}.")
    # tolerance = mean of the per-feature variances * tol (_kmeans.py:285-293), before centring
    tol_ = float(torch.var(X.double(), dim=0, unbiased=False).mean().item()) * tol
    Xc = X.contiguous().clone()
    mean = torch.empty(G, device=dev, dtype=torch.float32)
    check(L.cnmf_b200_col_center(ptr(Xc), Xc.stride(0), M, G, ptr(mean), _stream()), "col_center")
    if centers_init is not None:
        centers0 = torch.from_numpy(np.asarray(centers_init, dtype=np.float32)).to(dev)
        n_init = int(centers0.shape[0])
    else:
        centers0, _ = kmeans_plusplus_device(Xc, k, n_init, random_state)
    labels, inertia, centers, n_iter = kmeans_lloyd_batched(Xc, centers0, tol_, max_iter)
    best = None
    for i in range(n_init):   # best-of-n with the same-clustering guard (_kmeans.py:1531-1541)
        if best is None or (inertia[i] < inertia[best] and not _is_same_clustering(labels[i], labels[best], k)):
            best = i
    return dict(labels=labels[best], inertia=float(inertia[best]), n_iter=int(n_iter[best]),
                centers=(centers[best] + mean).cpu().numpy(), all_labels=labels, all_inertia=inertia,
                all_n_iter=n_iter, best=best)


def cluster_medians(X: torch.Tensor, labels0: np.ndarray, k: int):
    """groupby(labels).median() and the row normalisation (cnmf.py:1087-1090) on the device."""
    L = lib()
    dev = X.device
    G = int(X.shape[1])
    order = np.argsort(labels0, kind="stable").astype(np.int32)
    cnt = np.bincount(labels0, minlength=k)
    offsets = np.concatenate([[0], np.cumsum(cnt)]).astype(np.int32)
    med = torch.empty((k, G), device=dev, dtype=torch.float32)
    nrm = torch.empty((k, G), device=dev, dtype=torch.float32)
    order_d = torch.from_numpy(order).to(dev)      # named: a temporary would be freed (and its block
    offsets_d = torch.from_numpy(offsets).to(dev)  # reused) before the kernel reads it
    check(L.cnmf_b200_cluster_median(ptr(X), X.stride(0), G, ptr(order_d), ptr(offsets_d), k,
                                     int(max(1, cnt.max())), ptr(med), ptr(nrm), _stream()), "cluster_median")
    return med, nrm


def consensus_front_half(merged_spectra: np.ndarray, k: int, density_threshold: float, n_neighbors: int,
                         device="cuda:0", skip_density: bool = False, local_density=None,
                         return_neighbors: bool = False, centers_init_fn=None):
    """cnmf.py:1056-1090.  Returns dict(local_density, keep, labels (1-based), median_spectra,
    l2_spectra, inertia, neighbors).  ``centers_init_fn(kept_l2_spectra) -> [n_init, k, G]`` (centred space) is a
    test hook that replaces the device seeding."""
    lib()
    if not torch.cuda.is_available():
        raise _lib.Cnmfb200Error("CUDA device required: cnmf_torch_b200 has no CPU path")
    dev = torch.device(device)
    torch.cuda.set_device(dev)
    l2 = l2_normalize_rows(torch.from_numpy(np.ascontiguousarray(merged_spectra, dtype=np.float32)).to(dev))
    dens_h, nbr_h = None, None
    keep = np.ones(l2.shape[0], dtype=bool)
    if not skip_density:
        if local_density is None:
            dens, _, nbr = local_density_device(l2, n_neighbors, return_neighbors)
            dens_h = dens.cpu().numpy()
            nbr_h = nbr.cpu().numpy() if nbr is not None else None
        else:
            dens_h = np.asarray(local_density, dtype=np.float32)
        keep = dens_h < density_threshold                        # strict, cnmf.py:1077
        if not keep.any():
            raise RuntimeError("Zero components remain after density filtering. "
                               "Consider increasing density threshold")
    kept = l2[torch.from_numpy(keep).to(dev)].contiguous()
    km = kmeans_fit(kept, k, centers_init=None if centers_init_fn is None else centers_init_fn(kept))
    _, nrm = cluster_medians(kept, km["labels"], k)
    return dict(local_density=dens_h, keep=keep, labels=km["labels"] + 1, median_spectra=nrm.cpu().numpy(),
                l2_spectra=l2.cpu().numpy(), inertia=km["inertia"], neighbors=nbr_h, n_iter=km["n_iter"])

