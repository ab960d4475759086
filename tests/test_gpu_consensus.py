"""GPU parity of the consensus kernel set (through the C ABI) against the oracle (= the reference's
own sklearn/numpy/pandas calls) and against goldens produced by the reference's cNMF.consensus.

Bar (north_star): kNN index sets and cluster labels exact; local density / medians at fp32 round-off
(their last bit depends on numpy's summation order, which is not part of the contract)."""
import os
import time

import numpy as np
import pytest
import torch

from oracle import consensus_oracle as co

pytestmark = pytest.mark.gpu


def _sklearn_seeds(k, n_init=10, random_state=1):
    """centers_init_fn for the strict variant of a test: sklearn's own k-means++ seeds (host), all restarts
    from one RandomState in restart order, computed on the centred data exactly as KMeans.fit does."""
    def fn(kept):
        from sklearn.cluster._kmeans import _kmeans_plusplus
        from sklearn.utils.extmath import row_norms
        x = kept.cpu().numpy()
        xc = x - x.mean(axis=0)
        rs = np.random.RandomState(random_state)
        w = np.ones(xc.shape[0], dtype=xc.dtype)
        xsq = row_norms(xc, squared=True)
        return np.stack([_kmeans_plusplus(xc, k, xsq, w, rs)[0] for _ in range(n_init)])
    return fn


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name), allow_pickle=True)


def _spectra(k, n_rep, g, seed):
    from cnmf_torch_b200.synth import replicate_spectra
    return replicate_spectra(k, n_rep, g, seed=seed, outlier_frac=0.03)


@pytest.mark.parametrize("k,n_rep,g", [(5, 30, 80), (7, 60, 333), (9, 100, 2000)])
def test_pairwise_distances_match_sklearn(k, n_rep, g):
    from sklearn.metrics.pairwise import euclidean_distances
    from cnmf_torch_b200 import consensus as cs
    l2d = cs.l2_normalize_rows(_spectra(k, n_rep, g, 3))
    l2 = l2d.cpu().numpy()
    np.testing.assert_array_equal(l2, co.l2_normalize(_spectra(k, n_rep, g, 3)))   # bit-identical to pandas
    ref = euclidean_distances(l2)
    _, D, _ = cs.local_density_device(l2d, 3)
    got = D.cpu().numpy()
    # fp64 accumulation then one rounding to fp32: equal except where the fp64 sums straddle a rounding
    # boundary (different summation order than OpenBLAS) -- at most 1 ulp and vanishingly rare
    neq = got != ref
    assert neq.mean() < 1e-5, f"{neq.sum()} of {neq.size} entries differ"
    np.testing.assert_allclose(got, ref, rtol=3e-7, atol=1e-7)
    assert np.all(np.diag(got) == 0.0) and np.array_equal(got, got.T)


@pytest.mark.parametrize("k,n_rep,g,lns", [(5, 30, 80, 0.3), (8, 50, 300, 0.3), (20, 200, 500, 0.3), (4, 25, 64, 0.1)])
def test_local_density_and_knn_sets(k, n_rep, g, lns):
    from cnmf_torch_b200 import consensus as cs
    s = _spectra(k, n_rep, g, 5)
    l2d = cs.l2_normalize_rows(s)
    l2 = l2d.cpu().numpy()
    n = co.n_neighbors_for(lns, s.shape[0], k)
    dens_ref, dist_ref, part_ref = co.local_density(l2, n)
    dens, D, nbr = cs.local_density_device(l2d, n, return_neighbors=True)
    np.testing.assert_allclose(dens.cpu().numpy(), dens_ref, rtol=2e-6)
    nbr = nbr.cpu().numpy()
    Dh = D.cpu().numpy()
    for i in range(s.shape[0]):
        a, b = set(nbr[i].tolist()), set(part_ref[i].tolist())
        if a != b:   # only a tie at the selection boundary may swap members: same distance multiset
            np.testing.assert_array_equal(np.sort(Dh[i, sorted(a)]), np.sort(dist_ref[i, sorted(b)]))
        assert len(a) == n + 1 and i in a


@pytest.mark.parametrize("seeding", ["device", "sklearn"])
@pytest.mark.parametrize("k,n_rep,g", [(5, 30, 80), (8, 50, 300), (13, 100, 700)])
def test_kmeans_labels_match_sklearn(k, n_rep, g, seeding):
    from sklearn.cluster import KMeans
    from cnmf_torch_b200 import consensus as cs
    l2d = cs.l2_normalize_rows(_spectra(k, n_rep, g, 11))
    l2 = l2d.cpu().numpy()
    ref = KMeans(n_clusters=k, n_init=10, random_state=1).fit(l2.copy())
    out = cs.kmeans_fit(l2d, k, centers_init=None if seeding == "device" else _sklearn_seeds(k)(l2d))
    np.testing.assert_array_equal(out["labels"], ref.labels_)
    assert abs(out["inertia"] - ref.inertia_) / ref.inertia_ < 1e-5
    np.testing.assert_allclose(out["centers"], ref.cluster_centers_, rtol=1e-4, atol=1e-6)


@pytest.mark.parametrize("k,n_rep,g", [(5, 30, 80), (8, 50, 300), (20, 200, 2000)])
def test_device_seeding_picks_sklearns_centres(k, n_rep, g):
    """k-means++ on the device against sklearn's own _kmeans_plusplus fed from the same RandomState."""
    from sklearn.cluster._kmeans import _kmeans_plusplus
    from sklearn.utils.extmath import row_norms
    from cnmf_torch_b200 import consensus as cs
    from cnmf_torch_b200._lib import lib, check, ptr
    l2d = cs.l2_normalize_rows(_spectra(k, n_rep, g, 21))
    Xc = l2d.clone()
    mean = torch.empty(g, device="cuda")
    check(lib().cnmf_b200_col_center(ptr(Xc), Xc.stride(0), Xc.shape[0], g, ptr(mean),
                                     torch.cuda.current_stream().cuda_stream))
    Xh = Xc.cpu().numpy()
    ref_mean = l2d.cpu().numpy().mean(axis=0)
    np.testing.assert_array_equal(mean.cpu().numpy(), ref_mean)            # numpy's axis-0 order reproduced
    np.testing.assert_array_equal(Xh, l2d.cpu().numpy() - ref_mean)
    rs = np.random.RandomState(1)
    xsq = row_norms(Xh, squared=True)
    w = np.ones(Xh.shape[0], dtype=np.float32)
    ref_idx = np.stack([_kmeans_plusplus(Xh, k, xsq, w, rs)[1] for _ in range(10)])
    centers, chosen = cs.kmeans_plusplus_device(Xc, k, 10, 1)
    chosen = chosen.cpu().numpy()
    same = (chosen == ref_idx).all(axis=1)
    print(f"[k-means++ k={k} M={Xh.shape[0]}] restarts with identical centre indices: {same.sum()}/10")
    assert same.sum() >= 9        # a differing pick needs a potential within fp32 round-off of a boundary
    np.testing.assert_array_equal(centers.cpu().numpy()[same], Xh[ref_idx[same]])


def test_cluster_medians_match_pandas():
    import pandas as pd
    from cnmf_torch_b200 import consensus as cs
    rng = np.random.default_rng(0)
    x = rng.random((501, 77)).astype(np.float32)
    x[10] = x[11]                       # duplicates exercise the tie rule of the rank count
    labels = rng.integers(0, 6, size=501)
    labels[:3] = 5
    med, nrm = cs.cluster_medians(torch.from_numpy(x).cuda(), labels.astype(np.int64), 6)
    ref = pd.DataFrame(x).groupby(pd.Series(labels)).median()
    np.testing.assert_array_equal(med.cpu().numpy(), ref.values)
    ref_n = (ref.T / ref.sum(axis=1)).T.values
    np.testing.assert_allclose(nrm.cpu().numpy(), ref_n, rtol=2e-6)


@pytest.mark.parametrize("seeding", ["device", "sklearn"])
@pytest.mark.parametrize("tag", ["small", "nofilter", "tightfilter", "mid"])
def test_consensus_front_half_matches_reference_golden(golden_dir, tag, seeding):
    from cnmf_torch_b200 import consensus as cs
    g = _load(golden_dir, f"consensus_{tag}.npz")
    k = int(g["k"])
    n = co.n_neighbors_for(float(g["local_neighborhood_size"]), g["spectra"].shape[0], k)
    out = cs.consensus_front_half(g["spectra"], k, float(g["density_threshold"]), n,
                                  centers_init_fn=None if seeding == "device" else _sklearn_seeds(k))
    np.testing.assert_allclose(out["local_density"], g["local_density"], rtol=2e-6)
    np.testing.assert_array_equal(out["keep"], g["local_density"] < float(g["density_threshold"]))
    np.testing.assert_array_equal(out["labels"], g["labels"])
    np.testing.assert_allclose(out["median_spectra"], g["median_spectra"], rtol=5e-6, atol=1e-9)
    assert np.corrcoef(out["median_spectra"].ravel(), g["median_spectra"].ravel())[0, 1] >= 0.999


def test_consensus_empty_filter_raises(golden_dir):
    from cnmf_torch_b200 import consensus as cs
    g = _load(golden_dir, "consensus_empty.npz")
    with pytest.raises(RuntimeError) as e:
        cs.consensus_front_half(g["spectra"], 3, 1e-6, co.n_neighbors_for(0.3, g["spectra"].shape[0], 3))
    assert str(e.value) == str(g["message"])


def test_consensus_cfg4_shape_against_oracle():
    """BASELINE configs[3]: 200 replicates x K=20 over 2000 genes, n=60 neighbours."""
    from cnmf_torch_b200 import consensus as cs
    s = _spectra(20, 200, 2000, 14)
    n = co.n_neighbors_for(0.30, s.shape[0], 20)
    assert n == 60
    t0 = time.perf_counter()
    ref = co.consensus_front_half(s, 20, 0.5, 0.30)
    t_cpu = time.perf_counter() - t0
    cs.consensus_front_half(s, 20, 0.5, n)          # warm-up (module load, allocator)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = cs.consensus_front_half(s, 20, 0.5, n)
    torch.cuda.synchronize()
    t_gpu = time.perf_counter() - t0
    np.testing.assert_allclose(out["local_density"], ref["local_density"], rtol=2e-6)
    np.testing.assert_array_equal(out["keep"], ref["keep"])
    np.testing.assert_array_equal(out["labels"], ref["labels"])
    np.testing.assert_allclose(out["median_spectra"], ref["median_spectra"], rtol=5e-6, atol=1e-9)
    print(f"[consensus cfg-4 4000x2000 k=20] oracle (sklearn, host) {t_cpu:.3f}s  gpu path {t_gpu:.3f}s")
