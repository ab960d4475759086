"""GPU: the rows after the hot path against the reference's own outputs / library calls."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_ols_matches_reference_golden(golden_dir):
    """efficient_ols_all_cols against outputs of the reference's own function (tests/golden/ols.npz)."""
    import scipy.sparse as sps
    from cnmf_torch_b200.stats import efficient_ols_all_cols
    g = np.load(os.path.join(golden_dir, "ols.npz"))
    X, Y = g["X"], g["Y"]
    # the golden was produced from float32 X / Y: the reference forms each batch product X_b^T Y_b in float32
    # before adding it to its float64 accumulator, so it carries fp32 round-off (~1e-5 relative); the device
    # path accumulates every product in fp64.  Bar vs the reference: 2e-4; bar vs an all-fp64 restatement: 1e-9.
    X64, Y64 = X.astype(np.float64), Y.astype(np.float64)
    exact_raw = np.linalg.lstsq(X64.T @ X64, X64.T @ Y64, rcond=None)[0]
    out = efficient_ols_all_cols(X, Y, normalize_y=False)
    scale = np.abs(exact_raw).max()
    np.testing.assert_allclose(out, exact_raw, rtol=1e-9, atol=1e-11 * scale)
    np.testing.assert_allclose(out, g["beta_dense_raw"], rtol=2e-4, atol=2e-6 * scale)
    mean, var = Y64.mean(0), Y64.var(0)
    np.testing.assert_allclose(mean, g["mean"], rtol=1e-6, atol=1e-9)     # get_mean_var of the reference
    var[var < 1e-12] = 1e-12
    exact_norm = np.linalg.lstsq(X64.T @ X64, X64.T @ ((Y64 - mean) / np.sqrt(var)), rcond=None)[0]
    scale = np.abs(exact_norm).max()
    for Yin, key in ((Y, "beta_dense_norm"), (sps.csr_matrix(Y), "beta_sparse_norm")):
        out = efficient_ols_all_cols(X, Yin, batch_size=77, normalize_y=True)
        assert out.shape == g[key].shape and out.dtype == np.float64
        np.testing.assert_allclose(out, exact_norm, rtol=1e-7, atol=1e-9 * scale)
        np.testing.assert_allclose(out, g[key], rtol=2e-4, atol=2e-6 * scale)
    with pytest.raises(ValueError, match="same number of rows"):
        efficient_ols_all_cols(X[:-1], Y)


@pytest.mark.parametrize("k,n_rep,g", [(5, 30, 80), (8, 50, 300), (20, 200, 2000)])
def test_silhouette_matches_sklearn(k, n_rep, g):
    from sklearn.cluster import KMeans
    from sklearn.metrics import silhouette_score as sk_sil
    from cnmf_torch_b200 import consensus as cs
    from cnmf_torch_b200.stats import silhouette_score
    from cnmf_torch_b200.synth import replicate_spectra
    l2d = cs.l2_normalize_rows(replicate_spectra(k, n_rep, g, seed=9, outlier_frac=0.03))
    l2 = l2d.cpu().numpy()
    labels = KMeans(n_clusters=k, n_init=10, random_state=1).fit(l2.copy()).labels_
    ref = sk_sil(l2, labels, metric="euclidean")
    got = silhouette_score(l2d, labels, k)
    assert abs(got - ref) < 2e-6, (got, ref)


def test_k_selection_stats_through_cnmf(tmp_path):
    from cnmf_torch_b200 import cnmf as host
    from cnmf_torch_b200.synth import nb_counts
    x = nb_counts(300, 120, k_true=4, seed=14, libsize=500.0)
    obj = host.cNMF(output_dir=str(tmp_path), name="ks")
    obj.prepare_from_matrix(x, [3, 4], n_iter=5, seed=14, beta_loss="frobenius")
    obj.factorize(nmf_kwargs_override=dict(mode="batch", batch_max_iter=80))
    obj.combine()
    st = obj.consensus(4, skip_density_and_return_after_stats=True)
    assert list(st.index) == ["k", "local_density_threshold", "silhouette", "prediction_error"]   # cnmf.py:1106-1108
    assert st.loc["k", "stats"] == 4 and -1.0 <= st.loc["silhouette", "stats"] <= 1.0
    # prediction error = ||X - usages @ spectra||^2 for the refit usages: recompute it on the host
    merged = host.load_df_from_npz(obj.paths["merged_spectra"] % 4)
    from cnmf_torch_b200.consensus import consensus_front_half
    out = consensus_front_half(merged.values, 4, 2.0, 1, skip_density=True)
    H = obj.refit_usage(x, out["median_spectra"], seed=1)
    ref = float(((x.astype(np.float64) - H.astype(np.float64) @ out["median_spectra"].astype(np.float64)) ** 2).sum())
    assert abs(st.loc["prediction_error", "stats"] - ref) / ref < 1e-5
    table = obj.k_selection_stats()
    assert list(table.columns) == ["k", "local_density_threshold", "silhouette", "prediction_error"] and len(table) == 2
