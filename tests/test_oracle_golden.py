"""CPU: the oracle restatements against fixtures produced by the reference's own code
(tests/golden/make_golden.py).  Tolerances are written next to each check."""
import os

import numpy as np
import pytest

from oracle import consensus_oracle as co
from oracle import nmf_oracle as no


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name), allow_pickle=True)


@pytest.mark.parametrize("tag,kw", [
    ("default", dict(chunk_size=100, chunk_max_iter=200, h_tol=0.05)),
    ("tight", dict(chunk_size=64, chunk_max_iter=50, h_tol=1e-3)),
    ("reg", dict(chunk_size=100, chunk_max_iter=30, h_tol=1e-3, l1=0.1, l2=0.2)),
    ("one", dict(chunk_size=5000, chunk_max_iter=1, h_tol=0.0)),
])
def test_mu_h_update_matches_reference_fit_H_online(golden_dir, tag, kw):
    """Pins MU numer/denom, the eps rule and the relative-change stop (cnmf.py:332-383)."""
    g = _load(golden_dir, "fit_h_online.npz")
    out = no.fit_H_given_W(g["x"], g["w"], g["h_init"], **kw)
    # same torch-CPU fp32 ops in the same order: agree to fp32 round-off
    np.testing.assert_allclose(out, g["out_" + tag], rtol=2e-5, atol=1e-7)


def test_mu_dead_component_is_zeroed(golden_dir):
    g = _load(golden_dir, "fit_h_online.npz")
    out = no.fit_H_given_W(g["x"], g["w_dead"], g["h_init"], chunk_size=100, chunk_max_iter=20, h_tol=1e-3)
    np.testing.assert_allclose(out, g["out_dead"], rtol=2e-5, atol=1e-7)
    assert np.all(out[:, 4] == 0.0) and np.all(g["out_dead"][:, 4] == 0.0)


@pytest.mark.parametrize("tag", ["small", "nofilter", "tightfilter", "mid"])
def test_consensus_oracle_matches_reference(golden_dir, tag):
    g = _load(golden_dir, f"consensus_{tag}.npz")
    out = co.consensus_front_half(g["spectra"], int(g["k"]), float(g["density_threshold"]),
                                  float(g["local_neighborhood_size"]))
    np.testing.assert_array_equal(out["local_density"], g["local_density"])   # same library calls: exact
    np.testing.assert_array_equal(out["labels"], g["labels"])
    np.testing.assert_array_equal(out["median_spectra"], g["median_spectra"])


def test_consensus_empty_filter_raises(golden_dir):
    g = _load(golden_dir, "consensus_empty.npz")
    with pytest.raises(RuntimeError) as e:
        co.consensus_front_half(g["spectra"], 3, 1e-6)
    assert str(e.value) == str(g["message"])


def test_batch_oracle_objective_decreases_and_hals_beats_mu():
    rng = np.random.default_rng(0)
    x = rng.poisson(rng.gamma(1.0, 1.0, (120, 5)) @ rng.gamma(0.5, 1.0, (5, 70))).astype(np.float32)
    h0, w0 = no.init_random(x.mean(), 120, 70, 5, seed=7)
    mu = no.nmf_batch(x, h0, w0, algo="mu", tol=0, max_iter=60)
    hals = no.nmf_batch(x, h0, w0, algo="hals", tol=0, max_iter=60)
    kl = no.nmf_batch(x, h0, w0, beta_loss="kullback-leibler", tol=0, max_iter=60)
    for out in (mu, hals, kl):
        obj = [v for _, v in out["trace"]]
        assert all(b <= a * (1 + 1e-6) for a, b in zip(obj, obj[1:]))   # monotone (MU/HALS theory)
    assert hals["trace"][-1][1] <= mu["trace"][-1][1]


def test_sklearn_cross_check_frobenius_mu():
    """Independent cross-check (SURVEY 8c iii): sklearn's MU from the same custom init follows the
    same objective path (sklearn names W=usages, H=spectra and updates W first, like H here)."""
    from sklearn.decomposition import non_negative_factorization
    rng = np.random.default_rng(1)
    x = rng.poisson(rng.gamma(1.0, 1.0, (90, 4)) @ rng.gamma(0.5, 1.0, (4, 50))).astype(np.float64) + 0.0
    h0, w0 = no.init_random(x.mean(), 90, 50, 4, seed=3)
    ours = no.nmf_batch(x, h0, w0, algo="mu", tol=0, max_iter=30)
    Wsk, Hsk, _ = non_negative_factorization(x, W=h0.astype(np.float64).copy(), H=w0.astype(np.float64).copy(),
                                             n_components=4, init="custom", solver="mu", beta_loss="frobenius",
                                             tol=0, max_iter=30, update_H=True)
    obj_sk = 0.5 * np.sum((x - Wsk @ Hsk) ** 2)
    assert abs(obj_sk - ours["trace"][-1][1]) / obj_sk < 1e-3


def test_online_oracle_runs_and_improves():
    rng = np.random.default_rng(2)
    x = rng.poisson(rng.gamma(1.0, 1.0, (300, 4)) @ rng.gamma(0.5, 1.0, (4, 40))).astype(np.float32)
    h0, w0 = no.init_random(x.mean(), 300, 40, 4, seed=5)
    out = no.nmf_online(x, h0, w0, chunk_size=100, max_pass=5, chunk_max_iter=50)
    assert out["trace"][-1][1] < out["trace"][0][1]
    assert out["H"].shape == (300, 4) and out["W"].shape == (4, 40)
