"""GPU: fit_H_online (usage refit, cnmf.py:260-388) against outputs of the REFERENCE's own function
(tests/golden/fit_h_online.npz, made by importing /root/reference in the build container)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("tag,kw", [
    ("default", dict(chunk_size=100, chunk_max_iter=200, h_tol=0.05)),
    ("tight", dict(chunk_size=64, chunk_max_iter=50, h_tol=1e-3)),
    ("reg", dict(chunk_size=100, chunk_max_iter=30, h_tol=1e-3, l1_reg_H=0.1, l2_reg_H=0.2)),
    ("one", dict(chunk_size=5000, chunk_max_iter=1, h_tol=0.0)),
])
def test_fit_H_online_matches_reference(golden_dir, tag, kw):
    from cnmf_torch_b200.refit import fit_H_online
    g = np.load(os.path.join(golden_dir, "fit_h_online.npz"), allow_pickle=True)
    out = fit_H_online(g["x"], g["w"], H_init=g["h_init"].copy(), **kw)
    np.testing.assert_allclose(out, g["out_" + tag], rtol=2e-4, atol=1e-6)


def test_fit_H_online_dead_component_and_seeded_init(golden_dir):
    from cnmf_torch_b200.refit import fit_H_online
    g = np.load(os.path.join(golden_dir, "fit_h_online.npz"), allow_pickle=True)
    out = fit_H_online(g["x"], g["w_dead"], H_init=g["h_init"].copy(), chunk_size=100, chunk_max_iter=20, h_tol=1e-3)
    np.testing.assert_allclose(out, g["out_dead"], rtol=2e-4, atol=1e-6)
    assert np.all(out[:, 4] == 0.0)
    a = fit_H_online(g["x"], g["w"], seed=5, chunk_size=100)
    b = fit_H_online(g["x"], g["w"], seed=5, chunk_size=100)
    np.testing.assert_array_equal(a, b)          # reproducible when seeded (the reference's is not)
    assert a.shape == g["h_init"].shape and (a >= 0).all()


def test_refit_usage_and_spectra_through_cnmf(tmp_path, golden_dir):
    from cnmf_torch_b200 import cnmf as host
    g = np.load(os.path.join(golden_dir, "fit_h_online.npz"), allow_pickle=True)
    obj = host.cNMF(output_dir=str(tmp_path), name="rf")
    x = g["x"] + 0.0
    x[x.sum(1) == 0, 0] = 1.0
    obj.prepare_from_matrix(x, [5], n_iter=2, seed=1, batch_size=100, max_iter=200)
    H = obj.refit_usage(x, g["w"], usage=g["h_init"].copy())
    assert H.shape == g["h_init"].shape
    Wt = obj.refit_spectra(x, H, seed=3)
    assert Wt.shape == g["w"].shape and np.isfinite(Wt).all()
    # the refit lowers the Frobenius error of the pair it is given
    e0 = np.linalg.norm(x - g["h_init"] @ g["w"]); e1 = np.linalg.norm(x - H @ g["w"])
    assert e1 < e0
