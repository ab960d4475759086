"""GPU parity of the sparse KL path (through the C ABI) against the oracle from identical H0/W0:
objective within 1e-3 relative at every check, final factors Pearson >= 0.999 (north_star)."""
import numpy as np
import pytest
import torch

from oracle import nmf_oracle as no

pytestmark = pytest.mark.gpu


def _pearson(a, b):
    return float(np.corrcoef(a.ravel().astype(np.float64), b.ravel().astype(np.float64))[0, 1])


@pytest.fixture(autouse=True, params=["slab", "gather"])
def kl_impl(request, monkeypatch):
    """Every test of this file runs on both kernel families behind cnmf_b200_kl_iterate (include/cnmf_b200.h:
    CNMF_B200_KL_IMPL_SLAB, the default, and CNMF_B200_KL_IMPL_GATHER)."""
    monkeypatch.setenv("CNMF_B200_KL_IMPL", request.param)
    return request.param


@pytest.fixture(scope="module")
def problem():
    from cnmf_torch_b200.synth import nb_counts
    return nb_counts(421, 233, k_true=6, seed=14, libsize=500.0)


def test_kl_constants(problem):
    from cnmf_torch_b200.kl_engine import SparseKLEngine
    eng = SparseKLEngine(problem)
    x = problem.astype(np.float64)
    x = x * eng.c ** 2          # the engine's units: X scaled by a power of four to mean ~1
    assert 0.25 <= x.mean() <= 4.0 and np.log2(eng.c) == int(np.log2(eng.c))
    assert abs(eng.sumX - x.sum()) / x.sum() < 1e-12
    xl = (x[x > 0] * np.log(x[x > 0])).sum()
    assert abs(eng.xlogx - xl) <= 1e-9 * max(1.0, abs(xl))
    assert eng.nnz == int((problem > 0).sum())


@pytest.mark.parametrize("K", [3, 5, 9, 13])
def test_kl_single_iteration_matches_oracle(problem, K):
    """One full iteration (H then W) elementwise against the oracle's dense formula.  The gathered copy of
    the fixed factor is fp16 (11-bit significand, like the tf32 operands of the dense path), so single
    entries agree to ~1e-3 relative; the objective-level bar of north_star is checked in the next test."""
    from cnmf_torch_b200.kl_engine import SparseKLEngine
    eng = SparseKLEngine(problem)
    H0, W0 = eng.make_init(K, [7, 8])
    b = eng.new_batch(K, H0=H0, W0=W0)
    b.iterate(1)
    W, H = b.results(True)
    xt = torch.from_numpy(problem)
    for r in range(2):
        h = no.mu_kl_H(xt, H0[r], W0[r], 0.0, 0.0, 1e-16)
        w = no.mu_kl_W(xt, h, W0[r], 0.0, 0.0, 1e-16)
        np.testing.assert_allclose(H[r], h.numpy(), rtol=2e-3, atol=1e-6)
        np.testing.assert_allclose(W[r], w.numpy(), rtol=2e-3, atol=1e-6)
        assert np.abs(H[r] / np.maximum(h.numpy(), 1e-12) - 1).mean() < 2e-4    # unbiased, small on average


@pytest.mark.parametrize("K", [5, 12])
def test_kl_engine_trace_matches_oracle(problem, K):
    from cnmf_torch_b200.kl_engine import SparseKLEngine
    eng = SparseKLEngine(problem)
    seeds = [3, 4, 5]
    H0, W0 = eng.make_init(K, seeds)
    res = eng.factorize(K, H0=H0, W0=W0, tol=0.0, max_iter=40, return_H=True)
    assert res.launches > 0
    worst = 0.0
    for r in range(3):
        ora = no.nmf_batch(problem, H0[r].numpy(), W0[r].numpy(), beta_loss="kullback-leibler", tol=0.0, max_iter=40)
        for (it_o, obj_o), (it_g, obj_g) in zip(ora["trace"], res.trace):
            assert it_o == it_g
            rel = abs(obj_g[r] - obj_o) / obj_o
            worst = max(worst, rel)
            assert rel < 1e-3, (K, r, it_o, obj_g[r], obj_o)
        assert _pearson(res.W[r], ora["W"]) >= 0.999 and _pearson(res.H[r], ora["H"]) >= 0.999
        assert abs(res.err[r] - ora["err"]) / ora["err"] < 1e-3
    print(f"[kl engine K={K}] worst objective rel err {worst:.2e}")


def test_kl_convergence_freeze(problem):
    from cnmf_torch_b200.kl_engine import SparseKLEngine
    eng = SparseKLEngine(problem)
    H0, W0 = eng.make_init(4, [1, 2, 3])
    res = eng.factorize(4, H0=H0, W0=W0, tol=1e-3, max_iter=300)
    assert np.all(res.n_iter % 10 == 0)
    for r in range(3):
        ora = no.nmf_batch(problem, H0[r].numpy(), W0[r].numpy(), beta_loss="kullback-leibler", tol=1e-3, max_iter=300)
        assert abs(int(res.n_iter[r]) - ora["n_iter"]) <= 10
        assert abs(res.err[r] - ora["err"]) / ora["err"] < 1e-3


def test_run_nmf_dropin_kl_and_frobenius(problem):
    """The reference's call shape (cnmf.py:819): (usages, spectra, err) = run_nmf(X, **kwargs)."""
    from cnmf_torch_b200.nmf import run_nmf
    kw = dict(alpha_W=0.0, alpha_H=0.0, l1_ratio_H=0.0, l1_ratio_W=0.0, algo="mu", tol=1e-4, mode="batch",
              init="random", n_jobs=-1, use_gpu=True, random_state=11, n_components=5)
    for beta in ("frobenius", "kullback-leibler"):
        H, W, err = run_nmf(problem.astype(np.float64), beta_loss=beta, init_device="cpu", **kw)
        assert H.shape == (problem.shape[0], 5) and W.shape == (5, problem.shape[1]) and np.isfinite(err)
        ho, wo, eo = no.run_nmf_oracle(problem, beta_loss=beta, **{k: v for k, v in kw.items() if k != "use_gpu"})
        assert abs(err - eo) / eo < 1e-3


def test_run_nmf_with_the_reference_default_kwargs(problem):
    """The exact dict ``cnmf prepare`` writes without --use_gpu (cnmf.py:757-771: use_gpu=False, n_jobs=-1,
    mode='online') plus the two keys the worker loop adds (cnmf.py:885-886, numpy integers as pandas hands them
    over).  use_gpu=False selects the host generator's init stream -- the oracle's -- not a CPU path."""
    from cnmf_torch_b200.nmf import run_nmf
    for beta in ("frobenius", "kullback-leibler"):
        kw = dict(alpha_W=0.0, alpha_H=0.0, l1_ratio_H=0.0, l1_ratio_W=0.0, beta_loss=beta, algo="mu", tol=1e-4,
                  mode="online", online_chunk_max_iter=1000, online_chunk_size=5000, init="random", n_jobs=-1,
                  use_gpu=False)
        kw["random_state"] = np.int64(12345)
        kw["n_components"] = np.int64(5)
        H, W, err = run_nmf(problem, **kw)
        assert H.shape == (problem.shape[0], 5) and W.shape == (5, problem.shape[1])
        ho, wo, eo = no.run_nmf_oracle(problem, **{k: v for k, v in kw.items() if k not in ("use_gpu", "n_jobs")})
        assert abs(err - eo) / eo < 1e-3
        assert _pearson(W, wo) >= 0.999
        # use_gpu=True starts from the device generator's draws: another replicate, same quality
        H2, W2, err2 = run_nmf(problem, **{**kw, "use_gpu": True})
        assert not np.allclose(W2, W) and abs(err2 - err) / err < 0.05


def test_itakura_saito_is_refused_with_the_reason(problem, tmp_path):
    """A ``choices=`` value of the reference CLI (cnmf.py:1426).  Refused where the run is set up, with the reason."""
    from cnmf_torch_b200.cnmf import cNMF
    from cnmf_torch_b200.nmf import run_nmf
    with pytest.raises(NotImplementedError, match="zero"):
        run_nmf(problem, n_components=4, beta_loss="itakura-saito")
    obj = cNMF(output_dir=str(tmp_path), name="is")
    with pytest.raises(NotImplementedError, match="zero"):
        obj.prepare_from_matrix(problem, [4], n_iter=2, seed=1, beta_loss="itakura-saito")
    with pytest.raises(ValueError):
        run_nmf(problem, n_components=4, beta_loss="hellinger")


def test_cnmf_default_online_kwargs_with_kl_fall_back_to_batch(problem, capsys):
    """cNMF always writes mode='online' (cnmf.py:765); with beta_loss='kullback-leibler' there are no
    sufficient statistics, so -- like nmf-torch -- the run switches to batch mode instead of failing."""
    from cnmf_torch_b200.nmf import run_nmf
    kw = dict(alpha_W=0.0, alpha_H=0.0, l1_ratio_H=0.0, l1_ratio_W=0.0, beta_loss="kullback-leibler", algo="mu",
              tol=1e-4, mode="online", online_chunk_max_iter=1000, online_chunk_size=5000, init="random", n_jobs=-1,
              use_gpu=True, random_state=3, n_components=4)
    H, W, err = run_nmf(problem, init_device="cpu", **kw)     # the oracle draws from the host generator
    assert "Switch to batch" in capsys.readouterr().out
    ho, wo, eo = no.run_nmf_oracle(problem, **{k: v for k, v in kw.items() if k != "use_gpu"})
    assert abs(err - eo) / eo < 1e-3 and _pearson(W, wo) >= 0.999


@pytest.mark.parametrize("beta", ["frobenius", "kullback-leibler"])
def test_nndsvd_init_through_run_nmf(problem, beta):
    from cnmf_torch_b200.nmf import run_nmf
    kw = dict(beta_loss=beta, algo="mu", tol=1e-4, mode="batch", init="nndsvd", random_state=5, n_components=5,
              batch_max_iter=60)
    H, W, err = run_nmf(problem, use_gpu=True, **kw)
    ho, wo, eo = no.run_nmf_oracle(problem, **kw)
    assert abs(err - eo) / eo < 1e-3 and _pearson(W, wo) >= 0.999


def test_kl_cut_rows_empty_rows_and_many_segments(kl_impl):
    """Edge cases of the segment table: genes expressed in every cell (rows of CSR(X^T) cut into a dozen segments,
    finished by kl_finish_rows_kernel), a gene and a cell without any count (empty rows in either orientation), rows
    of exactly one segment length, and R not a divisor of the CTA size -- one iteration against the oracle's dense
    formula, and bit-identical results when the same replicates run in batches of another size (another packing of
    (segment, replicate) pairs over threads, another segment length)."""
    from cnmf_torch_b200.kl_engine import SparseKLEngine
    rng = np.random.default_rng(7)
    n, g = 3100, 90
    x = (rng.poisson(0.3, size=(n, g)) * (rng.random((n, g)) < 0.4)).astype(np.float32)
    x[:, 0] = rng.poisson(3.0, size=n) + 1          # expressed everywhere: 3100 entries in its CSR(X^T) row
    x[:, 1] = rng.poisson(2.0, size=n) + 1
    x[:, 2] = 0.0                                   # never expressed: empty row of CSR(X^T)
    x[:256, 3] = 1.0; x[256:, 3] = 0.0              # exactly 256 entries
    x[5, :] = 0.0; x[5, 0] = 0.0                    # a cell without counts: empty row of CSR(X)
    x[5, :3] = 0.0
    eng = SparseKLEngine(x)
    K, seeds = 5, [1, 2, 3, 4, 5, 6, 7]
    H0, W0 = eng.make_init(K, seeds)
    b = eng.new_batch(K, H0=H0, W0=W0)
    b.iterate(1)
    W, H = b.results(True)
    xt = torch.from_numpy(x)
    for r in (0, 3, 6):
        h = no.mu_kl_H(xt, H0[r], W0[r], 0.0, 0.0, 1e-16)
        w = no.mu_kl_W(xt, h, W0[r], 0.0, 0.0, 1e-16)
        np.testing.assert_allclose(H[r], h.numpy(), rtol=2e-3, atol=1e-6)
        np.testing.assert_allclose(W[r], w.numpy(), rtol=2e-3, atol=1e-6)
        assert np.all(W[r][:, 2] == 0.0) and np.all(H[r][5] == 0.0)       # nothing to explain: the rate is zero
    # the objective after 10 iterations, and independence of the batch composition
    full = eng.factorize(K, H0=H0, W0=W0, tol=0.0, max_iter=10, return_H=True)
    part = eng.factorize(K, H0=H0[2:5], W0=W0[2:5], tol=0.0, max_iter=10, return_H=True)
    for r in range(3):
        ora = no.nmf_batch(x, H0[2 + r].numpy(), W0[2 + r].numpy(), beta_loss="kullback-leibler", tol=0.0, max_iter=10)
        assert abs(part.trace[-1][1][r] - ora["trace"][-1][1]) / ora["trace"][-1][1] < 1e-3
    # (gather: a replicate's arithmetic does not depend on its neighbours; slab: its lane -- entry slot and rotation
    # of the operands -- does, so sums are taken in another order: round-off, amplified over 10 iterations)
    np.testing.assert_allclose(full.W[2:5], part.W, rtol=1e-5 if kl_impl == "gather" else 1e-3, atol=1e-6)
    np.testing.assert_allclose(full.err[2:5], part.err, rtol=1e-6 if kl_impl == "gather" else 1e-5)


def test_kl_column_blocked_gene_major_pass_matches_oracle(kl_impl):
    """The gene-major half-step of a large matrix walks CSR(X^T) cell block by cell block (so the rows of H gathered
    by the threads resident at one time fit in L2) and finishes every gene from per-segment partials.  The block
    size is chosen from the problem size when the library is first used; CNMF_B200_KL_BLOCK=128 forces 4-25 blocks
    on the small problems of this file, which then must pass the same oracle comparisons."""
    import os, subprocess, sys
    if os.environ.get("CNMF_B200_KL_BLOCK"):
        pytest.skip("already inside the forced-block run")
    here = os.path.abspath(__file__)
    # the slab path blocks the columns of BOTH matrices by its shared-memory size: 96 lines and 100-entry segments put
    # the same small problems through 3-33 column blocks per half-step, slab reloads and cut rows
    rc = subprocess.run([sys.executable, "-m", "pytest", here, "-x", "-q", "-k",
                         "(single_iteration or trace_matches or cut_rows or convergence_freeze or l1_and_l2) and "
                         + kl_impl],
                        env=dict(os.environ, CNMF_B200_KL_BLOCK="128", CNMF_B200_KS_LINES="96", CNMF_B200_KS_SEG="100"),
                        capture_output=True, text=True)
    assert rc.returncode == 0, rc.stdout[-2000:] + rc.stderr[-2000:]


def test_kl_with_l1_and_l2_regularisation_matches_oracle(problem):
    """alpha_usage / alpha_spectra reach run_nmf as pure L2 strengths (cNMF fixes l1_ratio = 0, cnmf.py:757-761):
    update denominators and the 0.5 * l2 * ||.||^2 terms of the objective."""
    from cnmf_torch_b200.kl_engine import SparseKLEngine
    eng = SparseKLEngine(problem)
    reg = dict(l1_H=0.02, l2_H=0.1, l1_W=0.01, l2_W=0.3)
    H0, W0 = eng.make_init(5, [3, 4])
    res = eng.factorize(5, H0=H0, W0=W0, tol=0.0, max_iter=30, return_H=True, **reg)
    for r in range(2):
        ora = no.nmf_batch(problem, H0[r].numpy(), W0[r].numpy(), beta_loss="kullback-leibler", tol=0.0, max_iter=30,
                           l1_h=reg["l1_H"], l2_h=reg["l2_H"], l1_w=reg["l1_W"], l2_w=reg["l2_W"])
        for (it_o, obj_o), (it_g, obj_g) in zip(ora["trace"], res.trace):
            assert it_o == it_g and abs(obj_g[r] - obj_o) / obj_o < 1e-3, (r, it_o, obj_g[r], obj_o)
        assert _pearson(res.W[r], ora["W"]) >= 0.999 and _pearson(res.H[r], ora["H"]) >= 0.999
    plain = eng.factorize(5, H0=H0, W0=W0, tol=0.0, max_iter=30)
    assert np.all(res.trace[-1][1] != plain.trace[-1][1])      # the terms are really there


def test_kl_data_scale_and_out_of_range_init_are_handled_not_clamped(problem):
    """The gathered copies are fp16.  (i) The units of X do not matter: the engine works on X scaled by a power of four
    to mean ~1, which KL multiplicative updates follow exactly -- X * 1e-6 and X * 1e6 give the factors of X scaled by
    1e-3 and 1e3.  (ii) An initial factor many orders of magnitude out of range is REFUSED (an error that says why),
    never silently saturated at 65504 as round 1's kernel did."""
    from cnmf_torch_b200 import _lib
    from cnmf_torch_b200.kl_engine import SparseKLEngine
    K, seeds = 5, [3, 4]
    base = SparseKLEngine(problem)
    H0, W0 = base.make_init(K, seeds)
    ref = base.factorize(K, H0=H0, W0=W0, tol=0.0, max_iter=20, return_H=True)
    for scale in (1e-6, 1e6):
        eng = SparseKLEngine(problem * np.float32(scale))
        s = np.float32(np.sqrt(scale))
        res = eng.factorize(K, H0=H0 * s, W0=W0 * s, tol=0.0, max_iter=20, return_H=True)
        for r in range(2):
            assert abs(res.trace[-1][1][r] / scale - ref.trace[-1][1][r]) / ref.trace[-1][1][r] < 1e-3
            assert _pearson(res.W[r], ref.W[r]) >= 0.9999 and _pearson(res.H[r], ref.H[r]) >= 0.9999
            np.testing.assert_allclose(res.W[r].sum() / s, ref.W[r].sum(), rtol=1e-2)
    # entries of H0 up to ~1e8 in the engine's units (W0 shrunk to keep the product): beyond fp16
    big = 3e7
    with pytest.raises(_lib.Cnmfb200Error, match="65504"):
        base.factorize(K, H0=H0 * np.float32(big), W0=W0 / np.float32(big), tol=0.0, max_iter=10)
    # ... while a merely large but representable start runs and reaches the same objective as the oracle
    mid = 64.0
    res = base.factorize(K, H0=H0 * np.float32(mid), W0=W0 / np.float32(mid), tol=0.0, max_iter=20)
    for r in range(2):
        ora = no.nmf_batch(problem, (H0[r] * mid).numpy(), (W0[r] / mid).numpy(), beta_loss="kullback-leibler", tol=0.0,
                           max_iter=20)
        assert abs(res.trace[-1][1][r] - ora["trace"][-1][1]) / ora["trace"][-1][1] < 1e-3


def test_kl_warm_calls_reuse_tables_and_copies(problem):
    """cnmf_b200_kl_problem.warm (include/cnmf_b200.h): after the first iterate call the batch reuses the segment
    tables, packed entries and fp16 copies its workspace holds -- 3 + 3 + 2 iterations with an objective in between
    are bit-identical to 8 iterations of a fresh batch; `invalidate()` makes a caller's write to a factor count."""
    from cnmf_torch_b200.kl_engine import SparseKLEngine
    eng = SparseKLEngine(problem)
    H0, W0 = eng.make_init(7, [11, 12, 13])
    a = eng.new_batch(7, H0=H0, W0=W0)
    assert a.p.warm == 0
    a.iterate(3)
    assert a.p.warm == 1
    a.objective()
    a.iterate(3)
    a.set_active(np.array([True, False, True]))
    a.iterate(2)
    b = eng.new_batch(7, H0=H0, W0=W0)
    b.iterate(6)
    b.invalidate()
    b.set_active(np.array([True, False, True]))
    b.iterate(2)
    Wa, Ha = a.results(True)
    Wb, Hb = b.results(True)
    np.testing.assert_array_equal(Wa, Wb)
    np.testing.assert_array_equal(Ha, Hb)
    np.testing.assert_array_equal(a.read_loss(), b.read_loss())
    # a write to the master without invalidate() would be ignored by the fp16 copies; with it, it counts
    c = eng.new_batch(7, H0=H0, W0=W0)
    c.iterate(1)
    c.Hr.mul_(0.5)
    c.Wt.mul_(2.0)
    c.invalidate()
    c.iterate(1)
    d = eng.new_batch(7, H0=H0, W0=W0)
    d.iterate(2)
    Wc, Hc = c.results(True)
    Wd, Hd = d.results(True)
    np.testing.assert_allclose(Hc @ Wc, Hd @ Wd, rtol=5e-3, atol=1e-6)   # H/2, 2W: the same product, fp16 rounding aside


def test_kl_large_ratios_are_held(problem):
    """Entries fitted four to five orders of magnitude too low (x / (WH) up to ~1e6: beyond fp16's 65504): the slab
    kernels stage x * 2^-8 so that the fp16 ratio holds up to 1.6e7 (kl_slab.cu), the gather kernels keep the ratio
    in fp32 -- one iteration from such a start still matches the oracle and raises no range flag."""
    from cnmf_torch_b200.kl_engine import SparseKLEngine
    eng = SparseKLEngine(problem)
    H0, W0 = eng.make_init(6, [21, 22])
    H0 = (H0 * 0.01).contiguous()
    W0 = (W0 * 0.01).contiguous()
    b = eng.new_batch(6, H0=H0, W0=W0)
    b.iterate(1)
    b.check_range()
    W, H = b.results(True)
    xt = torch.from_numpy(problem)
    ratio_max = float((xt / (H0[0] @ W0[0])).max())
    assert ratio_max > 2e5          # the case is what it claims to be
    for r in range(2):
        h = no.mu_kl_H(xt, H0[r], W0[r], 0.0, 0.0, 1e-16)
        w = no.mu_kl_W(xt, h, W0[r], 0.0, 0.0, 1e-16)
        np.testing.assert_allclose(H[r], h.numpy(), rtol=3e-3, atol=1e-6)
        np.testing.assert_allclose(W[r], w.numpy(), rtol=3e-3, atol=1e-6)


def test_kl_compaction_is_bit_identical(problem):
    """Once most replicates of a stack are frozen the live ones continue in a smaller batch (kl_engine._run_kl_batch);
    replicates are independent, so spectra, usages, objectives and iteration counts equal the uncompacted run's bit
    for bit -- on a mixed-K stack whose replicates stop at different checks."""
    from cnmf_torch_b200.kl_engine import SparseKLEngine
    eng = SparseKLEngine(problem)
    seeds = {5: list(range(100, 114)), 6: list(range(200, 213))}
    kw = dict(tol=2e-4, max_iter=300, return_H=True)
    a = eng.factorize_many([5, 6], seeds, compact=0.8, **kw)
    b = eng.factorize_many([5, 6], seeds, compact=0.0, **kw)
    stops = set()
    for k in (5, 6):
        np.testing.assert_array_equal(a[k].n_iter, b[k].n_iter)
        np.testing.assert_array_equal(a[k].W, b[k].W)
        np.testing.assert_array_equal(a[k].H, b[k].H)
        np.testing.assert_array_equal(a[k].err, b[k].err)
        stops |= set(a[k].n_iter.tolist())
    assert len(stops) >= 3           # replicates did stop at different checks, so the compaction was exercised
