"""GPU: size-independent properties at BASELINE.json's full sizes and the edge cases of the path.

The oracle finishes only small cases in seconds; at full size the engine is checked through properties the
algorithms guarantee: the objective never increases under MU / HALS, a run is bit-reproducible, a
replicate's result does not depend on which other replicates share its batch, frozen replicates stay
frozen, and the consensus front half is invariant under a permutation of the replicate rows."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cfg2():
    from cnmf_torch_b200.synth import nb_counts
    return nb_counts(2700, 2000, seed=14)


@pytest.mark.parametrize("algo", ["mu", "hals"])
def test_cfg2_full_size_monotone_deterministic_batch_independent(cfg2, algo):
    from cnmf_torch_b200.engine import DenseFrobeniusEngine
    from cnmf_torch_b200.synth import ledger_seeds
    eng = DenseFrobeniusEngine(cfg2)
    seeds = [int(s) for s in ledger_seeds(100, 14)]
    a = eng.factorize(10, seeds=seeds, algo=algo, tol=0.0, max_iter=60)
    obj = np.stack([v for _, v in a.trace])                       # [checks, R]
    # the objective is evaluated for the fp32 factors (tf32 copy + residual pass): noise ~1e-7 relative
    assert np.all(np.diff(obj, axis=0) <= 1e-6 * obj[:-1]), "objective must not increase"
    assert np.all(obj[-1] < 0.9 * obj[0])
    b = eng.factorize(10, seeds=seeds, algo=algo, tol=0.0, max_iter=60)
    assert np.array_equal(a.W, b.W) and np.array_equal(a.err, b.err), "bit-reproducible"
    # the same seeds inside a different batch (other companions, other position in the stack)
    sub = [seeds[7], seeds[42], seeds[99], seeds[0]]
    c = eng.factorize(10, seeds=sub, algo=algo, tol=0.0, max_iter=60)
    for j, i in enumerate([7, 42, 99, 0]):
        np.testing.assert_allclose(c.W[j], a.W[i], rtol=2e-3, atol=1e-6)   # tile shapes differ: fp32 sum order
        assert abs(c.err[j] - a.err[i]) / a.err[i] < 1e-5


def test_cfg2_convergence_freezes_replicates(cfg2):
    from cnmf_torch_b200.engine import DenseFrobeniusEngine
    eng = DenseFrobeniusEngine(cfg2)
    res = eng.factorize(7, seeds=list(range(1, 41)), algo="hals", tol=1e-3, max_iter=400)
    assert np.all(res.n_iter % 10 == 0) and res.n_iter.min() < res.n_iter.max()
    last = res.trace[-1][1]
    for r in range(40):                      # after its stop iteration a replicate's objective is constant
        vals = [v[r] for it, v in res.trace if it >= res.n_iter[r]]
        assert max(vals) - min(vals) <= 1e-9 * abs(last[r])


def test_cfg2_compaction_matches_uncompacted_run(cfg2):
    """run_batches moves the live replicates of a stack into a smaller batch as the others freeze
    (FrobBatch.subset); the factors come out as in the uncompacted run, up to the summation-order rounding of a
    different tile shape."""
    from cnmf_torch_b200.engine import DenseFrobeniusEngine
    eng = DenseFrobeniusEngine(cfg2)
    kw = dict(seeds=list(range(1, 65)), algo="hals", tol=1e-3, max_iter=400)
    a = eng.factorize(7, compact=0.5, **kw)
    b = eng.factorize(7, compact=0.0, **kw)
    assert a.n_iter.min() < a.n_iter.max()                       # replicates stopped at different checks
    assert np.mean(a.n_iter == b.n_iter) >= 0.9                  # (a borderline stop may move by one check)
    same = a.n_iter == b.n_iter
    assert np.all(np.abs(a.err[same] - b.err[same]) <= 1e-4 * b.err[same])
    for r in np.flatnonzero(same):
        c = np.corrcoef(a.W[r].ravel(), b.W[r].ravel())[0, 1]
        assert c >= 0.9999, (r, c)


def test_kl_large_monotone_and_edge_rows():
    from cnmf_torch_b200.kl_engine import SparseKLEngine
    from cnmf_torch_b200.synth import nb_counts
    x = nb_counts(6000, 2000, seed=3, libsize=170.0)
    x[:, 17] = 0.0                           # a gene that is never observed
    x[5, :] = 0.0; x[5, 3] = 2.5             # a cell with a single non-zero
    eng = SparseKLEngine(x)
    res = eng.factorize(32, seeds=[1, 2, 3], tol=0.0, max_iter=30, return_H=True)   # K = 32: the largest Kp
    obj = np.stack([v for _, v in res.trace])
    assert np.all(np.isfinite(obj)) and np.all(np.diff(obj, axis=0) <= 1e-6 * obj[:-1])
    assert np.all(res.W[:, :, 17] == 0.0)    # MU keeps an all-zero column at zero
    assert np.all(res.H >= 0) and np.all(res.W >= 0)


def test_single_replicate_and_k1():
    from cnmf_torch_b200.engine import DenseFrobeniusEngine
    from cnmf_torch_b200.synth import nb_counts
    from oracle import nmf_oracle as no
    x = nb_counts(150, 90, k_true=3, seed=1, libsize=300.0)
    eng = DenseFrobeniusEngine(x)
    H0, W0 = eng.make_init(1, [4])
    res = eng.factorize(1, H0=H0, W0=W0, algo="mu", tol=0.0, max_iter=20)
    ora = no.nmf_batch(x, H0[0].numpy(), W0[0].numpy(), algo="mu", tol=0.0, max_iter=20)
    assert abs(res.err[0] - ora["err"]) / ora["err"] < 1e-3
    with pytest.raises(ValueError):
        eng.factorize(33, seeds=[1])          # beyond the engine limit: refused, not truncated


def test_consensus_is_invariant_to_row_order_and_rejects_k_above_32():
    from cnmf_torch_b200 import consensus as cs
    from cnmf_torch_b200.synth import replicate_spectra
    s = replicate_spectra(20, 200, 2000, seed=5)
    n = int(0.3 * s.shape[0] / 20)
    a = cs.consensus_front_half(s, 20, 0.5, n)
    perm = np.random.default_rng(0).permutation(s.shape[0])
    b = cs.consensus_front_half(s[perm], 20, 0.5, n)
    np.testing.assert_allclose(b["local_density"], a["local_density"][perm], rtol=1e-6)
    assert np.array_equal(b["keep"], a["keep"][perm])
    # the partition is the same (cluster numbering depends on the seeding order, the medians do not)
    ma = np.sort(a["median_spectra"] @ np.arange(2000))
    mb = np.sort(b["median_spectra"] @ np.arange(2000))
    np.testing.assert_allclose(mb, ma, rtol=1e-5)
    with pytest.raises(NotImplementedError):
        cs.kmeans_fit(torch.zeros(100, 8, device="cuda"), 33)
