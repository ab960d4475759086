"""GPU parity of the tcgen05/TMA path (through the C ABI).

* the fused half-step against a float64 restatement built from the oracle's elementary updates, on
  operands that are exactly representable in tf32 (so the only difference is fp32 accumulation
  order): rtol 2e-4;
* the whole engine (tf32 operands, fp32 accumulate) against the fp32 oracle from identical H0/W0:
  objective within 1e-3 relative at every check, final spectra Pearson >= 0.999 (north_star).
"""
import numpy as np
import pytest
import torch

from oracle import nmf_oracle as no

pytestmark = pytest.mark.gpu


def _lib():
    from cnmf_torch_b200 import _lib as L
    return L, L.lib()


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _trunc_tf32(a: np.ndarray) -> np.ndarray:
    return (a.view(np.uint32) & np.uint32(0xFFFFE000)).view(np.float32)


def _pearson(a, b):
    return float(np.corrcoef(a.ravel().astype(np.float64), b.ravel().astype(np.float64))[0, 1])


def test_round_tf32_is_round_to_nearest():
    L, lib = _lib()
    x = torch.rand(10007, device="cuda") * 3 + 1e-3
    y = x.clone()
    L.check(lib.cnmf_b200_round_tf32(y.data_ptr(), y.numel(), _stream()))
    torch.cuda.synchronize()
    yb = y.cpu().numpy().view(np.uint32)
    assert np.all((yb & 0x1FFF) == 0)                        # representable in tf32
    rel = ((y - x).abs() / x).max().item()
    assert rel <= 2.0 ** -11 * 1.0001                        # nearest, not truncation (2^-10)
    assert abs(((y - x) / x).mean().item()) < 2e-5           # unbiased


@pytest.mark.parametrize("algo", ["mu", "hals"])
@pytest.mark.parametrize("M,Kd,R,K", [
    (128, 32, 4, 4), (300, 57, 5, 5), (2700, 2000, 23, 10), (2000, 2700, 9, 7), (1000, 200, 3, 16),
    (513, 96, 2, 32), (257, 40, 50, 12), (129, 2000, 7, 20),
])
def test_fused_half_step_matches_fp64_restatement(algo, M, Kd, R, K):
    L, lib = _lib()
    rng = np.random.default_rng(M + Kd + R + K)
    Kp = (K + 3) // 4 * 4
    lda = (Kd + 3) // 4 * 4
    ldm = (M + 3) // 4 * 4
    A = _trunc_tf32(rng.random((M, Kd)).astype(np.float32) * (rng.random((M, Kd)) < 0.3))
    F = np.zeros((R * Kp, Kd), dtype=np.float32)
    S = np.zeros((R * Kp, M), dtype=np.float32)
    gram = np.zeros((R, Kp, Kp), dtype=np.float32)
    for r in range(R):
        F[r * Kp:r * Kp + K] = _trunc_tf32(np.abs(rng.normal(size=(K, Kd))).astype(np.float32))
        S[r * Kp:r * Kp + K] = np.abs(rng.normal(size=(K, M))).astype(np.float32)
        f = F[r * Kp:(r + 1) * Kp].astype(np.float64)
        gram[r] = (f @ f.T).astype(np.float32)
    active = np.ones(R, dtype=np.uint8)
    if R > 2:
        active[1] = 0
    dA = torch.zeros(M, lda, device="cuda"); dA[:, :Kd] = torch.from_numpy(A).cuda()
    dF = torch.zeros(R * Kp, lda, device="cuda"); dF[:, :Kd] = torch.from_numpy(F).cuda()
    dS = torch.zeros(R * Kp, ldm, device="cuda"); dS[:, :M] = torch.from_numpy(S).cuda()
    dG = torch.from_numpy(gram).cuda(); dAct = torch.from_numpy(active).cuda()
    dOp = dS.clone()
    nblk = lib.cnmf_b200_row_update_blocks(M)
    part = torch.full((nblk, R, 2), float("nan"), dtype=torch.float64, device="cuda")
    l1, l2 = 0.01, 0.02
    algo_id = L.ALGO_MU if algo == "mu" else L.ALGO_HALS
    L.check(lib.cnmf_b200_dense_half_step(dA.data_ptr(), lda, M, Kd, dF.data_ptr(), lda, dS.data_ptr(),
                                          dOp.data_ptr(), ldm, dG.data_ptr(), dAct.data_ptr(), R, Kp, algo_id, l1, l2, 1e-16,
                                          part.data_ptr(), _stream()), "dense_half_step")
    torch.cuda.synchronize()
    got = dS[:, :M].cpu().numpy()
    op = dOp[:, :M].cpu().numpy()
    # operand copy = master rounded to tf32 (nearest) where updated
    assert np.all((op.view(np.uint32)[active.repeat(Kp).astype(bool)] & 0x1FFF) == 0)
    assert np.max(np.abs(op - got) / np.maximum(np.abs(got), 1e-30)) <= 2.0 ** -11 * 1.001
    P = A.astype(np.float64) @ F.astype(np.float64).T            # [M, R*Kp]
    worst = 0.0
    for r in range(R):
        p = torch.from_numpy(P[:, r * Kp:r * Kp + K])
        s = torch.from_numpy(S[r * Kp:r * Kp + K].T.astype(np.float64))
        g = torch.from_numpy(gram[r, :K, :K].astype(np.float64))
        if not active[r]:
            new = s
        elif algo == "mu":
            new = no.mu_apply(s, (p - l1).clamp(min=0), s @ g + l2 * s, 1e-16)
        else:
            new = no.hals_sweep_rows(p, s, g, l1, l2, 1e-16)
        exp = new.numpy()
        g_r = got[r * Kp:r * Kp + K].T
        err = np.abs(g_r - exp).max() / max(1e-30, np.abs(exp).max())
        worst = max(worst, err)
        np.testing.assert_allclose(g_r, exp, rtol=2e-4, atol=2e-5 * np.abs(exp).max(), err_msg=f"replicate {r}")
        assert np.all(got[r * Kp + K:(r + 1) * Kp] == 0.0)
        cross = part[:, r, 0].sum().item(); ssum = part[:, r, 1].sum().item()
        cross_ref = float((new * p).sum()); sum_ref = float(new.sum())
        assert abs(cross - cross_ref) <= 2e-4 * abs(cross_ref) + 1e-6
        assert abs(ssum - sum_ref) <= 2e-4 * abs(sum_ref) + 1e-6
    print(f"[tc half-step {algo} M={M} Kd={Kd} R={R} K={K}] worst rel err {worst:.2e}")


@pytest.fixture(scope="module")
def problem():
    from cnmf_torch_b200.synth import nb_counts
    return nb_counts(700, 420, k_true=6, seed=14, libsize=900.0)


@pytest.mark.parametrize("algo", ["mu", "hals"])
@pytest.mark.parametrize("K", [5, 9, 12])
def test_engine_tc_trace_matches_oracle(problem, algo, K):
    from cnmf_torch_b200.engine import DenseFrobeniusEngine
    L, _ = _lib()
    x = problem
    eng = DenseFrobeniusEngine(x, impl=L.IMPL_TCGEN05)
    seeds = [5, 6, 7, 8, 9]
    H0, W0 = eng.make_init(K, seeds)
    res = eng.factorize(K, H0=H0, W0=W0, algo=algo, tol=0.0, max_iter=50, return_H=True)
    worst = 0.0
    for r in (0, 2, 4):
        ora = no.nmf_batch(x, H0[r].numpy(), W0[r].numpy(), algo=algo, tol=0.0, max_iter=50)
        for (it_o, obj_o), (it_g, obj_g) in zip(ora["trace"], res.trace):
            assert it_o == it_g
            rel = abs(obj_g[r] - obj_o) / obj_o
            worst = max(worst, rel)
            assert rel < 1e-3, (algo, K, r, it_o, obj_g[r], obj_o)
        assert _pearson(res.W[r], ora["W"]) >= 0.999
        assert _pearson(res.H[r], ora["H"]) >= 0.999
    print(f"[tc engine {algo} K={K}] worst objective rel err {worst:.2e}")


def test_engine_tc_matches_simt_engine(problem):
    """Same init through both CUDA paths: tf32 tensor-core path vs fp32 CUDA-core path."""
    from cnmf_torch_b200.engine import DenseFrobeniusEngine
    L, _ = _lib()
    e1 = DenseFrobeniusEngine(problem, impl=L.IMPL_SIMT)
    e2 = DenseFrobeniusEngine(problem, impl=L.IMPL_TCGEN05)
    H0, W0 = e1.make_init(8, list(range(40)))
    a = e1.factorize(8, H0=H0, W0=W0, algo="hals", tol=1e-4, max_iter=200)
    b = e2.factorize(8, H0=H0, W0=W0, algo="hals", tol=1e-4, max_iter=200)
    # stopping is a threshold on a small difference (checked every 10 iterations): most replicates stop at the
    # same check, none more than three checks apart
    assert np.mean(a.n_iter == b.n_iter) >= 0.7 and np.all(np.abs(a.n_iter - b.n_iter) <= 30)
    assert np.max(np.abs(a.err - b.err) / a.err) < 1e-3
    same = a.n_iter == b.n_iter
    for r in np.where(same)[0][:10]:
        assert _pearson(a.W[r], b.W[r]) >= 0.999


@pytest.mark.parametrize("N,G,K,R", [(50, 37, 2, 1), (129, 33, 6, 3), (1000, 31, 3, 7), (77, 500, 32, 2), (300, 130, 17, 5)])
@pytest.mark.parametrize("algo", ["mu", "hals"])
def test_engine_tc_ragged_shapes_match_oracle(N, G, K, R, algo):
    """Ragged and extreme shapes on the tcgen05 path: fewer rows than one tile, leading dimensions that are not a
    multiple of the padding, one replicate, the largest K (32), odd K."""
    from cnmf_torch_b200.engine import DenseFrobeniusEngine
    from cnmf_torch_b200.synth import nb_counts
    L, _ = _lib()
    x = nb_counts(N, G, k_true=min(K, 5), seed=N + G, libsize=40.0 * G ** 0.5)
    eng = DenseFrobeniusEngine(x, impl=L.IMPL_TCGEN05)
    H0, W0 = eng.make_init(K, list(range(R)))
    res = eng.factorize(K, H0=H0, W0=W0, algo=algo, tol=0.0, max_iter=20, return_H=True)
    for r in range(R):
        ora = no.nmf_batch(x, H0[r].numpy(), W0[r].numpy(), algo=algo, tol=0.0, max_iter=20)
        for (it_o, obj_o), (it_g, obj_g) in zip(ora["trace"], res.trace):
            assert it_o == it_g and abs(obj_g[r] - obj_o) / obj_o < 1e-3, (r, it_o, obj_g[r], obj_o)
        assert _pearson(res.W[r], ora["W"]) >= 0.999 and _pearson(res.H[r], ora["H"]) >= 0.999


@pytest.mark.parametrize("algo", ["mu", "hals"])
def test_factorize_many_stacks_match_single_k_runs(problem, algo):
    """factorize_many stacks K values that share a padded Kp into one batch (K = 5..8 -> Kp 8, K = 9, 10 -> Kp 12) and
    runs the stacks on separate streams; every K must come out as if it had been run alone, with the same stopping
    iteration per replicate."""
    from cnmf_torch_b200.engine import DenseFrobeniusEngine
    L, _ = _lib()
    eng = DenseFrobeniusEngine(problem, impl=L.IMPL_TCGEN05)
    ks = [5, 6, 7, 8, 9, 10]
    seeds = {k: [100 * k + i for i in range(7 if k % 2 else 4)] for k in ks}
    many = eng.factorize_many(ks, seeds, algo=algo, tol=1e-4, max_iter=60, return_H=True)
    for k in ks:
        one = eng.factorize(k, seeds=seeds[k], algo=algo, tol=1e-4, max_iter=60, return_H=True)
        assert many[k].W.shape == one.W.shape == (len(seeds[k]), k, problem.shape[1])
        assert np.array_equal(many[k].n_iter, one.n_iter), (k, many[k].n_iter, one.n_iter)
        np.testing.assert_allclose(many[k].err, one.err, rtol=1e-6)
        np.testing.assert_allclose(many[k].W, one.W, rtol=1e-4, atol=1e-6 * np.abs(one.W).max())
        np.testing.assert_allclose(many[k].H, one.H, rtol=1e-4, atol=1e-6 * np.abs(one.H).max())
        # the stack keeps checking while any of its K still runs: the single run's trace is a prefix of it
        n = len(one.trace)
        assert [it for it, _ in many[k].trace][:n] == [it for it, _ in one.trace]
        np.testing.assert_allclose(many[k].trace[n - 1][1], one.trace[n - 1][1], rtol=1e-6)


@pytest.mark.parametrize("algo", ["mu", "hals"])
def test_exact_k_epilogue_is_bit_identical_to_the_padded_one(algo):
    """A stack of K = 9..12 padded to Kp = 12: with rep_k the epilogue updates, loads and stores only each replicate's
    own K components (cnmf_b200_dense_half_step_k); the padded components are exact zeros, so the result must be
    bit-identical to the run over all Kp components, and the padding rows must still be zero."""
    from cnmf_torch_b200 import _lib as L
    from cnmf_torch_b200.engine import DenseFrobeniusEngine, FrobBatch
    from cnmf_torch_b200.synth import nb_counts
    lib = L.lib()
    x = nb_counts(700, 420, k_true=8, seed=14, libsize=900.0)
    eng = DenseFrobeniusEngine(x)
    groups = [(k,) + tuple(eng.make_init(k, [10 * k + i for i in range(5)])) for k in (9, 10, 11, 12)]
    b = FrobBatch(eng, None, None, None, algo, 0.01, 0.02, 0.0, 0.0, 1e-16, groups=groups)
    assert b.rep_k is not None and b.rep_k.tolist() == [9] * 5 + [10] * 5 + [11] * 5 + [12] * 5
    p = b.p
    st = torch.cuda.current_stream().cuda_stream
    Wop = b.Wall.clone()
    L.check(lib.cnmf_b200_round_tf32(Wop.data_ptr(), Wop.numel(), st))
    outs = []
    for with_k in (False, True):
        Hs, Hop = b.Ht.clone(), torch.zeros_like(b.Ht)
        part = torch.zeros(64 * b.R * 2, dtype=torch.float64, device="cuda")
        if with_k:
            L.check(lib.cnmf_b200_dense_half_step_k(p.X, p.ldx, p.n_cells, p.n_genes, Wop.data_ptr(), p.ldw, Hs.data_ptr(),
                                                    Hop.data_ptr(), p.ldh, p.gramW, None, b.rep_k.data_ptr(), p.R, p.Kp,
                                                    p.algo, 0.01, 0.02, p.eps, part.data_ptr(), st))
        else:
            L.check(lib.cnmf_b200_dense_half_step(p.X, p.ldx, p.n_cells, p.n_genes, Wop.data_ptr(), p.ldw, Hs.data_ptr(),
                                                  Hop.data_ptr(), p.ldh, p.gramW, None, p.R, p.Kp, p.algo, 0.01, 0.02,
                                                  p.eps, part.data_ptr(), st))
        torch.cuda.synchronize()
        outs.append((Hs, Hop, part))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    torch.testing.assert_close(outs[0][2], outs[1][2], rtol=1e-12, atol=0)
    Hs = outs[1][0].view(b.R, b.Kp, -1)
    assert not torch.equal(Hs, b.Ht.view(b.R, b.Kp, -1))             # the step really changed the state
    for r, k in enumerate(b.rep_k.tolist()):
        assert torch.all(Hs[r, k:] == 0)
