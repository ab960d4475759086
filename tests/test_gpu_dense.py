"""GPU parity tests of the dense Frobenius path, all through the C ABI (ctypes).

Tolerances (north_star): objective relative error <= 1e-3 at every check, final spectra Pearson
>= 0.999 against the oracle from identical H0/W0.  Elementary kernels are compared to a plain
torch fp32 reference of the same op at fp32 round-off.
"""
import ctypes as C

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


def _pearson(a, b):
    a = a.ravel().astype(np.float64); b = b.ravel().astype(np.float64)
    return float(np.corrcoef(a, b)[0, 1])


@pytest.fixture(scope="module")
def small_problem():
    from cnmf_torch_b200.synth import nb_counts
    return nb_counts(333, 212, k_true=6, seed=14, libsize=800.0)


@pytest.mark.parametrize("M,N,Kd", [(128, 128, 16), (333, 212, 57), (1000, 36, 2000), (5, 3, 7), (257, 129, 130)])
def test_gemm_nt_matches_torch(M, N, Kd):
    L, lib = _lib()
    torch.manual_seed(0)
    lda = ldb = (Kd + 3) // 4 * 4
    ldc = (N + 3) // 4 * 4
    A = torch.rand(M, lda, device="cuda"); B = torch.rand(N, ldb, device="cuda")
    Cm = torch.full((M, ldc), -7.0, device="cuda")
    L.check(lib.cnmf_b200_gemm_nt(A.data_ptr(), lda, M, B.data_ptr(), ldb, N, Kd, Cm.data_ptr(), ldc, _stream()))
    torch.cuda.synchronize()
    ref = (A[:, :Kd].double() @ B[:, :Kd].double().T).float()
    torch.testing.assert_close(Cm[:, :N], ref, rtol=1e-5, atol=1e-4)
    if ldc > N:
        assert torch.all(Cm[:, N:] == -7.0)   # padding untouched


@pytest.mark.parametrize("Kp", [0, -4, 6, 10, 36])
def test_gram_rejects_unsupported_kp(Kp):
    """cnmf_b200_gram used to run the <12> kernel for any Kp <= 12 (wrong stride, out-of-bounds reads)."""
    L, lib = _lib()
    F = torch.rand(64, 64, device="cuda")
    out = torch.zeros(4, 32, 32, device="cuda")
    ws = torch.zeros(1 << 16, device="cuda")
    for fn, args in ((lib.cnmf_b200_gram, (ws.data_ptr(), _stream())), (lib.cnmf_b200_gram_static, (_stream(),))):
        rc = fn(F.data_ptr(), 64, 1, Kp, 64, out.data_ptr(), *args)
        assert rc == L.ERR_INVALID and b"unsupported" in lib.cnmf_b200_last_error()


def test_gram_rejects_misaligned_wide_kp():
    L, lib = _lib()
    F = torch.rand(16 * 67 + 8, device="cuda")
    out = torch.zeros(1, 16, 16, device="cuda")
    ws = torch.zeros(1 << 16, device="cuda")
    rc = lib.cnmf_b200_gram(F.data_ptr(), 67, 1, 16, 64, out.data_ptr(), ws.data_ptr(), _stream())
    assert rc == L.ERR_INVALID and b"16-byte aligned" in lib.cnmf_b200_last_error()
    rc = lib.cnmf_b200_gram(F.data_ptr() + 4, 68, 1, 16, 64, out.data_ptr(), ws.data_ptr(), _stream())
    assert rc == L.ERR_INVALID
    # Kp <= 12 reads scalars: any ld is fine
    out12 = torch.zeros(1, 12, 12, device="cuda")
    L.check(lib.cnmf_b200_gram(F.data_ptr(), 67, 1, 12, 64, out12.data_ptr(), ws.data_ptr(), _stream()))
    torch.cuda.synchronize()
    Fr = F[:12 * 67].reshape(12, 67)[:, :64].double()
    torch.testing.assert_close(out12[0], (Fr @ Fr.T).float(), rtol=2e-5, atol=1e-5)


def test_gemm_nt_rejects_bad_alignment():
    L, lib = _lib()
    A = torch.rand(8, 8, device="cuda")
    rc = lib.cnmf_b200_gemm_nt(A.data_ptr(), 7, 8, A.data_ptr(), 8, 8, 7, A.data_ptr(), 8, _stream())
    assert rc == L.ERR_INVALID and b"multiples of 4" in lib.cnmf_b200_last_error()


@pytest.mark.parametrize("R,Kp,Lc", [(3, 4, 100), (5, 12, 2700), (2, 32, 5000), (7, 8, 63), (1, 16, 4097)])
def test_gram_matches_torch(R, Kp, Lc):
    L, lib = _lib()
    torch.manual_seed(1)
    ld = (Lc + 3) // 4 * 4
    F = torch.rand(R * Kp, ld, device="cuda")
    gram = torch.zeros(R, Kp, Kp, device="cuda")
    ws = torch.empty(max(1, lib.cnmf_b200_gram_workspace_bytes(R, Kp, Lc)), dtype=torch.uint8, device="cuda")
    L.check(lib.cnmf_b200_gram(F.data_ptr(), ld, R, Kp, Lc, gram.data_ptr(), ws.data_ptr(), _stream()))
    torch.cuda.synchronize()
    Fr = F[:, :Lc].reshape(R, Kp, Lc).double()
    ref = torch.bmm(Fr, Fr.transpose(1, 2)).float()
    torch.testing.assert_close(gram, ref, rtol=2e-5, atol=1e-5)


def test_sumsq_fp64():
    L, lib = _lib()
    x = torch.rand(777, 132, device="cuda")
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    ws = torch.empty(lib.cnmf_b200_sumsq_workspace_bytes(), dtype=torch.uint8, device="cuda")
    L.check(lib.cnmf_b200_sumsq(x.data_ptr(), 777, 130, 132, out.data_ptr(), ws.data_ptr(), _stream()))
    ref = (x[:, :130].double() ** 2).sum().item()
    assert abs(out.item() - ref) / ref < 1e-12


@pytest.mark.parametrize("algo", ["mu", "hals"])
@pytest.mark.parametrize("K", [3, 5, 9, 16])
def test_row_update_matches_oracle(algo, K):
    """One half-step (H update) on random data vs the oracle's elementary update."""
    L, lib = _lib()
    rng = np.random.default_rng(K)
    M, G, R = 301, 90, 3
    Kp = (K + 3) // 4 * 4
    x = rng.random((M, G)).astype(np.float32)
    ldm = (M + 3) // 4 * 4
    state = torch.zeros(R * Kp, ldm, device="cuda"); Pt = torch.zeros(R * Kp, ldm, device="cuda")
    gram = torch.zeros(R, Kp, Kp, device="cuda")
    expect, crosses = [], []
    for r in range(R):
        h = np.abs(rng.normal(size=(M, K))).astype(np.float32)
        w = np.abs(rng.normal(size=(K, G))).astype(np.float32)
        if r == 1:
            h[5, 0] = 0.0
        xt, ht, wt = map(torch.from_numpy, (x, h, w))
        p = xt @ wt.T
        s = wt @ wt.T
        state[r * Kp:r * Kp + K, :M] = ht.T.cuda()
        Pt[r * Kp:r * Kp + K, :M] = p.T.cuda()
        gram[r, :K, :K] = s.cuda()
        if algo == "mu":
            denom = ht @ s + 0.1 * ht
            new = no.mu_apply(ht, (p - 0.05).clamp(min=0), denom, 1e-16)
        else:
            new = no.hals_sweep_rows(p, ht, s, 0.05, 0.1, 1e-16)
        expect.append(new.numpy())
        crosses.append(float((new.double() * p.double()).sum()))
    nblk = lib.cnmf_b200_row_update_blocks(M)
    part = torch.zeros(nblk, R, 2, dtype=torch.float64, device="cuda")
    algo_id = L.ALGO_MU if algo == "mu" else L.ALGO_HALS
    L.check(lib.cnmf_b200_row_update(Pt.data_ptr(), ldm, state.data_ptr(), ldm, gram.data_ptr(), None, M, R, Kp,
                                     algo_id, 0.05, 0.1, 1e-16, part.data_ptr(), _stream()))
    torch.cuda.synchronize()
    got = state.cpu().numpy()
    for r in range(R):
        np.testing.assert_allclose(got[r * Kp:r * Kp + K, :M].T, expect[r], rtol=2e-4, atol=1e-6)
        assert np.all(got[r * Kp + K:(r + 1) * Kp] == 0.0)          # padding stays exactly zero
        cross = part[:, r, 0].sum().item()
        assert abs(cross - crosses[r]) / abs(crosses[r]) < 1e-5


@pytest.mark.parametrize("algo", ["mu", "hals"])
@pytest.mark.parametrize("K", [5, 7, 12])
def test_engine_trace_matches_oracle(small_problem, algo, K):
    """R replicates from identical H0/W0: objective at every check within 1e-3 relative, final W
    Pearson >= 0.999 (tol=0 so both run exactly max_iter iterations)."""
    from cnmf_torch_b200.engine import DenseFrobeniusEngine
    from cnmf_torch_b200 import _lib as L
    x = small_problem
    eng = DenseFrobeniusEngine(x, impl=L.IMPL_SIMT)
    seeds = [11, 22, 33]
    H0, W0 = eng.make_init(K, seeds)
    res = eng.factorize(K, H0=H0, W0=W0, algo=algo, tol=0.0, max_iter=40, return_H=True)
    assert res.launches > 0
    for r in range(len(seeds)):
        ora = no.nmf_batch(x, H0[r].numpy(), W0[r].numpy(), algo=algo, tol=0.0, max_iter=40)
        for (it_o, obj_o), (it_g, obj_g) in zip(ora["trace"], res.trace):
            assert it_o == it_g
            assert abs(obj_g[r] - obj_o) / obj_o < 1e-3, (algo, K, r, it_o, obj_g[r], obj_o)
        assert _pearson(res.W[r], ora["W"]) >= 0.999
        assert _pearson(res.H[r], ora["H"]) >= 0.999
        assert abs(res.err[r] - ora["err"]) / ora["err"] < 1e-3


def test_engine_init_matches_oracle_init(small_problem):
    from cnmf_torch_b200.engine import DenseFrobeniusEngine
    eng = DenseFrobeniusEngine(small_problem)
    H0, W0 = eng.make_init(6, [123])
    h, w = no.init_random(float(small_problem.mean(dtype=np.float64)), *small_problem.shape, 6, 123)
    np.testing.assert_allclose(H0[0].numpy(), h, rtol=1e-6)
    np.testing.assert_allclose(W0[0].numpy(), w, rtol=1e-6)


def test_engine_convergence_and_freeze(small_problem):
    """tol > 0: replicates stop at their own iteration (multiple of 10), frozen replicates keep their
    state; iteration counts agree with the oracle within one check period."""
    from cnmf_torch_b200.engine import DenseFrobeniusEngine
    from cnmf_torch_b200 import _lib as L
    x = small_problem
    eng = DenseFrobeniusEngine(x, impl=L.IMPL_SIMT)
    seeds = [1, 2, 3, 4]
    H0, W0 = eng.make_init(5, seeds)
    res = eng.factorize(5, H0=H0, W0=W0, algo="mu", tol=1e-3, max_iter=300)
    assert np.all(res.n_iter % 10 == 0)
    for r in range(len(seeds)):
        ora = no.nmf_batch(x, H0[r].numpy(), W0[r].numpy(), algo="mu", tol=1e-3, max_iter=300)
        assert abs(int(res.n_iter[r]) - ora["n_iter"]) <= 10
        assert abs(res.err[r] - ora["err"]) / ora["err"] < 1e-3
        assert _pearson(res.W[r], ora["W"]) >= 0.999
