"""GPU parity of online mode (the mode cNMF itself configures, cnmf.py:757-771) against the oracle from
identical H0/W0.  Operands are tf32 on the device, so an inner loop can stop one iteration earlier or
later than the fp32 oracle when its relative change sits on the tolerance; the bar is therefore the
north_star one -- objective after every pass within 1e-3 relative, final factors Pearson >= 0.999."""
import numpy as np
import pytest
import torch

from oracle import nmf_oracle as no

pytestmark = pytest.mark.gpu


def _pearson(a, b):
    return float(np.corrcoef(a.ravel().astype(np.float64), b.ravel().astype(np.float64))[0, 1])


@pytest.fixture(scope="module")
def problem():
    from cnmf_torch_b200.synth import nb_counts
    return nb_counts(500, 260, k_true=6, seed=14, libsize=700.0)


def test_gemm_store_and_accumulate_modes():
    from cnmf_torch_b200 import _lib as L
    lib = L.lib()
    st = torch.cuda.current_stream().cuda_stream
    torch.manual_seed(0)
    M, Kd, R, Kp = 300, 164, 5, 12
    A = torch.rand(M, 192, device="cuda"); F = torch.rand(R * Kp, 192, device="cuda")
    L.check(lib.cnmf_b200_round_tf32(A.data_ptr(), A.numel(), st)); L.check(lib.cnmf_b200_round_tf32(F.data_ptr(), F.numel(), st))
    out = torch.full((R * Kp, 320), 2.0, device="cuda")
    active = torch.ones(R, dtype=torch.uint8, device="cuda"); active[3] = 0
    L.check(lib.cnmf_b200_dense_gemm_tf32(A.data_ptr(), 192, M, Kd, F.data_ptr(), 192, out.data_ptr(), 320,
                                          active.data_ptr(), R, Kp, 0, st))
    ref = (F[:, :Kd].double() @ A[:, :Kd].double().T).float()
    mask = active.bool().repeat_interleave(Kp)
    torch.testing.assert_close(out[mask, :M], ref[mask], rtol=1e-5, atol=1e-4)
    assert torch.all(out[~mask] == 2.0) and torch.all(out[:, M:] == 2.0)
    L.check(lib.cnmf_b200_dense_gemm_tf32(A.data_ptr(), 192, M, Kd, F.data_ptr(), 192, out.data_ptr(), 320,
                                          active.data_ptr(), R, Kp, 1, st))
    torch.testing.assert_close(out[mask, :M], 2 * ref[mask], rtol=1e-5, atol=2e-4)


def test_online_inner_loop_matches_reference_fit_H_online(golden_dir):
    """The inner h loop against the REFERENCE's own fit_H_online output (tests/golden/fit_h_online.npz):
    same update, same epsilon rule, same stopping rule, one chunk = all rows."""
    import os
    from cnmf_torch_b200 import _lib as L
    lib = L.lib()
    st = torch.cuda.current_stream().cuda_stream
    g = np.load(os.path.join(golden_dir, "fit_h_online.npz"), allow_pickle=True)
    x, w, h0 = g["x"], g["w"], g["h_init"]
    n, k = h0.shape
    Kp = 8
    for tag, chunk, tol, max_it in (("default", 100, 0.05, 200), ("tight", 64, 1e-3, 50)):
        out = np.zeros_like(h0)
        for lo in range(0, n, chunk):
            c = min(chunk, n - lo)
            ld = (c + 3) // 4 * 4
            P = torch.zeros(Kp, ld, device="cuda"); S = torch.zeros(Kp, ld, device="cuda")
            P[:k, :c] = torch.from_numpy((x[lo:lo + c] @ w.T).T.copy()).cuda()
            S[:k, :c] = torch.from_numpy(np.clip(h0[lo:lo + c], 0, None).T.copy()).cuda()
            gram = torch.zeros(1, Kp, Kp, device="cuda"); gram[0, :k, :k] = torch.from_numpy(w @ w.T).cuda()
            done = torch.zeros(1, dtype=torch.uint8, device="cuda"); iters = torch.zeros(1, dtype=torch.int32, device="cuda")
            npend = torch.zeros(1, dtype=torch.int32, device="cuda")
            ws = torch.empty(int(lib.cnmf_b200_online_inner_workspace_bytes(c, 1)) + 256, dtype=torch.uint8, device="cuda")
            L.check(lib.cnmf_b200_online_inner(P.data_ptr(), ld, S.data_ptr(), ld, gram.data_ptr(), done.data_ptr(), None,
                                               c, 1, Kp, 0.0, 0.0, 1e-16, tol, max_it, iters.data_ptr(),
                                               npend.data_ptr(), ws.data_ptr(), st))
            out[lo:lo + c] = S[:k, :c].T.cpu().numpy()
        np.testing.assert_allclose(out, g["out_" + tag], rtol=2e-4, atol=1e-6)


def test_persistent_loop_matches_reference_fit_H_online_and_the_stepwise_form(golden_dir):
    """cnmf_b200_online_loop (the whole inner loop in one cooperative launch, stopping decided on the device) against
    the REFERENCE's fit_H_online output, and against the two-kernels-per-iteration form for many replicates at once:
    several column slabs per replicate (R small), one slab (R large), more replicates than resident CTAs."""
    import os
    from cnmf_torch_b200 import _lib as L
    lib = L.lib()
    st = torch.cuda.current_stream().cuda_stream
    g = np.load(os.path.join(golden_dir, "fit_h_online.npz"), allow_pickle=True)
    x, w, h0 = g["x"], g["w"], g["h_init"]
    n, k = h0.shape
    Kp = 8
    for tag, chunk, tol, max_it in (("default", 100, 0.05, 200), ("tight", 64, 1e-3, 50)):
        out = np.zeros_like(h0)
        for lo in range(0, n, chunk):
            c = min(chunk, n - lo)
            ld = (c + 3) // 4 * 4
            P = torch.zeros(Kp, ld, device="cuda"); S = torch.zeros(Kp, ld, device="cuda")
            P[:k, :c] = torch.from_numpy((x[lo:lo + c] @ w.T).T.copy()).cuda()
            S[:k, :c] = torch.from_numpy(np.clip(h0[lo:lo + c], 0, None).T.copy()).cuda()
            gram = torch.zeros(1, Kp, Kp, device="cuda"); gram[0, :k, :k] = torch.from_numpy(w @ w.T).cuda()
            iters = torch.zeros(1, dtype=torch.int32, device="cuda")
            ws = torch.empty(int(lib.cnmf_b200_online_loop_workspace_bytes(1)) + 256, dtype=torch.uint8, device="cuda")
            L.check(lib.cnmf_b200_online_loop(P.data_ptr(), ld, S.data_ptr(), ld, gram.data_ptr(), None, c, 1, Kp, 0.0, 0.0,
                                              1e-16, tol, max_it, iters.data_ptr(), None, ws.data_ptr(), st))
            out[lo:lo + c] = S[:k, :c].T.cpu().numpy()
        np.testing.assert_allclose(out, g["out_" + tag], rtol=2e-4, atol=1e-6)
    # many replicates: identical iteration counts and states as the stepwise form
    torch.manual_seed(3)
    for R, M, Kp in ((3, 5000, 16), (50, 2000, 16), (130, 777, 12), (700, 300, 4)):
        ld = (M + 3) // 4 * 4
        P = torch.rand(R * Kp, ld, device="cuda") * 3.0
        S0 = torch.rand(R * Kp, ld, device="cuda") + 0.05
        Gm = torch.rand(R, Kp, Kp, device="cuda")
        Gm = Gm @ Gm.transpose(1, 2) + 0.1 * torch.eye(Kp, device="cuda")
        active = torch.ones(R, dtype=torch.uint8, device="cuda"); active[R // 2] = 0
        tol, max_it = 2e-2, 40
        # stepwise
        S1 = S0.clone()
        done = torch.zeros(R, dtype=torch.uint8, device="cuda"); it1 = torch.zeros(R, dtype=torch.int32, device="cuda")
        npend = torch.zeros(1, dtype=torch.int32, device="cuda")
        ws1 = torch.empty(int(lib.cnmf_b200_online_inner_workspace_bytes(M, R)) + 256, dtype=torch.uint8, device="cuda")
        L.check(lib.cnmf_b200_online_inner(P.data_ptr(), ld, S1.data_ptr(), ld, Gm.data_ptr(), done.data_ptr(),
                                           active.data_ptr(), M, R, Kp, 0.01, 0.02, 1e-16, tol, max_it, it1.data_ptr(),
                                           npend.data_ptr(), ws1.data_ptr(), st))
        # one launch
        S2 = S0.clone()
        it2 = torch.zeros(R, dtype=torch.int32, device="cuda")
        ws2 = torch.empty(int(lib.cnmf_b200_online_loop_workspace_bytes(R)) + 256, dtype=torch.uint8, device="cuda")
        L.check(lib.cnmf_b200_online_loop(P.data_ptr(), ld, S2.data_ptr(), ld, Gm.data_ptr(), active.data_ptr(), M, R, Kp,
                                          0.01, 0.02, 1e-16, tol, max_it, it2.data_ptr(), None, ws2.data_ptr(), st))
        torch.cuda.synchronize()
        assert torch.equal(it1, it2), (R, M, Kp, it1[:8].tolist(), it2[:8].tolist())
        assert int(it2[R // 2]) == 0 and torch.equal(S2[(R // 2) * Kp:(R // 2 + 1) * Kp], S0[(R // 2) * Kp:(R // 2 + 1) * Kp])
        assert 1 < int(it2.max()) and int(it2[it2 > 0].min()) < max_it      # replicates really stop at different times
        torch.testing.assert_close(S2, S1, rtol=0, atol=0)


@pytest.mark.parametrize("K,chunk", [(5, 100), (9, 5000), (6, 248)])
def test_online_engine_matches_oracle(problem, K, chunk):
    from cnmf_torch_b200.engine import DenseFrobeniusEngine
    x = problem
    eng = DenseFrobeniusEngine(x)
    seeds = [3, 4, 5]
    H0, W0 = eng.make_init(K, seeds)
    res = eng.factorize_online(K, H0=H0, W0=W0, tol=0.0, max_pass=4, chunk_size=chunk, chunk_max_iter=60, return_H=True)
    worst = 0.0
    for r in range(3):
        ora = no.nmf_online(x, H0[r].numpy(), W0[r].numpy(), tol=0.0, max_pass=4, chunk_size=chunk, chunk_max_iter=60)
        for (p_o, obj_o), (p_g, obj_g) in zip(ora["trace"], res.trace):
            assert p_o == p_g
            rel = abs(obj_g[r] - obj_o) / obj_o
            worst = max(worst, rel)
            assert rel < 1e-3, (K, chunk, r, p_o, obj_g[r], obj_o)
        assert _pearson(res.W[r], ora["W"]) >= 0.999 and _pearson(res.H[r], ora["H"]) >= 0.999
    print(f"[online K={K} chunk={chunk}] worst objective rel err {worst:.2e}; inner iterations {res.inner_iterations}")


def test_online_with_l1_and_l2_regularisation_matches_oracle(problem):
    """alpha_usage / alpha_spectra (cnmf.py:757-761) in online mode: the L1 / L2 terms of both inner loops and of the
    objective (whose sum(H), sum(W) terms do not come from the refresh kernels)."""
    from cnmf_torch_b200.engine import DenseFrobeniusEngine
    eng = DenseFrobeniusEngine(problem)
    reg = dict(l1_H=0.02, l2_H=0.1, l1_W=0.01, l2_W=0.3)
    H0, W0 = eng.make_init(5, [3, 4])
    kw = dict(tol=0.0, max_pass=3, chunk_size=200, chunk_max_iter=60)
    res = eng.factorize_online(5, H0=H0, W0=W0, return_H=True, **kw, **reg)
    for r in range(2):
        ora = no.nmf_online(problem, H0[r].numpy(), W0[r].numpy(), l1_h=reg["l1_H"], l2_h=reg["l2_H"], l1_w=reg["l1_W"],
                            l2_w=reg["l2_W"], **kw)
        for (p_o, obj_o), (p_g, obj_g) in zip(ora["trace"], res.trace):
            assert p_o == p_g and abs(obj_g[r] - obj_o) / obj_o < 1e-3, (r, p_o, obj_g[r], obj_o)
        assert _pearson(res.W[r], ora["W"]) >= 0.999 and _pearson(res.H[r], ora["H"]) >= 0.999
    plain = eng.factorize_online(5, H0=H0, W0=W0, **kw)
    assert np.all(res.trace[-1][1] != plain.trace[-1][1])


def test_online_frozen_replicates_keep_their_objective(problem):
    """A replicate that converged is skipped by every kernel of later passes: its state, spectra and reported error
    are those of the pass it was frozen at (run to convergence vs the same replicates run alone)."""
    from cnmf_torch_b200.engine import DenseFrobeniusEngine
    eng = DenseFrobeniusEngine(problem)
    seeds = [1, 2, 3, 4, 5, 6]
    res = eng.factorize_online(6, seeds=seeds, tol=2e-3, max_pass=12, chunk_size=100, chunk_max_iter=40)
    assert len(set(res.n_iter.tolist())) > 1 or res.n_iter.max() < 12      # they do not all run to the cap together
    for r in (0, 3):
        one = eng.factorize_online(6, seeds=[seeds[r]], tol=2e-3, max_pass=12, chunk_size=100, chunk_max_iter=40)
        assert one.n_iter[0] == res.n_iter[r]
        assert abs(one.err[0] - res.err[r]) / one.err[0] < 1e-5
        assert _pearson(one.W[0], res.W[r]) >= 0.99999


def test_online_convergence_and_run_nmf_default_kwargs(problem):
    """The exact kwargs cNMF writes to its YAML (cnmf.py:757-771) through the run_nmf drop-in."""
    from cnmf_torch_b200.nmf import run_nmf
    kw = dict(alpha_W=0.0, alpha_H=0.0, l1_ratio_H=0.0, l1_ratio_W=0.0, beta_loss="frobenius", algo="mu", tol=1e-4,
              mode="online", online_chunk_max_iter=1000, online_chunk_size=5000, init="random", n_jobs=-1, use_gpu=True,
              random_state=59886188, n_components=7)
    H, W, err = run_nmf(problem.astype(np.float64), init_device="cpu", **kw)   # host generator, as the oracle
    ho, wo, eo = no.run_nmf_oracle(problem, **{k: v for k, v in kw.items() if k != "use_gpu"})
    assert H.shape == ho.shape and W.shape == wo.shape
    assert abs(err - eo) / eo < 1e-3
    assert _pearson(W, wo) >= 0.999


def test_online_stack_of_several_k_matches_single_k_runs(problem):
    """K = 5, 6, 7 share Kp = 8: run as one stack, each replicate must come out as in its own single-K run."""
    from cnmf_torch_b200.engine import DenseFrobeniusEngine
    eng = DenseFrobeniusEngine(problem)
    seeds = {5: [1, 2], 6: [3], 7: [4, 5]}
    many = eng.factorize_online_many([5, 6, 7], seeds, tol=0.0, max_pass=3, chunk_size=100, chunk_max_iter=40)
    for k, sd in seeds.items():
        one = eng.factorize_online(k, seeds=sd, tol=0.0, max_pass=3, chunk_size=100, chunk_max_iter=40)
        assert many[k].W.shape == one.W.shape
        for r in range(len(sd)):
            assert abs(many[k].err[r] - one.err[r]) / one.err[r] < 1e-4
            assert _pearson(many[k].W[r], one.W[r]) >= 0.9999
