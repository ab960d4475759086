"""GPU test of cnmf_b200_gram_static, the Gram entry point for callers that cannot allocate.

Who needs it: a host program that has already handed all device memory to its own pool (so there is
nothing left for a workspace), or one that must not pay cudaMalloc's device-wide synchronisation in
its loop.  The entry point keeps its split partials in an array built into the library, so the test
checks (through the C ABI, like every other GPU test) that it gives what cnmf_b200_gram gives with a
caller workspace -- twice in a row with different inputs, because the built-in array is shared by all
calls and still holds the previous call's partials -- and that it refuses a problem that does not fit.

Collected last on purpose (file name): a run that stops at the first failure has then already run
every other test.  It passes alone in a fresh process and after the rest of the suite.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _gram_both(L, lib, F, ld, R, Kp, Lc):
    st = torch.cuda.current_stream().cuda_stream
    ref = torch.full((R, Kp, Kp), float("nan"), device="cuda")
    out = torch.full((R, Kp, Kp), float("nan"), device="cuda")
    ws = torch.empty(max(1, lib.cnmf_b200_gram_workspace_bytes(R, Kp, Lc)), dtype=torch.uint8, device="cuda")
    L.check(lib.cnmf_b200_gram(F.data_ptr(), ld, R, Kp, Lc, ref.data_ptr(), ws.data_ptr(), st), "gram")
    L.check(lib.cnmf_b200_gram_static(F.data_ptr(), ld, R, Kp, Lc, out.data_ptr(), st), "gram_static")
    torch.cuda.synchronize()
    return ref, out


def test_gram_static_matches_gram_twice_in_a_row():
    from cnmf_torch_b200 import _lib as L
    lib = L.lib()
    torch.manual_seed(7)
    # first call: larger problem (several column splits) leaves its partials in the built-in array
    R, Kp, Lc = 9, 12, 5000
    ld = (Lc + 3) // 4 * 4
    F1 = torch.rand(R * Kp, ld, device="cuda")
    ref1, out1 = _gram_both(L, lib, F1, ld, R, Kp, Lc)
    # same tolerance as tests/test_gpu_dense.py::test_gram_matches_torch
    torch.testing.assert_close(out1, ref1, rtol=2e-5, atol=1e-5)
    Fr = F1[:, :Lc].reshape(R, Kp, Lc).double()
    torch.testing.assert_close(out1, torch.bmm(Fr, Fr.transpose(1, 2)).float(), rtol=2e-5, atol=1e-5)
    # second call straight after, different shape and data: stale partials must not leak in
    R2, Kp2, Lc2 = 4, 8, 2100
    ld2 = (Lc2 + 3) // 4 * 4
    F2 = torch.rand(R2 * Kp2, ld2, device="cuda") * 3.0
    ref2, out2 = _gram_both(L, lib, F2, ld2, R2, Kp2, Lc2)
    torch.testing.assert_close(out2, ref2, rtol=2e-5, atol=1e-5)
    # and the first problem again gives the first answer again
    _, out1b = _gram_both(L, lib, F1, ld, R, Kp, Lc)
    assert torch.equal(out1b, out1)


def test_gram_static_rejects_oversize_problem():
    from cnmf_torch_b200 import _lib as L
    lib = L.lib()
    cap = lib.cnmf_b200_gram_static_max_floats()
    assert cap == 1 << 20
    R, Kp = 64, 32                      # 64 * ceil(L/256) * 1024 floats > 2^20 once L > 16 * 256
    Lc = 17 * 256
    F = torch.zeros(8, 8, device="cuda")  # never read: the size check comes first
    out = torch.zeros(R, Kp, Kp, device="cuda")
    rc = lib.cnmf_b200_gram_static(F.data_ptr(), Lc, R, Kp, Lc, out.data_ptr(), None)
    assert rc == L.ERR_WORKSPACE
    msg = lib.cnmf_b200_last_error().decode()
    assert "built-in array holds" in msg and "cnmf_b200_gram" in msg
