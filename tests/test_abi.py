"""CPU: the C-ABI library loads and exports every symbol include/cnmf_b200.h declares; the ctypes
table covers the same set.  No compute calls (there is no GPU in this run)."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    names = set()
    for fn in os.listdir(os.path.join(ROOT, "include")):
        if fn.endswith(".h"):
            text = open(os.path.join(ROOT, "include", fn)).read()
            text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
            names |= set(re.findall(r"\b(cnmf_b200_[a-z0-9_]+)\s*\(", text))
    return names


@pytest.fixture(scope="module")
def built():
    from cnmf_torch_b200 import build
    return build.build()


def test_header_symbols_exported(built):
    import ctypes
    handle = ctypes.CDLL(built)
    declared = _declared()
    assert len(declared) >= 10
    for name in sorted(declared):
        assert hasattr(handle, name), f"{name} declared in include/ but not exported"


def test_ctypes_table_matches_header(built):
    from cnmf_torch_b200 import _lib
    assert set(_lib.SIGNATURES) == _declared()
    assert _lib.lib().cnmf_b200_abi_version() == 1


def test_argument_validation_without_gpu(built):
    """Validation happens before any CUDA call, so bad arguments are reported even here."""
    from cnmf_torch_b200 import _lib
    import ctypes as C
    lib = _lib.lib()
    rc = lib.cnmf_b200_gemm_nt(None, 4, 4, None, 4, 4, 4, None, 4, None)
    assert rc == _lib.ERR_INVALID and b"null" in lib.cnmf_b200_last_error()
    p = _lib.FrobProblem(n_cells=10, n_genes=10, R=1, K=5, Kp=6)
    assert lib.cnmf_b200_frob_workspace_bytes(C.byref(p)) == _lib.ERR_INVALID
    assert b"multiple of 4" in lib.cnmf_b200_last_error()
    with pytest.raises(_lib.Cnmfb200Error):
        _lib.check(rc, "gemm_nt")


def test_struct_layout_matches_header():
    """sizeof / offsets of the ctypes mirror equal what the C compiler lays out."""
    import subprocess, tempfile, ctypes as C
    from cnmf_torch_b200 import _lib
    src = r'''
#include <stdio.h>
#include <stddef.h>
#include "cnmf_b200.h"
int main(void){printf("%zu %zu %zu %zu %zu\n", sizeof(cnmf_b200_frob_problem),
  offsetof(cnmf_b200_frob_problem, Wall), offsetof(cnmf_b200_frob_problem, active),
  offsetof(cnmf_b200_frob_problem, normX2), offsetof(cnmf_b200_frob_problem, workspace_bytes));
  printf("%zu %zu %zu %zu\n", sizeof(cnmf_b200_kl_problem), offsetof(cnmf_b200_kl_problem, impl),
  offsetof(cnmf_b200_kl_problem, sumX), offsetof(cnmf_b200_kl_problem, workspace_bytes));
  printf("%zu %zu %zu %zu %zu\n", sizeof(cnmf_b200_online_chunk_args), offsetof(cnmf_b200_online_chunk_args, Wall),
  offsetof(cnmf_b200_online_chunk_args, active), offsetof(cnmf_b200_online_chunk_args, max_iter),
  offsetof(cnmf_b200_online_chunk_args, ws_gram));return 0;}
'''
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "t.c"), "w").write(src)
        subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), os.path.join(d, "t.c"), "-o",
                        os.path.join(d, "t")], check=True)
        out = subprocess.run([os.path.join(d, "t")], capture_output=True, text=True, check=True).stdout.split()
    P, Q, O = _lib.FrobProblem, _lib.KlProblem, _lib.OnlineChunkArgs
    assert [int(v) for v in out] == [C.sizeof(P), P.Wall.offset, P.active.offset, P.normX2.offset,
                                     P.workspace_bytes.offset,
                                     C.sizeof(Q), Q.impl.offset, Q.sumX.offset, Q.workspace_bytes.offset,
                                     C.sizeof(O), O.Wall.offset, O.active.offset, O.max_iter.offset, O.ws_gram.offset]


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    """No fallback: without the built CUDA library the product refuses to run."""
    from cnmf_torch_b200 import _lib
    monkeypatch.setattr(_lib, "_LIB", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "absent.so"))
    with pytest.raises(_lib.Cnmfb200Error, match="no CPU or PyTorch fallback"):
        _lib.lib()


def test_engines_refuse_to_run_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip("needs a machine without a GPU")
    import numpy as np
    from cnmf_torch_b200 import _lib
    from cnmf_torch_b200.engine import DenseFrobeniusEngine
    from cnmf_torch_b200.kl_engine import SparseKLEngine
    from cnmf_torch_b200.consensus import consensus_front_half
    x = np.ones((8, 6), dtype=np.float32)
    for fn in (lambda: DenseFrobeniusEngine(x), lambda: SparseKLEngine(x), lambda: consensus_front_half(x, 2, 0.5, 1)):
        with pytest.raises(_lib.Cnmfb200Error, match="CUDA device required"):
            fn()
