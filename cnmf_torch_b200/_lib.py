"""ctypes binding of libcnmf_b200.so (the C ABI declared in include/cnmf_b200.h).

There is no fallback: if the library has not been built (``python -m cnmf_torch_b200.build``)
importing this module's ``lib()`` raises.
"""
from __future__ import annotations

import ctypes as C
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "lib", "libcnmf_b200.so")

ALGO_MU, ALGO_HALS = 0, 1
IMPL_AUTO, IMPL_SIMT, IMPL_TCGEN05 = 0, 1, 2
KL_IMPL_GATHER, KL_IMPL_SLAB = 0, 1
MAX_KP = 32

ERR_INVALID, ERR_CUDA, ERR_WORKSPACE, ERR_UNSUPPORTED = -1, -2, -3, -4

i32, i64, f32, f64, vp = C.c_int32, C.c_int64, C.c_float, C.c_double, C.c_void_p


class FrobProblem(C.Structure):
    """Mirror of ``cnmf_b200_frob_problem``."""
    _fields_ = [
        ("n_cells", i64), ("n_genes", i64),
        ("X", vp), ("ldx", i64), ("Xt", vp), ("ldxt", i64),
        ("R", i32), ("K", i32), ("Kp", i32), ("algo", i32), ("impl", i32), ("reserved", i32),
        ("Wall", vp), ("ldw", i64), ("Ht", vp), ("ldh", i64),
        ("gramW", vp), ("gramH", vp), ("active", vp),
        ("l1_H", f32), ("l2_H", f32), ("l1_W", f32), ("l2_W", f32), ("eps", f32), ("pad0", f32),
        ("normX2", f64), ("err2", vp), ("workspace", vp), ("workspace_bytes", i64), ("rep_k", vp),
    ]


class KlProblem(C.Structure):
    """Mirror of ``cnmf_b200_kl_problem``."""
    _fields_ = [
        ("n_cells", i64), ("n_genes", i64), ("nnz", i64),
        ("row_ptr", vp), ("col_idx", vp), ("val", vp),
        ("t_row_ptr", vp), ("t_col_idx", vp), ("t_val", vp),
        ("R", i32), ("K", i32), ("Kp", i32), ("impl", i32),
        ("Hr", vp), ("Wt", vp),
        ("active", vp),
        ("l1_H", f32), ("l2_H", f32), ("l1_W", f32), ("l2_W", f32), ("eps", f32), ("warm", i32),
        ("sumX", f64), ("xlogx", f64), ("err2", vp), ("workspace", vp), ("workspace_bytes", i64),
    ]


class OnlineChunkArgs(C.Structure):
    """Mirror of ``cnmf_b200_online_chunk_args``."""
    _fields_ = [
        ("x", vp), ("ldx", i64), ("xt", vp), ("ldxt", i64), ("c", i64), ("n_genes", i64), ("R", i32), ("Kp", i32),
        ("Wall", vp), ("Wop", vp), ("ldw", i64), ("Ht", vp), ("ldh", i64), ("Hs", vp), ("Hopc", vp), ("Pt", vp),
        ("ldp", i64), ("Wlo", vp), ("cross", vp), ("gramH_acc", vp),
        ("gramW", vp), ("A", vp), ("Ac", vp), ("Bst", vp), ("Bc", vp), ("active", vp),
        ("l1_H", f32), ("l2_H", f32), ("l1_W", f32), ("l2_W", f32), ("eps", f32), ("h_tol", f32), ("w_tol", f32),
        ("max_iter", i32), ("iters_h", vp), ("iters_w", vp), ("ws_loop", vp), ("ws_gram", vp),
    ]


ONLINE_REFRESH, ONLINE_USAGE, ONLINE_SPECTRA = 0, 1, 2

# name -> (restype, argtypes).  Every symbol include/cnmf_b200.h declares must be listed here;
# tests/test_abi.py checks the header against this table and the built library.
SIGNATURES = {
    "cnmf_b200_abi_version": (i32, []),
    "cnmf_b200_last_error": (C.c_char_p, []),
    "cnmf_b200_last_launch_count": (i64, []),
    "cnmf_b200_sumsq_workspace_bytes": (i64, []),
    "cnmf_b200_sumsq": (i32, [vp, i64, i64, i64, vp, vp, vp]),
    "cnmf_b200_gram_workspace_bytes": (i64, [i32, i32, i64]),
    "cnmf_b200_gram": (i32, [vp, i64, i32, i32, i64, vp, vp, vp]),
    "cnmf_b200_gram_static_max_floats": (i64, []),
    "cnmf_b200_gram_static": (i32, [vp, i64, i32, i32, i64, vp, vp]),
    "cnmf_b200_gemm_nt": (i32, [vp, i64, i64, vp, i64, i64, i64, vp, i64, vp]),
    "cnmf_b200_row_update_blocks": (i64, [i64]),
    "cnmf_b200_row_update": (i32, [vp, i64, vp, i64, vp, vp, i64, i32, i32, i32, f32, f32, f32, vp, vp]),
    "cnmf_b200_round_tf32": (i32, [vp, i64, vp]),
    "cnmf_b200_init_random": (i32, [vp, i32, i32, i32, i64, i64, i64, i64, f32, vp, i64, vp, i64, vp]),
    "cnmf_b200_dense_half_step": (i32, [vp, i64, i64, i64, vp, i64, vp, vp, i64, vp, vp, i32, i32, i32, f32, f32,
                                        f32, vp, vp]),
    "cnmf_b200_dense_half_step_k": (i32, [vp, i64, i64, i64, vp, i64, vp, vp, i64, vp, vp, vp, i32, i32, i32, f32, f32,
                                          f32, vp, vp]),
    "cnmf_b200_frob_objective_parts": (i32, [C.POINTER(FrobProblem), vp, vp]),
    "cnmf_b200_frob_objective_combine": (i32, [f64, vp, vp, vp, i32, i32, f32, f32, f32, f32, vp, vp]),
    "cnmf_b200_dense_gemm_tf32": (i32, [vp, i64, i64, i64, vp, i64, vp, i64, vp, i32, i32, i32, vp]),
    "cnmf_b200_round_tf32_2d": (i32, [vp, vp, i64, i64, i64, vp]),
    "cnmf_b200_axpy": (i32, [vp, vp, i64, vp]),
    "cnmf_b200_online_inner_workspace_bytes": (i64, [i64, i32]),
    "cnmf_b200_online_inner": (i32, [vp, i64, vp, i64, vp, vp, vp, i64, i32, i32, f32, f32, f32, f32, i32, vp, vp,
                                     vp, vp]),
    "cnmf_b200_online_loop_workspace_bytes": (i64, [i32]),
    "cnmf_b200_online_loop": (i32, [vp, i64, vp, i64, vp, vp, i64, i32, i32, f32, f32, f32, f32, i32, vp, vp, vp, vp]),
    "cnmf_b200_online_chunk": (i32, [C.POINTER(OnlineChunkArgs), i32, vp]),
    "cnmf_b200_kl_constants": (i32, [vp, i64, vp, vp]),
    "cnmf_b200_to_half": (i32, [vp, vp, i64, vp]),
    "cnmf_b200_write_npz_batch": (i32, [i64, vp, vp, vp, vp, vp, vp, vp, C.c_char_p, i64, i64, C.c_uint32, i32, i32]),
    "cnmf_b200_write_tsv": (i32, [C.c_char_p, C.c_char_p, i64, C.c_char_p, vp, vp, i32, i64, i64, i64, i32]),
    "cnmf_b200_kl_half_step": (i32, [vp, vp, vp, i64, vp, vp, vp, vp, vp, i32, i32, f32, f32, f32, vp]),
    "cnmf_b200_kl_workspace_bytes": (i64, [C.POINTER(KlProblem)]),
    "cnmf_b200_kl_iterate": (i32, [C.POINTER(KlProblem), i32, vp]),
    "cnmf_b200_kl_objective": (i32, [C.POINTER(KlProblem), vp]),
    "cnmf_b200_kl_range_flag": (i32, [C.POINTER(KlProblem), C.POINTER(i32), vp]),
    "cnmf_b200_kl_range_flag_clear": (i32, [C.POINTER(KlProblem), vp]),
    "cnmf_b200_l2_normalize": (i32, [vp, i64, i32, i32, vp, i64, vp]),
    "cnmf_b200_kmeanspp_workspace_bytes": (i64, [i32, i32, i32]),
    "cnmf_b200_kmeanspp": (i32, [vp, i64, i32, i32, i32, i32, i32, vp, vp, vp, vp, vp, vp]),
    "cnmf_b200_pairwise_dist": (i32, [vp, i64, i32, i32, vp, i64, vp, vp]),
    "cnmf_b200_topk_density": (i32, [vp, i64, i32, i32, vp, vp, vp]),
    "cnmf_b200_col_center": (i32, [vp, i64, i32, i32, vp, vp]),
    "cnmf_b200_kmeans_workspace_bytes": (i64, [i32, i32, i32]),
    "cnmf_b200_kmeans_assign": (i32, [vp, i64, i32, i32, vp, i32, i32, vp, vp, vp, vp]),
    "cnmf_b200_kmeans_update": (i32, [vp, i64, i32, i32, vp, i32, i32, vp, vp, vp, vp, vp, vp, vp]),
    "cnmf_b200_kmeans_inertia_blocks": (i64, [i32]),
    "cnmf_b200_kmeans_inertia": (i32, [vp, i64, i32, i32, vp, i32, i32, vp, vp, vp]),
    "cnmf_b200_cluster_median": (i32, [vp, i64, i32, vp, vp, i32, i32, vp, vp, vp]),
    "cnmf_b200_silhouette": (i32, [vp, i64, i32, vp, vp, i32, vp, vp]),
    "cnmf_b200_ols_xty": (i32, [vp, vp, vp, i32, vp, i32, vp, vp, vp, vp]),
    "cnmf_b200_frob_workspace_bytes": (i64, [C.POINTER(FrobProblem)]),
    "cnmf_b200_frob_prepare": (i32, [C.POINTER(FrobProblem), vp]),
    "cnmf_b200_frob_iterate": (i32, [C.POINTER(FrobProblem), i32, vp]),
    "cnmf_b200_frob_objective": (i32, [C.POINTER(FrobProblem), vp]),
}

_LIB = None


class Cnmfb200Error(RuntimeError):
    pass


def lib() -> C.CDLL:
    """Load the shared library once; raise loudly if it is absent."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise Cnmfb200Error(
                f"{LIB_PATH} is missing: build it with `python -m cnmf_torch_b200.build` "
                "(nvcc, sm_100a). There is no CPU or PyTorch fallback for this path.")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _LIB = handle
    return _LIB


def check(rc: int, what: str = "") -> int:
    """Raise on a negative return code, with the library's own message."""
    if rc < 0:
        msg = lib().cnmf_b200_last_error().decode(errors="replace")
        raise Cnmfb200Error(f"{what or 'libcnmf_b200'} failed (rc={rc}): {msg}")
    return rc


def ptr(t) -> int:
    """Device pointer of a torch tensor (or None -> NULL)."""
    return None if t is None else t.data_ptr()
