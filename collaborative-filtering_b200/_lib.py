"""ctypes binding of libb200als.so (the C-ABI in include/b200als.h).

There is no CPU fallback: if the shared library is missing, or a compute entry
point is called without a B200 current, this raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200als.so")
ABI_VERSION = 8
MAX_K = 160
MAX_FEATURES = 8
STATUS_WORDS = 4
MAX_PEERS = 7

STEP_SOLVE, STEP_FACTOR = 0, 1
PATH_AUTO, PATH_FFMA, PATH_UMMA, PATH_F64 = 0, 1, 2, 3

c_i64p = C.c_void_p   # device pointers travel as integers
vp = C.c_void_p


class B200ALSError(RuntimeError):
    pass


class SplitPlan(C.Structure):
    """mirror of struct b200als_split_plan (include/b200als.h)"""
    _fields_ = [
        ("items", vp), ("hdr", vp), ("slot_np", vp), ("slot_ctr", vp), ("slot_acc", vp),
        ("n_list", C.c_int64),
        ("slot_floats", C.c_int32), ("thresh", C.c_int32), ("piece", C.c_int32), ("reserved", C.c_int32),
    ]


class StepArgs(C.Structure):
    """Mirror of `struct b200als_step_args` (include/b200als.h)."""
    _fields_ = [
        ("struct_bytes", C.c_uint32), ("k", C.c_int32), ("mode", C.c_int32), ("path", C.c_int32),
        ("n_rows", C.c_int64),
        ("indptr", vp), ("indices", vp), ("vals", vp),
        ("rows", vp), ("n_list", C.c_int64),
        ("F", vp), ("ldf", C.c_int64), ("n_f_rows", C.c_int64),
        ("bias_other", vp), ("mu", vp),
        ("reg", C.c_float), ("reg_row", vp),
        ("lam_bias", C.c_float), ("lam_bias_row", vp),
        ("X", vp), ("ldx", C.c_int64),
        ("bias_own", vp),
        ("chol_out", vp), ("aux_out", vp),
        ("status", vp), ("stats_acc", vp), ("norm_acc", vp),
        ("split", vp),
        ("workspace", vp), ("workspace_bytes", C.c_size_t),
        ("row_ids", vp), ("n_peers", C.c_int32), ("reserved0", C.c_int32),
        ("peer_X", vp * MAX_PEERS), ("peer_bias", vp * MAX_PEERS),
    ]


# name -> argtypes (restype is int unless listed in _RESTYPES)
_SIGNATURES = {
    "b200als_abi_version": [],
    "b200als_last_error": [],
    "b200als_device_ok": [],
    "b200als_peer_enable": [C.c_int32],
    "b200als_ipc_export": [vp, C.c_char_p, C.POINTER(C.c_int64)],
    "b200als_ipc_import": [C.c_char_p, C.c_int64, C.POINTER(vp)],
    "b200als_ipc_release": [C.c_char_p],
    "b200als_last_launch_count": [],
    "b200als_last_path": [],
    "b200als_dense_indptr": [vp, C.c_int64, C.c_int64, vp, vp, vp],
    "b200als_dense_fill": [vp, C.c_int64, C.c_int64, vp, vp, vp, vp, vp, vp, vp, vp],
    "b200als_coo_workspace_bytes": [C.c_int64, C.c_int64, C.c_int64, C.POINTER(C.c_size_t)],
    "b200als_coo_to_csr_csc": [vp, vp, C.c_int64, C.c_int64, C.c_int64, C.c_int,
                               vp, vp, vp, vp, vp, vp, vp, vp, C.c_size_t, vp, vp],
    "b200als_row_order_workspace_bytes": [C.c_int64, C.POINTER(C.c_size_t)],
    "b200als_row_order": [vp, C.c_int64, vp, vp, vp, C.c_size_t, vp],
    "b200als_row_order_static_max_rows": [],
    "b200als_row_order_static": [vp, C.c_int64, vp, vp, vp],
    "b200als_compose_z": [vp, C.c_int64, C.c_int64, C.c_int32, C.c_int32,
                          C.POINTER(vp), C.POINTER(C.c_int32), C.POINTER(vp),
                          C.c_int64, C.c_int32, vp, C.c_int64, vp],
    "b200als_step_workspace_bytes": [C.c_int32, C.c_int32, C.c_int64, C.POINTER(C.c_size_t)],
    "b200als_split_default": [C.c_int32],
    "b200als_split_plan_bytes": [C.c_int32, C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_int64,
                                 C.POINTER(C.c_size_t)],
    "b200als_split_plan_build": [vp, vp, C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_int64, vp,
                                 C.c_size_t, C.POINTER(SplitPlan), vp],
    "b200als_row_step": [C.POINTER(StepArgs), vp],
    "b200als_graph_level_solve": [C.c_int32, vp, C.c_int64, vp, vp, vp, vp, vp,
                                  C.c_float, C.c_float, vp, vp, C.c_int64, vp, vp],
    "b200als_graph_sweep": [C.c_int32, vp, C.c_int32, C.c_int64, vp, vp, vp, vp, vp, vp,
                            C.c_float, C.c_float, vp, vp, C.c_int64, vp, vp],
    "b200als_residual_stats": [C.c_int64, vp, vp, vp, C.c_int32, vp, C.c_int64, vp, C.c_int64,
                               vp, vp, vp, vp],
    "b200als_sumsq": [vp, C.c_int64, C.c_int64, C.c_int64, vp, vp],
    "b200als_finalize_iter": [vp, C.c_int64, vp, vp, vp],
    "b200als_finalize_iter_fused": [vp, vp, C.c_int64, vp, vp, vp],
    "b200als_item_gram_f64": [C.c_int64, vp, vp, vp, C.c_int32, vp, C.c_int64, vp, vp, vp, vp, vp, vp],
    "b200als_wstep_system": [C.c_int64, C.c_int32, C.c_int32, vp, vp, vp, vp, C.c_int64,
                             C.c_double, vp, vp, vp],
    "b200als_cholesky_solve_f64": [vp, vp, C.c_int32, vp, vp],
    "b200als_f64_to_f32": [vp, vp, C.c_int64, vp],
    "b200als_predict_dense": [vp, C.c_int64, vp, C.c_int64, vp, vp, vp,
                              C.c_int64, C.c_int64, C.c_int32, vp, vp],
    "b200als_predict_pairs": [vp, C.c_int64, vp, C.c_int64, vp, vp, vp, vp, vp,
                              C.c_int64, C.c_int32, vp, vp],
    "b200als_fold_split_workspace_bytes": [C.c_int64, C.POINTER(C.c_size_t)],
    "b200als_fold_split": [vp, vp, C.c_int64, vp, C.c_int64, vp, vp, vp, vp, vp, vp, C.c_size_t, vp],
    "b200als_sse_on_flat": [vp, C.c_int64, vp, C.c_int64, vp, vp, vp, vp, vp, C.c_int64, C.c_int64, C.c_int32,
                            vp, C.c_int32, vp, vp, vp],
    "b200als_normalize_workspace_bytes": [C.c_int64, C.POINTER(C.c_size_t)],
    "b200als_normalize_feature": [vp, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_double, vp, vp, vp, C.c_size_t, vp],
    "b200als_similarity_topk_rows": [vp, C.c_int64, C.c_int32, C.c_int32, C.c_float, vp, vp, C.c_int64, vp, vp, vp],
}
_RESTYPES = {"b200als_last_error": C.c_char_p, "b200als_row_order_static_max_rows": C.c_int64}

EXPORTED_SYMBOLS = tuple(_SIGNATURES)

_lib = None


def load():
    """Load libb200als.so (idempotent).  Raises B200ALSError if it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B200ALSError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            f"(or `make -C collaborative-filtering_b200/csrc`). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, argtypes in _SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is missing
        fn.argtypes = argtypes
        fn.restype = _RESTYPES.get(name, C.c_int)
    got = lib.b200als_abi_version()
    if got != ABI_VERSION:
        raise B200ALSError(f"libb200als ABI {got} != binding ABI {ABI_VERSION}: rebuild the library")
    _lib = lib
    return lib


def last_error() -> str:
    return (load().b200als_last_error() or b"").decode("utf-8", "replace")


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        raise B200ALSError(f"{what or 'libb200als'} failed (code {rc}): {last_error()}")


def require_device() -> None:
    """Fail loudly unless a B200 (sm_100) is the current CUDA device."""
    lib = load()
    if not lib.b200als_device_ok():
        raise B200ALSError("no usable B200 device: " + last_error() + " — the ALS path has no CPU fallback")


def ptr(t):
    """Device pointer of a torch tensor (or None)."""
    return None if t is None else t.data_ptr()
