"""ctypes binding of libb200als_host.so (include/b200als_host.h): host-side array passes of the path on all cores.

This is plumbing around the CUDA path, not a substitute for it: every function here has a numpy expression of the
same values, which callers use when the library has not been built.
"""
from __future__ import annotations

import ctypes as C
import logging
import os

logger = logging.getLogger(__name__)

_HERE = os.path.dirname(os.path.abspath(__file__))
HOST_LIB_PATH = os.path.join(_HERE, "libb200als_host.so")
_lib = None

vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
_SIGNATURES = {
    "b200als_host_normal_fill": [C.c_uint64] * 4 + [i32, vp, vp, vp, C.c_double, i32, vp],
    "b200als_host_symmetrize_max": [vp, i64, i32, i32],
    "b200als_host_dense_row_nnz": [vp, i64, i64, i32, vp, i32],
    "b200als_host_dense_to_csr": [vp, i64, i64, i32, vp, vp, vp, i32],
    "b200als_host_graph_levels": [vp, vp, vp, i64, vp],
    "b200als_host_widen_f32": [vp, i64, vp, i64, i64, i32],
}
EXPORTED_SYMBOLS = tuple(_SIGNATURES)


def load():
    """The library, or None if it has not been built (callers then take their numpy expression)."""
    global _lib
    if _lib is None:
        if not os.path.exists(HOST_LIB_PATH):
            _lib = False
            logger.info("libb200als_host.so not built: host set-up runs in numpy on one core")
        else:
            lib = C.CDLL(HOST_LIB_PATH)
            for name, argtypes in _SIGNATURES.items():
                fn = getattr(lib, name)
                fn.restype = C.c_int
                fn.argtypes = argtypes
            _lib = lib
    return _lib or None


def host_threads(shared: bool = True) -> int:
    """Threads for one host pass: the cores this process may run on (at most 32), shared evenly between the
    ranks of a node under torchrun (LOCAL_WORLD_SIZE) when every rank runs the same pass at the same time — an
    oversubscribed OpenMP team spins at its barriers.  shared=False: the pass runs on one rank only while the
    others wait for its result.  B200ALS_HOST_THREADS overrides."""
    env = os.environ.get("B200ALS_HOST_THREADS")
    if env:
        return max(1, int(env))
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    try:
        local_ranks = max(1, int(os.environ.get("LOCAL_WORLD_SIZE", "1")))
    except ValueError:
        local_ranks = 1
    return max(1, min(n, 32) // (local_ranks if shared else 1))
