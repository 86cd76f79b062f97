"""Item-similarity graph for the Laplacian term.

The similarity matrix is built on the HOST with the same numpy call sequence
as the reference (`scripts/als.py:220-240`): the per-row top-k uses
`np.argpartition`, whose choice among tied similarities (binary genre vectors
tie massively) is implementation-defined, so only the identical calls give the
identical graph.  It is one-shot setup per `fit`.  What goes to the GPU is
CSR(S), the degree vector and a level schedule that lets the Gauss–Seidel item
sweep (`als.py:453-461`, live V) run level-parallel with sequential semantics.
"""
from __future__ import annotations

from concurrent.futures import ThreadPoolExecutor

import numpy as np

from . import _hostlib

_SMALL = 1 << 20          # below this many entries of S a thread team costs more than the pass


def _threads(S: np.ndarray, n_threads) -> int:
    return 1 if S.size < _SMALL else int(n_threads or _hostlib.host_threads())


def _elem_ok(S: np.ndarray) -> bool:
    return S.ndim == 2 and S.flags.c_contiguous and S.dtype in (np.dtype(np.float32), np.dtype(np.float64))


def similarity_topk(X: np.ndarray, topk, eps: float, n_threads: int | None = None) -> np.ndarray:
    """Dense symmetric top-k cosine similarity, zero diagonal (als.py:220-240).

    dtype follows X (float32 features → float32 S), as in the reference.  Which of the tied similarities
    survive is decided by numpy's own `argpartition` (introselect), as in the reference's
    `np.argpartition(S[i], -topk)`; along the last axis of a 2-D block numpy runs that same selection row by
    row, so blocks of rows go through one call each on a thread pool (the GIL is released inside), and the
    final `np.maximum(S, S.T)` is one in-place pass of libb200als_host.so over tile pairs.
    tests/test_host_graph.py holds this against the row-by-row loop and the reference's fixtures.
    """
    row_norm = np.sqrt((X * X).sum(axis=1, keepdims=True)) + eps
    unit = X / row_norm
    S = unit @ unit.T
    nth = _threads(S, n_threads)
    np.fill_diagonal(S, 0.0)
    n = S.shape[0]
    if topk is not None and topk < n:
        rb = max(16, min(256, (1 << 19) // n))       # rows per call: long enough to amortise the call, short enough to balance

        def prune(lo):
            blk = S[lo:lo + rb]
            order = np.argpartition(blk, -topk, axis=1)                 # row by row the reference's 1-D call
            np.put_along_axis(blk, order[:, :-topk], 0.0, axis=1)       # S[i, weakest] = 0.0

        if nth > 1:
            with ThreadPoolExecutor(nth) as pool:
                list(pool.map(prune, range(0, n, rb)))
        else:
            for lo in range(0, n, rb):
                prune(lo)
    lib = _hostlib.load()
    if lib is None or not _elem_ok(S):
        return np.maximum(S, S.T)
    if lib.b200als_host_symmetrize_max(S.ctypes.data, n, S.dtype.itemsize, nth) != 0:
        raise ValueError("b200als_host_symmetrize_max: bad arguments")
    return S


def to_csr(S: np.ndarray, n_threads: int | None = None):
    """CSR of the nonzeros of S: (sptr int64, sidx int32, sval float32)."""
    n = S.shape[0]
    lib = _hostlib.load()
    if lib is None or not _elem_ok(S):
        r, c = np.nonzero(S)                       # row-major → ascending columns per row
        sptr = np.zeros(n + 1, dtype=np.int64)
        np.cumsum(np.bincount(r, minlength=n), out=sptr[1:])
        return sptr, c.astype(np.int32), S[r, c].astype(np.float32)
    nth = _threads(S, n_threads)
    cols, eb = S.shape[1], S.dtype.itemsize
    sptr = np.zeros(n + 1, dtype=np.int64)
    if lib.b200als_host_dense_row_nnz(S.ctypes.data, n, cols, eb, sptr[1:].ctypes.data, nth) != 0:
        raise ValueError("b200als_host_dense_row_nnz: bad arguments")
    np.cumsum(sptr[1:], out=sptr[1:])
    sidx = np.empty(int(sptr[-1]), dtype=np.int32)
    sval = np.empty(int(sptr[-1]), dtype=np.float32)
    if lib.b200als_host_dense_to_csr(S.ctypes.data, n, cols, eb, sptr.ctypes.data, sidx.ctypes.data, sval.ctypes.data,
                                     nth) != 0:
        raise ValueError("b200als_host_dense_to_csr: bad arguments")
    return sptr, sidx, sval


def level_schedule(sptr: np.ndarray, sidx: np.ndarray, rated: np.ndarray):
    """Levels of the lower-triangular dependency DAG restricted to rated items.

    Item i (rated) must run after every rated neighbour j < i (it reads their
    freshly written V_j, als.py:458-461).  Unrated items are never updated
    (als.py:440-441) and impose no order.  Because S is symmetric, a rated
    neighbour j > i lands in a strictly later level, so i still reads its old
    value — the write-after-read hazard is covered by the same schedule.

    Returns (level_ptr int64 [L+1], level_rows int32) with rows ascending
    inside a level.
    """
    n = sptr.size - 1
    level = np.full(n, -1, dtype=np.int64)
    lib = _hostlib.load()
    if lib is not None:
        sp = np.ascontiguousarray(sptr, dtype=np.int64)
        si = np.ascontiguousarray(sidx, dtype=np.int32)
        ra = np.ascontiguousarray(rated, dtype=np.uint8)
        if lib.b200als_host_graph_levels(sp.ctypes.data, si.ctypes.data, ra.ctypes.data, n, level.ctypes.data) != 0:
            raise ValueError("b200als_host_graph_levels: bad CSR")
    else:
        for i in range(n):
            if not rated[i]:
                continue
            nb = sidx[sptr[i]:sptr[i + 1]]
            nb = nb[nb < i]
            lv = level[nb]
            level[i] = (lv.max() + 1) if lv.size and lv.max() >= 0 else 0
    rows = np.flatnonzero(level >= 0)
    if rows.size == 0:
        return np.zeros(1, dtype=np.int64), np.zeros(0, dtype=np.int32)
    order = np.argsort(level[rows], kind="stable")
    rows = rows[order]
    nlev = int(level.max()) + 1
    level_ptr = np.zeros(nlev + 1, dtype=np.int64)
    np.cumsum(np.bincount(level[rows], minlength=nlev), out=level_ptr[1:])
    return level_ptr, rows.astype(np.int32)
