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

import numpy as np


def similarity_topk(X: np.ndarray, topk, eps: float) -> np.ndarray:
    """Dense symmetric top-k cosine similarity, zero diagonal (als.py:220-240).

    dtype follows X (float32 features → float32 S), as in the reference.
    """
    row_norm = np.sqrt((X * X).sum(axis=1, keepdims=True)) + eps
    unit = X / row_norm
    S = unit @ unit.T
    np.fill_diagonal(S, 0.0)
    n = S.shape[0]
    if topk is not None and topk < n:
        for i in range(n):
            weakest = np.argpartition(S[i], -topk)[:-topk]
            S[i, weakest] = 0.0
    return np.maximum(S, S.T)


def to_csr(S: np.ndarray):
    """CSR of the nonzeros of S: (sptr int64, sidx int32, sval float32)."""
    n = S.shape[0]
    r, c = np.nonzero(S)                       # row-major → ascending columns per row
    sptr = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(np.bincount(r, minlength=n), out=sptr[1:])
    return sptr, c.astype(np.int32), S[r, c].astype(np.float32)


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
