"""A4 / A8 host set-up (scripts/als.py:220-240): the threaded similarity build and the C passes of
libb200als_host.so give exactly the arrays of the reference's row-by-row numpy sequence."""
import numpy as np
import pytest

from b200als import _hostlib, graph


def _reference_sequence(X, topk, eps):
    """The reference's statements, one row at a time (scripts/als.py:220-240)."""
    unit = X / (np.sqrt((X * X).sum(axis=1, keepdims=True)) + eps)
    S = unit @ unit.T
    np.fill_diagonal(S, 0.0)
    n = S.shape[0]
    if topk is not None and topk < n:
        for i in range(n):
            weakest = np.argpartition(S[i], -topk)[:-topk]
            S[i, weakest] = 0.0
    return np.maximum(S, S.T)


def _numpy_csr(S):
    n = S.shape[0]
    r, c = np.nonzero(S)
    sptr = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(np.bincount(r, minlength=n), out=sptr[1:])
    return sptr, c.astype(np.int32), S[r, c].astype(np.float32)


def _features(n, d, dtype, signed, seed):
    rng = np.random.default_rng(seed)
    X = (rng.random((n, d)) < 0.15).astype(dtype)          # binary genre-like rows: massive ties, some all-zero rows
    if signed:
        X = X * rng.standard_normal((n, d)).astype(dtype)
    return X


@pytest.mark.parametrize("n,d,dtype,signed", [
    (1, 3, np.float32, False), (37, 5, np.float32, False), (131, 7, np.float64, True),
    (700, 19, np.float32, False), (1100, 19, np.float32, False),      # 1100² > 2^20: thread pool + OpenMP passes
    (1300, 12, np.float64, True),
])
@pytest.mark.parametrize("topk", [1, 5, 40, None])
def test_similarity_and_csr_equal_the_row_by_row_sequence(n, d, dtype, signed, topk):
    X = _features(n, d, dtype, signed, seed=n + (topk or 0))
    want = _reference_sequence(X, topk, 1e-8)
    for nth in (1, 3, 8):
        S = graph.similarity_topk(X, topk, 1e-8, n_threads=nth)
        assert S.dtype == want.dtype and np.array_equal(S, want), (n, topk, nth)
        assert np.array_equal(S, S.T)
        got, ref = graph.to_csr(S, n_threads=nth), _numpy_csr(want)
        assert all(g.dtype == r.dtype and np.array_equal(g, r) for g, r in zip(got, ref))


def test_threading_is_forced_for_large_inputs(monkeypatch):
    """With the size threshold lowered every pass takes its threaded branch on a small case."""
    monkeypatch.setattr(graph, "_SMALL", 1)
    X = _features(257, 9, np.float32, False, seed=5)
    want = _reference_sequence(X, 7, 1e-8)
    S = graph.similarity_topk(X, 7, 1e-8, n_threads=4)
    assert np.array_equal(S, want)
    assert all(np.array_equal(g, r) for g, r in zip(graph.to_csr(S, n_threads=4), _numpy_csr(want)))


def test_numpy_expressions_are_used_without_the_library(monkeypatch):
    monkeypatch.setattr(_hostlib, "load", lambda: None)
    X = _features(90, 6, np.float32, False, seed=2)
    want = _reference_sequence(X, 5, 1e-8)
    S = graph.similarity_topk(X, 5, 1e-8)
    assert np.array_equal(S, want)
    assert all(np.array_equal(g, r) for g, r in zip(graph.to_csr(S), _numpy_csr(want)))


def test_symmetrize_propagates_nan_like_numpy():
    lib = _hostlib.load()
    assert lib is not None, "libb200als_host.so missing: run __graft_entry__.build()"
    rng = np.random.default_rng(0)
    for dt in (np.float32, np.float64):
        A = rng.standard_normal((150, 150)).astype(dt)
        A[rng.random(A.shape) < 0.05] = np.nan
        want = np.maximum(A, A.T)
        assert lib.b200als_host_symmetrize_max(A.ctypes.data, 150, A.dtype.itemsize, 4) == 0
        assert np.array_equal(A, want, equal_nan=True)
    assert lib.b200als_host_symmetrize_max(None, 5, 4, 1) == -1
    assert lib.b200als_host_symmetrize_max(A.ctypes.data, 5, 2, 1) == -1


def test_level_schedule_equals_the_python_loop(monkeypatch):
    rng = np.random.default_rng(4)
    for n, topk, p_rated in ((1, 1, 1.0), (60, 3, 0.7), (900, 5, 0.9), (1200, 40, 0.5), (300, None, 1.0)):
        X = _features(n, 8, np.float32, False, seed=n)
        sptr, sidx, _ = graph.to_csr(graph.similarity_topk(X, topk, 1e-8))
        rated = rng.random(n) < p_rated
        got = graph.level_schedule(sptr, sidx, rated)
        with monkeypatch.context() as mp:
            mp.setattr(_hostlib, "load", lambda: None)
            want = graph.level_schedule(sptr, sidx, rated)
        assert all(g.dtype == w.dtype and np.array_equal(g, w) for g, w in zip(got, want)), n
        # the schedule's defining property: every rated neighbour with a smaller index sits in an earlier level
        level = np.full(n, -1)
        for lv in range(got[0].size - 1):
            level[got[1][got[0][lv]:got[0][lv + 1]]] = lv
        assert np.array_equal(level >= 0, rated)
        for i in np.flatnonzero(rated)[:200]:
            nb = sidx[sptr[i]:sptr[i + 1]]
            nb = nb[(nb < i) & rated[nb]]
            assert (level[nb] < level[i]).all()


def test_readback_float64_copies(monkeypatch):
    """readback.f64_from_device == t.double().cpu().numpy() in both modes (CPU tensors stand in for the device)."""
    import torch
    from b200als import readback
    monkeypatch.delenv("B200ALS_READBACK", raising=False)
    monkeypatch.setattr(readback, "MIN_PIPELINED_ELEMS", 1)
    g = torch.Generator().manual_seed(0)
    base = torch.randn(1003, 72, generator=g)
    for t in (base, base[:901, :64], base[:, 0].contiguous(), base[:5, :3], torch.randn(4096 * 3, generator=g)):
        want = t.double().numpy()
        assert np.array_equal(readback.f64_from_device(t), want)                 # default: direct
        for chunk in (1 << 10, 4096 * 4, 1 << 16, 1 << 24):
            got = readback.f64_from_device(t, mode="pipelined", chunk_bytes=chunk, n_threads=3)
            assert got.dtype == np.float64 and got.shape == want.shape and np.array_equal(got, want), (t.shape, chunk)
    with pytest.raises(ValueError):
        readback.f64_from_device(base, mode="lazy")
    monkeypatch.setenv("B200ALS_READBACK", "pipelined")
    assert np.array_equal(readback.f64_from_device(base), base.double().numpy())


def test_similarity_graph_is_reused_for_identical_inputs():
    """graph.similarity_topk_cached: the graph depends only on (feature matrix, top-k, eps) (scripts/als.py:220-240), so
    the folds of a trial share one build; a different matrix, top-k or eps is a different entry; at most two are kept."""
    from b200als import graph as G
    G._S_CACHE.clear()
    rng = np.random.default_rng(2)
    X = (rng.random((300, 7)) < 0.3).astype(np.float32)
    S1, csr1, D1 = G.similarity_topk_cached(X, 20, 1e-8)
    S2, csr2, D2 = G.similarity_topk_cached(X.copy(), 20, 1e-8)           # same bytes, another object
    assert S2 is S1 and csr2 is csr1 and D2 is D1
    assert np.array_equal(S1, G.similarity_topk(X, 20, 1e-8))
    assert np.array_equal(D1, S1.sum(axis=1)) and D1.dtype == S1.dtype      # als.py:357: D in S's dtype
    sp, si, sv = G.to_csr(S1)
    assert np.array_equal(csr1[0], sp) and np.array_equal(csr1[1], si) and np.array_equal(csr1[2], sv)
    Y = X.copy(); Y[0, 0] = 1.0 - Y[0, 0]
    assert G.similarity_topk_cached(Y, 20, 1e-8)[0] is not S1
    assert G.similarity_topk_cached(X, 21, 1e-8)[0] is not S1
    assert len(G._S_CACHE) <= G._S_CACHE_MAX
    assert G.similarity_topk_cached(X.astype(np.float64), 20, 1e-8)[0].dtype == np.float64
    G._S_CACHE.clear()
