"""The reference's factor-initialisation draw on all host cores, bit-identical to numpy.

The reference (scripts/als.py:329, 367-376) seeds ONE `np.random.default_rng(random_state)` and draws
`rng.normal(scale=0.1, size=(m, k))` for U, then V, then every W_f.  Seed parity needs exactly those values;
numpy walks the stream on one core (half of the end-to-end fit time at the Netflix shape).
`normal_stream` returns the same arrays from `libb200als_host.so` (include/b200als_host.h,
csrc/host_rng.c: numpy's own ziggurat code over a random-access PCG64, chunks decoded speculatively in
parallel and stitched).  If that library has not been built the draw falls back to numpy itself — the same
values, one core — since this is host-side plumbing, not the CUDA path.
"""
from __future__ import annotations

import numpy as np

from ._hostlib import HOST_LIB_PATH, host_threads, load as _load   # noqa: F401  (re-exported)

_M64 = (1 << 64) - 1


def normal_stream(seed, shapes, dtypes, scale: float, n_threads: int | None = None, force_numpy: bool = False,
                  alloc=None):
    """[rng.normal(scale=scale, size=s).astype(d) for s, d in zip(shapes, dtypes)] with rng = default_rng(seed),
    dtypes float32 / float64.  `alloc(shape, dtype)` (optional) supplies the output arrays — e.g. views of pinned
    host buffers, so that the upload that follows is a DMA from where the values were written."""
    dtypes = [np.dtype(d) for d in dtypes]
    assert all(d in (np.dtype(np.float32), np.dtype(np.float64)) for d in dtypes)
    lib = None if force_numpy else _load()
    rng = np.random.default_rng(seed)
    st = rng.bit_generator.state
    # a caller-owned Generator / BitGenerator must be left advanced exactly as numpy would leave it, and only PCG64
    # has the walk below: both go through numpy itself
    caller_owned = isinstance(seed, (np.random.Generator, np.random.BitGenerator))
    if lib is None or caller_owned or st["bit_generator"] != "PCG64" or st["has_uint32"]:
        res = [rng.normal(scale=scale, size=s).astype(d, copy=False) for s, d in zip(shapes, dtypes)]
        if alloc is not None:
            outs = [alloc(s, d) for s, d in zip(shapes, dtypes)]
            for o, r in zip(outs, res):
                o[...] = r
            return outs
        return res
    state, inc = int(st["state"]["state"]), int(st["state"]["inc"])
    outs = [(alloc(s, d) if alloc is not None else np.empty(s, dtype=d)) for s, d in zip(shapes, dtypes)]
    assert all(o.flags.c_contiguous and o.dtype == d and o.shape == tuple(s) for o, s, d in zip(outs, shapes, dtypes))
    counts = np.array([o.size for o in outs], dtype=np.int64)
    ptrs = np.array([o.ctypes.data for o in outs], dtype=np.uint64)
    as_f32 = np.array([d == np.dtype(np.float32) for d in dtypes], dtype=np.int32)
    rc = lib.b200als_host_normal_fill(state >> 64, state & _M64, inc >> 64, inc & _M64, len(outs), counts.ctypes.data,
                                      ptrs.ctypes.data, as_f32.ctypes.data, float(scale),
                                      int(n_threads or host_threads()), None)
    if rc != 0:
        raise MemoryError(f"b200als_host_normal_fill failed (code {rc})")
    return outs
