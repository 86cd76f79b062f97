"""float64 host copies of the learned float32 device arrays (scripts/als.py:404-407: U, V, b_u, b_i are float64).

Called when a public attribute of a fitted model is first read (ALS._get_param).  Two ways to the same array:

`direct`    widen on the device, one pageable D2H of the float64 array.
`pipelined` (default for arrays of >= 4 Mi elements) float32 D2H in chunks through two small pinned buffers, each
            chunk widened into the float64 result by libb200als_host.so on the host threads while the next chunk's
            copy is in flight.  Half the bytes cross PCIe and the copies are pinned.

Either way the result array is advised MADV_HUGEPAGE before its first touch: at the Netflix shape the 492 MB result
is 120 k first-touch page faults with 4 KB pages (they, not PCIe, were the 0.23 s of round 1's read-back) and 240
with 2 MB pages.  B200ALS_READBACK=direct|pipelined overrides the choice; B200ALS_HUGEPAGE=0 skips the advice.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np
import torch

from . import _hostlib

CHUNK_BYTES = 16 << 20
MIN_PIPELINED_ELEMS = 1 << 22
_HUGE = 2 << 20
_MADV_HUGEPAGE = 14
_libc = None


def empty_f64(shape) -> np.ndarray:
    """np.empty(shape, float64) whose interior is advised to be backed by 2 MB pages (Linux THP in `madvise` mode)."""
    global _libc
    out = np.empty(shape, dtype=np.float64)
    if out.nbytes < 2 * _HUGE or os.environ.get("B200ALS_HUGEPAGE", "1") == "0":
        return out
    try:
        if _libc is None:
            _libc = C.CDLL(None, use_errno=True)
            _libc.madvise.argtypes = [C.c_void_p, C.c_size_t, C.c_int]
            _libc.madvise.restype = C.c_int
        a = out.ctypes.data
        lo = (a + _HUGE - 1) & ~(_HUGE - 1)
        hi = (a + out.nbytes) & ~(_HUGE - 1)
        if hi > lo:
            _libc.madvise(C.c_void_p(lo), C.c_size_t(hi - lo), _MADV_HUGEPAGE)      # advisory: failure changes nothing
    except Exception:   # noqa: BLE001 — no libc / not Linux: plain pages
        pass
    return out


def _widen_into(src: np.ndarray, dst: np.ndarray, nth: int) -> None:
    lib = _hostlib.load()
    if lib is None or src.strides[1] != 4 or src.strides[0] % 4:
        dst[...] = src
        return
    if lib.b200als_host_widen_f32(src.ctypes.data, src.strides[0] // 4, dst.ctypes.data, src.shape[0], src.shape[1],
                                  nth) != 0:
        raise ValueError("b200als_host_widen_f32: bad arguments")


def f64_from_device(t: torch.Tensor, mode: str | None = None, chunk_bytes: int = CHUNK_BYTES,
                    n_threads: int | None = None) -> np.ndarray:
    """`t.double().cpu().numpy()` for a float32 vector or matrix (row-strided views allowed)."""
    mode = mode or os.environ.get("B200ALS_READBACK", "auto")
    if mode not in ("auto", "direct", "pipelined"):
        raise ValueError(f"B200ALS_READBACK={mode!r}: expected 'auto', 'direct' or 'pipelined'")
    small = t.numel() < MIN_PIPELINED_ELEMS
    if mode == "direct" or t.dtype != torch.float32 or t.dim() not in (1, 2) or (small and mode == "auto") or t.numel() == 0:
        out = empty_f64(tuple(t.shape))
        if t.numel():
            torch.from_numpy(out).copy_(t.double() if t.is_cuda else t)      # device-side widening, one D2H into `out`
        return out
    t2 = t.reshape(1, -1) if t.dim() == 1 else t
    if t2.shape[0] == 1 and t2.shape[1] * 4 > chunk_bytes:      # a long vector: chunk along it
        width = chunk_bytes // 4
        if t2.shape[1] % width:
            out = empty_f64(tuple(t.shape))
            torch.from_numpy(out).copy_(t.double() if t.is_cuda else t)      # ragged: not worth a third path
            return out
        t2 = t2.reshape(-1, width)
    rows, cols = t2.shape
    nth = int(n_threads or _hostlib.host_threads())
    out = empty_f64((rows, cols))
    chunk_rows = max(1, chunk_bytes // (4 * cols))
    on_gpu = t2.is_cuda
    bufs = [torch.empty((chunk_rows, cols), dtype=torch.float32, pin_memory=on_gpu) for _ in range(2)]
    events = [torch.cuda.Event() if on_gpu else None for _ in range(2)]
    starts = list(range(0, rows, chunk_rows))

    def issue(i):
        lo = starts[i]
        hi = min(lo + chunk_rows, rows)
        bufs[i % 2][:hi - lo].copy_(t2[lo:hi], non_blocking=on_gpu)
        if on_gpu:
            events[i % 2].record()

    issue(0)
    for i, lo in enumerate(starts):
        hi = min(lo + chunk_rows, rows)
        if i + 1 < len(starts):
            issue(i + 1)            # buffer (i + 1) % 2 was consumed by the widening of chunk i - 1, already done
        if on_gpu:
            events[i % 2].synchronize()
        _widen_into(bufs[i % 2][:hi - lo].numpy(), out[lo:hi], nth)
    return out.reshape(t.shape)
