"""float64 host copies of the learned float32 device arrays (scripts/als.py:404-407: U, V, b_u, b_i are float64).

`direct` (the default, the path every measured number of round 1 used): widen on the device, one pageable D2H.
`pipelined` (B200ALS_READBACK=pipelined; written at the end of round 1 after the GPU budget was spent, so NOT yet
run on a GPU and therefore not the default): float32 D2H in chunks through two small pinned buffers, each chunk
widened into the float64 result by libb200als_host.so on the host threads while the next chunk's copy is in
flight.  Half the bytes cross PCIe, the copies are pinned, and the first-touch page faults of the result are
spread over the cores.  The host part is covered on CPU tensors by tests/test_host_graph.py.
"""
from __future__ import annotations

import os

import numpy as np
import torch

from . import _hostlib

CHUNK_BYTES = 16 << 20
MIN_PIPELINED_ELEMS = 1 << 22


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
    mode = mode or os.environ.get("B200ALS_READBACK", "direct")
    if mode not in ("direct", "pipelined"):
        raise ValueError(f"B200ALS_READBACK={mode!r}: expected 'direct' or 'pipelined'")
    if mode == "direct" or t.dtype != torch.float32 or t.dim() not in (1, 2) or t.numel() < MIN_PIPELINED_ELEMS:
        return t.double().cpu().numpy()
    t2 = t.reshape(1, -1) if t.dim() == 1 else t
    if t2.dim() == 2 and t2.shape[0] == 1 and t2.shape[1] * 4 > chunk_bytes:      # a long vector: chunk along it
        width = chunk_bytes // 4
        pad = (-t2.shape[1]) % width
        if pad:
            return t.double().cpu().numpy()                                      # ragged: not worth a third path
        t2 = t2.reshape(-1, width)
    rows, cols = t2.shape
    nth = int(n_threads or _hostlib.host_threads())
    out = np.empty((rows, cols), dtype=np.float64)
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
