"""Device-resident CSR + CSC of the rating matrix.

Index-bit-exact with `scripts/als.py:335-340`: CSR column ids ascending per
user (`np.flatnonzero(mask[u])`), CSC user ids ascending per item
(`np.flatnonzero(mask[:, i])`), COO order == CSR order (`np.where(mask)`).
torch only owns the buffers; every array is produced by libb200als kernels.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np
import torch

from . import _lib
from ._lib import check, ptr


@dataclass
class Ratings:
    """CSR (users) and CSC (items) views of the same observed entries, in HBM."""
    m: int
    n: int
    nnz: int
    indptr: torch.Tensor      # int64 [m+1]
    indices: torch.Tensor     # int32 [nnz]   item ids, ascending per user
    vals: torch.Tensor        # float32 [nnz]
    cindptr: torch.Tensor     # int64 [n+1]
    cindices: torch.Tensor    # int32 [nnz]   user ids, ascending per item
    cvals: torch.Tensor       # float32 [nnz]
    perm: torch.Tensor | None = None   # int64 [nnz]  perm[p_csc] = p_csr

    def to_host(self):
        """numpy copies (tests / oracle comparison)."""
        out = {k: getattr(self, k).cpu().numpy() for k in
               ("indptr", "indices", "vals", "cindptr", "cindices", "cvals")}
        if self.perm is not None:
            out["perm"] = self.perm.cpu().numpy()
        return out


def _stream():
    return torch.cuda.current_stream().cuda_stream


def ratings_from_dense(R: np.ndarray, device, want_perm: bool = False) -> Ratings:
    """(m, n) float64 with NaN = missing → Ratings (als.py:335-340)."""
    lib = _lib.load()
    _lib.require_device()
    if R.ndim != 2:
        raise ValueError("R must be a 2-D matrix")
    m, n = R.shape
    Rd = torch.from_numpy(np.ascontiguousarray(R, dtype=np.float64)).to(device, non_blocking=False)
    indptr = torch.empty(m + 1, dtype=torch.int64, device=device)
    cindptr = torch.empty(n + 1, dtype=torch.int64, device=device)
    check(lib.b200als_dense_indptr(ptr(Rd), m, n, ptr(indptr), ptr(cindptr), _stream()), "dense_indptr")
    nnz = int(indptr[-1].item())
    z = max(nnz, 1)
    indices = torch.empty(z, dtype=torch.int32, device=device)
    vals = torch.empty(z, dtype=torch.float32, device=device)
    cindices = torch.empty(z, dtype=torch.int32, device=device)
    cvals = torch.empty(z, dtype=torch.float32, device=device)
    perm = torch.empty(z, dtype=torch.int64, device=device) if want_perm else None
    if nnz:
        check(lib.b200als_dense_fill(ptr(Rd), m, n, ptr(indptr), ptr(cindptr), ptr(indices), ptr(vals),
                                     ptr(cindices), ptr(cvals), ptr(perm), _stream()), "dense_fill")
    return Ratings(m, n, nnz, indptr, indices[:nnz], vals[:nnz], cindptr, cindices[:nnz], cvals[:nnz],
                   perm[:nnz] if perm is not None else None)


def ratings_from_flat(flat_idx, vals, shape, device, sorted_input: bool = False,
                      want_perm: bool = False) -> Ratings:
    """flat_idx = u*n + i (create_folds.py:73-84), unique → Ratings.

    `flat_idx` / `vals` may be numpy arrays (copied H2D here) or torch tensors
    already on `device` (int64 / float32).

    The arrays are checked on the device before anything is built from them: an index outside the matrix raises
    IndexError (as the reference's matrix_from_indices does), duplicates (which the reference's scatter would
    resolve last-wins), a `sorted_input` promise that does not hold, or a non-finite rating raise ValueError.
    """
    lib = _lib.load()
    _lib.require_device()
    m, n = int(shape[0]), int(shape[1])
    if isinstance(flat_idx, torch.Tensor):
        fd = flat_idx.to(device=device, dtype=torch.int64).contiguous()
    else:
        fd = torch.from_numpy(np.ascontiguousarray(flat_idx, dtype=np.int64)).to(device)
    if isinstance(vals, torch.Tensor):
        vd = vals.to(device=device, dtype=torch.float32).contiguous()
    else:
        vd = torch.from_numpy(np.ascontiguousarray(vals, dtype=np.float32)).to(device)
    nnz = fd.numel()
    if vd.numel() != nnz:
        raise ValueError("flat_idx and vals differ in length")
    need = C.c_size_t(0)
    check(lib.b200als_coo_workspace_bytes(nnz, m, n, C.byref(need)), "coo_workspace_bytes")
    ws = torch.empty(max(need.value, 1), dtype=torch.uint8, device=device)
    z = max(nnz, 1)
    indptr = torch.empty(m + 1, dtype=torch.int64, device=device)
    cindptr = torch.empty(n + 1, dtype=torch.int64, device=device)
    indices = torch.empty(z, dtype=torch.int32, device=device)
    ovals = torch.empty(z, dtype=torch.float32, device=device)
    cindices = torch.empty(z, dtype=torch.int32, device=device)
    cvals = torch.empty(z, dtype=torch.float32, device=device)
    perm = torch.empty(z, dtype=torch.int64, device=device) if want_perm else None
    status = torch.zeros(3, dtype=torch.int64, device=device)
    check(lib.b200als_coo_to_csr_csc(ptr(fd), ptr(vd), nnz, m, n, 1 if sorted_input else 0,
                                     ptr(indptr), ptr(indices), ptr(ovals), ptr(cindptr), ptr(cindices),
                                     ptr(cvals), ptr(perm), ptr(ws), ws.numel(), ptr(status), _stream()),
          "coo_to_csr_csc")
    st = status.cpu().numpy()                   # also the synchronisation after which `ws` may be released
    if st[0]:
        # matrix_from_indices (create_folds.py:169-174) raises IndexError on these
        raise IndexError(f"{int(st[0])} flat indices lie outside the ({m}, {n}) matrix")
    if st[1]:
        what = "is not strictly ascending (sorted_input=True was promised)" if sorted_input else "holds duplicates"
        raise ValueError(f"flat_idx {what}: {int(st[1])} adjacent pairs out of order")
    if st[2]:
        raise ValueError(f"{int(st[2])} ratings are not finite (a list of observed entries has no 'missing' marker)")
    return Ratings(m, n, nnz, indptr, indices[:nnz], ovals[:nnz], cindptr, cindices[:nnz], cvals[:nnz],
                   perm[:nnz] if perm is not None else None)


def upload_flat_sharded(flat_idx, vals, device, shard):
    """Host (flat_idx, vals) -> device tensors on EVERY rank, each rank copying only its 1/N slice over PCIe and the
    ranks all-gathering the slices over NVLink.  Contract of the sharded `fit_coo`: every rank is given the same
    arrays (checked: same length on all ranks).  At the Netflix shape the whole list is 1.2 GB per rank; with 8 ranks
    the host copy drops from ~48 ms to ~6 ms + a 2 ms all-gather."""
    import torch.distributed as td
    nnz = int(flat_idx.numel() if isinstance(flat_idx, torch.Tensor) else np.asarray(flat_idx).shape[0])
    w, r = shard.world, shard.rank
    lens = torch.tensor([nnz, -nnz], dtype=torch.int64, device=device)
    td.all_reduce(lens, op=td.ReduceOp.MAX, group=shard.group)
    if int(lens[0].item()) != nnz or int(-lens[1].item()) != nnz:
        raise ValueError("sharded fit_coo: the ranks were given rating lists of different lengths")
    per = max((nnz + w - 1) // w, 1)
    lo, hi = min(r * per, nnz), min((r + 1) * per, nnz)

    def part(x, dtype, np_dtype):
        buf = torch.zeros(per, dtype=dtype, device=device)
        if hi > lo:
            if isinstance(x, torch.Tensor):
                buf[:hi - lo].copy_(x[lo:hi].to(dtype), non_blocking=True)
            else:
                buf[:hi - lo].copy_(torch.from_numpy(np.ascontiguousarray(np.asarray(x)[lo:hi], dtype=np_dtype)))
        full = torch.empty(w * per, dtype=dtype, device=device)
        td.all_gather_into_tensor(full, buf, group=shard.group)
        return full[:nnz]
    return part(flat_idx, torch.int64, np.int64), part(vals, torch.float32, np.float32)


def row_schedule(indptr: torch.Tensor):
    """Longest-first row order and the number of non-empty rows."""
    lib = _lib.load()
    n_rows = indptr.numel() - 1
    device = indptr.device
    need = C.c_size_t(0)
    check(lib.b200als_row_order_workspace_bytes(n_rows, C.byref(need)), "row_order_workspace_bytes")
    ws = torch.empty(need.value, dtype=torch.uint8, device=device)
    order = torch.empty(n_rows, dtype=torch.int32, device=device)
    nne = torch.zeros(1, dtype=torch.int64, device=device)
    check(lib.b200als_row_order(ptr(indptr), n_rows, ptr(order), ptr(nne), ptr(ws), ws.numel(), _stream()),
          "row_order")
    return order, int(nne.item())
