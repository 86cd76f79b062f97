"""Slab-only storage of the ratings for the sharded fit ("each GPU holds only its CSR row slab and CSC
column slab plus full replicas of U, V, b_u, b_i").

`sparse.Ratings` keeps the whole CSR and the whole CSC on every rank — fine up to the Netflix shape (1.6 GB), not at
BASELINE.json's C5 (10 M x 1 M, 2e9 ratings: 32 GB of CSR + CSC, and more than 2^31 positions).  A `RatingSlabs`
holds, on rank r of w, the CSR of the users u with u % w == r and the CSC of the items i with i % w == r, both with
LOCAL row pointers (every slab stays far below 2^31 ratings) and GLOBAL ids inside (`indices` are item ids,
`cindices` user ids, as in the reference's `user_items` / `item_users` lists, scripts/als.py:338-339).  Local row l
of the user slab is user l * w + r (`row_ids_u`), likewise for items; the half-step kernels take that map
(b200als_step_args.row_ids) and write the solved row where the replicas keep it.

Ownership by residue is what makes the slabs constructible shard-wise: whoever generates or reads a rating knows
its two owners from the pair alone.  With activity / popularity randomly permuted over the ids (synth.py) the
residue classes are statistically balanced; the split plans (include/b200als.h) take care of the few very long rows.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import torch


@dataclass
class RatingSlabs:
    m: int                     # users, items and observed entries of the WHOLE matrix
    n: int
    nnz: int
    rank: int
    world: int
    indptr: torch.Tensor       # int64 [users of this rank + 1], local
    indices: torch.Tensor      # int32, global item ids, ascending per user
    vals: torch.Tensor         # float32
    row_ids_u: torch.Tensor    # int32 [users of this rank]: global id of local row l (= l * world + rank)
    cindptr: torch.Tensor      # int64 [items of this rank + 1], local
    cindices: torch.Tensor     # int32, global user ids, ascending per item
    cvals: torch.Tensor
    row_ids_i: torch.Tensor
    vals_sum: float            # sum of ALL ratings (for mu0 = nanmean(R), scripts/als.py:360)

    @property
    def nnz_users(self) -> int:
        return int(self.indices.numel())

    @property
    def nnz_items(self) -> int:
        return int(self.cindices.numel())


def owned_count(total: int, rank: int, world: int) -> int:
    """How many ids in [0, total) have id % world == rank."""
    return (total - rank + world - 1) // world if total > rank else 0


def slabs_from_coo(rows: torch.Tensor, cols: torch.Tensor, vals: torch.Tensor, m: int, n: int, rank: int, world: int,
                   nnz_total: int | None = None, vals_sum: float | None = None) -> RatingSlabs:
    """Slabs of rank `rank` from (a superset of) its ratings as COO on the device: keeps the pairs whose user it owns
    for the CSR slab and the pairs whose item it owns for the CSC slab.  Pairs must be unique.  torch only sorts
    and scans here (one-shot set-up, like sparse_build.cu's radix sorts); the arrays feed the same kernels."""
    dev = rows.device
    rows, cols = rows.to(torch.int64), cols.to(torch.int64)
    mu, ni = owned_count(m, rank, world), owned_count(n, rank, world)
    nnz_total = int(rows.numel()) if nnz_total is None else int(nnz_total)
    vals_sum = float(vals.sum(dtype=torch.float64).item()) if vals_sum is None else float(vals_sum)

    def one_side(own, other, n_own, n_other):
        keep = (own % world) == rank
        lo, ot, vv = own[keep] // world, other[keep], vals[keep]
        order = torch.argsort(lo * n_other + ot)                       # (local row, global other id) ascending
        lo, ot, vv = lo[order], ot[order], vv[order]
        ptr = torch.zeros(n_own + 1, dtype=torch.int64, device=dev)
        if lo.numel():
            ptr[1:] = torch.cumsum(torch.bincount(lo, minlength=n_own), 0)
        ids = (torch.arange(n_own, device=dev, dtype=torch.int64) * world + rank).to(torch.int32)
        return ptr, ot.to(torch.int32).contiguous(), vv.to(torch.float32).contiguous(), ids

    indptr, indices, v_u, ids_u = one_side(rows, cols, mu, n)
    cindptr, cindices, v_i, ids_i = one_side(cols, rows, ni, m)
    return RatingSlabs(m, n, nnz_total, rank, world, indptr, indices, v_u, ids_u, cindptr, cindices, v_i, ids_i, vals_sum)


def synth_slabs_torch(m: int, n: int, nnz: int, seed: int, rank: int, world: int, device, n_chunks: int = 32,
                      oversample: float = 1.04) -> RatingSlabs:
    """The power-law recipe of synth.py generated SHARD-WISE: the users are cut into `n_chunks` contiguous ranges, every
    rank walks all chunks with the same counter-based generator state (seed, chunk), draws the chunk's pairs
    (user by the activity law restricted to the chunk, item by the popularity law), de-duplicates inside the chunk
    (pairs of different chunks cannot collide) and keeps only what it owns — so no rank ever holds the 2e9 pairs,
    and all ranks agree on the data without exchanging any.  The total after de-duplication is close to, not
    exactly, `nnz` (the exact count is `RatingSlabs.nnz`); ratings come from the same planted rank-16 model."""
    from .synth import PLANT_RANK
    dev = torch.device(device)
    g0 = torch.Generator(device=dev)
    g0.manual_seed(seed)

    def powerlaw(size, expo):
        p = torch.arange(1, size + 1, dtype=torch.float64, device=dev) ** (-expo)
        p = p[torch.randperm(size, generator=g0, device=dev)]
        return p / p.sum()

    pu, pi = powerlaw(m, 0.7), powerlaw(n, 0.9)
    ci = torch.cumsum(pi, 0)
    ci[-1] = 1.0
    del pi
    P = torch.randn(m, PLANT_RANK, generator=g0, device=dev) * 0.35
    Q = torch.randn(n, PLANT_RANK, generator=g0, device=dev) * 0.35
    a = torch.randn(m, generator=g0, device=dev) * 0.4
    c = torch.randn(n, generator=g0, device=dev) * 0.5

    mu_own, ni_own = owned_count(m, rank, world), owned_count(n, rank, world)
    bounds = np.linspace(0, m, n_chunks + 1).astype(np.int64)
    u_parts, i_parts = [], []
    total, vsum = 0, 0.0
    for ch in range(n_chunks):
        u0, u1 = int(bounds[ch]), int(bounds[ch + 1])
        if u1 <= u0:
            continue
        g = torch.Generator(device=dev)
        g.manual_seed(seed * 1_000_003 + ch + 1)
        pc = pu[u0:u1]
        mass = float(pc.sum().item())
        draws = int(nnz * mass * oversample) + 16
        cu = torch.cumsum(pc / pc.sum(), 0)
        cu[-1] = 1.0
        u = torch.searchsorted(cu, torch.rand(draws, generator=g, device=dev, dtype=torch.float64), right=True)
        u = u.clamp_(max=u1 - u0 - 1) + u0
        i = torch.searchsorted(ci, torch.rand(draws, generator=g, device=dev, dtype=torch.float64), right=True)
        i = i.clamp_(max=n - 1)
        flat = torch.unique(u * n + i)                               # sorted, de-duplicated
        del u, i, cu
        keep_target = int(round(nnz * mass))
        if flat.numel() > keep_target:                               # drop a random surplus, keep the order
            drop = torch.randperm(flat.numel(), generator=g, device=dev)[: flat.numel() - keep_target]
            mask = torch.ones(flat.numel(), dtype=torch.bool, device=dev)
            mask[drop] = False
            flat = flat[mask]
            del mask, drop
        ru = flat // n
        ri = flat - ru * n
        del flat
        raw = 3.5 + a[ru] + c[ri] + (P[ru] * Q[ri]).sum(1) + torch.randn(ru.numel(), generator=g, device=dev) * 0.6
        vv = torch.clamp(torch.round(raw * 2.0) / 2.0, 0.5, 5.0).to(torch.float32)
        del raw
        total += int(ru.numel())
        vsum += float(vv.sum(dtype=torch.float64).item())
        ku = (ru % world) == rank
        u_parts.append((ru[ku], ri[ku].to(torch.int32), vv[ku]))     # already (user, item) ascending
        ki = (ri % world) == rank
        i_parts.append((ri[ki], ru[ki].to(torch.int32), vv[ki]))
        del ru, ri, vv, ku, ki
    del P, Q, a, c, pu, ci

    # user slab: chunks are ascending user ranges, each sorted — concatenation is the CSR order
    ru = torch.cat([p[0] for p in u_parts])
    indices = torch.cat([p[1] for p in u_parts])
    v_u = torch.cat([p[2] for p in u_parts])
    del u_parts
    indptr = torch.zeros(mu_own + 1, dtype=torch.int64, device=dev)
    indptr[1:] = torch.cumsum(torch.bincount(ru // world, minlength=mu_own), 0)
    del ru
    # item slab: stable sort by local item id keeps ascending users inside an item (chunks ascend in users)
    ri = torch.cat([p[0] for p in i_parts])
    cu_ids = torch.cat([p[1] for p in i_parts])
    v_i = torch.cat([p[2] for p in i_parts])
    del i_parts
    li = (ri // world).to(torch.int32)
    del ri
    li_sorted, order = torch.sort(li, stable=True)
    del li
    cindices, cvals = cu_ids[order].contiguous(), v_i[order].contiguous()
    del cu_ids, v_i, order
    cindptr = torch.zeros(ni_own + 1, dtype=torch.int64, device=dev)
    cindptr[1:] = torch.cumsum(torch.bincount(li_sorted, minlength=ni_own), 0)
    del li_sorted
    ids_u = (torch.arange(mu_own, device=dev, dtype=torch.int64) * world + rank).to(torch.int32)
    ids_i = (torch.arange(ni_own, device=dev, dtype=torch.int64) * world + rank).to(torch.int32)
    return RatingSlabs(m, n, total, rank, world, indptr, indices.contiguous(), v_u.contiguous(), ids_u,
                       cindptr, cindices, cvals, ids_i, vsum)
