"""
Synthetic power-law rating generators.

The reference's `data/ratings.npy` is not shipped, so every named shape is
generated: item popularity p_i ∝ rank^-0.9, user activity p_u ∝ rank^-0.7
(both randomly permuted), (u, i) pairs de-duplicated on the flat index
`u*n + i` (the convention of create_folds.py:73-84) and topped up to exactly
`nnz`; ratings come from a planted rank-16 model on the half-star grid
{0.5, ..., 5.0}, so RMSE is meaningful.

Two generators with the same recipe:
  * `synth_ratings`       numpy, host, deterministic in (m, n, nnz, seed) —
                          tests, golden fixtures, small shapes.
  * `synth_ratings_torch` torch, any device — the 20 M / 100 M / 2 B shapes,
                          generated on the GPU in seconds.
Data generation is not part of the ALS path; torch ops are fine here.
"""
from __future__ import annotations

import numpy as np

PLANT_RANK = 16


def _powerlaw(rng, size, expo):
    p = np.arange(1, size + 1, dtype=np.float64) ** (-expo)
    p /= p.sum()
    return p[rng.permutation(size)]


def synth_ratings(m, n, nnz, seed=1234):
    """Return (flat_idx int64 sorted ascending, vals float64 on half-star grid)."""
    if nnz > m * n:
        raise ValueError("nnz exceeds m*n")
    rng = np.random.default_rng(seed)
    pu = _powerlaw(rng, m, 0.7)
    pi = _powerlaw(rng, n, 0.9)
    cu, ci = np.cumsum(pu), np.cumsum(pi)
    cu[-1] = ci[-1] = 1.0
    flat = np.empty(0, dtype=np.int64)
    while flat.size < nnz:
        need = nnz - flat.size
        draw = int(need * 1.3) + 16
        u = np.searchsorted(cu, rng.random(draw), side="right")
        i = np.searchsorted(ci, rng.random(draw), side="right")
        cand = np.unique(np.concatenate([flat, u.astype(np.int64) * n + i]))
        if cand.size > nnz:
            # keep all previously accepted, drop a random subset of the new ones
            new = np.setdiff1d(cand, flat, assume_unique=True)
            new = new[rng.permutation(new.size)[:need]]
            cand = np.union1d(flat, new)
        flat = cand
    flat.sort()
    u = flat // n
    i = flat - u * n
    # planted model
    P = rng.normal(scale=0.35, size=(m, PLANT_RANK))
    Q = rng.normal(scale=0.35, size=(n, PLANT_RANK))
    a = rng.normal(scale=0.4, size=m)
    c = rng.normal(scale=0.5, size=n)
    raw = 3.5 + a[u] + c[i] + np.einsum("ij,ij->i", P[u], Q[i]) + rng.normal(scale=0.6, size=flat.size)
    vals = np.clip(np.round(raw * 2.0) / 2.0, 0.5, 5.0)
    return flat, vals


def synth_dense(m, n, nnz, seed=1234):
    """Dense (m, n) float64 matrix with NaN for missing — the reference's input form."""
    flat, vals = synth_ratings(m, n, nnz, seed)
    R = np.full(m * n, np.nan)
    R[flat] = vals
    return R.reshape(m, n)


def synth_ratings_torch(m, n, nnz, seed=1234, device="cuda"):
    """Same recipe with torch ops on `device`.

    Returns (rows int64, cols int32-range int64, vals float32), sorted by flat
    index u*n+i (row-major), exactly nnz unique pairs.
    """
    import torch

    g = torch.Generator(device=device)
    g.manual_seed(seed)

    def powerlaw_cdf(size, expo):
        p = torch.arange(1, size + 1, dtype=torch.float64, device=device) ** (-expo)
        p = p[torch.randperm(size, generator=g, device=device)]
        c = torch.cumsum(p / p.sum(), 0)
        c[-1] = 1.0
        return c

    cu = powerlaw_cdf(m, 0.7)
    ci = powerlaw_cdf(n, 0.9)
    flat = torch.empty(0, dtype=torch.int64, device=device)
    chunk = 1 << 26
    while flat.numel() < nnz:
        need = nnz - flat.numel()
        draw = min(int(need * 1.3) + 1024, chunk * 4)
        r1 = torch.rand(draw, generator=g, device=device, dtype=torch.float64)
        u = torch.searchsorted(cu, r1, right=True).clamp_(max=m - 1)
        del r1
        r2 = torch.rand(draw, generator=g, device=device, dtype=torch.float64)
        i = torch.searchsorted(ci, r2, right=True).clamp_(max=n - 1)
        del r2
        new = u * n + i
        del u, i
        cand = torch.unique(torch.cat([flat, new]))
        del new
        if cand.numel() > nnz:
            # drop a random subset of the *new* entries only
            is_old = torch.isin(cand, flat, assume_unique=True)
            new_pos = torch.nonzero(~is_old).squeeze(1)
            drop = new_pos[torch.randperm(new_pos.numel(), generator=g, device=device)[: cand.numel() - nnz]]
            keep = torch.ones(cand.numel(), dtype=torch.bool, device=device)
            keep[drop] = False
            cand = cand[keep]
        flat = cand
    flat, _ = torch.sort(flat)
    rows = flat // n
    cols = flat - rows * n
    P = torch.randn(m, PLANT_RANK, generator=g, device=device) * 0.35
    Q = torch.randn(n, PLANT_RANK, generator=g, device=device) * 0.35
    a = torch.randn(m, generator=g, device=device) * 0.4
    c = torch.randn(n, generator=g, device=device) * 0.5
    vals = torch.empty(nnz, dtype=torch.float32, device=device)
    for s in range(0, nnz, chunk):
        e = min(nnz, s + chunk)
        ru, ri = rows[s:e], cols[s:e]
        raw = 3.5 + a[ru] + c[ri] + (P[ru] * Q[ri]).sum(1) \
            + torch.randn(e - s, generator=g, device=device) * 0.6
        vals[s:e] = torch.clamp(torch.round(raw * 2.0) / 2.0, 0.5, 5.0)
    return rows, cols, vals
