"""Fold evaluation without dense matrices.

The reference's drivers materialise a dense NaN-filled (m, n) train matrix per fold
(`make_train_valid_split`, create_folds.py:177-208), fit, build the dense (m, n) prediction and index it with the
fold's flat indices (`_rmse_on_indices`, tune_params.py:147-167; popularity bins, evaluate_models.py:131-191,
263-274).  Here the observed entries stay one sorted (flat_idx, vals) pair on the device: a fold is a membership
mask over it, the fit consumes the surviving pairs directly, and the validation error is one fused kernel.
Index conventions are the reference's: flat = u * n + i, int64, row-major (create_folds.py:73-84).
"""
from __future__ import annotations

import numpy as np
import torch

N_POP_BINS = 5                   # evaluate_models.py:107
POP_BIN_STRATEGY = "quantile"    # evaluate_models.py:108


def popularity_bins(counts, n_bins: int = N_POP_BINS, strategy: str = POP_BIN_STRATEGY):
    """Item popularity bins from per-item rating counts — `_popularity_bins_from_R` (evaluate_models.py:131-170)
    with `counts = (~isnan(R)).sum(0)` supplied by the caller (the CSC row lengths).  Same numpy calls, so the
    edges and the bin of every item are identical.  Returns (item_bin int[n], edges float[n_bins + 1])."""
    counts = np.asarray(counts).astype(float)
    if strategy == "quantile":
        edges = np.quantile(counts, np.linspace(0, 1, n_bins + 1))
    elif strategy == "uniform":
        edges = np.linspace(float(np.min(counts)), float(np.max(counts)), n_bins + 1)
    else:
        raise ValueError(f"Unknown popularity binning strategy '{strategy}'")
    edges = np.array(edges, dtype=float)
    for i in range(1, len(edges)):                       # strictly increasing edges (evaluate_models.py:161-164)
        if edges[i] <= edges[i - 1]:
            edges[i] = edges[i - 1] + 1e-9
    item_bin = np.clip(np.searchsorted(edges, counts, side="right") - 1, 0, n_bins - 1)
    return item_bin.astype(int), edges


def split_fold(flat_sorted: torch.Tensor, vals: torch.Tensor, val_idx):
    """`make_train_valid_split` (create_folds.py:177-208) on the resident observed set.

    flat_sorted: int64 device tensor, ascending flat indices of every observed entry (`flatnonzero(~isnan(R))`);
    vals: their ratings; val_idx: the fold's validation flat indices (any order, a subset of flat_sorted).
    Returns (train_flat, train_vals, val_flat, val_vals): the train pair keeps ascending order — the sorted
    `setdiff1d` of create_folds.py:200 — and the validation pair follows `val_idx`'s order.
    Raises ValueError if a validation index is not an observed entry (the reference would read NaN there).
    Duplicated validation indices are removed from the train set once, like `setdiff1d`."""
    from . import _lib
    from ._lib import check, ptr
    import ctypes as C
    lib = _lib.load()
    _lib.require_device()
    dev = flat_sorted.device
    nnz = int(flat_sorted.numel())
    if nnz == 0:
        raise ValueError("no observed ratings")
    flat_sorted = flat_sorted.to(torch.int64).contiguous()
    vals = vals.to(torch.float32).contiguous()
    v = torch.as_tensor(val_idx).to(device=dev, dtype=torch.int64).contiguous()
    with torch.cuda.device(dev):
        need = C.c_size_t(0)
        check(lib.b200als_fold_split_workspace_bytes(nnz, C.byref(need)), "fold_split_workspace_bytes")
        ws = torch.empty(need.value, dtype=torch.uint8, device=dev)
        train_flat = torch.empty(nnz, dtype=torch.int64, device=dev)
        train_vals = torch.empty(nnz, dtype=torch.float32, device=dev)
        val_vals = torch.empty(max(v.numel(), 1), dtype=torch.float32, device=dev)
        n_train = torch.zeros(1, dtype=torch.int64, device=dev)
        status = torch.zeros(2, dtype=torch.int32, device=dev)
        check(lib.b200als_fold_split(ptr(flat_sorted), ptr(vals), nnz, ptr(v), v.numel(), ptr(train_flat),
                                     ptr(train_vals), ptr(val_vals), ptr(n_train), ptr(status), ptr(ws), ws.numel(),
                                     torch.cuda.current_stream().cuda_stream), "fold_split")
        st = status.cpu().numpy()
        if st[0]:
            raise ValueError(f"{int(st[0])} validation indices are not observed entries")
        nt = int(n_train.item())
    return train_flat[:nt], train_vals[:nt], v, val_vals[:v.numel()]


def cv_score(make_model, flat_sorted: torch.Tensor, vals: torch.Tensor, shape, folds, features=None,
             tol=1e-4, min_iters=10, verbose=0, item_bin=None, n_bins: int = N_POP_BINS):
    """K-fold loop of `_cv_score_single_trial` / `_eval_variant_cv` (tune_params.py:341-421,
    evaluate_models.py:194-276) on the device: per fold split -> `fit_coo` -> fused validation RMSE
    (+ popularity-bin RMSEs when `item_bin` is given).  `make_model()` returns a fresh ALS per fold.
    Returns a list of dicts {rmse, iters, train_rmse, bin_rmse}."""
    out = []
    for val_idx in folds:
        tf, tv, vf, vv = split_fold(flat_sorted, vals, val_idx)
        model = make_model()
        model.fit_coo(tf, tv, shape, features=features, tol=tol, min_iters=min_iters, verbose=verbose,
                      sorted_input=True)
        if item_bin is not None:
            rmse, per_bin = model.rmse_by_item_bin(vf, vv, item_bin, n_bins, features)
        else:
            rmse, per_bin = model.rmse_on_flat(vf, vv, features), None
        hist = list(model.history.get("train_rmse", []))
        out.append({"rmse": rmse, "iters": len(hist), "train_rmse": hist, "bin_rmse": per_bin})
    return out


def folds_of_rank(n_folds: int, rank: int, world: int):
    """Fold indices rank `rank` of `world` evaluates: rank, rank + world, ... (every fold exactly once)."""
    return list(range(rank, n_folds, world))


def cv_score_fold_parallel(make_model, flat_sorted, vals, shape, folds, features=None, tol=1e-4, min_iters=10,
                           verbose=0, item_bin=None, n_bins: int = N_POP_BINS, group=None):
    """The same K-fold loop with the folds spread over the GPUs of the box — one fold per GPU at a time, each fit
    un-sharded on its own device (second half: the reference's real throughput lever is
    trials x folds, tune_params.py:341-421, and a bundled-shape fit cannot fill one B200, let alone eight).

    Call it from every rank of an initialised torch.distributed group (torchrun, one process per GPU) with the
    same arguments; `flat_sorted` / `vals` may be host arrays (each rank uploads its copy) or tensors on the rank's
    device.  Fold f is evaluated by rank f % world; the per-fold dicts are exchanged with one all_gather_object and
    every rank returns the full list in fold order — identical to `cv_score`'s on one GPU, fold by fold, because
    each fold's fit is the same single-GPU fit.  Without a process group it is `cv_score`."""
    import torch.distributed as td
    if not (td.is_available() and td.is_initialized()):
        return cv_score(make_model, torch.as_tensor(flat_sorted).cuda(), torch.as_tensor(vals).cuda(), shape, folds,
                        features, tol, min_iters, verbose, item_bin, n_bins)
    rank, world = td.get_rank(group), td.get_world_size(group)
    mine = folds_of_rank(len(folds), rank, world)
    res = []
    if mine:
        dev = torch.device("cuda", torch.cuda.current_device())
        f_d = torch.as_tensor(flat_sorted).to(device=dev, dtype=torch.int64)
        v_d = torch.as_tensor(vals).to(device=dev, dtype=torch.float32)
        res = cv_score(make_model, f_d, v_d, shape, [folds[f] for f in mine], features, tol, min_iters, verbose,
                       item_bin, n_bins)
    return merge_fold_results(list(zip(mine, res)), len(folds), group)


def merge_fold_results(mine, n_folds: int, group=None):
    """all_gather_object of [(fold index, result)] lists -> results in fold order on every rank."""
    import torch.distributed as td
    box = [None] * td.get_world_size(group)
    td.all_gather_object(box, mine, group=group)
    out = [None] * n_folds
    for part in box:
        for f, r in part:
            if out[f] is not None:
                raise RuntimeError(f"fold {f} was evaluated twice")
            out[f] = r
    if any(r is None for r in out):
        raise RuntimeError("a fold was not evaluated by any rank")
    return out
