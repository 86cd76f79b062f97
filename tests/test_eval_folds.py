"""Fold split and fold metrics without dense matrices (N1, N2) against outputs of the
reference's own create_folds.py / evaluate_models.py / tune_params.py (tests/golden/eval_folds.npz, written by
tests/golden/make_golden_eval.py)."""
import os

import numpy as np
import pytest

from _util import GOLDEN

RMSE_TOL = 1e-4


def golden():
    return np.load(os.path.join(GOLDEN, "eval_folds.npz"))


@pytest.mark.parametrize("strategy", ["quantile", "uniform"])
def test_popularity_bins_match_reference(strategy):
    """`_popularity_bins_from_R` (evaluate_models.py:131-170) from CSC counts: bins and edges identical."""
    from b200als.evaluate import popularity_bins
    z = golden()
    n = int(z["n"])
    counts = np.bincount(z["flat_idx"] % n, minlength=n)
    item_bin, edges = popularity_bins(counts, int(z["n_bins"]), strategy)
    assert np.array_equal(item_bin, z[f"bin_{strategy}"])
    assert np.array_equal(edges, z[f"edges_{strategy}"])
    with pytest.raises(ValueError):
        popularity_bins(counts, 5, "nope")


def _model_factory(z):
    import b200als
    k, iters, lu, lv, seed = z["cfg_core"]
    def make():
        return b200als.ALS(b200als.ALSConfig(
            core=b200als.CoreConfig(n_factors=int(k), n_iters=int(iters), lambda_u=float(lu), lambda_v=float(lv),
                                    random_state=int(seed)),
            biases=b200als.BiasesConfig(lambda_bu=1.5, lambda_bi=1.5)))
    return make


@pytest.mark.gpu
def test_fold_split_bit_exact(cuda_device):
    """`make_train_valid_split` (create_folds.py:177-208): the train set is the reference's, index for index."""
    import torch
    from b200als.evaluate import split_fold
    z = golden()
    flat = torch.from_numpy(z["flat_idx"]).to(cuda_device)
    vals = torch.from_numpy(z["vals"].astype(np.float32) / 2.0).to(cuda_device)
    lookup = dict(zip(z["flat_idx"].tolist(), (z["vals"].astype(np.float32) / 2.0).tolist()))
    for k in range(4):
        vi = z[f"val_idx_{k}"]
        tf, tv, vf, vv = split_fold(flat, vals, vi)
        assert np.array_equal(tf.cpu().numpy(), z[f"train_flat_{k}"])
        assert np.array_equal(tv.cpu().numpy(), np.asarray([lookup[f] for f in z[f"train_flat_{k}"]], np.float32))
        assert np.array_equal(vf.cpu().numpy(), vi)
        assert np.array_equal(vv.cpu().numpy(), np.asarray([lookup[f] for f in vi], np.float32))
    # edge cases: empty validation set, everything held out, an index that is not an observed entry
    tf, tv, vf, vv = split_fold(flat, vals, np.empty(0, np.int64))
    assert np.array_equal(tf.cpu().numpy(), z["flat_idx"]) and vv.numel() == 0
    tf, tv, vf, vv = split_fold(flat, vals, z["flat_idx"][::-1].copy())
    assert tf.numel() == 0 and np.array_equal(vv.cpu().numpy(), (z["vals"].astype(np.float32) / 2.0)[::-1])
    missing = np.setdiff1d(np.arange(int(z["m"]) * int(z["n"])), z["flat_idx"])[:1]
    with pytest.raises(ValueError):
        split_fold(flat, vals, missing)


@pytest.mark.gpu
def test_fold_split_large_ragged(cuda_device):
    """Sizes that are not multiples of the 32-entry words / 8192-entry blocks of the compaction."""
    import torch
    from b200als.evaluate import split_fold
    rng = np.random.default_rng(5)
    for nnz, nval in [(1, 1), (31, 7), (33, 33), (8191, 100), (8193, 8000), (250_007, 50_001)]:
        flat = np.sort(rng.choice(10 * nnz + 50, size=nnz, replace=False)).astype(np.int64)
        vals = (rng.integers(1, 11, size=nnz) / 2.0).astype(np.float32)
        vi = rng.permutation(flat)[:nval]
        tf, tv, vf, vv = split_fold(torch.from_numpy(flat).to(cuda_device), torch.from_numpy(vals).to(cuda_device), vi)
        keep = ~np.isin(flat, vi)
        assert np.array_equal(tf.cpu().numpy(), np.setdiff1d(flat, vi))          # create_folds.py:200
        assert np.array_equal(tv.cpu().numpy(), vals[keep])
        assert np.array_equal(vv.cpu().numpy(), vals[np.searchsorted(flat, vi)])


@pytest.mark.gpu
def test_cv_score_matches_reference(cuda_device):
    """Per-fold validation RMSE, popularity-bin RMSEs and the train-RMSE curve of the reference's fold loop
    (tune_params.py:376-392, evaluate_models.py:246-274), without dense matrices."""
    import torch
    from b200als.evaluate import cv_score
    z = golden()
    m, n = int(z["m"]), int(z["n"])
    flat = torch.from_numpy(z["flat_idx"]).to(cuda_device)
    vals = torch.from_numpy(z["vals"].astype(np.float32) / 2.0).to(cuda_device)
    folds = [z[f"val_idx_{k}"] for k in range(4)]
    res = cv_score(_model_factory(z), flat, vals, (m, n), folds, tol=None, min_iters=0,
                   item_bin=z["bin_quantile"], n_bins=int(z["n_bins"]))
    for k, r in enumerate(res):
        assert abs(r["rmse"] - float(z[f"rmse_{k}"])) < RMSE_TOL
        want = z[f"bin_rmse_{k}"]
        got = np.asarray(r["bin_rmse"])
        assert np.array_equal(np.isnan(got), np.isnan(want))                     # empty bins are NaN in both
        assert np.nanmax(np.abs(got - want)) < RMSE_TOL
        assert np.abs(np.asarray(r["train_rmse"]) - z[f"train_rmse_{k}"]).max() < RMSE_TOL
    # sse_on_flat agrees with predict_pairs on the same model
    import b200als
    model = _model_factory(z)()
    model.fit_coo(z["flat_idx"], z["vals"].astype(np.float64) / 2.0, (m, n), tol=None, min_iters=0, verbose=0,
                  sorted_input=True)
    vi = folds[0]
    yy = np.asarray([dict(zip(z["flat_idx"].tolist(), (z["vals"] / 2.0).tolist()))[f] for f in vi])
    pp = model.predict_pairs(vi // n, vi % n)
    assert abs(model.rmse_on_flat(vi, yy) - float(np.sqrt(np.mean((yy - pp) ** 2)))) < 1e-6
    assert np.isnan(model.rmse_on_flat(np.empty(0, np.int64), np.empty(0)))
    with pytest.raises(IndexError):
        model.sse_on_flat(np.asarray([m * n]), np.asarray([1.0]))
