#!/usr/bin/env python
"""Golden fixture for the fold / evaluation rows, from the UNMODIFIED reference.

    python tests/golden/make_golden_eval.py          # build container only (/root/reference)

Runs the reference's own `make_entrywise_folds`, `make_train_valid_split` (create_folds.py),
`_popularity_bins_from_R`, `_split_val_indices_by_popularity` (evaluate_models.py), `_rmse_on_indices`
(tune_params.py) and `ALS.fit/predict` on a small seeded matrix and stores inputs + outputs in eval_folds.npz.
`optuna` and `matplotlib` are absent from this image; the two driver modules only need them at import time, so
empty stand-in modules are injected before importing them (nothing of them is called).
"""
import os
import sys
import types

import numpy as np


def _dummy_attr(attr):
    """Stand-in modules: any public attribute is a dummy class (nothing of them is ever called)."""
    if attr.startswith("__"):
        raise AttributeError(attr)
    return type(attr, (), {})

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("B200ALS_REFERENCE", "/root/reference")
sys.path.insert(0, REF)

for name in ("optuna", "optuna.samplers", "optuna.pruners", "optuna.visualization", "optuna.trial",
             "matplotlib", "matplotlib.pyplot"):
    if name not in sys.modules:
        try:
            __import__(name)
        except Exception:
            mod = types.ModuleType(name)
            mod.__getattr__ = _dummy_attr
            sys.modules[name] = mod

from scripts.als import ALS as RefALS                                                    # noqa: E402
from scripts.als_config import ALSConfig, BiasesConfig, CoreConfig                       # noqa: E402
from scripts.create_folds import make_entrywise_folds, make_train_valid_split           # noqa: E402
from scripts.tune_params import _rmse_on_indices                                        # noqa: E402
from scripts.evaluate_models import (_popularity_bins_from_R,                           # noqa: E402
                                     _split_val_indices_by_popularity)

assert os.path.realpath(sys.modules["scripts.evaluate_models"].__file__).startswith(os.path.realpath(REF))

import importlib.util                                                                   # noqa: E402
spec = importlib.util.spec_from_file_location("b200als_synth",
                                              os.path.join(REPO, "collaborative-filtering_b200", "synth.py"))
synth = importlib.util.module_from_spec(spec)
spec.loader.exec_module(synth)


def main():
    m, n, nnz, K, n_bins = 70, 50, 1400, 4, 5
    R = synth.synth_dense(m, n, nnz, 77)
    R[:, 3] = np.nan                                   # an item nobody rated (count 0 in the bins)
    obs = np.flatnonzero(~np.isnan(R))
    out = dict(m=m, n=n, flat_idx=obs.astype(np.int64), vals=(R.ravel()[obs] * 2).astype(np.uint8), n_bins=n_bins)
    folds = make_entrywise_folds(R, n_splits=K, seed=7)
    for strategy in ("quantile", "uniform"):
        item_bin, edges = _popularity_bins_from_R(R, n_bins=n_bins, strategy=strategy)
        out[f"bin_{strategy}"] = item_bin.astype(np.int64)
        out[f"edges_{strategy}"] = edges
    item_bin, _ = _popularity_bins_from_R(R, n_bins=n_bins, strategy="quantile")
    cfg = dict(n_factors=8, n_iters=6, lambda_u=2.0, lambda_v=2.0, random_state=3)
    out["cfg_core"] = np.asarray(list(cfg.values()), dtype=np.float64)
    for k in range(K):
        R_train, R_valid, val_idx = make_train_valid_split(R, folds, k)
        out[f"val_idx_{k}"] = np.asarray(val_idx, dtype=np.int64)
        out[f"train_flat_{k}"] = np.flatnonzero(~np.isnan(R_train)).astype(np.int64)
        model = RefALS(ALSConfig(core=CoreConfig(**cfg), biases=BiasesConfig(lambda_bu=1.5, lambda_bi=1.5)))
        model.fit(R_train, features=None, tol=None, min_iters=0, verbose=0)
        R_hat = model.predict(None)
        out[f"rmse_{k}"] = _rmse_on_indices(R_valid, R_hat, val_idx)
        bins = _split_val_indices_by_popularity(val_idx, R.shape, item_bin, n_bins)
        out[f"bin_rmse_{k}"] = np.asarray([_rmse_on_indices(R_valid, R_hat, b) for b in bins], dtype=np.float64)
        out[f"bin_cnt_{k}"] = np.asarray([len(b) for b in bins], dtype=np.int64)
        out[f"train_rmse_{k}"] = np.asarray(model.history["train_rmse"])
    np.savez_compressed(os.path.join(HERE, "eval_folds.npz"), **out)
    print("wrote eval_folds.npz:", {k: (np.asarray(v).shape) for k, v in out.items() if k.startswith(("rmse", "bin_cnt"))})


if __name__ == "__main__":
    main()
