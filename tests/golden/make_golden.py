#!/usr/bin/env python
"""Generate the golden fixtures in this directory from the UNMODIFIED reference.

Run in the build container (the only place /root/reference exists):

    python tests/golden/make_golden.py

It imports `scripts.als.ALS`, `scripts.create_folds` and
`scripts.prepare_features` from /root/reference, runs them on small seeded
inputs, and stores inputs + outputs as .npz.  Per-iteration state is observed
(not modified) through a subclass whose `_converged` override snapshots the
attributes before delegating; with tol=-1e9, min_iters=0 the reference calls it
after every iteration (als.py:520) and never stops early.

Nothing at test / bench time reads /root/reference; only these fixtures travel.
"""
import importlib.util
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("B200ALS_REFERENCE", "/root/reference")
sys.path.insert(0, REF)            # the reference's `scripts` package, not this repo's shim

from scripts.als import ALS as RefALS                                   # noqa: E402
from scripts.als_config import (ALSConfig, BiasesConfig, CoreConfig,    # noqa: E402
                                GraphConfig, GraphSimConfig)
from scripts.create_folds import make_entrywise_folds, make_train_valid_split   # noqa: E402
from scripts.prepare_features import normalize_feature                  # noqa: E402

assert os.path.realpath(sys.modules["scripts.als"].__file__).startswith(os.path.realpath(REF))


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


synth = _load(os.path.join(REPO, "collaborative-filtering_b200", "synth.py"), "b200als_synth")


class SpyALS(RefALS):
    """Observes per-iteration state; the fit loop itself is the reference's."""

    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        self.snaps = []

    def _converged(self, tol, window=2):
        self.snaps.append(dict(U=self.U.copy(), V=self.V.copy(), b_u=self.b_u.copy(), b_i=self.b_i.copy(),
                               mu=self.mu, W={k: v.copy() for k, v in self.W.items()}))
        return super()._converged(tol, window)


def make_cfg(c):
    sim = GraphSimConfig(**c["sim"]) if c.get("sim") else None
    return ALSConfig(core=CoreConfig(**c["core"]), biases=BiasesConfig(**c.get("biases", {})),
                     graph=GraphConfig(alpha=c.get("alpha", 0.0), sim=sim))


def small_features(n, rng):
    g = (rng.random((n, 6)) < 0.3).astype(np.int64)       # multi-hot with many ties, a few all-zero rows
    y = 1950.0 + np.floor(rng.random(n) * 60.0)
    return g, y


def run_case(name, cfgd, m, n, nnz, features, lambda_w, seed, spy=True, tol=None, min_iters=5,
             empty_users=(), empty_items=(), store_f32=False, R=None):
    if R is None:
        R = synth.synth_dense(m, n, nnz, seed)
        for u in empty_users:
            R[u, :] = np.nan
        for i in empty_items:
            R[:, i] = np.nan
    flat = np.flatnonzero(~np.isnan(R))
    vals = R.ravel()[flat]
    cfg = make_cfg(cfgd)
    if spy:
        model = SpyALS(config=cfg, lambda_w=lambda_w)
        model.fit(R, features=features, tol=-1e9, min_iters=0, verbose=0)
    else:
        model = RefALS(config=cfg, lambda_w=lambda_w)
        model.fit(R, features=features, tol=tol, min_iters=min_iters, verbose=0)
    out = dict(m=m, n=n, flat_idx=flat.astype(np.int64 if not store_f32 else np.int32),
               vals=vals if not store_f32 else (vals * 2).astype(np.uint8),
               cfg=json.dumps(cfgd), lambda_w=json.dumps(lambda_w or {}),
               fit=json.dumps(dict(tol=tol, min_iters=min_iters, spy=spy)),
               feature_names=json.dumps(list((features or {}).keys())))
    cast = (lambda a: a.astype(np.float32)) if store_f32 else (lambda a: a)
    if not store_f32:
        for fn, X in (features or {}).items():
            out["X_" + fn] = X
    out.update(U=cast(model.U), V=cast(model.V), b_u=cast(model.b_u), b_i=cast(model.b_i), mu=model.mu)
    for fn, W in model.W.items():
        out["W_" + fn] = W
    for key, h in model.history.items():
        out["hist_" + key] = np.asarray(h, dtype=np.float64)
    if spy and not store_f32:
        out["snap_U"] = np.stack([s["U"] for s in model.snaps])
        out["snap_V"] = np.stack([s["V"] for s in model.snaps])
        out["snap_b_u"] = np.stack([s["b_u"] for s in model.snaps])
        out["snap_b_i"] = np.stack([s["b_i"] for s in model.snaps])
        out["snap_mu"] = np.asarray([s["mu"] for s in model.snaps])
        for fn in model.W:
            out["snap_W_" + fn] = np.stack([s["W"][fn] for s in model.snaps])
    pred = model.predict(features=features)
    if store_f32:
        rs = np.random.default_rng(7).integers(0, m * n, size=4096)
        out["pred_idx"] = rs.astype(np.int64)
        out["pred_val"] = pred.ravel()[rs]
    else:
        out["pred"] = pred
    if model.S is not None and not store_f32:
        out["S"] = model.S
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **out)
    print(f"{name}: iters={len(model.history['train_rmse'])} rmse={model.history['train_rmse'][-1]:.6f} "
          f"{os.path.getsize(path) / 1024:.0f} KiB")


def main():
    rng = np.random.default_rng(99)
    m, n, nnz = 48, 64, 800
    g, y = small_features(n, rng)
    Gn = normalize_feature(g.astype(float), "row_l2")            # float32, as the drivers pass them
    Yn = normalize_feature(y, "col_zscore")
    core = dict(n_factors=8, n_iters=4, lambda_u=5.0, lambda_v=6.0)

    run_case("small_plain", dict(core=core), m, n, nnz, None, None, 1)
    run_case("small_biases", dict(core=core, biases=dict(lambda_bu=3.0, lambda_bi=2.0)), m, n, nnz, None, None, 2)
    run_case("small_popreg", dict(core=dict(core, pop_reg_mode="inverse_sqrt"),
                                  biases=dict(lambda_bu=3.0, lambda_bi=2.0)), m, n, nnz, None, None, 3)
    run_case("small_empty", dict(core=core), m, n, nnz, None, None, 4,
             empty_users=(0, 17, 47), empty_items=(3, 63))
    run_case("small_k1", dict(core=dict(core, n_factors=1)), m, n, nnz, None, None, 5)
    run_case("small_k20", dict(core=dict(core, n_factors=20, lambda_u=0.5, lambda_v=0.7)), m, n, nnz, None, None, 6)
    run_case("small_feat1", dict(core=dict(core, update_w_every=1)), m, n, nnz,
             {"genres": Gn}, {"genres": 5.0}, 7)
    run_case("small_feat2", dict(core=dict(core, update_w_every=1)), m, n, nnz,
             {"genres": Gn, "years": Yn}, {"genres": 5.0, "years": 10.0}, 8)
    run_case("small_feat2_every3", dict(core=dict(core, n_iters=6, update_w_every=3)), m, n, nnz,
             {"genres": Gn, "years": Yn}, {"genres": 5.0}, 9)        # years: lambda_w missing → 0.0 + EPS
    run_case("small_feat_rawint", dict(core=dict(core, update_w_every=2)), m, n, nnz,
             {"genres": g}, {"genres": 1.0}, 10)                      # int64 multi-hot passed raw
    gsim = dict(source="feature", feature_name="genres", metric="cosine", topk=5, eps=1e-8)
    run_case("small_graph_topk5", dict(core=dict(core, update_w_every=1), alpha=0.5, sim=gsim), m, n, nnz,
             {"genres": Gn}, {"genres": 5.0}, 11)
    run_case("small_graph_topk40", dict(core=dict(core, update_w_every=2), alpha=2.0, sim=dict(gsim, topk=40)),
             m, n, nnz, {"genres": Gn, "years": Yn}, {"genres": 5.0, "years": 10.0}, 12,
             empty_items=(5,))
    run_case("small_graph_missing_feature", dict(core=core, alpha=0.5, sim=dict(gsim, feature_name="nope")),
             m, n, nnz, {"genres": Gn}, {"genres": 5.0}, 13)
    run_case("small_full", dict(core=dict(core, n_iters=6, pop_reg_mode="inverse_sqrt", update_w_every=5),
                                biases=dict(lambda_bu=3.0, lambda_bi=2.0), alpha=0.5, sim=dict(gsim, topk=10)),
             m, n, nnz, {"genres": Gn, "years": Yn}, {"genres": 5.0, "years": 10.0}, 14)
    run_case("small_earlystop", dict(core=dict(core, n_iters=40)), m, n, nnz, None, None, 15,
             spy=False, tol=1e-3, min_iters=3)

    # ---- bundled shape (C1 / C2): 610 x 4980, ~1e5 ratings, real genres / years
    genres = np.load(os.path.join(REF, "data", "genres.npy"))
    years = np.load(os.path.join(REF, "data", "years.npy"))
    G = normalize_feature(genres, "row_l2", impute="col_median")
    Y = normalize_feature(years, "col_zscore", impute="col_median")
    np.savez_compressed(os.path.join(HERE, "bundled_features.npz"), genres=G, years=Y)
    Rfull = synth.synth_dense(610, 4980, 100_000, 1234)
    folds = make_entrywise_folds(Rfull, n_splits=5, seed=42)
    R_train, R_valid, val_idx = make_train_valid_split(Rfull, folds, 0)
    np.savez_compressed(os.path.join(HERE, "bundled_fold0.npz"),
                        val_idx=val_idx.astype(np.int32), val_vals=(Rfull.ravel()[val_idx] * 2).astype(np.uint8),
                        fold_sizes=np.asarray([len(f) for f in folds]),
                        fold_heads=np.stack([f[:8] for f in folds]))
    c1 = dict(core=dict(n_factors=32, n_iters=4, lambda_u=5.0, lambda_v=6.0),
              biases=dict(lambda_bu=3.0, lambda_bi=2.0))
    run_case("c1_k32", c1, 610, 4980, 0, None, None, 0, spy=False, tol=None, store_f32=True, R=R_train)
    c2 = dict(core=dict(n_factors=32, n_iters=7, lambda_u=5.0, lambda_v=6.0, pop_reg_mode="inverse_sqrt",
                        random_state=42, update_w_every=5),
              biases=dict(lambda_bu=3.0, lambda_bi=2.0), alpha=0.5,
              sim=dict(source="feature", feature_name="genres", metric="cosine", topk=50, eps=1e-8))
    run_case("c2_full", c2, 610, 4980, 0, {"genres": G, "years": Y}, {"genres": 5.0, "years": 10.0}, 0,
             spy=False, tol=None, store_f32=True, R=R_train)


if __name__ == "__main__":
    main()
