"""Shared helpers for the test-suite: golden-fixture loading and oracle runs."""
import json
import os
import sys
from types import SimpleNamespace

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(REPO, "tests", "golden")
sys.path.insert(0, os.path.join(REPO, "oracle"))

SMALL_CASES = [
    "small_plain", "small_biases", "small_popreg", "small_empty", "small_k1", "small_k20",
    "small_feat1", "small_feat2", "small_feat2_every3", "small_feat_rawint",
    "small_graph_topk5", "small_graph_topk40", "small_graph_missing_feature", "small_full",
]


def load_case(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)
    c = SimpleNamespace()
    c.name = name
    c.m, c.n = int(z["m"]), int(z["n"])
    c.flat = z["flat_idx"].astype(np.int64)
    vals = z["vals"]
    c.vals = vals.astype(np.float64) / 2.0 if vals.dtype == np.uint8 else vals.astype(np.float64)
    c.cfgd = json.loads(str(z["cfg"]))
    c.lambda_w = json.loads(str(z["lambda_w"]))
    c.fit = json.loads(str(z["fit"]))
    c.feature_names = json.loads(str(z["feature_names"]))
    c.features = {fn: z["X_" + fn] for fn in c.feature_names if "X_" + fn in z.files}
    c.z = z
    return c


def bundled_features():
    z = np.load(os.path.join(GOLDEN, "bundled_features.npz"))
    return {"genres": z["genres"], "years": z["years"]}


def make_cfg(cfgd, mod=None):
    """Config dataclasses from the fixture's dict (this repo's by default)."""
    if mod is None:
        import b200als as mod
    sim = mod.GraphSimConfig(**cfgd["sim"]) if cfgd.get("sim") else None
    return mod.ALSConfig(core=mod.CoreConfig(**cfgd["core"]), biases=mod.BiasesConfig(**cfgd.get("biases", {})),
                         graph=mod.GraphConfig(alpha=cfgd.get("alpha", 0.0), sim=sim))


def dense_from_case(c):
    R = np.full(c.m * c.n, np.nan)
    R[c.flat] = c.vals
    return R.reshape(c.m, c.n)


def run_oracle(c, features=None, spy=True):
    import als_oracle
    cfg = make_cfg(c.cfgd)
    o = als_oracle.SparseALSOracle(cfg, c.lambda_w)
    sp = als_oracle.csr_csc_from_flat(c.flat, c.vals, c.m, c.n)
    snaps = []

    def on_iter(it, s):
        snaps.append(dict(U=s.U.copy(), V=s.V.copy(), b_u=s.b_u.copy(), b_i=s.b_i.copy(), mu=s.mu,
                          W={k: v.copy() for k, v in s.W.items()}))
    feats = c.features if features is None else features
    if c.fit["spy"]:
        o.fit_sparse(sp, c.m, c.n, feats, tol=None, on_iter=on_iter)
    else:
        o.fit_sparse(sp, c.m, c.n, feats, tol=c.fit["tol"], min_iters=c.fit["min_iters"], on_iter=on_iter)
    return o, snaps


def rel_err(a, b):
    """Frobenius relative error of a vs reference b."""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    den = np.linalg.norm(b)
    return float(np.linalg.norm(a - b) / den) if den > 0 else float(np.linalg.norm(a - b))
