"""The reference's own drivers, imported UNMODIFIED through the `scripts/` shim, bind this repo's ALS class
(`from scripts.als import ALS`, tune_params.py:86, evaluate_models.py:95) and drive it
through exactly the calls they make on the reference's — constructor, `fit(R_train, features=, tol=, min_iters=,
verbose=)`, `history.get("train_rmse")`, `predict(features=)`.

Needs the reference checkout (build container only; skipped elsewhere) and no GPU: the device layer below the class
is substituted by the CPU oracle for this test, so what is checked is the host-side surface the drivers touch and the
numbers they compute from it, against the reference's own ALS run through the same driver functions.
`optuna` / `matplotlib` are absent from the image; empty stand-in modules satisfy the drivers' imports.
"""
import importlib
import importlib.util
import os
import sys
import types

import numpy as np


def _dummy_attr(attr):
    """Stand-in modules: any public attribute is a dummy class (nothing of them is ever called)."""
    if attr.startswith("__"):
        raise AttributeError(attr)
    return type(attr, (), {})
import pytest

REF = os.environ.get("B200ALS_REFERENCE", "/root/reference")
pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "scripts")),
                                reason="reference checkout not present (build container only)")


def _stub_missing_modules():
    for name in ("optuna", "optuna.samplers", "optuna.pruners", "optuna.visualization", "optuna.trial",
                 "matplotlib", "matplotlib.pyplot"):
        if name in sys.modules:
            continue
        try:
            importlib.import_module(name)
        except Exception:
            mod = types.ModuleType(name)
            mod.__getattr__ = _dummy_attr
            sys.modules[name] = mod


@pytest.fixture()
def drivers(monkeypatch):
    """(tune_params, evaluate_models) of the reference, loaded through this repo's scripts/ shim."""
    _stub_missing_modules()
    import b200als                                      # noqa: F401  (puts the repo root first on sys.path)
    for k in [k for k in sys.modules if k == "scripts" or k.startswith("scripts.")]:
        monkeypatch.delitem(sys.modules, k)
    monkeypatch.setenv("B200ALS_REFERENCE_SCRIPTS", os.path.join(REF, "scripts"))
    tp = importlib.import_module("scripts.tune_params")
    em = importlib.import_module("scripts.evaluate_models")
    assert os.path.realpath(tp.__file__).startswith(os.path.realpath(REF))      # the reference's file, unmodified
    assert os.path.realpath(em.__file__).startswith(os.path.realpath(REF))
    return tp, em


def _oracle_backed(monkeypatch):
    """Substitute the device layer of b200als.ALS by the CPU oracle (same algorithm, als.py:300-574)."""
    import b200als
    import als_oracle

    def fit(self, R, features=None, tol=1e-3, min_iters=5, verbose=1):
        o = als_oracle.SparseALSOracle(self.cfg, self.lambda_w)
        o.fit(np.asarray(R), features=features, tol=tol, min_iters=min_iters)
        for a in ("U", "V", "W", "b_u", "b_i", "mu", "S", "history"):
            setattr(self, a, getattr(o, a))
        self._oracle = o
        return self

    def predict(self, features=None):
        if self.U is None or self.V is None:
            raise RuntimeError("Model must be fitted before prediction.")
        return self._oracle.predict(features)

    monkeypatch.setattr(b200als.ALS, "fit", fit)
    monkeypatch.setattr(b200als.ALS, "predict", predict)


def _data():
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle"))
    from b200als.synth import synth_dense
    m, n = 40, 30
    R = synth_dense(m, n, 500, 5)
    rng = np.random.default_rng(1)
    feats = {"genres": (rng.random((n, 4)) < 0.4).astype(np.float32)}
    feats["genres"][feats["genres"].sum(1) == 0, 0] = 1.0
    return R, feats


def test_drivers_bind_this_als(drivers):
    import b200als
    tp, em = drivers
    assert tp.ALS is b200als.ALS and em.ALS is b200als.ALS
    assert tp.ALSConfig is b200als.ALSConfig and tp.GraphSimConfig is b200als.GraphSimConfig
    cfg = tp._make_config(dict(n_factors=8, n_iters=5, lambda_u=2.0, lambda_v=3.0, alpha=0.5, graph_feature="genres",
                               S_topk=7, update_w_every=2, lambda_bu=0.0, lambda_bi=1.0))
    assert isinstance(cfg, b200als.ALSConfig)
    model = tp.ALS(config=cfg, lambda_w={"genres": 5.0})
    assert model.lambda_bu == 2.0            # `lambda_bu or lambda_u` fallback (als.py:166-167)
    assert model.S_topk == 7 and model.alpha == 0.5 and model.update_w_every == 2


def test_cv_score_single_trial_runs_on_this_class(drivers, monkeypatch):
    """tune_params._cv_score_single_trial (:341-421) and evaluate_models._eval_variant_cv (:194-276), unmodified,
    against this class == the same functions against the reference's own ALS."""
    tp, em = drivers
    R, feats = _data()
    folds = importlib.import_module("scripts.create_folds").make_entrywise_folds(R, n_splits=3, seed=42)
    params = tp._normalize_params(dict(n_factors=6, n_iters=12, lambda_u=2.0, lambda_v=2.5, lambda_bu=1.0, lambda_bi=1.5,
                                       alpha=0.3, graph_feature="genres", S_topk=5, update_w_every=3,
                                       pop_reg_mode="inverse_sqrt", lambda_w_genres=4.0), R.shape, list(feats))
    item_bin, _ = em._popularity_bins_from_R(R)

    def run_both():
        score = tp._cv_score_single_trial(dict(params), R, feats, folds, verbose_fit=0, trial=None)
        curves = {}
        fold_rmse, _, fold_bin_rmse, fold_iters = em._eval_variant_cv(
            "full", R, feats, folds, dict(params), item_bin, em.N_POP_BINS, tp.ES_TOL, tp.ES_MIN_ITERS, curves,
            verbose_fit=0)
        return score, fold_rmse, fold_bin_rmse, fold_iters, curves

    # (1) the reference's own ALS through the reference's drivers
    spec = importlib.util.spec_from_file_location("ref_als_module", os.path.join(REF, "scripts", "als.py"))
    ref_mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref_mod)
    with monkeypatch.context() as mp:
        mp.setattr(tp, "ALS", ref_mod.ALS)
        mp.setattr(em, "ALS", ref_mod.ALS)
        want = run_both()
    # (2) this repo's class through the same, unmodified driver functions
    import b200als
    assert tp.ALS is b200als.ALS and em.ALS is b200als.ALS
    _oracle_backed(monkeypatch)
    got = run_both()
    assert abs(got[0] - want[0]) < 1e-9
    assert np.allclose(got[1], want[1], atol=1e-9) and got[3] == want[3]
    for g, w in zip(got[2], want[2]):
        assert g.keys() == w.keys()
        for key in g:
            assert (np.isnan(g[key]) and np.isnan(w[key])) or abs(g[key] - w[key]) < 1e-9
    assert np.allclose(got[4]["full"], want[4]["full"], atol=1e-9)
