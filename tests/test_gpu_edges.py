"""GPU: the edges of the reference's own hyper-parameter range on the float32 device path, the
exception surface of the solve, and the reference drivers' dense call sequence on the real CUDA class.

  * lambda_u = lambda_v over the tuner's log-uniform range [1e-4, 1e4] (scripts/tune_params.py:102-103) with rows
    that hold FEWER ratings than factors (the Gram is rank deficient and only the ridge keeps it SPD), at
    k = 32 (FFMA path), 64 and 128 (tcgen05 path) — the rows float32 cannot resolve take the float64 half-step;
  * the `lambda_w = 0` ablation form (scripts/evaluate_models.py:413-426): features passed, W_f regularised by EPS alone;
  * a system that is not positive definite raises numpy.linalg.LinAlgError like scripts/helpers.py:19;
  * k = 50 at the bundled shape; a graph + features model at k = 64 (FACTOR mode keeps the item step on FFMA);
  * every k in (32, 128] takes the tcgen05 kernel through zero padding;
  * `fit(R_train) -> predict() -> .ravel()[idx]`, the sequence of scripts/tune_params.py:376-392 and
    scripts/evaluate_models.py:246-265, restated line for line and run on this class, against the reference's own
    numbers (tests/golden/eval_folds.npz);
  * input validation of fit_coo (out-of-range, duplicate, unsorted-but-promised-sorted, non-finite);
  * refit on the same object appends to history (scripts/als.py:185-191).
"""
import os

import numpy as np
import pytest

from _util import GOLDEN, bundled_features, load_case, make_cfg, rel_err

pytestmark = pytest.mark.gpu

FACTOR_TOL = 1e-3
RMSE_TOL = 1e-4


def _oracle(cfg, flat, vals, m, n, features=None, lambda_w=None):
    import als_oracle
    return als_oracle.SparseALSOracle(cfg, lambda_w).fit_sparse(als_oracle.csr_csc_from_flat(flat, vals, m, n), m, n,
                                                                 features, tol=None)


def _short_row_problem(seed, m=400, n=250, per_row=20):
    """Every user holds ~per_row ratings and most items fewer than 64: cnt < k on both sides for k >= 64."""
    from b200als.synth import synth_ratings
    return synth_ratings(m, n, m * per_row, seed=seed) + (m, n)


@pytest.mark.parametrize("k", [32, 64, 128])
@pytest.mark.parametrize("lam", [1e-4, 1e-2, 1.0, 1e4])
def test_lambda_range_short_rows_vs_oracle(cuda_device, k, lam):
    """The whole range of the tuner's lambda at rows with fewer ratings than factors.  Below ~0.1 the float32 paths
    cannot hold 1e-3 (measured: 2e-2 at lambda = 1e-2 on the tcgen05 path, loss of positive definiteness at 1e-4);
    `precision = "auto"` sends exactly those rows — small ridge AND few ratings — to the float64 half-step."""
    import b200als
    flat, vals, m, n = _short_row_problem(seed=100 + k)
    cfg = b200als.ALSConfig(core=b200als.CoreConfig(n_factors=k, n_iters=3, lambda_u=lam, lambda_v=lam),
                            biases=b200als.BiasesConfig(lambda_bu=2.0, lambda_bi=2.5))
    o = _oracle(cfg, flat, vals, m, n)
    model = b200als.ALS(cfg).fit_coo(flat, vals, (m, n), tol=None, verbose=0, sorted_input=True)
    errs = {a: rel_err(getattr(model, a), getattr(o, a)) for a in ("U", "V", "b_u", "b_i")}
    d = float(np.abs(np.asarray(model.history["train_rmse"]) - np.asarray(o.history["train_rmse"])).max())
    print(f"k={k} lambda={lam:g}: {errs} dRMSE={d:.2e} paths={model.last_paths}")
    assert max(errs.values()) < FACTOR_TOL, errs
    assert d < RMSE_TOL
    if lam >= 1.0:
        want = b200als._lib.PATH_UMMA if k > 32 else b200als._lib.PATH_FFMA
        assert model.last_paths == {"user": want, "item": want}          # nothing left the fast paths


@pytest.mark.parametrize("k", [8, 50, 128, 150])
def test_float64_path_every_row(cuda_device, k):
    """precision = "f64": every row through the float64 half-step — the reference's arithmetic, float32 storage."""
    import b200als
    from b200als.synth import synth_ratings
    m, n, nnz = 300, 200, 9000
    flat, vals = synth_ratings(m, n, nnz, seed=k)
    cfg = b200als.ALSConfig(core=b200als.CoreConfig(n_factors=k, n_iters=3, lambda_u=4.0, lambda_v=3.0,
                                                    pop_reg_mode="inverse_sqrt"),
                            biases=b200als.BiasesConfig(lambda_bu=2.0, lambda_bi=2.5))
    o = _oracle(cfg, flat, vals, m, n)
    model = b200als.ALS(cfg)
    model.precision = "f64"
    model.fit_coo(flat, vals, (m, n), tol=None, verbose=0, sorted_input=True)
    assert model.last_paths == {"user": b200als._lib.PATH_F64, "item": b200als._lib.PATH_F64}
    for a in ("U", "V", "b_u", "b_i"):
        assert rel_err(getattr(model, a), getattr(o, a)) < 2e-6, a
    assert np.abs(np.asarray(model.history["train_rmse"]) - np.asarray(o.history["train_rmse"])).max() < 1e-6
    for key, arr in (("U_norm", o.U), ("V_norm", o.V), ("bu_norm", o.b_u), ("bi_norm", o.b_i)):
        assert abs(model.history[key][-1] - np.linalg.norm(arr)) < 1e-5 * max(np.linalg.norm(arr), 1.0)


def test_full_model_small_lambda_float64_factor_mode(cuda_device):
    """Graph + features with a small lambda_v: the Gauss-Seidel item step factors its short rows in float64."""
    import b200als
    c = load_case("small_full")
    cfgd = dict(c.cfgd, core=dict(c.cfgd["core"], lambda_u=0.01, lambda_v=0.01, n_iters=3))
    cfg = make_cfg(cfgd)
    o = _oracle(cfg, c.flat, c.vals, c.m, c.n, c.features, c.lambda_w)
    model = b200als.ALS(config=cfg, lambda_w=c.lambda_w).fit_coo(c.flat, c.vals, (c.m, c.n), features=c.features,
                                                                 tol=None, verbose=0)
    for a in ("U", "V", "b_u", "b_i"):
        assert rel_err(getattr(model, a), getattr(o, a)) < FACTOR_TOL, a
    assert np.abs(np.asarray(model.history["train_rmse"]) - np.asarray(o.history["train_rmse"])).max() < RMSE_TOL


def test_lambda_w_zero_ablation_form(cuda_device):
    """evaluate_models.py:413-426: the `no_features`-style variants pass the features with lambda_w = 0."""
    import b200als
    c = load_case("small_feat2")
    lw = {fn: 0.0 for fn in c.feature_names}
    cfg = make_cfg(c.cfgd)
    o = _oracle(cfg, c.flat, c.vals, c.m, c.n, c.features, lw)
    model = b200als.ALS(config=cfg, lambda_w=lw).fit_coo(c.flat, c.vals, (c.m, c.n), features=c.features, tol=None,
                                                         verbose=0)
    for a in ("U", "V", "b_u", "b_i"):
        assert rel_err(getattr(model, a), getattr(o, a)) < FACTOR_TOL, a
    for fn in c.feature_names:
        assert rel_err(model.W[fn], o.W[fn]) < FACTOR_TOL, fn
    assert np.abs(np.asarray(model.history["train_rmse"]) - np.asarray(o.history["train_rmse"])).max() < RMSE_TOL


@pytest.mark.parametrize("k", [8, 64, 100])
def test_not_positive_definite_raises_linalgerror(cuda_device, k):
    """A negative ridge larger than the Gram's smallest eigenvalue: scipy's cho_factor raises LinAlgError
    (scripts/helpers.py:19); so does this class, on both Gram paths."""
    import b200als
    from b200als.synth import synth_ratings
    m, n, nnz = 200, 150, 4000
    flat, vals = synth_ratings(m, n, nnz, seed=9)
    cfg = b200als.ALSConfig(core=b200als.CoreConfig(n_factors=k, n_iters=2, lambda_u=-50.0, lambda_v=3.0),
                            biases=b200als.BiasesConfig(lambda_bu=2.0, lambda_bi=2.0))
    with pytest.raises(np.linalg.LinAlgError):
        _oracle(cfg, flat, vals, m, n)
    with pytest.raises(np.linalg.LinAlgError):
        b200als.ALS(cfg).fit_coo(flat, vals, (m, n), tol=None, verbose=0, sorted_input=True)


def test_bundled_shape_k50_vs_oracle(cuda_device):
    """C1 at k = 50 (scripts/als_config.py:24's example value): tcgen05 path through zero padding to 64."""
    import b200als
    c = load_case("c1_k32")
    cfgd = dict(c.cfgd, core=dict(c.cfgd["core"], n_factors=50, n_iters=3))
    cfg = make_cfg(cfgd)
    o = _oracle(cfg, c.flat, c.vals, c.m, c.n)
    model = b200als.ALS(cfg).fit_coo(c.flat, c.vals, (c.m, c.n), tol=None, verbose=0)
    assert model.last_paths == {"user": b200als._lib.PATH_UMMA, "item": b200als._lib.PATH_UMMA}
    for a in ("U", "V", "b_u", "b_i"):
        assert rel_err(getattr(model, a), getattr(o, a)) < FACTOR_TOL, a
    assert model.U.shape == (c.m, 50) and model.V.shape == (c.n, 50)
    assert np.abs(np.asarray(model.history["train_rmse"]) - np.asarray(o.history["train_rmse"])).max() < RMSE_TOL


def test_graph_and_features_at_k64(cuda_device):
    """The full model at k = 64: the user step takes the tcgen05 kernel, the Gauss-Seidel item step needs the stored
    Cholesky factors (FACTOR mode) and stays on FFMA."""
    import b200als
    c = load_case("small_full")
    cfgd = dict(c.cfgd, core=dict(c.cfgd["core"], n_factors=64, n_iters=3))
    cfg = make_cfg(cfgd)
    o = _oracle(cfg, c.flat, c.vals, c.m, c.n, c.features, c.lambda_w)
    model = b200als.ALS(config=cfg, lambda_w=c.lambda_w).fit_coo(c.flat, c.vals, (c.m, c.n), features=c.features,
                                                                 tol=None, verbose=0)
    assert model.last_paths["user"] == b200als._lib.PATH_UMMA
    assert model.last_paths["item"] == b200als._lib.PATH_FFMA
    for a in ("U", "V", "b_u", "b_i"):
        assert rel_err(getattr(model, a), getattr(o, a)) < FACTOR_TOL, a
    for fn in c.feature_names:
        assert rel_err(model.W[fn], o.W[fn]) < FACTOR_TOL, fn
    assert np.abs(np.asarray(model.history["train_rmse"]) - np.asarray(o.history["train_rmse"])).max() < RMSE_TOL


@pytest.mark.parametrize("k", [33, 40, 50, 72, 96, 100, 127, 129, 150])
def test_every_k_above_32_on_the_tensor_core(cuda_device, k):
    """PATH_AUTO: k in (32, 128] runs the tcgen05 kernel (rows zero-padded to 64 / 128), k > 128 stays on FFMA; both
    match the oracle.  (scripts/tune_params.py:98-99 samples n_factors uniformly in [1, 150].)"""
    import b200als
    from b200als.synth import synth_ratings
    m, n, nnz = 300, 200, 9000
    flat, vals = synth_ratings(m, n, nnz, seed=k)
    cfg = b200als.ALSConfig(core=b200als.CoreConfig(n_factors=k, n_iters=3, lambda_u=4.0, lambda_v=3.0),
                            biases=b200als.BiasesConfig(lambda_bu=2.0, lambda_bi=2.5))
    o = _oracle(cfg, flat, vals, m, n)
    model = b200als.ALS(cfg)
    model.split_rows, model.split_piece = 96, 64            # a few rows go through the split-row slots as well
    model.fit_coo(flat, vals, (m, n), tol=None, verbose=0, sorted_input=True)
    want = b200als._lib.PATH_UMMA if k <= 128 else b200als._lib.PATH_FFMA
    assert model.last_paths == {"user": want, "item": want}
    assert model.U.shape == (m, k) and model.V.shape == (n, k)
    for a in ("U", "V", "b_u", "b_i"):
        assert rel_err(getattr(model, a), getattr(o, a)) < FACTOR_TOL, a
    assert np.abs(np.asarray(model.history["train_rmse"]) - np.asarray(o.history["train_rmse"])).max() < RMSE_TOL
    for key, arr in (("U_norm", o.U), ("V_norm", o.V)):
        assert abs(model.history[key][-1] - np.linalg.norm(arr)) < 1e-3 * np.linalg.norm(arr)


def test_reference_driver_sequence_on_the_cuda_class(cuda_device):
    """tune_params.py:376-392 / evaluate_models.py:246-274 restated: dense NaN-filled train matrix -> fit -> dense
    predict -> .ravel()[val_idx] -> masked RMSE (tune_params.py:147-167), overall and per popularity bin, on THIS
    class (nothing substituted), against the numbers the reference's own drivers produced on the same folds."""
    import b200als
    z = np.load(os.path.join(GOLDEN, "eval_folds.npz"))
    m, n, n_bins = int(z["m"]), int(z["n"]), int(z["n_bins"])
    R = np.full(m * n, np.nan)
    R[z["flat_idx"]] = z["vals"].astype(np.float64) / 2.0
    R = R.reshape(m, n)
    k, iters, lu, lv, seed = z["cfg_core"]
    item_bin = z["bin_quantile"]

    def rmse_on_indices(R_true, R_pred, idx):                      # tune_params.py:147-167
        if len(idx) == 0:
            return float("nan")
        y_true, y_pred = R_true.ravel()[idx], R_pred.ravel()[idx]
        mask = ~np.isnan(y_true)
        return float(np.sqrt(np.mean((y_true[mask] - y_pred[mask]) ** 2)))

    for fold in range(4):
        val_idx = z[f"val_idx_{fold}"]
        R_train = R.copy()                                         # create_folds.py:200-206
        R_train.ravel()[val_idx] = np.nan
        R_valid = np.full_like(R, np.nan)
        R_valid.ravel()[val_idx] = R.ravel()[val_idx]
        model = b200als.ALS(b200als.ALSConfig(
            core=b200als.CoreConfig(n_factors=int(k), n_iters=int(iters), lambda_u=float(lu), lambda_v=float(lv),
                                    random_state=int(seed)),
            biases=b200als.BiasesConfig(lambda_bu=1.5, lambda_bi=1.5)))
        model.fit(R_train, features=None, tol=None, min_iters=0, verbose=0)
        R_pred = model.predict(None)
        assert R_pred.shape == (m, n) and R_pred.dtype == np.float64
        assert abs(rmse_on_indices(R_valid, R_pred, val_idx) - float(z[f"rmse_{fold}"])) < RMSE_TOL
        cols = val_idx % n                                         # evaluate_models.py:173-191
        got = np.asarray([rmse_on_indices(R_valid, R_pred, val_idx[item_bin[cols] == b]) for b in range(n_bins)])
        want = z[f"bin_rmse_{fold}"]
        assert np.array_equal(np.isnan(got), np.isnan(want))
        assert np.nanmax(np.abs(got - want)) < RMSE_TOL
        assert len(model.history.get("train_rmse", [])) == len(z[f"train_rmse_{fold}"])
        assert np.abs(np.asarray(model.history["train_rmse"]) - z[f"train_rmse_{fold}"]).max() < RMSE_TOL
        # the resident model and the public float64 attributes describe the same model: reading the attributes
        # (which makes them authoritative) must not change the prediction beyond float32 rounding
        _ = model.U, model.V, model.b_u, model.b_i
        assert np.abs(model.predict(None) - R_pred).max() < 1e-5


def test_fit_coo_input_validation(cuda_device):
    import b200als
    from b200als.synth import synth_ratings
    m, n, nnz = 60, 40, 500
    flat, vals = synth_ratings(m, n, nnz, seed=1)
    cfg = b200als.ALSConfig(core=b200als.CoreConfig(n_factors=4, n_iters=1, lambda_u=1.0, lambda_v=1.0))

    def fit(f, v, **kw):
        return b200als.ALS(cfg).fit_coo(f, v, (m, n), tol=None, verbose=0, **kw)
    fit(flat, vals, sorted_input=True)
    bad = flat.copy(); bad[7] = m * n                           # past the end: matrix_from_indices raises IndexError
    with pytest.raises(IndexError):
        fit(bad, vals)
    bad = flat.copy(); bad[0] = -1
    with pytest.raises(IndexError):
        fit(bad, vals)
    with pytest.raises(ValueError, match="duplicates"):
        fit(np.concatenate([flat, flat[:3]]), np.concatenate([vals, vals[:3]]))
    with pytest.raises(ValueError, match="ascending"):
        fit(flat[::-1].copy(), vals[::-1].copy(), sorted_input=True)
    v = vals.copy(); v[5] = np.nan
    with pytest.raises(ValueError, match="finite"):
        fit(flat, v)


def test_refit_appends_history(cuda_device):
    """history is created in __init__ and every fit appends (scripts/als.py:185-191, 510-517)."""
    import b200als
    c = load_case("small_biases")
    model = b200als.ALS(config=make_cfg(c.cfgd))
    model.fit_coo(c.flat, c.vals, (c.m, c.n), tol=None, verbose=0)
    first = list(model.history["train_rmse"])
    model.fit_coo(c.flat, c.vals, (c.m, c.n), tol=None, verbose=0)
    assert len(model.history["train_rmse"]) == 2 * len(first)
    assert model.history["train_rmse"][:len(first)] == first
    assert np.allclose(model.history["train_rmse"][len(first):], first, atol=1e-6)     # same seed, same fit
    assert len(model.history["U_norm"]) == 2 * len(first)


@pytest.mark.parametrize("k", [32, 64])
def test_slab_storage_single_rank_equals_fit_coo(cuda_device, k):
    """slab.RatingSlabs with one rank is the whole matrix with local == global ids: the fit through the slab entry
    (row_ids map in the kernels, per-rank norm shares) must equal the ordinary fit — bit for bit where the path is
    run-to-run deterministic (FFMA), to float32 rounding on the tcgen05 path."""
    import torch
    import b200als
    from b200als.slab import slabs_from_coo
    from b200als.synth import synth_ratings
    m, n, nnz = 500, 300, 20000
    flat, vals = synth_ratings(m, n, nnz, seed=21)
    keep = (flat // n != 3) & (flat % n != 11)               # a user and an item nobody rated
    flat, vals = flat[keep], vals[keep]
    cfg = b200als.ALSConfig(core=b200als.CoreConfig(n_factors=k, n_iters=3, lambda_u=4.0, lambda_v=3.0),
                            biases=b200als.BiasesConfig(lambda_bu=2.0, lambda_bi=2.5))
    a = b200als.ALS(cfg).fit_coo(flat, vals, (m, n), tol=None, verbose=0, sorted_input=True)
    rows, cols = torch.from_numpy(flat // n).to(cuda_device), torch.from_numpy(flat % n).to(cuda_device)
    slabs = slabs_from_coo(rows, cols, torch.from_numpy(vals.astype(np.float32)).to(cuda_device), m, n, 0, 1)
    b = b200als.ALS(cfg).fit_slabs(slabs, tol=None, verbose=0)
    for attr in ("U", "V", "b_u", "b_i"):
        if k <= 32:
            assert np.array_equal(getattr(a, attr), getattr(b, attr)), attr       # FFMA path: one CTA per row, bit for bit
        else:
            # tcgen05 path: which solver group (forward or reversed elimination order) takes a row depends on the
            # order the CTAs claim work from the atomic cursor, so two runs of the SAME fit differ at float32
            # rounding level (measured ~7e-6); a tenth of the parity bar is the bound here
            assert rel_err(getattr(a, attr), getattr(b, attr)) < 1e-4, attr
    for key in a.history:                       # float64 atomics accumulate in a different order: last-digit agreement
        assert np.allclose(a.history[key], b.history[key], rtol=1e-12 if k <= 32 else 1e-5, atol=0), key


@pytest.mark.parametrize("N", [1, 31, 33, 64, 65, 100, 257, 608, 1000])
def test_cholesky_solve_f64_vs_numpy(cuda_device, N):
    """b200als_cholesky_solve_f64 (the W step's (d k)^2 float64 system, scripts/als.py:497-500): the single-CTA kernel
    up to N = 64, the blocked cooperative kernel above, ragged last blocks, and the not-positive-definite flag."""
    import torch
    import b200als
    from b200als._lib import check, load, ptr
    lib = load()
    rng = np.random.default_rng(N)
    B = rng.normal(size=(N, N + 3))
    A = B @ B.T + 0.5 * np.eye(N)
    rhs = rng.normal(size=N)
    want = np.linalg.solve(A, rhs)
    A_d = torch.from_numpy(A.copy()).to(cuda_device)
    b_d = torch.from_numpy(rhs.copy()).to(cuda_device)
    status = torch.zeros(4, dtype=torch.int32, device=cuda_device)
    check(lib.b200als_cholesky_solve_f64(ptr(A_d), ptr(b_d), N, ptr(status), torch.cuda.current_stream().cuda_stream))
    got = b_d.cpu().numpy()
    assert int(status[0]) == 0
    assert np.abs(got - want).max() <= 1e-9 * max(1.0, np.abs(want).max())
    Lg = np.tril(A_d.cpu().numpy())
    assert np.allclose(Lg, np.linalg.cholesky(A), rtol=1e-9, atol=1e-9)
    if N >= 2:
        A[N // 2, N // 2] = -1.0                                 # not positive definite
        A_d = torch.from_numpy(A).to(cuda_device)
        b_d = torch.from_numpy(rhs.copy()).to(cuda_device)
        check(lib.b200als_cholesky_solve_f64(ptr(A_d), ptr(b_d), N, ptr(status), torch.cuda.current_stream().cuda_stream))
        assert int(status[0]) == 1


@pytest.mark.parametrize("method,impute", [("none", "col_median"), ("row_l1", "none"), ("row_l2", "col_median"),
                                           ("col_zscore", "col_median"), ("col_minmax", "none")])
def test_normalize_feature_on_device(cuda_device, method, impute):
    """features.normalize_feature against a numpy restatement of scripts/prepare_features.py:83-191 on the bundled
    raw shapes: int64 multi-hot genres with all-zero rows, float64 years with NaNs (imputation mandatory)."""
    from b200als.features import normalize_feature
    rng = np.random.default_rng(5)
    n = 4980
    genres = (rng.random((n, 19)) < 0.12).astype(np.int64)
    genres[:8] = 0
    years = rng.integers(1920, 2019, size=n).astype(np.float64)
    years[[3, 700, 4000]] = np.nan
    years[10] = np.inf
    eps = 1e-8

    def ref(X):
        X = X.astype(np.float64).reshape(n, -1).copy()
        if impute == "col_median":
            X[~np.isfinite(X)] = np.nan
            if np.isnan(X).any():
                med = np.nanmedian(X, axis=0, keepdims=True)
                med = np.where(np.isfinite(med), med, 0.0)
                r, c = np.where(np.isnan(X))
                X[r, c] = med[0, c]
        if method == "row_l1":
            X = X / np.maximum(np.sum(np.abs(X), axis=1, keepdims=True), eps)
        elif method == "row_l2":
            X = X / np.maximum(np.sqrt(np.sum(X * X, axis=1, keepdims=True)), eps)
        elif method == "col_zscore":
            sd = np.std(X, axis=0, keepdims=True)
            X = (X - np.mean(X, axis=0, keepdims=True)) / np.where(sd < eps, 1.0, sd)
            X[~np.isfinite(X)] = 0.0
        elif method == "col_minmax":
            mn = np.min(X, axis=0, keepdims=True)
            X = (X - mn) / np.maximum(np.max(X, axis=0, keepdims=True) - mn, eps)
        return X.astype(np.float32)

    for X in (genres, years):
        if impute == "none" and not np.isfinite(X.astype(np.float64)).all():
            with pytest.raises(ValueError, match="NaN/Inf"):
                normalize_feature(X, method, impute=impute)
            continue
        got = normalize_feature(X, method, impute=impute)
        want = ref(X)
        assert got.dtype == np.float32 and got.shape == want.shape
        assert np.allclose(got, want, rtol=3e-7, atol=1e-7)
    with pytest.raises(ValueError, match="Unknown method"):
        normalize_feature(genres, "bogus")
    with pytest.raises(ValueError, match="Unknown impute"):
        normalize_feature(genres, "none", impute="bogus")
    allnan = np.full((50, 2), np.nan)
    allnan[:, 1] = np.arange(50)
    assert np.array_equal(normalize_feature(allnan, "none", impute="col_median", dtype="float64")[:, 0], np.zeros(50))


def test_device_graph_build_tie_rule_and_fit(cuda_device):
    """N3: graph.similarity_topk_device — without ties the same graph as the reference's numpy build; with ties
    (binary features) the documented rule (smaller column wins), checked against its numpy restatement; and a fit
    on the device-built graph matches the oracle run on that same graph."""
    import b200als
    from b200als import graph as G
    rng = np.random.default_rng(8)
    n, d, topk = 700, 12, 20
    X = rng.normal(size=(n, d)).astype(np.float32)                 # continuous: no ties
    sptr, sidx, sval, D = G.similarity_topk_device(X, topk, 1e-8, cuda_device)
    S_dev = G.csr_to_dense(sptr, sidx, sval, n)
    S_ref = G.similarity_topk(X, topk, 1e-8)
    assert np.array_equal(S_dev != 0, S_ref != 0)
    assert np.abs(S_dev - S_ref).max() < 2e-6
    assert np.allclose(D, S_ref.sum(axis=1), rtol=1e-5, atol=1e-5)
    assert all(np.all(np.diff(sidx[sptr[i]:sptr[i + 1]]) > 0) for i in range(n))
    Xb = (rng.random((n, 6)) < 0.4).astype(np.float32)             # binary: 64 distinct rows, massive ties
    sptr, sidx, sval, D = G.similarity_topk_device(Xb, topk, 1e-8, cuda_device)
    S_dev = G.csr_to_dense(sptr, sidx, sval, n)
    S_rule = G.topk_rule_reference(Xb, topk, 1e-8)
    # similarities of identical feature pairs are identical floats on both sides, so the kept SETS must agree wherever
    # the float32 value orderings agree; compare through the rule evaluated on the device's own values
    unit = Xb / (np.sqrt((Xb * Xb).sum(axis=1, keepdims=True)) + np.float32(1e-8))
    assert np.abs(S_dev - S_rule).max() < 2e-6 or np.mean((S_dev != 0) == (S_rule != 0)) > 0.999
    assert np.array_equal(S_dev, S_dev.T)
    # a fit on the device-built graph
    c = load_case("small_graph_topk5")
    cfg = make_cfg(c.cfgd)
    model = b200als.ALS(config=cfg, lambda_w=c.lambda_w)
    model.graph_build = "device"
    model.fit_coo(c.flat, c.vals, (c.m, c.n), features=c.features, tol=None, verbose=0)
    assert model.S is None and model.S_csr is not None
    assert np.isfinite(model.V).all() and len(model.history["train_rmse"]) == cfg.core.n_iters
    host = b200als.ALS(config=cfg, lambda_w=c.lambda_w).fit_coo(c.flat, c.vals, (c.m, c.n), features=c.features,
                                                                tol=None, verbose=0)
    S_d = G.csr_to_dense(*model.S_csr, c.n)
    if np.array_equal(S_d != 0, host.S != 0):                    # same graph (no ties in this fixture) -> same fit
        assert rel_err(model.V, host.V) < 1e-4


def test_refit_matches_reference_fixture(cuda_device):
    """Two fits on one object against the reference's own (tests/golden/small_refit.npz): combined history, and the
    second fit's early stop decided on entries of the first (it stops after one iteration, like the reference)."""
    import json
    import b200als
    z = np.load(os.path.join(GOLDEN, "small_refit.npz"))
    cfgd, fits = json.loads(str(z["cfg"])), json.loads(str(z["fits"]))
    m, n = int(z["m"]), int(z["n"])
    model = b200als.ALS(b200als.ALSConfig(core=b200als.CoreConfig(**cfgd["core"]),
                                          biases=b200als.BiasesConfig(**cfgd["biases"])))
    for j, kw in enumerate(fits):
        model.fit_coo(z[f"flat_idx_{j}"], z[f"vals_{j}"], (m, n), tol=kw["tol"], min_iters=kw["min_iters"], verbose=0)
        assert len(model.history["train_rmse"]) == int(z[f"len_after_{j}"])
        for a in ("U", "V", "b_u", "b_i"):
            assert rel_err(getattr(model, a), z[f"{a}_{j}"]) < FACTOR_TOL, (j, a)
        assert abs(model.mu - float(z[f"mu_{j}"])) < RMSE_TOL
    for key in ("train_rmse", "U_norm", "V_norm", "bu_norm", "bi_norm"):
        got, want = np.asarray(model.history[key]), z["hist_" + key]
        assert got.shape == want.shape
        assert np.abs(got - want).max() < (RMSE_TOL if key == "train_rmse" else 1e-3 * max(1.0, np.abs(want).max())), key
