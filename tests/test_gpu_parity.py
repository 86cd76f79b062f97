"""GPU: the CUDA path (through the C-ABI, via the drop-in ALS class) against the
reference's golden fixtures and against the oracle on the same seeded inputs.

Tolerances are BASELINE.json's: CSR / CSC / fold indices bit-exact; factors within
1e-3 relative (Frobenius, float32 device vs float64 reference) per iteration; RMSE
history and predictions within 1e-4; same early-stop iteration.
"""
import numpy as np
import pytest

from _util import (SMALL_CASES, bundled_features, dense_from_case, load_case, make_cfg, rel_err, run_oracle)

pytestmark = pytest.mark.gpu

FACTOR_TOL = 1e-3
RMSE_TOL = 1e-4


def fit_b200(c, features=None, n_iters=None, dense=True, **fit_kw):
    import b200als
    cfgd = c.cfgd if n_iters is None else dict(c.cfgd, core=dict(c.cfgd["core"], n_iters=n_iters))
    model = b200als.ALS(config=make_cfg(cfgd), lambda_w=c.lambda_w)
    feats = c.features if features is None else features
    kw = dict(tol=None, verbose=0)
    if not c.fit["spy"]:
        kw.update(tol=c.fit["tol"], min_iters=c.fit["min_iters"])
    kw.update(fit_kw)
    if dense:
        model.fit(dense_from_case(c), features=feats, **kw)
    else:
        model.fit_coo(c.flat, c.vals, (c.m, c.n), features=feats, **kw)
    return model


def check_against(model, ref, c, feature_names):
    assert rel_err(model.U, ref["U"]) < FACTOR_TOL
    assert rel_err(model.V, ref["V"]) < FACTOR_TOL
    assert rel_err(model.b_u, ref["b_u"]) < FACTOR_TOL
    assert rel_err(model.b_i, ref["b_i"]) < FACTOR_TOL
    assert abs(model.mu - float(ref["mu"])) < RMSE_TOL
    for fn in feature_names:
        assert rel_err(model.W[fn], ref["W_" + fn]) < FACTOR_TOL, fn


@pytest.mark.parametrize("name", SMALL_CASES)
def test_fit_matches_reference_golden(cuda_device, name):
    c = load_case(name)
    model = fit_b200(c)
    z = c.z
    check_against(model, z, c, c.feature_names)
    for key in ("train_rmse", "U_norm", "V_norm", "bu_norm", "bi_norm"):
        got, want = np.asarray(model.history[key]), z["hist_" + key]
        assert got.shape == want.shape
        tol = RMSE_TOL if key == "train_rmse" else 1e-3 * max(1.0, float(np.abs(want).max()))
        assert np.abs(got - want).max() < tol, key
    pred = model.predict(c.features)
    assert pred.dtype == np.float64 and pred.shape == (c.m, c.n)
    assert np.abs(pred - z["pred"]).max() < 2e-3
    obs = z["pred"].ravel()[c.flat]
    assert abs(np.sqrt(np.mean((pred.ravel()[c.flat] - c.vals) ** 2)) - np.sqrt(np.mean((obs - c.vals) ** 2))) < RMSE_TOL
    if "S" in z.files:
        assert np.array_equal(model.S, z["S"])


@pytest.mark.parametrize("name", ["small_biases", "small_feat2", "small_full", "small_graph_topk40"])
def test_per_iteration_trajectory(cuda_device, name):
    """State after t iterations for every t (the reference's own snapshots)."""
    c = load_case(name)
    z = c.z
    T = z["snap_U"].shape[0]
    for t in range(1, T + 1):
        if c.feature_names and t != T:
            continue    # `it == n_iters - 1` (als.py:468) makes shorter runs a different schedule
        model = fit_b200(c, n_iters=t)
        assert rel_err(model.U, z["snap_U"][t - 1]) < FACTOR_TOL
        assert rel_err(model.V, z["snap_V"][t - 1]) < FACTOR_TOL
        assert rel_err(model.b_u, z["snap_b_u"][t - 1]) < FACTOR_TOL
        assert rel_err(model.b_i, z["snap_b_i"][t - 1]) < FACTOR_TOL
        assert abs(model.mu - float(z["snap_mu"][t - 1])) < RMSE_TOL


@pytest.mark.parametrize("name", ["small_plain", "small_empty", "small_feat2", "small_full"])
def test_coo_entry_equals_dense_entry(cuda_device, name):
    c = load_case(name)
    a = fit_b200(c, dense=True)
    rng = np.random.default_rng(3)
    p = rng.permutation(c.flat.size)       # unsorted COO input
    import b200als
    b = b200als.ALS(config=make_cfg(c.cfgd), lambda_w=c.lambda_w)
    b.fit_coo(c.flat[p], c.vals[p], (c.m, c.n), features=c.features, tol=None, verbose=0)
    assert np.array_equal(a.U, b.U) and np.array_equal(a.V, b.V)
    assert np.array_equal(a.b_u, b.b_u) and np.array_equal(a.b_i, b.b_i)


def test_early_stop_same_iteration(cuda_device):
    c = load_case("small_earlystop")
    model = fit_b200(c)
    want = c.z["hist_train_rmse"]
    assert len(model.history["train_rmse"]) == want.shape[0]
    assert np.abs(np.asarray(model.history["train_rmse"]) - want).max() < RMSE_TOL
    check_against(model, c.z, c, [])


@pytest.mark.parametrize("name", ["c1_k32", "c2_full"])
def test_bundled_shape_vs_reference(cuda_device, name):
    """C1 / C2: 610 x 4980, 80 k train ratings, real genres / years."""
    c = load_case(name)
    feats = bundled_features() if c.feature_names else {}
    model = fit_b200(c, features=feats, dense=(name == "c1_k32"))
    z = c.z
    check_against(model, z, c, c.feature_names)
    assert np.abs(np.asarray(model.history["train_rmse"]) - z["hist_train_rmse"]).max() < RMSE_TOL
    rows, cols = z["pred_idx"] // c.n, z["pred_idx"] % c.n
    pp = model.predict_pairs(rows, cols, feats)
    assert np.abs(pp - z["pred_val"]).max() < 2e-3
    # held-out fold: test RMSE within 1e-4 of the reference's
    f = np.load(__import__("os").path.join(__import__("_util").GOLDEN, "bundled_fold0.npz"))
    vi, vv = f["val_idx"].astype(np.int64), f["val_vals"].astype(np.float64) / 2.0
    mine = model.rmse_on_flat(vi, vv, feats)
    o, _ = run_oracle(c, features=feats)
    theirs = float(np.sqrt(np.mean((vv - o.predict_pairs(vi // c.n, vi % c.n, feats)) ** 2)))
    assert abs(mine - theirs) < RMSE_TOL


@pytest.mark.parametrize("k", [8, 32, 50, 64, 96, 128, 150])
def test_factor_sizes_vs_oracle(cuda_device, k):
    """k sweep over every FFMA tile configuration, checked against the oracle."""
    import b200als
    from b200als.synth import synth_ratings
    import als_oracle
    m, n, nnz = 300, 200, 9000
    flat, vals = synth_ratings(m, n, nnz, seed=k)
    cfg = b200als.ALSConfig(core=b200als.CoreConfig(n_factors=k, n_iters=3, lambda_u=4.0, lambda_v=3.0),
                            biases=b200als.BiasesConfig(lambda_bu=2.0, lambda_bi=2.5))
    o = als_oracle.SparseALSOracle(cfg).fit_sparse(als_oracle.csr_csc_from_flat(flat, vals, m, n), m, n, tol=None)
    for path in (b200als._lib.PATH_FFMA, b200als._lib.PATH_AUTO):
        model = b200als.ALS(cfg)
        model.gram_path = path
        model.fit_coo(flat, vals, (m, n), tol=None, verbose=0, sorted_input=True)
        assert rel_err(model.U, o.U) < FACTOR_TOL and rel_err(model.V, o.V) < FACTOR_TOL
        assert rel_err(model.b_u, o.b_u) < FACTOR_TOL and rel_err(model.b_i, o.b_i) < FACTOR_TOL
        assert np.abs(np.asarray(model.history["train_rmse"]) - np.asarray(o.history["train_rmse"])).max() < RMSE_TOL


def test_heavy_rows_vs_oracle(cuda_device):
    """A few very long rows (multi-chunk path) next to many short ones."""
    import b200als
    import als_oracle
    rng = np.random.default_rng(5)
    m, n = 4000, 50
    mask = rng.random((m, n)) < 0.02
    mask[:, 0] = True
    mask[:, 1] = rng.random(m) < 0.7
    mask[7, :] = True
    flat = np.flatnonzero(mask)
    vals = rng.integers(1, 11, size=flat.size) / 2.0
    cfg = b200als.ALSConfig(core=b200als.CoreConfig(n_factors=16, n_iters=2, lambda_u=3.0, lambda_v=3.0))
    o = als_oracle.SparseALSOracle(cfg).fit_sparse(als_oracle.csr_csc_from_flat(flat, vals, m, n), m, n, tol=None)
    model = b200als.ALS(cfg).fit_coo(flat, vals, (m, n), tol=None, verbose=0)
    assert rel_err(model.U, o.U) < FACTOR_TOL and rel_err(model.V, o.V) < FACTOR_TOL
    assert np.abs(np.asarray(model.history["train_rmse"]) - np.asarray(o.history["train_rmse"])).max() < RMSE_TOL


@pytest.mark.parametrize("k", [64, 128])
@pytest.mark.parametrize("split", [(64, 32), (200, 96), (1000, 0), (-1, 0)])
def test_split_rows_vs_oracle(cuda_device, k, split):
    """tcgen05 path: rows longer than the threshold are cut into pieces accumulated by different SMs
    (partial normal matrices added in global memory, the last piece solves).  Includes rows whose pieces
    pair up with whole rows / other pieces in one k=64 accumulator pair, ragged tails, and the odd tail."""
    import b200als
    import als_oracle
    rng = np.random.default_rng(11 + k)
    m, n = 1500, 61
    mask = rng.random((m, n)) < 0.03
    mask[:, 0] = True                                   # 1500 ratings
    mask[:, 1] = rng.random(m) < 0.7
    mask[:, 2] = rng.random(m) < 0.31
    mask[:, 3] = rng.random(m) < 0.1
    mask[7, :] = True                                   # one 61-rating user (just under / over the thresholds)
    mask[11, :40] = True
    flat = np.flatnonzero(mask)
    vals = rng.integers(1, 11, size=flat.size) / 2.0
    cfg = b200als.ALSConfig(core=b200als.CoreConfig(n_factors=k, n_iters=3, lambda_u=3.0, lambda_v=2.0),
                            biases=b200als.BiasesConfig(lambda_bu=2.0, lambda_bi=2.5))
    o = als_oracle.SparseALSOracle(cfg).fit_sparse(als_oracle.csr_csc_from_flat(flat, vals, m, n), m, n, tol=None)
    model = b200als.ALS(cfg)
    model.split_rows, model.split_piece = split
    model.fit_coo(flat, vals, (m, n), tol=None, verbose=0)
    assert rel_err(model.U, o.U) < FACTOR_TOL and rel_err(model.V, o.V) < FACTOR_TOL
    assert rel_err(model.b_u, o.b_u) < FACTOR_TOL and rel_err(model.b_i, o.b_i) < FACTOR_TOL
    assert np.abs(np.asarray(model.history["train_rmse"]) - np.asarray(o.history["train_rmse"])).max() < RMSE_TOL


def test_errors_match_reference_surface(cuda_device):
    import b200als
    c = load_case("small_feat1")
    with pytest.raises(ValueError):
        b200als.ALS(config=None)
    model = b200als.ALS(config=make_cfg(c.cfgd), lambda_w=c.lambda_w)
    with pytest.raises(RuntimeError):
        model.predict()
    R = dense_from_case(c)
    with pytest.raises(ValueError, match="rows"):
        model.fit(R, features={"genres": c.features["genres"][:-1]}, verbose=0)
    bad = c.features["genres"].copy()
    bad[0, 0] = np.nan
    with pytest.raises(ValueError, match="infinite"):
        model.fit(R, features={"genres": bad}, verbose=0)
    cfgd = dict(c.cfgd, core=dict(c.cfgd["core"], pop_reg_mode="bogus"))
    with pytest.raises(ValueError, match="pop_reg_mode"):
        b200als.ALS(config=make_cfg(cfgd)).fit(R, verbose=0)
