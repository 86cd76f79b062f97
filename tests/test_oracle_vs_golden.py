"""CPU: the oracle (oracle/als_oracle.py) against fixtures produced by the unmodified
reference (tests/golden/make_golden.py).  Same numpy/scipy calls on the same operands →
bit-identical, including every per-iteration snapshot."""
import os

import numpy as np
import pytest

from _util import GOLDEN, SMALL_CASES, bundled_features, load_case, run_oracle


@pytest.mark.parametrize("name", SMALL_CASES)
def test_oracle_bit_identical_to_reference(name):
    c = load_case(name)
    o, snaps = run_oracle(c)
    z = c.z
    assert np.array_equal(o.U, z["U"]) and np.array_equal(o.V, z["V"])
    assert np.array_equal(o.b_u, z["b_u"]) and np.array_equal(o.b_i, z["b_i"])
    assert o.mu == float(z["mu"])
    for fn in c.feature_names:
        assert np.array_equal(o.W[fn], z["W_" + fn])
    for key in ("train_rmse", "U_norm", "V_norm", "bu_norm", "bi_norm"):
        assert np.array_equal(np.asarray(o.history[key]), z["hist_" + key]), key
    # per-iteration state (observed in the reference through the _converged hook)
    assert len(snaps) == z["snap_U"].shape[0]
    for it, s in enumerate(snaps):
        assert np.array_equal(s["U"], z["snap_U"][it]) and np.array_equal(s["V"], z["snap_V"][it])
        assert np.array_equal(s["b_u"], z["snap_b_u"][it]) and np.array_equal(s["b_i"], z["snap_b_i"][it])
        assert s["mu"] == float(z["snap_mu"][it])
        for fn in c.feature_names:
            assert np.array_equal(s["W"][fn], z["snap_W_" + fn][it])
    assert np.array_equal(o.predict(c.features), z["pred"])
    if "S" in z.files:
        assert np.array_equal(o.S, z["S"])
    else:
        assert o.S is None


def test_oracle_early_stop_iteration_matches():
    c = load_case("small_earlystop")
    o, _ = run_oracle(c)
    assert len(o.history["train_rmse"]) == c.z["hist_train_rmse"].shape[0] == 15
    assert np.array_equal(np.asarray(o.history["train_rmse"]), c.z["hist_train_rmse"])
    assert np.array_equal(o.U, c.z["U"])


def test_oracle_predict_pairs_matches_dense():
    c = load_case("small_feat2")
    o, _ = run_oracle(c)
    rows, cols = c.flat // c.n, c.flat % c.n
    dense = o.predict(c.features)
    assert np.allclose(o.predict_pairs(rows, cols, c.features), dense[rows, cols], rtol=0, atol=1e-12)


@pytest.mark.parametrize("name,tol", [("c1_k32", 5e-7), ("c2_full", 5e-7)])
def test_oracle_bundled_shape(name, tol):
    """610 x 4980 fixtures are stored as float32 to keep the repo small: compare at float32 ulp."""
    c = load_case(name)
    feats = bundled_features() if c.feature_names else {}
    o, _ = run_oracle(c, features=feats)
    z = c.z
    assert np.allclose(o.U, z["U"], rtol=tol, atol=1e-7) and np.allclose(o.V, z["V"], rtol=tol, atol=1e-7)
    assert np.allclose(o.b_u, z["b_u"], rtol=tol, atol=1e-7) and np.allclose(o.b_i, z["b_i"], rtol=tol, atol=1e-7)
    assert np.array_equal(np.asarray(o.history["train_rmse"]), z["hist_train_rmse"])
    for fn in c.feature_names:
        assert np.array_equal(o.W[fn], z["W_" + fn])
    rows, cols = z["pred_idx"] // c.n, z["pred_idx"] % c.n
    assert np.allclose(o.predict_pairs(rows, cols, feats), z["pred_val"], rtol=0, atol=1e-12)


def test_oracle_csr_csc_match_numpy_flatnonzero():
    """als.py:338-340 ordering on a dense matrix."""
    import als_oracle
    rng = np.random.default_rng(0)
    R = rng.random((37, 53))
    R[rng.random(R.shape) < 0.8] = np.nan
    R[5, :] = np.nan
    R[:, 7] = np.nan
    sp = als_oracle.csr_csc_from_dense(R)
    mask = ~np.isnan(R)
    for u in range(R.shape[0]):
        assert np.array_equal(sp["indices"][sp["indptr"][u]:sp["indptr"][u + 1]], np.flatnonzero(mask[u]))
        assert np.array_equal(sp["vals"][sp["indptr"][u]:sp["indptr"][u + 1]], R[u, mask[u]])
    for i in range(R.shape[1]):
        assert np.array_equal(sp["cindices"][sp["cindptr"][i]:sp["cindptr"][i + 1]], np.flatnonzero(mask[:, i]))
        assert np.array_equal(sp["cvals"][sp["cindptr"][i]:sp["cindptr"][i + 1]], R[mask[:, i], i])
    ru, ri = np.where(mask)
    assert np.array_equal(sp["rows_u"], ru) and np.array_equal(sp["indices"], ri)
    # perm maps CSC positions to CSR positions
    assert np.array_equal(sp["vals"][sp["perm"]], sp["cvals"])


def test_refit_on_the_same_object_matches_reference():
    """history is created once and every fit() appends (scripts/als.py:185-191, 510-517); `_converged` of a second fit
    looks at entries of the first — tests/golden/small_refit.npz (make_golden_refit.py) pins the reference's
    behaviour, including the second fit stopping after ONE iteration."""
    import json
    import als_oracle
    import b200als
    z = np.load(os.path.join(GOLDEN, "small_refit.npz"))
    cfgd, fits = json.loads(str(z["cfg"])), json.loads(str(z["fits"]))
    m, n = int(z["m"]), int(z["n"])
    cfg = b200als.ALSConfig(core=b200als.CoreConfig(**cfgd["core"]), biases=b200als.BiasesConfig(**cfgd["biases"]))
    o = als_oracle.SparseALSOracle(cfg)
    for j, kw in enumerate(fits):
        sp = als_oracle.csr_csc_from_flat(z[f"flat_idx_{j}"], z[f"vals_{j}"], m, n)
        o.fit_sparse(sp, m, n, tol=kw["tol"], min_iters=kw["min_iters"])
        assert len(o.history["train_rmse"]) == int(z[f"len_after_{j}"])
        assert np.array_equal(o.U, z[f"U_{j}"]) and np.array_equal(o.V, z[f"V_{j}"])
        assert np.array_equal(o.b_u, z[f"b_u_{j}"]) and np.array_equal(o.b_i, z[f"b_i_{j}"])
    assert int(z["len_after_1"]) == int(z["len_after_0"]) + 1          # the quirk the fixture exists for
    for key in o.history:
        assert np.array_equal(np.asarray(o.history[key]), z["hist_" + key]), key
