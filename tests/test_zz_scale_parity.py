"""GPU, parity AT THE BENCHMARKED SCALE: one ALS iteration at BASELINE.json's C3 (138 493 x 26 744, 2e7 ratings,
k = 64) and C4 (480 189 x 17 770, 1e8 ratings, k = 128) sizes from a given start, compared row by row with the
float64 oracle (oracle/als_oracle.py, the restatement of scripts/als.py:414-466 pinned to the reference's fixtures)
on a ROW SAMPLE — the whole sweep is out of the oracle's reach, a sample of rows is not:

  * user half-step: every user row depends only on the start (V0, b_i = 0, mu0), so ~2 000 random user rows plus the
    32 longest are solved by the oracle and compared with the device's rows;
  * item half-step: the oracle gets the DEVICE's own U, b_u after the user half-step as its start and solves the 32
    heaviest item rows (up to ~4.8e5 ratings: a float64 Gram of ~8 GFLOP each) plus 200 random ones.

Bar: 1e-3 relative per row (north_star's factor tolerance), biases to 1e-3 of their scale.  The worst row's error is
reported against its rating count in gpurun_out/scale_parity_<shape>.json — the record that shows whether the error
grows with the length of the float32 accumulation chain in the tensor core.
"""
import json
import os
import warnings

import numpy as np
import pytest

from _util import REPO

pytestmark = pytest.mark.gpu

SHAPES = {
    # name: (m, n, nnz, k) — bench.py WORKLOADS
    "c3": (138_493, 26_744, 20_000_000, 64),
    "c4": (480_189, 17_770, 100_000_000, 128),
}
HYPER = dict(lambda_u=5.0, lambda_v=5.0, lambda_bu=3.0, lambda_bi=3.0)      # bench.py HYPER
ROW_TOL = 1e-3
N_USER_RANDOM, N_HEAVY, N_ITEM_RANDOM = 2000, 32, 200


def _gather_rows(indptr_h, idx_d, val_d, rows):
    """Host copies of the listed rows of a device CSR: (local indptr, indices int64, vals float64)."""
    import torch
    lo, hi = indptr_h[rows], indptr_h[rows + 1]
    lens = hi - lo
    loc = np.zeros(rows.size + 1, dtype=np.int64)
    np.cumsum(lens, out=loc[1:])
    pos = np.concatenate([np.arange(a, b, dtype=np.int64) for a, b in zip(lo, hi)]) if rows.size else np.zeros(0, np.int64)
    pos_d = torch.from_numpy(pos).to(idx_d.device)
    return loc, idx_d[pos_d].cpu().numpy().astype(np.int64), val_d[pos_d].double().cpu().numpy()


def _row_errors(got, want):
    num = np.linalg.norm(got - want, axis=1)
    den = np.maximum(np.linalg.norm(want, axis=1), 1e-30)
    return num / den


@pytest.mark.parametrize("shape", ["c3", "c4"])
def test_one_iteration_row_sample_vs_oracle(cuda_device, shape):
    import torch
    import als_oracle
    import b200als
    from b200als.sparse import ratings_from_flat
    from b200als.synth import synth_ratings_torch

    m, n, nnz, k = SHAPES[shape]
    dev = cuda_device
    rows, cols, vals = synth_ratings_torch(m, n, nnz, seed=1234, device=dev)
    flat = rows * n + cols
    del rows, cols
    rt = ratings_from_flat(flat, vals, (m, n), dev, sorted_input=True)
    del flat
    mu0 = float(vals.sum(dtype=torch.float64).item() / nnz)
    rng = np.random.default_rng(42)
    U0 = rng.normal(scale=0.1, size=(m, k)).astype(np.float32)
    V0 = rng.normal(scale=0.1, size=(n, k)).astype(np.float32)

    cfg = b200als.ALSConfig(core=b200als.CoreConfig(n_factors=k, n_iters=1, lambda_u=HYPER["lambda_u"],
                                                    lambda_v=HYPER["lambda_v"]),
                            biases=b200als.BiasesConfig(HYPER["lambda_bu"], HYPER["lambda_bi"]))
    model = b200als.ALS(cfg)
    model.fit_ratings(rt, tol=None, verbose=0, init=dict(U=U0, V=V0, W={}))
    assert model.last_paths["user"] == b200als._lib.PATH_UMMA and model.last_paths["item"] == b200als._lib.PATH_UMMA
    U1, V1, bu1, bi1 = model.U, model.V, model.b_u, model.b_i          # float64 copies of the device's float32 result

    indptr_h, cindptr_h = rt.indptr.cpu().numpy(), rt.cindptr.cpu().numpy()
    ucnt, icnt = np.diff(indptr_h), np.diff(cindptr_h)

    # ---- user rows: start state only -------------------------------------------------------------------------
    heavy_u = np.argsort(-ucnt, kind="stable")[:N_HEAVY]
    nonempty_u = np.flatnonzero(ucnt > 0)
    rand_u = rng.choice(nonempty_u, size=min(N_USER_RANDOM, nonempty_u.size), replace=False)
    su = np.unique(np.concatenate([heavy_u, rand_u]))
    loc, idx, vv = _gather_rows(indptr_h, rt.indices, rt.vals, su)
    sp = dict(indptr=loc, indices=idx, vals=vv, cindptr=np.zeros(n + 1, dtype=np.int64),
              cindices=np.zeros(0, dtype=np.int64), cvals=np.zeros(0), rows_u=np.repeat(np.arange(su.size), np.diff(loc)))
    o = als_oracle.SparseALSOracle(cfg)
    o.fit_sparse(sp, su.size, n, tol=None, init=dict(U=np.zeros((su.size, k)), V=V0.astype(np.float64)), mu0=mu0)
    eu = _row_errors(U1[su], o.U)
    ebu = np.abs(bu1[su] - o.b_u) / max(float(np.abs(o.b_u).max()), 1e-30)

    # ---- item rows: the device's own U, b_u as the start -------------------------------------------------------
    heavy_i = np.argsort(-icnt, kind="stable")[:N_HEAVY]
    nonempty_i = np.flatnonzero(icnt > 0)
    rand_i = rng.choice(nonempty_i, size=min(N_ITEM_RANDOM, nonempty_i.size), replace=False)
    si = np.unique(np.concatenate([heavy_i, rand_i]))
    cloc, cidx, cvv = _gather_rows(cindptr_h, rt.cindices, rt.cvals, si)
    spi = dict(indptr=np.zeros(m + 1, dtype=np.int64), indices=np.zeros(0, dtype=np.int64), vals=np.zeros(0),
               cindptr=cloc, cindices=cidx, cvals=cvv, rows_u=np.zeros(0, dtype=np.int64))
    oi = als_oracle.SparseALSOracle(cfg)
    with warnings.catch_warnings(), np.errstate(all="ignore"):
        warnings.simplefilter("ignore")                      # the mu / RMSE step of the sub-problem has no ratings
        oi.fit_sparse(spi, m, si.size, tol=None, init=dict(U=U1, V=np.zeros((si.size, k)), b_u=bu1), mu0=mu0)
    ei = _row_errors(V1[si], oi.V)
    ebi = np.abs(bi1[si] - oi.b_i) / max(float(np.abs(oi.b_i).max()), 1e-30)

    wu, wi = int(np.argmax(eu)), int(np.argmax(ei))
    order_i = np.argsort(-icnt[si], kind="stable")[:N_HEAVY]
    report = dict(
        shape=shape, users=m, items=n, ratings=nnz, k=k, tol=ROW_TOL,
        split_rows=model.split_rows, split_piece=model.split_piece,
        user_rows=int(su.size), user_worst=dict(row=int(su[wu]), ratings=int(ucnt[su[wu]]), rel_err=float(eu[wu])),
        user_median=float(np.median(eu)), user_bias_worst=float(ebu.max()),
        item_rows=int(si.size), item_worst=dict(row=int(si[wi]), ratings=int(icnt[si[wi]]), rel_err=float(ei[wi])),
        item_median=float(np.median(ei)), item_bias_worst=float(ebi.max()),
        heaviest_items=[dict(row=int(si[j]), ratings=int(icnt[si[j]]), rel_err=float(ei[j])) for j in order_i],
        heaviest_users=[dict(row=int(r), ratings=int(ucnt[r]), rel_err=float(eu[np.searchsorted(su, r)]))
                        for r in heavy_u[:8]],
    )
    out_dir = os.path.join(REPO, "gpurun_out")
    os.makedirs(out_dir, exist_ok=True)
    tag = os.environ.get("B200ALS_PARITY_TAG", "")
    with open(os.path.join(out_dir, f"scale_parity_{shape}{tag}.json"), "w") as f:
        json.dump(report, f, indent=1)
    print(json.dumps({key: report[key] for key in ("shape", "user_worst", "user_median", "item_worst", "item_median",
                                                    "user_bias_worst", "item_bias_worst")}))
    assert eu.max() < ROW_TOL, report["user_worst"]
    assert ei.max() < ROW_TOL, report["item_worst"]
    assert ebu.max() < ROW_TOL and ebi.max() < ROW_TOL
