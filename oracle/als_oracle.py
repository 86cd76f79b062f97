"""
TEST INFRASTRUCTURE — CPU oracle for the ALS block-update hot path.

This file is a numpy/scipy restatement of the reference's `ALS.fit` loop
(`/root/reference/scripts/als.py:300-529`) that takes the rating matrix as
CSR + CSC arrays instead of a dense NaN-filled (m, n) float64 matrix, so it
can stand in for the reference at shapes whose dense matrix cannot be
allocated.  It makes the *same numpy / scipy calls on the
same operands in the same order* as the reference, row by row, so on inputs
the reference can hold it is bit-identical to it (pinned by
`tests/test_oracle_vs_golden.py` against fixtures produced by the unmodified
reference, see `tests/golden/make_golden.py`).

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline /
`--impl reference` legs may import this module.  The product path
(`collaborative-filtering_b200/`) never does.

Parity status: the reference ships no tests / golden vectors of its own
(SURVEY.md §4), so parity is pinned by outputs of the reference itself run in
the build container and committed under `tests/golden/`.
"""
from __future__ import annotations

import numpy as np
from scipy.linalg import cho_factor, cho_solve

SCALE_FACTOR = 0.1   # als.py:93
EPS = 1e-10          # als.py:94


def cholesky_solve(A, b):
    """helpers.py:5-20 — LAPACK potrf + potrs, check_finite=False."""
    L, lower = cho_factor(A, check_finite=False)
    return cho_solve((L, lower), b, check_finite=False)


# --------------------------------------------------------------------------
# Index construction (als.py:335-340) from flat indices
# --------------------------------------------------------------------------
def csr_csc_from_flat(flat_idx, vals, m, n):
    """Build CSR and CSC exactly as als.py:338-340 orders them.

    `flat_idx = u*n + i` (create_folds.py:73-84 convention), any order, unique.
    CSR: per user ascending item ids (== np.flatnonzero(mask[u]));
    CSC: per item ascending user ids (== np.flatnonzero(mask[:, i]));
    COO (rows_u, rows_i) == CSR order (np.where is row-major).
    Returns dict with indptr/indices/vals, cindptr/cindices/cvals, perm
    (perm[p_csc] = p_csr).
    """
    flat_idx = np.asarray(flat_idx, dtype=np.int64)
    vals = np.asarray(vals, dtype=np.float64)
    order = np.argsort(flat_idx, kind="stable")
    f = flat_idx[order]
    v = vals[order]
    rows = f // n
    cols = f - rows * n
    indptr = np.zeros(m + 1, dtype=np.int64)
    np.cumsum(np.bincount(rows, minlength=m), out=indptr[1:])
    # CSC: stable sort by column keeps ascending rows inside a column
    perm = np.argsort(cols, kind="stable").astype(np.int64)
    cindptr = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(np.bincount(cols, minlength=n), out=cindptr[1:])
    return dict(
        indptr=indptr, indices=cols.astype(np.int64), vals=v,
        cindptr=cindptr, cindices=rows[perm].astype(np.int64), cvals=v[perm],
        perm=perm, rows_u=rows.astype(np.int64),
    )


def csr_csc_from_dense(R):
    """als.py:335-340 on a dense matrix, returned in the same dict form."""
    mask = ~np.isnan(R)
    m, n = R.shape
    flat = np.flatnonzero(mask)
    return csr_csc_from_flat(flat, R.ravel()[flat], m, n)


# --------------------------------------------------------------------------
# One-shot setup pieces
# --------------------------------------------------------------------------
def build_item_similarity(X, topk, eps):
    """als.py:220-240 — cosine similarity, zero diagonal, per-row top-k via
    np.argpartition, symmetrised with max.  Same expression sequence so that
    tie-breaking inside argpartition is the reference's."""
    norms = np.sqrt((X * X).sum(axis=1, keepdims=True)) + eps
    Xn = X / norms
    S = Xn @ Xn.T
    np.fill_diagonal(S, 0.0)
    if topk is not None and topk < S.shape[0]:
        nn = S.shape[0]
        for i in range(nn):
            idx = np.argpartition(S[i], -topk)[:-topk]
            S[i, idx] = 0.0
    S = np.maximum(S, S.T)
    return S


def item_reg(counts, lambda_v, pop_reg_mode):
    """als.py:243-259."""
    if not pop_reg_mode:
        return np.full_like(counts, lambda_v, dtype=float)
    elif pop_reg_mode == "inverse_sqrt":
        return lambda_v / np.sqrt(counts + 1.0)
    raise ValueError(f"Unknown pop_reg_mode '{pop_reg_mode}'")


class SparseALSOracle:
    """Sparse-input restatement of `scripts.als.ALS` (als.py:104-574).

    `config` is duck-typed: anything with `.core/.biases/.graph` shaped like
    als_config.py:57-95 (the reference's or this repo's dataclasses).
    """

    def __init__(self, config, lambda_w=None):
        if config is None:
            raise ValueError("ALSConfig must be provided.")
        self.cfg = config
        self.W = {}
        self.lambda_w = dict(lambda_w or {})
        core = config.core
        self.n_factors = core.n_factors
        self.n_iters = core.n_iters
        self.lambda_u = core.lambda_u
        self.lambda_v = core.lambda_v
        self.random_state = core.random_state
        self.update_w_every = core.update_w_every
        self.pop_reg_mode = core.pop_reg_mode
        self.lambda_bu = config.biases.lambda_bu or self.lambda_u      # als.py:166
        self.lambda_bi = config.biases.lambda_bi or self.lambda_v      # als.py:167
        self.alpha = config.graph.alpha
        self.S_topk = config.graph.sim.topk if config.graph.sim is not None else None
        self.S_eps = config.graph.sim.eps if config.graph.sim is not None else EPS
        self.U = self.V = self.b_u = self.b_i = None
        self.mu = 0.0
        self.S = None
        self.history = {"train_rmse": [], "U_norm": [], "V_norm": [],
                        "bu_norm": [], "bi_norm": []}

    # als.py:262-281
    def _compose_Z(self, V, features):
        Z = V.copy()
        for name, X in features.items():
            W = self.W.get(name)
            if W is not None:
                Z += X @ W
        return Z

    # als.py:283-297
    def _converged(self, tol, window=2):
        h = self.history["train_rmse"]
        return len(h) >= window + 1 and (h[-window - 1] - h[-1]) <= tol

    def fit_sparse(self, sp, m, n, features=None, tol=1e-3, min_iters=5,
                   on_iter=None, init=None, mu0=None, max_rows=None):
        """als.py:300-529 over the dict returned by csr_csc_from_*.

        on_iter(it, self) is called after each iteration's history append.
        init: optional dict(U=, V=, W={}[, b_u=, b_i=]) to skip RNG draws (big shapes).
        max_rows: (CPU-baseline sampling only) int or (users, items): process only
        the first so-many user rows and item rows per sweep and skip the W / mu
        steps; `self.sampled_ratings` reports the ratings touched.  Never used
        for parity.
        """
        rng = np.random.default_rng(self.random_state)                 # als.py:329
        indptr, indices, vals = sp["indptr"], sp["indices"], sp["vals"]
        cindptr, cindices, cvals = sp["cindptr"], sp["cindices"], sp["cvals"]
        rows_u, rows_i = sp["rows_u"], indices                         # als.py:340
        features = features or {}
        for name, X in features.items():                               # als.py:346-351
            if X.shape[0] != n:
                raise ValueError(f"Feature '{name}' has {X.shape[0]} rows; "
                                 f"expected {n} (number of items).")
            if not np.isfinite(X).all():
                raise ValueError(f"Feature '{name}' contains infinite values.")

        use_graph = (self.alpha > 0.0) and (self.cfg.graph.sim is not None)   # als.py:354
        self.S = None
        if use_graph:
            X = features.get(self.cfg.graph.sim.feature_name)
            if X is not None:
                self.S = build_item_similarity(X, self.S_topk, self.S_eps)
        use_graph = bool(self.S is not None)
        D = self.S.sum(axis=1) if use_graph else None                  # als.py:357

        # als.py:360 — nanmean(R) == sum(observed)/count; identical bits when
        # the observed values sum exactly (half-star grid), else within 1 ulp.
        self.mu = float(vals.sum() / vals.size) if mu0 is None else float(mu0)
        self.b_u = np.zeros(m, dtype=float)
        self.b_i = np.zeros(n, dtype=float)
        k = self.n_factors
        I = np.eye(k)
        if init is None:
            self.U = rng.normal(scale=SCALE_FACTOR, size=(m, k))       # als.py:367
            self.V = rng.normal(scale=SCALE_FACTOR, size=(n, k))       # als.py:368
            for name, X in features.items():                           # als.py:371-376
                self.W[name] = rng.normal(scale=SCALE_FACTOR, size=(X.shape[1], k))
        else:
            self.U = np.array(init["U"], dtype=float)
            self.V = np.array(init["V"], dtype=float)
            self.W = {kk: np.array(vv, dtype=float) for kk, vv in init.get("W", {}).items()}
            # (tests at full size restart a half-step from a given state: biases are then given too)
            if init.get("b_u") is not None:
                self.b_u = np.array(init["b_u"], dtype=float)
            if init.get("b_i") is not None:
                self.b_i = np.array(init["b_i"], dtype=float)

        item_counts = np.diff(cindptr).astype(float)                   # als.py:379
        lambda_v_i = item_reg(item_counts, self.lambda_v, self.pop_reg_mode).astype(float)
        lambda_bi_i = np.full(n, float(self.lambda_bi), dtype=float)   # als.py:384

        if max_rows is None:
            mu_rows, mi_rows = m, n
        else:
            mr_u, mr_i = (max_rows, max_rows) if np.isscalar(max_rows) else max_rows
            mu_rows, mi_rows = min(m, int(mr_u)), min(n, int(mr_i))
        self.sampled_ratings = (int(indptr[mu_rows] - indptr[0]), int(cindptr[mi_rows] - cindptr[0]))

        for it in range(self.n_iters):                                 # als.py:408
            Z_full = self._compose_Z(self.V, features)                 # als.py:411
            # (1) users — als.py:414-433
            for u in range(mu_rows):
                lo, hi = indptr[u], indptr[u + 1]
                if hi == lo:
                    continue
                idx = indices[lo:hi]
                R_u = vals[lo:hi]
                Z_u = Z_full[idx]
                r_u = R_u - (self.mu + self.b_u[u] + self.b_i[idx])
                A = Z_u.T @ Z_u + (self.lambda_u + EPS) * I
                b = Z_u.T @ r_u
                self.U[u] = cholesky_solve(A, b)
                denom = idx.size + self.lambda_bu + EPS
                pred_wo_bu = (Z_u @ self.U[u]) + self.mu + self.b_i[idx]
                self.b_u[u] = float(np.sum(R_u - pred_wo_bu) / denom)
            # (2) items — als.py:436-466
            for i in range(mi_rows):
                lo, hi = cindptr[i], cindptr[i + 1]
                if hi == lo:
                    continue
                idx = cindices[lo:hi]
                R_i = cvals[lo:hi]
                U_i = self.U[idx]
                r = R_i - (self.mu + self.b_u[idx] + self.b_i[i])
                reg_i = lambda_v_i[i] + EPS
                if use_graph:
                    reg_i += float(self.alpha) * float(D[i])
                A = U_i.T @ U_i + reg_i * I
                b = U_i.T @ r
                if use_graph:
                    b += float(self.alpha) * (self.S[i] @ self.V)      # live V (Gauss-Seidel)
                self.V[i] = cholesky_solve(A, b)
                denom = idx.size + lambda_bi_i[i] + EPS
                pred_wo_bi = (U_i @ self.V[i]) + self.mu + self.b_u[idx]
                self.b_i[i] = float(np.sum(R_i - pred_wo_bi) / denom)
            if max_rows is not None:
                continue      # CPU-baseline sampling: only the two sweeps are timed
            # (3) W_f — als.py:468-501
            if features and ((it % self.update_w_every == 0) or (it == self.n_iters - 1)):
                r_obs = vals - (self.mu + self.b_u[rows_u] + self.b_i[rows_i])
                r_obs = r_obs - np.sum(self.U[rows_u] * self.V[rows_i], axis=1)
                residual = r_obs.copy()
                for name, X in features.items():
                    W = self.W.get(name)
                    if W is not None:
                        WX = X[rows_i] @ W
                        residual -= np.sum(self.U[rows_u] * WX, axis=1)
                for name, X in features.items():
                    d = X.shape[1]
                    if name in self.W:
                        residual_plus_self = residual.copy()
                        WX = X[rows_i] @ self.W[name]
                        residual_plus_self += np.sum(self.U[rows_u] * WX, axis=1)
                    else:
                        residual_plus_self = residual
                    U_obs = self.U[rows_u]
                    X_obs = X[rows_i]
                    X_design = (X_obs[:, :, None] * U_obs[:, None, :]).reshape(len(rows_u), d * k)
                    lam = float(self.lambda_w.get(name, 0.0))
                    A = X_design.T @ X_design + (lam + EPS) * np.eye(d * k)
                    b = X_design.T @ residual_plus_self
                    vec = cholesky_solve(A, b)
                    self.W[name] = vec.reshape(d, k)
            # (4) mu, history — als.py:503-517
            Z_full = self._compose_Z(self.V, features)
            pred_wo_mu = (np.sum(self.U[rows_u] * Z_full[rows_i], axis=1)
                          + self.b_u[rows_u] + self.b_i[rows_i])
            self.mu = float(np.mean(vals - pred_wo_mu))
            pred = pred_wo_mu + self.mu
            err = vals - pred
            rmse = float(np.sqrt(np.mean(err ** 2)))
            self.history["train_rmse"].append(rmse)
            self.history["U_norm"].append(float(np.linalg.norm(self.U)))
            self.history["V_norm"].append(float(np.linalg.norm(self.V)))
            self.history["bu_norm"].append(float(np.linalg.norm(self.b_u)))
            self.history["bi_norm"].append(float(np.linalg.norm(self.b_i)))
            if on_iter is not None:
                on_iter(it, self)
            if tol is not None and it + 1 >= min_iters and self._converged(tol):   # als.py:520
                break
        return self

    def fit(self, R, features=None, tol=1e-3, min_iters=5, on_iter=None):
        """Dense-input convenience wrapper with the reference's signature."""
        m, n = R.shape
        sp = csr_csc_from_dense(R)
        return self.fit_sparse(sp, m, n, features, tol, min_iters, on_iter,
                               mu0=float(np.nanmean(R)))

    # als.py:532-574
    def predict(self, features=None):
        if self.U is None or self.V is None:
            raise RuntimeError("Model must be fitted before prediction.")
        n = self.V.shape[0]
        features = features or {}
        for name, X in features.items():
            if X.shape[0] != n:
                raise ValueError(f"Feature '{name}' has {X.shape[0]} rows. "
                                 f"Expected number of rows: {n}.")
            if not np.isfinite(X).all():
                raise ValueError(f"Feature '{name}' contains infinite values.")
        Z = self.V.copy()
        for name, X in features.items():
            W = self.W.get(name)
            if W is not None:
                Z += X @ W
        return self.U @ Z.T + self.mu + self.b_u[:, None] + self.b_i[None, :]

    def predict_pairs(self, rows, cols, features=None):
        """Entries of predict() at (rows, cols) without the dense matrix."""
        Z = self._compose_Z(self.V, features or {})
        return (np.sum(self.U[rows] * Z[cols], axis=1) + self.mu
                + self.b_u[rows] + self.b_i[cols])
