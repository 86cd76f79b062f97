"""ALS matrix factorisation with side features on B200 — host orchestration.

Same public surface as the reference model (`scripts/als.py:104-574`):

    model = ALS(config=ALSConfig(...), lambda_w={"genres": 5.0})
    model.fit(R, features={...}, tol=1e-3, min_iters=5, verbose=1)   # -> self
    R_hat = model.predict(features={...})                            # (m, n) float64

plus sparse-first entry points for shapes whose dense matrix cannot exist:
`fit_coo(flat_idx, vals, shape, ...)` and `predict_pairs(rows, cols, ...)`.

The model is  R ≈ U (V + Σ_f X_f W_f)ᵀ + μ + b_u + b_i  (als.py:10).  Python
only sequences the work; every array operation of the fit loop
(als.py:408-523) runs in libb200als kernels on device tensors that torch merely
owns.  There is no CPU fallback: without the library or a B200 this raises.
"""
from __future__ import annotations

import ctypes as C
import logging
import os
import time

import numpy as np
import torch

from . import _lib, graph as _graph, hostrng as _hostrng, readback as _readback
from ._lib import StepArgs, check, ptr
from .config import ALSConfig
from .dist import Shard
from .sparse import Ratings, ratings_from_dense, ratings_from_flat, row_schedule, upload_flat_sharded

SCALE_FACTOR = 0.1      # als.py:93 — std of the Gaussian factor init
EPS = 1e-10             # als.py:94 — added to every ridge term

logger = logging.getLogger(__name__)


def _round4(k: int) -> int:
    return (k + 3) & ~3


def factor_ld(k: int, gram_path: int) -> int:
    """Row stride (floats) of the device factor matrices.  The tcgen05 half-step exists for widths 64 and 128
    and covers every k in (32, 128] by zero padding (include/b200als.h): rows are then exactly 64 / 128 floats
    with zeros behind column k.  Other k (<= 32: a 64-wide tensor-core tile would be mostly padding; 129..160)
    take the FFMA path with 16-byte aligned rows."""
    if gram_path != _lib.PATH_FFMA:
        if 32 < k <= 64 or (gram_path == _lib.PATH_UMMA and k <= 64):
            return 64
        if 64 < k <= 128:
            return 128
    return _round4(k)


# Ridge below which a row with fewer than PRECISE_COUNT_FACTOR * k ratings goes to the float64 half-step, per Gram
# path.  Measured on B200 against the float64 oracle on rows of ~20 ratings (tests/test_gpu_edges.py): the relative
# factor error grows like lambda^-1.5 from 1.3e-6 (FFMA) / 2e-5 (tcgen05, bf16 hi/lo operands) at lambda = 1 and
# crosses the 1e-3 bar near 0.03 / 0.1; rows with many more ratings than factors have a full-rank Gram whose smallest
# eigenvalue, not the ridge, bounds the condition number, and stay on the fast paths whatever lambda is.
PRECISE_RIDGE = {_lib.PATH_FFMA: 0.1, _lib.PATH_UMMA: 0.3}
PRECISE_COUNT_FACTOR = 4


def draw_initial(seed, m, n, k, feature_dims, pin: bool):
    """als.py:329, 367-376: default_rng(seed).normal(scale=0.1, ...) for U, V (float32 copies) and every W_f (float64),
    the same values, drawn on all host cores (hostrng.py).  Large draws land in pinned host memory (torch caches the
    buffers between fits), so the upload that follows is a DMA.  Returns (arrays, pinned tensors to keep alive)."""
    pinned = []

    def alloc(shape, dtype):
        if int(np.prod(shape)) < (1 << 20) or not pin:
            return np.empty(shape, dtype=dtype)
        t = torch.empty(tuple(shape), dtype=torch.float32 if np.dtype(dtype) == np.float32 else torch.float64,
                        pin_memory=True)
        pinned.append(t)
        return t.numpy()
    draws = _hostrng.normal_stream(seed, [(m, k), (n, k)] + [(d, k) for d in feature_dims],
                                   [np.float32, np.float32] + [np.float64] * len(feature_dims), SCALE_FACTOR,
                                   n_threads=_hostrng.host_threads(shared=False), alloc=alloc)
    return draws, pinned


def _env_int(name: str, default: int) -> int:
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


class ALS:
    """Drop-in for `scripts.als.ALS` (constructor contract: als.py:134-191)."""

    def __init__(self, config: ALSConfig, lambda_w: dict[str, float] | None = None) -> None:
        if config is None:
            raise ValueError("ALSConfig must be provided.")
        self.cfg = config
        self.W: dict[str, np.ndarray] = {}
        self.lambda_w: dict[str, float] = dict(lambda_w or {})

        core = self.cfg.core
        self.n_factors = core.n_factors
        self.n_iters = core.n_iters
        self.lambda_u = core.lambda_u
        self.lambda_v = core.lambda_v
        self.random_state = core.random_state
        self.update_w_every = core.update_w_every
        self.pop_reg_mode = core.pop_reg_mode
        # falsy (None or 0.0) bias penalties fall back to the factor penalties (als.py:166-167)
        self.lambda_bu = self.cfg.biases.lambda_bu or self.lambda_u
        self.lambda_bi = self.cfg.biases.lambda_bi or self.lambda_v
        self.alpha = self.cfg.graph.alpha
        sim = self.cfg.graph.sim
        self.S_topk = sim.topk if sim is not None else None
        self.S_eps = sim.eps if sim is not None else EPS

        # U, V, b_u, b_i (als.py:176-179) are properties: float64 numpy arrays like the reference's, copied from
        # the device on first read after a fit (see `_host` / `_dev_model` below)
        self._host: dict[str, np.ndarray | None] = {"U": None, "V": None, "b_u": None, "b_i": None}
        self._dev_model: dict | None = None    # device-resident float32 model of the last fit
        self.mu: float = 0.0
        self.S: np.ndarray | None = None
        self.history: dict[str, list[float]] = {
            "train_rmse": [], "U_norm": [], "V_norm": [], "bu_norm": [], "bi_norm": []}

        # build-side knobs (not part of the reference surface)
        self.device = None                 # torch device; default: current CUDA device
        self.gram_path = _lib.PATH_AUTO    # PATH_FFMA / PATH_UMMA force one Gram path
        # tcgen05 path: rows longer than split_rows are cut into pieces of split_piece ratings (0: the library's
        # absolute defaults, include/b200als.h; < 0: never)
        self.split_rows = _env_int("B200ALS_SPLIT_ROWS", 0)
        self.split_piece = _env_int("B200ALS_SPLIT_PIECE", 0)
        # "auto": rows whose ridge is too small for float32 (few ratings AND small lambda, see PRECISE_RIDGE) take
        # the float64 half-step, the rest the fast paths; "f32": never; "f64": every row (the reference's arithmetic)
        self.precision = os.environ.get("B200ALS_PRECISION", "auto")
        # "host": the similarity graph is built with the reference's own numpy calls (index parity with np.argpartition's
        # tie-breaking); "device": cosine + top-k on the GPU with a documented tie rule (smaller column wins) — the same
        # graph whenever similarities do not tie, a different but deterministic one when they do (binary genres)
        self.graph_build = os.environ.get("B200ALS_GRAPH_BUILD", "host")
        self.shard: Shard | None = None    # row sharding across ranks (dist.Shard)
        self.timings: dict[str, float] = {}
        self.launches_per_iter: int = 0
        self.last_paths: dict[str, int] = {}   # Gram path the last iteration's half-steps took (PATH_FFMA / PATH_UMMA)

    # ---- learned parameters: float64 numpy on the host, fetched lazily -------------------------------
    # The fit leaves the model in HBM.  predict / predict_pairs / sse_on_flat read it there, and the float64
    # host copies the reference exposes (als.py:404-407) are made when an attribute is first read: a driver that
    # only calls predict() after fit() (tune_params.py:376-392, evaluate_models.py:246-265) never pays for them.
    # Once an attribute has been read (the caller may edit the array in place) or assigned, the host arrays are
    # authoritative again and predictions upload them, as before.
    def _get_param(self, name: str):
        v = self._host[name]
        if v is None and self._dev_model is not None:
            d = self._dev_model
            t = {"U": d["U"][:d["m"], :d["k"]], "V": d["V"][:d["n"], :d["k"]],
                 "b_u": d["b_u"][:d["m"]], "b_i": d["b_i"][:d["n"]]}[name]
            with torch.cuda.device(t.device):
                v = _readback.f64_from_device(t)
            self._host[name] = v
            d["host_touched"] = True
        return v

    def _set_param(self, name: str, value) -> None:
        self._host[name] = value
        if self._dev_model is not None:
            self._dev_model["host_touched"] = True

    U = property(lambda self: self._get_param("U"), lambda self, v: self._set_param("U", v))
    V = property(lambda self: self._get_param("V"), lambda self, v: self._set_param("V", v))
    b_u = property(lambda self: self._get_param("b_u"), lambda self, v: self._set_param("b_u", v))
    b_i = property(lambda self: self._get_param("b_i"), lambda self, v: self._set_param("b_i", v))

    @property
    def fitted(self) -> bool:
        return self._dev_model is not None or (self._host["U"] is not None and self._host["V"] is not None)

    # ------------------------------------------------------------------ helpers
    def _item_reg(self, counts: np.ndarray) -> np.ndarray:
        """λ_{v,i}: uniform, or λ_v / sqrt(c_i + 1) (als.py:243-259)."""
        if not self.pop_reg_mode:
            return np.full_like(counts, self.lambda_v, dtype=float)
        if self.pop_reg_mode == "inverse_sqrt":
            return self.lambda_v / np.sqrt(counts + 1.0)
        raise ValueError(f"Unknown pop_reg_mode '{self.pop_reg_mode}'")

    def _converged(self, tol: float, window: int = 2) -> bool:
        """RMSE gain over the last `window` iterations ≤ tol (als.py:283-297)."""
        h = self.history["train_rmse"]
        return len(h) >= window + 1 and (h[-window - 1] - h[-1]) <= tol

    @staticmethod
    def _check_features(features, n, where):
        for name, X in features.items():
            if X.shape[0] != n:
                if where == "fit":
                    raise ValueError(f"Feature '{name}' has {X.shape[0]} rows; "
                                     f"expected {n} (number of items).")
                raise ValueError(f"Feature '{name}' has {X.shape[0]} rows. "
                                 f"Expected number of rows: {n}.")
            if not np.isfinite(X).all():
                raise ValueError(f"Feature '{name}' contains infinite values.")

    def _n_items(self) -> int:
        if self._host["V"] is not None:
            return int(self._host["V"].shape[0])
        return int(self._dev_model["n"])

    def _dev(self):
        if self.device is not None:
            return torch.device(self.device)
        if not torch.cuda.is_available():
            raise _lib.B200ALSError("CUDA is not available: the ALS path has no CPU fallback")
        return torch.device("cuda", torch.cuda.current_device())

    # ------------------------------------------------------------------ fit
    def fit(self, R: np.ndarray, features: dict[str, np.ndarray] | None = None,
            tol: float = 1e-3, min_iters: int = 5, verbose: int = 1) -> "ALS":
        """Fit on a dense (m, n) matrix with NaN = missing (als.py:300-529)."""
        _lib.load()
        _lib.require_device()
        dev = self._dev()
        with torch.cuda.device(dev):
            ratings = ratings_from_dense(np.asarray(R), dev)
            mu0 = float(np.nanmean(R))                                  # als.py:360
            return self._fit_ratings(ratings, mu0, features, tol, min_iters, verbose, None)

    def fit_coo(self, flat_idx, vals, shape, features=None, tol=1e-3, min_iters=5, verbose=1,
                sorted_input=False, init=None) -> "ALS":
        """Fit on observed entries `R.ravel()[flat_idx] = vals`, flat_idx = u*n + i
        (the fold convention of create_folds.py:73-84, 200-206) — the same model as
        `fit(matrix_from_indices(shape, flat_idx, vals))` without the dense matrix.

        init: None → reference RNG stream on the host (als.py:329, 367-376);
              "device" → N(0, 0.1²) drawn on the GPU (benchmarks at shapes where the
              host draw would dominate; not seed-compatible with the reference);
              dict(U=, V=, W={}) → given arrays.
        """
        _lib.load()
        _lib.require_device()
        dev = self._dev()
        with torch.cuda.device(dev):
            self._start_predraw(shape, features, init)
            try:
                sh = self.shard
                on_dev = isinstance(flat_idx, torch.Tensor) and flat_idx.is_cuda
                if sh is not None and sh.active and not on_dev and os.environ.get("B200ALS_SHARDED_UPLOAD", "1") != "0":
                    # every rank holds the same host arrays: each copies 1/N over PCIe, NVLink does the rest
                    flat_idx, vals = upload_flat_sharded(flat_idx, vals, dev, sh)
                ratings = ratings_from_flat(flat_idx, vals, shape, dev, sorted_input=sorted_input)
            except Exception:
                self.__dict__.pop("_predraw", None)
                raise
            if ratings.nnz == 0:
                raise ValueError("no observed ratings")
            mu0 = float(ratings.vals.sum(dtype=torch.float64).item() / ratings.nnz)
            return self._fit_ratings(ratings, mu0, features, tol, min_iters, verbose, init)

    def _start_predraw(self, shape, features, init) -> None:
        """The initial draw is host work on all cores (~65 ms at the Netflix shape), the H2D copy and the CSR/CSC build
        are device work: start the draw on a thread now, the session picks it up (FitSession.__init__)."""
        import threading
        shard = self.shard
        if init is not None or (shard is not None and shard.active and shard.rank != 0):
            return
        m, n, k = int(shape[0]), int(shape[1]), int(self.n_factors)
        feats = features or {}
        if m * k < (1 << 22) or not (1 <= k <= _lib.MAX_K) or any(getattr(X, "ndim", 0) != 2 for X in feats.values()):
            return                                     # small draws are not worth a thread
        names = tuple(feats.keys())
        dims = [int(feats[nm].shape[1]) for nm in names]
        holder = {}
        seed = self.random_state

        dev = self._dev()

        def work():
            try:
                torch.cuda.set_device(dev)             # the pinned allocations belong to this fit's device context
                holder["draws"], holder["pinned"] = draw_initial(seed, m, n, k, dims, True)
            except Exception as e:   # noqa: BLE001 — re-raised by the session
                holder["error"] = e
        th = threading.Thread(target=work, daemon=True)
        th.start()
        self._predraw = (th, holder, (m, n, k, names))

    def fit_ratings(self, ratings: Ratings, features=None, tol=1e-3, min_iters=5, verbose=1,
                    init=None) -> "ALS":
        """Fit on an already-built device CSR/CSC (`sparse.Ratings`)."""
        _lib.load()
        _lib.require_device()
        with torch.cuda.device(ratings.vals.device):
            mu0 = float(ratings.vals.sum(dtype=torch.float64).item() / max(ratings.nnz, 1))
            return self._fit_ratings(ratings, mu0, features, tol, min_iters, verbose, init)

    def fit_slabs(self, slabs, tol=1e-3, min_iters=5, verbose=1, init=None) -> "ALS":
        """Sharded fit on slab-only storage (`slab.RatingSlabs`: this rank's CSR rows and CSC columns, nothing else;
        U/V + biases model).  Every rank of the job calls it with its own slabs and `model.shard` set."""
        _lib.load()
        _lib.require_device()
        with torch.cuda.device(slabs.vals.device):
            mu0 = float(slabs.vals_sum / max(slabs.nnz, 1))                 # als.py:360 over ALL ratings
            return self._fit_ratings(slabs, mu0, None, tol, min_iters, verbose, init)

    # ------------------------------------------------------------------ core loop
    def session(self, rt: Ratings, mu0: float, features=None, init=None, profile: bool = False) -> "FitSession":
        """Set up device state for a fit and return the stepping handle (bench / tests)."""
        return FitSession(self, rt, mu0, features, init, profile)

    def _fit_ratings(self, rt: Ratings, mu0, features, tol, min_iters, verbose, init) -> "ALS":
        if verbose == 0:
            logging.disable(logging.CRITICAL)          # als.py:325-326 (global, as in the reference)
        # history is created once per object and every fit() appends to it (als.py:185-191, 510-517): a second fit
        # on the same object continues the lists, and _converged() looks at the combined list
        self._hist_prefix = {key: list(v) for key, v in self.history.items()}
        ses = FitSession(self, rt, mu0, features, init, False)
        n_iters = int(self.n_iters)
        if verbose > 0:
            logger.info(f"Starting ALS training: n_factors={ses.k}, n_iters={n_iters}, "
                        f"lambda_u={self.lambda_u}, lambda_v={self.lambda_v}, "
                        f"pop_reg_mode={self.pop_reg_mode}, features={ses.names}, "
                        f"lambda_w={self.lambda_w}, random_state={self.random_state}, "
                        f"graph_alpha={self.alpha}, graph_topk={self.S_topk}, "
                        f"graph_eps={self.S_eps}, update_w_every={self.update_w_every}")
        done = 0
        for it in range(n_iters):                                       # als.py:408
            ses.iteration(it)
            done = it + 1
            # early stopping needs this iteration's RMSE on the host (als.py:520-523)
            if tol is not None and it + 1 >= min_iters:
                self._set_history(ses.read_history(done))
                if self._converged(tol):
                    if verbose > 0:
                        logger.info(f"Early stopping at iter {it + 1}; ΔRMSE ≤ {tol:.3g}")
                    break
        ses.finish(done)
        if verbose > 0 and self.history["train_rmse"]:
            logger.info(f"ALS training finished. Final train RMSE: {self.history['train_rmse'][-1]:.4f}")
        return self

    def _set_history(self, h: np.ndarray) -> None:
        keys = ("train_rmse", "U_norm", "V_norm", "bu_norm", "bi_norm")
        prefix = getattr(self, "_hist_prefix", None) or {}
        for j, key in enumerate(keys):
            self.history[key] = list(prefix.get(key, [])) + [float(x) for x in h[:, j]]

    # ------------------------------------------------------------------ predict
    def _resident_model(self, dev):
        """The device-resident model of the last fit, if it still is what the public attributes say."""
        d = self._dev_model
        if d is None or d.get("host_touched") or d["U"].device != dev:
            return None
        if float(self.mu) != d["mu_value"] or set(self.W) != set(d["W"]):
            return None
        if any(self.W[nm] is not d["W_host"][nm] for nm in d["W"]):
            return None
        return d

    def _upload_model(self, features, dev):
        """Device float32 model for predict / predict_pairs / sse_on_flat: the arrays the fit left in HBM, or —
        when the caller has read or assigned U, V, b_u, b_i, mu, W since — an upload of the public attributes."""
        lib = _lib.load()
        f32 = torch.float32
        res = self._resident_model(dev)
        if res is not None:
            m, n, k, ld = res["m"], res["n"], res["k"], res["ld"]
            U_d, V_d, b_u_d, b_i_d, mu_d = res["U"], res["V"], res["b_u"], res["b_i"], res["mu"]
        else:
            U, V = self.U, self.V
            m, k = U.shape
            n = V.shape[0]
            ld = _round4(k)
            U_d = torch.zeros(m, ld, dtype=f32, device=dev)
            V_d = torch.zeros(n, ld, dtype=f32, device=dev)
            U_d[:, :k] = torch.from_numpy(np.asarray(U, dtype=np.float32)).to(dev)
            V_d[:, :k] = torch.from_numpy(np.asarray(V, dtype=np.float32)).to(dev)
            b_u_d = torch.from_numpy(np.asarray(self.b_u, dtype=np.float32)).to(dev)
            b_i_d = torch.from_numpy(np.asarray(self.b_i, dtype=np.float32)).to(dev)
            mu_d = torch.full((1,), float(self.mu), dtype=f32, device=dev)
        used = [(nm, X) for nm, X in features.items() if self.W.get(nm) is not None]   # als.py:569-572
        if len(used) > _lib.MAX_FEATURES:
            raise ValueError(f"at most {_lib.MAX_FEATURES} feature matrices are supported")
        if used:
            X_d = [torch.from_numpy(np.ascontiguousarray(X, dtype=np.float32)).to(dev) for _, X in used]
            W_d = []
            for nm, X in used:
                if res is not None:
                    W_d.append(res["W"][nm])
                    continue
                w = torch.zeros(X.shape[1], ld, dtype=f32, device=dev)
                w[:, :k] = torch.from_numpy(np.asarray(self.W[nm], dtype=np.float32)).to(dev)
                W_d.append(w)
            xptr = (C.c_void_p * len(used))(*[x.data_ptr() for x in X_d])
            wptr = (C.c_void_p * len(used))(*[w.data_ptr() for w in W_d])
            dims = (C.c_int32 * len(used))(*[int(x.shape[1]) for x in X_d])
            Z_d = torch.empty(V_d.shape[0], ld, dtype=f32, device=dev)
            check(lib.b200als_compose_z(ptr(V_d), ld, n, k, len(used), xptr, dims, wptr, ld, -1, ptr(Z_d), ld,
                                        torch.cuda.current_stream().cuda_stream), "compose_z")
            torch.cuda.current_stream().synchronize()
        else:
            Z_d = V_d
        return dict(U=U_d, Z=Z_d, b_u=b_u_d, b_i=b_i_d, mu=mu_d, ld=ld, m=m, n=n, k=k)

    def predict(self, features: dict[str, np.ndarray] | None = None) -> np.ndarray:
        """Completed (m, n) float64 matrix  U Zᵀ + μ + b_u + b_i  (als.py:532-574)."""
        if not self.fitted:
            raise RuntimeError("Model must be fitted before prediction.")
        features = features or {}
        self._check_features(features, self._n_items(), "predict")      # als.py:558-565
        lib = _lib.load()
        _lib.require_device()
        dev = self._dev()
        with torch.cuda.device(dev):
            s = self._upload_model(features, dev)
            out = torch.empty(s["m"], s["n"], dtype=torch.float64, device=dev)
            check(lib.b200als_predict_dense(ptr(s["U"]), s["ld"], ptr(s["Z"]), s["ld"], ptr(s["b_u"]), ptr(s["b_i"]),
                                            ptr(s["mu"]), s["m"], s["n"], s["k"], ptr(out),
                                            torch.cuda.current_stream().cuda_stream), "predict_dense")
            return out.cpu().numpy()

    def predict_pairs(self, rows, cols, features: dict[str, np.ndarray] | None = None) -> np.ndarray:
        """Entries of `predict(features)` at (rows, cols) — `R_pred.ravel()[u*n+i]` of
        tune_params.py:165-167 without the dense matrix.  Returns float64."""
        if not self.fitted:
            raise RuntimeError("Model must be fitted before prediction.")
        features = features or {}
        self._check_features(features, self._n_items(), "predict")
        lib = _lib.load()
        _lib.require_device()
        dev = self._dev()
        with torch.cuda.device(dev):
            s = self._upload_model(features, dev)
            r = torch.as_tensor(np.asarray(rows)).to(device=dev, dtype=torch.int32).contiguous()
            c = torch.as_tensor(np.asarray(cols)).to(device=dev, dtype=torch.int32).contiguous()
            if r.numel() != c.numel():
                raise ValueError("rows and cols differ in length")
            out = torch.empty(max(r.numel(), 1), dtype=torch.float32, device=dev)
            check(lib.b200als_predict_pairs(ptr(s["U"]), s["ld"], ptr(s["Z"]), s["ld"], ptr(s["b_u"]), ptr(s["b_i"]),
                                            ptr(s["mu"]), ptr(r), ptr(c), r.numel(), s["k"], ptr(out),
                                            torch.cuda.current_stream().cuda_stream), "predict_pairs")
            return out[:r.numel()].double().cpu().numpy()

    def sse_on_flat(self, flat_idx, y_true, features=None, item_bin=None, n_bins: int = 1):
        """Squared-error sums and counts over flat indices u*n+i, split by `item_bin[i]` when given, computed on
        the device without the dense (m, n) prediction (b200als_sse_on_flat).  `flat_idx` / `y_true` may be numpy
        arrays or tensors already on the device.  Returns (sse float64[n_bins], cnt int64[n_bins]) as numpy."""
        if not self.fitted:
            raise RuntimeError("Model must be fitted before prediction.")
        features = features or {}
        self._check_features(features, self._n_items(), "predict")
        if not (1 <= int(n_bins) <= 32):
            raise ValueError("n_bins must be in [1, 32]")
        lib = _lib.load()
        _lib.require_device()
        dev = self._dev()
        with torch.cuda.device(dev):
            s = self._upload_model(features, dev)
            f = torch.as_tensor(flat_idx).to(device=dev, dtype=torch.int64).contiguous()
            y = torch.as_tensor(y_true).to(device=dev, dtype=torch.float32).contiguous()
            if f.numel() != y.numel():
                raise ValueError("flat_idx and y_true differ in length")
            if f.numel() and (int(f.min()) < 0 or int(f.max()) >= s["m"] * s["n"]):
                raise IndexError("flat index outside the (m, n) matrix")
            b = None
            if item_bin is not None:
                b = torch.as_tensor(np.asarray(item_bin)).to(device=dev, dtype=torch.int32).contiguous()
                if b.numel() != s["n"]:
                    raise ValueError("item_bin must have one entry per item")
            sse = torch.zeros(n_bins, dtype=torch.float64, device=dev)
            cnt = torch.zeros(n_bins, dtype=torch.int64, device=dev)
            check(lib.b200als_sse_on_flat(ptr(s["U"]), s["ld"], ptr(s["Z"]), s["ld"], ptr(s["b_u"]), ptr(s["b_i"]),
                                          ptr(s["mu"]), ptr(f), ptr(y), f.numel(), s["n"], s["k"], ptr(b),
                                          int(n_bins), ptr(sse), ptr(cnt), torch.cuda.current_stream().cuda_stream),
                  "sse_on_flat")
            return sse.cpu().numpy(), cnt.cpu().numpy()

    def rmse_on_flat(self, flat_idx, y_true, features=None) -> float:
        """RMSE over flat indices u*n+i (`_rmse_on_indices`, tune_params.py:147-167); NaN for an empty set."""
        sse, cnt = self.sse_on_flat(flat_idx, y_true, features)
        return float(np.sqrt(sse[0] / cnt[0])) if cnt[0] else float("nan")

    def rmse_by_item_bin(self, flat_idx, y_true, item_bin, n_bins: int, features=None):
        """(overall RMSE, [RMSE of bin 0, ..., bin n_bins-1]) in one pass — the overall and `rmse_pop_b` fold
        metrics of evaluate_models.py:263-274 (`_split_val_indices_by_popularity` + `_rmse_on_indices`)."""
        sse, cnt = self.sse_on_flat(flat_idx, y_true, features, item_bin, n_bins)
        with np.errstate(invalid="ignore", divide="ignore"):
            per_bin = [float(np.sqrt(sse[b] / cnt[b])) if cnt[b] else float("nan") for b in range(n_bins)]
        tot = int(cnt.sum())
        return (float(np.sqrt(sse.sum() / tot)) if tot else float("nan")), per_bin


class FitSession:
    """Device state of one fit plus `iteration(it)`: one pass of als.py:408-517.

    Everything is enqueued on torch's current CUDA stream and is asynchronous;
    `read_history` / `finish` are the only host synchronisation points.
    """

    PHASES = ("compose", "user", "item", "wstep", "stats")

    def __init__(self, model: ALS, rt: Ratings, mu0: float, features, init, profile: bool):
        self.model = model
        self.rt = rt
        self.lib = lib = _lib.load()
        self.profile = profile
        self.phase_ms = {p: [] for p in self.PHASES}
        self._events = []
        dev = self.dev = rt.vals.device
        m, n, nnz = rt.m, rt.n, rt.nnz
        k = self.k = int(model.n_factors)
        if not (1 <= k <= _lib.MAX_K):
            raise ValueError(f"n_factors={k} outside the supported range [1, {_lib.MAX_K}]")
        if nnz == 0:
            raise ValueError("no observed ratings")
        shard = self.shard = model.shard or Shard()
        features = features or {}
        model._check_features(features, n, "fit")                       # als.py:346-351
        if len(features) > _lib.MAX_FEATURES:
            raise ValueError(f"at most {_lib.MAX_FEATURES} feature matrices are supported")
        names = self.names = list(features.keys())
        ld = self.ld = factor_ld(k, model.gram_path)
        model._dev_model = None                    # a new fit replaces the resident model
        model._host = {"U": None, "V": None, "b_u": None, "b_i": None}
        f32, f64 = torch.float32, torch.float64

        # ---- graph (host build for argpartition tie parity, als.py:354-357) ----
        use_graph = (model.alpha > 0.0) and (model.cfg.graph.sim is not None)
        model.S = None
        if use_graph:
            fname = model.cfg.graph.sim.feature_name
            Xg = features.get(fname)
            if Xg is None:
                logger.warning(f"GraphSim feature '{fname}' not found in features dict. "
                               f"Graph regularization disabled.")
            else:
                t0 = time.perf_counter()
                self.S_csr = None
                if (model.graph_build == "device" and model.S_topk is not None and 1 <= model.S_topk < n):
                    # opt-in: cosine + top-k on the GPU with a documented tie rule (graph.similarity_topk_device)
                    sp_, si_, sv_, D_ = _graph.similarity_topk_device(Xg, model.S_topk, model.S_eps, dev)
                    self.S_csr = (sp_, si_, sv_, D_)
                    model.S_csr = (sp_, si_, sv_)      # the graph as CSR; the dense `S` of the reference is not formed
                                                       # (graph.csr_to_dense(*model.S_csr, n) gives it)
                elif model.graph_build not in ("host", "device"):
                    raise ValueError(f"graph_build={model.graph_build!r}: expected 'host' or 'device'")
                else:
                    # the reference's own numpy calls (tie parity); the same graph is reused across the folds of a trial
                    model.S, csr_, D_ = _graph.similarity_topk_cached(Xg, model.S_topk, model.S_eps)
                    self.S_host = (csr_, D_)
                model.timings["S_build_s"] = time.perf_counter() - t0
        use_graph = self.use_graph = (model.S is not None) or (getattr(self, "S_csr", None) is not None)
        if use_graph and getattr(self, "S_csr", None) is not None:
            D = self.S_csr[3]
        elif use_graph and getattr(self, "S_host", None) is not None:
            D = self.S_host[1]                                          # als.py:357 (S's dtype), cached with S
        else:
            D = model.S.sum(axis=1) if use_graph else None

        # ---- init (als.py:360-376): same PCG64 stream, order U, V, W_f ----------
        m_pad, n_pad = shard.padded_rows(m), shard.padded_rows(n)
        # U, V, b_u, b_i live in ONE allocation (views below): under the fused exchange a rank then maps one block per
        # peer instead of four (28 -> 7 handle opens per fit on an 8-GPU box)
        n_u, n_v = m_pad * ld, n_pad * ld
        self._rep = torch.zeros(n_u + n_v + m_pad + n_pad, dtype=f32, device=dev)
        self._rep_off = dict(U=0, V=n_u, b_u=n_u + n_v, b_i=n_u + n_v + m_pad)         # in floats
        U_d = self.U_d = self._rep[:n_u].view(m_pad, ld)
        V_d = self.V_d = self._rep[n_u:n_u + n_v].view(n_pad, ld)
        W_d = self.W_d = [torch.zeros(features[nm].shape[1], ld, dtype=f32, device=dev) for nm in names]
        W64_d = self.W64_d = [torch.zeros(features[nm].shape[1] * k, dtype=f64, device=dev) for nm in names]
        if init is None:
            # als.py:329, 367-376: default_rng(random_state).normal(scale=0.1, ...) — the same values, drawn on
            # all host cores (hostrng.py).  Sharded: one rank draws with every core and the device arrays are
            # broadcast, instead of N ranks walking the same stream on 1/N of the cores each.
            if shard.rank == 0 or not shard.active:
                pre = model.__dict__.pop("_predraw", None)
                if pre is not None and pre[2] == (m, n, k, tuple(names)):
                    pre[0].join()                      # started by fit_coo under the H2D copy and the CSR/CSC build
                    if "error" in pre[1]:
                        raise pre[1]["error"]
                    draws, pinned = pre[1]["draws"], pre[1]["pinned"]
                else:
                    draws, pinned = draw_initial(model.random_state, m, n, k, [features[nm].shape[1] for nm in names],
                                                 dev.type == "cuda")
                U_d[:m, :k].copy_(torch.from_numpy(draws[0]), non_blocking=True)
                V_d[:n, :k].copy_(torch.from_numpy(draws[1]), non_blocking=True)
                torch.cuda.current_stream().synchronize() if dev.type == "cuda" else None
                del pinned
                for f, nm in enumerate(names):
                    w0 = draws[2 + f]
                    W64_d[f].copy_(torch.from_numpy(w0.reshape(-1)).to(dev))
                    W_d[f][:, :k] = torch.from_numpy(w0.astype(np.float32)).to(dev)
            for t in [U_d, V_d] + W_d + W64_d:
                shard.broadcast_from_first(t)
        elif isinstance(init, str) and init == "device":
            g = torch.Generator(device=dev)
            g.manual_seed(int(model.random_state))
            U_d[:m, :k].normal_(0.0, SCALE_FACTOR, generator=g)
            V_d[:n, :k].normal_(0.0, SCALE_FACTOR, generator=g)
            for f in range(len(names)):
                W_d[f][:, :k].normal_(0.0, SCALE_FACTOR, generator=g)
                W64_d[f].copy_(W_d[f][:, :k].reshape(-1).double())
        else:
            U_d[:m, :k] = torch.as_tensor(np.asarray(init["U"], dtype=np.float32)).to(dev)
            V_d[:n, :k] = torch.as_tensor(np.asarray(init["V"], dtype=np.float32)).to(dev)
            for f, nm in enumerate(names):
                w0 = np.asarray(init["W"][nm], dtype=np.float64)
                W64_d[f].copy_(torch.from_numpy(w0.reshape(-1)).to(dev))
                W_d[f][:, :k] = torch.from_numpy(w0.astype(np.float32)).to(dev)
        self.b_u_d = self._rep[n_u + n_v:n_u + n_v + m_pad]
        self.b_i_d = self._rep[n_u + n_v + m_pad:]
        self.mu_d = torch.full((1,), float(mu0), dtype=f32, device=dev)
        model.mu = float(mu0)
        self.X_d = [torch.from_numpy(np.ascontiguousarray(features[nm], dtype=np.float32)).to(dev) for nm in names]
        self.Z_d = torch.zeros(n_pad, ld, dtype=f32, device=dev) if names else V_d

        # ---- regularisers (als.py:379-384, 426, 450-454), folded in float64 ------
        slab = self.slab = hasattr(rt, "row_ids_u")       # slab.RatingSlabs: local row pointers, global ids
        if slab:
            if not shard.active and rt.world != 1:
                raise ValueError("RatingSlabs of a multi-rank job need model.shard to be the job's Shard")
            if (rt.rank, rt.world) != (shard.rank, shard.world):
                raise ValueError(f"slabs were built for rank {rt.rank}/{rt.world}, the fit runs as {shard.rank}/{shard.world}")
            if names or use_graph:
                raise ValueError("slab storage covers the U/V + biases model (the W step and the graph sweep are not sharded)")
            cnt_d = torch.zeros(n, dtype=torch.int64, device=dev)
            cnt_d[rt.row_ids_i.long()] = rt.cindptr[1:] - rt.cindptr[:-1]
            shard.all_reduce_sum(cnt_d)
            item_counts = cnt_d.cpu().numpy().astype(float)
        else:
            item_counts = (rt.cindptr[1:] - rt.cindptr[:-1]).cpu().numpy().astype(float)
        lambda_v_i = model._item_reg(item_counts).astype(float)
        reg_item = lambda_v_i + EPS
        if use_graph:
            reg_item = reg_item + float(model.alpha) * D.astype(np.float64)
        self.reg_item_d = torch.from_numpy(reg_item.astype(np.float32)).to(dev)
        self.reg_user = float(np.float32(model.lambda_u + EPS))
        self.lam_bu = float(np.float32(model.lambda_bu + EPS))
        self.lam_bi = float(np.float32(float(model.lambda_bi) + EPS))

        # ---- schedules: longest rows first, dealt round-robin over the ranks (dist.py) ----
        u_order, u_nonempty = row_schedule(rt.indptr)
        i_order, i_nonempty = row_schedule(rt.cindptr)
        self.ulo, self.uhi = shard.my_range(m)          # contiguous split, residual reduction only
        self.ilo, self.ihi = shard.my_range(n)
        self.u_ids = rt.row_ids_u if slab else None     # local row -> id in U / b_u (slab storage)
        self.i_ids = rt.row_ids_i if slab else None
        if slab:
            # the slab IS this rank's share: its non-empty rows, longest first (local ids)
            self.u_list, self.i_list = u_order[:u_nonempty], i_order[:i_nonempty]
            if shard.active:
                self.u_plan = shard.plan_exchange_ids(rt.row_ids_u.long()[self.u_list.long()], m)
                self.i_plan = shard.plan_exchange_ids(rt.row_ids_i.long()[self.i_list.long()], n)
        elif shard.active:
            uo, io = u_order[:u_nonempty], i_order[:i_nonempty]
            self.u_list, self.i_list = shard.deal(uo), shard.deal(io)
            self.u_plan = shard.plan_exchange(uo, m)    # row m / n of the replicas is the dummy row
            self.i_plan = shard.plan_exchange(io, n)
        else:
            self.u_list, self.i_list = u_order[:u_nonempty], i_order[:i_nonempty]
        if use_graph:
            self.i_list = i_order[:i_nonempty]      # the Gauss–Seidel sweep is not sharded (§8e)

        # fused exchange: the other ranks' replicas mapped here; the half-steps store solved rows into all of them
        self.peers = None
        if shard.active and not use_graph:
            shared = shard.share_replicas(dict(rep=self._rep))
            if shared is not None:
                self.peers = {nm: [base + 4 * off for base in shared["ptrs"]["rep"]]
                              for nm, off in self._rep_off.items()}
                self._peer_handles = shared["handles"]  # released by close()
                self._token = torch.zeros(1, dtype=torch.float32, device=dev)
        # rows that need float64 (small ridge, few ratings) leave the fast lists
        self.u_list, self.u_f64 = self._precision_split(rt.indptr, self.u_list, self.reg_user, None, self.u_ids, False)
        self.i_list, self.i_f64 = self._precision_split(rt.cindptr, self.i_list, None, self.reg_item_d, self.i_ids,
                                                        use_graph)
        self.u_split = self._split_plan(rt.indptr, self.u_list)
        self.i_split = None if use_graph else self._split_plan(rt.cindptr, self.i_list)

        need = C.c_size_t(0)
        check(lib.b200als_step_workspace_bytes(k, model.gram_path, max(m, n), C.byref(need)), "step_workspace_bytes")
        self.ws = torch.empty(need.value, dtype=torch.uint8, device=dev)
        self.status = torch.zeros(_lib.STATUS_WORDS, dtype=torch.int32, device=dev)
        self.n_iters = int(model.n_iters)
        self.acc = torch.zeros(6, dtype=f64, device=dev)
        self.hist = torch.zeros(max(self.n_iters, 1), 6, dtype=f64, device=dev)

        if use_graph:
            t0 = time.perf_counter()
            if getattr(self, "S_csr", None) is not None:
                sptr, sidx, sval = self.S_csr[:3]
            elif getattr(self, "S_host", None) is not None:
                sptr, sidx, sval = self.S_host[0]
            else:
                sptr, sidx, sval = _graph.to_csr(model.S)
            self.level_ptr, level_rows = _graph.level_schedule(sptr, sidx, item_counts > 0)
            model.timings["S_csr_levels_s"] = time.perf_counter() - t0
            self.sptr_d = torch.from_numpy(sptr).to(dev)
            self.sidx_d = torch.from_numpy(sidx).to(dev)
            self.sval_d = torch.from_numpy(sval).to(dev)
            self.level_rows_d = torch.from_numpy(level_rows).to(dev)
            self.level_ptr_d = torch.from_numpy(np.ascontiguousarray(self.level_ptr, dtype=np.int64)).to(dev)
            self.max_level = int(np.diff(self.level_ptr).max()) if self.level_ptr.size > 1 else 0
            self.sweep_fused = os.environ.get("B200ALS_GRAPH_SWEEP", "fused") == "fused"
            self.chol_d = torch.empty(n * k * k, dtype=f32, device=dev)
            self.aux_d = torch.empty(n * (2 * k + 2), dtype=f32, device=dev)
        else:
            self.chol_d = self.aux_d = None
        if names:
            self.G_d = torch.empty(n * k * k, dtype=f64, device=dev)
            self.cvec_d = torch.empty(n * k, dtype=f64, device=dev)
            self.Zmf_d = torch.empty(n_pad, ld, dtype=f32, device=dev)
            dmax = max(int(x.shape[1]) for x in self.X_d)
            self.A_d = torch.empty((dmax * k) ** 2, dtype=f64, device=dev)
            self.Wnew64 = [torch.empty_like(w) for w in W64_d]
            self.xptr = (C.c_void_p * len(names))(*[x.data_ptr() for x in self.X_d])
            self.wptr = (C.c_void_p * len(names))(*[w.data_ptr() for w in W_d])
            self.dims = (C.c_int32 * len(names))(*[int(x.shape[1]) for x in self.X_d])
        self.launches = 0
        # A10 fused into the steps (residual sums by the item step, norms by both): possible when Z == V and no
        # Gauss-Seidel sweep follows, i.e. the item step sees every residual of the iteration's final state.
        # Rows without ratings are never solved: their constant share of the norms is added up front (by rank 0).
        self.fused_stats = (not names) and (not use_graph)
        if self.fused_stats:
            base = torch.zeros(6, dtype=f64, device=dev)
            if slab:
                # every rank adds the share of the rows IT owns and nobody rated
                ue = rt.row_ids_u.long()[u_order[u_nonempty:].long()]
                ie = rt.row_ids_i.long()[i_order[i_nonempty:].long()]
            else:
                ue, ie = u_order[u_nonempty:].long(), i_order[i_nonempty:].long()
            if slab or shard.rank == 0 or not shard.active:
                base[2] = U_d.index_select(0, ue).double().pow(2).sum()
                base[3] = V_d.index_select(0, ie).double().pow(2).sum()
                base[4] = self.b_u_d.index_select(0, ue).double().pow(2).sum()
                base[5] = self.b_i_d.index_select(0, ie).double().pow(2).sum()
            self.acc_base = base
            self.acc.copy_(base)        # every later iteration: b200als_finalize_iter_fused puts the base back

    # ---- small wrappers ---------------------------------------------------------
    def _stream(self):
        return torch.cuda.current_stream().cuda_stream

    def _compose(self, out, skip=-1):
        check(self.lib.b200als_compose_z(ptr(self.V_d), self.ld, self.rt.n, self.k, len(self.names), self.xptr,
                                         self.dims, self.wptr, self.ld, skip, ptr(out), self.ld, self._stream()),
              "compose_z")
        self.launches += 1

    def _fast_path(self, factor_mode: bool) -> int:
        """Gram path the library's AUTO rule / the caller's choice gives the fast list (include/b200als.h)."""
        gp = self.model.gram_path
        if factor_mode or gp == _lib.PATH_FFMA or self.ld not in (64, 128):
            return _lib.PATH_FFMA
        return _lib.PATH_UMMA if (gp == _lib.PATH_UMMA or 2 * self.k > self.ld) else _lib.PATH_FFMA

    def _precision_split(self, indptr, rows, reg, reg_row, ids, factor_mode):
        """(fast list, float64 list or None) of a half-step's schedule — see PRECISE_RIDGE."""
        prec = self.model.precision
        if prec not in ("auto", "f32", "f64"):
            raise ValueError(f"precision={prec!r}: expected 'auto', 'f32' or 'f64'")
        if prec == "f32" or rows.numel() == 0:
            return rows, None
        if prec == "f64":
            return rows[:0].contiguous(), rows
        thr = PRECISE_RIDGE[self._fast_path(factor_mode)]
        r64 = rows.to(torch.int64)
        few = (indptr[r64 + 1] - indptr[r64]) < PRECISE_COUNT_FACTOR * self.k
        if reg_row is None:
            if not reg < thr:
                return rows, None
            mask = few
        else:
            g = ids.to(torch.int64)[r64] if ids is not None else r64
            mask = few & (reg_row[g] < thr)
        if not bool(mask.any()):
            return rows, None
        return rows[~mask].contiguous(), rows[mask].contiguous()

    def _split_plan(self, indptr, rows):
        """Long rows of a schedule cut into pieces for the tcgen05 path (b200als_split_plan_build); None if
        the path is not in use or no row can exceed the threshold."""
        model, lib = self.model, self.lib
        if model.gram_path == _lib.PATH_FFMA or self.ld not in (64, 128) or model.split_rows < 0 or rows.numel() == 0:
            return None
        if model.gram_path == _lib.PATH_AUTO and 2 * self.k <= self.ld:
            return None
        r64 = rows.to(torch.int64)
        cnt = indptr[r64 + 1] - indptr[r64]
        thresh = model.split_rows or lib.b200als_split_default(0)
        piece = model.split_piece or (lib.b200als_split_default(1) if not model.split_rows else 0)
        nnz_list, n_long = (int(x) for x in torch.stack([cnt.sum(), (cnt > thresh).sum()]).tolist())
        if n_long == 0:
            return None
        need = C.c_size_t(0)
        check(lib.b200als_split_plan_bytes(self.k, rows.numel(), nnz_list, thresh, piece, n_long, C.byref(need)),
              "split_plan_bytes")
        mem = torch.empty(max(need.value, 256), dtype=torch.uint8, device=self.dev)
        plan = _lib.SplitPlan()
        check(lib.b200als_split_plan_build(ptr(indptr), ptr(rows), rows.numel(), nnz_list, self.k, thresh, piece,
                                           n_long, ptr(mem), mem.numel(), C.byref(plan), self._stream()),
              "split_plan_build")
        if not plan.items:
            return None
        plan._mem = mem                     # keeps the device memory alive
        return plan

    def _step(self, indptr, indices, vals, n_rows, rows, F, n_f_rows, bias_other, reg, reg_row, lamb, X, bias_own,
              mode, stats=None, split=None, norms=None, peers=None, row_ids=None, path=None):
        a = StepArgs()
        a.row_ids = ptr(row_ids)
        a.struct_bytes = C.sizeof(StepArgs)
        a.k, a.mode, a.path = self.k, mode, (self.model.gram_path if path is None else path)
        a.n_rows = n_rows
        a.indptr, a.indices, a.vals = ptr(indptr), ptr(indices), ptr(vals)
        a.rows, a.n_list = ptr(rows), rows.numel()
        a.F, a.ldf, a.n_f_rows = ptr(F), self.ld, n_f_rows
        a.bias_other, a.mu = ptr(bias_other), ptr(self.mu_d)
        a.reg, a.reg_row = reg, ptr(reg_row)
        a.lam_bias, a.lam_bias_row = lamb, None
        a.X, a.ldx = ptr(X), self.ld
        a.bias_own = ptr(bias_own)
        a.chol_out = ptr(self.chol_d) if mode == _lib.STEP_FACTOR else None
        a.aux_out = ptr(self.aux_d) if mode == _lib.STEP_FACTOR else None
        a.status = ptr(self.status)
        a.stats_acc = ptr(stats)
        a.norm_acc = norms
        a.split = C.addressof(split) if split is not None else None
        a.workspace, a.workspace_bytes = ptr(self.ws), self.ws.numel()
        if peers is not None:
            a.n_peers = len(peers[0])
            for j, (px, pb) in enumerate(zip(*peers)):
                a.peer_X[j], a.peer_bias[j] = px, pb
        check(self.lib.b200als_row_step(C.byref(a), self._stream()), "row_step")
        self.launches += self.lib.b200als_last_launch_count() + 1      # + the cursor memset
        return self.lib.b200als_last_path()

    def _mark(self, phase):
        """Profile mode: CUDA events on the launching stream around each phase."""
        if self.profile:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            self._events.append((phase, ev))

    def collect_phase_times(self):
        """Resolve recorded events into per-phase milliseconds (call after a sync)."""
        for (p0, e0), (p1, e1) in zip(self._events[:-1], self._events[1:]):
            if p0 != "end":
                self.phase_ms[p0].append(e0.elapsed_time(e1))
        self._events = []
        return self.phase_ms

    # ---- one ALS iteration (als.py:408-517) ------------------------------------------
    def iteration(self, it: int) -> None:
        lib, rt, model, shard = self.lib, self.rt, self.model, self.shard
        m, n, k, ld = rt.m, rt.n, self.k, self.ld
        names = self.names
        stream = self._stream()
        self.launches = 0
        fused = self.fused_stats
        self._mark("compose")
        if names:
            self._compose(self.Z_d)                                     # als.py:411
        # (1) users (als.py:414-433)
        self._mark("user")
        peers = self.peers
        model.last_paths["user"] = self._step(
            rt.indptr, rt.indices, rt.vals, rt.indptr.numel() - 1, self.u_list, self.Z_d, n, self.b_i_d, self.reg_user, None,
            self.lam_bu, self.U_d, self.b_u_d, _lib.STEP_SOLVE, split=self.u_split,
            norms=(self.acc.data_ptr() + 16) if fused else None,
            peers=(peers["U"], peers["b_u"]) if peers else None, row_ids=self.u_ids)
        if self.u_f64 is not None:                      # the rows float32 cannot resolve: same update in float64
            p64 = self._step(
                rt.indptr, rt.indices, rt.vals, rt.indptr.numel() - 1, self.u_f64, self.Z_d, n, self.b_i_d,
                self.reg_user, None, self.lam_bu, self.U_d, self.b_u_d, _lib.STEP_SOLVE,
                norms=(self.acc.data_ptr() + 16) if fused else None,
                peers=(peers["U"], peers["b_u"]) if peers else None, row_ids=self.u_ids, path=_lib.PATH_F64)
            if self.u_list.numel() == 0:
                model.last_paths["user"] = p64
        if shard.active:
            if peers:
                shard.order_ranks(self._token)      # every rank's rows of U, b_u have landed in every replica
            else:
                shard.exchange_rows(self.U_d, self.u_plan)
                shard.exchange_rows(self.b_u_d, self.u_plan)
        # (2) items (als.py:436-466); the item step ignores the feature term
        self._mark("item")
        if self.use_graph:
            model.last_paths["item"] = self._step(
                rt.cindptr, rt.cindices, rt.cvals, n, self.i_list, self.U_d, m, self.b_u_d, 0.0,
                self.reg_item_d, self.lam_bi, self.V_d, self.b_i_d, _lib.STEP_FACTOR)
            if self.i_f64 is not None:
                p64 = self._step(rt.cindptr, rt.cindices, rt.cvals, n, self.i_f64, self.U_d, m, self.b_u_d, 0.0,
                                 self.reg_item_d, self.lam_bi, self.V_d, self.b_i_d, _lib.STEP_FACTOR,
                                 path=_lib.PATH_F64)
                if self.i_list.numel() == 0:
                    model.last_paths["item"] = p64
            lp = self.level_ptr
            if self.sweep_fused:
                # all levels in one cooperative launch, a grid barrier between them
                check(lib.b200als_graph_sweep(
                    k, ptr(self.level_ptr_d), lp.size - 1, self.max_level, ptr(self.level_rows_d), ptr(self.chol_d),
                    ptr(self.aux_d), ptr(self.sptr_d), ptr(self.sidx_d), ptr(self.sval_d), float(model.alpha),
                    self.lam_bi, None, ptr(self.V_d), ld, ptr(self.b_i_d), stream), "graph_sweep")
                self.launches += 1
            else:
                for lv in range(lp.size - 1):
                    lo, hi = int(lp[lv]), int(lp[lv + 1])
                    check(lib.b200als_graph_level_solve(
                        k, self.level_rows_d.data_ptr() + 4 * lo, hi - lo, ptr(self.chol_d), ptr(self.aux_d),
                        ptr(self.sptr_d), ptr(self.sidx_d), ptr(self.sval_d), float(model.alpha), self.lam_bi, None,
                        ptr(self.V_d), ld, ptr(self.b_i_d), stream), "graph_level_solve")
                    self.launches += 1
        else:
            # A10 fused: with Z == V and no sweep after it, the item step sees every residual of
            # the iteration's final state, so it accumulates sum d / sum d^2 itself
            model.last_paths["item"] = self._step(
                rt.cindptr, rt.cindices, rt.cvals, rt.cindptr.numel() - 1, self.i_list, self.U_d, m, self.b_u_d, 0.0,
                self.reg_item_d, self.lam_bi, self.V_d, self.b_i_d, _lib.STEP_SOLVE,
                stats=self.acc if fused else None, split=self.i_split,
                norms=(self.acc.data_ptr() + 24) if fused else None,
                peers=(peers["V"], peers["b_i"]) if peers else None, row_ids=self.i_ids)
            if self.i_f64 is not None:
                p64 = self._step(
                    rt.cindptr, rt.cindices, rt.cvals, rt.cindptr.numel() - 1, self.i_f64, self.U_d, m, self.b_u_d,
                    0.0, self.reg_item_d, self.lam_bi, self.V_d, self.b_i_d, _lib.STEP_SOLVE,
                    stats=self.acc if fused else None, norms=(self.acc.data_ptr() + 24) if fused else None,
                    peers=(peers["V"], peers["b_i"]) if peers else None, row_ids=self.i_ids, path=_lib.PATH_F64)
                if self.i_list.numel() == 0:
                    model.last_paths["item"] = p64
            if shard.active and not peers:
                shard.exchange_rows(self.V_d, self.i_plan)
                shard.exchange_rows(self.b_i_d, self.i_plan)
            # (fused exchange: the residual all-reduce below orders the ranks after the item step; without fused
            # statistics an explicit token does)
            if shard.active and peers and not fused:
                shard.order_ranks(self._token)
        # (3) W_f (als.py:468-501)
        self._mark("wstep")
        if names and ((it % model.update_w_every == 0) or (it == self.n_iters - 1)):
            check(lib.b200als_item_gram_f64(n, ptr(rt.cindptr), ptr(rt.cindices), ptr(rt.cvals), k,
                                            ptr(self.U_d), ld, ptr(self.b_u_d), ptr(self.b_i_d), ptr(self.mu_d),
                                            ptr(self.G_d), ptr(self.cvec_d), stream), "item_gram_f64")
            self.launches += 1
            for f, nm in enumerate(names):
                d = int(self.X_d[f].shape[1])
                self._compose(self.Zmf_d, skip=f)   # every W still old: `residual` is never refreshed (als.py:474-489)
                lam = float(model.lambda_w.get(nm, 0.0)) + EPS
                check(lib.b200als_wstep_system(n, k, d, ptr(self.X_d[f]), ptr(self.G_d), ptr(self.cvec_d),
                                               ptr(self.Zmf_d), ld, lam, ptr(self.A_d), ptr(self.Wnew64[f]), stream),
                      "wstep_system")
                check(lib.b200als_cholesky_solve_f64(ptr(self.A_d), ptr(self.Wnew64[f]), d * k, ptr(self.status),
                                                     stream), "cholesky_solve_f64")
                self.launches += 6
            for f in range(len(names)):
                self.W64_d[f].copy_(self.Wnew64[f])
                self.W_d[f][:, :k] = self.Wnew64[f].view(-1, k).to(torch.float32)
        # (4) mu, RMSE, norms (als.py:503-517)
        self._mark("stats")
        if names:
            self._compose(self.Z_d)
        acc = self.acc
        ulo, uhi = self.ulo, self.uhi
        if not fused:
            acc.zero_()
            check(lib.b200als_residual_stats(uhi - ulo, rt.indptr.data_ptr() + 8 * ulo, ptr(rt.indices), ptr(rt.vals),
                                             k, self.U_d.data_ptr() + 4 * ld * ulo, ld, ptr(self.Z_d), ld,
                                             self.b_u_d.data_ptr() + 4 * ulo, ptr(self.b_i_d), ptr(acc), stream),
                  "residual_stats")
            self.launches += 1
        if shard.active:
            shard.all_reduce_sum(acc if fused else acc[:2])
        if not fused:
            check(lib.b200als_sumsq(ptr(self.U_d), m, k, ld, acc.data_ptr() + 16, stream), "sumsq")
            check(lib.b200als_sumsq(ptr(self.V_d), n, k, ld, acc.data_ptr() + 24, stream), "sumsq")
            check(lib.b200als_sumsq(ptr(self.b_u_d), 1, m, m, acc.data_ptr() + 32, stream), "sumsq")
            check(lib.b200als_sumsq(ptr(self.b_i_d), 1, n, n, acc.data_ptr() + 40, stream), "sumsq")
            self.launches += 4
        slot = it % self.hist.shape[0]
        if fused:       # also resets the accumulators to their start values for the next iteration
            check(lib.b200als_finalize_iter_fused(ptr(acc), ptr(self.acc_base), rt.nnz, ptr(self.mu_d),
                                                  self.hist.data_ptr() + 48 * slot, stream), "finalize_iter")
        else:
            check(lib.b200als_finalize_iter(ptr(acc), rt.nnz, ptr(self.mu_d), self.hist.data_ptr() + 48 * slot,
                                            stream), "finalize_iter")
        self.launches += 1
        self._mark("end")

    def close(self) -> None:
        """Drop the mappings of the other ranks' replicas (fused exchange).  Every rank has stopped storing into this
        rank's replicas once the last iteration's all-reduce has completed on its stream; a barrier after the release
        keeps the owners from freeing the memory under a mapping that is still open."""
        handles = getattr(self, "_peer_handles", None)
        if handles:
            torch.cuda.current_stream().synchronize()
            Shard.release_replicas(handles, self.dev)
            self.peers = None
            self.shard.barrier()

    def __del__(self):
        try:
            handles = getattr(self, "_peer_handles", None)
            if handles:
                Shard.release_replicas(handles, self.dev)
        except Exception:   # noqa: BLE001 — interpreter shutdown
            pass

    # ---- host sync points -----------------------------------------------------------------
    def raise_if_bad(self) -> None:
        if self.shard.active:
            # every rank must take the same branch here (the ranks meet again in the next collective): the counts
            # are summed over the ranks, so a system that is not positive definite on ONE rank raises on all of them
            cnt = self.status[:1].to(torch.float32)
            self.shard.all_reduce_sum(cnt)
            if float(cnt.item()) != 0.0 and int(self.status[0].item()) == 0:
                raise np.linalg.LinAlgError(
                    f"{int(cnt.item())} normal matrices were not positive definite (on another rank)")
        st = self.status.cpu().numpy()
        if st[0] != 0:
            raise np.linalg.LinAlgError(
                f"{int(st[0])} normal matrices were not positive definite (first row {int(st[1])})")

    def read_history(self, done: int) -> np.ndarray:
        h = self.hist[:done].cpu().numpy()
        self.raise_if_bad()
        return h

    def finish(self, done: int) -> None:
        """Copy the learned parameters back to the model's public float64 attributes."""
        model, rt, k = self.model, self.rt, self.k
        try:
            h = self.read_history(done)
        finally:
            self.close()        # also on the LinAlgError path: every rank releases its mappings and meets the barrier
        model._set_history(h)
        if done:
            model.mu = float(h[-1, 5])
        for f, nm in enumerate(self.names):
            model.W[nm] = self.W64_d[f].view(-1, k).cpu().numpy().copy()
        # U, V, b_u, b_i stay in HBM; the float64 host copies are made when the attributes are read (ALS._get_param)
        model._host = {"U": None, "V": None, "b_u": None, "b_i": None}
        model._dev_model = dict(U=self.U_d, V=self.V_d, b_u=self.b_u_d, b_i=self.b_i_d, mu=self.mu_d,
                                mu_value=float(model.mu), W={nm: self.W_d[f] for f, nm in enumerate(self.names)},
                                W_host={nm: model.W[nm] for nm in self.names}, ld=self.ld, m=rt.m, n=rt.n, k=k,
                                host_touched=False)
        if os.environ.get("B200ALS_EAGER_READBACK"):
            model.U, model.V, model.b_u, model.b_i     # noqa: B018 — reads materialise the host copies
        model.launches_per_iter = self.launches
