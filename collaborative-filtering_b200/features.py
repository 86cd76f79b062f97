"""Item-feature preparation on the device — `scripts/prepare_features.py`.

Same functions, arguments, defaults and error messages as the reference's `normalize_feature` /
`normalize_features_dict` (prepare_features.py:131-233); the array work (column-median imputation of NaN / Inf,
row L1 / L2, column z-score / min-max) runs in libb200als kernels in float64 — the dtype numpy computes in for the
bundled int64 / float64 inputs — and is cast to `dtype` at the end, like prepare_features.py:51.  Sums are taken in
a different order than numpy's pairwise reduction, so values agree to float64 rounding (and, after the cast to
float32, to within one unit in the last place), not bit for bit.
"""
from __future__ import annotations

import ctypes as C
from typing import Any, Dict, Mapping

import numpy as np
import torch

from . import _lib
from ._lib import check, ptr

DEFAULT_DTYPE = "float32"      # prepare_features.py:50
DEFAULT_EPS = 1e-8             # prepare_features.py:51
_METHODS = {"none": 0, "row_l1": 1, "row_l2": 2, "col_zscore": 3, "col_minmax": 4}
_IMPUTES = {"none": 0, "col_median": 1}


def normalize_feature(X: np.ndarray, method: str = "none", *, impute: str = "none", eps: float = DEFAULT_EPS,
                      dtype: str = DEFAULT_DTYPE, copy: bool = True, device=None) -> np.ndarray:
    """prepare_features.py:131-191 on the device.  `copy` is accepted for signature parity (the input is never
    modified: the work happens on a device copy)."""
    if method not in _METHODS:
        raise ValueError(f"Unknown method '{method}'.")
    if impute not in _IMPUTES:
        raise ValueError(f"Unknown impute '{impute}'.")
    lib = _lib.load()
    _lib.require_device()
    X = np.asarray(X)
    if X.ndim == 1:
        X = X.reshape(-1, 1)                                          # prepare_features.py:55-59
    n, d = X.shape
    dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
    with torch.cuda.device(dev):
        X_d = torch.from_numpy(np.ascontiguousarray(X, dtype=np.float64)).to(dev)
        bad = torch.zeros(1, dtype=torch.int64, device=dev)
        med = torch.zeros(d + 1, dtype=torch.float64, device=dev)
        need = C.c_size_t(0)
        check(lib.b200als_normalize_workspace_bytes(n, C.byref(need)), "normalize_workspace_bytes")
        ws = torch.empty(need.value, dtype=torch.uint8, device=dev)
        # impute == "none": the library only counts the non-finite entries; they are an error (prepare_features.py:74-77)
        if impute == "none":
            check(lib.b200als_normalize_feature(ptr(X_d), n, d, 0, 0, float(eps), ptr(bad), None, None, 0,
                                                torch.cuda.current_stream().cuda_stream), "normalize_feature")
            if int(bad.item()):
                raise ValueError("Input feature contains NaN/Inf and impute='none'.")
        check(lib.b200als_normalize_feature(ptr(X_d), n, d, _METHODS[method], _IMPUTES[impute], float(eps), ptr(bad),
                                            ptr(med), ptr(ws), ws.numel(), torch.cuda.current_stream().cuda_stream),
              "normalize_feature")
        out = X_d.cpu().numpy()
    return out.astype(dtype, copy=False)


def normalize_features_dict(features: Mapping[str, np.ndarray], *, method: str = "none", impute: str = "none",
                            eps: float = DEFAULT_EPS, dtype: str = DEFAULT_DTYPE, copy: bool = True,
                            per_feature_overrides: Mapping[str, Mapping[str, Any]] | None = None) -> Dict[str, np.ndarray]:
    """prepare_features.py:194-233: shared defaults with per-feature overrides."""
    out: Dict[str, np.ndarray] = {}
    per_feature_overrides = per_feature_overrides or {}
    for name, X in features.items():
        o = dict(per_feature_overrides.get(name, {}))
        out[name] = normalize_feature(X, method=o.pop("method", method), impute=o.pop("impute", impute),
                                      eps=o.pop("eps", eps), dtype=o.pop("dtype", dtype), copy=o.pop("copy", copy), **o)
    return out
