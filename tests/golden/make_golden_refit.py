#!/usr/bin/env python
"""Golden fixture for a SECOND fit() on the same object, from the UNMODIFIED reference (build container only).

    python tests/golden/make_golden_refit.py

The reference creates `history` once in __init__ (scripts/als.py:185-191) and every fit() appends to it
(:510-517), so after two fits the lists hold both runs, and `_converged` (:283-297) of the second fit looks at entries
of the first: with min_iters = 1 the second fit stops after its FIRST iteration, because h[-3] (the converged end of
fit 1) minus h[-1] (a fresh first iteration) is negative and therefore <= tol.  The fixture pins that behaviour.
"""
import importlib.util
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("B200ALS_REFERENCE", "/root/reference")
sys.path.insert(0, REF)

from scripts.als import ALS as RefALS                                   # noqa: E402
from scripts.als_config import ALSConfig, BiasesConfig, CoreConfig      # noqa: E402

assert os.path.realpath(sys.modules["scripts.als"].__file__).startswith(os.path.realpath(REF))
spec = importlib.util.spec_from_file_location("b200als_synth", os.path.join(REPO, "collaborative-filtering_b200", "synth.py"))
synth = importlib.util.module_from_spec(spec)
spec.loader.exec_module(synth)


def main():
    m, n = 40, 56
    R1 = synth.synth_dense(m, n, 700, 5)
    R2 = synth.synth_dense(m, n, 650, 6)
    core = dict(n_factors=6, n_iters=12, lambda_u=2.0, lambda_v=2.5, random_state=11)
    biases = dict(lambda_bu=1.5, lambda_bi=1.0)
    fits = [dict(tol=1e-3, min_iters=3), dict(tol=1e-3, min_iters=1)]
    model = RefALS(ALSConfig(core=CoreConfig(**core), biases=BiasesConfig(**biases)))
    out = dict(m=m, n=n, cfg=json.dumps(dict(core=core, biases=biases)), fits=json.dumps(fits))
    for j, (R, kw) in enumerate(zip((R1, R2), fits)):
        model.fit(R, features=None, verbose=0, **kw)
        flat = np.flatnonzero(~np.isnan(R))
        out[f"flat_idx_{j}"] = flat.astype(np.int64)
        out[f"vals_{j}"] = R.ravel()[flat]
        out[f"len_after_{j}"] = len(model.history["train_rmse"])
        out[f"U_{j}"], out[f"V_{j}"] = model.U.copy(), model.V.copy()
        out[f"b_u_{j}"], out[f"b_i_{j}"], out[f"mu_{j}"] = model.b_u.copy(), model.b_i.copy(), model.mu
    for key, h in model.history.items():
        out["hist_" + key] = np.asarray(h, dtype=np.float64)
    np.savez_compressed(os.path.join(HERE, "small_refit.npz"), **out)
    print("small_refit: history lengths after fit 1 / 2:", out["len_after_0"], out["len_after_1"])


if __name__ == "__main__":
    main()
