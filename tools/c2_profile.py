"""One C2 fit (full model at the bundled shape, 20 iterations) + predict, for an ncu launch list:
python tools/c2_profile.py [iterations]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
import b200als
from _util import bundled_features, load_case, make_cfg
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 20
c = load_case("c2_full")
feats = bundled_features()
cfgd = dict(c.cfgd, core=dict(c.cfgd["core"], n_iters=iters))
for rep in range(2):
    t0 = time.perf_counter()
    model = b200als.ALS(config=make_cfg(cfgd), lambda_w=c.lambda_w)
    model.fit_coo(c.flat, c.vals, (c.m, c.n), features=feats, tol=None, verbose=0)
    pred = model.predict(feats)
    print(f"rep {rep}: fit + predict {time.perf_counter() - t0:.4f} s, rmse {model.history['train_rmse'][-1]:.6f}")
