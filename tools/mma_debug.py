"""Debug (TIMING=1 build): window of row RPW before / after the first tensor-core trailing update."""
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import b200als
from b200als.synth import synth_ratings
k = int(sys.argv[1]) if len(sys.argv) > 1 else 64
m, n, nnz = 40, 30, 600
flat, vals = synth_ratings(m, n, nnz, seed=1)
cfg = b200als.ALSConfig(core=b200als.CoreConfig(n_factors=k, n_iters=1, lambda_u=4.0, lambda_v=3.0),
                        biases=b200als.BiasesConfig(lambda_bu=2.0, lambda_bi=2.5))
mdl = b200als.ALS(cfg)
try:
    mdl.fit_coo(flat, vals, (m, n), tol=None, verbose=0, sorted_input=True)
except Exception as e:
    print("fit:", type(e).__name__, e)
lib = C.CDLL(b200als._lib.LIB_PATH)
out = (C.c_float * 64)()
lib.b200als_debug_floats(out)
v = np.array(out[:])
print("before", v[:16])
print("after ", v[16:32])
print("diff  ", v[16:32] - v[:16])
