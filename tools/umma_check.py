"""Scratch check: tcgen05 path vs FFMA path vs oracle on small shapes (run on the GPU box)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import numpy as np
import b200als, als_oracle
from b200als.synth import synth_ratings

def rel(a, b): return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))

for k in (64, 128):
    for (m, n, nnz, iters) in ((40, 30, 600, 1), (300, 200, 9000, 3), (3000, 500, 120000, 2)):
        flat, vals = synth_ratings(m, n, nnz, seed=k + m)
        cfg = b200als.ALSConfig(core=b200als.CoreConfig(n_factors=k, n_iters=iters, lambda_u=4.0, lambda_v=3.0),
                                biases=b200als.BiasesConfig(lambda_bu=2.0, lambda_bi=2.5))
        o = als_oracle.SparseALSOracle(cfg).fit_sparse(als_oracle.csr_csc_from_flat(flat, vals, m, n), m, n, tol=None)
        res = {}
        for name, path in (("ffma", 1), ("umma", 2)):
            mdl = b200als.ALS(cfg); mdl.gram_path = path
            t0 = time.time()
            try:
                mdl.fit_coo(flat, vals, (m, n), tol=None, verbose=0, sorted_input=True)
            except Exception as e:
                print(f"k={k} {m}x{n} {name}: {type(e).__name__} {e}", flush=True)
                continue
            res[name] = mdl
            print(f"k={k} {m}x{n} nnz={nnz} it={iters} {name}: U {rel(mdl.U,o.U):.2e} V {rel(mdl.V,o.V):.2e} "
                  f"bu {rel(mdl.b_u,o.b_u):.2e} bi {rel(mdl.b_i,o.b_i):.2e} "
                  f"rmse {abs(mdl.history['train_rmse'][-1]-o.history['train_rmse'][-1]):.2e} ({time.time()-t0:.2f}s)", flush=True)
