"""Where the end-to-end time of `ALS.fit_coo` goes at a bench workload's size (host timers with a device
synchronise after each stage): python tools/e2e_breakdown.py [c3|c4] [iterations] -> one JSON line."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import b200als, bench
from b200als.sparse import ratings_from_flat
from b200als.synth import synth_ratings_torch

wl = sys.argv[1] if len(sys.argv) > 1 else "c4"
K = int(sys.argv[2]) if len(sys.argv) > 2 else 10
m, n, nnz, k = bench.WORKLOADS[wl]
dev = torch.device("cuda", 0)
rows, cols, vals = synth_ratings_torch(m, n, nnz, seed=bench.DATA_SEED, device=dev)
flat_h = (rows * n + cols).cpu().pin_memory()
vals_h = vals.cpu().pin_memory()
del rows, cols, vals
cfg = b200als.ALSConfig(core=b200als.CoreConfig(n_factors=k, n_iters=K, lambda_u=5.0, lambda_v=5.0),
                        biases=b200als.BiasesConfig(3.0, 3.0))
out = {}
def stage(name, fn):
    torch.cuda.synchronize(); t = time.perf_counter(); r = fn(); torch.cuda.synchronize()
    out[name] = round(time.perf_counter() - t, 4); return r
for rep in range(2):                       # second pass: warm allocator, the numbers reported
    model = b200als.ALS(cfg)
    rt = stage("h2d_and_csr_csc_build", lambda: ratings_from_flat(flat_h, vals_h, (m, n), dev, sorted_input=True))
    rng = np.random.default_rng(42)
    t = time.perf_counter(); rng.normal(scale=0.1, size=(m, k)); rng.normal(scale=0.1, size=(n, k))
    out["numpy_rng_alone"] = round(time.perf_counter() - t, 4)
    t = time.perf_counter(); b200als.hostrng.normal_stream(42, [(m, k), (n, k)], [np.float32] * 2, 0.1)
    out["host_rng_alone"] = round(time.perf_counter() - t, 4)
    out["host_threads"] = b200als.hostrng.host_threads()
    mu0 = float(rt.vals.sum(dtype=torch.float64).item() / nnz)
    ses = stage("session_init_incl_host_rng_and_upload", lambda: model.session(rt, mu0))
    stage("iterations", lambda: [ses.iteration(it) for it in range(K)])
    stage("finish_history_only", lambda: ses.finish(K))
    from b200als import readback
    for mode in ("direct", "pipelined"):
        for hp in ("0", "1"):
            os.environ["B200ALS_HUGEPAGE"] = hp
            stage(f"readback_{mode}_hugepage{hp}", lambda: [readback.f64_from_device(t, mode=mode) for t in
                                                            (ses.U_d[:m, :k], ses.V_d[:n, :k], ses.b_u_d[:m], ses.b_i_d[:n])])
    os.environ["B200ALS_HUGEPAGE"] = "1"
    stage("read_attributes_default", lambda: (model.U, model.V, model.b_u, model.b_i))
t = time.perf_counter()
m2 = b200als.ALS(cfg); m2.fit_coo(flat_h, vals_h, (m, n), tol=None, verbose=0, sorted_input=True)
out["fit_coo_total"] = round(time.perf_counter() - t, 4)
_ = m2.U, m2.V, m2.b_u, m2.b_i
out["fit_coo_plus_attribute_reads"] = round(time.perf_counter() - t, 4)
out.update(workload=wl, n_iterations=K)
print(json.dumps(out))
