"""Debug (TIMING=1 build only): per-phase cycle counts of solver group 0, warp 3 on CTA 0 for ONE half-step
of a bench workload.  usage: solver_timing.py [c3|c4] [user|item]"""
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import b200als
from b200als import _lib
from b200als.sparse import ratings_from_flat
from b200als.synth import synth_ratings_torch
import bench

wl = sys.argv[1] if len(sys.argv) > 1 else "c3"
half = sys.argv[2] if len(sys.argv) > 2 else "user"
m, n, nnz, k = bench.WORKLOADS[wl]
dev = torch.device("cuda", 0)
rows, cols, vals = synth_ratings_torch(m, n, nnz, seed=bench.DATA_SEED, device=dev)
cfg = b200als.ALSConfig(core=b200als.CoreConfig(n_factors=k, n_iters=4, lambda_u=5.0, lambda_v=5.0),
                        biases=b200als.BiasesConfig(3.0, 3.0))
rt = ratings_from_flat(rows * n + cols, vals, (m, n), dev, sorted_input=True)
mdl = b200als.ALS(cfg)
ses = mdl.session(rt, float(vals.double().mean().item()), None, "device", True)
fn = C.CDLL(_lib.LIB_PATH).b200als_debug_timing
names = ["wait_rd", "wait_acc", "tmem_ld", "factor", "backsolve", "epilogue", "pairs"]
sub_names = ["win_ld", "start_bar", "left", "diag", "wait_diag", "subst"]
for it in range(3):
    ses.iteration(it)
    if half == "user":            # run the user half-step once more so that its counters are the last written
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ses._step(rt.indptr, rt.indices, rt.vals, m, ses.u_list, ses.Z_d, n, ses.b_i_d, ses.reg_user, None,
                  ses.lam_bu, ses.U_d, ses.b_u_d, _lib.STEP_SOLVE, split=ses.u_split)
        e1.record()
    torch.cuda.synchronize()
    out = (C.c_ulonglong * 16)()
    fn(out)
    v = [out[i] for i in range(7)]
    sub = [out[i] for i in range(8, 14)]
    pairs = max(v[6], 1)
    print(wl, half, "iter", it, {nm: round(v[i] / pairs) for i, nm in enumerate(names[:6])}, "pairs", v[6],
          "sum", round(sum(v[:6]) / pairs), "factor split", dict(zip(sub_names, [round(x / pairs) for x in sub])),
          ("ms %.3f" % e0.elapsed_time(e1)) if half == "user" else "")
ses.collect_phase_times()
print({p_: float(np.mean(t)) for p_, t in ses.phase_ms.items() if t})
