"""Debug (TIMING=1 build only): per-phase cycle counts of solver group 0, warp 3 on CTA 0 for one half-step."""
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import b200als
from b200als.synth import synth_ratings
k = int(sys.argv[1]) if len(sys.argv) > 1 else 64
m, n, nnz = (int(x) for x in sys.argv[2:5]) if len(sys.argv) > 4 else (60000, 8000, 4000000)
flat, vals = synth_ratings(m, n, nnz, seed=1)
cfg = b200als.ALSConfig(core=b200als.CoreConfig(n_factors=k, n_iters=2, lambda_u=4.0, lambda_v=3.0),
                        biases=b200als.BiasesConfig(lambda_bu=2.0, lambda_bi=2.5))
mdl = b200als.ALS(cfg)
from b200als.sparse import ratings_from_flat
rt = ratings_from_flat(flat, vals, (m, n), torch.device("cuda"))
ses = mdl.session(rt, float(vals.mean()), None, "device", True)
lib = b200als._lib.load()
lib._handle  # noqa
fn = C.CDLL(b200als._lib.LIB_PATH).b200als_debug_timing
names = ["wait_rd", "wait_acc", "tmem_ld", "factor", "backsolve", "epilogue", "pairs"]
for it in range(2):
    ses.iteration(it)
    torch.cuda.synchronize()
    out = (C.c_ulonglong * 16)()
    fn(out)
    v = [out[i] for i in range(7)]
    sub = [out[i] for i in range(8, 12)]
    pairs = max(v[6], 1)
    print("iter", it, "(item step timing, last launch)", {nm: round(v[i] / pairs) for i, nm in enumerate(names[:6])}, "pairs", v[6],
          "factor split", dict(zip(["win_ld", "start_bar", "left", "inblock"], [round(x / pairs) for x in sub])))
ses.collect_phase_times()
print({p_: np.mean(t) for p_, t in ses.phase_ms.items() if t})
