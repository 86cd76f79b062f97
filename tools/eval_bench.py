"""Measure the fold rows N1 / N2 on one GPU at a bench workload's size, against the
reference's host operations on the same arrays: python tools/eval_bench.py [c3|c4] -> one JSON line.

N2 fold split: algorithmic bytes = nnz * 12 (flat + rating read) + n_val * 8 (validation index read)
               + n_train * 12 (compacted pair written) + n_val * 4 (validation ratings written).
N1 fused SSE : per validation entry 8 (flat) + 4 (rating) + 2 * 4k (the two factor rows) + 8 (two biases).
CPU side: np.setdiff1d(all_obs, val_idx) + fancy-indexed ratings (create_folds.py:200-206 without the dense
scatter) and the pairwise prediction + RMSE of tune_params.py:165-167 over the same indices (numpy einsum)."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import b200als, bench
from b200als.evaluate import split_fold, popularity_bins
from b200als.synth import synth_ratings_torch

wl = sys.argv[1] if len(sys.argv) > 1 else "c3"
m, n, nnz, k = bench.WORKLOADS[wl]
dev = torch.device("cuda", 0)
peak = bench.measured_peaks()[0] if hasattr(bench, "measured_peaks") else 6452.5
rows, cols, vals = synth_ratings_torch(m, n, nnz, seed=bench.DATA_SEED, device=dev)
flat = rows * n + cols                                        # ascending (generator returns row-major order)
g = torch.Generator(device=dev); g.manual_seed(42)
perm = torch.randperm(nnz, device=dev, generator=g)
val = flat[perm[: nnz // 5]].contiguous()                      # one of 5 entry-wise folds (create_folds.py:79-84)
n_val = val.numel()

def timed(fn, reps=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): out = fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, out

tf, tv, vf, vv = split_fold(flat, vals, val)                   # the public call (allocates its outputs)
# timed: the C-ABI call alone on preallocated buffers (what a fold loop that reuses its buffers pays)
import ctypes as C
from b200als._lib import check, ptr
lib = b200als._lib.load()
need = C.c_size_t(0)
check(lib.b200als_fold_split_workspace_bytes(nnz, C.byref(need)), "ws")
ws = torch.empty(need.value, dtype=torch.uint8, device=dev)
o_flat = torch.empty(nnz, dtype=torch.int64, device=dev); o_vals = torch.empty(nnz, dtype=torch.float32, device=dev)
o_vv = torch.empty(n_val, dtype=torch.float32, device=dev)
o_n = torch.zeros(1, dtype=torch.int64, device=dev); o_st = torch.zeros(2, dtype=torch.int32, device=dev)
def run_split():
    check(lib.b200als_fold_split(ptr(flat), ptr(vals), nnz, ptr(val), n_val, ptr(o_flat), ptr(o_vals), ptr(o_vv), ptr(o_n),
                                 ptr(o_st), ptr(ws), ws.numel(), torch.cuda.current_stream().cuda_stream), "fold_split")
ms_split, _ = timed(run_split, reps=10)
assert int(o_n.item()) == tf.numel() and torch.equal(o_flat[: tf.numel()], tf) and torch.equal(o_vv, vv)
bytes_split = nnz * 12 + n_val * 8 + tf.numel() * 12 + n_val * 4

# a model to score with: random factors of the workload's k (the kernel's cost does not depend on the values)
cfg = b200als.ALSConfig(core=b200als.CoreConfig(n_factors=k, n_iters=1, lambda_u=5.0, lambda_v=5.0),
                        biases=b200als.BiasesConfig(3.0, 3.0))
model = b200als.ALS(cfg)
rng = np.random.default_rng(0)
model.U, model.V = rng.normal(scale=0.1, size=(m, k)), rng.normal(scale=0.1, size=(n, k))
model.b_u, model.b_i, model.mu = np.zeros(m), np.zeros(n), 3.5
counts = torch.bincount(cols, minlength=n).cpu().numpy()
item_bin, _ = popularity_bins(counts)
s = model._upload_model({}, dev)
ib = torch.from_numpy(item_bin.astype(np.int32)).to(dev)
sse = torch.zeros(5, dtype=torch.float64, device=dev); cnt = torch.zeros(5, dtype=torch.int64, device=dev)
def run_sse():
    sse.zero_(); cnt.zero_()
    check(lib.b200als_sse_on_flat(ptr(s["U"]), s["ld"], ptr(s["Z"]), s["ld"], ptr(s["b_u"]), ptr(s["b_i"]), ptr(s["mu"]),
                                  ptr(vf), ptr(vv), n_val, n, k, ptr(ib), 5, ptr(sse), ptr(cnt),
                                  torch.cuda.current_stream().cuda_stream), "sse_on_flat")
ms_sse, _ = timed(run_sse)
bytes_sse = n_val * (8 + 4 + 2 * 4 * k + 8)
rmse_gpu = float(torch.sqrt(sse.sum() / cnt.sum()))

# host reference operations on the same arrays (bounded: the full arrays, once)
fh, vh, valh = flat.cpu().numpy(), vals.cpu().numpy(), val.cpu().numpy()
t0 = time.perf_counter()
train_h = np.setdiff1d(fh, valh)                               # create_folds.py:200
tv_h = vh[np.searchsorted(fh, train_h)]
t1 = time.perf_counter()
sub = valh[: min(n_val, 2_000_000)]                            # bounded sample of the validation set
u, i = sub // n, sub % n
pred = np.einsum("ij,ij->i", model.U[u], model.V[i]) + model.mu + model.b_u[u] + model.b_i[i]
y = vh[np.searchsorted(fh, sub)]
rmse_h = float(np.sqrt(np.mean((y - pred) ** 2)))
t2 = time.perf_counter()
assert np.array_equal(train_h, tf.cpu().numpy()) and np.array_equal(tv_h, tv.cpu().numpy())
print(json.dumps({
    "workload": wl, "nnz": nnz, "n_val": n_val, "k": k,
    "fold_split": {"ms": ms_split, "GBps": bytes_split / ms_split / 1e6, "frac_of_hbm_peak": bytes_split / ms_split / 1e6 / peak,
                   "algorithmic_bytes": bytes_split, "cpu_numpy_s": t1 - t0, "bit_exact_vs_numpy": True,
                   "note": "C-ABI call on preallocated buffers, CUDA events"},
    "sse_on_flat": {"ms": ms_sse, "GBps": bytes_sse / ms_sse / 1e6, "frac_of_hbm_peak": bytes_sse / ms_sse / 1e6 / peak,
                    "algorithmic_bytes": bytes_sse, "rmse": rmse_gpu,
                    "cpu_numpy_s_extrapolated": (t2 - t1) * n_val / sub.size, "cpu_sample": int(sub.size), "cpu_rmse_sample": rmse_h},
    "peak_GBps": peak}))
