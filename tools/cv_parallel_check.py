"""torchrun --nproc-per-node N tools/cv_parallel_check.py: K folds spread over N GPUs (evaluate.cv_score_fold_parallel)
must give, fold by fold, what the serial loop on one GPU gives (evaluate.cv_score), and the golden numbers of the
reference's own drivers (tests/golden/eval_folds.npz).  Prints one JSON line with both wall times at a larger shape."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as td
import b200als
from b200als.evaluate import cv_score, cv_score_fold_parallel
from b200als.synth import synth_ratings

rank = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(rank)
td.init_process_group("nccl", device_id=torch.device("cuda", rank))
dev = torch.device("cuda", rank)
z = np.load(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "eval_folds.npz"))
m, n = int(z["m"]), int(z["n"])
k, iters, lu, lv, seed = z["cfg_core"]


def make():
    return b200als.ALS(b200als.ALSConfig(core=b200als.CoreConfig(n_factors=int(k), n_iters=int(iters), lambda_u=float(lu),
                                                                 lambda_v=float(lv), random_state=int(seed)),
                                         biases=b200als.BiasesConfig(lambda_bu=1.5, lambda_bi=1.5)))


folds = [z[f"val_idx_{f}"] for f in range(4)]
vals = z["vals"].astype(np.float32) / 2.0
par = cv_score_fold_parallel(make, z["flat_idx"], vals, (m, n), folds, tol=None, min_iters=0,
                             item_bin=z["bin_quantile"], n_bins=int(z["n_bins"]))
ser = cv_score(make, torch.from_numpy(z["flat_idx"]).to(dev), torch.from_numpy(vals).to(dev), (m, n), folds, tol=None,
               min_iters=0, item_bin=z["bin_quantile"], n_bins=int(z["n_bins"]))
for f in range(4):
    assert abs(par[f]["rmse"] - ser[f]["rmse"]) < 1e-6 and par[f]["iters"] == ser[f]["iters"], (f, par[f], ser[f])
    assert abs(par[f]["rmse"] - float(z[f"rmse_{f}"])) < 1e-4

# timing at the bundled shape: 8 folds, U/V + biases, k = 32 (C1)
M, N, NNZ = 610, 4980, 100_000
flat, v = synth_ratings(M, N, NNZ, seed=5)
rng = np.random.default_rng(42)
perm = rng.permutation(flat)
F = 8
big_folds = [np.sort(p) for p in np.array_split(perm, F)]


def make_c1():
    return b200als.ALS(b200als.ALSConfig(core=b200als.CoreConfig(n_factors=32, n_iters=20, lambda_u=5.0, lambda_v=6.0),
                                         biases=b200als.BiasesConfig(lambda_bu=3.0, lambda_bi=2.0)))


out = {}
for name, fn in (("serial_one_gpu", lambda: cv_score(make_c1, torch.from_numpy(flat).to(dev), torch.from_numpy(v.astype(np.float32)).to(dev),
                                                      (M, N), big_folds, tol=1e-4, min_iters=10)),
                 ("fold_parallel", lambda: cv_score_fold_parallel(make_c1, flat, v.astype(np.float32), (M, N), big_folds,
                                                                  tol=1e-4, min_iters=10))):
    fn()                                                  # warm-up
    td.barrier(); torch.cuda.synchronize(); t = time.perf_counter()
    r = fn()
    torch.cuda.synchronize(); td.barrier()
    out[name + "_s"] = round(time.perf_counter() - t, 4)
    out[name + "_rmse"] = [round(x["rmse"], 6) for x in r]
assert np.allclose(out["serial_one_gpu_rmse"], out["fold_parallel_rmse"], atol=1e-5)
if td.get_rank() == 0:
    out.update(world=td.get_world_size(), folds=F, shape=[M, N, NNZ], k=32)
    print(json.dumps(out))
td.destroy_process_group()
