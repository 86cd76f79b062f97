"""torchrun --nproc-per-node N tools/slab_check.py: the sharded fit on slab-only storage (slab.RatingSlabs: every rank
holds the CSR of its users and the CSC of its items, nothing else) must reproduce the single-GPU fit on the whole
matrix, k = 64 / 128 (tcgen05, split rows) and 32 (FFMA), rows without ratings included."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as td
import b200als
from b200als.slab import slabs_from_coo
from b200als.synth import synth_ratings

rank = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(rank)
dev = torch.device("cuda", rank)
td.init_process_group("nccl", device_id=dev)
shard = b200als.Shard.from_env()
worst = 0.0
for k, (m, n, nnz) in ((64, (3000, 500, 120000)), (128, (700, 900, 60000)), (32, (400, 300, 9000))):
    flat, vals = synth_ratings(m, n, nnz, seed=7)
    keep = (flat // n != 5) & (flat % n != 7)
    flat, vals = flat[keep], vals[keep]
    cfg = b200als.ALSConfig(core=b200als.CoreConfig(n_factors=k, n_iters=3, lambda_u=4.0, lambda_v=3.0),
                            biases=b200als.BiasesConfig(lambda_bu=2.0, lambda_bi=2.5))
    one = b200als.ALS(cfg)
    one.split_rows, one.split_piece = 512, 256
    one.fit_coo(flat, vals, (m, n), verbose=0)
    rows, cols = torch.from_numpy(flat // n).to(dev), torch.from_numpy(flat % n).to(dev)
    slabs = slabs_from_coo(rows, cols, torch.from_numpy(vals.astype(np.float32)).to(dev), m, n, shard.rank, shard.world)
    many = b200als.ALS(cfg)
    many.split_rows, many.split_piece = 512, 256
    many.shard = shard
    many.fit_slabs(slabs, verbose=0)
    for nm in ("U", "V", "b_u", "b_i"):
        a, b = np.asarray(getattr(one, nm)), np.asarray(getattr(many, nm))
        err = float(np.abs(a - b).max() / max(np.abs(a).max(), 1e-30))
        worst = max(worst, err)
        assert err < 5e-5, (k, nm, err)       # see dist_check.py: run-to-run rounding of the tcgen05 path
    for key in ("train_rmse", "U_norm", "V_norm", "bu_norm", "bi_norm"):
        h1, h2 = np.asarray(one.history[key]), np.asarray(many.history[key])
        assert np.abs(h1 - h2).max() < 1e-5 * max(np.abs(h1).max(), 1.0), (k, key, h1, h2)
    if td.get_rank() == 0:
        print(f"slab k={k} world={td.get_world_size()} ok: rmse {one.history['train_rmse'][-1]:.6f} vs {many.history['train_rmse'][-1]:.6f}")
if td.get_rank() == 0:
    print("slab_check ok, worst rel diff", worst)
td.destroy_process_group()
