"""GPU, at BASELINE.json's C3 size (138 493 users x 26 744 items, 20 M ratings, k = 64) — too large for the oracle,
so the path is held to properties that do not depend on the size:

  * the train RMSE the half-steps accumulate on the fly (fused residual statistics, scripts/als.py:503-512) equals
    the RMSE a different kernel recomputes from the returned model over all 20 M training entries;
  * the Frobenius norms accumulated inside the solver (scripts/als.py:513-516) equal the norms of the returned arrays;
  * a fold split followed by a merge gives back the observed set (train ∪ validation, disjoint, train ascending —
    scripts/create_folds.py:200-206), and squared errors are additive over the two parts and over popularity bins.

(File name: runs after the parity tests.)
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

M, N, NNZ, K = 138_493, 26_744, 20_000_000, 64          # BASELINE.json configs[2] (bench workload c3)
DATA_SEED = 1234


@pytest.fixture(scope="module")
def c3(cuda_device):
    import torch
    import b200als
    from b200als.synth import synth_ratings_torch
    dev = cuda_device
    rows, cols, vals = synth_ratings_torch(M, N, NNZ, seed=DATA_SEED, device=dev)
    flat = rows * N + cols                                # ascending: the generator returns row-major order
    counts = torch.bincount(cols, minlength=N).cpu().numpy()
    del rows, cols
    cfg = b200als.ALSConfig(core=b200als.CoreConfig(n_factors=K, n_iters=3, lambda_u=5.0, lambda_v=5.0),
                            biases=b200als.BiasesConfig(3.0, 3.0))
    model = b200als.ALS(cfg)
    model.fit_coo(flat, vals, (M, N), tol=None, verbose=0, sorted_input=True)
    return dict(model=model, flat=flat, vals=vals, counts=counts, dev=dev)


def test_fused_train_rmse_equals_recomputed(c3):
    model = c3["model"]
    hist = model.history["train_rmse"]
    assert len(hist) == 3 and all(np.isfinite(hist)) and hist[-1] < hist[0]
    sse, cnt = model.sse_on_flat(c3["flat"], c3["vals"])
    assert int(cnt[0]) == NNZ
    recomputed = float(np.sqrt(sse[0] / cnt[0]))
    assert abs(recomputed - hist[-1]) < 2e-5, (recomputed, hist[-1])


def test_fused_norms_equal_norms_of_the_result(c3):
    model = c3["model"]
    assert model.U.shape == (M, K) and model.V.shape == (N, K) and model.U.dtype == np.float64
    for key, arr in (("U_norm", model.U), ("V_norm", model.V), ("bu_norm", model.b_u), ("bi_norm", model.b_i)):
        want = float(np.linalg.norm(arr))
        got = model.history[key][-1]
        assert abs(got - want) <= 1e-5 * max(want, 1.0), (key, got, want)


def test_fold_split_round_trip_and_additive_errors(c3):
    import torch
    from b200als.evaluate import popularity_bins, split_fold
    model, flat, vals, dev = c3["model"], c3["flat"], c3["vals"], c3["dev"]
    g = torch.Generator(device=dev)
    g.manual_seed(42)
    perm = torch.randperm(NNZ, device=dev, generator=g)
    val = flat[perm[: NNZ // 5]].contiguous()             # one of five entry-wise folds (create_folds.py:79-84)
    tf, tv, vf, vv = split_fold(flat, vals, val)
    assert tf.numel() + vf.numel() == NNZ and torch.equal(vf, val)
    assert bool((tf[1:] > tf[:-1]).all())                 # ascending and unique: the sorted setdiff1d
    merged, order = torch.sort(torch.cat([tf, vf]))
    assert torch.equal(merged, flat)                      # train ∪ validation == observed set, disjoint
    assert torch.equal(torch.cat([tv, vv])[order], vals)  # and every rating travelled with its index
    del merged, order, perm

    sse_all, cnt_all = model.sse_on_flat(flat, vals)
    sse_t, cnt_t = model.sse_on_flat(tf, tv)
    sse_v, cnt_v = model.sse_on_flat(vf, vv)
    assert int(cnt_t[0]) + int(cnt_v[0]) == int(cnt_all[0]) == NNZ
    assert abs((sse_t[0] + sse_v[0]) - sse_all[0]) <= 1e-7 * sse_all[0]

    item_bin, _ = popularity_bins(c3["counts"])
    n_bins = int(item_bin.max()) + 1
    sse_b, cnt_b = model.sse_on_flat(vf, vv, item_bin=item_bin, n_bins=n_bins)
    assert int(cnt_b.sum()) == int(cnt_v[0])
    assert abs(sse_b.sum() - sse_v[0]) <= 1e-7 * sse_v[0]
    bin_of_entry = torch.from_numpy(item_bin).to(dev)[vf % N]
    assert np.array_equal(cnt_b, torch.bincount(bin_of_entry, minlength=n_bins).cpu().numpy())
