"""GPU: CSR / CSC construction is index-bit-exact with als.py:335-340 and with the
fold convention of create_folds.py (flat = u*n + i)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def numpy_csr_csc(R):
    mask = ~np.isnan(R)
    ui = [np.flatnonzero(mask[u]) for u in range(R.shape[0])]       # als.py:338
    iu = [np.flatnonzero(mask[:, i]) for i in range(R.shape[1])]    # als.py:339
    return mask, ui, iu


def check(rt, R):
    h = rt.to_host()
    mask, ui, iu = numpy_csr_csc(R)
    m, n = R.shape
    assert h["indptr"][0] == 0 and h["cindptr"][0] == 0
    assert np.array_equal(np.diff(h["indptr"]), mask.sum(1)) and np.array_equal(np.diff(h["cindptr"]), mask.sum(0))
    assert np.array_equal(h["indices"], np.concatenate(ui) if ui else [])
    assert np.array_equal(h["cindices"], np.concatenate(iu) if iu else [])
    ru, ri = np.where(mask)                                          # als.py:340
    assert np.array_equal(h["indices"], ri)
    assert np.array_equal(h["vals"], R[ru, ri].astype(np.float32))
    cu, ci = np.where(mask.T)
    assert np.array_equal(h["cvals"], R[ci, cu].astype(np.float32))
    if "perm" in h:
        assert np.array_equal(h["vals"][h["perm"]], h["cvals"])
        assert np.array_equal(h["indices"][h["perm"]], cu)


@pytest.mark.parametrize("shape,density", [((37, 53), 0.2), ((1, 70), 0.5), ((70, 1), 0.5),
                                           ((257, 129), 0.03), ((64, 64), 1.0), ((33, 31), 0.0)])
def test_dense_and_coo_builders(cuda_device, shape, density):
    from b200als.sparse import ratings_from_dense, ratings_from_flat
    rng = np.random.default_rng(11)
    R = np.round(rng.random(shape) * 9 + 1) / 2.0
    R[rng.random(shape) >= density] = np.nan
    if shape[0] > 8 and density > 0:
        R[3, :] = np.nan          # empty user
        R[:, 0] = np.nan          # empty item
    rt = ratings_from_dense(R, cuda_device, want_perm=True)
    check(rt, R)
    flat = np.flatnonzero(~np.isnan(R))
    vals = R.ravel()[flat]
    for sorted_input in (True, False):
        if sorted_input:
            f, v = flat, vals
        else:
            p = rng.permutation(flat.size)
            f, v = flat[p], vals[p]
        rt2 = ratings_from_flat(f, v, shape, cuda_device, sorted_input=sorted_input, want_perm=True)
        check(rt2, R)


def test_row_schedule_longest_first(cuda_device):
    import torch
    from b200als.sparse import row_schedule
    counts = np.array([3, 0, 7, 7, 1, 0, 12, 3], dtype=np.int64)
    indptr = torch.from_numpy(np.concatenate([[0], np.cumsum(counts)])).to(cuda_device)
    order, nne = row_schedule(indptr)
    assert nne == 6
    assert order.cpu().numpy().tolist() == [6, 2, 3, 0, 7, 4, 1, 5]


def test_fold_indices_roundtrip(cuda_device):
    """Train split given as flat indices (create_folds.py:200-206) builds the same CSR as the
    dense R_train the reference materialises."""
    from b200als.sparse import ratings_from_dense, ratings_from_flat
    from b200als.synth import synth_dense
    R = synth_dense(60, 90, 1500, seed=2)
    obs = np.flatnonzero(~np.isnan(R))
    rng = np.random.default_rng(42)
    rng.shuffle(obs)
    folds = [np.asarray(c, dtype=np.int64) for c in np.array_split(obs, 5)]      # create_folds.py:80-84
    val = folds[1]
    train = np.setdiff1d(np.flatnonzero(~np.isnan(R)), val)                       # create_folds.py:200
    R_train = np.full(R.size, np.nan)
    R_train[train] = R.ravel()[train]
    R_train = R_train.reshape(R.shape)
    a = ratings_from_dense(R_train, cuda_device).to_host()
    b = ratings_from_flat(train, R.ravel()[train], R.shape, cuda_device, sorted_input=True).to_host()
    for key in a:
        assert np.array_equal(a[key], b[key]), key
