"""GPU: b200als_row_order_static — the row schedule without any allocation (scratch in a file-scope static __device__
array inside libb200als.so, include/b200als.h) — against the existing b200als_row_order, through the C-ABI, bit for bit
(the tolerance of every index test in this suite), on two calls in a row that share the static scratch.

No counterpart in the reference.
(File name: collected LAST, so that under `pytest -x` a failure here hides no other test.)
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _indptr(rng, n_rows, empty_frac):
    cnt = rng.integers(0, 200, size=n_rows)
    cnt[rng.random(n_rows) < empty_frac] = 0
    cnt[rng.integers(0, n_rows, size=max(1, n_rows // 50))] = rng.integers(1000, 100_000, size=max(1, n_rows // 50))
    ip = np.zeros(n_rows + 1, dtype=np.int64)
    np.cumsum(cnt, out=ip[1:])
    return ip


def _both(lib, torch, dev, ip_host):
    """(order, n_nonempty) of the existing entry point and of the static one on the same indptr."""
    from b200als._lib import check, ptr
    n_rows = ip_host.size - 1
    ip = torch.from_numpy(ip_host).to(dev)
    stream = torch.cuda.current_stream().cuda_stream
    # existing entry point: caller-owned workspace
    need = C.c_size_t(0)
    check(lib.b200als_row_order_workspace_bytes(n_rows, C.byref(need)), "row_order_workspace_bytes")
    ws = torch.empty(need.value, dtype=torch.uint8, device=dev)
    order_a = torch.full((n_rows,), -7, dtype=torch.int32, device=dev)
    nne_a = torch.full((1,), -7, dtype=torch.int64, device=dev)
    check(lib.b200als_row_order(ptr(ip), n_rows, ptr(order_a), ptr(nne_a), ptr(ws), ws.numel(), stream), "row_order")
    # static entry point: inputs and outputs only
    order_b = torch.full((n_rows,), -9, dtype=torch.int32, device=dev)
    nne_b = torch.full((1,), -9, dtype=torch.int64, device=dev)
    check(lib.b200als_row_order_static(ptr(ip), n_rows, ptr(order_b), ptr(nne_b), stream), "row_order_static")
    torch.cuda.synchronize()
    return order_a.cpu().numpy(), int(nne_a.item()), order_b.cpu().numpy(), int(nne_b.item())


def test_row_order_static_equals_row_order_on_two_calls_in_a_row(cuda_device):
    import torch
    import b200als
    from b200als._lib import last_error, load, ptr
    lib = load()
    limit = int(lib.b200als_row_order_static_max_rows())
    assert limit == 1 << 18
    rng = np.random.default_rng(17)
    # two calls in a row on the shared scratch: a long one, then a shorter one that finds the first call's contents
    # behind its own rows; then the limit itself
    for n_rows, empty_frac in ((200_003, 0.1), (4_980, 0.3), (limit, 0.0), (1, 0.0)):
        ip = _indptr(rng, n_rows, empty_frac)
        oa, na, ob, nb = _both(lib, torch, cuda_device, ip)
        cnt = np.diff(ip)
        assert na == nb == int((cnt > 0).sum())
        assert np.array_equal(oa, ob)                                   # bit-exact: same schedule
        assert np.array_equal(ob, np.argsort(-cnt, kind="stable"))      # and it IS longest first, ties by ascending id
    # beyond the limit: a clear error, nothing launched, nothing written
    n_big = limit + 1
    ip = torch.zeros(n_big + 1, dtype=torch.int64, device=cuda_device)
    order = torch.full((n_big,), -3, dtype=torch.int32, device=cuda_device)
    nne = torch.full((1,), -3, dtype=torch.int64, device=cuda_device)
    rc = lib.b200als_row_order_static(ptr(ip), n_big, ptr(order), ptr(nne), torch.cuda.current_stream().cuda_stream)
    assert rc == -4                                                     # B200ALS_E_UNSUPPORTED
    assert "exceeds the static scratch" in last_error()
    torch.cuda.synchronize()
    assert int(nne.item()) == -3 and int(order[0].item()) == -3
    assert lib.b200als_row_order_static(None, 5, ptr(order), ptr(nne), None) == -1      # B200ALS_E_INVALID
