"""A2 (scripts/als.py:329, 367-376): libb200als_host.so draws the reference's initial factors — one PCG64 stream,
`rng.normal(scale=0.1, size=...)` for U, V, W_f in order — on several threads, BIT-identical to numpy."""
import ctypes
import os
import re

import numpy as np
import pytest

from _util import REPO

import b200als
from b200als import hostrng


def _same_bits(a, b):
    return a.dtype == b.dtype and a.shape == b.shape and np.array_equal(a.view(np.uint8), b.view(np.uint8))


def _numpy_draw(seed, shapes, dtypes, scale):
    rng = np.random.default_rng(seed)                                     # what the reference does
    return [rng.normal(scale=scale, size=s).astype(d) for s, d in zip(shapes, dtypes)]


def test_host_library_is_built_and_exports_its_header():
    assert os.path.exists(hostrng.HOST_LIB_PATH), "libb200als_host.so missing: run __graft_entry__.build()"
    hdr = open(os.path.join(REPO, "include", "b200als_host.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(b200als_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(b200als._hostlib.EXPORTED_SYMBOLS) and "b200als_host_normal_fill" in declared
    lib = ctypes.CDLL(hostrng.HOST_LIB_PATH)
    for sym in declared:
        assert hasattr(lib, sym)


@pytest.mark.parametrize("seed", [0, 42, 2**40 + 17])
@pytest.mark.parametrize("shapes", [
    [(1, 1)], [(0, 4), (3, 4)], [(7, 3), (5, 3), (2, 3)],                 # serial walk
    [(1000, 64), (777, 64), (19, 64)],
    [(131072, 1), (1, 1), (5, 1)],                                        # exactly at the chunked-path threshold
    [(300001, 1), (12, 8), (19, 8)],                                      # chunked: several waves end inside a chunk
    [(1500, 128), (2_000_003, 1), (21, 128), (3, 128)],
])
def test_bit_identical_to_numpy(seed, shapes):
    dtypes = ([np.float32, np.float32] + [np.float64] * 4)[:len(shapes)]
    want = _numpy_draw(seed, shapes, dtypes, 0.1)
    for nth in (1, 2, 3, 8):
        got = hostrng.normal_stream(seed, shapes, dtypes, 0.1, n_threads=nth)
        assert all(_same_bits(g, w) for g, w in zip(got, want)), (seed, shapes, nth)


def test_other_scales_and_all_float64():
    shapes = [(400_000,), (100_000,)]
    for scale in (1.0, 0.37):
        want = _numpy_draw(5, shapes, [np.float64, np.float64], scale)
        got = hostrng.normal_stream(5, shapes, [np.float64, np.float64], scale, n_threads=4)
        assert all(_same_bits(g, w) for g, w in zip(got, want))


def test_generator_object_seed_and_numpy_fallback_agree():
    shapes, dtypes = [(50_000, 4), (30, 4)], [np.float32, np.float64]
    a = hostrng.normal_stream(11, shapes, dtypes, 0.1, n_threads=4)
    b = hostrng.normal_stream(11, shapes, dtypes, 0.1, force_numpy=True)
    assert all(_same_bits(x, y) for x, y in zip(a, b))
    # a caller-owned generator is drawn from by numpy itself and is left advanced as the reference leaves it
    g1, g2 = np.random.default_rng(11), np.random.default_rng(11)
    c = hostrng.normal_stream(g1, shapes, dtypes, 0.1, n_threads=4)
    want = _numpy_draw(g2, shapes, dtypes, 0.1)
    assert all(_same_bits(x, y) for x, y in zip(c, want)) and all(_same_bits(x, y) for x, y in zip(c, a))
    assert g1.bit_generator.state == g2.bit_generator.state
    # another bit generator: numpy's own walk
    d = hostrng.normal_stream(np.random.MT19937(3), [(1000,)], [np.float64], 0.1)
    assert _same_bits(d[0], np.random.Generator(np.random.MT19937(3)).normal(scale=0.1, size=(1000,)))


def test_session_init_uses_the_stream(monkeypatch):
    """FitSession draws through hostrng with the reference's order and dtypes (U, V float32; W_f float64)."""
    import inspect
    src = inspect.getsource(b200als.als)
    assert "_hostrng.normal_stream(seed" in src and "rng.normal(" not in src
    # the function the session (and fit_coo's overlapped pre-draw) calls: reference order and dtypes
    draws, pinned = b200als.als.draw_initial(42, 30, 20, 4, [3, 1], False)
    rng = np.random.default_rng(42)
    for got, (shape, dt) in zip(draws, [((30, 4), np.float32), ((20, 4), np.float32), ((3, 4), np.float64),
                                        ((1, 4), np.float64)]):
        assert got.dtype == dt and _same_bits(got, rng.normal(scale=0.1, size=shape).astype(dt))
    assert pinned == []


def test_tiny_chunks_exercise_the_stitch(tmp_path):
    """Rebuild with 64-word chunks: thousands of chunk boundaries, so calls that spill over a boundary (the case the
    serial stitch exists for) occur hundreds of times instead of about once per 80 chunks."""
    import shutil
    import subprocess
    if shutil.which("gcc") is None:
        pytest.skip("no gcc")
    npdir = os.path.dirname(np.__file__)
    so = str(tmp_path / "libhostrng_tiny.so")
    subprocess.run(["gcc", "-O1", "-fopenmp", "-fPIC", "-shared", "-DCHUNK_LOG=6", f"-I{npdir}/_core/include",
                    f"-I{REPO}/include", "-o", so, os.path.join(REPO, "collaborative-filtering_b200", "csrc", "host_rng.c"),
                    f"{npdir}/random/lib/libnpyrandom.a", f"{npdir}/_core/lib/libnpymath.a", "-lm"], check=True)
    lib = ctypes.CDLL(so)
    fn = lib.b200als_host_normal_fill
    fn.restype = ctypes.c_int
    fn.argtypes = [ctypes.c_uint64] * 4 + [ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                           ctypes.c_double, ctypes.c_int32, ctypes.c_void_p]
    M = (1 << 64) - 1
    for seed in range(6):
        st = np.random.default_rng(seed).bit_generator.state["state"]
        out = np.empty(200_000 + seed, np.float64)
        cnt, ptrs, f32 = np.array([out.size], np.int64), np.array([out.ctypes.data], np.uint64), np.zeros(1, np.int32)
        rc = fn(st["state"] >> 64, st["state"] & M, st["inc"] >> 64, st["inc"] & M, 1, cnt.ctypes.data, ptrs.ctypes.data,
                f32.ctypes.data, 1.0, 4, None)
        assert rc == 0
        assert _same_bits(out, np.random.default_rng(seed).standard_normal(out.size))


def test_host_threads_are_shared_between_local_ranks(monkeypatch):
    monkeypatch.delenv("B200ALS_HOST_THREADS", raising=False)
    monkeypatch.delenv("LOCAL_WORLD_SIZE", raising=False)
    alone = hostrng.host_threads()
    assert 1 <= alone <= 32
    monkeypatch.setenv("LOCAL_WORLD_SIZE", "8")
    assert hostrng.host_threads() == max(1, alone // 8)
    monkeypatch.setenv("B200ALS_HOST_THREADS", "5")
    assert hostrng.host_threads() == 5
