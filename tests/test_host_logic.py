"""CPU: host-side logic — config surface, drop-in shim, C-ABI symbol table, graph
schedule, row sharding over gloo (world_size 2), loud failure without a GPU."""
import ctypes
import dataclasses
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from _util import REPO, load_case, make_cfg


def test_config_surface_matches_reference_fields():
    import b200als
    core = [f.name for f in dataclasses.fields(b200als.CoreConfig)]
    assert core == ["n_factors", "n_iters", "lambda_u", "lambda_v", "pop_reg_mode", "random_state", "update_w_every"]
    c = b200als.CoreConfig(8, 3, 1.0, 2.0)                      # positional, als_config.py:57-66
    assert (c.pop_reg_mode, c.random_state, c.update_w_every) == (None, 42, 5)
    assert [f.name for f in dataclasses.fields(b200als.BiasesConfig)] == ["lambda_bu", "lambda_bi"]
    s = b200als.GraphSimConfig()
    assert (s.source, s.feature_name, s.metric, s.topk, s.eps) == ("feature", "genres", "cosine", 50, 1e-8)
    g = b200als.GraphConfig()
    assert g.alpha == 0.0 and g.sim is None
    a = b200als.ALSConfig(core=c)
    assert a.biases.lambda_bu is None and a.graph.alpha == 0.0


def test_constructor_semantics():
    import b200als
    cfg = b200als.ALSConfig(core=b200als.CoreConfig(8, 3, 1.5, 2.5), biases=b200als.BiasesConfig(0.0, None))
    m = b200als.ALS(cfg, lambda_w={"g": 1.0})
    assert m.lambda_bu == 1.5 and m.lambda_bi == 2.5            # falsy → fallback (als.py:166-167)
    assert m.S_topk is None and m.S_eps == 1e-10                # EPS default (als.py:173-174)
    assert set(m.history) == {"train_rmse", "U_norm", "V_norm", "bu_norm", "bi_norm"}
    assert m.U is None and m.mu == 0.0 and m.W == {} and m.lambda_w == {"g": 1.0}
    with pytest.raises(ValueError):
        b200als.ALS(None)
    with pytest.raises(RuntimeError):
        m.predict()


def test_scripts_shim_resolves_to_b200():
    from scripts.als import ALS
    from scripts.als_config import ALSConfig, CoreConfig, BiasesConfig, GraphConfig, GraphSimConfig  # noqa: F401
    import b200als
    assert ALS is b200als.ALS and ALSConfig is b200als.ALSConfig


def test_c_abi_exports_every_declared_symbol():
    """The library loads without a GPU and exports exactly what include/b200als.h declares."""
    import b200als
    hdr = open(os.path.join(REPO, "include", "b200als.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(b200als_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    lib = ctypes.CDLL(b200als._lib.LIB_PATH)
    for sym in declared:
        assert hasattr(lib, sym), f"{sym} declared in b200als.h but not exported"
    assert declared == set(b200als._lib.EXPORTED_SYMBOLS), declared ^ set(b200als._lib.EXPORTED_SYMBOLS)
    assert b200als._lib.load().b200als_abi_version() == b200als._lib.ABI_VERSION
    m = re.search(r"#define\s+B200ALS_ABI_VERSION\s+(\d+)", hdr)
    assert int(m.group(1)) == b200als._lib.ABI_VERSION
    # struct layout agreed between C and ctypes: size is checked by the library at call time
    assert ctypes.sizeof(b200als._lib.StepArgs) % 8 == 0


def test_invalid_arguments_are_rejected_without_a_gpu():
    import b200als
    lib = b200als._lib.load()
    need = ctypes.c_size_t(0)
    assert lib.b200als_step_workspace_bytes(0, 0, 10, ctypes.byref(need)) == -1
    assert b"k out of range" in lib.b200als_last_error()
    a = b200als._lib.StepArgs()
    a.struct_bytes = 4
    assert lib.b200als_row_step(ctypes.byref(a), None) == -1
    assert b"ABI" in lib.b200als_last_error()


def test_product_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import b200als
    c = load_case("small_plain")
    model = b200als.ALS(make_cfg(c.cfgd))
    with pytest.raises(b200als._lib.B200ALSError):
        model.fit_coo(c.flat, c.vals, (c.m, c.n), verbose=0)
    assert model.U is None


def test_product_never_imports_oracle():
    pkg = os.path.join(REPO, "collaborative-filtering_b200")
    for root, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(root, fn)).read()
                assert "als_oracle" not in src, f"{fn} references the oracle module"
                assert not re.search(r"^\s*(from|import)\s+[\w.]*oracle", src, flags=re.M), fn
                assert not re.search(r"#include\s+[\"<][^\n]*oracle", src), fn


def test_similarity_graph_matches_reference_fixture():
    from b200als import graph
    for name in ("small_graph_topk5", "small_graph_topk40"):
        c = load_case(name)
        sim = c.cfgd["sim"]
        S = graph.similarity_topk(c.features["genres"], sim["topk"], sim["eps"])
        assert S.dtype == c.z["S"].dtype and np.array_equal(S, c.z["S"])
        sptr, sidx, sval = graph.to_csr(S)
        dense = np.zeros_like(S)
        for i in range(S.shape[0]):
            dense[i, sidx[sptr[i]:sptr[i + 1]]] = sval[sptr[i]:sptr[i + 1]]
        assert np.array_equal(dense, S.astype(np.float32))


def test_level_schedule_reproduces_gauss_seidel():
    """Level-parallel execution (all items of a level read, then write) == sequential sweep."""
    from b200als import graph
    rng = np.random.default_rng(1)
    n, k = 120, 5
    X = (rng.random((n, 6)) < 0.35).astype(np.float32)
    S = graph.similarity_topk(X, 7, 1e-8)
    rated = rng.random(n) < 0.85
    sptr, sidx, sval = graph.to_csr(S)
    lp, rows = graph.level_schedule(sptr, sidx, rated)
    assert sorted(rows.tolist()) == np.flatnonzero(rated).tolist()
    V0 = rng.normal(size=(n, k))
    Vs = V0.copy()
    for i in range(n):                                   # als.py:436-461 shape of the dependency
        if rated[i]:
            Vs[i] = 0.5 * (S[i] @ Vs) + 0.1
    Vl = V0.copy()
    for lv in range(lp.size - 1):
        lvl = rows[lp[lv]:lp[lv + 1]]
        new = [0.5 * (S[i] @ Vl) + 0.1 for i in lvl]     # whole level reads before any write
        for i, v in zip(lvl, new):
            Vl[i] = v
    assert np.allclose(Vs, Vl, rtol=0, atol=1e-12)


def test_synth_is_deterministic_and_well_formed():
    from b200als.synth import synth_ratings
    f1, v1 = synth_ratings(50, 80, 700, seed=3)
    f2, v2 = synth_ratings(50, 80, 700, seed=3)
    assert np.array_equal(f1, f2) and np.array_equal(v1, v2)
    assert f1.size == 700 and np.unique(f1).size == 700 and np.all(np.diff(f1) > 0)
    assert f1.min() >= 0 and f1.max() < 50 * 80
    assert np.all((v1 * 2) == np.round(v1 * 2)) and v1.min() >= 0.5 and v1.max() <= 5.0


def test_shard_partition_covers_rows():
    import torch
    from b200als import Shard
    for n_rows in (1, 7, 8, 9, 1000):
        for world in (1, 2, 3, 8):
            seen = []
            for r in range(world):
                s = Shard(r, world)
                lo, hi = s.my_range(n_rows)
                seen += list(range(lo, hi))
                assert s.padded_rows(n_rows) == n_rows + (1 if world > 1 else 0)
            assert seen == list(range(n_rows))
            # the dealt schedule: every entry exactly once, counts differ by at most one,
            # and the exchange permutation lists rank r's rows in its slots (padding -> dummy row)
            order = torch.randperm(n_rows)
            dealt = [Shard(r, world).deal(order) for r in range(world)]
            assert sorted(torch.cat(dealt).tolist()) == list(range(n_rows))
            assert max(d.numel() for d in dealt) - min(d.numel() for d in dealt) <= 1
            for r in range(world):
                plan = Shard(r, world).plan_exchange(order, n_rows)
                assert plan.perm.numel() == world * plan.cnt
                for q in range(world):
                    slots = plan.perm[q * plan.cnt:(q + 1) * plan.cnt]
                    assert slots[:dealt[q].numel()].tolist() == dealt[q].tolist()
                    assert (slots[dealt[q].numel():] == n_rows).all()
                assert torch.equal(plan.mine, plan.perm[r * plan.cnt:(r + 1) * plan.cnt])


_GLOO_WORKER = r'''
import os, sys
sys.path.insert(0, os.environ["B200ALS_REPO"])
import torch, torch.distributed as td
import b200als
td.init_process_group("gloo", init_method="tcp://127.0.0.1:" + os.environ["PORT"],
                      rank=int(os.environ["RANK"]), world_size=2)
s = b200als.Shard.from_env()
m, ld = 7, 4
order = torch.tensor([5, 0, 3, 6, 1])                  # rows 2 and 4 are empty: nobody solves them
for shape in ((s.padded_rows(m), ld), (s.padded_rows(m),)):
    full = torch.full(shape, -1.0)
    full[m] = 0.0
    mine = s.deal(order)
    val = mine.to(torch.float32) + 0.25 * s.rank
    full[mine] = val[:, None] if len(shape) == 2 else val
    s.exchange_rows(full, s.plan_exchange(order, m))
    want = torch.full(shape, -1.0)
    want[m] = 0.0
    for r in range(2):
        rows = b200als.Shard(r, 2).deal(order)
        v = rows.to(torch.float32) + 0.25 * r
        want[rows] = v[:, None] if len(shape) == 2 else v
    assert torch.equal(full, want), (full, want)
# initial factors: rank 0 draws, everyone ends with its arrays (als.py FitSession init)
import numpy as np
from b200als import hostrng
shapes, dts = [(300, 8), (50, 8), (3, 8)], [np.float32, np.float32, np.float64]
bufs = [torch.zeros(sh, dtype=torch.float32 if d is np.float32 else torch.float64) for sh, d in zip(shapes, dts)]
if s.rank == 0:
    for b, a in zip(bufs, hostrng.normal_stream(42, shapes, dts, 0.1, n_threads=hostrng.host_threads(shared=False))):
        b.copy_(torch.from_numpy(a))
for b in bufs:
    s.broadcast_from_first(b)
rng = np.random.default_rng(42)
for b, sh, d in zip(bufs, shapes, dts):
    assert np.array_equal(b.numpy(), rng.normal(scale=0.1, size=sh).astype(d)), s.rank
# fold-per-GPU cross-validation: every fold evaluated once, results in fold order on every rank
from b200als.evaluate import folds_of_rank, merge_fold_results
mine = folds_of_rank(5, s.rank, 2)
assert mine == ([0, 2, 4] if s.rank == 0 else [1, 3])
got = merge_fold_results([(f, {"rmse": 0.5 + f, "by": s.rank}) for f in mine], 5)
assert [g["rmse"] for g in got] == [0.5, 1.5, 2.5, 3.5, 4.5] and [g["by"] for g in got] == [0, 1, 0, 1, 0]
# peer-mapped exchange is a CUDA facility: on CPU tensors the collective decision is "all-gather"
assert s.share_replicas(dict(U=torch.zeros(4, 4))) is None
tok = torch.zeros(1); s.order_ranks(tok)
# sharded upload of the rating list: each rank copies 1/N, all ranks end with the whole list (ragged length)
from b200als.sparse import upload_flat_sharded
flat_all = np.arange(0, 3 * 1001, 3, dtype=np.int64)
vals_all = (np.arange(1001) % 10 / 2.0 + 0.5)
f_d, v_d = upload_flat_sharded(flat_all, vals_all, torch.device("cpu"), s)
assert f_d.dtype == torch.int64 and v_d.dtype == torch.float32
assert np.array_equal(f_d.numpy(), flat_all) and np.array_equal(v_d.numpy(), vals_all.astype(np.float32))
f_d, v_d = upload_flat_sharded(torch.from_numpy(flat_all[:1]), torch.from_numpy(vals_all[:1]).float(), torch.device("cpu"), s)
assert f_d.tolist() == [0] and v_d.tolist() == [0.5]
try:
    upload_flat_sharded(flat_all[: 10 + s.rank], vals_all[: 10 + s.rank], torch.device("cpu"), s)
    raise AssertionError("lists of different lengths were accepted")
except ValueError:
    pass
acc = torch.tensor([1.0 + s.rank, 10.0], dtype=torch.float64)
s.all_reduce_sum(acc)
assert acc.tolist() == [3.0, 20.0]
s.barrier()
td.destroy_process_group()
print("ok", s.rank)
'''


def test_row_sharding_collectives_gloo_world2(tmp_path):
    import socket
    sock = socket.socket()
    sock.bind(("127.0.0.1", 0))
    port = sock.getsockname()[1]
    sock.close()
    script = tmp_path / "worker.py"
    script.write_text(_GLOO_WORKER)
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), PORT=str(port), B200ALS_REPO=REPO)
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.STDOUT))
    for p in procs:
        out, _ = p.communicate(timeout=180)
        assert p.returncode == 0, out.decode()


def test_rating_slabs_partition_matches_csr_csc():
    """slab.slabs_from_coo / synth_slabs_torch (CPU tensors): the slabs of all ranks together are the CSR and CSC of
    the whole matrix (scripts/als.py:338-339 order inside every row), local pointers, global ids."""
    import torch
    import als_oracle
    from b200als.slab import owned_count, slabs_from_coo, synth_slabs_torch
    from b200als.synth import synth_ratings
    m, n, nnz = 53, 37, 600
    flat, vals = synth_ratings(m, n, nnz, seed=4)
    sp = als_oracle.csr_csc_from_flat(flat, vals, m, n)
    rows, cols = torch.from_numpy(flat // n), torch.from_numpy(flat % n)
    for world in (1, 2, 3, 8):
        seen_u = seen_i = 0
        for rank in range(world):
            s = slabs_from_coo(rows, cols, torch.from_numpy(vals.astype(np.float32)), m, n, rank, world)
            assert s.nnz == nnz and abs(s.vals_sum - vals.sum()) < 1e-3
            assert s.row_ids_u.numel() == owned_count(m, rank, world) == s.indptr.numel() - 1
            assert s.row_ids_i.numel() == owned_count(n, rank, world) == s.cindptr.numel() - 1
            for l, u in enumerate(s.row_ids_u.tolist()):
                assert u % world == rank
                a, b = int(s.indptr[l]), int(s.indptr[l + 1])
                lo, hi = sp["indptr"][u], sp["indptr"][u + 1]
                assert np.array_equal(s.indices[a:b].numpy(), sp["indices"][lo:hi])
                assert np.array_equal(s.vals[a:b].numpy(), sp["vals"][lo:hi].astype(np.float32))
                seen_u += b - a
            for l, i in enumerate(s.row_ids_i.tolist()):
                a, b = int(s.cindptr[l]), int(s.cindptr[l + 1])
                lo, hi = sp["cindptr"][i], sp["cindptr"][i + 1]
                assert np.array_equal(s.cindices[a:b].numpy(), sp["cindices"][lo:hi])
                assert np.array_equal(s.cvals[a:b].numpy(), sp["cvals"][lo:hi].astype(np.float32))
                seen_i += b - a
        assert seen_u == nnz and seen_i == nnz
    # shard-wise generation: all ranks agree on one data set without exchanging it
    world = 3
    parts = [synth_slabs_torch(3000, 1200, 4000, seed=11, rank=r, world=world, device="cpu", n_chunks=5) for r in range(world)]
    total = parts[0].nnz
    assert all(p.nnz == total and p.vals_sum == parts[0].vals_sum for p in parts)
    assert abs(total - 4000) < 400          # de-duplication loses a few per cent at this density
    pairs_u, pairs_i = set(), set()
    for p in parts:
        for l, u in enumerate(p.row_ids_u.tolist()):
            seg = p.indices[int(p.indptr[l]):int(p.indptr[l + 1])].tolist()
            assert seg == sorted(set(seg))
            pairs_u.update((u, i) for i in seg)
        for l, i in enumerate(p.row_ids_i.tolist()):
            seg = p.cindices[int(p.cindptr[l]):int(p.cindptr[l + 1])].tolist()
            assert seg == sorted(set(seg))
            pairs_i.update((u, i) for u in seg)
    assert pairs_u == pairs_i and len(pairs_u) == total


def test_library_sass_holds_the_sm100_instructions_the_design_claims():
    """The built library is sm_100a code that really uses what DESIGN.md §4.1 says it uses: tcgen05 MMAs (SASS UTCHMMA)
    and TMEM loads (LDTM) in both instantiations of the Gram kernel, and packed fp32 FMAs (FFMA2) in the solvers of the
    tensor-core and of the FFMA path.  (cuobjdump ships with the CUDA toolkit; skipped where it is absent.)"""
    import re
    import shutil
    import subprocess
    import b200als
    exe = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(exe):
        pytest.skip("cuobjdump not available")
    sass = subprocess.run([exe, "-sass", b200als._lib.LIB_PATH], capture_output=True, text=True, timeout=300).stdout
    assert "sm_100a" in sass
    bodies = {}
    for m in re.finditer(r"Function : (\S+)", sass):
        end = sass.find("Function :", m.end())
        bodies[m.group(1)] = sass[m.end(): end if end > 0 else len(sass)]
    for K in (64, 128):
        name = next(n for n in bodies if f"row_step_umma_kernelILi{K}E" in n)
        body = bodies[name]
        assert len(re.findall(r"\bUTCHMMA\b", body)) >= 6, name        # Gram (3 products) + trailing updates of the LDL^T
        assert "LDTM" in body and "UTCBAR" in body, name
        assert len(re.findall(r"\bFFMA2\b", body)) >= 60, name
    ffma = [n for n in bodies if "row_step_ffma_kernel" in n]
    assert len(ffma) >= 5
    for n in ffma:
        assert len(re.findall(r"\bFFMA2\b", bodies[n])) >= 32, n


def test_factor_row_stride_and_float64_routing():
    """Host rules that decide WHICH kernel a row runs on (als.py): the factor row stride that selects the zero-padded
    tcgen05 kernels, and the rows that go to the float64 half-step (small ridge AND fewer than 4 k ratings)."""
    import types
    import torch
    import b200als
    from b200als import als as A
    from b200als._lib import PATH_AUTO, PATH_FFMA, PATH_UMMA
    # stride: (32, 64] -> 64, (64, 128] -> 128 on the tensor-core path; everything else 16-byte aligned rows
    want = {(8, PATH_AUTO): 8, (32, PATH_AUTO): 32, (33, PATH_AUTO): 64, (50, PATH_AUTO): 64, (64, PATH_AUTO): 64,
            (65, PATH_AUTO): 128, (96, PATH_AUTO): 128, (128, PATH_AUTO): 128, (129, PATH_AUTO): 132, (150, PATH_AUTO): 152,
            (20, PATH_UMMA): 64, (64, PATH_UMMA): 64, (100, PATH_UMMA): 128, (50, PATH_FFMA): 52, (96, PATH_FFMA): 96}
    for (k, path), ld in want.items():
        assert A.factor_ld(k, path) == ld, (k, path)

    cnt = torch.tensor([1, 300, 5, 1000, 0, 10], dtype=torch.int64)
    indptr = torch.zeros(7, dtype=torch.int64)
    indptr[1:] = torch.cumsum(cnt, 0)
    rows = torch.tensor([3, 1, 5, 2, 0], dtype=torch.int32)                  # a schedule: longest first, row 4 is empty

    def session(precision, path):
        s = types.SimpleNamespace(model=types.SimpleNamespace(precision=precision), k=64)
        s._fast_path = lambda factor_mode: path
        return s

    split = A.FitSession._precision_split
    # scalar ridge below the tensor-core threshold (0.3): the rows with fewer than 4 k = 256 ratings go to float64
    fast, slow = split(session("auto", PATH_UMMA), indptr, rows, 0.05, None, None, False)
    assert fast.tolist() == [3, 1] and slow.tolist() == [5, 2, 0]
    # 0.2 is below the tcgen05 threshold but above the FFMA one (0.1)
    fast, slow = split(session("auto", PATH_FFMA), indptr, rows, 0.2, None, None, False)
    assert fast.tolist() == rows.tolist() and slow is None
    fast, slow = split(session("auto", PATH_UMMA), indptr, rows, 5.0, None, None, False)
    assert fast.tolist() == rows.tolist() and slow is None
    # per-row ridge (popularity-scaled lambda_v): only short rows whose OWN ridge is small
    reg_row = torch.tensor([0.01, 0.01, 5.0, 0.01, 0.01, 0.2], dtype=torch.float32)
    fast, slow = split(session("auto", PATH_UMMA), indptr, rows, 0.0, reg_row, None, False)
    assert fast.tolist() == [3, 1, 2] and slow.tolist() == [5, 0]
    # ... looked up through the slab's global row ids when the schedule is local
    ids = torch.tensor([5, 4, 3, 2, 1, 0], dtype=torch.int64)
    reg_glob = torch.tensor([5.0, 0.01, 0.01, 0.01, 0.01, 0.01], dtype=torch.float32)
    fast, slow = split(session("auto", PATH_UMMA), indptr, rows, 0.0, reg_glob, ids, False)
    assert fast.tolist() == [3, 1, 5] and slow.tolist() == [2, 0]          # local row 5 is global row 0: ridge 5.0
    fast, slow = split(session("f32", PATH_UMMA), indptr, rows, 1e-4, None, None, False)
    assert fast.tolist() == rows.tolist() and slow is None
    fast, slow = split(session("f64", PATH_UMMA), indptr, rows, 5.0, None, None, False)
    assert fast.numel() == 0 and slow.tolist() == rows.tolist()
    with pytest.raises(ValueError):
        split(session("double", PATH_UMMA), indptr, rows, 1.0, None, None, False)


def test_bench_reference_arm_prints_the_contract_line():
    """`bench.py --impl reference` needs no GPU: it times the CPU port (oracle/) on a bounded sample.  The line must carry
    the keys the driver reads, with the same `metric` / `unit` / `config` naming as the CUDA arm."""
    import json
    out = subprocess.run([sys.executable, os.path.join(REPO, "bench.py"), "--impl", "reference", "--workload", "tiny",
                          "--steps", "1", "--warmup", "1"], capture_output=True, text=True, timeout=600, cwd=REPO)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    for key in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert key in line, key
    assert line["impl"] == "reference" and line["unit"] == "ratings/s" and line["higher_is_better"] is True
    assert line["config"]["workload"] == "tiny_debug" and "model" in line["config"]
    assert line["vs_baseline"] is None and line["value"] > 0
    cb = line["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == line["value"] and "sample" in cb
    e2e = line["e2e"]
    assert e2e["value"] == line["value"] and e2e["unit"] == line["unit"]
    assert e2e["h2d_bytes_per_step"] == 0 and e2e["d2h_bytes_per_step"] == 0
