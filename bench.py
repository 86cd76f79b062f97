#!/usr/bin/env python
"""bench.py — ALS iterations/s and ratings/s/iter on synthetic power-law ratings.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c4] [--impl b200|reference]

A "step" is one full ALS iteration (user half-step, item half-step, mu / RMSE /
norm reductions; als.py:408-517) over the whole rating set.  Default workload is
C4 (Netflix-prize shape, 480 189 x 17 770, 1e8 ratings, k = 128): the largest
single-GPU configuration of BASELINE.json and the one named for 1/2/4/8-GPU
row sharding (strong scaling: the rating set is fixed as N grows).

Prints ONE JSON line (rank 0).  `value` = ratings/s/iter with ratings resident in
HBM (CUDA events, max over ranks); `e2e` = the same metric through the public
`ALS.fit_coo` call with pinned HOST buffers (H2D copy, CSR/CSC build, K
iterations, D2H of the factors all inside the timed region); `roofline` = the
dominant kernel's algorithmic bytes / its event-timed duration against the
measured copy bandwidth; `cpu_baseline` = the oracle port of the reference's
numpy loop timed on this box's host cores over a bounded row sample.
`--impl reference` runs only that CPU arm and prints the same line shape.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

REPO = os.path.dirname(os.path.abspath(__file__))
if REPO not in sys.path:
    sys.path.insert(0, REPO)

WORKLOADS = {
    # name: (m, n, nnz, k)   — SURVEY.md §8 config shorthand
    "c1": (610, 4980, 100_000, 32),
    "c3": (138_493, 26_744, 20_000_000, 64),
    "c4": (480_189, 17_770, 100_000_000, 128),
    "c4k64": (480_189, 17_770, 100_000_000, 64),
    "tiny": (3000, 2000, 200_000, 64),
}
# DRAM bytes (read + write) of one row_step launch, from `ncu --set full` on a single GPU; None = not captured.
# Sources: profiles/r01_ncu_c3_umma_k64_v8.txt, profiles/r01_ncu_c4_umma_k128_v6.txt (at c3 the algorithmic bytes
# are mostly served by the 126 MB L2; at c4 the item half-step's gathered matrix no longer fits).
NCU_DRAM_BYTES = {("c3", "user"): 204.5e6, ("c3", "item"): 211.2e6, ("c4", "user"): 1.195e9, ("c4", "item"): 22.23e9}
WORKLOAD_NAMES = {"c1": "c1_bundled_shape_uv_biases", "c3": "c3_movielens20m_shape",
                  "c4": "c4_netflix_shape", "c4k64": "c4_netflix_shape_k64", "tiny": "tiny_debug"}
HYPER = dict(lambda_u=5.0, lambda_v=5.0, lambda_bu=3.0, lambda_bi=3.0)
DATA_SEED = 1234


def algorithmic_bytes_half(nnz, rows, k):
    """per rating one gathered factor row (4k) + index + rating + opposite bias;
    per row: factor row out (4k), own bias r/w (8), indptr (4), lambda (4)."""
    return nnz * (4 * k + 12) + rows * (4 * k + 16)


def measured_peaks():
    try:
        with open(os.path.join(REPO, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc, self.th = index, [], None, None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            return
        self.th = threading.Thread(target=self._read, daemon=True)
        self.th.start()

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line.strip()))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, reasons, n = [], None, set(), 0
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        rows = [r for (t, r) in self.rows if t0 - 0.05 <= t <= t1 + 0.15] or [r for (_, r) in self.rows]
        for r in rows:
            parts = [x.strip() for x in r.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0]))
                smax = float(parts[1])
            except ValueError:
                continue
            n += 1
            for nm, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smax,
                "reasons": sorted(reasons), "samples": n}


# --------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference's numpy loop on a bounded row sample
# --------------------------------------------------------------------------------------
def cpu_reference_arm(wl, steps, warmup, budget_s, data=None):
    """Times oracle/als_oracle.py (kind "port": the reference is pure Python and cannot travel
    to the GPU box; the port makes the same numpy/scipy calls per row and is pinned
    bit-identical to it by tests/test_oracle_vs_golden.py) on the first phi*m user rows and
    phi*n item rows of the workload.  Returns (ratings_per_s, info)."""
    sys.path.insert(0, os.path.join(REPO, "oracle"))
    import als_oracle
    import b200als
    m, n, nnz, k = WORKLOADS[wl]
    if data is None:
        data = make_host_sample_source(wl)
    indptr, indices, vals, cindptr, cindices, cvals = data
    cfg = b200als.ALSConfig(core=b200als.CoreConfig(n_factors=k, n_iters=1, lambda_u=HYPER["lambda_u"],
                                                    lambda_v=HYPER["lambda_v"]),
                            biases=b200als.BiasesConfig(HYPER["lambda_bu"], HYPER["lambda_bi"]))
    rng = np.random.default_rng(42)
    U = rng.normal(scale=0.1, size=(m, k))
    V = rng.normal(scale=0.1, size=(n, k))

    def run(phi):
        nu, ni = max(1, int(round(phi * m))), max(1, int(round(phi * n)))
        nu, ni = min(nu, indptr.size - 1), min(ni, cindptr.size - 1)
        sp = dict(indptr=indptr[:nu + 1], indices=indices[:indptr[nu]], vals=vals[:indptr[nu]].astype(np.float64),
                  cindptr=cindptr[:ni + 1], cindices=cindices[:cindptr[ni]],
                  cvals=cvals[:cindptr[ni]].astype(np.float64), rows_u=np.zeros(0, dtype=np.int64))
        o = als_oracle.SparseALSOracle(cfg)
        t0 = time.perf_counter()
        o.fit_sparse(sp, m, n, tol=None, init=dict(U=U, V=V), mu0=3.5, max_rows=(nu, ni))
        dt = time.perf_counter() - t0
        su, si = o.sampled_ratings
        return dt, 0.5 * (su + si), nu, ni

    # calibrate on a small fraction, then size the sample to the per-step budget
    phi_max = min(1.0, (indptr.size - 1) / m, (cindptr.size - 1) / n)
    phi0 = min(phi_max, max(2e-4, 64.0 / n))
    dt0, r0, _, _ = run(phi0)
    per_step = budget_s / max(1, steps + warmup)
    phi = min(phi_max, max(phi0, phi0 * per_step / max(dt0, 1e-3)))
    for _ in range(warmup):
        run(phi)
    times, ratings = [], 0.0
    for _ in range(steps):
        dt, r, nu, ni = run(phi)
        times.append(dt)
        ratings = r
    t = float(np.mean(times))
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    info = dict(kind="port", cores=cores, unit="ratings/s",
                sample=f"first {nu} of {m} user rows + first {ni} of {n} item rows per step "
                       f"({ratings:.3g} rating-updates/step, phi={phi:.2e}); numpy/scipy row loop, "
                       f"OpenBLAS threads default")
    return ratings / t, t, info


def make_host_sample_source(wl):
    """Host CSR/CSC prefix arrays for the CPU arm.  Data are generated with the same recipe
    (on the GPU when one is present — generation is not the measured path — else on the CPU)."""
    import torch
    from b200als.synth import synth_ratings_torch
    m, n, nnz, k = WORKLOADS[wl]
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    rows, cols, vals = synth_ratings_torch(m, n, nnz, seed=DATA_SEED, device=dev)
    return host_prefix_from_coo(rows, cols, vals, m, n)


def host_prefix_from_coo(rows, cols, vals, m, n, frac=0.05):
    """CSR of the first users and CSC of the first items (a prefix is all the CPU arm reads).
    Uses torch ops for the data shuffle only (test-data preparation, not the ALS path)."""
    import torch
    nu, ni = max(64, int(m * frac)), max(64, int(n * frac))
    nu, ni = min(nu, m), min(ni, n)
    ru = rows < nu
    r_u, c_u, v_u = rows[ru], cols[ru], vals[ru]             # already sorted by (row, col)
    indptr = torch.zeros(nu + 1, dtype=torch.int64, device=rows.device)
    indptr[1:] = torch.cumsum(torch.bincount(r_u, minlength=nu), 0)
    ci = cols < ni
    r_i, c_i, v_i = rows[ci], cols[ci], vals[ci]
    order = torch.sort(c_i, stable=True).indices             # stable: ascending users inside an item
    cindptr = torch.zeros(ni + 1, dtype=torch.int64, device=rows.device)
    cindptr[1:] = torch.cumsum(torch.bincount(c_i, minlength=ni), 0)
    out = (indptr, c_u, v_u, cindptr, r_i[order], v_i[order])
    return tuple(t.cpu().numpy() for t in out)


# --------------------------------------------------------------------------------------
# B200 arm
# --------------------------------------------------------------------------------------
def b200_arm(args):
    import torch
    import torch.distributed as td
    import b200als
    from b200als.sparse import ratings_from_flat
    from b200als.synth import synth_ratings_torch

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.gpus != world and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        td.init_process_group("nccl", device_id=dev)
    b200als._lib.require_device()
    shard = b200als.Shard.from_env()
    wl = args.workload
    m, n, nnz, k = WORKLOADS[wl]
    K, W = args.steps, args.warmup

    rows, cols, vals = synth_ratings_torch(m, n, nnz, seed=DATA_SEED, device=dev)
    flat = rows * n + cols
    cfg = b200als.ALSConfig(core=b200als.CoreConfig(n_factors=k, n_iters=W + K, lambda_u=HYPER["lambda_u"],
                                                    lambda_v=HYPER["lambda_v"]),
                            biases=b200als.BiasesConfig(HYPER["lambda_bu"], HYPER["lambda_bi"]))

    # ---- resident arm: ratings already in HBM ---------------------------------------
    rt = ratings_from_flat(flat, vals, (m, n), dev, sorted_input=True)
    model = b200als.ALS(cfg)
    model.shard = shard
    model.gram_path = {"auto": 0, "ffma": 1, "umma": 2}[args.path]
    mu0 = float(vals.sum(dtype=torch.float64).item() / nnz)
    ses = model.session(rt, mu0, init="device", profile=True)
    for it in range(W):
        ses.iteration(it)
    torch.cuda.synchronize()
    ses.collect_phase_times()
    for p in ses.phase_ms:
        ses.phase_ms[p] = []
    shard.barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.25)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    shard.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e0.record()
    for it in range(W, W + K):
        ses.iteration(it)
    e1.record()
    torch.cuda.synchronize()
    shard.barrier()
    t1 = time.perf_counter()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop(t0, t1) if rank == 0 else None
    launches = ses.launches * K
    phases = {p: (float(np.mean(v)) if v else 0.0) for p, v in ses.collect_phase_times().items()}
    hist = ses.read_history(W + K)
    tms = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        td.all_reduce(tms, op=td.ReduceOp.MAX)
    ms = float(tms.item())
    ms_per_step = ms / K
    value = nnz / (ms_per_step * 1e-3)
    rmse_last = float(hist[-1, 0])
    del ses

    # ---- end-to-end arm: public fit_coo with pinned HOST buffers ----------------------
    flat_h = flat.cpu().pin_memory()
    vals_h = vals.cpu().pin_memory()
    del rows, cols, flat, vals, rt
    torch.cuda.empty_cache()
    cfg_e = b200als.ALSConfig(core=b200als.CoreConfig(n_factors=k, n_iters=K, lambda_u=HYPER["lambda_u"],
                                                      lambda_v=HYPER["lambda_v"]),
                              biases=b200als.BiasesConfig(HYPER["lambda_bu"], HYPER["lambda_bi"]))
    m2 = b200als.ALS(cfg_e)
    m2.shard = shard
    m2.gram_path = model.gram_path
    shard.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    m2.fit_coo(flat_h, vals_h, (m, n), tol=None, verbose=0, sorted_input=True)
    torch.cuda.synchronize()
    te = time.perf_counter() - t0
    tet = torch.tensor([te], dtype=torch.float64, device=dev)
    if world > 1:
        td.all_reduce(tet, op=td.ReduceOp.MAX)
    te = float(tet.item())
    e2e_value = nnz * K / te
    h2d = (nnz * 12 + (m + n) * k * 4) / K
    d2h = ((m + n) * (k + 1) * 8 + K * 48) / K

    if rank != 0:
        if world > 1:
            td.destroy_process_group()
        return
    peak, peak_kind = measured_peaks()
    # dominant kernel = the slower of the two half-step launches (same kernel, two data sets)
    rows_u, rows_i = (m + world - 1) // world, (n + world - 1) // world
    cand = {"user": algorithmic_bytes_half(nnz / world, rows_u, k), "item": algorithmic_bytes_half(nnz / world, rows_i, k)}
    dom = max(("user", "item"), key=lambda p: phases[p])
    ach = cand[dom] / (phases[dom] * 1e-3) / 1e9 if phases[dom] > 0 else 0.0
    b_iter = cand["user"] + cand["item"]
    # dram__bytes_read.sum + dram__bytes_write.sum of that launch from one `ncu --set full` capture (profiles/)
    traffic = NCU_DRAM_BYTES.get((wl, dom)) if world == 1 and args.path == "auto" else None
    out = {
        "metric": "ALS ratings/sec/iter (= nnz x ALS iterations/sec)", "value": value, "unit": "ratings/s",
        "iterations_per_sec": 1e3 / ms_per_step,
        "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD_NAMES[wl], "users": m, "items": n, "ratings": nnz, "k": k,
                   "model": "U/V + biases", "hyper": HYPER, "sharding": f"rows/{world}",
                   "l2_policy": "inputs larger than L2 (CSR+CSC %.1f GB)" % (nnz * 16 / 1e9),
                   "gram_path": args.path, "train_rmse_last": rmse_last},
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": "ratings/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "seconds_total": te, "iterations": K,
                "what": "ALS.fit_coo(pinned host flat_idx+vals) incl. H2D, CSR/CSC build, host RNG init, "
                        "K iterations, D2H of U,V,b_u,b_i"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": f"row_step ({dom} half-step)", "achieved": ach, "peak": peak,
                     "unit": "GB/s", "frac": ach / peak, "traffic": traffic, "peak_source": peak_kind,
                     "algorithmic_bytes_per_launch": cand[dom],
                     "iteration_frac": b_iter / (ms_per_step * 1e-3) / 1e9 / peak},
        "phase_ms": phases,
    }
    if world == 1 and not args.no_cpu:
        v, t, info = cpu_reference_arm(wl, steps=2, warmup=1, budget_s=args.cpu_budget)
        out["cpu_baseline"] = dict(value=v, **info)
    print(json.dumps(out), flush=True)
    if world > 1:
        td.destroy_process_group()


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    wl = args.workload
    m, n, nnz, k = WORKLOADS[wl]
    v, t, info = cpu_reference_arm(wl, steps=args.steps, warmup=args.warmup, budget_s=args.ref_budget)
    out = {
        "impl": "reference",
        "metric": "ALS ratings/sec/iter (= nnz x ALS iterations/sec)", "value": v, "unit": "ratings/s",
        "iterations_per_sec": v / nnz,
        "n_gpus": int(os.environ.get("WORLD_SIZE", "1")), "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD_NAMES[wl], "users": m, "items": n, "ratings": nnz, "k": k,
                   "model": "U/V + biases", "hyper": HYPER},
        "cpu_baseline": dict(value=v, **info),
        "e2e": {"value": v, "unit": "ratings/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c4", choices=sorted(WORKLOADS))
    ap.add_argument("--path", default="auto", choices=["auto", "ffma", "umma"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--cpu-budget", type=float, default=15.0)
    ap.add_argument("--ref-budget", type=float, default=90.0)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    if args.impl == "reference":
        reference_arm(args)
    else:
        if args.warmup < 3:
            print(f"note: warmup {args.warmup} < 3", file=sys.stderr)
        b200_arm(args)


if __name__ == "__main__":
    main()
