#!/bin/bash
# round 2, GPU call 23 (1 GPU): (1) A/B of the FFMA path's Gram tile on FFMA2 (C3 shape at k = 96 forced onto the FFMA path, C1);
# (2) the new default build (FFMA2 + tapered diagonal steps): GPU suite with -x, smoke, default bench
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
P=collaborative-filtering_b200
cp $P/libb200als.so /tmp/base.so
for v in gramscalar gramf2 gramscalar gramf2; do
  cp $P/variants/libb200als_$v.so $P/libb200als.so
  for wl in "c3k96 --path ffma" "c1"; do
    tag=$(echo $wl | cut -d' ' -f1)
    timeout 300 python bench.py --workload $wl --steps 10 --warmup 3 --no-secondary --no-cpu --no-parity --no-e2e > gpurun_out/c23_${v}_${tag}.json 2> gpurun_out/c23_${v}_${tag}.err
    python - <<PY
import json
try:
    r = json.loads(open("gpurun_out/c23_${v}_${tag}.json").read().strip().splitlines()[-1])
    print("$v $tag", round(r["ms_per_step"], 4), {k: round(x, 4) for k, x in r["phase_ms"].items() if x > 0.01}, r["run"]["paths_taken"], r["run"]["train_rmse_last"])
except Exception as e:
    print("$v $tag failed", e)
PY
  done
done
cp /tmp/base.so $P/libb200als.so
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/c23_pytest.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/c23_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c23_smoke.log 2>&1; echo "smoke exit $?"; tail -4 gpurun_out/c23_smoke.log
timeout 600 python bench.py > gpurun_out/c23_bench_c4.json 2> gpurun_out/c23_bench_c4.err; echo "bench exit $?"
python - <<PY
import json
r = json.loads(open("gpurun_out/c23_bench_c4.json").read().strip().splitlines()[-1])
print("c4", round(r["ms_per_step"], 2), {k: round(v, 2) for k, v in r["phase_ms"].items() if v > 0.01}, "e2e", round(r["e2e"]["value"] / 1e9, 3), "frac", round(r["roofline"]["frac"], 3))
for k, v in r["secondary"].items():
    print(k, {a: v[a] for a in ("ms_per_step", "seconds_fit_predict", "train_rmse_last") if a in v})
PY
echo "call23 done"
