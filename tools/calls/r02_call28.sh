#!/bin/bash
# round 2, GPU call 28 (1 GPU): A/B of the diagonal step's vector result stores (B200ALS_DIAG_VEC) on C4, C3;
# then the default bench of the current library (e2e leg = median of three fits)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
P=collaborative-filtering_b200
cp $P/libb200als.so /tmp/base.so
for v in cur vec cur vec; do
  cp $P/variants/libb200als_$v.so $P/libb200als.so
  for wl in c4 c3; do
    timeout 300 python bench.py --workload $wl --steps 10 --warmup 3 --no-secondary --no-cpu --no-parity --no-e2e > gpurun_out/c28_${v}_${wl}.json 2> gpurun_out/c28_${v}_${wl}.err
    python - <<PY
import json
try:
    r = json.loads(open("gpurun_out/c28_${v}_${wl}.json").read().strip().splitlines()[-1])
    print("$v $wl", round(r["ms_per_step"], 3), {k: round(x, 3) for k, x in r["phase_ms"].items() if x > 0.01}, r["run"]["train_rmse_last"])
except Exception as e:
    print("$v $wl failed", e)
PY
  done
done
cp /tmp/base.so $P/libb200als.so
timeout 600 python bench.py > gpurun_out/c28_bench_c4.json 2> gpurun_out/c28_bench_c4.err; echo "bench exit $?"
python - <<PY
import json
r = json.loads(open("gpurun_out/c28_bench_c4.json").read().strip().splitlines()[-1])
print("c4", round(r["ms_per_step"], 2), "e2e", round(r["e2e"]["value"] / 1e9, 3), r["e2e"]["seconds_all"], "traffic", r["roofline"]["traffic_source"])
PY
echo "call28 done"
