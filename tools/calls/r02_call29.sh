#!/bin/bash
# round 2, GPU call 29 (1 GPU): A/B of the lean right-looking substitution (B200ALS_SUBST_LEAN) on C4, C3
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
P=collaborative-filtering_b200
cp $P/libb200als.so /tmp/base.so
for v in cur lean cur lean; do
  cp $P/variants/libb200als_$v.so $P/libb200als.so
  for wl in c4 c3; do
    timeout 300 python bench.py --workload $wl --steps 10 --warmup 3 --no-secondary --no-cpu --no-parity --no-e2e > gpurun_out/c29_${v}_${wl}.json 2> gpurun_out/c29_${v}_${wl}.err
    python - <<PY
import json
try:
    r = json.loads(open("gpurun_out/c29_${v}_${wl}.json").read().strip().splitlines()[-1])
    print("$v $wl", round(r["ms_per_step"], 3), {k: round(x, 3) for k, x in r["phase_ms"].items() if x > 0.01}, r["run"]["train_rmse_last"])
except Exception as e:
    print("$v $wl failed", e)
PY
  done
done
cp /tmp/base.so $P/libb200als.so
echo "call29 done"
