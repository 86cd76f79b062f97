#!/bin/bash
# round 2, GPU call 21 (1 GPU): FFMA / FFMA2 micro-benchmark, then A/B of the solver's rank-4 updates on packed FFMA2
# (B200ALS_FFMA2) with and without the tapered diagonal steps, on C4, C3, C4 at k = 64
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 60 tools/ffma2_bench.bin > gpurun_out/c21_ffma2_bench.txt 2>&1; echo "microbench exit $?"; cat gpurun_out/c21_ffma2_bench.txt
P=collaborative-filtering_b200
cp $P/libb200als.so /tmp/base.so
for v in plain f2 f2t plain f2 f2t; do
  cp $P/variants/libb200als_$v.so $P/libb200als.so
  for wl in c4 c3 c4k64; do
    timeout 300 python bench.py --workload $wl --steps 10 --warmup 3 --no-secondary --no-cpu --no-parity --no-e2e > gpurun_out/c21_${v}_${wl}.json 2> gpurun_out/c21_${v}_${wl}.err
    python - <<PY
import json
try:
    r = json.loads(open("gpurun_out/c21_${v}_${wl}.json").read().strip().splitlines()[-1])
    print("$v $wl", round(r["ms_per_step"], 3), {k: round(x, 3) for k, x in r["phase_ms"].items() if x > 0.01}, r["run"]["train_rmse_last"])
except Exception as e:
    print("$v $wl failed", e)
PY
  done
done
cp /tmp/base.so $P/libb200als.so
echo "call21 done"
