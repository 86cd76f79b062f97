#!/bin/bash
# round 2, GPU call 18 (1 GPU): default bench with the C1 record and the captured traffic figure; then the launch list
# of the final build (ncu time-only pass, after the same command exited 0 without ncu)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python bench.py > gpurun_out/c18_bench_c4.json 2> gpurun_out/c18_bench_c4.err; echo "bench exit $?"
python - <<PY
import json
r = json.loads(open("gpurun_out/c18_bench_c4.json").read().strip().splitlines()[-1])
print("c4", round(r["ms_per_step"], 2), "traffic", r["roofline"]["traffic"], "frac", round(r["roofline"]["frac"], 3))
for k, v in r["secondary"].items():
    print(k, {a: v[a] for a in ("ms_per_step", "paths_taken", "train_rmse_last") if a in v})
PY
timeout 300 python bench.py --steps 2 --warmup 3 --no-secondary --no-cpu --no-parity > gpurun_out/c18_short.json 2> gpurun_out/c18_short.err; rc=$?; echo "short bench exit $rc"
if [ $rc -eq 0 ]; then
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/c18_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-secondary --no-cpu --no-parity > gpurun_out/c18_ncu.log 2>&1; echo "ncu exit $?"
fi
echo "call18 done"
