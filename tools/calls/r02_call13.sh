#!/bin/bash
# round 2, GPU call 13 (1 GPU): solver variant — parity tests of the tcgen05 paths, then C4 / C3 timings
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edges.py tests/test_zz_scale_parity.py -m gpu -x -q > gpurun_out/c13_pytest.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/c13_pytest.log
for wl in c4 c3 c4k64; do
  timeout 300 python bench.py --workload $wl --no-cpu --no-secondary --steps 10 --warmup 3 > gpurun_out/c13_bench_$wl.json 2> gpurun_out/c13_bench_$wl.err; echo "$wl exit $?"
  python - <<PY
import json
r=json.loads(open("gpurun_out/c13_bench_$wl.json").read().strip().splitlines()[-1])
print("$wl", round(r["ms_per_step"],3), {k:round(v,3) for k,v in r["phase_ms"].items() if v>0.01}, r["run"]["train_rmse_last"])
PY
done
echo "call13 done"
