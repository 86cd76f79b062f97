#!/bin/bash
# round 2, GPU call 19 (1 GPU): launch list of the final build restricted to the kernels of the iteration (the synthetic-data
# generation in torch alone exceeds a 400-launch cap); the same command first exits 0 without ncu
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-secondary --no-cpu --no-parity"
timeout 300 $CMD > gpurun_out/c19_short.json 2> gpurun_out/c19_short.err; rc=$?; echo "short bench exit $rc"
if [ $rc -eq 0 ]; then
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k "regex:row_step|split_|finalize_iter|compose|stats_" -c 600 --csv \
    --log-file gpurun_out/c19_launches.csv $CMD > gpurun_out/c19_ncu.log 2>&1; echo "ncu exit $?"
fi
echo "call19 done"
