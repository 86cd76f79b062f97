#!/bin/bash
# round 2, GPU call 26 (1 GPU): the build with FFMA2 + taper + right-looking substitution + packed panel split:
# GPU suite with -x, smoke, default bench; then ncu --set full of the two C4 launches (after the same command exited 0)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/c26_pytest.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/c26_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c26_smoke.log 2>&1; echo "smoke exit $?"; tail -4 gpurun_out/c26_smoke.log
timeout 600 python bench.py > gpurun_out/c26_bench_c4.json 2> gpurun_out/c26_bench_c4.err; echo "bench exit $?"
python - <<PY
import json
r = json.loads(open("gpurun_out/c26_bench_c4.json").read().strip().splitlines()[-1])
print("c4", round(r["ms_per_step"], 2), {k: round(v, 2) for k, v in r["phase_ms"].items() if v > 0.01}, "e2e", round(r["e2e"]["value"] / 1e9, 3), "frac", round(r["roofline"]["frac"], 3))
for k, v in r["secondary"].items():
    print(k, {a: v[a] for a in ("ms_per_step", "seconds_fit_predict", "train_rmse_last") if a in v})
PY
CMD="python bench.py --no-cpu --no-secondary --no-e2e --steps 2 --warmup 1"
timeout 300 $CMD > gpurun_out/c26_plain_c4.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:row_step_umma -s 2 -c 2 -f -o gpurun_out/c26_c4_umma $CMD > gpurun_out/c26_ncu_c4.log 2>&1
echo "ncu c4 exit $?"; tail -2 gpurun_out/c26_ncu_c4.log
ls -la gpurun_out/*.ncu-rep
echo "call26 done"
