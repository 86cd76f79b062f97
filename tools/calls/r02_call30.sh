#!/bin/bash
# round 2, GPU call 30 (1 GPU): the final build (lean substitution at k = 128 on top of call 26) — GPU suite with -x, smoke, default bench
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/c30_pytest.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/c30_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c30_smoke.log 2>&1; echo "smoke exit $?"; tail -4 gpurun_out/c30_smoke.log
timeout 600 python bench.py > gpurun_out/c30_bench_c4.json 2> gpurun_out/c30_bench_c4.err; echo "bench exit $?"
python - <<PY
import json
r = json.loads(open("gpurun_out/c30_bench_c4.json").read().strip().splitlines()[-1])
print("c4", round(r["ms_per_step"], 2), {k: round(v, 2) for k, v in r["phase_ms"].items() if v > 0.01}, "e2e", round(r["e2e"]["value"] / 1e9, 3), "c2", r["secondary"]["c2"]["seconds_fit_predict"])
PY
echo "call30 done"
