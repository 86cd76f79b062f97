#!/bin/bash
# round 2, GPU call 2 (1 GPU): full GPU suite with the float64 path / fused sweep / coalesced slots, default bench, scaled C5 on one GPU
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -s > gpurun_out/c2_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/c2_pytest.log
tail -4 gpurun_out/c2_pytest.log
timeout 600 python bench.py > gpurun_out/c2_bench_c4.json 2> gpurun_out/c2_bench_c4.err; echo "bench exit $?"
B200ALS_GRAPH_SWEEP=levels timeout 200 python - > gpurun_out/c2_c2_levels.json 2>> gpurun_out/c2_bench_c4.err <<'PY'
import json, sys, torch
sys.path.insert(0, ".")
import b200als, bench
print(json.dumps(bench.c2_run(b200als, torch.device("cuda", 0), 20, with_cpu=False)))
PY
timeout 400 python bench.py --workload c5 --c5-scale 0.02 --steps 5 --warmup 3 > gpurun_out/c2_c5_scaled.json 2> gpurun_out/c2_c5.err; echo "c5 exit $?"; tail -3 gpurun_out/c2_c5.err
timeout 300 python tools/e2e_breakdown.py c4 20 > gpurun_out/c2_e2e_breakdown_c4.json 2> gpurun_out/c2_e2e.err
echo "call2 done"
