#!/bin/bash
# round 2, GPU call 4 (1 GPU): whole GPU suite, default bench, C2 timing
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -s > gpurun_out/c4_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/c4_pytest.log
tail -12 gpurun_out/c4_pytest.log
timeout 600 python bench.py > gpurun_out/c4_bench_c4.json 2> gpurun_out/c4_bench_c4.err; echo "bench exit $?"
timeout 200 python tools/c2_profile.py 20 > gpurun_out/c4_c2_plain.log 2>&1
B200ALS_GRAPH_BUILD=device timeout 200 python tools/c2_profile.py 20 > gpurun_out/c4_c2_devgraph.log 2>&1
echo "call4 done"
