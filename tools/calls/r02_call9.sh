#!/bin/bash
# round 2, GPU call 9 (1 GPU): whole GPU suite (sweep kernel changed), default bench, C2 timing
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/c9_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/c9_pytest.log
timeout 600 python bench.py > gpurun_out/c9_bench_c4.json 2> gpurun_out/c9_bench_c4.err; echo "bench exit $?"; tail -2 gpurun_out/c9_bench_c4.err
timeout 200 python tools/c2_profile.py 20 > gpurun_out/c9_c2_plain.log 2>&1; cat gpurun_out/c9_c2_plain.log
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/c9_bench_reference.json 2> gpurun_out/c9_bench_reference.err; echo "reference exit $?"
echo "call9 done"
