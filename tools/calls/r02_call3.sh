#!/bin/bash
# round 2, GPU call 3 (1 GPU): GPU suite after the alignment fix, C2 launch list, bench
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -s -x > gpurun_out/c3_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/c3_pytest.log
tail -4 gpurun_out/c3_pytest.log
timeout 600 python bench.py > gpurun_out/c3_bench_c4.json 2> gpurun_out/c3_bench_c4.err; echo "bench exit $?"
timeout 300 python tools/e2e_breakdown.py c4 20 > gpurun_out/c3_e2e_breakdown_c4.json 2> gpurun_out/c3_e2e.err
timeout 200 python tools/c2_profile.py 20 > gpurun_out/c3_c2_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c3_c2_launches.csv python tools/c2_profile.py 20 > gpurun_out/c3_c2_ncu.log 2>&1
echo "call3 done"
