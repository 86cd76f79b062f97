#!/bin/bash
# round 2, GPU call 10 (1 GPU): GPU suite (refit fixture test), then per-section solver cycles from a TIMING=1 build made in a scratch copy
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
OUT=$PWD/gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/c10_pytest.log 2>&1; echo "pytest exit $?"; tail -2 $OUT/c10_pytest.log
rm -rf /tmp/timing && mkdir -p /tmp/timing && cp -r bench.py b200als collaborative-filtering_b200 include tools tests oracle /tmp/timing/ && cd /tmp/timing
touch collaborative-filtering_b200/csrc/row_step.cu
(time make -C collaborative-filtering_b200/csrc -j8 TIMING=1) > $OUT/c10_timing_build.log 2>&1; echo "build exit $?"
timeout 300 python tools/solver_timing.py c4 user > $OUT/c10_solver_timing_c4_user.txt 2>&1; tail -4 $OUT/c10_solver_timing_c4_user.txt
timeout 300 python tools/solver_timing.py c3 user > $OUT/c10_solver_timing_c3_user.txt 2>&1; tail -2 $OUT/c10_solver_timing_c3_user.txt
echo "call10 done"
