#!/bin/bash
# round 2, GPU call 8 (1 GPU): ncu --set full of the dominant kernel (C4 user + item half-step) and of the C2 graph sweep,
# each only after the same command has exited 0 without ncu
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
CMD="python bench.py --no-cpu --no-secondary --steps 2 --warmup 1"
timeout 300 $CMD > gpurun_out/c8_plain_c4.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:row_step_umma -s 2 -c 2 -f -o gpurun_out/c8_c4_umma $CMD > gpurun_out/c8_ncu_c4.log 2>&1
echo "ncu c4 exit $?"; tail -2 gpurun_out/c8_ncu_c4.log
timeout 200 python tools/c2_profile.py 20 > gpurun_out/c8_plain_c2.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:graph_sweep -s 25 -c 1 -f -o gpurun_out/c8_c2_sweep python tools/c2_profile.py 20 > gpurun_out/c8_ncu_c2.log 2>&1
echo "ncu c2 exit $?"; tail -2 gpurun_out/c8_ncu_c2.log; cat gpurun_out/c8_plain_c2.log
ls -la gpurun_out/*.ncu-rep
echo "call8 done"
