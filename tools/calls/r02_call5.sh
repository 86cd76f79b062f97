#!/bin/bash
# round 2, GPU call 5 (2 GPUs): sharded parity (fused peer-store exchange, slab storage, fold-parallel CV), C4 at N = 2
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29511 tools/dist_check.py > gpurun_out/c5_dist_check.log 2>&1; echo "dist_check exit $?"; tail -3 gpurun_out/c5_dist_check.log
timeout 300 $TR --master-port 29512 tools/slab_check.py > gpurun_out/c5_slab_check.log 2>&1; echo "slab_check exit $?"; tail -3 gpurun_out/c5_slab_check.log
timeout 300 $TR --master-port 29513 tools/cv_parallel_check.py > gpurun_out/c5_cv_parallel.log 2>&1; echo "cv exit $?"; tail -2 gpurun_out/c5_cv_parallel.log
timeout 400 $TR --master-port 29514 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/c5_bench_c4_n2.json 2> gpurun_out/c5_bench_c4_n2.err; echo "bench n2 exit $?"; tail -3 gpurun_out/c5_bench_c4_n2.err
B200ALS_EXCHANGE=allgather timeout 400 $TR --master-port 29515 bench.py --gpus 2 --steps 10 --warmup 3 --no-parity > gpurun_out/c5_bench_c4_n2_allgather.json 2> gpurun_out/c5_bench_c4_n2_allgather.err; echo "bench n2 allgather exit $?"
timeout 400 $TR --master-port 29516 bench.py --gpus 2 --workload c5 --c5-scale 0.05 --steps 5 --warmup 3 --no-cpu > gpurun_out/c5_c5_scaled_n2.json 2> gpurun_out/c5_c5_scaled_n2.err; echo "c5 n2 exit $?"; tail -3 gpurun_out/c5_c5_scaled_n2.err
echo "call5 done"
