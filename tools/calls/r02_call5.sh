#!/bin/bash
# round 2, GPU call 5 (2 GPUs): sharded parity in both exchange modes (packed all-gather = default, fused peer stores = opt-in),
# slab storage, fold-parallel CV, C4 at N = 2 in both modes, scaled C5
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
P=29511
run() { name=$1; shift; P=$((P+1)); timeout 400 "$@" > gpurun_out/c5_$name.log 2>&1; echo "$name exit $?"; grep -v "^\[W\|^W1\|^\*\*\*" gpurun_out/c5_$name.log | tail -2 | cut -c1-300; }
run dist_check_allgather $TR --master-port $P tools/dist_check.py
B200ALS_EXCHANGE=peer run dist_check_peer $TR --master-port $P tools/dist_check.py
run slab_check_allgather $TR --master-port $P tools/slab_check.py
B200ALS_EXCHANGE=peer run slab_check_peer $TR --master-port $P tools/slab_check.py
run cv_parallel $TR --master-port $P tools/cv_parallel_check.py
run bench_c4_n2_allgather $TR --master-port $P bench.py --gpus 2 --steps 10 --warmup 3
B200ALS_EXCHANGE=peer run bench_c4_n2_peer $TR --master-port $P bench.py --gpus 2 --steps 10 --warmup 3
run c5_scaled_n2 $TR --master-port $P bench.py --gpus 2 --workload c5 --c5-scale 0.05 --steps 5 --warmup 3 --no-cpu
echo "call5 done"
