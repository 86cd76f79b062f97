#!/bin/bash
# round 2, GPU call 12 (N GPUs, default 4): sharded parity and C4 in both exchange modes
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
N=${1:-4}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
P=29711
run() { name=$1; shift; P=$((P+1)); timeout 600 "$@" > gpurun_out/c12_$name.log 2>&1; echo "$name exit $?"; grep -v "^\[W\|^W1\|^\*\*\*" gpurun_out/c12_$name.log | tail -1 | cut -c1-260; }
run n${N}_dist_check_peer $TR --master-port $P tools/dist_check.py
run n${N}_slab_check_peer $TR --master-port $P tools/slab_check.py
run n${N}_bench_c4_peer $TR --master-port $P bench.py --gpus $N --steps 10 --warmup 3
B200ALS_EXCHANGE=allgather run n${N}_bench_c4_allgather $TR --master-port $P bench.py --gpus $N --steps 10 --warmup 3
run n${N}_bench_reference $TR --master-port $P bench.py --impl reference --gpus $N --steps 2 --warmup 1
echo "call12 done"
