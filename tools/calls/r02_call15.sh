#!/bin/bash
# round 2, GPU call 15 (8 GPUs): the final code at N = 8 — sharded parity (incl. the collective LinAlgError path), C4, full C5
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
N=8
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
P=29911
run() { name=$1; shift; P=$((P+1)); timeout 600 "$@" > gpurun_out/c15_$name.log 2>&1; echo "$name exit $?"; grep -v "^\[W\|^W1\|^\*\*\*" gpurun_out/c15_$name.log | tail -1 | cut -c1-200; }
run n8_dist_check $TR --master-port $P tools/dist_check.py
run n8_bench_c4 $TR --master-port $P bench.py --gpus $N --steps 10 --warmup 3
run n8_c5_full $TR --master-port $P bench.py --gpus $N --workload c5 --steps 5 --warmup 3 --no-cpu
echo "call15 done"
