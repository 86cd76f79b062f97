#!/bin/bash
# round 2, GPU call 27 (N GPUs, default 2): the final build on the sharded path — N-GPU == 1-GPU parity (dist_check),
# then the default bench line at N GPUs (fused peer-store exchange, the default)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
P=29711
run() { name=$1; shift; P=$((P+1)); timeout 600 "$@" > gpurun_out/c27_$name.log 2>&1; echo "$name exit $?"; grep -v "^\[W\|^W1\|^\*\*\*" gpurun_out/c27_$name.log | tail -1 | cut -c1-400; }
run n${N}_dist_check $TR --master-port $P tools/dist_check.py
run n${N}_bench_c4 $TR --master-port $P bench.py --gpus $N --steps 10 --warmup 3
echo "call27 done"
