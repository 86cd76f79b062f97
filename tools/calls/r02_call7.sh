#!/bin/bash
# round 2, GPU call 7 (8 GPUs): sharded parity + C4 in both exchange modes, the full C5 shape (10 M x 1 M, 2e9 ratings, k = 128)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
P=29611
run() { name=$1; shift; P=$((P+1)); timeout 600 "$@" > gpurun_out/c7_$name.log 2>&1; echo "$name exit $?"; grep -v "^\[W\|^W1\|^\*\*\*" gpurun_out/c7_$name.log | tail -1 | cut -c1-260; }
B200ALS_EXCHANGE=peer run n${N}_dist_check_peer $TR --master-port $P tools/dist_check.py
run n${N}_bench_c4_allgather $TR --master-port $P bench.py --gpus $N --steps 10 --warmup 3
B200ALS_EXCHANGE=peer run n${N}_bench_c4_peer $TR --master-port $P bench.py --gpus $N --steps 10 --warmup 3
B200ALS_EXCHANGE=peer run n${N}_c5_full_peer $TR --master-port $P bench.py --gpus $N --workload c5 --steps 5 --warmup 3
run n${N}_c5_full_allgather $TR --master-port $P bench.py --gpus $N --workload c5 --steps 5 --warmup 3 --no-cpu
echo "call7 done"
