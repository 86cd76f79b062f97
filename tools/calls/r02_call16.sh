#!/bin/bash
# round 2, GPU call 16 (N GPUs, default 2): sharded upload of the rating list — parity and end-to-end
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
P=30011
run() { name=$1; shift; P=$((P+1)); timeout 600 "$@" > gpurun_out/c16_$name.log 2>&1; echo "$name exit $?"; grep -v "^\[W\|^W1\|^\*\*\*" gpurun_out/c16_$name.log | tail -1 | cut -c1-200; }
run n${N}_dist_check $TR --master-port $P tools/dist_check.py
run n${N}_bench_c4 $TR --master-port $P bench.py --gpus $N --steps 10 --warmup 3
B200ALS_SHARDED_UPLOAD=0 run n${N}_bench_c4_fullupload $TR --master-port $P bench.py --gpus $N --steps 10 --warmup 3 --no-parity
python - <<PY
import json
def last_json(f):
    for line in reversed(open(f).read().strip().splitlines()):
        if line.startswith('{'): return json.loads(line)
for f in ("gpurun_out/c16_n${N}_bench_c4.log", "gpurun_out/c16_n${N}_bench_c4_fullupload.log"):
    r = last_json(f); print(f, round(r["ms_per_step"], 3), "e2e", round(r["e2e"]["value"] / 1e9, 3), round(r["e2e"]["seconds_total"], 4))
PY
echo "call16 done"
