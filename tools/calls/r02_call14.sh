#!/bin/bash
# round 2, GPU call 14 (2 GPUs): final checks — sharded parity incl. the collective LinAlgError path, N=2 bench, then on GPU 0
# the whole GPU suite (-x, as the round end runs it), smoke, the default bench and the reference arm
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29811 tools/dist_check.py > gpurun_out/c14_dist_check.log 2>&1; echo "dist_check exit $?"; grep -v "^\[W\|^W1\|^\*\*\*" gpurun_out/c14_dist_check.log | tail -2
B200ALS_EXCHANGE=allgather timeout 400 $TR --master-port 29812 tools/dist_check.py > gpurun_out/c14_dist_check_allgather.log 2>&1; echo "dist_check allgather exit $?"
timeout 400 $TR --master-port 29813 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/c14_bench_n2.log 2>&1; echo "bench n2 exit $?"
export CUDA_VISIBLE_DEVICES=0
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/c14_pytest.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/c14_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c14_smoke.log 2>&1; echo "smoke exit $?"
(time timeout 900 python bench.py) > gpurun_out/c14_bench_c4.json 2> gpurun_out/c14_bench_c4.err; echo "bench exit $?"; tail -4 gpurun_out/c14_bench_c4.err
(time timeout 900 python bench.py --impl reference) > gpurun_out/c14_bench_reference.json 2> gpurun_out/c14_bench_reference.err; echo "reference exit $?"; tail -4 gpurun_out/c14_bench_reference.err
echo "call14 done"
