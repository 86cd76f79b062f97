#!/bin/bash
# round 2, GPU call 1: full GPU test-suite, default bench (c4 + secondary), split-piece sweep, e2e breakdown, launch list
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/c1_smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/c1_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/c1_pytest.log
tail -5 gpurun_out/c1_pytest.log
timeout 600 python bench.py > gpurun_out/c1_bench_c4.json 2> gpurun_out/c1_bench_c4.err; echo "bench exit $?"
# accumulation-chain length against accuracy and speed (C4, heaviest item rows)
for cfg in "-1 0 nosplit" "4096 2048 p2048"; do
  set -- $cfg
  B200ALS_SPLIT_ROWS=$1 B200ALS_SPLIT_PIECE=$2 B200ALS_PARITY_TAG=_$3 timeout 400 python -m pytest tests/test_zz_scale_parity.py -m gpu -q -s -k c4 > gpurun_out/c1_parity_$3.log 2>&1
done
for cfg in "4096 2048 p2048" "65536 32768 p32768"; do
  set -- $cfg
  B200ALS_SPLIT_ROWS=$1 B200ALS_SPLIT_PIECE=$2 timeout 300 python bench.py --no-cpu --no-secondary --steps 5 > gpurun_out/c1_bench_c4_$3.json 2>> gpurun_out/c1_bench_c4.err
done
timeout 300 python tools/e2e_breakdown.py c4 10 > gpurun_out/c1_e2e_breakdown_c4.json 2> gpurun_out/c1_e2e.err
timeout 300 python bench.py --no-cpu --no-secondary --steps 2 --warmup 1 > gpurun_out/c1_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/c1_launches.csv python bench.py --no-cpu --no-secondary --steps 2 --warmup 1 > gpurun_out/c1_ncu.log 2>&1
echo "call1 done"
