#!/bin/bash
# round 2, GPU call 6 (1 GPU): the no-alloc entry point's test alone in a fresh process, then the whole GPU suite (it runs last)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zzz_no_alloc.py -m gpu -x -q > gpurun_out/c6_noalloc_alone.log 2>&1; echo "alone exit $?" | tee -a gpurun_out/c6_noalloc_alone.log
tail -3 gpurun_out/c6_noalloc_alone.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/c6_pytest_all.log 2>&1; echo "all exit $?" | tee -a gpurun_out/c6_pytest_all.log
tail -4 gpurun_out/c6_pytest_all.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c6_smoke.log 2>&1; echo "smoke exit $?"; tail -4 gpurun_out/c6_smoke.log
echo "call6 done"
