#!/bin/bash
# round 2, GPU call 11 (1 GPU): whole GPU suite + smoke after the single-allocation replicas
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/c11_pytest.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/c11_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c11_smoke.log 2>&1; echo "smoke exit $?"; tail -4 gpurun_out/c11_smoke.log
echo "call11 done"
