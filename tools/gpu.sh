#!/bin/bash
# usage: tools/gpu.sh "<make args>" "<gpurun timeout>" '<command>' — builds first, runs on the GPU box only if the build succeeded
cd /root/repo
touch collaborative-filtering_b200/csrc/row_step.cu
if ! make -C collaborative-filtering_b200/csrc -j8 $1 > /tmp/build.log 2>&1; then grep -i "error" /tmp/build.log | head -5; echo "BUILD FAILED"; exit 1; fi
/usr/local/graft/bin/gpurun --timeout "$2" -- "$3" 2>&1 | tail -${4:-8}
