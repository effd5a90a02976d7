#!/bin/bash
# Full ncu capture of K1 at the headline shape (one launch), after a plain run of the same command.
#   gpurun --timeout 900 -- 'bash tools/ncu_k1.sh [ENV=VAL ...]'
set -u
mkdir -p gpurun_out
CMD="python tools/perf_topk.py --batches ${BATCHES:-1024} --iters 1"
env "$@" $CMD > gpurun_out/ncu_k1_plain.log 2>&1 &&
env "$@" ncu --set full --clock-control none --import-source on -k regex:topk_gemm_kernel -s 2 -c 1 \
    -f -o gpurun_out/prof_k1 $CMD > gpurun_out/ncu_k1.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_k1.log
