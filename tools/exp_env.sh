#!/bin/bash
# C3 timing under a list of environment settings:  gpurun -- 'bash tools/exp_env.sh "A=1" "A=2 B=3"'
set -u
mkdir -p gpurun_out
for cfg in "$@"; do
  echo "=== $cfg"
  env $cfg timeout 600 python tools/bench_hybrid.py --docs ${DOCS:-10000000} --skip-c4 --check 4 2>gpurun_out/exp_env.err | grep -o '"ms_per_batch": [0-9.]*\|"frac": [0-9.]*\|"restatement_check_ok": [a-z]*'
  tail -1 gpurun_out/exp_env.err
done
