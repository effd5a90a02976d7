#!/bin/bash
# K2 A/B: C3 timing for a list of library variants (tags built with tools/build_variant.sh; "default" = the in-tree lib).
#   gpurun --timeout 900 -- 'bash tools/exp_bm25.sh default d3w24'
set -u
mkdir -p gpurun_out
for tag in "$@"; do
  lib=""; [ "$tag" != "default" ] && lib="$PWD/super_rag_b200/lib/libsrag_b200_$tag.so"
  echo "=== $tag"
  SRAG_B200_LIB=$lib timeout 600 python tools/bench_hybrid.py --docs ${DOCS:-10000000} --skip-c4 --check 2 --steps ${STEPS:-5} 2>gpurun_out/exp_bm25_$tag.err | grep -o '"ms_per_batch": [0-9.]*\|"restatement_check_ok": [a-z]*'
  tail -1 gpurun_out/exp_bm25_$tag.err
done
