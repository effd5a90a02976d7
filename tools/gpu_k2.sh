#!/bin/bash
# K2 round: BM25-related GPU tests, then C3 with the warp-granular phase 1 (variants) and without it.
#   gpurun --timeout 1500 -- 'bash tools/gpu_k2.sh [variant tags...]'
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_bm25_fuse.py tests/test_gpu_filter.py tests/test_shardio.py tests/test_gpu_nodes.py "tests/test_gpu_full_size.py::test_c3_two_phase_selection_matches_restatement" -m gpu -x -q > gpurun_out/pytest_k2.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_k2.log
run() {  # tag, SRAG_BM25_SUBKERNEL, lib
  echo "=== $1"
  SRAG_BM25_SUBKERNEL=$2 SRAG_B200_LIB=$3 timeout 600 python tools/bench_hybrid.py --docs 10000000 --skip-c4 --check 4 2>gpurun_out/c3_$1.err | tee gpurun_out/c3_$1.jsonl | grep -o '"ms_per_batch": [0-9.]*\|"frac": [0-9.]*\|"restatement_check_ok": [a-z]*'
  tail -2 gpurun_out/c3_$1.err
}
run sub 1 ""
for tag in "$@"; do run sub_$tag 1 "$PWD/super_rag_b200/lib/libsrag_b200_$tag.so"; done
run items 0 ""
