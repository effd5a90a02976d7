#!/bin/bash
# Build an A/B variant of the library with extra nvcc flags:  tools/build_variant.sh <tag> [-DNAME=VALUE ...]
# -> super_rag_b200/lib/libsrag_b200_<tag>.so (git-ignored, travels to the GPU box); load it with SRAG_B200_LIB=<path>.
set -eu
tag=$1; shift
cd "$(dirname "$0")/.."
out=super_rag_b200/lib/libsrag_b200_$tag.so
tmp=$(mktemp -d)
for f in super_rag_b200/csrc/*.cu; do
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xptxas -v "$@" -c "$f" -o "$tmp/$(basename "$f" .cu).o" 2> "$tmp/$(basename "$f" .cu).log" &
done
wait
nvcc -shared -o "$out" "$tmp"/*.o -gencode arch=compute_100a,code=sm_100a
grep -A1 "bm25_items_kernel" "$tmp/bm25.log" | grep -E "registers|spill" | head -4
rm -rf "$tmp"
echo "$out"
