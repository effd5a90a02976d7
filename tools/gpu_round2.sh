#!/bin/bash
# One gpurun call: GPU parity tests, smoke, bench, then C3 (BM25) with ascending vs bank-ordered postings.
#   gpurun --timeout 1500 -- 'bash tools/gpu_round2.sh'
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json
SRAG_BM25_BANK_ORDER=0 timeout 600 python tools/bench_hybrid.py --docs 10000000 --skip-c4 --check 2 > gpurun_out/c3_plain.jsonl 2> gpurun_out/c3_plain.err; echo "c3 plain rc=$?"; cut -c1-700 gpurun_out/c3_plain.jsonl
timeout 600 python tools/bench_hybrid.py --docs 10000000 --skip-c4 --check 2 > gpurun_out/c3_bank.jsonl 2> gpurun_out/c3_bank.err; echo "c3 bank rc=$?"; cut -c1-700 gpurun_out/c3_bank.jsonl
tail -3 gpurun_out/c3_bank.err
