#!/bin/bash
# One gpurun call: GPU parity tests, smoke, bench, K1 batch sweep, ncu launch list + full capture of K1.
#   gpurun --timeout 1500 -- 'bash tools/gpu_round.sh'
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
nproc > gpurun_out/nproc.txt; lscpu | head -20 >> gpurun_out/nproc.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/bench_ref.json
timeout 600 python tools/perf_topk.py --batches 1024,512,256,128,64,8,1 > gpurun_out/perf_topk.log 2>&1; cat gpurun_out/perf_topk.log
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'topk|ingest|scale_bf16|merge|bm25|fuse' -c 60 \
    --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:topk_gemm_kernel -s 3 -c 2 \
    -o gpurun_out/prof_topk $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la gpurun_out
# C3 / C4 lines, the K2 launch list and a full capture of the BM25 phase-1 kernel (bm25_sub_kernel)
timeout 600 python tools/bench_hybrid.py --docs 10000000 --check 4 > gpurun_out/c3_c4.jsonl 2> gpurun_out/c3_c4.err; cat gpurun_out/c3_c4.jsonl | cut -c1-600
CMD2="python tools/bench_hybrid.py --docs 4000000 --skip-c4 --check 0 --steps 1 --warmup 0"
$CMD2 > gpurun_out/plain3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:bm25_sub_kernel -s 1 -c 1 -f -o gpurun_out/prof_bm25 $CMD2 > gpurun_out/ncu_bm25.log 2>&1
echo "ncu bm25 rc=$?"
CMD3="python tools/bench_hybrid.py --docs 10000000 --skip-c4 --check 0 --steps 2 --warmup 1"
$CMD3 > gpurun_out/plain4.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'bm25|finalize' -c 40 --csv --log-file gpurun_out/c3_launches.csv $CMD3 > gpurun_out/ncu_c3_launches.log 2>&1
echo "ncu c3 launches rc=$?"
