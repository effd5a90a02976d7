#!/bin/bash
# One gpurun call (1 GPU): GPU parity tests, smoke, both bench arms, K1 batch sweeps, ncu launch lists and full captures of
# K1 (C2 shape) and K2 (the C3 index itself, 10 M documents). Every ncu run follows a plain run of the same command.
#   gpurun --timeout 2400 -- 'bash tools/gpu_round.sh'
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
nproc > gpurun_out/nproc.txt; lscpu | head -20 >> gpurun_out/nproc.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cut -c1-600 gpurun_out/bench_ref.json
timeout 600 python tools/perf_topk.py --batches 1024,512,256,128,64,8,1 > gpurun_out/batch_sweep.jsonl 2>&1; cat gpurun_out/batch_sweep.jsonl | cut -c1-140
timeout 600 python tools/perf_topk.py --batches 1024,256,128 --k 100 > gpurun_out/batch_sweep_k100.jsonl 2>&1; cat gpurun_out/batch_sweep_k100.jsonl | cut -c1-140
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-hybrid --no-c1"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'topk|ingest|scale_bf16|merge|bm25|fuse' -c 60 \
    --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:topk_gemm_kernel -s 3 -c 1 -f \
    -o gpurun_out/prof_topk $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu K1 full rc=$?"
# K2 on the C3 index itself: launch list, then a full capture of the MAIN pass (the 2nd bm25_sub_kernel launch of a search)
CMD3="python tools/bench_hybrid.py --docs 10000000 --skip-c4 --skip-stopwords --check 0 --steps 2 --warmup 1"
$CMD3 > gpurun_out/plain3.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'bm25|finalize' -c 40 --csv --log-file gpurun_out/c3_launches.csv $CMD3 > gpurun_out/ncu_c3_launches.log 2>&1
echo "ncu c3 launches rc=$?"
$CMD3 > gpurun_out/plain4.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:bm25_sub_kernel -s 3 -c 1 -f -o gpurun_out/prof_bm25 $CMD3 > gpurun_out/ncu_bm25.log 2>&1
echo "ncu K2 full rc=$?"
ls -la gpurun_out
