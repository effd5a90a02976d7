#!/bin/bash
# K1 experiment: timing for a list of env settings, then a light ncu metrics pass per setting.
#   gpurun --timeout 900 -- 'bash tools/exp_k1.sh "SRAG_TOPK_WINDOW=0" "SRAG_TOPK_WINDOW=3"'
set -u
mkdir -p gpurun_out
: > gpurun_out/exp_k1.log
M=dram__bytes_read.sum,gpu__time_duration.sum,sm__cycles_elapsed.avg.per_second,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,lts__t_sector_hit_rate.pct
for cfg in "$@"; do
  echo "=== $cfg" | tee -a gpurun_out/exp_k1.log
  env $cfg timeout 300 python tools/perf_topk.py --batches ${BATCHES:-1024} --iters 10 >> gpurun_out/exp_k1.log 2>&1
  echo "rc=$?" >> gpurun_out/exp_k1.log
done
if [ "${NCU:-1}" = "1" ]; then
for cfg in "$@"; do
  tag=$(echo "$cfg" | tr -c 'A-Za-z0-9=\n' '_')
  env $cfg python tools/perf_topk.py --batches ${BATCHES:-1024} --iters 1 > gpurun_out/plain_$tag.log 2>&1 &&
  env $cfg ncu --metrics $M --clock-control none -k regex:topk_gemm_kernel -s 2 -c 1 --csv \
      --log-file gpurun_out/ncu_$tag.csv python tools/perf_topk.py --batches ${BATCHES:-1024} --iters 1 > /dev/null 2>&1
  echo "=== ncu $cfg rc=$?" >> gpurun_out/exp_k1.log
  grep -E "dram__bytes_read|gpu__time_duration|per_second|tensor|hit_rate" gpurun_out/ncu_$tag.csv | awk -F'","' '{print $(NF-2), $(NF-1), $NF}' >> gpurun_out/exp_k1.log
done
fi
cat gpurun_out/exp_k1.log
