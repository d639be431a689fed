#!/bin/bash
# DRAM traffic and duration of every kernel of ONE decode step (eager launches), for roofline.traffic and the launch list
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu --no-pp --no-graph > gpurun_out/plain_traffic.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
   -k regex:"k_gemv|k_quantize" -s 1644 -c 274 --csv --log-file gpurun_out/traffic.csv \
   python bench.py --steps 1 --warmup 3 --no-cpu --no-pp --no-graph > gpurun_out/ncu_traffic.log 2>&1
tail -c 600 gpurun_out/plain_traffic.log; tail -2 gpurun_out/ncu_traffic.log; wc -l gpurun_out/traffic.csv
