#!/bin/bash
# after clean runs: (1) launch list of one eager decode step + one pp512 step, (2) ncu --set full of the persistent GEMM
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 300 python tools/prof_mmq.py q4_K 28672 4096 512 2 > gpurun_out/plain_mmq.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_mmq_p -s 2 -c 1 -f -o gpurun_out/prof_mmq_p python tools/prof_mmq.py q4_K 28672 4096 512 2 > gpurun_out/ncu_mmq_p.log 2>&1
tail -1 gpurun_out/plain_mmq.log; tail -1 gpurun_out/ncu_mmq_p.log
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu --no-graph > gpurun_out/plain2.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"k_gemv|k_quantize|k_mmq|k_plane|k_mmid" -c 4000 --csv --log-file gpurun_out/launches_r01b.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-graph > gpurun_out/ncu2.log 2>&1
tail -1 gpurun_out/ncu2.log | cut -c1-300; wc -l gpurun_out/launches_r01b.csv
