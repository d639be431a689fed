#!/bin/bash
# one ncu --set full capture of the tensor-core GEMM (BN=128 path, the one the pp512 step mostly runs), after a clean run
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 120 python tools/prof_mmq.py q4_K 4096 14336 512 2 > gpurun_out/plain_mmq.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_mmq -s 2 -c 1 -f -o gpurun_out/prof_mmq_r01 python tools/prof_mmq.py q4_K 4096 14336 512 2 > gpurun_out/ncu_mmq.log 2>&1
tail -2 gpurun_out/plain_mmq.log; tail -1 gpurun_out/ncu_mmq.log
