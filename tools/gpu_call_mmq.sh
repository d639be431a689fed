#!/bin/bash
mkdir -p gpurun_out
timeout 120 python tools/mmq_check.py quick 2>&1 | tail -12
for bn in 128 256; do for shape in "4096 4096" "14336 4096" "4096 14336"; do
  B200_MMQ_BN=$bn timeout 120 python tools/prof_mmq.py q4_K $shape 512 4
done; done 2>&1 | tee gpurun_out/mmq_graph.log
BN=${PROF_BN:-128}
B200_MMQ_BN=$BN timeout 120 python tools/prof_mmq.py q4_K 14336 4096 512 2 > gpurun_out/plain_mmq.log 2>&1 && \
B200_MMQ_BN=$BN timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_mmq -s 2 -c 1 -o gpurun_out/prof_mmq python tools/prof_mmq.py q4_K 14336 4096 512 2 > gpurun_out/ncu_mmq.log 2>&1
tail -1 gpurun_out/ncu_mmq.log
