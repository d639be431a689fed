#!/bin/bash
# ncu --set full of the persistent tensor-core GEMM (k_mmq_p, 28672 x 4096 x 512, Q4_K) after a clean run of the same command,
# plus the per-shape times of the 8B prefill matmuls
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
for shape in "28672 4096 512" "4096 14336 512" "6144 4096 512" "4096 4096 512"; do
  set -- $shape
  timeout 300 python tools/prof_mmq.py q4_K $1 $2 $3 4 2>&1 | tail -1
done | tee gpurun_out/mmq_shapes.log
tag=ncu_mmq_p_v3_tmem_bias_q4_K_28672x4096x512
timeout 300 python tools/prof_mmq.py q4_K 28672 4096 512 2 > gpurun_out/${tag}_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_mmq_p -s 2 -c 1 -f -o gpurun_out/$tag python tools/prof_mmq.py q4_K 28672 4096 512 2 > gpurun_out/${tag}_ncu.log 2>&1
ncu -i gpurun_out/$tag.ncu-rep --page details > gpurun_out/${tag}_details.txt 2>/dev/null
ncu -i gpurun_out/$tag.ncu-rep --page source --csv --launch-skip 0 --launch-count 1 > /tmp/src.csv 2>/dev/null
python tools/ncu_src_summary.py /tmp/src.csv 30 > gpurun_out/${tag}_source_summary.txt 2>&1
rm -f gpurun_out/$tag.ncu-rep
grep -E "Tensor|Issue Slots Busy|Local|Registers Per|Duration|Executed Ipc" gpurun_out/${tag}_details.txt | head -12
head -40 gpurun_out/${tag}_source_summary.txt
