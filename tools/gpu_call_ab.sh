#!/bin/bash
mkdir -p gpurun_out
L=llama.cpp.dsp_b200/lib
{
for shp in "18944 4096 256" "14336 4096 512" "8192 4096 256"; do
B200_MMQ_PERSIST=0 B200_MMQ_KSPLIT=1 B200_QMM_LIB=$PWD/$L/libb200qmm_trace.so timeout 120 python tools/mmq_trace.py q4_K $shp
done
} 2>&1 | tee gpurun_out/mmq_trace2.log
