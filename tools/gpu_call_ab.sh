#!/bin/bash
mkdir -p gpurun_out
L=$PWD/llama.cpp.dsp_b200/lib
{
for shp in "28672 4096 512" "18944 4096 256"; do
B200_QMM_LIB=$L/libb200qmm_trace.so timeout 120 python tools/mmq_trace.py q4_K $shp
done
} 2>&1 | tee gpurun_out/mmq_trace_persist.log
