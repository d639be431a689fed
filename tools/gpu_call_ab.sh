#!/bin/bash
mkdir -p gpurun_out
L=llama.cpp.dsp_b200/lib
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
{
timeout 600 python tools/pp_sweep.py B200_MMQ_KSPLIT=1 B200_MMQ_EPI=2 B200_MMQ_KSPLIT=1,B200_MMQ_BN=128 B200_MMQ_EPI=0
} 2>&1 | tee gpurun_out/pp_ab.log
for cfg in "B200_MMQ_KSPLIT=1" "B200_MMQ_EPI=2"; do for shp in "14336 4096" "4096 14336" "4096 4096" "6144 4096"; do
  echo -n "[$cfg] "; env $cfg timeout 120 python tools/prof_mmq.py q4_K $shp 512 4 2>&1 | tail -1
done; done | tee gpurun_out/mmq_shapes.log
{
echo "=== base (HEAD) lib"; B200_QMM_LIB=$PWD/$L/libb200qmm_base.so timeout 300 python tools/decode_sweep.py HINTS=0
echo "=== new lib"; timeout 600 python tools/decode_sweep.py HINTS=0 HINTS=0,B200_GEMV_DEBUG=32
} 2>&1 | tee gpurun_out/decode_ab.log
