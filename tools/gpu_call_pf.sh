#!/bin/bash
# parity + L2-prefetch sweep (decode) + epilogue sweep (prefill)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 600 python tools/decode_sweep.py HINTS=0 B200_GEMV_PF_MB=4 B200_GEMV_PF_MB=8 B200_GEMV_PF_MB=16 B200_GEMV_PF_MB=32 B200_GEMV_PF_MB=64 HINTS=0 2>&1 | tee gpurun_out/decode_sweep.log
timeout 600 python tools/pp_sweep.py B200_MMQ_EPI=0 B200_MMQ_EPI=1 B200_MMQ_EPI=2 B200_MMQ_EPI=2,B200_MMQ_BN=128 B200_MMQ_EPI=2,B200_MMQ_BN=256 B200_MMQ_EPI=0 2>&1 | tee gpurun_out/pp_sweep.log
for cfg in "B200_MMQ_EPI=0" "B200_MMQ_EPI=2" "B200_MMQ_PAIR=1"; do for shp in "14336 4096" "4096 14336" "4096 4096"; do
  echo -n "[$cfg] "; env $cfg timeout 120 python tools/prof_mmq.py q4_K $shp 512 4 2>&1 | tail -1
done; done | tee gpurun_out/mmq_shapes.log
