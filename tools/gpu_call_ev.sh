#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/decode_sweep.py HINTS=0,B200_GEMV_EVICT_FIRST=0 HINTS=0 HINTS=0,B200_GEMV_EVICT_FIRST=0 HINTS=0 2>&1 | tee gpurun_out/decode_evict.log
timeout 600 python -m pytest tests -m gpu -x -q -k "mul_mat and not mmq" 2>&1 | tail -3
