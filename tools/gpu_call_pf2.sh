#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/pp_sweep.py B200_MMQ_PF=0 B200_MMQ_PF=4 B200_MMQ_PF=8 B200_MMQ_PF=16 B200_MMQ_PF=0 B200_MMQ_PF=8,B200_MMQ_RASTER=1 2>&1 | tee gpurun_out/pp_sweep_pf.log
