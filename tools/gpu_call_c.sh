#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 600 python tools/gemv_sweep.py q4_K > gpurun_out/sweep_q4k.log 2>&1; grep "n=1 " gpurun_out/sweep_q4k.log
timeout 600 python bench.py --no-cpu --no-pp > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python -c "
import json;d=json.load(open('gpurun_out/bench.json'));print('tok/s',d['value'],'ms',d['ms_per_step'],'frac',d['roofline']['frac'],'e2e',d['e2e']['value'],'launches',d['roofline']['launches_per_step'])"
for shape in "4096 14336" "4096 4096"; do set -- $shape
timeout 300 python tools/prof_gemv.py q4_K $1 $2 1 8 > gpurun_out/plain_$2.log 2>&1 && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_gemv -s 8 -c 1 -o gpurun_out/prof_gemv_q4k_$2 python tools/prof_gemv.py q4_K $1 $2 1 8 > gpurun_out/ncu_$2.log 2>&1
tail -1 gpurun_out/ncu_$2.log; done
