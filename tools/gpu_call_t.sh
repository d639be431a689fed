#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
{
echo "=== fused default"; python tools/gemv_trace.py q4_K 4096 4096 5
python tools/gemv_trace.py q4_K 1024 4096 5
python tools/gemv_trace.py q4_K 4096 14336 5
echo "=== separate quantizer"; B200_FUSED_QUANT_MAX_K=0 python tools/gemv_trace.py q4_K 4096 4096 5
B200_FUSED_QUANT_MAX_K=0 python tools/gemv_trace.py q4_K 4096 14336 5
} > gpurun_out/trace.log 2>&1
cat gpurun_out/trace.log
timeout 600 python tools/gemv_sweep.py q4_K > gpurun_out/sweep_q4k.log 2>&1; grep "n=1 " gpurun_out/sweep_q4k.log
timeout 600 python bench.py --no-cpu --no-pp > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python -c "
import json;d=json.load(open('gpurun_out/bench.json'));print('tok/s',d['value'],'ms',d['ms_per_step'],'frac',d['roofline']['frac'],'e2e',d['e2e']['value'],'launches',d['roofline']['launches_per_step'])"
