#!/bin/bash
# GEMV iteration loop: parity tests, quick sweep, bench (decode only), ncu capture of two GEMV shapes
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 600 python tools/gemv_sweep.py --quick ${SWEEP_TYPES:-q4_K q6_K q4_0 q8_0 iq4_xs} > gpurun_out/sweep_quick.log 2>&1
grep "n=1 " gpurun_out/sweep_quick.log
timeout 600 python bench.py --no-cpu --no-pp > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json
if [ -n "$FULL_SWEEP" ]; then timeout 600 python tools/gemv_sweep.py q4_K > gpurun_out/sweep_q4k.log 2>&1; grep "n=1 " gpurun_out/sweep_q4k.log; fi
timeout 300 python tools/prof_gemv.py q4_K 4096 14336 1 8 > gpurun_out/plain.log 2>&1 && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_gemv -s 8 -c 1 -o gpurun_out/prof_gemv_q4k python tools/prof_gemv.py q4_K 4096 14336 1 8 > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/ncu.log
timeout 300 python tools/prof_gemv.py q4_K 4096 4096 1 8 > gpurun_out/plain_b.log 2>&1 && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_gemv -s 8 -c 1 -o gpurun_out/prof_gemv_q4k_4096 python tools/prof_gemv.py q4_K 4096 4096 1 8 > gpurun_out/ncu_b.log 2>&1
tail -2 gpurun_out/ncu_b.log
