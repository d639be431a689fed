#!/bin/bash
# full bench line + parity + ncu launch list of our kernels (shares, not absolutes)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 900 python bench.py ${BENCH_ARGS} > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json
timeout 300 python tools/gemv_sweep.py --quick q4_K q6_K q4_0 q8_0 iq4_xs > gpurun_out/sweep_quick.log 2>&1; grep "n=1 " gpurun_out/sweep_quick.log
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu --no-pp --no-graph > gpurun_out/plain2.log 2>&1 && \
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"k_gemv|k_quantize|k_mmq|k_plane" -s 1200 -c 500 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-pp --no-graph > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/ncu2.log
