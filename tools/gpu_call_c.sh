#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
for res in 3 2; do for v in 0 2 4 8; do
  echo "== RESIDENT=$res V=$v"
  B200_GEMV_RESIDENT=$res B200_GEMV_V=$v timeout 300 python tools/gemv_sweep.py --quick q4_K 2>&1 | grep "n=1 "
done; done > gpurun_out/sweep_res.log 2>&1
cat gpurun_out/sweep_res.log
timeout 600 python bench.py --no-cpu --no-pp > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json
