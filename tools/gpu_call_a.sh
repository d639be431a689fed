#!/bin/bash
# first GPU call of the round: parity tests, bench line, GEMV sweep, ncu launch list + full capture of the GEMV
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
lscpu | head -25 > gpurun_out/lscpu.txt; nproc >> gpurun_out/lscpu.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json
timeout 300 python bench.py --impl reference --steps 20 --warmup 2 > gpurun_out/bench_ref.json 2>&1
cat gpurun_out/bench_ref.json
timeout 600 python tools/gemv_sweep.py --quick q4_K q6_K q4_0 q8_0 iq4_xs > gpurun_out/sweep_quick.log 2>&1
cat gpurun_out/sweep_quick.log
timeout 600 python tools/gemv_sweep.py q4_K > gpurun_out/sweep_q4k.log 2>&1
timeout 300 python tools/prof_gemv.py q4_K 4096 14336 1 8 > gpurun_out/plain.log 2>&1 && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_gemv -s 8 -c 2 -o gpurun_out/prof_gemv_q4k python tools/prof_gemv.py q4_K 4096 14336 1 8 > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/ncu.log
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu --no-pp --no-graph > gpurun_out/plain2.log 2>&1 && \
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 700 -c 460 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-pp --no-graph > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/ncu2.log
