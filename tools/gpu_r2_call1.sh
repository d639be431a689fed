#!/bin/bash
# round 2, call 1: baseline on this box -- GPU tests, TMEM read-out micro-benchmark, bench line, ncu of the SHIPPED GEMV
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/c1_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/c1_pytest.log
timeout 120 tools/ubench/bin/tmem_read_bench > gpurun_out/c1_tmem_read.log 2>&1; cat gpurun_out/c1_tmem_read.log
timeout 900 python bench.py --steps 64 --warmup 4 > gpurun_out/c1_bench.json 2> gpurun_out/c1_bench.err; echo "bench rc=$?"; cut -c1-600 gpurun_out/c1_bench.json
for spec in "q4_K 14336 4096" "q4_K 4096 14336" "q6_K 4096 14336" "q4_K 4096 4096"; do
  set -- $spec
  tag=$1_$2x$3
  timeout 300 python tools/prof_gemv.py $1 $2 $3 1 8 > gpurun_out/c1_plain_$tag.log 2>&1 && \
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_gemv -s 8 -c 2 -f -o gpurun_out/c1_gemv_$tag python tools/prof_gemv.py $1 $2 $3 1 8 > gpurun_out/c1_ncu_$tag.log 2>&1
  tail -1 gpurun_out/c1_plain_$tag.log; tail -1 gpurun_out/c1_ncu_$tag.log | cut -c1-200
done
ls -la gpurun_out | head -40
