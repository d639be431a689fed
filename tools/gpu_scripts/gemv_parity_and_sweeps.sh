#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "gemv or sequence or mul_mat or noalloc" > gpurun_out/c22_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/c22_pytest.log
timeout 600 python tools/gemv_sweep.py --70b > gpurun_out/c22_gemv_sweep_70b.log 2>&1; echo "sweep rc=$?"; tail -9 gpurun_out/c22_gemv_sweep_70b.log
timeout 600 python tools/gemv_sweep.py > gpurun_out/c22_gemv_sweep_8b.log 2>&1; echo "sweep rc=$?"; tail -9 gpurun_out/c22_gemv_sweep_8b.log
for m in llama3-8b-q4_k_m llama3-70b-q4_k_m; do
  timeout 600 python tools/seq_bench.py $m 30 > gpurun_out/c22_seq_bench_$m.log 2>&1; echo "seq_bench $m rc=$?"; grep "graph replay\|bit-identical after" gpurun_out/c22_seq_bench_$m.log
done
