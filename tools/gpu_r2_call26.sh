#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "sequence" > gpurun_out/c26_pytest_seq.log 2>&1; echo "pytest seq rc=$?"; tail -3 gpurun_out/c26_pytest_seq.log
for m in llama3-8b-q4_k_m llama3-70b-q4_k_m; do
  timeout 600 python tools/seq_bench.py $m 30 > gpurun_out/c26_seq_bench_$m.log 2>&1; echo "seq_bench $m rc=$?"; grep "graph replay\|bit-identical after" gpurun_out/c26_seq_bench_$m.log
done
