#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 180 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "sequence" > gpurun_out/c5_pytest_seq.log 2>&1; echo "seq tests rc=$?"; tail -3 gpurun_out/c5_pytest_seq.log
for sb in 24576 16384; do
  echo "== B200_SEQ_SLOT_BYTES=$sb"
  B200_SEQ_SLOT_BYTES=$sb timeout 300 python tools/seq_bench.py > gpurun_out/c5_seq_bench_$sb.log 2>&1; echo "rc=$?"; grep -E "sequence, |grouped|identical after" gpurun_out/c5_seq_bench_$sb.log
  B200_SEQ_SLOT_BYTES=$sb timeout 300 python tools/seq_trace.py 40 5 > gpurun_out/c5_seq_trace_$sb.log 2>&1; echo "rc=$?"; cat gpurun_out/c5_seq_trace_$sb.log
done
