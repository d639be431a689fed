#!/bin/bash
# round 2, call 4: the sequence kernel -- parity tests, then timing against the grouped calls (slot-size sweep)
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "sequence" > gpurun_out/c4_pytest_seq.log 2>&1; echo "seq tests rc=$?"; tail -25 gpurun_out/c4_pytest_seq.log
for sb in 9600 16384 24576; do
  echo "== B200_SEQ_SLOT_BYTES=$sb"
  B200_SEQ_SLOT_BYTES=$sb timeout 600 python tools/seq_bench.py > gpurun_out/c4_seq_bench_$sb.log 2>&1; echo "seq bench rc=$?"; tail -7 gpurun_out/c4_seq_bench_$sb.log
done
