#!/bin/bash
# Runs the UNMODIFIED reference test-backend-ops (built by oracle/Makefile.ref) against the B200 plug-in.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export GGML_BACKEND_PATH=$PWD/llama.cpp.dsp_b200/lib/libggml-b200.so
T=oracle/_ref/test-backend-ops
for op in MUL_MAT MUL_MAT_ID; do
  timeout 600 $T test -o $op -b B200_0 > gpurun_out/tbo_test_$op.log 2>&1
  echo "test $op rc=$?"; grep -c "OK" gpurun_out/tbo_test_$op.log; grep -v "OK\|not supported" gpurun_out/tbo_test_$op.log | tail -15
done
timeout 600 $T perf -o MUL_MAT -b B200_0 > gpurun_out/tbo_perf_MUL_MAT.log 2>&1
echo "perf rc=$?"; grep -i "q4_K\|q4_0\|q8_0\|q6_K\|iq4_xs" gpurun_out/tbo_perf_MUL_MAT.log | tail -60
