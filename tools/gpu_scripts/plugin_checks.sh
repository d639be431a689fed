#!/bin/bash
# the ggml plug-in on one GPU: the unmodified reference harnesses + the token through the ggml API (e2e_ggml), 8B and 70B
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
timeout 1200 python -m pytest tests/test_ggml_plugin.py -x -q -m gpu > gpurun_out/p_pytest_plugin.log 2>&1; echo "plugin tests rc=$?"; tail -3 gpurun_out/p_pytest_plugin.log
PLUG=$PWD/llama.cpp.dsp_b200/lib/libggml-b200.so
for cfg in "--layers 32 --embd 4096 --ff 14336 --kv 1024 --vocab 128256" "--layers 80 --embd 8192 --ff 28672 --kv 1024 --vocab 128256"; do
  for mode in "" "--sync"; do
    GGML_BACKEND_PATH=$PLUG timeout 600 oracle/_ref/ggml-graph-bench --backend B200_0 --steps 20 --warmup 3 $cfg $mode 2>/dev/null | tail -1
  done
  B200_HOST_ZEROCOPY=0 GGML_BACKEND_PATH=$PLUG timeout 600 oracle/_ref/ggml-graph-bench --backend B200_0 --steps 20 --warmup 3 $cfg 2>/dev/null | tail -1
done | tee gpurun_out/p_e2e_ggml.log | cut -c1-330
