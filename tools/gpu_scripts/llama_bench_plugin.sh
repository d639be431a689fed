#!/bin/bash
# the UNMODIFIED reference llama-bench on a random-init Llama-3-8B-shaped Q4_K_M GGUF: ggml-cpu alone, then with the plug-in
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
python tools/make_gguf.py /tmp/llama3-8b-q4_k_m.gguf > gpurun_out/lb_make_gguf.log 2>&1; tail -1 gpurun_out/lb_make_gguf.log
export LD_LIBRARY_PATH=$PWD/oracle/_ref:$LD_LIBRARY_PATH
GGML_BACKEND_PATH=$PWD/llama.cpp.dsp_b200/lib/libggml-b200.so timeout 900 oracle/_ref/llama-bench -m /tmp/llama3-8b-q4_k_m.gguf -ngl 99 -nkvo 1 -p 512 -n 128 -r 3 > gpurun_out/lb_b200.log 2>&1; echo "b200 rc=$?"; tail -6 gpurun_out/lb_b200.log
timeout 900 oracle/_ref/llama-bench -m /tmp/llama3-8b-q4_k_m.gguf -ngl 0 -p 512 -n 128 -r 2 > gpurun_out/lb_cpu.log 2>&1; echo "cpu rc=$?"; tail -5 gpurun_out/lb_cpu.log
