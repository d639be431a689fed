#!/bin/bash
# round 2, call 9 (2 GPUs): split buffer type parity, llama-bench --split-mode row, test-backend-ops on device 1, bench N=2
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
export GGML_BACKEND_PATH=$PWD/llama.cpp.dsp_b200/lib/libggml-b200.so
nvidia-smi -L
timeout 600 oracle/_ref/ggml-split-test > gpurun_out/c9_split_test.log 2>&1; echo "split test rc=$?"; tail -8 gpurun_out/c9_split_test.log
timeout 600 oracle/_ref/test-backend-ops test -o MUL_MAT -b B200_1 > gpurun_out/c9_tbo_dev1.log 2>&1; echo "tbo dev1 rc=$?"; tail -3 gpurun_out/c9_tbo_dev1.log
timeout 600 python -m pytest tests/test_ggml_plugin.py tests/test_rowsplit.py -m gpu -x -q > gpurun_out/c9_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/c9_pytest.log
timeout 600 python tools/make_gguf.py /tmp/llama3-8b-q4_k_m.gguf > /dev/null 2>&1
GGML_SCHED_DEBUG=2 timeout 600 oracle/_ref/llama-bench -v -m /tmp/llama3-8b-q4_k_m.gguf -ngl 99 -nkvo 1 -sm row -p 0 -n 1 -r 1 > /tmp/sched_row.log 2>&1; echo "sm row debug rc=$?"
grep -E "offloaded|buffer size" /tmp/sched_row.log | sort | uniq -c | head -12 | tee gpurun_out/c9_llama_row_load_summary.txt
timeout 900 oracle/_ref/llama-bench -m /tmp/llama3-8b-q4_k_m.gguf -ngl 99 -nkvo 1 -sm row -p 512 -n 64 -r 2 > gpurun_out/c9_llama_bench_row.log 2>&1; echo "llama-bench -sm row rc=$?"; tail -5 gpurun_out/c9_llama_bench_row.log
timeout 900 oracle/_ref/llama-bench -m /tmp/llama3-8b-q4_k_m.gguf -ngl 99 -nkvo 1 -sm none -p 512 -n 64 -r 2 > gpurun_out/c9_llama_bench_none.log 2>&1; tail -4 gpurun_out/c9_llama_bench_none.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/c9_bench_n2.json 2> gpurun_out/c9_bench_n2.err; echo "bench n2 rc=$?"; tail -3 gpurun_out/c9_bench_n2.err; cut -c1-500 gpurun_out/c9_bench_n2.json
