#!/bin/bash
# round 2, call 6: the ggml boundary -- test-backend-ops through the plug-in, e2e_ggml driver, llama-bench on a synthetic GGUF
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
export GGML_BACKEND_PATH=$PWD/llama.cpp.dsp_b200/lib/libggml-b200.so
timeout 900 python -m pytest tests/test_ggml_plugin.py -m gpu -x -q > gpurun_out/c6_pytest_plugin.log 2>&1; echo "plugin tests rc=$?"; tail -3 gpurun_out/c6_pytest_plugin.log
timeout 600 oracle/_ref/ggml-graph-bench --backend B200_0 --steps 50 --warmup 5 > gpurun_out/c6_e2e_ggml_b200.json 2> gpurun_out/c6_e2e_ggml_b200.err; echo "e2e_ggml rc=$?"; cat gpurun_out/c6_e2e_ggml_b200.json; tail -2 gpurun_out/c6_e2e_ggml_b200.err
B200_NO_GROUPING=1 timeout 600 oracle/_ref/ggml-graph-bench --backend B200_0 --steps 50 --warmup 5 2>/dev/null | tee gpurun_out/c6_e2e_ggml_b200_nogroup.json
timeout 600 oracle/_ref/ggml-graph-bench --backend CPU --steps 3 --warmup 1 2>/dev/null | tee gpurun_out/c6_e2e_ggml_cpu.json
timeout 600 python tools/make_gguf.py /tmp/llama3-8b-q4_k_m.gguf > gpurun_out/c6_make_gguf.log 2>&1; tail -1 gpurun_out/c6_make_gguf.log
timeout 900 oracle/_ref/llama-bench -m /tmp/llama3-8b-q4_k_m.gguf -ngl 99 -nkvo 1 -p 512 -n 32 -r 2 > gpurun_out/c6_llama_bench_b200.log 2>&1; echo "llama-bench rc=$?"; tail -12 gpurun_out/c6_llama_bench_b200.log
GGML_SCHED_DEBUG=2 timeout 600 oracle/_ref/llama-bench -m /tmp/llama3-8b-q4_k_m.gguf -ngl 99 -nkvo 1 -p 0 -n 1 -r 1 > gpurun_out/c6_sched_debug_full.log 2>&1; echo "sched debug rc=$?"
grep -c "MUL_MAT" gpurun_out/c6_sched_debug_full.log
python - <<'PY'
import re, collections
c = collections.Counter()
lines = open("gpurun_out/c6_sched_debug_full.log", errors="replace").read().splitlines()
for ln in lines:
    m = re.search(r"node #\s*\d+ \(\s*(\w+)\):\s+\S+ \(.*?\) \[\s*(\w+)\s*\]", ln)
    if m:
        c[(m.group(1), m.group(2))] += 1
print("node (op, backend) counts in the LAST scheduled graphs:")
for k, v in sorted(c.items()):
    print("  %-14s %-8s %d" % (k[0], k[1], v))
open("gpurun_out/c6_sched_debug_summary.txt", "w").write("\n".join("%s %s %d" % (k[0], k[1], v) for k, v in sorted(c.items())) + "\n")
# keep the dump small: the first split listing only
keep = [ln for ln in lines if "SPLIT" in ln or "node #" in ln][:400]
open("gpurun_out/c6_sched_debug_head.txt", "w").write("\n".join(keep) + "\n")
PY
rm -f gpurun_out/c6_sched_debug_full.log
timeout 900 oracle/_ref/llama-bench -m /tmp/llama3-8b-q4_k_m.gguf -ngl 0 -p 64 -n 8 -r 1 > gpurun_out/c6_llama_bench_cpu.log 2>&1; tail -6 gpurun_out/c6_llama_bench_cpu.log
