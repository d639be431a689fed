#!/bin/bash
# round 2, call 8: llama-bench through the plug-in after the weight-placement fix (+ scheduler assignment dump)
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
export GGML_BACKEND_PATH=$PWD/llama.cpp.dsp_b200/lib/libggml-b200.so
timeout 900 python -m pytest tests/test_ggml_plugin.py -m gpu -x -q > gpurun_out/c8_pytest_plugin.log 2>&1; echo "plugin tests rc=$?"; tail -2 gpurun_out/c8_pytest_plugin.log
timeout 600 python tools/make_gguf.py /tmp/llama3-8b-q4_k_m.gguf > gpurun_out/c8_make_gguf.log 2>&1; tail -1 gpurun_out/c8_make_gguf.log
GGML_SCHED_DEBUG=2 timeout 600 oracle/_ref/llama-bench -v -m /tmp/llama3-8b-q4_k_m.gguf -ngl 99 -nkvo 1 -p 0 -n 1 -r 1 > /tmp/sched_full.log 2>&1; echo "sched debug rc=$?"
grep -E "offloaded|buffer size" /tmp/sched_full.log | sort | uniq -c | head -12 | tee gpurun_out/c8_llama_load_summary.txt
python - <<'PY'
import re, collections
lines = open("/tmp/sched_full.log", errors="replace").read().splitlines()
idx = [i for i, ln in enumerate(lines) if "## SPLIT #0" in ln]
tail = lines[idx[-1]:] if idx else lines
c, splits = collections.Counter(), 0
for ln in tail:
    if "## SPLIT" in ln:
        splits += 1
    m = re.search(r"node #\s*\d+ \(\s*(\w+)\):\s+(\S+) \(.*?\) \[\s*(\w+)\s*\]", ln)
    if m:
        c[(m.group(1), m.group(3))] += 1
out = ["last scheduled graph (one generated token): %d splits" % splits] + ["  %-14s %-8s %d" % (k[0], k[1], v) for k, v in sorted(c.items())]
print("\n".join(out))
open("gpurun_out/c8_sched_debug_summary.txt", "w").write("\n".join(out) + "\n")
open("gpurun_out/c8_sched_debug_head.txt", "w").write("\n".join(tail[:200]) + "\n")
PY
timeout 900 oracle/_ref/llama-bench -m /tmp/llama3-8b-q4_k_m.gguf -ngl 99 -nkvo 1 -p 512 -n 128 -r 2 > gpurun_out/c8_llama_bench_b200.log 2>&1; echo "llama-bench rc=$?"; tail -6 gpurun_out/c8_llama_bench_b200.log
timeout 900 oracle/_ref/llama-bench -m /tmp/llama3-8b-q4_k_m.gguf -ngl 0 -p 512 -n 32 -r 1 > gpurun_out/c8_llama_bench_cpu.log 2>&1; tail -5 gpurun_out/c8_llama_bench_cpu.log
