#!/bin/bash
# round 2, call 10 (2 GPUs): split buffer type parity + llama-bench --split-mode row after the event fix
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
export GGML_BACKEND_PATH=$PWD/llama.cpp.dsp_b200/lib/libggml-b200.so
timeout 600 oracle/_ref/ggml-split-test > gpurun_out/c10_split_test.log 2>&1; echo "split test rc=$?"; tail -6 gpurun_out/c10_split_test.log
timeout 600 python -m pytest tests/test_ggml_plugin.py tests/test_rowsplit.py -m gpu -x -q > gpurun_out/c10_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/c10_pytest.log
timeout 600 python tools/make_gguf.py /tmp/llama3-8b-q4_k_m.gguf > /dev/null 2>&1
GGML_SCHED_DEBUG=2 timeout 600 oracle/_ref/llama-bench -v -m /tmp/llama3-8b-q4_k_m.gguf -ngl 99 -nkvo 1 -sm row -p 0 -n 1 -r 1 > /tmp/sched_row.log 2>&1; echo "sm row debug rc=$?"
grep -E "offloaded|buffer size" /tmp/sched_row.log | sort | uniq -c | head -12 > gpurun_out/c10_llama_row_load_summary.txt
python - <<'PY'
import re, collections
lines = open("/tmp/sched_row.log", errors="replace").read().splitlines()
idx = [i for i, ln in enumerate(lines) if "## SPLIT #0" in ln]
tail = lines[idx[-1]:] if idx else lines
c = collections.Counter()
for ln in tail:
    m = re.search(r"node #\s*\d+ \(\s*(\w+)\):\s+(\S+) \(.*?\) \[\s*(\w+)\s*\]", ln)
    if m:
        c[(m.group(1), m.group(3))] += 1
out = ["-sm row, last scheduled graph:"] + ["  %-14s %-8s %d" % (k[0], k[1], v) for k, v in sorted(c.items())]
print("\n".join(out))
open("gpurun_out/c10_sched_row_summary.txt", "w").write("\n".join(out) + "\n")
PY
timeout 900 oracle/_ref/llama-bench -m /tmp/llama3-8b-q4_k_m.gguf -ngl 99 -nkvo 1 -sm row -p 512 -n 64 -r 2 > gpurun_out/c10_llama_bench_row.log 2>&1; echo "llama-bench -sm row rc=$?"; tail -5 gpurun_out/c10_llama_bench_row.log
