#!/bin/bash
# round 2, call 7: e2e_ggml (async calls), llama-bench -v with the scheduler's assignment dump, the new bench.py line
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
export GGML_BACKEND_PATH=$PWD/llama.cpp.dsp_b200/lib/libggml-b200.so
timeout 600 oracle/_ref/ggml-graph-bench --backend B200_0 --steps 50 --warmup 5 2>/dev/null | tee gpurun_out/c7_e2e_ggml_b200_async.json
timeout 600 oracle/_ref/ggml-graph-bench --backend B200_0 --steps 50 --warmup 5 --sync 2>/dev/null | tee gpurun_out/c7_e2e_ggml_b200_sync.json
timeout 600 python tools/make_gguf.py /tmp/llama3-8b-q4_k_m.gguf > gpurun_out/c7_make_gguf.log 2>&1; tail -1 gpurun_out/c7_make_gguf.log
GGML_SCHED_DEBUG=2 timeout 600 oracle/_ref/llama-bench -v -m /tmp/llama3-8b-q4_k_m.gguf -ngl 99 -nkvo 1 -p 0 -n 1 -r 1 > /tmp/sched_full.log 2>&1; echo "sched debug rc=$?"
grep -E "offloaded|offloading|assigned to device|buffer size|model buffer" /tmp/sched_full.log | sort | uniq -c | sort -rn | head -20 > gpurun_out/c7_llama_load_summary.txt; cat gpurun_out/c7_llama_load_summary.txt
python - <<'PY'
import re, collections
lines = open("/tmp/sched_full.log", errors="replace").read().splitlines()
# the LAST scheduled graph dump (the generated token): from the last "## SPLIT #0" on
idx = [i for i, ln in enumerate(lines) if "## SPLIT #0" in ln]
tail = lines[idx[-1]:] if idx else lines
c, splits = collections.Counter(), 0
for ln in tail:
    if "## SPLIT" in ln:
        splits += 1
    m = re.search(r"node #\s*\d+ \(\s*(\w+)\):\s+(\S+) \(.*?\) \[\s*(\w+)\s*\]", ln)
    if m:
        c[(m.group(1), m.group(3))] += 1
out = ["last scheduled graph: %d splits" % splits] + ["  %-14s %-8s %d" % (k[0], k[1], v) for k, v in sorted(c.items())]
print("\n".join(out))
open("gpurun_out/c7_sched_debug_summary.txt", "w").write("\n".join(out) + "\n")
open("gpurun_out/c7_sched_debug_head.txt", "w").write("\n".join(tail[:260]) + "\n")
PY
timeout 900 oracle/_ref/llama-bench -m /tmp/llama3-8b-q4_k_m.gguf -ngl 99 -nkvo 1 -p 512 -n 32 -r 2 > gpurun_out/c7_llama_bench_b200.log 2>&1; echo "llama-bench rc=$?"; tail -6 gpurun_out/c7_llama_bench_b200.log
timeout 1500 python bench.py --steps 20 --warmup 3 > gpurun_out/c7_bench.json 2> gpurun_out/c7_bench.err; echo "bench rc=$?"; tail -5 gpurun_out/c7_bench.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/c7_bench.json"))
print("headline:", d["metric"], d["value"], "frac", d["roofline"]["frac"], "e2e", d["e2e"] and d["e2e"]["value"], "e2e_ggml", d["e2e_ggml"], "\nseq", d["sequence"], "\ncpu", d["cpu_baseline"])
s = d.get("llama3_8b_q4_k_m") or {}
print("8B sub:", s.get("value"), (s.get("roofline") or {}).get("frac"), "seq", (s.get("sequence") or {}).get("value"), "e2e", (s.get("e2e") or {}).get("value"), "e2e_ggml", (s.get("e2e_ggml") or {}).get("value"), "pp512", (s.get("pp512") or {}).get("value"), s.get("error"))
PY
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 | cut -c1-400
