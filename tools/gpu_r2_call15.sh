#!/bin/bash
# round 2, call 15: planner rule for long rows -- regression sweep on 8B shapes (defaults only), 70B shapes, GPU suite, bench
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 600 python tools/gemv_sweep.py --quick --big q4_K q6_K 2>&1 | grep "m=" | tee gpurun_out/c15_sweep_8b_defaults.log
timeout 600 python tools/gemv_sweep.py --quick --70b q4_K q6_K 2>&1 | grep "m=" | tee gpurun_out/c15_sweep_70b_defaults.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/c15_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/c15_pytest.log
timeout 1500 python bench.py --steps 32 --warmup 4 > gpurun_out/c15_bench.json 2> gpurun_out/c15_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/c15_bench.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/c15_bench.json"))
print("70B:", d["value"], "frac", d["roofline"]["frac"], "e2e", d["e2e"]["value"], "e2e_ggml", (d["e2e_ggml"] or {}).get("value"), "seq", (d["sequence"] or {}).get("value"), "cpu", (d["cpu_baseline"] or {}).get("value"))
s = d.get("llama3_8b_q4_k_m") or {}
print("8B :", s.get("value"), (s.get("roofline") or {}).get("frac"), "e2e", (s.get("e2e") or {}).get("value"), "e2e_ggml", (s.get("e2e_ggml") or {}).get("value"), "seq", (s.get("sequence") or {}).get("value"), "pp512", (s.get("pp512") or {}).get("value"), (s.get("pp512") or {}).get("roofline", {}).get("frac_of_2x_bf16_burst"), s.get("error"))
PY
