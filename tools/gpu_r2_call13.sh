#!/bin/bash
# round 2, call 13: MUL_MAT_ID on the tensor cores -- parity, TBO through the plug-in, bench mul_mat_id cases
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "mul_mat_id" > gpurun_out/c13_pytest_mmid.log 2>&1; echo "mmid tests rc=$?"; tail -15 gpurun_out/c13_pytest_mmid.log
timeout 600 python -m pytest tests/test_ggml_plugin.py -m gpu -x -q > gpurun_out/c13_pytest_plugin.log 2>&1; echo "plugin rc=$?"; tail -2 gpurun_out/c13_pytest_plugin.log
timeout 900 python bench.py --workload llama3-8b-q4_k_m --sub --no-cpu --steps 20 --warmup 3 > gpurun_out/c13_bench_8b.json 2> gpurun_out/c13_bench_8b.err; echo "bench rc=$?"; tail -3 gpurun_out/c13_bench_8b.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/c13_bench_8b.json"))
for c in d["mul_mat_id"]["cases"]:
    print("%-18s n_tokens %4d  %-38s %8.1f us  %7.1f GB/s algorithmic  %.2f of HBM peak" % (c["tensor"], c["n_tokens"], c["path"], c["us_per_call"], c["algorithmic_gbs"], c["frac_of_hbm_peak"]))
PY
