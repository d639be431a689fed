#!/bin/bash
# A/B of the persistent GEMM's epilogue: conversion bias kept in TMEM (default) against the epilogue that adds it itself
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "mmq or realistic or mul_mat_id or exact_dequant" > gpurun_out/t_pytest_mmq.log 2>&1; echo "mmq tests rc=$?"; tail -3 gpurun_out/t_pytest_mmq.log
for b in 1 0; do
  B200_MMQ_TMEM_BIAS=$b timeout 900 python bench.py --workload llama3-8b-q4_k_m --no-cpu --steps 10 --warmup 3 > gpurun_out/t_bench_8b_bias$b.json 2> gpurun_out/t_bench_8b_bias$b.err; echo "bias=$b rc=$?"
  python - $b <<'PY'
import json, sys
r = json.loads([l for l in open("gpurun_out/t_bench_8b_bias%s.json" % sys.argv[1]) if l.startswith("{")][-1])
pp = r["pp512"]
print("pp512", pp["value"], pp["ms_per_step"], pp["launch"], pp["roofline"]["frac_of_2x_bf16_burst"], pp["roofline"]["frac"])
for c in r["mul_mat_id"]["cases"] if isinstance(r["mul_mat_id"], dict) else r["mul_mat_id"]:
    if c["n_tokens"] >= 32: print(c["tensor"], c["n_tokens"], round(c["us_per_call"], 1))
PY
done
for b in 1 0; do for t in q4_K q8_0; do B200_MMQ_TMEM_BIAS=$b timeout 300 python tools/prof_mmq.py $t 28672 4096 512 5 2>&1 | tail -1; done; done
