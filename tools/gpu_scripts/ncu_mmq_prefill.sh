#!/bin/bash
# round 2, call 12: pipelined MMQ epilogue -- parity, per-shape time, pp512 via the 8B sub-run, ncu of k_mmq_p
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "mmq or realistic or gemv or sequence or mul_mat" > gpurun_out/c12_pytest_mmq.log 2>&1; echo "mmq tests rc=$?"; tail -3 gpurun_out/c12_pytest_mmq.log
for shape in "28672 4096 512" "4096 14336 512" "6144 4096 512" "4096 4096 512"; do
  set -- $shape
  timeout 300 python tools/prof_mmq.py q4_K $1 $2 $3 4 2>&1 | tail -1
done | tee gpurun_out/c12_mmq_shapes.log
timeout 900 python bench.py --workload llama3-8b-q4_k_m --sub --no-cpu --steps 20 --warmup 3 > gpurun_out/c12_bench_8b.json 2> gpurun_out/c12_bench_8b.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.load(open("gpurun_out/c12_bench_8b.json"))
p = d["pp512"]; r = p["roofline"]
print("pp512", p["value"], p["ms_per_step"], "TOPS", r["achieved"], "frac(cublaslt)", r["frac"], "frac(2xbf16 burst)", r["frac_of_2x_bf16_burst"], "sustained", r["frac_of_2x_bf16_sustained"])
print("tg", d["value"], d["roofline"]["frac"])
PY
tag=c12_mmq_p_q4_K_28672x4096x512
timeout 300 python tools/prof_mmq.py q4_K 28672 4096 512 2 > gpurun_out/${tag}_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_mmq_p -s 2 -c 1 -f -o gpurun_out/$tag python tools/prof_mmq.py q4_K 28672 4096 512 2 > gpurun_out/${tag}_ncu.log 2>&1
ncu -i gpurun_out/$tag.ncu-rep --page details > gpurun_out/${tag}_details.txt 2>/dev/null
ncu -i gpurun_out/$tag.ncu-rep --page source --csv --launch-skip 0 --launch-count 1 > /tmp/src.csv 2>/dev/null
python tools/ncu_src_summary.py /tmp/src.csv 30 > gpurun_out/${tag}_source_summary.txt 2>&1
rm -f gpurun_out/$tag.ncu-rep
grep -E "Tensor|Issue Slots Busy|Local|Registers Per|Duration" gpurun_out/${tag}_details.txt | head -12
head -16 gpurun_out/${tag}_source_summary.txt
