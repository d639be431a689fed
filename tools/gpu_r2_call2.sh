#!/bin/bash
# round 2, call 2: residual-plane MMQ at KC=256 -- GPU tests, bench line, ncu of the SHIPPED GEMV and of k_mmq_p
# (text summaries are produced on the box; the .ncu-rep files are too big to travel back)
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/c2_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/c2_pytest.log
timeout 900 python bench.py --steps 64 --warmup 4 > gpurun_out/c2_bench.json 2> gpurun_out/c2_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/c2_bench.err
python - <<'PY'
import json
d=json.load(open("gpurun_out/c2_bench.json"))
print("tg", d["value"], d["roofline"]["frac"], "e2e", d["e2e"]["value"], "cpu", d["cpu_baseline"]["value"])
p=d["pp512"]; print("pp512", p["value"], p["ms_per_step"], p["roofline"]["achieved"], p["roofline"]["peak"], p["roofline"]["frac"])
PY
summ() {  # $1 = report stem
  ncu -i gpurun_out/$1.ncu-rep --page details > gpurun_out/$1_details.txt 2>/dev/null
  ncu -i gpurun_out/$1.ncu-rep --page source --csv --launch-skip 0 --launch-count 1 > /tmp/$1_src.csv 2>/dev/null
  python tools/ncu_src_summary.py /tmp/$1_src.csv 40 > gpurun_out/$1_source_summary.txt 2>&1
  ncu -i gpurun_out/$1.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin)); H=rows[0]
keep=['Kernel Name','gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','launch__registers_per_thread','launch__grid_size','launch__block_size','sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active','smsp__inst_executed.sum','sm__inst_executed_pipe_lsu.sum','smsp__issue_active.avg.pct_of_peak_sustained_active','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed','sm__warps_active.avg.pct_of_peak_sustained_active']
ix=[H.index(k) for k in keep if k in H]
for r in rows[:1]+rows[2:]: print(' | '.join(r[i] for i in ix))
" > gpurun_out/$1_raw.txt
  rm -f gpurun_out/$1.ncu-rep
}
for spec in "q4_K 14336 4096" "q4_K 4096 14336" "q6_K 4096 14336" "q4_K 4096 4096"; do
  set -- $spec
  tag=c2_gemv_$1_$2x$3
  timeout 300 python tools/prof_gemv.py $1 $2 $3 1 8 > gpurun_out/${tag}_plain.log 2>&1 && \
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_gemv -s 8 -c 1 -f -o gpurun_out/$tag python tools/prof_gemv.py $1 $2 $3 1 8 > gpurun_out/${tag}_ncu.log 2>&1
  summ $tag; head -3 gpurun_out/${tag}_raw.txt | cut -c1-400
done
tag=c2_mmq_p_q4_K_28672x4096x512
timeout 300 python tools/prof_mmq.py q4_K 28672 4096 512 2 > gpurun_out/${tag}_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_mmq_p -s 2 -c 1 -f -o gpurun_out/$tag python tools/prof_mmq.py q4_K 28672 4096 512 2 > gpurun_out/${tag}_ncu.log 2>&1
summ $tag; tail -1 gpurun_out/${tag}_plain.log; head -3 gpurun_out/${tag}_raw.txt | cut -c1-400
du -sh gpurun_out
