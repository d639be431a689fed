#!/bin/bash
# ncu --set full of the shipped GEMV (quantizer fused for k <= 8192) on two 70B-token shapes, each after a clean run
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
for spec in "q4_K 57344 8192" "q4_K 8192 28672"; do
  set -- $spec
  tag=ncu_gemv_final_$1_$2x$3
  timeout 300 python tools/prof_gemv.py $1 $2 $3 1 6 > gpurun_out/${tag}_plain.log 2>&1 || { echo "plain run failed"; continue; }
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_gemv -s 8 -c 1 -f -o gpurun_out/$tag python tools/prof_gemv.py $1 $2 $3 1 6 > gpurun_out/${tag}_ncu.log 2>&1
  ncu -i gpurun_out/$tag.ncu-rep --page details > gpurun_out/${tag}_details.txt 2>/dev/null
  ncu -i gpurun_out/$tag.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv, sys
rows = list(csv.reader(sys.stdin))
h = rows[0]; v = rows[-1]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'dram__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread']
for w in want:
    if w in h: print(w, v[h.index(w)])
" > gpurun_out/${tag}_raw_summary.txt
  ncu -i gpurun_out/$tag.ncu-rep --page source --csv --launch-skip 0 --launch-count 1 > /tmp/src.csv 2>/dev/null
  python tools/ncu_src_summary.py /tmp/src.csv 25 > gpurun_out/${tag}_source_summary.txt 2>&1
  rm -f gpurun_out/$tag.ncu-rep
  tail -1 gpurun_out/${tag}_plain.log; cat gpurun_out/${tag}_raw_summary.txt
  grep -E "Duration|DRAM Throughput|Issue Slots Busy|Executed Ipc Active|Warp Cycles Per Issued" gpurun_out/${tag}_details.txt | head -6
  sed -n 2,12p gpurun_out/${tag}_source_summary.txt
done
