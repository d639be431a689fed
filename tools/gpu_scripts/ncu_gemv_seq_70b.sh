#!/bin/bash
# ncu of the sequence kernel on the 70B token (one launch), after the same command ran clean
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
timeout 600 python tools/seq_bench.py llama3-70b-q4_k_m 3 > gpurun_out/c21_plain.log 2>&1; echo "plain rc=$?"; grep "graph replay" gpurun_out/c21_plain.log
tag=c21_ncu_gemv_seq_70b
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_gemv_seq -s 2 -c 1 -f -o gpurun_out/$tag python tools/seq_bench.py llama3-70b-q4_k_m 3 > gpurun_out/${tag}_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/$tag.ncu-rep --page details > gpurun_out/${tag}_details.txt 2>/dev/null
ncu -i gpurun_out/$tag.ncu-rep --page source --csv --launch-skip 0 --launch-count 1 > /tmp/src.csv 2>/dev/null
python tools/ncu_src_summary.py /tmp/src.csv 60 gpurun_out/${tag}_lines.txt > gpurun_out/${tag}_source_summary.txt 2>&1
rm -f gpurun_out/$tag.ncu-rep
grep -E "Duration|DRAM Throughput|Issue Slots Busy|Executed Ipc|No Eligible|Registers Per|Local|L2 Cache Throughput|Warp Cycles Per Issued|Stall" gpurun_out/${tag}_details.txt | head -30
head -70 gpurun_out/${tag}_source_summary.txt
