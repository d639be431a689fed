#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/c18_seq_knobs.log; : > $L
for pub in 0 1 3; do for la in 0 1 4 8; do
  echo "== 8B publish=$pub lookahead=$la" >> $L
  B200_SEQ_PUBLISH=$pub B200_SEQ_LOOKAHEAD=$la timeout 300 python tools/seq_bench.py llama3-8b-q4_k_m 30 2>&1 | grep "graph replay\|bit-identical after" >> $L
done; done
for pub in 0 1 3; do for la in 0 4; do
  echo "== 70B publish=$pub lookahead=$la" >> $L
  B200_SEQ_PUBLISH=$pub B200_SEQ_LOOKAHEAD=$la timeout 300 python tools/seq_bench.py llama3-70b-q4_k_m 20 2>&1 | grep "graph replay\|bit-identical after" >> $L
done; done
cat $L
B200_SEQ_PUBLISH=1 timeout 600 python tools/seq_trace.py 40 8 llama3-8b-q4_k_m > gpurun_out/c18_seq_trace_8b_pub1.log 2>&1; tail -20 gpurun_out/c18_seq_trace_8b_pub1.log
B200_SEQ_PUBLISH=3 B200_SEQ_LOOKAHEAD=4 timeout 600 python tools/seq_trace.py 40 8 llama3-8b-q4_k_m > gpurun_out/c18_seq_trace_8b_pub3_la4.log 2>&1; tail -20 gpurun_out/c18_seq_trace_8b_pub3_la4.log
