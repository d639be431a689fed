#!/bin/bash
mkdir -p gpurun_out
L=gpurun_out/c20_seq_experiments.log; : > $L
for m in llama3-70b-q4_k_m llama3-8b-q4_k_m; do
for cfg in "" "SEQ_INDEPENDENT=1" "B200_SEQ_DEBUG=1" "SEQ_INDEPENDENT=1 B200_SEQ_DEBUG=1" "B200_SEQ_PUBLISH=3" "SEQ_INDEPENDENT=1 B200_SEQ_DEBUG=1 B200_SEQ_PUBLISH=3"; do
  echo "== $m $cfg" >> $L
  env $cfg timeout 300 python tools/seq_bench.py $m 20 2>&1 | grep "sequence, graph replay" >> $L
done; done
cat $L
