#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/seq_trace.py 40 8 llama3-70b-q4_k_m > gpurun_out/c17_seq_trace_70b.log 2>&1; echo "rc=$?"; cat gpurun_out/c17_seq_trace_70b.log | tail -22
timeout 600 python tools/seq_trace.py 40 8 llama3-8b-q4_k_m > gpurun_out/c17_seq_trace_8b.log 2>&1; echo "rc=$?"; cat gpurun_out/c17_seq_trace_8b.log | tail -22
