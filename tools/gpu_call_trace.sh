#!/bin/bash
mkdir -p gpurun_out
{ echo "=== no hints"; timeout 300 python tools/layer_trace.py 2; echo "=== hints"; HINTS=1 timeout 300 python tools/layer_trace.py 2; } 2>&1 | tee gpurun_out/layer_trace.log
