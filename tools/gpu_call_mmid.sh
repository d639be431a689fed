#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "mul_mat_id" 2>&1 | tail -8
timeout 600 python tools/mmid_bench.py 0 1 2>&1 | tee gpurun_out/mmid_bench.log
