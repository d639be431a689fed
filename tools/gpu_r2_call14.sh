#!/bin/bash
# round 2, call 14: MUL_MAT_ID tensor-core parity after the pitch fix; GEMV ring geometry on the 70B row lengths
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "mul_mat_id" > gpurun_out/c14_pytest_mmid.log 2>&1; echo "mmid tests rc=$?"; tail -4 gpurun_out/c14_pytest_mmid.log
timeout 900 python tools/gemv_sweep.py --70b q4_K q6_K > gpurun_out/c14_gemv_sweep_70b.log 2>&1; echo "sweep rc=$?"; cat gpurun_out/c14_gemv_sweep_70b.log
