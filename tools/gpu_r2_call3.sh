#!/bin/bash
# round 2, call 3: the no-alloc entry point -- its test alone in a fresh process, then the whole GPU suite (it runs last)
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/c3_noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -5 gpurun_out/c3_noalloc_alone.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/c3_pytest.log 2>&1; echo "suite rc=$?"; tail -5 gpurun_out/c3_pytest.log
