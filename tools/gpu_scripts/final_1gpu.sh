#!/bin/bash
# final 1-GPU measurements of the round
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/f1_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/f1_pytest_gpu.log
timeout 900 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/f1_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/f1_smoke.log
# DRAM traffic + launch list of one 70B step (plain run first, then the same command under ncu)
B200_BENCH_NVTX_STEP=1 timeout 600 python bench.py --steps 1 --warmup 3 --no-cpu --no-pp --no-sub > gpurun_out/f1_step_plain.json 2> gpurun_out/f1_step_plain.err; echo "plain step rc=$?"
B200_BENCH_NVTX_STEP=1 timeout 1500 ncu --nvtx --nvtx-include "b200_step/" --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum \
   --clock-control none --csv --log-file gpurun_out/f1_traffic.csv python bench.py --steps 1 --warmup 3 --no-cpu --no-pp --no-sub > gpurun_out/f1_step_ncu.log 2>&1; echo "ncu rc=$?"
python tools/traffic_from_csv.py gpurun_out/f1_traffic.csv gpurun_out/f1_step_plain.json gpurun_out/decode_step_dram_traffic_llama3-70b-q4_k_m.json
mkdir -p profiles/r02; cp gpurun_out/decode_step_dram_traffic_llama3-70b-q4_k_m.json profiles/r02/
gzip -f gpurun_out/f1_traffic.csv
timeout 1500 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/f1_bench_reference.json 2> gpurun_out/f1_bench_reference.err; echo "ref rc=$?"; tail -c 600 gpurun_out/f1_bench_reference.json
timeout 2400 python bench.py > gpurun_out/f1_bench.json 2> gpurun_out/f1_bench.err; echo "bench rc=$?"; tail -c 300 gpurun_out/f1_bench.json; tail -3 gpurun_out/f1_bench.err
timeout 900 python bench.py --workload llama2-7b-q4_0 --no-sub --steps 20 --warmup 3 > gpurun_out/f1_bench_llama2_7b_q4_0.json 2> gpurun_out/f1_bench_llama2.err; echo "llama2 rc=$?"; tail -c 300 gpurun_out/f1_bench_llama2_7b_q4_0.json
timeout 1500 python bench.py --sweep > gpurun_out/f1_config3_sweep.json 2> gpurun_out/f1_sweep.err; echo "sweep rc=$?"; tail -c 400 gpurun_out/f1_config3_sweep.json
