#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "host" 2>&1 | tail -4
for zc in 1 0; do
B200_HOST_ZEROCOPY=$zc timeout 600 python bench.py --no-cpu --no-pp --steps 64 > gpurun_out/bench_zc$zc.json 2> gpurun_out/bench_zc$zc.err; echo "rc=$?"
python -c "
import json;d=json.load(open('gpurun_out/bench_zc$zc.json'));print('zc=$zc tok/s',d['value'],'e2e',d['e2e']['value'],d['e2e']['ms_per_step'],'eager',d['e2e']['ms_per_step_eager'])"
done
