#!/bin/bash
# row-split regression on 2 GPUs: tests + a short bench (fused peer-store GEMV)
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "rowsplit or row_split" 2>&1 | tail -3
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 16 --warmup 3 --no-cpu > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/bench_n2.json') if l.startswith('{')][-1])
print('N=2', d['value'], 'tok/s', d['ms_per_step'], 'ms', d['config']['workload']); print('single', d.get('single_gpu_same_workload')); print('nccl', d.get('nccl_all_gather_baseline'))
PY
