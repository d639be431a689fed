#!/bin/bash
# row-split bench on N GPUs of one box (N from $1): fused peer-store GEMV vs NCCL all-gather vs the same token on one GPU
N=${1:-4}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 16 --warmup 3 --no-cpu > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/bench_n$N.json') if l.startswith('{')][-1])
print('N=$N', d['value'], 'tok/s', d['ms_per_step'], 'ms'); print('single', (d.get('single_gpu_same_workload') or {}).get('value')); print('nccl', (d.get('nccl_all_gather_baseline') or {}).get('value'))
PY
