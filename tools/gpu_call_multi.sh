#!/bin/bash
# N-GPU row-split bench (N = $1, default 2).  Never run under ncu.  $3 = extra env assignments for a second run.
cd "$(dirname "$0")/.."
N=${1:-2}
STEPS=${2:-32}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_rowsplit.py -m gpu -x -q 2>&1 | tail -5
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
    --master-port 29517 bench.py --gpus $N --steps $STEPS --warmup 3 > gpurun_out/bench_n${N}.json 2> gpurun_out/bench_n${N}.err
echo "N=$N rc=$?"; tail -c 3500 gpurun_out/bench_n${N}.json; grep -v "^W1\|^\*\*\*\|OMP_NUM" gpurun_out/bench_n${N}.err | tail -12
if [ -n "$3" ]; then
  env $3 timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
    --master-port 29518 bench.py --gpus $N --steps $STEPS --warmup 3 --no-single > gpurun_out/bench_n${N}_alt.json 2> gpurun_out/bench_n${N}_alt.err
  echo "alt ($3) rc=$?"; python - <<PY
import json
d=json.loads(open("gpurun_out/bench_n${N}_alt.json").read().strip().splitlines()[-1])
print("alt value", d["value"], "ms", d["ms_per_step"], "nccl", d["nccl_all_gather_baseline"]["ms_per_step"])
PY
fi
