#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
B200_BENCH_SEQ_TRACE=1 timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 3 --no-pp > gpurun_out/c23_bench_n$N.json 2> gpurun_out/c23_bench_n$N.err
echo "bench n$N rc=$?"; python - $N <<'PY'
import json, sys
try:
    r = json.loads([l for l in open("gpurun_out/c23_bench_n%s.json" % sys.argv[1]) if l.startswith("{")][-1])
    print("value", r["value"], r["config"]["workload"])
    print("seq_split", r.get("sequence_row_split"))
except Exception as e:
    print("no json", e)
PY
tail -5 gpurun_out/c23_bench_n$N.err
