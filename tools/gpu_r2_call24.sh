#!/bin/bash
N=${1:-4}
mkdir -p gpurun_out
for la in 1 4 16; do
B200_SEQ_LOOKAHEAD=$la B200_BENCH_SEQ_TRACE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 3 --no-pp --no-single > gpurun_out/c24_bench_n${N}_la$la.json 2> gpurun_out/c24_bench_n${N}_la$la.err
echo "la=$la bench n$N rc=$?"; python - $N $la <<'PY'
import json, sys
try:
    r = json.loads([l for l in open("gpurun_out/c24_bench_n%s_la%s.json" % (sys.argv[1], sys.argv[2])) if l.startswith("{")][-1])
    q = r.get("sequence_row_split")
    print(q["value"], q["ms_per_step"], q.get("trace_us_median"))
except Exception as e:
    print("no json", e)
PY
done
