#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
if [ "$N" = "2" ]; then
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "sequence" > gpurun_out/c25_pytest_seq.log 2>&1; echo "pytest seq rc=$?"; tail -3 gpurun_out/c25_pytest_seq.log
for ll in 1 0; do
SEQ_SPLIT_LL=$ll timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/seq_rowsplit_check.py > gpurun_out/c25_seqsplit_ll$ll.log 2>&1
echo "seqsplit ll=$ll rc=$?"; grep "^rank" gpurun_out/c25_seqsplit_ll$ll.log | tail -6
done
fi
for ll in 1 0; do
B200_SEQ_SPLIT_LL=$ll B200_BENCH_SEQ_TRACE=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 3 --no-pp --no-single > gpurun_out/c25_bench_n${N}_ll$ll.json 2> gpurun_out/c25_bench_n${N}_ll$ll.err
echo "ll=$ll bench n$N rc=$?"; python - $N $ll <<'PY'
import json, sys
try:
    r = json.loads([l for l in open("gpurun_out/c25_bench_n%s_ll%s.json" % (sys.argv[1], sys.argv[2])) if l.startswith("{")][-1])
    q = r.get("sequence_row_split")
    print(q.get("value"), q.get("ms_per_step"), q.get("trace_us_median"), q.get("error"))
    print("e2e", r["e2e"]["value"], "headline", r["value"], r["config"]["workload"])
except Exception as e:
    print("no json", e)
PY
tail -3 gpurun_out/c25_bench_n${N}_ll$ll.err
done
