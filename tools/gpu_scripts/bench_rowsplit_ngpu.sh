#!/bin/bash
# N GPUs: row-split sequence check + bench (N from $1)
N=${1:-2}
mkdir -p gpurun_out
if [ "$N" = "2" ]; then
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "sequence" > gpurun_out/c19_pytest_seq.log 2>&1; echo "pytest seq rc=$?"; tail -2 gpurun_out/c19_pytest_seq.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/seq_rowsplit_check.py > gpurun_out/c19_seqsplit.log 2>&1
echo "seqsplit rc=$?"; grep "rank" gpurun_out/c19_seqsplit.log | grep -v "^\[rank\|^W" | tail -12
for m in llama3-8b-q4_k_m llama3-70b-q4_k_m; do
  timeout 600 python tools/seq_bench.py $m 30 > gpurun_out/c19_seq_bench_$m.log 2>&1; echo "seq_bench $m rc=$?"; grep "graph replay\|bit-identical after" gpurun_out/c19_seq_bench_$m.log
done
fi
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/c19_bench_n$N.json 2> gpurun_out/c19_bench_n$N.err
echo "bench n$N rc=$?"; python - $N <<'PY'
import json, sys
try:
    r = json.loads([l for l in open("gpurun_out/c19_bench_n%s.json" % sys.argv[1]) if l.startswith("{")][-1])
    print("value", r["value"], r["config"]["workload"])
    print("seq_split", r.get("sequence_row_split"))
    print("single", r.get("single_gpu_same_workload"))
    print("nccl", r.get("nccl_all_gather_baseline"))
    print("pp512", r.get("pp512"))
    print("e2e", r.get("e2e"))
except Exception as e:
    print("no json", e)
PY
tail -5 gpurun_out/c19_bench_n$N.err
