#!/bin/bash
# sharded-table mode on N GPUs: NCCL + peer-exchange tests, then the bench in both exchange modes
N=${1:-2}
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_shard_gpu.py -x -q -k nccl > gpurun_out/pytest_shard.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_shard.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
for mode in "peer 1"; do
  set -- $mode
  timeout 300 $TR --master-port 295$2$N bench.py --gpus $N --sharded --exchange $1 --chunks $2 --steps 5 --warmup 2 > gpurun_out/bench_sharded_$1$2_n$N.json 2> gpurun_out/bench_sharded_$1$2_n$N.err
  echo "$1 chunks=$2 rc=$?"; python - <<PY
import json
try:
    d=[json.loads(l) for l in open("gpurun_out/bench_sharded_$1$2_n$N.json") if l.startswith("{")][0]
    print(d["ms_per_step"], d["value"]/1e9, d["e2e"]["ms_per_step"], d["roofline"]["probe_ms"])
except Exception as e:
    print("no line", e); print(open("gpurun_out/bench_sharded_$1$2_n$N.err").read()[-1500:])
PY
done
