#!/bin/bash
# peer-sharded bench on N GPUs (+ the 2-process parity test)
N=${1:-2}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_shard_gpu.py -x -q -k nccl 2>&1 | tail -2
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29513 bench.py --gpus $N --sharded --exchange peer --steps 5 --warmup 2 > gpurun_out/bench_sharded_peer_n$N.json 2> gpurun_out/bench_sharded_peer_n$N.err; echo "sharded rc=$?"
python - <<PY
import json
try:
    d=[json.loads(l) for l in open("gpurun_out/bench_sharded_peer_n$N.json") if l.startswith("{")][0]
    print(d["n_gpus"], d["ms_per_step"], d["value"]/1e9, d["e2e"]["ms_per_step"], d["config"]["phases_rank0"])
except Exception as e:
    print("no line", e); print(open("gpurun_out/bench_sharded_peer_n$N.err").read()[-1500:])
PY
