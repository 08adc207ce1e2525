#!/bin/bash
# peer-sharded bench on N GPUs: one round per step, two rounds back to back, two rounds overlapped on two streams
N=${1:-2}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_shard_gpu.py -x -q 2>&1 | tail -2
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
i=0
for extra in "" "--chunks 2" "--chunks 2 --overlap" "--chunks 4 --overlap"; do
  tag=$(echo "$extra" | tr -d ' -'); i=$((i+1))
  timeout 300 $TR --master-port $((29513+i)) bench.py --gpus $N --sharded --exchange peer --steps 5 --warmup 2 $extra > gpurun_out/bench_sharded_peer_n${N}_$tag.json 2> gpurun_out/bench_sharded_peer_n${N}_$tag.err; echo "sharded [$extra] rc=$?"
  python - <<PY
import json
try:
    d=[json.loads(l) for l in open("gpurun_out/bench_sharded_peer_n${N}_$tag.json") if l.startswith("{")][0]
    print(d["n_gpus"], "ms/step", d["ms_per_step"], "G lookups/s", d["value"]/1e9, "e2e ms", d["e2e"]["ms_per_step"], d["config"].get("phases_rank0"))
except Exception as e:
    print("no line", e); print(open("gpurun_out/bench_sharded_peer_n${N}_$tag.err").read()[-1200:])
PY
done
