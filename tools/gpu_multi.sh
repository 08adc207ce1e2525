#!/bin/bash
# N-GPU validation: the bench through torchrun (both arms), then the sharded mode with the peer exchange
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L | head -8
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d=[json.loads(l) for l in open("gpurun_out/bench_n$N.json") if l.startswith("{")][0]
    print(d["n_gpus"], d["ms_per_step"], d["value"]/1e9, d["e2e"]["value"]/1e9, d["clocks"])
except Exception as e:
    print("no line", e); print(open("gpurun_out/bench_n$N.err").read()[-1500:])
PY
timeout 300 $TR --master-port 29512 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err; echo "benchref rc=$?"; cut -c1-160 gpurun_out/bench_ref_n$N.json
timeout 400 $TR --master-port 29513 bench.py --gpus $N --sharded --exchange peer --steps 5 --warmup 2 > gpurun_out/bench_sharded_peer_n$N.json 2> gpurun_out/bench_sharded_peer_n$N.err; echo "sharded rc=$?"
python - <<PY
import json
try:
    d=[json.loads(l) for l in open("gpurun_out/bench_sharded_peer_n$N.json") if l.startswith("{")][0]
    print(d["n_gpus"], d["ms_per_step"], d["value"]/1e9, d["e2e"]["ms_per_step"], d["config"]["phases_rank0"])
except Exception as e:
    print("no line", e); print(open("gpurun_out/bench_sharded_peer_n$N.err").read()[-1500:])
PY
