#!/bin/bash
# sharded-table mode on N GPUs: NCCL test, then the bench (sharded vs replicated on the same box)
N=${1:-2}
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_shard_gpu.py -x -q -k nccl > gpurun_out/pytest_shard.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_shard.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29521 bench.py --gpus $N --sharded --steps 5 --warmup 2 > gpurun_out/bench_sharded_n$N.json 2> gpurun_out/bench_sharded_n$N.err; echo "sharded rc=$?"; cut -c1-1500 gpurun_out/bench_sharded_n$N.json; tail -3 gpurun_out/bench_sharded_n$N.err
timeout 300 $TR --master-port 29522 bench.py --gpus $N --sharded --chunks 4 --steps 5 --warmup 2 > gpurun_out/bench_sharded4_n$N.json 2> gpurun_out/bench_sharded4_n$N.err; echo "sharded4 rc=$?"; cut -c1-400 gpurun_out/bench_sharded4_n$N.json
