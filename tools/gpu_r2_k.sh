#!/bin/bash
# round 2, two GPUs: gpurun --gpus 2 -- 'bash tools/gpu_r2_k.sh'
#   the driver's bench line at N=2; staged against direct peer stores in the sharded route at 2 shards;
#   bin/cuCLARK on two GPUs, table replicated (-d 2) and sharded (--sharded -d 2), CSVs compared; the 2-GPU tests
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29555"
timeout 900 python -m pytest tests/test_shard_gpu.py tests/test_group_gpu.py tests/test_cli_gpu.py -m gpu -q --tb=short -rs > gpurun_out/r02_k_pytest_2gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r02_k_pytest_2gpu.log
timeout 900 $TR bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r02_bench_n2.json 2> gpurun_out/r02_bench_n2.err; echo "bench n=2 rc=$? $(cut -c1-300 gpurun_out/r02_bench_n2.json)"
SH="bench.py --gpus 2 --sharded --exchange peer --steps 5 --warmup 3 --no-cpu-baseline"
for st in 1 0; do
  CCB_ROUTE_STAGED=$st timeout 600 $TR $SH > gpurun_out/r02_k_sharded_staged${st}_n2.json 2> gpurun_out/r02_k_sharded_staged$st.err
  echo "staged=$st rc=$? $(python - <<P
import json
try:
    d=json.load(open("gpurun_out/r02_k_sharded_staged${st}_n2.json")); print("ms", round(d["ms_per_step"],2), d["config"]["phases_rank0"])
except Exception as e: print("no line", e)
P
)"
done
timeout 900 python tools/bench_cli.py --genomes 128 --threads 16 --also "-d 2;--sharded -d 2" > gpurun_out/r02_k_cli_2gpu.json 2> gpurun_out/r02_k_cli_2gpu.err; echo "cli rc=$? $(cut -c1-1500 gpurun_out/r02_k_cli_2gpu.json)"
