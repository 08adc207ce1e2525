#!/bin/bash
# round 2, multi-GPU call: gpurun --gpus N -- 'bash tools/gpu_r2_n8.sh N'
#   1. the driver's own bench line at N GPUs (replicated headline + the sharded leg with results_equal_replicated_path)
#   2. N = 8 only: BASELINE.json's scale-out database (configs[4]): 20 000 genomes x 4 Mbp = 80 G 31-mers, at load
#      factor 0.6 = 1.07 TB of table = 133 GB per GPU (0.5 would be 160 GB + inboxes: too close to the 192 GB there are);
#      peer exchange.  Falls back to 16 000 genomes at load 0.5 (128 GB per GPU) if the denser table does not build.
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
nvidia-smi --query-gpu=index,name,memory.total,clocks.max.sm --format=csv > gpurun_out/r02_n${N}_gpus.csv
timeout 900 $TR bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r02_bench_n$N.json 2> gpurun_out/r02_bench_n$N.err
echo "bench n=$N rc=$? $(cut -c1-1500 gpurun_out/r02_bench_n$N.json)"; tail -3 gpurun_out/r02_bench_n$N.err
if [ "$N" = 8 ]; then
  timeout 900 $TR bench.py --gpus 8 --sharded --exchange peer --genomes 20000 --load-factor 0.6 --steps 5 --warmup 3 --no-cpu-baseline \
      > gpurun_out/r02_sharded_c5_n8.json 2> gpurun_out/r02_sharded_c5_n8.err
  rc=$?; echo "c5 rc=$rc $(cut -c1-1500 gpurun_out/r02_sharded_c5_n8.json)"; tail -3 gpurun_out/r02_sharded_c5_n8.err
  if [ $rc != 0 ]; then
    timeout 600 $TR bench.py --gpus 8 --sharded --exchange peer --genomes 16000 --steps 5 --warmup 3 --no-cpu-baseline \
        > gpurun_out/r02_sharded_16k_n8.json 2> gpurun_out/r02_sharded_16k_n8.err
    echo "16k rc=$? $(cut -c1-1500 gpurun_out/r02_sharded_16k_n8.json)"; tail -3 gpurun_out/r02_sharded_16k_n8.err
  fi
fi
