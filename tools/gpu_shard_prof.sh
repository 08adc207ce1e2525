#!/bin/bash
# per-kernel times of the sharded pipeline on ONE GPU (1 shard: the exchange is local, the kernels are the same)
mkdir -p gpurun_out
CMD="python bench.py --sharded --exchange peer --steps 3 --warmup 2 --genomes 500 --reads 5000000"
timeout 600 $CMD > gpurun_out/shard1_plain.json 2> gpurun_out/shard1_plain.err || { echo plain failed; tail -5 gpurun_out/shard1_plain.err; exit 1; }
cut -c1-300 gpurun_out/shard1_plain.json
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'k_query|k_probe|k_vote|k_post|DeviceScan' -c 60 --csv --log-file gpurun_out/launches_sharded.csv $CMD > gpurun_out/ncu_shard.log 2>&1
echo "ncu rc=$?"
