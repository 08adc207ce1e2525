#!/bin/bash
mkdir -p gpurun_out
CMD2="python bench.py --kernel-only --steps 2 --warmup 3 --genomes 250 --reads 2500000"
timeout 600 $CMD2 > gpurun_out/plain_small2.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_query -s 4 -c 1 -f -o gpurun_out/prof_r01_kquery_v2 $CMD2 > gpurun_out/ncu_full2.log 2>&1
echo "ncu full rc=$?"; cat gpurun_out/plain_small2.log | tail -2
tail -3 gpurun_out/ncu_full2.log
