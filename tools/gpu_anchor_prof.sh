#!/bin/bash
# ncu --set full of the thread-per-read kernel with the anchor index, quarter-size workload
mkdir -p gpurun_out
CMD2="python bench.py --kernel-only --lane --anchor --steps 2 --warmup 3 --genomes 250 --reads 2500000"
timeout 600 $CMD2 > gpurun_out/plain_small_lane.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_query_lane -s 4 -c 1 -f -o gpurun_out/prof_lane_anchor $CMD2 > gpurun_out/ncu_full_lane.log 2>&1
echo "ncu full rc=$?"; tail -1 gpurun_out/plain_small_lane.log
