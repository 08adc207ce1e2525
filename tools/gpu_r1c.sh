#!/bin/bash
# evidence for the anchor index + block kernel beside the table-only path:  tools/gpu_r1c.sh a|b
mkdir -p gpurun_out
case "$1" in
a)
    timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_r1c.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_r1c.log
    timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_r1c.json 2> gpurun_out/bench_r1c.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench_r1c.json
    timeout 300 python bench.py --kernel-only --lane --steps 10 --warmup 3 > gpurun_out/kernel_only_lane_plain.log 2>&1; echo "lane plain rc=$?"; tail -1 gpurun_out/kernel_only_lane_plain.log
    CMD="python bench.py --kernel-only --lane --anchor --steps 3 --warmup 3"
    timeout 300 $CMD > gpurun_out/kernel_only_lane_anchor.log 2>&1 && \
    timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_lane_anchor.csv $CMD > gpurun_out/ncu_launches_lane_anchor.log 2>&1
    echo "launch list rc=$?"; tail -1 gpurun_out/kernel_only_lane_anchor.log
    timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,dram__sectors_read.sum --clock-control none -k regex:k_query_block -s 4 -c 2 --csv --log-file gpurun_out/traffic_lane_anchor.csv $CMD > gpurun_out/ncu_traffic_lane_anchor.log 2>&1
    echo "traffic rc=$?" ;;
b)  # full-section capture of the block kernel with the anchor index, quarter-size workload
    CMD2="python bench.py --kernel-only --lane --anchor --steps 2 --warmup 3 --genomes 250 --reads 2500000"
    timeout 600 $CMD2 > gpurun_out/plain_small_lane.log 2>&1 && \
    timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_query_block -s 4 -c 1 -f -o gpurun_out/prof_block_anchor $CMD2 > gpurun_out/ncu_full_block.log 2>&1
    echo "ncu full rc=$?"; tail -1 gpurun_out/plain_small_lane.log ;;
esac
