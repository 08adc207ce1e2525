#!/bin/bash
# round-end evidence, one ncu pass per gpurun call:  tools/gpu_final.sh a|b|c
mkdir -p gpurun_out
case "$1" in
a)  # full tests, smoke, bench (both arms), launch list
    timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_final.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_final.log
    timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/smoke_final.log 2>&1; echo "smoke rc=$?"
    timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo "bench rc=$?"; cut -c1-700 gpurun_out/bench_final.json
    timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref_final.json 2> gpurun_out/bench_ref_final.err; echo "benchref rc=$?"
    CMD="python bench.py --kernel-only --steps 3 --warmup 3"
    timeout 600 $CMD > gpurun_out/plain_full.log 2>&1 && \
    timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_final.csv $CMD > gpurun_out/ncu_launches.log 2>&1
    echo "launch list rc=$?" ;;
b)  # DRAM traffic of k_query on the full configuration
    CMD="python bench.py --kernel-only --steps 3 --warmup 3"
    timeout 600 $CMD > gpurun_out/plain_full.log 2>&1 && \
    timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,dram__sectors_read.sum --clock-control none -k regex:k_query -s 4 -c 2 --csv --log-file gpurun_out/traffic_final.csv $CMD > gpurun_out/ncu_traffic.log 2>&1
    echo "traffic rc=$?" ;;
c)  # full-section capture on a quarter-size workload (replay of a 64 GB kernel x40 is too slow)
    CMD2="python bench.py --kernel-only --steps 2 --warmup 3 --genomes 250 --reads 2500000"
    timeout 600 $CMD2 > gpurun_out/plain_small.log 2>&1 && \
    timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_query -s 4 -c 1 -f -o gpurun_out/prof_final_kquery $CMD2 > gpurun_out/ncu_full.log 2>&1
    echo "ncu full rc=$?" ;;
esac
