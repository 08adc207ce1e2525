#!/bin/bash
# round-end evidence:  tools/gpu_final.sh a|b|c   (each ncu pass only after the same command exited 0 without ncu)
mkdir -p gpurun_out
KO="python bench.py --kernel-only --steps 3 --warmup 3"
case "$1" in
a)  # full tests, smoke, bench (both arms), launch lists
    timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_final.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_final.log
    timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/smoke_final.log 2>&1; echo "smoke rc=$?"
    timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo "bench rc=$?"; cut -c1-300 gpurun_out/bench_final.json
    timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref_final.json 2> gpurun_out/bench_ref_final.err; echo "benchref rc=$?"
    timeout 300 python bench.py --kernel-only --warp --steps 10 --warmup 3 > gpurun_out/ko_warp.log 2>&1; echo "warp kernel: $(tail -1 gpurun_out/ko_warp.log)"
    timeout 300 python bench.py --kernel-only --read-len 10000 --reads 125000 --steps 10 --warmup 3 > gpurun_out/ko_long.log 2>&1; echo "10-kbp reads (C4 shape, 125 k reads per GPU): $(tail -1 gpurun_out/ko_long.log)"
    timeout 600 $KO > gpurun_out/ko_block.log 2>&1 && \
    timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_final.csv $KO > gpurun_out/ncu_launches.log 2>&1
    echo "launch list rc=$?"
    timeout 600 $KO --anchor > gpurun_out/ko_block_anchor.log 2>&1 && \
    timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_final_anchor.csv $KO --anchor > gpurun_out/ncu_launches_anchor.log 2>&1
    echo "launch list (anchor) rc=$?" ;;
b)  # DRAM traffic of k_query_block on the full configuration, table only and with the anchor index
    M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,dram__sectors_read.sum
    timeout 600 $KO > gpurun_out/ko_block.log 2>&1 && \
    timeout 900 ncu --metrics $M --clock-control none -k regex:k_query_block -s 4 -c 2 --csv --log-file gpurun_out/traffic_final.csv $KO > gpurun_out/ncu_traffic.log 2>&1
    echo "traffic rc=$?"
    timeout 600 $KO --anchor > gpurun_out/ko_block_anchor.log 2>&1 && \
    timeout 900 ncu --metrics $M --clock-control none -k regex:k_query_block -s 4 -c 2 --csv --log-file gpurun_out/traffic_final_anchor.csv $KO --anchor > gpurun_out/ncu_traffic_anchor.log 2>&1
    echo "traffic (anchor) rc=$?" ;;
c)  # full-section captures on a quarter-size workload (replaying a 64 GB kernel x40 is too slow)
    Q="python bench.py --kernel-only --steps 2 --warmup 3 --genomes 250 --reads 2500000"
    timeout 600 $Q > gpurun_out/ko_small.log 2>&1 && \
    timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_query_block -s 4 -c 1 -f -o gpurun_out/prof_final_kblock $Q > gpurun_out/ncu_full.log 2>&1
    echo "ncu full rc=$?"
    timeout 600 $Q --anchor > gpurun_out/ko_small_anchor.log 2>&1 && \
    timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_query_block -s 4 -c 1 -f -o gpurun_out/prof_final_kblock_anchor $Q --anchor > gpurun_out/ncu_full_anchor.log 2>&1
    echo "ncu full (anchor) rc=$?" ;;
esac
