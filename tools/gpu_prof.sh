#!/bin/bash
# gather probe + launch list (full config) + one full ncu capture (reduced config)
mkdir -p gpurun_out
timeout 600 python tools/probe_gather.py > gpurun_out/gather_probe.log 2>&1; echo "probe rc=$?"
cat gpurun_out/gather_probe.log | tail -20
CMD="python bench.py --kernel-only --steps 3 --warmup 3"
timeout 600 $CMD > gpurun_out/plain_full.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_r01.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
CMD2="python bench.py --kernel-only --steps 2 --warmup 3 --genomes 250 --reads 2500000"
timeout 600 $CMD2 > gpurun_out/plain_small.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_query -s 4 -c 1 -f -o gpurun_out/prof_r01_kquery $CMD2 > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
tail -5 gpurun_out/ncu_full.log
ls -la gpurun_out
