#!/bin/bash
mkdir -p gpurun_out
python bench.py --anchor --steps 10 --warmup 3 --no-cpu-baseline --no-gather > gpurun_out/bench_anchor.json 2> gpurun_out/bench_anchor.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/bench_anchor.json")); print(d["ms_per_step"], d["value"]/1e9, d["config"]["sectors_per_lookup"], d["e2e"]["ms_per_step"], d["roofline"]["frac"])
PY
CMD2="python bench.py --kernel-only --anchor --steps 2 --warmup 3 --genomes 250 --reads 2500000"
timeout 600 $CMD2 > gpurun_out/plain_small_anchor.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_query -s 4 -c 1 -f -o gpurun_out/prof_anchor_kquery $CMD2 > gpurun_out/ncu_full_anchor.log 2>&1
echo "ncu full rc=$?"
