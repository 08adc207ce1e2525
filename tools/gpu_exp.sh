#!/bin/bash
mkdir -p gpurun_out
for sl in 4 16 32; do
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-gather --slots $sl > gpurun_out/bench_slots$sl.json 2> gpurun_out/bench_slots$sl.err; echo "slots $sl rc=$?"
python - <<P
import json
d=json.load(open('gpurun_out/bench_slots$sl.json'))
print('plain',d['ms_per_step'],'e2e',d['e2e']['ms_per_step'],'anchor',d['anchor_index']['ms_per_step'],'e2e',d['anchor_index']['e2e']['ms_per_step'])
P
done
