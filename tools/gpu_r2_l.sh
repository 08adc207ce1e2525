#!/bin/bash
# round 2, one GPU: what a range pass of the owner-side probe costs -- a 128 GB table (2 000 genomes) on one B200 through the
# sharded pipeline (world = 1), probed in 1 / 2 / 3 / 4 / 6 ranges (CCB_PROBE_RANGE_MB)
mkdir -p gpurun_out
for mb in 200000 65536 49152 32768 22000; do
  CCB_PROBE_RANGE_MB=$mb timeout 600 python bench.py --sharded --exchange peer --genomes 2000 --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/r02_l_range$mb.json 2> gpurun_out/r02_l_range$mb.err
  echo "range $mb MB rc=$? $(python - <<P
import json
try:
    d=json.load(open("gpurun_out/r02_l_range$mb.json")); print("ms", round(d["ms_per_step"],2), d["config"]["phases_rank0"])
except Exception as e: print("no line", e)
P
)"; tail -1 gpurun_out/r02_l_range$mb.err | cut -c1-200
done
