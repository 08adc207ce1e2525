#!/bin/bash
# round 2, one GPU: where exactly does the translation cliff start?  table-only kernel on 70 / 77 GB tables, and a 134 GB
# table probed in two ranges of 67 GB against three of 45 GB (CCB_PROBE_RANGE_MB)
mkdir -p gpurun_out
for g in 1100 1200; do
  timeout 300 python bench.py --kernel-only --no-gather --no-anchor --steps 5 --warmup 3 --genomes $g > gpurun_out/r02_m_table_$g.json 2>/dev/null; echo "table-only, $g genomes: $(cat gpurun_out/r02_m_table_$g.json)"
done
for mb in 70000 49152; do
  CCB_PROBE_RANGE_MB=$mb timeout 600 python bench.py --sharded --exchange peer --genomes 2100 --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/r02_m_range$mb.json 2> gpurun_out/r02_m_range$mb.err
  echo "134 GB table, range $mb MB rc=$? $(python - <<P
import json
try:
    d=json.load(open("gpurun_out/r02_m_range$mb.json")); print("ms", round(d["ms_per_step"],2), d["config"]["phases_rank0"])
except Exception as e: print("no line", e)
P
)"; tail -1 gpurun_out/r02_m_range$mb.err | cut -c1-200
done
