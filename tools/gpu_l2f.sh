#!/bin/bash
mkdir -p gpurun_out
for g in 128 64 32; do
  CCB_L2_FETCH=$g timeout 300 python tools/probe_gather.py 1000 quick > gpurun_out/gather_l2f$g.log 2>&1; echo "probe $g rc=$?"; cat gpurun_out/gather_l2f$g.log
  CCB_L2_FETCH=$g timeout 300 python bench.py --kernel-only --steps 5 --warmup 3 2>&1 | tail -1
done
