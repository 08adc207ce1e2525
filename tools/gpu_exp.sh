#!/bin/bash
B="python bench.py --kernel-only --steps 10 --warmup 3"
for c in 16 32 64 128 256 1024; do echo cap$c; CCB_KQ_CAP=$c $B 2>&1 | tail -1; done
echo cap64lf04; CCB_KQ_CAP=64 $B --load-factor 0.4 2>&1 | tail -1
echo cap64lf06; CCB_KQ_CAP=64 $B --load-factor 0.6 2>&1 | tail -1
