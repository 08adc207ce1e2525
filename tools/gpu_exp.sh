#!/bin/bash
mkdir -p gpurun_out
for shape in 256x2 256x3 128x6 128x4; do
  CCB_BLOCK_SHAPE=$shape timeout 300 python bench.py --kernel-only --steps 10 --warmup 3 > gpurun_out/ko_plain_$shape.log 2>&1; echo "$shape plain rc=$? $(tail -1 gpurun_out/ko_plain_$shape.log)"
done
