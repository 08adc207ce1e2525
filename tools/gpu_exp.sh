#!/bin/bash
mkdir -p gpurun_out
CCB_BLOCK_SHAPE=128x4s timeout 600 python -m pytest tests/test_anchor_gpu.py tests/test_parity_gpu.py -x -q 2>&1 | tail -2
for shape in 128x4 128x4s 64x8s 256x2s; do
  CCB_BLOCK_SHAPE=$shape timeout 200 python bench.py --kernel-only --anchor --steps 10 --warmup 3 > gpurun_out/ko_anchor_$shape.log 2>&1; echo "$shape anchor rc=$? $(tail -1 gpurun_out/ko_anchor_$shape.log)"
done
for shape in 256x3 128x4 128x4s; do
  CCB_BLOCK_SHAPE=$shape timeout 200 python bench.py --kernel-only --steps 10 --warmup 3 > gpurun_out/ko_plain_$shape.log 2>&1; echo "$shape plain rc=$? $(tail -1 gpurun_out/ko_plain_$shape.log)"
done
