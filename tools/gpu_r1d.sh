#!/bin/bash
# block kernel v2 (direct window extraction, two sampled k-mers per round, interleaved text): parity + shapes
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_anchor_gpu.py tests/test_cli_gpu.py tests/test_parity_gpu.py -x -q > gpurun_out/pytest_r1d.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_r1d.log
for shape in 256x2 128x4 256x3 128x6; do
  CCB_BLOCK_SHAPE=$shape timeout 300 python bench.py --kernel-only --lane --anchor --steps 10 --warmup 3 > gpurun_out/ko_anchor_$shape.log 2>&1; echo "$shape anchor rc=$? $(tail -1 gpurun_out/ko_anchor_$shape.log)"
done
for shape in 256x2 256x3; do
  CCB_BLOCK_SHAPE=$shape timeout 300 python bench.py --kernel-only --lane --steps 10 --warmup 3 > gpurun_out/ko_plain_$shape.log 2>&1; echo "$shape plain rc=$? $(tail -1 gpurun_out/ko_plain_$shape.log)"
done
