#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_anchor_gpu.py tests/test_parity_gpu.py tests/test_cli_gpu.py -x -q 2>&1 | tail -2
for i in 1 2; do
timeout 300 python bench.py --kernel-only --anchor --steps 10 --warmup 3 > gpurun_out/ko_exp_anchor.log 2>&1; echo "anchor $(tail -1 gpurun_out/ko_exp_anchor.log)"
timeout 300 python bench.py --kernel-only --steps 10 --warmup 3 > gpurun_out/ko_exp_plain.log 2>&1; echo "plain $(tail -1 gpurun_out/ko_exp_plain.log)"
done
