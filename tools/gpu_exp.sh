#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_anchor_gpu.py tests/test_cli_gpu.py -x -q 2>&1 | tail -2
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo "bench rc=$?"; cut -c1-160 gpurun_out/bench_final.json
