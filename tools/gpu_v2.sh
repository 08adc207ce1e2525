#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q > gpurun_out/pytest_v2.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_v2.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_v2.log 2> gpurun_out/bench_v2.err; echo "bench rc=$?"; tail -c 2500 gpurun_out/bench_v2.log; tail -5 gpurun_out/bench_v2.err
