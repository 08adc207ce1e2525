#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q > gpurun_out/pytest_v1.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_v1.log
timeout 600 python bench.py --kernel-only --steps 10 --warmup 3 2>&1 | tail -1
