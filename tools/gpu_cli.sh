#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu2.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu2.log
timeout 1500 python tests/golden/make_golden.py gpu gpurun_out/ref_gpu > gpurun_out/make_golden_gpu.log 2>&1; echo "golden rc=$?"
ls -la gpurun_out/ref_gpu; tail -40 gpurun_out/ref_gpu/log.txt
