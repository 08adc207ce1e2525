#!/bin/bash
# round 2, one GPU: the no-allocation entry point's test alone in a fresh process, then the whole GPU suite (it runs last there)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --tb=short > gpurun_out/r02_h_noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -5 gpurun_out/r02_h_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -q --tb=short > gpurun_out/r02_h_pytest_gpu.log 2>&1; echo "suite rc=$?"; tail -8 gpurun_out/r02_h_pytest_gpu.log
