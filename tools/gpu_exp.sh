#!/bin/bash
mkdir -p gpurun_out
timeout 1200 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_anchor_gpu.py -x -q > gpurun_out/memcheck_anchor.log 2>&1; echo "memcheck anchor rc=$?"; grep -c "Invalid\|ERROR SUMMARY" gpurun_out/memcheck_anchor.log; tail -4 gpurun_out/memcheck_anchor.log
timeout 1200 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_parity_gpu.py -x -q -k "not synthetic" > gpurun_out/memcheck_parity.log 2>&1; echo "memcheck parity rc=$?"; tail -4 gpurun_out/memcheck_parity.log
