#!/bin/bash
# round 2, one GPU: the command-line tests once more after the FASTQ chunk walker became a function of its own
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_cli_gpu.py -m gpu -q --tb=short -x > gpurun_out/r02_q_pytest_cli.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r02_q_pytest_cli.log
