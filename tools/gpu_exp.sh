#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_exp.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_exp.log
timeout 300 python bench.py --kernel-only --steps 10 --warmup 3 > gpurun_out/ko_exp_plain.log 2>&1; echo "plain $(tail -1 gpurun_out/ko_exp_plain.log)"
timeout 300 python bench.py --kernel-only --anchor --steps 10 --warmup 3 > gpurun_out/ko_exp_anchor.log 2>&1; echo "anchor $(tail -1 gpurun_out/ko_exp_anchor.log)"
