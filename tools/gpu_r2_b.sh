#!/bin/bash
# round 2, second GPU call (2-GPU box): the new tests, A/B of the 64-byte fill hint, first N=2 bench with the sharded leg
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_shard_gpu.py tests/test_group_gpu.py tests/test_parity_gpu.py tests/test_ref_shim_gpu.py tests/test_callers_gpu.py \
    tests/test_cli_gpu.py tests/test_anchor_gpu.py tests/test_build_gpu.py -m gpu -q -rs --tb=short > gpurun_out/r02_b_pytest.log 2>&1; echo "pytest rc=$?"; tail -30 gpurun_out/r02_b_pytest.log
for v in "" _ld0; do
  CCB_LIBRARY=$PWD/jn_cuclark_b200/libcuclark_b200$v.so timeout 600 python bench.py --kernel-only --steps 10 --warmup 3 > gpurun_out/r02_b_ko$v.json 2> gpurun_out/r02_b_ko$v.err; echo "ko$v rc=$? $(cat gpurun_out/r02_b_ko$v.json)"
done
timeout 900 python -m pytest tests/test_bench_gpu.py -m gpu -q -rs --tb=short > gpurun_out/r02_b_pytest_bench.log 2>&1; echo "pytest bench rc=$?"; tail -15 gpurun_out/r02_b_pytest_bench.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 \
    > gpurun_out/r02_b_bench_n2.json 2> gpurun_out/r02_b_bench_n2.err; echo "bench n2 rc=$?"; tail -5 gpurun_out/r02_b_bench_n2.err; cut -c1-1500 gpurun_out/r02_b_bench_n2.json
