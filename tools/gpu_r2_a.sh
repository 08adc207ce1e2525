#!/bin/bash
# round 2, first GPU call (2-GPU box): full GPU suite with no skips, e2e-gap bisect at N=1 (VERDICT weak #3),
# fill-granularity probe by request flavour (tools/probe_fill.cu) on the second GPU meanwhile
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r02_a_gpus.txt; nproc >> gpurun_out/r02_a_gpus.txt
timeout 900 python -m pytest tests -m gpu -q -rs > gpurun_out/r02_pytest_gpu_2gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r02_pytest_gpu_2gpu.log
( export CUDA_VISIBLE_DEVICES=1
  timeout 300 tools/probe_fill 4 28 > gpurun_out/r02_probe_fill.jsonl 2> gpurun_out/r02_probe_fill.err
  timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,lts__t_sectors_srcunit_tex_op_read.sum --clock-control none --csv \
      --log-file gpurun_out/r02_probe_fill_ncu.csv tools/probe_fill 4 26 > gpurun_out/r02_probe_fill_ncu.log 2>&1 ) &
# the driver's order: reference arm first, then our arm, with nvidia-smi polled at 4 Hz meanwhile
( while true; do nvidia-smi --query-gpu=index,clocks.sm,power.draw,utilization.gpu --format=csv,noheader -i 0 >> gpurun_out/r02_a_smi.log; sleep 0.25; done ) &
POLL=$!
export CUDA_VISIBLE_DEVICES=0
timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_a_ref.json 2> gpurun_out/r02_a_ref.err; echo "ref rc=$?"
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_a_bench_default.json 2> gpurun_out/r02_a_bench_default.err; echo "default rc=$?"
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r02_a_bench_nocpu.json 2> gpurun_out/r02_a_bench_nocpu.err; echo "nocpu rc=$?"
OMP_NUM_THREADS=1 timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_a_bench_omp1.json 2> gpurun_out/r02_a_bench_omp1.err; echo "omp1 rc=$?"
kill $POLL
wait
for f in default nocpu omp1; do python - "$f" <<'P'
import json,sys
f=sys.argv[1]
try:
    d=json.loads(open(f"gpurun_out/r02_a_bench_{f}.json").read().strip().splitlines()[-1])
    print(f, "kernel", round(d["ms_per_step"],2), "e2e", round(d["e2e"]["ms_per_step"],2), "steps", d["e2e"].get("host_step_ms"))
    a=d.get("anchor_index",{}); print("   anchor", a.get("ms_per_step"), a.get("e2e",{}).get("ms_per_step"))
except Exception as e: print(f, "ERR", e)
P
done
cat gpurun_out/r02_probe_fill.jsonl
