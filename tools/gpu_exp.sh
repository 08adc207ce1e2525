#!/bin/bash
mkdir -p gpurun_out
for shape in 64x8 128x5 128x4; do
  CCB_BLOCK_SHAPE=$shape timeout 300 python bench.py --kernel-only --anchor --steps 10 --warmup 3 > gpurun_out/ko_anchor_$shape.log 2>&1; echo "$shape anchor rc=$? $(tail -1 gpurun_out/ko_anchor_$shape.log)"
done
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_exp.json 2> gpurun_out/bench_exp.err; echo "bench rc=$?"
python - <<'P'
import json
d=json.load(open('gpurun_out/bench_exp.json'))
print('plain',d['ms_per_step'],'e2e',d['e2e']['ms_per_step'],'anchor',d['anchor_index']['ms_per_step'],'e2e',d['anchor_index']['e2e']['ms_per_step'])
P
timeout 600 python -m pytest tests/test_parity_gpu.py tests/test_anchor_gpu.py -x -q 2>&1 | tail -2
