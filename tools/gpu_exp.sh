#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_cli_gpu.py tests/test_parity_gpu.py -x -q 2>&1 | tail -2
timeout 900 python tools/bench_cli.py --threads 16,16,4 --anchor > gpurun_out/cli_e2e.json 2> gpurun_out/cli_e2e.err; echo "cli rc=$?"
python - <<'P'
import json
d=json.load(open('gpurun_out/cli_e2e.json'))
for r in d['runs']: print(r['threads'], r['wall_s'], r['classify_s'], r['phases'][-1:])
print(d.get('anchor_run')['phases'], d['anchor_run']['csv_identical_to_table_only'], d['csv_identical_across_threads'])
P
