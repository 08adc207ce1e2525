#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_final.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_final.log
timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/smoke_final.log 2>&1; echo "smoke rc=$?"
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo "bench rc=$?"; cut -c1-200 gpurun_out/bench_final.json
timeout 900 python tools/bench_cli.py --threads 16,16,4 --anchor > gpurun_out/cli_e2e.json 2> gpurun_out/cli_e2e.err; echo "cli rc=$?"
python - <<'P'
import json
d=json.load(open('gpurun_out/cli_e2e.json'))
for r in d['runs']: print(r['threads'], r['wall_s'], r['classify_s'], r['phases'][-1:])
print(d.get('anchor_run')['phases'], d['anchor_run']['csv_identical_to_table_only'], d['csv_identical_across_threads'])
P
