#!/bin/bash
# round 2, one GPU: the FASTQ streaming path walks the text once (index + pack + name copy per record, no survey pass):
# every test that goes through the host pipeline, then the command line end to end
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_cli_gpu.py tests/test_callers_gpu.py tests/test_ref_shim_gpu.py -m gpu -q --tb=short > gpurun_out/r02_o_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r02_o_pytest.log
timeout 900 python tools/bench_cli.py --genomes 128 --threads 16,8,4 > gpurun_out/r02_o_cli.json 2> gpurun_out/r02_o_cli.err; echo "cli rc=$?"
python - <<'P'
import json
d=json.load(open("gpurun_out/r02_o_cli.json"))
for r in d["runs"]: print(r["threads"], r["rc"], r["classify_s"], r["reads_per_min"], r["phases"][-1][:260] if r["phases"] else r["stderr_tail"])
print("identical across threads", d["csv_identical_across_threads"])
P
