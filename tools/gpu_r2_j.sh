#!/bin/bash
# round 2, one GPU: EXPERIMENT (code removed again, profiles/r02_scan_overlap_ab.json): the anchor pass of most of a batch on a second stream beside the probes of its first piece -- tests, then A/B over CCB_SCAN_OVERLAP
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_anchor_gpu.py tests/test_parity_gpu.py tests/test_cli_gpu.py tests/test_zz_no_alloc.py -m gpu -q --tb=short > gpurun_out/r02_j_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r02_j_pytest.log
for ov in 0 8 4 16 2 0 8; do
  CCB_SCAN_OVERLAP=$ov timeout 300 python bench.py --kernel-only --no-gather --steps 10 --warmup 3 > gpurun_out/r02_j_ko_ov$ov.json 2>/dev/null; echo "overlap 1/$ov: $(cat gpurun_out/r02_j_ko_ov$ov.json)"
done
for ov in 0 8; do
  CCB_SCAN_OVERLAP=$ov timeout 600 python bench.py --steps 10 --warmup 3 --no-workloads --no-gather --no-cpu-baseline --no-cli-e2e > gpurun_out/r02_j_line_ov$ov.json 2>/dev/null
  python - <<P
import json
d=json.load(open("gpurun_out/r02_j_line_ov$ov.json")); print("overlap 1/$ov line: ms", round(d["ms_per_step"],3), "e2e ms", round(d["e2e"]["ms_per_step"],3), "frac", round(d["roofline"]["frac"],3), "equal", d["config"]["anchor_index"]["results_equal_table_only_path"], d["e2e"]["results_equal_resident_path"])
P
done
