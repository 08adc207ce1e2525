#!/bin/bash
# round 2, sixth GPU call (one GPU): the split anchor path -- parity tests, A/B against the fused kernel, launch list
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_anchor_gpu.py tests/test_parity_gpu.py "tests/test_cli_gpu.py::test_single_end" tests/test_bench_gpu.py -m gpu -q -rs --tb=short > gpurun_out/r02_f_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r02_f_pytest.log
for v in 0 1; do
  CCB_FUSED_ANCHOR=$v timeout 600 python bench.py --kernel-only --no-gather --steps 10 --warmup 3 > gpurun_out/r02_f_ko_fused$v.json 2>/dev/null; echo "fused=$v $(cat gpurun_out/r02_f_ko_fused$v.json)"
done
KO="python bench.py --kernel-only --steps 3 --warmup 3 --no-gather"
timeout 600 $KO > gpurun_out/r02_f_ko.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 40 --csv --log-file gpurun_out/r02_f_launches.csv $KO > gpurun_out/r02_f_ncu.log 2>&1
echo "launch list rc=$?"
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cli-e2e > gpurun_out/r02_f_bench_n1.json 2> gpurun_out/r02_f_bench_n1.err; echo "bench rc=$?"
python - <<'P'
import json,csv
d=json.loads(open("gpurun_out/r02_f_bench_n1.json").read().strip().splitlines()[-1])
print("n1 ms", round(d["ms_per_step"],2), "e2e", round(d["e2e"]["ms_per_step"],2), "frac", round(d["roofline"]["frac"],3), "table_only", round(d["table_only"]["ms_per_step"],2), {k:(round(v["ms_per_step"],2), v.get("parity_sample")) for k,v in d["workloads"].items()}, d.get("parity_sample"))
rows=list(csv.reader(open("gpurun_out/r02_f_launches.csv"))); hdr=None; out={}
for r in rows:
    if "Kernel Name" in r: hdr=r; continue
    if hdr and len(r)==len(hdr):
        x=dict(zip(hdr,r)); out.setdefault((int(x["ID"]),x["Kernel Name"].split("(")[0][-42:]),{})[x["Metric Name"]]=float(x["Metric Value"].replace(",",""))
for (i,k),m in sorted(out.items())[:24]: print(i,k,"ms %.3f"%(m["gpu__time_duration.sum"]/1e6),"GB %.1f"%((m["dram__bytes_read.sum"]+m["dram__bytes_write.sum"])/1e9))
P
