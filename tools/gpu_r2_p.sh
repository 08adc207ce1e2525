#!/bin/bash
# round 2, one GPU: k_anchor_scan with the sampled-k-mer search in lockstep (bitmask over the first 64 windows) and the
# k-mer length as a compile-time constant: anchor + parity tests, then the kernel-only A/B numbers and the launch list
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_anchor_gpu.py tests/test_parity_gpu.py tests/test_zz_no_alloc.py -m gpu -q --tb=short > gpurun_out/r02_p_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r02_p_pytest.log
KO="python bench.py --kernel-only --no-gather --steps 10 --warmup 3"
timeout 300 $KO > gpurun_out/r02_p_ko.json 2>/dev/null; echo "split: $(cat gpurun_out/r02_p_ko.json)"
CCB_FUSED_ANCHOR=1 timeout 300 $KO > gpurun_out/r02_p_ko_fused.json 2>/dev/null; echo "fused: $(cat gpurun_out/r02_p_ko_fused.json)"
KO3="python bench.py --kernel-only --no-gather --steps 3 --warmup 3"
timeout 300 $KO3 > gpurun_out/r02_p_ko3.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__thread_inst_executed_per_inst_executed.ratio,sm__inst_executed.sum --clock-control none -k "regex:k_anchor_scan|k_query_block" -s 6 -c 4 --csv --log-file gpurun_out/r02_p_launches.csv $KO3 > gpurun_out/r02_p_ncu.log 2>&1
python - <<'P'
import csv
rows=list(csv.reader(open("gpurun_out/r02_p_launches.csv"))); hdr=None; out={}
for r in rows:
    if "Kernel Name" in r: hdr=r; continue
    if hdr and len(r)==len(hdr):
        x=dict(zip(hdr,r)); out.setdefault((int(x["ID"]),x["Kernel Name"].split("(")[0][-40:]),{})[x["Metric Name"]]=float(x["Metric Value"].replace(",",""))
for (i,k),m in sorted(out.items()): print(i,k,{a.split("__")[-1][:22]:round(v/1e6,3) if v>1e4 else round(v,2) for a,v in m.items()})
P
