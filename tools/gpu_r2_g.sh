#!/bin/bash
# round 2, seventh GPU call (one GPU): look-ahead in the anchor scan, denser tables
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_anchor_gpu.py -m gpu -q --tb=short > gpurun_out/r02_g_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02_g_pytest.log
timeout 600 python bench.py --kernel-only --no-gather --steps 10 --warmup 3 > gpurun_out/r02_g_ko.json 2>/dev/null; echo "ko $(cat gpurun_out/r02_g_ko.json)"
for lf in 0.6 0.75 0.85; do
  timeout 600 python bench.py --kernel-only --no-gather --steps 5 --warmup 3 --load-factor $lf > gpurun_out/r02_g_ko_lf$lf.json 2> gpurun_out/r02_g_ko_lf$lf.err; echo "load $lf: $(cat gpurun_out/r02_g_ko_lf$lf.json) $(tail -1 gpurun_out/r02_g_ko_lf$lf.err | cut -c1-200)"
done
KO="python bench.py --kernel-only --steps 3 --warmup 3 --no-gather"
timeout 600 $KO > gpurun_out/r02_g_ko.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k "regex:k_anchor_scan|k_query_block" -c 10 --csv --log-file gpurun_out/r02_g_launches.csv $KO > gpurun_out/r02_g_ncu.log 2>&1
python - <<'P'
import csv
rows=list(csv.reader(open("gpurun_out/r02_g_launches.csv"))); hdr=None; out={}
for r in rows:
    if "Kernel Name" in r: hdr=r; continue
    if hdr and len(r)==len(hdr):
        x=dict(zip(hdr,r)); out.setdefault((int(x["ID"]),x["Kernel Name"].split("(")[0][-42:]),{})[x["Metric Name"]]=float(x["Metric Value"].replace(",",""))
for (i,k),m in sorted(out.items())[:10]: print(i,k,"ms %.3f"%(m["gpu__time_duration.sum"]/1e6),"GB %.1f"%((m["dram__bytes_read.sum"]+m["dram__bytes_write.sum"])/1e9))
P
