#!/bin/bash
# round 2, final evidence on one GPU:  tools/gpu_r2_final.sh a|b   (each ncu pass only after the same command exited 0 without ncu)
mkdir -p gpurun_out
KO="python bench.py --kernel-only --steps 3 --warmup 3 --no-gather"
case "$1" in
a)  # the whole GPU suite, smoke, both bench arms, A/B of the fused anchor kernel and of denser tables, launch list with DRAM bytes
    timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02_final_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02_final_pytest_gpu.log
    timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/r02_final_smoke.log 2>&1; echo "smoke rc=$? $(tail -1 gpurun_out/r02_final_smoke.log | cut -c1-200)"
    t0=$(date +%s)
    timeout 1200 python bench.py > gpurun_out/r02_final_bench_n1.json 2> gpurun_out/r02_final_bench_n1.err; echo "bench rc=$? wall $(( $(date +%s) - t0 )) s"; cut -c1-400 gpurun_out/r02_final_bench_n1.json
    t0=$(date +%s)
    timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r02_final_bench_reference.json 2> gpurun_out/r02_final_bench_reference.err; echo "reference arm rc=$? wall $(( $(date +%s) - t0 )) s"; cut -c1-300 gpurun_out/r02_final_bench_reference.json
    timeout 300 python bench.py --kernel-only --no-gather --steps 10 --warmup 3 > gpurun_out/r02_final_ko_split.json 2>/dev/null; echo "split: $(cat gpurun_out/r02_final_ko_split.json)"
    CCB_FUSED_ANCHOR=1 timeout 300 python bench.py --kernel-only --no-gather --steps 10 --warmup 3 > gpurun_out/r02_final_ko_fused.json 2>/dev/null; echo "fused: $(cat gpurun_out/r02_final_ko_fused.json)"
    for lf in 0.6 0.75 0.85; do
      timeout 300 python bench.py --kernel-only --no-gather --steps 5 --warmup 3 --load-factor $lf > gpurun_out/r02_final_ko_lf$lf.json 2> gpurun_out/r02_final_ko_lf$lf.err
      echo "load $lf: $(cat gpurun_out/r02_final_ko_lf$lf.json) $(grep -i -m1 "too full\|error" gpurun_out/r02_final_ko_lf$lf.err | cut -c1-160)"
    done
    timeout 600 $KO > gpurun_out/r02_final_ko.log 2>&1 && \
    timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k "regex:k_anchor_scan|k_query" -c 24 --csv --log-file gpurun_out/r02_final_launches.csv $KO > gpurun_out/r02_final_ncu_launches.log 2>&1
    echo "launch list rc=$?" ;;
b)  # full-section captures of the two kernels of the default path on a quarter-size workload (replaying a 64 GB kernel x40 is too slow)
    Q="python bench.py --kernel-only --no-gather --steps 2 --warmup 3 --genomes 250 --reads 2500000"
    timeout 600 $Q > gpurun_out/r02_final_ko_small.log 2>&1 && \
    timeout 1500 ncu --set full --clock-control none --import-source on -k "regex:k_anchor_scan|k_query_block" -s 8 -c 2 -f -o gpurun_out/r02_final_prof $Q > gpurun_out/r02_final_ncu_full.log 2>&1
    echo "ncu full rc=$?"
    ncu -i gpurun_out/r02_final_prof.ncu-rep --page details --csv > gpurun_out/r02_final_ncu_details.csv 2>/dev/null
    ncu -i gpurun_out/r02_final_prof.ncu-rep --page raw --csv > gpurun_out/r02_final_ncu_raw.csv 2>/dev/null
    ls -la gpurun_out/r02_final_prof.ncu-rep ;;
esac
