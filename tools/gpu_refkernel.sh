#!/bin/bash
# kernel time of the reference's own queryKernel on this B200 (baseline #2 of SURVEY section 8 D4), same
# workload through our binary beside it.  128 targets: the reference kernel needs a multiple of 32.
mkdir -p gpurun_out; W=/tmp/ccb_refk
timeout 900 python tools/bench_cli.py --genomes 128 --reads 2000000 --threads 16 --reference 16:32 --dir $W > gpurun_out/refk_plain.json 2> gpurun_out/refk_plain.err || { echo plain failed; tail -3 gpurun_out/refk_plain.err; exit 1; }
REF="oracle/_ref/cuCLARK -T $W/targets.txt -D $W/db/ -O $W/reads.fq -R $W/refncu -n 16 -b 32"
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum --clock-control none -k regex:'queryKernel|resultKernel|mergeKernel' -c 100 --csv --log-file gpurun_out/ref_kernels.csv $REF > gpurun_out/ref_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ref_ncu.log
