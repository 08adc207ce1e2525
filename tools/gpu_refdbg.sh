#!/bin/bash
# where does the reference's GPU binary stall on B200?  its own DEBUG_* macros, tiny input, hard timeout
mkdir -p gpurun_out/ref_dbg; W=tests/golden/world; D=$(mktemp -d)
python - <<PY
import os
labels=open("$W/labels.txt").read().split()
open("$D/targets.txt","w").write("".join(f"{os.path.abspath('$W')}/genomes/g{i}.fa\t{l}\n" for i,l in enumerate(labels)))
PY
head -40 $W/reads.fa > $D/few.fa
timeout -s KILL 60 oracle/_ref/cuCLARK-l_dbg -T $D/targets.txt -D $D/ -O $D/few.fa -R $D/out -n 1 -b 1 > gpurun_out/ref_dbg/stdout.txt 2> gpurun_out/ref_dbg/stderr.txt
echo "rc=$?"; tail -c 1500 gpurun_out/ref_dbg/stderr.txt; echo; tail -c 600 gpurun_out/ref_dbg/stdout.txt; ls -la $D/out.csv 2>/dev/null && head -5 $D/out.csv && cp $D/out.csv gpurun_out/ref_dbg/
nvidia-smi --query-gpu=memory.used,utilization.gpu --format=csv,noheader
