#!/bin/bash
echo quarter; python bench.py --kernel-only --steps 20 --warmup 3 --reads 2500000 2>&1 | tail -1
for s in 2 8 16; do echo slots$s; python bench.py --steps 10 --warmup 3 --slots $s --no-cpu-baseline --no-gather 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['e2e']['ms_per_step'])"; done
