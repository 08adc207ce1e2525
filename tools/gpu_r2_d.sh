#!/bin/bash
# round 2, fourth GPU call (one GPU): does the table size slow a probe down (64 / 96 / 128 GB)?  which grid sizes for the
# sharded kernels?  then the N=1 line on the current build
mkdir -p gpurun_out
for g in 1000 1500 2000; do
  timeout 600 python bench.py --kernel-only --no-anchor --no-gather --genomes $g --steps 5 --warmup 3 > gpurun_out/r02_d_tablesize_$g.json 2> gpurun_out/r02_d_tablesize_$g.err; echo "table-only genomes=$g rc=$? $(cat gpurun_out/r02_d_tablesize_$g.json)"
done
run() { # name, env...
  n=$1; shift
  env "$@" timeout 600 python bench.py --sharded --exchange peer --steps 3 --warmup 3 > gpurun_out/r02_d_sh_$n.json 2> gpurun_out/r02_d_sh_$n.err
  python - "$n" <<'P'
import json,sys
n=sys.argv[1]
try:
    d=json.loads(open(f"gpurun_out/r02_d_sh_{n}.json").read().strip().splitlines()[-1])
    print(n, "ms/step", round(d["ms_per_step"],1), d["config"]["phases_rank0"])
except Exception as e: print(n, "ERR", e)
P
}
run r2p2v8 CCB_ROUTE_CTAS_PER_SM=2 CCB_PROBE_CTAS_PER_SM=2 CCB_VOTE_CTAS_PER_SM=8
run r3p4v8 CCB_ROUTE_CTAS_PER_SM=3 CCB_PROBE_CTAS_PER_SM=4 CCB_VOTE_CTAS_PER_SM=8
run r3p8v8 CCB_ROUTE_CTAS_PER_SM=3 CCB_PROBE_CTAS_PER_SM=8 CCB_VOTE_CTAS_PER_SM=8
run r3p32v32 CCB_ROUTE_CTAS_PER_SM=3 CCB_PROBE_CTAS_PER_SM=32 CCB_VOTE_CTAS_PER_SM=32
run r64p32v64 CCB_ROUTE_CTAS_PER_SM=64 CCB_PROBE_CTAS_PER_SM=32 CCB_VOTE_CTAS_PER_SM=64
env CCB_ROUTE_CTAS_PER_SM=2 CCB_PROBE_CTAS_PER_SM=2 timeout 600 python bench.py --sharded --exchange peer --chunks 4 --overlap --steps 3 --warmup 3 > gpurun_out/r02_d_sh_overlap_r2p2.json 2>/dev/null; echo "overlap r2p2: $(cut -c1-230 gpurun_out/r02_d_sh_overlap_r2p2.json)"
env CCB_ROUTE_CTAS_PER_SM=3 CCB_PROBE_CTAS_PER_SM=32 timeout 600 python bench.py --sharded --exchange peer --chunks 4 --overlap --steps 3 --warmup 3 > gpurun_out/r02_d_sh_overlap_r3p32.json 2>/dev/null; echo "overlap r3p32: $(cut -c1-230 gpurun_out/r02_d_sh_overlap_r3p32.json)"
env CCB_PROBE_CTAS_PER_SM=32 timeout 600 python bench.py --sharded --exchange peer --genomes 1500 --steps 3 --warmup 3 > gpurun_out/r02_d_sh_96gb.json 2>/dev/null; echo "sharded 96 GB on one GPU: $(python -c "import json;d=json.loads(open('gpurun_out/r02_d_sh_96gb.json').read().strip().splitlines()[-1]);print(d['ms_per_step'], d['config']['phases_rank0'])")"
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_d_bench_n1.json 2> gpurun_out/r02_d_bench_n1.err; echo "bench n1 rc=$?"
python - <<'P'
import json
d=json.loads(open("gpurun_out/r02_d_bench_n1.json").read().strip().splitlines()[-1])
print("n1 ms", round(d["ms_per_step"],2), "e2e", round(d["e2e"]["ms_per_step"],2), "frac", round(d["roofline"]["frac"],3), "table_only", round(d["table_only"]["ms_per_step"],2), {k:(round(v["ms_per_step"],2), v.get("parity_sample")) for k,v in d["workloads"].items()}, d["cli_e2e"].get("reads_per_s"), d["cli_e2e"].get("pipeline"))
P
