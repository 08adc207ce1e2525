#!/bin/bash
# round 2, fifth GPU call (2-GPU box): re-test what changed, N=2 bench with the pipelined sharded variant, a database that
# does not fit one HBM (3,000 genomes = 12 G 31-mers = 192 GB of table) sharded over the two GPUs, then on one GPU each:
# the N=1 line + reference arm + reference GPU binary (GPU 0) and the ncu evidence (GPU 1)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 1200 python -m pytest tests/test_cli_gpu.py tests/test_shard_gpu.py tests/test_group_gpu.py tests/test_abi.py tests/test_callers_gpu.py tests/test_bench_gpu.py -q -rs --tb=short > gpurun_out/r02_e_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/r02_e_pytest.log
timeout 900 $TR --master-port 29521 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r02_e_bench_n2.json 2> gpurun_out/r02_e_bench_n2.err; echo "bench n2 rc=$?"
timeout 900 $TR --master-port 29522 bench.py --gpus 2 --sharded --exchange peer --genomes 3000 --steps 5 --warmup 3 > gpurun_out/r02_e_sharded_bigdb_n2.json 2> gpurun_out/r02_e_sharded_bigdb_n2.err; echo "bigdb rc=$? $(cut -c1-400 gpurun_out/r02_e_sharded_bigdb_n2.json)"
timeout 900 $TR --master-port 29523 bench.py --gpus 2 --sharded --exchange peer --genomes 3000 --chunks 4 --overlap --steps 5 --warmup 3 > gpurun_out/r02_e_sharded_bigdb_n2_overlap.json 2> gpurun_out/r02_e_sharded_bigdb_n2_overlap.err; echo "bigdb overlap rc=$? $(cut -c1-300 gpurun_out/r02_e_sharded_bigdb_n2_overlap.json)"
(   export CUDA_VISIBLE_DEVICES=1
    KO="python bench.py --kernel-only --steps 3 --warmup 3 --no-gather"
    M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,dram__sectors_read.sum
    timeout 600 $KO > gpurun_out/r02_e_ko.log 2>&1 && \
    timeout 900 ncu --metrics $M --clock-control none -k regex:k_query_block -c 14 --csv --log-file gpurun_out/r02_e_traffic_kblock.csv $KO > gpurun_out/r02_e_ncu1.log 2>&1
    echo "traffic rc=$?"
    timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/r02_e_launches_ko.csv $KO > gpurun_out/r02_e_ncu2.log 2>&1
    echo "launches rc=$?"
    Q="python bench.py --kernel-only --steps 2 --warmup 3 --no-gather --genomes 250 --reads 2500000"
    timeout 600 $Q > gpurun_out/r02_e_ko_small.log 2>&1 && \
    timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_query_block -s 2 -c 1 -f -o gpurun_out/r02_prof_kblock_anchor_ld64 $Q > gpurun_out/r02_e_ncu3.log 2>&1
    echo "full rc=$?"
    S="python bench.py --sharded --exchange peer --genomes 250 --reads 2500000 --steps 2 --warmup 3"
    timeout 600 $S > gpurun_out/r02_e_sharded_1gpu.json 2> gpurun_out/r02_e_sharded_1gpu.err && \
    timeout 900 ncu --metrics $M --clock-control none -k "regex:k_route_block|k_probe_inbox|k_vote_block|k_query<|k_scan|k_post" -c 60 --csv --log-file gpurun_out/r02_e_traffic_sharded.csv $S > gpurun_out/r02_e_ncu4.log 2>&1
    echo "sharded traffic rc=$?"
) > gpurun_out/r02_e_gpu1.log 2>&1 &
export CUDA_VISIBLE_DEVICES=0
timeout 900 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_e_bench_ref.json 2> gpurun_out/r02_e_bench_ref.err; echo "ref rc=$?"
timeout 1200 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_e_bench_n1.json 2> gpurun_out/r02_e_bench_n1.err; echo "bench n1 rc=$?"; tail -3 gpurun_out/r02_e_bench_n1.err
wait
cat gpurun_out/r02_e_gpu1.log
python - <<'P'
import json
for f in ("r02_e_bench_n2","r02_e_bench_n1"):
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, "ms", round(d["ms_per_step"],2), "e2e", round(d["e2e"]["ms_per_step"],2), "frac", round(d["roofline"]["frac"],3), "table_only", round(d["table_only"]["ms_per_step"],2))
        if "sharded" in d:
            print("   sharded", {k: (round(v["ms_per_step"],1), v["results_equal_replicated_path"]) for k,v in d["sharded"]["exchanges"].items()}, d["sharded"]["exchanges"]["peer"].get("phases_rank0_ms"))
        if "workloads" in d: print("   workloads", {k:(round(v["ms_per_step"],2), v.get("parity_sample")) for k,v in d["workloads"].items()})
        if "cli_e2e" in d: print("   cli", d["cli_e2e"])
        if "cpu_baseline" in d: print("   cpu", d["cpu_baseline"]["value"], d["cpu_baseline"].get("value_1core"), d.get("parity_sample"))
    except Exception as e: print(f, "ERR", e)
P
