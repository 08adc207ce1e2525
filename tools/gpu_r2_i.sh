#!/bin/bash
# round 2, multi-GPU A/B call: gpurun --gpus N -- 'bash tools/gpu_r2_i.sh N'
#   topology of the box; what its PCIe gives all GPUs at once; pinned slots bound to the GPU's memory node or not;
#   the sharded route with staged (256-byte) or direct (32-byte at 8 GPUs) stores into the owners' inboxes
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544"
{ nvidia-smi topo -m; lscpu | grep -i -E "numa|socket|model name|^cpu\(s\)"; free -g | head -2; } > gpurun_out/r02_i_topo_n$N.txt 2>&1
timeout 300 $TR tools/probe_h2d.py > gpurun_out/r02_i_h2d_n$N.json 2> gpurun_out/r02_i_h2d.err; echo "h2d $(cat gpurun_out/r02_i_h2d_n$N.json)"
timeout 300 $TR tools/probe_h2d.py --numa > gpurun_out/r02_i_h2d_numa_n$N.json 2>> gpurun_out/r02_i_h2d.err; echo "h2d numa $(cat gpurun_out/r02_i_h2d_numa_n$N.json)"
E2E="bench.py --gpus $N --steps 10 --warmup 3 --no-workloads --no-sharded-leg --no-gather --no-cpu-baseline --no-cli-e2e"
for numa in 1 0; do
  CCB_NUMA=$numa timeout 600 $TR $E2E > gpurun_out/r02_i_e2e_numa${numa}_n$N.json 2> gpurun_out/r02_i_e2e_numa$numa.err
  echo "numa=$numa rc=$? $(python - <<P
import json
try:
    d=json.load(open("gpurun_out/r02_i_e2e_numa${numa}_n$N.json")); print("kernel ms", round(d["ms_per_step"],2), "e2e ms", round(d["e2e"]["ms_per_step"],2), d["e2e"]["host_step_ms"], "feed GB/s", round(d["e2e"]["host_feed_gbs"],1))
except Exception as e: print("no line", e)
P
)"; tail -2 gpurun_out/r02_i_e2e_numa$numa.err | cut -c1-300
done
SH="bench.py --gpus $N --sharded --exchange peer --steps 5 --warmup 3 --no-cpu-baseline"
for st in 1 0; do
  CCB_ROUTE_STAGED=$st timeout 600 $TR $SH > gpurun_out/r02_i_sharded_staged${st}_n$N.json 2> gpurun_out/r02_i_sharded_staged$st.err
  echo "staged=$st rc=$? $(python - <<P
import json
try:
    d=json.load(open("gpurun_out/r02_i_sharded_staged${st}_n$N.json")); print("ms", round(d["ms_per_step"],2), d["config"]["phases_rank0"])
except Exception as e: print("no line", e)
P
)"; tail -2 gpurun_out/r02_i_sharded_staged$st.err | cut -c1-300
done
