"""What the box gives the host pipeline: pinned host -> device copies (and device -> host beside them) on all GPUs at once.
   python -m torch.distributed.run --nproc-per-node N tools/probe_h2d.py [--numa]
prints one JSON line: GB/s per rank with the ranks copying one at a time and all together.  --numa binds each rank to the
CPUs of its GPU's memory node before the pinned buffers are allocated (what ccb_batch_alloc does, csrc/hostnuma.hpp).
Measurement tool, not on the product path."""
import json
import os
import sys

import torch
import torch.distributed as dist


def main():
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    node = None
    if "--numa" in sys.argv:
        bus = torch.cuda.get_device_properties(local).pci_bus_id
        dom = torch.cuda.get_device_properties(local).pci_domain_id
        dev = torch.cuda.get_device_properties(local).pci_device_id
        p = f"/sys/bus/pci/devices/{dom:04x}:{bus:02x}:{dev:02x}.0/numa_node"
        try:
            node = int(open(p).read())
            if node >= 0:
                cpus = set()
                for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
                    a, _, b = part.partition("-")
                    cpus |= set(range(int(a), int(b or a) + 1))
                os.sched_setaffinity(0, cpus & os.sched_getaffinity(0) or os.sched_getaffinity(0))
        except OSError:
            node = None
    NB = 440 << 20
    h_in = torch.empty(NB, dtype=torch.uint8).pin_memory(); h_in.fill_(1)
    h_out = torch.empty(100 << 20, dtype=torch.uint8).pin_memory()
    d_in = torch.empty(NB, dtype=torch.uint8, device="cuda"); d_out = torch.ones(100 << 20, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def run(reps, both):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record(s1)
        for _ in range(reps):
            with torch.cuda.stream(s1):
                d_in.copy_(h_in, non_blocking=True)
            if both:
                with torch.cuda.stream(s2):
                    h_out.copy_(d_out, non_blocking=True)
        e1.record(s1)
        torch.cuda.synchronize()
        return NB * reps / (e0.elapsed_time(e1) * 1e-3) / 1e9

    def gather(x):
        if world == 1:
            return [x]
        t = torch.zeros(world, dtype=torch.float64, device="cuda"); t[rank] = x
        dist.all_reduce(t)
        return [round(float(v), 1) for v in t]

    run(3, True)
    alone = 0.0
    for r in range(world):                     # one rank at a time
        if world > 1:
            dist.barrier()
        if r == rank:
            alone = run(10, False)
    if world > 1:
        dist.barrier()
    together = run(20, False)
    if world > 1:
        dist.barrier()
    together_both = run(20, True)
    a, b, c = gather(alone), gather(together), gather(together_both)
    nodes = gather(float(-9 if node is None else node))
    if rank == 0:
        print(json.dumps({"what": "pinned host -> device, 440 MB per copy, GB/s per rank", "n_gpus": world, "numa_bound": "--numa" in sys.argv,
                          "gpu_numa_node": nodes, "cpus": os.cpu_count(), "one_rank_at_a_time": a, "all_ranks_together": b,
                          "all_ranks_together_with_d2h_beside": c, "sum_together": round(sum(b), 1)}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
