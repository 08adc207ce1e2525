#!/usr/bin/env python
"""Random-access ceiling of one B200: 32-byte gathers over the resident table at several
footprints and access shapes (ccb_bench_gather_ex).  Writes one JSON object."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from jn_cuclark_b200 import CuClarkDB

G = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
db = CuClarkDB(1, 31, 1, min(G, 65534), load_factor=0.5)
db.synth_db(0xC1A2C000 + 2000, G, 4_000_000, min(G, 65534))
info = db.db_info()
stream = torch.cuda.current_stream().cuda_stream
NP = 1 << 30
res = {"l2_fetch": db.l2_fetch_granularity(), "table_gb": info.bytes / 1e9, "n_probes": NP, "rows": []}


def run(nbytes, mode, fp):
    db.bench_gather(NP // 4, nbytes, stream, mode, fp); torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); db.bench_gather(NP, nbytes, stream, mode, fp); b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    row = {"bytes": nbytes, "mode": mode, "footprint_gb": (fp or info.bytes) / 1e9, "ms": best,
           "gprobes_s": NP / best / 1e6, "gbs": NP * nbytes / best / 1e6}
    res["rows"].append(row)
    print(row, flush=True)


QUICK = len(sys.argv) > 2
print("l2 fetch granularity", res["l2_fetch"], flush=True)
for fp in ((0,) if QUICK else (32 << 20, 64 << 20, 128 << 20, 256 << 20, 512 << 20, 1 << 30, 4 << 30, 16 << 30, 0)):
    if fp and fp > info.bytes:
        continue
    run(32, 0, fp)
for mode in (1, 2, 3):
    run(32, mode, 0)
run(64, 0, 0)
run(128, 0, 0)
json.dump(res, open(os.path.join(ROOT, "gpurun_out", "gather_probe_l2f%d.json" % res["l2_fetch"]), "w"), indent=1)
db.close()
