#!/usr/bin/env python
"""Does any load flavour avoid the 128-byte L2 line fill?  Small table (ncu replays stay cheap)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from jn_cuclark_b200 import CuClarkDB
db = CuClarkDB(1, 31, 1, 100, load_factor=0.5)
db.synth_db(1, 100, 4_000_000, 100)            # 6.4 GB table
st = torch.cuda.current_stream().cuda_stream
NP = 1 << 28
for mode, nb in ((0, 32), (2, 32), (3, 32), (4, 32), (5, 32), (6, 32), (7, 32), (0, 64)):
    db.bench_gather(NP, nb, st, mode, 0); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); db.bench_gather(NP, nb, st, mode, 0); b.record(); torch.cuda.synchronize()
    print({"mode": mode, "bytes": nb, "gprobes_s": NP / a.elapsed_time(b) / 1e6}, flush=True)
db.close()
