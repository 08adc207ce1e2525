#!/usr/bin/env python
"""Line-reuse probes: how cheap are the other three sectors of a line once one has been fetched?"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from jn_cuclark_b200 import CuClarkDB
db = CuClarkDB(1, 31, 1, 500, load_factor=0.5)
db.synth_db(1, 500, 4_000_000, 500)            # 32 GB table
st = torch.cuda.current_stream().cuda_stream
NP = 1 << 28
for mode in (3, 11, 12):
    n = NP * 4 if mode == 3 else NP
    db.bench_gather(n, 32, st, mode, 0); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); db.bench_gather(n, 32, st, mode, 0); b.record(); torch.cuda.synchronize()
    print({"mode": mode, "glines_s": NP / a.elapsed_time(b) / 1e6, "gsectors_s": 4 * NP / a.elapsed_time(b) / 1e6}, flush=True)
db.close()
