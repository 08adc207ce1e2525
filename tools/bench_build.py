#!/usr/bin/env python
"""Database construction throughput (row F1): synthetic iid genomes as FASTA text in host memory ->
ccb_build_* -> (optionally) reference files + resident table.  Prints one JSON line."""
import argparse, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from jn_cuclark_b200 import CuClarkDB

ap = argparse.ArgumentParser()
ap.add_argument("--genomes", type=int, default=100)
ap.add_argument("--genome-len", type=int, default=4_000_000)
ap.add_argument("--k", type=int, default=31)
ap.add_argument("--light", action="store_true")
ap.add_argument("--write", default=None, help="prefix for .sz/.ky/.lb (default: no files)")
ap.add_argument("--verbose", action="store_true")
a = ap.parse_args()
HT = 57777779 if a.light else 1610612741
k = 27 if a.light else a.k
rng = np.random.default_rng(1)
lut = np.frombuffer(b"ACGT", dtype=np.uint8)
texts = []
t0 = time.time()
for g in range(a.genomes):
    seq = lut[rng.integers(0, 4, a.genome_len, dtype=np.uint8)]
    rows = seq[: a.genome_len // 80 * 80].reshape(-1, 80)
    body = np.concatenate([rows, np.full((rows.shape[0], 1), 10, np.uint8)], axis=1).tobytes() + seq[a.genome_len // 80 * 80:].tobytes() + b"\n"
    texts.append((b">g%d\n" % g + body, g % 65000))
t_gen = time.time() - t0
db = CuClarkDB(1, k, 1, min(a.genomes, 65000), verbose=a.verbose)
t0 = time.time()
info = db.build(texts, HT, 4, a.light, 4, 0, a.write, load_table=True)
t_build = time.time() - t0
print(json.dumps({"genomes": a.genomes, "genome_len": a.genome_len, "k": k, "light": a.light, "bases": info.bases,
                  "occurrences": info.occurrences, "distinct": info.distinct, "stored": info.stored, "max_bucket": info.max_bucket,
                  "build_s": t_build, "bases_per_s": info.bases / t_build, "host_generate_s": t_gen, "files": bool(a.write),
                  "table_gb": db.db_info().bytes / 1e9}))
db.close()
