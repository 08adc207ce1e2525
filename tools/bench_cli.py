#!/usr/bin/env python
"""End-to-end wall time of the drop-in binary (file -> CSV), rows A1/A2/A7/F2: synthetic genomes as FASTA
targets (database built by the CLI itself on first use), reads as a FASTQ file, then `bin/cuCLARK -n T`."""
import argparse, json, os, re, subprocess, sys, tempfile, time
import numpy as np

ap = argparse.ArgumentParser()
ap.add_argument("--genomes", type=int, default=100)
ap.add_argument("--genome-len", type=int, default=4_000_000)
ap.add_argument("--reads", type=int, default=10_000_000)
ap.add_argument("--threads", default="1,4,16")
ap.add_argument("--dir", default=None)
ap.add_argument("--anchor", action="store_true", help="one more run with CCB_ANCHOR=1 (target texts on the device), CSV compared")
ap.add_argument("--reference", default=None,
                help="also run the reference GPU binary (oracle/_ref/cuCLARK) with these '-n,-b' pairs, e.g. 1:4,16:32 "
                     "(needs a number of targets that is a multiple of 32, see tests/golden/ref_gpu/README.md)")
ap.add_argument("--also", default=None,
                help="more runs of bin/cuCLARK at the largest thread count with extra flags, ';'-separated, e.g. '-d 2;--sharded -d 2' "
                     "(two GPUs with the table replicated / sharded over them); CSVs compared with the first run's")
a = ap.parse_args()
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
d = a.dir or tempfile.mkdtemp(prefix="ccb_cli_")
os.makedirs(d, exist_ok=True)
rng = np.random.default_rng(7)
lut = np.frombuffer(b"ACGT", dtype=np.uint8)
t0 = time.time()
genomes = []
with open(os.path.join(d, "targets.txt"), "w") as tf:
    for g in range(a.genomes):
        seq = lut[rng.integers(0, 4, a.genome_len, dtype=np.uint8)]
        genomes.append(seq)
        p = os.path.join(d, f"g{g}.fa")
        rows = seq[: a.genome_len // 80 * 80].reshape(-1, 80)
        with open(p, "wb") as f:
            f.write(b">g%d\n" % g)
            f.write(np.concatenate([rows, np.full((rows.shape[0], 1), 10, np.uint8)], axis=1).tobytes())
            f.write(seq[a.genome_len // 80 * 80:].tobytes() + b"\n")
        tf.write(f"{p}\tT{g}\n")
# reads: 90 % windows of a genome with 1 % substitutions, 10 % random; fixed-width FASTQ records
N, L = a.reads, 150
rec = np.empty((N, 15 + 2 * L), dtype=np.uint8)
ids = np.arange(N)
rec[:, 0] = ord("@"); rec[:, 1] = ord("r")
for j in range(8):
    rec[:, 9 - j] = 48 + (ids // 10 ** j) % 10
rec[:, 10 - 0] = 10  # newline after the 8-digit id (columns 2..9)
allg = np.stack(genomes) if a.genomes * a.genome_len < 2_000_000_000 else None
gi = rng.integers(0, a.genomes, N); pos = rng.integers(0, a.genome_len - L, N)
CH = 1_000_000
for s in range(0, N, CH):
    e = min(N, s + CH)
    idx = pos[s:e, None] + np.arange(L)[None, :]
    seqs = (allg[gi[s:e, None], idx] if allg is not None else np.stack([genomes[g][p:p + L] for g, p in zip(gi[s:e], pos[s:e])]))
    sub = rng.random(seqs.shape) < 0.01
    seqs = np.where(sub, lut[rng.integers(0, 4, seqs.shape, dtype=np.uint8)], seqs)
    rnd = rng.random(e - s) < 0.10
    seqs[rnd] = lut[rng.integers(0, 4, (int(rnd.sum()), L), dtype=np.uint8)]
    rec[s:e, 11:11 + L] = seqs
rec[:, 11 + L] = 10; rec[:, 12 + L] = ord("+"); rec[:, 13 + L] = 10
rec[:, 14 + L:14 + 2 * L] = ord("I"); rec[:, 14 + 2 * L] = 10
fq = os.path.join(d, "reads.fq")
rec.tofile(fq)
t_gen = time.time() - t0
dbdir = os.path.join(d, "db"); os.makedirs(dbdir, exist_ok=True)
out = {"genomes": a.genomes, "genome_len": a.genome_len, "reads": N, "fastq_bytes": os.path.getsize(fq), "generate_s": t_gen, "runs": []}
exe = os.path.join(ROOT, "bin", "cuCLARK")
for i, T in enumerate([int(x) for x in a.threads.split(",")]):
    t0 = time.time()
    r = subprocess.run([exe, "-T", os.path.join(d, "targets.txt"), "-D", dbdir + "/", "-O", fq, "-R", os.path.join(d, f"res{T}"), "-n", str(T), "--verbose"],
                       capture_output=True, text=True)
    wall = time.time() - t0
    m = re.search(r"Done in ([0-9.]+)s \((\d+) reads/min, (\d+) reads\)", r.stderr.replace("\n", " "))
    out["runs"].append({"threads": T, "rc": r.returncode, "wall_s": wall, "built_db": "successfully stored" in r.stderr,
                        "classify_s": float(m.group(1)) if m else None, "reads_per_min": int(m.group(2)) if m else None,
                        "stderr_tail": r.stderr[-600:] if (r.returncode or not m) else None,
                        "phases": [l.strip() for l in r.stderr.splitlines() if "build:" in l or "host pipeline" in l or "Loading time" in l][:12]})
import hashlib
if a.anchor:
    T = max(int(x) for x in a.threads.split(","))
    t0 = time.time()
    r = subprocess.run([exe, "-T", os.path.join(d, "targets.txt"), "-D", dbdir + "/", "-O", fq, "-R", os.path.join(d, "res_anchor"), "-n", str(T), "--verbose"],
                       capture_output=True, text=True, env=dict(os.environ, CCB_ANCHOR="1"))
    wall = time.time() - t0
    m = re.search(r"Done in ([0-9.]+)s \((\d+) reads/min, (\d+) reads\)", r.stderr.replace("\n", " "))
    first = os.path.join(d, f"res{a.threads.split(',')[0]}.csv")
    out["anchor_run"] = {"threads": T, "rc": r.returncode, "wall_s": wall, "classify_s": float(m.group(1)) if m else None,
                         "reads_per_min": int(m.group(2)) if m else None,
                         "csv_identical_to_table_only": r.returncode == 0 and open(first, "rb").read() == open(os.path.join(d, "res_anchor.csv"), "rb").read(),
                         "phases": [l.strip() for l in r.stderr.splitlines() if "Anchor index" in l or "host pipeline" in l or "Loading time" in l][:6],
                         "stderr_tail": r.stderr[-600:] if (r.returncode or not m) else None}
if a.also:
    T = max(int(x) for x in a.threads.split(","))
    first = os.path.join(d, f"res{a.threads.split(',')[0]}.csv")
    out["also"] = []
    for i, extra in enumerate(x.strip() for x in a.also.split(";") if x.strip()):
        res = os.path.join(d, f"res_also{i}")
        t0 = time.time()
        r = subprocess.run([exe, "-T", os.path.join(d, "targets.txt"), "-D", dbdir + "/", "-O", fq, "-R", res, "-n", str(T), "--verbose"] + extra.split(),
                           capture_output=True, text=True)
        wall = time.time() - t0
        m = re.search(r"Done in ([0-9.]+)s \((\d+) reads/min, (\d+) reads\)", r.stderr.replace("\n", " "))
        out["also"].append({"flags": extra, "threads": T, "rc": r.returncode, "wall_s": wall, "classify_s": float(m.group(1)) if m else None,
                            "reads_per_min": int(m.group(2)) if m else None,
                            "csv_identical_to_first_run": r.returncode == 0 and open(first, "rb").read() == open(res + ".csv", "rb").read(),
                            "phases": [l.strip() for l in r.stderr.splitlines() if "host pipeline" in l or "Loading time" in l or "shard" in l.lower()][:8],
                            "stderr_tail": r.stderr[-600:] if (r.returncode or not m) else None})
if a.reference:
    ref = os.path.join(ROOT, "oracle", "_ref", "cuCLARK")
    out["reference_runs"] = []
    for pair in a.reference.split(","):
        n, b = pair.split(":")
        t0 = time.time()
        try:
            r = subprocess.run([ref, "-T", os.path.join(d, "targets.txt"), "-D", dbdir + "/", "-O", fq, "-R", os.path.join(d, f"ref{n}_{b}"),
                                "-n", n, "-b", b], capture_output=True, text=True, timeout=600)
            rc, err = r.returncode, r.stderr
        except subprocess.TimeoutExpired as e:
            rc, err = -999, "TIMEOUT " + str(e.stderr)[-300:]
        wall = time.time() - t0
        m = re.search(r"Done in ([0-9.]+)s \((\d+) reads/min, (\d+) reads\)", err.replace("\n", " "))
        same = None
        if rc == 0 and os.path.exists(os.path.join(d, f"ref{n}_{b}.csv")):
            mine = os.path.join(d, f"res{out['runs'][0]['threads']}.csv")
            same = open(mine, "rb").read() == open(os.path.join(d, f"ref{n}_{b}.csv"), "rb").read()
        out["reference_runs"].append({"n": int(n), "b": int(b), "rc": rc, "wall_s": wall, "classify_s": float(m.group(1)) if m else None,
                                      "reads_per_min": int(m.group(2)) if m else None, "csv_identical_to_ours": same,
                                      "stderr_tail": err[-400:] if (rc or not m) else None})
hs = {T["threads"]: hashlib.sha256(open(os.path.join(d, f"res{T['threads']}.csv"), "rb").read()).hexdigest() for T in out["runs"] if T["rc"] == 0}
out["csv_identical_across_threads"] = len(set(hs.values())) == 1
print(json.dumps(out))
