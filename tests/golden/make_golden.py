#!/usr/bin/env python
"""Generates the golden fixtures under tests/golden/ by running the UNMODIFIED reference
(oracle/_ref/cuCLARK, cuCLARK-l built by oracle/Makefile from /root/reference).

  python tests/golden/make_golden.py world     # deterministic inputs (numpy) -> tests/golden/world/
  python tests/golden/make_golden.py db        # reference CPU database build (runs anywhere; the
                                               # binary stops at its CUDA probe after writing the DB)
                                               # -> tests/golden/ref_db/{light,full}.npz
  python tests/golden/make_golden.py csv       # reference host engine (unmodified: parse, pack, merge, CSV)
                                               # + reference hash-table lookup + CPU stand-in for the
                                               # device kernels (oracle/ref_cpu_device.cc) -> tests/golden/ref_csv/
  python tests/golden/make_golden.py gpu OUT   # on a GPU box: reference classification CSVs -> OUT/
                                               # (copied to tests/golden/ref_gpu/ afterwards)
"""
import hashlib
import os
import shutil
import subprocess
import sys
import tempfile
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import synth  # noqa: E402

WORLD = os.path.join(HERE, "world")
REF = os.path.join(ROOT, "oracle", "_ref")


def make_world(n_genomes=6, length=12000, seed=21):
    shutil.rmtree(WORLD, ignore_errors=True)
    tp, genomes, labels = synth.make_genomes(os.path.join(WORLD, "genomes"), n_genomes, length, seed=seed)
    os.remove(tp)                                    # absolute paths: regenerated where used
    with open(os.path.join(WORLD, "labels.txt"), "w") as f:
        f.write("\n".join(labels) + "\n")
    reads = synth.make_reads(genomes, 400, 150, seed=seed + 1, k=31)
    synth.write_reads_fasta(os.path.join(WORLD, "reads.fa"), reads)
    synth.write_reads_fasta(os.path.join(WORLD, "reads_ml.fa"), reads, width=60)
    synth.write_reads_fastq(os.path.join(WORLD, "reads.fq"), [r for r in reads if " " not in r[0] and len(r[1])])
    r1, r2 = synth.make_pairs(genomes, 150, seed=seed + 2)
    synth.write_reads_fastq(os.path.join(WORLD, "p1.fq"), r1, "/1")
    synth.write_reads_fastq(os.path.join(WORLD, "p2.fq"), r2, "/2")


def write_targets(dst_dir):
    labels = open(os.path.join(WORLD, "labels.txt")).read().split()
    tp = os.path.join(dst_dir, "targets.txt")
    with open(tp, "w") as f:
        for i, lab in enumerate(labels):
            f.write(f"{os.path.join(WORLD, 'genomes', f'g{i}.fa')}\t{lab}\n")
    return tp


def run_ref(binary, args, cwd, timeout=600):
    try:
        return subprocess.run([os.path.join(REF, binary)] + args, cwd=cwd, capture_output=True, text=True, timeout=timeout)
    except subprocess.TimeoutExpired as e:
        class R:
            returncode = -999
            stdout = (e.stdout or b"").decode(errors="ignore") if isinstance(e.stdout, bytes) else (e.stdout or "")
            stderr = "TIMEOUT " + ((e.stderr or b"").decode(errors="ignore") if isinstance(e.stderr, bytes) else (e.stderr or ""))
        return R()


def make_db():
    out = os.path.join(HERE, "ref_db")
    os.makedirs(out, exist_ok=True)
    for mode, binary, k in (("light", "cuCLARK-l", 27), ("full", "cuCLARK", 31)):
        with tempfile.TemporaryDirectory() as d:
            tp = write_targets(d)
            os.makedirs(os.path.join(d, "db"))
            r = run_ref(binary, ["-k", str(k), "-T", tp, "-D", os.path.join(d, "db") + "/", "-O",
                                 os.path.join(WORLD, "reads.fa"), "-R", os.path.join(d, "out")], d)
            files = sorted(os.listdir(os.path.join(d, "db")))
            assert len(files) == 3, (files, r.stderr[-2000:])
            prefix = os.path.join(d, "db", files[0][:-3])
            sz = np.fromfile(prefix + ".sz", dtype=np.uint8)
            nz = np.flatnonzero(sz)
            sha = {e: hashlib.sha256(open(prefix + "." + e, "rb").read()).hexdigest() for e in ("sz", "ky", "lb")}
            np.savez_compressed(os.path.join(out, mode + ".npz"), name=os.path.basename(prefix), htsize=sz.size,
                                nz_bucket=nz.astype(np.uint32), nz_size=sz[nz],
                                ky=np.fromfile(prefix + ".ky", dtype=np.uint8), lb=np.fromfile(prefix + ".lb", dtype=np.uint16),
                                sha_sz=sha["sz"], sha_ky=sha["ky"], sha_lb=sha["lb"])
            print(mode, os.path.basename(prefix), "k-mers:", int(sz.sum()), sha)


def make_gpu(out, suffix=""):
    os.makedirs(out, exist_ok=True)
    log = open(os.path.join(out, "log.txt"), "w")
    for mode, binary, k in (("light", "cuCLARK-l" + suffix, 27), ("full", "cuCLARK" + suffix, 31)):
        d = tempfile.mkdtemp()
        tp = write_targets(d)
        os.makedirs(os.path.join(d, "db"))
        base = ["-k", str(k), "-T", tp, "-D", os.path.join(d, "db") + "/"]
        runs = [("fa", ["-O", os.path.join(WORLD, "reads.fa")]),
                ("ml", ["-O", os.path.join(WORLD, "reads_ml.fa")]),
                ("fq", ["-O", os.path.join(WORLD, "reads.fq")]),
                ("pe", ["-P", os.path.join(WORLD, "p1.fq"), os.path.join(WORLD, "p2.fq")]),
                ("fa_ext", ["-O", os.path.join(WORLD, "reads.fa"), "--extended"]),
                ("pe_ext", ["-P", os.path.join(WORLD, "p1.fq"), os.path.join(WORLD, "p2.fq"), "--extended"]),
                # -n 4 -b 4 crashes inside the reference's own engine on the CPU stand-in (its printer thread
                # races with the packers); the batch split is exercised single-threaded there
                ("fa_b4n4", ["-O", os.path.join(WORLD, "reads.fa"), "-n", "4", "-b", "4"]) if not suffix else
                ("fa_b4", ["-O", os.path.join(WORLD, "reads.fa"), "-b", "4"])]
        if mode == "full":                           # each full-mode run reloads a 1.6 GB .sz: keep to three
            runs = [r for r in runs if r[0] in ("fq", "fa_ext", "pe_ext")]
        for name, extra in runs:
            res = os.path.join(d, name)
            t0 = time.time()
            r = run_ref(binary, base + extra + ["-R", res], d)
            log.write(f"-- {time.time() - t0:.1f} s\n")
            log.write(f"== {mode} {name} rc={r.returncode}\n{r.stdout[-1500:]}\n{r.stderr[-1500:]}\n")
            if os.path.exists(res + ".csv"):
                shutil.copy(res + ".csv", os.path.join(out, f"{mode}_{name}.csv"))
            log.flush()
        shutil.rmtree(d, ignore_errors=True)
    log.close()


def make_gpu32(out):
    """The reference's GPU binaries on this box's GPU, world32 (32 targets).  With a number of targets that
    is not a multiple of 32 their queryKernel never returns on sm_70+ (lanes beyond numTargets skip the
    loop around `__ballot_sync(0xFFFFFFFF, ...)`, src/CuClarkDB.cu:1131-1136); with 32 it completes.  The
    databases are written by bin/cuCLARK[-l] first (byte-identical to the reference's, test_build_gpu.py),
    so the reference binaries only load them."""
    os.makedirs(out, exist_ok=True)
    log = open(os.path.join(out, "log32.txt"), "w")
    fa, fq, ml = (["-O", os.path.join(WORLD, f)] for f in ("reads.fa", "reads.fq", "reads_ml.fa"))
    pe = ["-P", os.path.join(WORLD, "p1.fq"), os.path.join(WORLD, "p2.fq")]
    plans = [("light", "cuCLARK-l", 27, [("fa", fa), ("fq", fq), ("ml", ml), ("pe", pe), ("fa_ext", fa + ["--extended"]), ("pe_ext", pe + ["--extended"])]),
             ("full", "cuCLARK", 31, [("fa", fa), ("fq", fq), ("pe", pe), ("fa_ext", fa + ["--extended"]),
                                    ("fa_s2", fa + ["-s", "2"])]),      # kept as refbug_full_fa_s2.csv, see ref_gpu/README.md
             ("fullk21", "cuCLARK", 21, [("fa", fa), ("pe", pe)])]
    for mode, binary, k, runs in plans:
        d = tempfile.mkdtemp()
        tp = write_targets(d)
        os.makedirs(os.path.join(d, "db"))
        base = ["-k", str(k), "-T", tp, "-D", os.path.join(d, "db") + "/"]
        r = subprocess.run([os.path.join(ROOT, "bin", binary)] + base + ["-O", os.path.join(WORLD, "reads.fa"), "-R", os.path.join(d, "mine")],
                           capture_output=True, text=True, timeout=600)
        log.write(f"== {mode} database by bin/{binary} rc={r.returncode}\n{r.stderr[-600:]}\n")
        for name, extra in runs:
            res = os.path.join(d, name)
            t0 = time.time()
            r = run_ref(binary, base + extra + ["-R", res], d, timeout=150)
            log.write(f"-- {time.time() - t0:.1f} s\n== {mode} {name} rc={r.returncode}\n{r.stdout[-800:]}\n{r.stderr[-800:]}\n")
            if r.returncode == 0 and os.path.exists(res + ".csv"):
                shutil.copy(res + ".csv", os.path.join(out, f"w32_{mode}_{name}.csv"))
            log.flush()
            if r.returncode == -999:
                break                                 # a hang: do not spend more GPU time on this mode
        shutil.rmtree(d, ignore_errors=True)
    log.close()


if __name__ == "__main__":
    cmd = sys.argv[1]
    if cmd.endswith("32"):
        WORLD = os.path.join(HERE, "world32")
    if cmd == "world32":
        make_world(33, 8000, seed=31)
    elif cmd == "gpu32":
        make_gpu32(sys.argv[2])
        sys.exit(0)
    if cmd == "world32":
        pass
    elif cmd == "world":
        make_world()
    elif cmd == "db":
        make_db()
    elif cmd == "gpu":
        make_gpu(sys.argv[2])
    elif cmd == "csv":
        # reference host engine + CPU stand-in for its device layer (oracle/ref_cpu_device.cc): runs anywhere
        make_gpu(os.path.join(HERE, "ref_csv"), "_cpu")
