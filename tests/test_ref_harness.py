"""A4/A5 pin: the REFERENCE's own CPU lookup (EHashtable::Read + queryElement = hTable::find, compiled in place
into oracle/_ref/ref_harness[-l], oracle/ref_harness.cc) under the oracle's read walk and vote gives the same
u16[5] per read as the oracle's own lookup (ccbo_db_find) on database files the unmodified reference wrote.
Runs wherever oracle/_ref exists (this container; the GPU box gets it with the snapshot).  No GPU needed."""
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
WORLD = os.path.join(HERE, "golden", "world")
HARNESS_L = os.path.join(ROOT, "oracle", "_ref", "ref_harness-l")


def _write_ref_db(npz, folder):
    g = np.load(npz)
    prefix = os.path.join(folder, str(g["name"]))
    sz = np.zeros(int(g["htsize"]), np.uint8)
    sz[g["nz_bucket"]] = g["nz_size"]
    sz.tofile(prefix + ".sz")
    g["ky"].tofile(prefix + ".ky")
    g["lb"].tofile(prefix + ".lb")
    return prefix


@pytest.mark.skipif(not os.path.exists(HARNESS_L), reason="oracle/_ref not built (make -C oracle ref needs /root/reference)")
@pytest.mark.parametrize("reads", ["reads.fa", "reads_ml.fa", "reads.fq"])
def test_reference_lookup_equals_oracle_lookup_light(oracle, tmp_path, reads):
    labels = open(os.path.join(WORLD, "labels.txt")).read().split()
    tp = str(tmp_path / "targets.txt")
    with open(tp, "w") as f:
        for i, lab in enumerate(labels):
            f.write(f"{os.path.join(WORLD, 'genomes', f'g{i}.fa')}\t{lab}\n")
    db = tmp_path / "db"
    db.mkdir()
    prefix = _write_ref_db(os.path.join(HERE, "golden", "ref_db", "light.npz"), str(db))
    dump = str(tmp_path / "final5.bin")
    src = os.path.join(WORLD, reads)
    r = subprocess.run([HARNESS_L, "-k", "27", "-T", tp, "-D", str(db) + "/", "-O", src, "--dump", dump, "-n", "2"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-1500:]
    got = np.fromfile(dump, np.uint16).reshape(-1, 5)
    odb = oracle.OracleDb.read(prefix, oracle.HTSIZE_LIGHT, 27)
    ir = oracle.IndexedReads(open(src, "rb").read())
    off, cont = ir.pack(27)
    want, _, nl = odb.query_packed(27, off, cont)
    assert got.shape == want.shape and np.array_equal(got, want)
    assert f"lookups={nl} " in r.stdout
    assert (want[:, 1] != 0).mean() > 0.3              # the case is not vacuous: reads do get classified


HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")


def _free_gb():
    try:
        import psutil
        return psutil.virtual_memory().available / 2 ** 30
    except Exception:
        return 0.0


@pytest.mark.skipif(not os.path.exists(HARNESS), reason="oracle/_ref not built (make -C oracle ref needs /root/reference)")
@pytest.mark.skipif(_free_gb() < 40, reason="the reference's full-mode hTable takes 25.8 GB for its 1 610 612 741 buckets")
def test_reference_lookup_equals_oracle_lookup_full(oracle, tmp_path):
    """Same pin in full mode (k = 31, u32 keys, HTSIZE 1 610 612 741): about 20 s, most of it the reference loading."""
    labels = open(os.path.join(WORLD, "labels.txt")).read().split()
    tp = str(tmp_path / "targets.txt")
    with open(tp, "w") as f:
        for i, lab in enumerate(labels):
            f.write(f"{os.path.join(WORLD, 'genomes', f'g{i}.fa')}\t{lab}\n")
    db = tmp_path / "db"
    db.mkdir()
    prefix = _write_ref_db(os.path.join(HERE, "golden", "ref_db", "full.npz"), str(db))
    dump = str(tmp_path / "final5.bin")
    src = os.path.join(WORLD, "reads.fq")
    r = subprocess.run([HARNESS, "-k", "31", "-T", tp, "-D", str(db) + "/", "-O", src, "--dump", dump, "-n", "2"],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-1500:]
    got = np.fromfile(dump, np.uint16).reshape(-1, 5)
    odb = oracle.OracleDb.read(prefix, oracle.HTSIZE_FULL, 31)
    ir = oracle.IndexedReads(open(src, "rb").read())
    off, cont = ir.pack(31)
    want, _, nl = odb.query_packed(31, off, cont)
    assert got.shape == want.shape and np.array_equal(got, want)
    assert f"lookups={nl} " in r.stdout
    assert (want[:, 1] != 0).mean() > 0.3
