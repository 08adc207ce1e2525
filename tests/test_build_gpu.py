"""Database construction on the GPU (rows A8/A9/F1 of SURVEY.md section 8): the files ccb_build_end writes
are byte-identical to the ones the UNMODIFIED reference binary wrote for the same targets
(tests/golden/ref_db/*.npz hold their SHA-256), the oracle's builder agrees on synthetic edge cases, and
the table left resident classifies like the oracle."""
import hashlib
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
WORLD = os.path.join(HERE, "golden", "world")


def _world_targets():
    labels = open(os.path.join(WORLD, "labels.txt")).read().split()
    order = []
    for lab in labels:
        if lab not in order:
            order.append(lab)
    return [(open(os.path.join(WORLD, "genomes", f"g{i}.fa"), "rb").read(), order.index(lab)) for i, lab in enumerate(labels)], order


def _sha(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        for blk in iter(lambda: f.read(1 << 24), b""):
            h.update(blk)
    return h.hexdigest()


@pytest.mark.parametrize("mode,env", [("light", {}), ("full", {}), ("full", {"CCB_BUILD_STAGE": "4099"}),
                                      ("light", {"CCB_BUILD_WIDE": "1"}), ("full", {"CCB_BUILD_WIDE": "1", "CCB_BUILD_STAGE": "70001"})])
def test_built_files_identical_to_reference(tmp_path, mode, env, monkeypatch):
    from jn_cuclark_b200 import CuClarkDB
    for k_, v in env.items():
        monkeypatch.setenv(k_, v)
    g = np.load(os.path.join(HERE, "golden", "ref_db", mode + ".npz"))
    light = mode == "light"
    k, ht = (27 if light else 31), int(g["htsize"])
    targets, order = _world_targets()
    db = CuClarkDB(1, k, 2, len(order))
    prefix = str(tmp_path / "db")
    info = db.build(targets, ht, 4, light, 4, 0, prefix, load_table=True)
    assert info.stored == g["lb"].size and info.max_bucket <= 255 and info.distinct >= info.stored
    for e in ("sz", "ky", "lb"):
        assert _sha(prefix + "." + e) == str(g["sha_" + e]), e
    assert db.db_info().entries == info.stored
    db.close()


def _edge_fasta(rng, n):
    """FASTA text with the quirks the scan rules care about."""
    seq = rng.choice(list(b"ACGT"), n).astype(np.uint8)
    seq[rng.integers(0, n, n // 50)] = ord("N")
    low = rng.integers(0, n, n // 20)
    seq[low] |= 0x20                                   # lower case
    s = seq.tobytes()
    out = [b">rec1 some text > with a marker\n"]
    w = 70
    for i in range(0, n // 2, w):
        out.append(s[i:i + w] + (b"\r\n" if i % 3 == 0 else b"\n"))
    out.append(b">rec2\n")
    out.append(s[n // 2:] + b"\n")
    out.append(b">empty\n>short\nACGT\n")
    return b"".join(out)


@pytest.mark.parametrize("light,k,min_count", [(False, 31, 0), (False, 21, 1), (True, 27, 0), (False, 32, 0), (False, 12, 0)])
def test_edge_cases_equal_oracle(oracle, tmp_path, light, k, min_count):
    from jn_cuclark_b200 import CuClarkDB
    rng = np.random.default_rng(5 + k)
    shared = _edge_fasta(rng, 3000)
    texts = [(_edge_fasta(rng, 6000), 0), (_edge_fasta(rng, 5000) + shared[1:], 1), (shared, 2),
             (_edge_fasta(rng, 4000), 0), (shared, 2)]                 # a label seen twice; a sequence under two labels
    ht = 1000003 if k < 32 else 4000037
    key_bytes = oracle.key_bytes(k, ht)
    paths = []
    for i, (t, lab) in enumerate(texts):
        p = str(tmp_path / f"t{i}.fa"); open(p, "wb").write(t); paths.append((p, f"L{lab}"))
    tp = str(tmp_path / "targets.txt")
    open(tp, "w").write("".join(f"{p}\t{l}\n" for p, l in paths))
    odb = oracle.OracleDb.build(tp, k, ht, light=light, gap=4, min_t=min_count)
    odb.write(str(tmp_path / "o"))
    db = CuClarkDB(1, k, 2, 3)
    info = db.build(texts, ht, key_bytes, light, 4, min_count, str(tmp_path / "g"), load_table=True)
    assert info.stored == odb.n
    for e in ("sz", "ky", "lb"):
        assert open(str(tmp_path / "g") + "." + e, "rb").read() == open(str(tmp_path / "o") + "." + e, "rb").read(), e
    db.close()


def test_cli_builds_missing_database_then_classifies(tmp_path):
    """B1: 'If the DB files are missing it builds them first' -- same files, same CSV as with the reference-built DB."""
    g = np.load(os.path.join(HERE, "golden", "ref_db", "light.npz"))
    labels = open(os.path.join(WORLD, "labels.txt")).read().split()
    tp = str(tmp_path / "targets.txt")
    with open(tp, "w") as f:
        for i, lab in enumerate(labels):
            f.write(f"{os.path.join(WORLD, 'genomes', f'g{i}.fa')}\t{lab}\n")
    dbdir = tmp_path / "db"; dbdir.mkdir()
    res = str(tmp_path / "res")
    cmd = [os.path.join(ROOT, "bin", "cuCLARK-l"), "-T", tp, "-D", str(dbdir) + "/", "-O", os.path.join(WORLD, "reads.fa"), "-R", res, "-n", "2"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "successfully stored in database" in r.stderr
    name = str(dbdir / str(g["name"]))
    for e in ("sz", "ky", "lb"):
        assert _sha(name + "." + e) == str(g["sha_" + e]), e
    want = open(os.path.join(HERE, "golden", "ref_csv", "light_fa.csv"), "rb").read()
    assert open(res + ".csv", "rb").read() == want
    # second run loads the files just written
    r2 = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r2.returncode == 0 and "successfully stored" not in r2.stderr
    assert open(res + ".csv", "rb").read() == want
