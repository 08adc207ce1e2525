"""Process boundary: bin/cuCLARK and bin/cuCLARK-l on reference-format
databases, CSV compared byte for byte with the oracle's CSV (and with the CSVs the reference's own
host engine wrote, tests/golden/ref_csv/); u16[5] dump compared bit-exactly."""
import glob
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
WORLD = os.path.join(HERE, "golden", "world")
REFCSV = os.path.join(HERE, "golden", "ref_csv")   # written by the reference's own host engine (make_golden.py csv)


def _write_ref_db(npz, folder):
    """Re-create the .sz/.ky/.lb files the unmodified reference wrote (tests/golden/make_golden.py db)."""
    g = np.load(npz)
    prefix = os.path.join(folder, str(g["name"]))
    sz = np.zeros(int(g["htsize"]), np.uint8)
    sz[g["nz_bucket"]] = g["nz_size"]
    sz.tofile(prefix + ".sz")
    g["ky"].tofile(prefix + ".ky")
    g["lb"].tofile(prefix + ".lb")
    return prefix


@pytest.fixture(scope="module", params=["light", "full"])
def setup(request, tmp_path_factory, oracle):
    mode = request.param
    d = str(tmp_path_factory.mktemp("cli_" + mode))
    labels = open(os.path.join(WORLD, "labels.txt")).read().split()
    tp = os.path.join(d, "targets.txt")
    with open(tp, "w") as f:
        for i, lab in enumerate(labels):
            f.write(f"{os.path.join(WORLD, 'genomes', f'g{i}.fa')}\t{lab}\n")
    os.makedirs(os.path.join(d, "db"))
    prefix = _write_ref_db(os.path.join(HERE, "golden", "ref_db", mode + ".npz"), os.path.join(d, "db"))
    light = mode == "light"
    k = 27 if light else 31
    ht = oracle.HTSIZE_LIGHT if light else oracle.HTSIZE_FULL
    odb = oracle.OracleDb.read(prefix, ht, k)
    names, seen = ["NA"], set()
    for lab in labels:
        if lab not in seen:
            seen.add(lab); names.append(lab)
    return {"mode": mode, "dir": d, "targets": tp, "k": k, "odb": odb, "names": names,
            "bin": os.path.join(ROOT, "bin", "cuCLARK-l" if light else "cuCLARK"),
            "row_stride": 2 * (23 if light else 15) + 2}


def _run(s, extra, res, env=None):
    cmd = [s["bin"], "-k", "31", "-T", s["targets"], "-D", os.path.join(s["dir"], "db"), "-R", res] + extra
    e = dict(os.environ)
    if env:
        e.update(env)
    r = subprocess.run(cmd, capture_output=True, text=True, env=e, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    return r


def _oracle_csv(s, oracle, data, path, paired=False, extended=False):
    ir = oracle.IndexedReads(data)
    off, cont = ir.pack(s["k"])
    out5, rows, _ = s["odb"].query_packed(s["k"], off, cont, row_stride=s["row_stride"] if extended else 0)
    ir.write_csv(path, s["k"], paired, s["names"], out5, rows if extended else None)
    return out5


CASES = [("fa", ["-O", "reads.fa"], {}), ("ml", ["-O", "reads_ml.fa"], {}), ("fq", ["-O", "reads.fq"], {}),
         ("fa_ext", ["-O", "reads.fa", "--extended"], {"extended": True}),
         ("fa_b4n4", ["-O", "reads.fa", "-n", "4", "-b", "4"], {}),
         ("fq_m", ["-O", "reads.fq", "-m", "0", "--verbose"], {}),
         # CCB_ANCHOR=1: the target texts on the device in front of the table -- same bytes out
         ("fa_anchor", ["-O", "reads.fa", "--verbose"], {"env": {"CCB_ANCHOR": "1"}}),
         ("fa_ext_anchor", ["-O", "reads.fa", "--extended", "-n", "3"], {"extended": True, "env": {"CCB_ANCHOR": "1"}})]


@pytest.mark.parametrize("name,args,kw", CASES, ids=[c[0] for c in CASES])
def test_single_end(setup, oracle, tmp_path, name, args, kw):
    s = setup
    args = [os.path.join(WORLD, a) if a.startswith("reads") else a for a in args]
    res = str(tmp_path / "out")
    dump = str(tmp_path / "final5.bin")
    r = _run(s, args, res, dict({"CCB_DUMP_FINAL5": dump}, **kw.get("env", {})))
    if kw.get("env", {}).get("CCB_ANCHOR") and "--verbose" in args:
        assert "Anchor index: " in r.stderr
    assert "Classifying: " in r.stderr and "reads/min" in r.stderr and "Results: " + res + ".csv" in r.stderr
    data = open(args[1], "rb").read()
    want5 = _oracle_csv(s, oracle, data, str(tmp_path / "want.csv"), extended=kw.get("extended", False))
    assert open(res + ".csv", "rb").read() == open(str(tmp_path / "want.csv"), "rb").read()
    got5 = np.fromfile(dump, np.uint16).reshape(-1, 5)
    assert np.array_equal(got5, want5)                   # includes second best, which the CSV does not print
    ref = os.path.join(REFCSV, f"{s['mode']}_{'fa_b4' if name == 'fa_b4n4' else name.replace('_anchor', '')}.csv")
    if os.path.exists(ref):
        assert open(res + ".csv", "rb").read() == open(ref, "rb").read(), "differs from the reference host engine's CSV"


@pytest.mark.parametrize("ext", [False, True])
def test_paired(setup, oracle, tmp_path, ext):
    s = setup
    p1, p2 = os.path.join(WORLD, "p1.fq"), os.path.join(WORLD, "p2.fq")
    res = str(tmp_path / "out")
    _run(s, ["-P", p1, p2] + (["--extended"] if ext else []), res)
    m = str(tmp_path / "m.fa")
    assert oracle.merge_paired(p1, p2, m) == 0
    _oracle_csv(s, oracle, open(m, "rb").read(), str(tmp_path / "want.csv"), paired=True, extended=ext)
    assert open(res + ".csv", "rb").read() == open(str(tmp_path / "want.csv"), "rb").read()
    assert not glob.glob(p1 + "_Concatenated*")          # no temp file is left (none is written)
    ref = os.path.join(REFCSV, f"{s['mode']}_{'pe_ext' if ext else 'pe'}.csv")
    if os.path.exists(ref):
        assert open(res + ".csv", "rb").read() == open(ref, "rb").read()


def test_list_of_files_mode(setup, oracle, tmp_path):
    """-O and -R name lists of files when -R exists and -O is not a reads file (CuCLARK_hh.hh:386-428)."""
    s = setup
    if s["mode"] != "light":
        pytest.skip("one variant is enough")
    objs, ress = tmp_path / "objs.txt", tmp_path / "ress.txt"
    objs.write_text(os.path.join(WORLD, "reads.fa") + "\n" + os.path.join(WORLD, "reads.fq") + "\n")
    ress.write_text(str(tmp_path / "a") + "\n" + str(tmp_path / "b") + "\n")
    _run(s, ["-O", str(objs)], str(ress))
    for out, src in (("a", "reads.fa"), ("b", "reads.fq")):
        _oracle_csv(s, oracle, open(os.path.join(WORLD, src), "rb").read(), str(tmp_path / "w.csv"))
        assert open(str(tmp_path / out) + ".csv", "rb").read() == open(str(tmp_path / "w.csv"), "rb").read()


def test_missing_db_is_built_like_reference(setup, tmp_path):
    """CuCLARK ctor (src/CuCLARK_hh.hh:306-311): no .sz/.ky/.lb for these targets -> build them, then classify.
    Byte-level checks of the built files live in test_build_gpu.py."""
    s = setup
    r = subprocess.run([s["bin"], "-T", s["targets"], "-D", str(tmp_path), "-O", os.path.join(WORLD, "reads.fa"), "-R", str(tmp_path / "o")],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-1500:]
    assert "Starting the creation of the database of targets specific" in r.stderr and "successfully stored in database." in r.stderr
    assert sorted(os.path.splitext(f)[1] for f in os.listdir(tmp_path) if f.startswith("db_central")) == [".ky", ".lb", ".sz"]


def test_missing_db_and_unbuildable_targets_fail_like_reference(setup, tmp_path):
    """A three-column targets file (label_c) is outside the rebuilt path: nothing to build from ->
    the reference's 'Failed to find the database.' exit (src/CuCLARK_hh.hh:635-639)."""
    s = setup
    tp = tmp_path / "targets3.txt"
    tp.write_text("".join(line.rstrip("\n") + "\tC0\n" for line in open(s["targets"])))
    r = subprocess.run([s["bin"], "-T", str(tp), "-D", str(tmp_path), "-O", os.path.join(WORLD, "reads.fa"), "-R", str(tmp_path / "o")],
                       capture_output=True, text=True)
    assert r.returncode == 255 and "Failed to find the database." in r.stderr


# ---- against CSVs the reference's own CUDA kernels wrote on a B200 (tests/golden/ref_gpu/w32_*.csv) ----
WORLD32 = os.path.join(HERE, "golden", "world32")
REFGPU = os.path.join(HERE, "golden", "ref_gpu")
W32 = {"fa": ["-O", "reads.fa"], "fq": ["-O", "reads.fq"], "ml": ["-O", "reads_ml.fa"], "fa_ext": ["-O", "reads.fa", "--extended"],
       "pe": ["-P", "p1.fq", "p2.fq"], "pe_ext": ["-P", "p1.fq", "p2.fq", "--extended"]}


@pytest.mark.parametrize("anchor", [False, True], ids=["table", "anchor"])
@pytest.mark.parametrize("mode", ["light", "full", "fullk21"])
def test_cli_equals_reference_gpu_binary(tmp_path, mode, anchor):
    """Same targets, same reads: bin/cuCLARK[-l] (database built by ccb_build_*) writes the CSV that the unmodified
    reference binary wrote with its own kernels on a B200 (make_golden.py gpu32)."""
    labels = open(os.path.join(WORLD32, "labels.txt")).read().split()
    tp = str(tmp_path / "targets.txt")
    with open(tp, "w") as f:
        for i, lab in enumerate(labels):
            f.write(f"{os.path.join(WORLD32, 'genomes', f'g{i}.fa')}\t{lab}\n")
    (tmp_path / "db").mkdir()
    exe = os.path.join(ROOT, "bin", "cuCLARK-l" if mode == "light" else "cuCLARK")
    kflag = ["-k", "21"] if mode == "fullk21" else []
    golden = sorted(glob.glob(os.path.join(REFGPU, f"w32_{mode}_*.csv")))
    assert golden
    for g in golden:
        case = os.path.basename(g)[len(f"w32_{mode}_"):-4]
        args = [a if a.startswith("-") else os.path.join(WORLD32, a) for a in W32[case]]
        res = str(tmp_path / case)
        env = dict(os.environ, CCB_ANCHOR="1") if anchor else None
        r = subprocess.run([exe, "-T", tp, "-D", str(tmp_path / "db") + "/", "-R", res, "-n", "3"] + kflag + args, capture_output=True, text=True, timeout=600, env=env)
        assert r.returncode == 0, r.stderr[-1500:]
        assert open(res + ".csv", "rb").read() == open(g, "rb").read(), case


@pytest.mark.parametrize("anchor", [False, True], ids=["table", "anchor"])
def test_two_devices_give_the_same_csv(setup, oracle, tmp_path, anchor):
    """-d 2 (row E1 at the process boundary): the table is replicated, batches are dealt round-robin over the
    devices, rows come out in input order -- byte-identical to the one-device CSV and to the oracle's."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    s = setup
    reads = os.path.join(WORLD, "reads.fq")
    env = {"CCB_ANCHOR": "1"} if anchor else None
    r = _run(s, ["-O", reads, "-d", "2", "-n", "4", "-b", "8", "--verbose"], str(tmp_path / "two"), env)
    assert "Using 2 CUDA devices as requested." in r.stderr
    _run(s, ["-O", reads, "-n", "4"], str(tmp_path / "one"), env)
    _oracle_csv(s, oracle, open(reads, "rb").read(), str(tmp_path / "want.csv"))
    two = open(str(tmp_path / "two.csv"), "rb").read()
    assert two == open(str(tmp_path / "one.csv"), "rb").read() == open(str(tmp_path / "want.csv"), "rb").read()


@pytest.mark.parametrize("case", ["fq", "pe_ext"])
def test_sharded_table_gives_the_same_csv(setup, oracle, tmp_path, case):
    """--sharded -d 2 (row E2 at the process boundary): ONE table cut into two shards (on two GPUs, or both on the one GPU
    of the box), k-mers and labels exchanged through peer memory -- byte-identical to the one-device CSV and the oracle's."""
    import torch
    s = setup
    env = {"CCB_SHARD_SAME_DEVICE": "1"} if torch.cuda.device_count() < 2 else {}
    if case == "fq":
        reads = os.path.join(WORLD, "reads.fq")
        r = _run(s, ["-O", reads, "-d", "2", "--sharded", "-n", "4", "--verbose"], str(tmp_path / "sh"), env)
        _oracle_csv(s, oracle, open(reads, "rb").read(), str(tmp_path / "want.csv"))
    else:
        p1, p2 = os.path.join(WORLD, "p1.fq"), os.path.join(WORLD, "p2.fq")
        r = _run(s, ["-P", p1, p2, "--extended", "-d", "2", "--sharded", "-n", "2", "--verbose"], str(tmp_path / "sh"), env)
        m = str(tmp_path / "m.fa")
        assert oracle.merge_paired(p1, p2, m) == 0
        _oracle_csv(s, oracle, open(m, "rb").read(), str(tmp_path / "want.csv"), paired=True, extended=True)
    assert "Database sharded over 2 devices" in r.stderr
    assert open(str(tmp_path / "sh.csv"), "rb").read() == open(str(tmp_path / "want.csv"), "rb").read()


def test_gzipped_in_process(setup, tmp_path):
    """--gzipped handed to the binary itself (classify_metagenome.sh strips it and gunzips first): inflated in memory."""
    import gzip
    s = setup
    for f in ("reads.fq", "p1.fq", "p2.fq"):
        with open(os.path.join(WORLD, f), "rb") as src, gzip.open(str(tmp_path / (f + ".gz")), "wb") as dst:
            dst.write(src.read())
    _run(s, ["-O", str(tmp_path / "reads.fq.gz"), "--gzipped"], str(tmp_path / "gz"))
    _run(s, ["-O", os.path.join(WORLD, "reads.fq")], str(tmp_path / "plain"))
    assert open(str(tmp_path / "gz.csv"), "rb").read() == open(str(tmp_path / "plain.csv"), "rb").read()
    _run(s, ["-P", str(tmp_path / "p1.fq.gz"), str(tmp_path / "p2.fq.gz"), "--gzipped"], str(tmp_path / "gzp"))
    _run(s, ["-P", os.path.join(WORLD, "p1.fq"), os.path.join(WORLD, "p2.fq")], str(tmp_path / "plainp"))
    assert open(str(tmp_path / "gzp.csv"), "rb").read() == open(str(tmp_path / "plainp.csv"), "rb").read()


def test_anchor_is_the_default_and_can_be_turned_off(setup, tmp_path):
    s = setup
    reads = os.path.join(WORLD, "reads.fa")
    r1 = _run(s, ["-O", reads, "--verbose"], str(tmp_path / "on"))
    r0 = _run(s, ["-O", reads, "--verbose"], str(tmp_path / "off"), {"CCB_ANCHOR": "0"})
    assert "Anchor index: " in r1.stderr and "Anchor index" not in r0.stderr
    assert open(str(tmp_path / "on.csv"), "rb").read() == open(str(tmp_path / "off.csv"), "rb").read()


@pytest.mark.parametrize("chunk", ["3000", "20000", "1000000"])
def test_streaming_pipeline_gives_the_same_csv(setup, tmp_path, chunk):
    """Large inputs go through the streaming host pipeline (hostpipe.cpp classify_stream: chunks indexed, packed and
    submitted by the worker that owns them); forced onto the small golden files here with small chunks.  FASTQ, multi-line
    FASTA, pairs (joined in memory = FASTA), extended rows; a file with a marker alone on its line falls back."""
    s = setup
    env = {"CCB_STREAM_MIN_BYTES": "1", "CCB_STREAM_CHUNK": chunk}
    off = {"CCB_NO_STREAM": "1"}
    for i, args in enumerate((["-O", os.path.join(WORLD, "reads.fq"), "-n", "3"], ["-O", os.path.join(WORLD, "reads_ml.fa"), "-n", "4", "--extended"],
                              ["-P", os.path.join(WORLD, "p1.fq"), os.path.join(WORLD, "p2.fq"), "-n", "2"])):
        r = _run(s, args + ["--verbose"], str(tmp_path / f"s{i}"), env)
        _run(s, args, str(tmp_path / f"t{i}"), off)
        assert open(str(tmp_path / f"s{i}.csv"), "rb").read() == open(str(tmp_path / f"t{i}.csv"), "rb").read(), args
    odd = tmp_path / "odd.fq"
    data = open(os.path.join(WORLD, "reads.fq"), "rb").read()
    cut = data.index(b"\n@", len(data) // 2) + 1
    odd.write_bytes(data[:cut] + b"@\nACGT\n+\nIIII\n" + data[cut:])
    _run(s, ["-O", str(odd), "-n", "3"], str(tmp_path / "so"), env)
    _run(s, ["-O", str(odd), "-n", "3"], str(tmp_path / "to"), off)
    assert open(str(tmp_path / "so.csv"), "rb").read() == open(str(tmp_path / "to.csv"), "rb").read()
    # FASTQ chunks find their first record by themselves (a line starting with '@' two lines above one starting with '+')
    # and the pieces are checked against each other at the end.  Files that try to fool that: every quality line starts
    # with '@'; the same with some sequences starting with '+' (then a quality line looks exactly like a header: the check
    # at the end must catch it and the file must go through the two-phase path); a last record cut short.
    lines = data.split(b"\n")
    assert lines[-1] == b"" and (len(lines) - 1) % 4 == 0
    recs = [lines[i:i + 4] for i in range(0, len(lines) - 1, 4)]
    at_quals = b"".join(b"\n".join([h, sq, pl, b"@" + q[1:]]) + b"\n" for h, sq, pl, q in recs)
    fooled = b"".join(b"\n".join([h, (b"+" + sq[1:]) if i % 3 == 1 else sq, pl, b"@" + q[1:]]) + b"\n" for i, (h, sq, pl, q) in enumerate(recs))
    cut_short = data[: len(data) - len(recs[-1][3]) - 3]
    for name, text in (("atq", at_quals), ("fooled", fooled), ("short", cut_short)):
        f = tmp_path / (name + ".fq")
        f.write_bytes(text)
        _run(s, ["-O", str(f), "-n", "3", "--extended"], str(tmp_path / ("s_" + name)), env)
        _run(s, ["-O", str(f), "-n", "3", "--extended"], str(tmp_path / ("t_" + name)), off)
        assert open(str(tmp_path / ("s_" + name + ".csv")), "rb").read() == open(str(tmp_path / ("t_" + name + ".csv")), "rb").read(), name
