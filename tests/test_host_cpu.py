"""Host logic (C++ in libcuclark_b200.so, include/ccb_host.h) against the CPU oracle: read indexing
(A1), 2-bit packing (A2), paired-end join (A10) and CSV text (A7).  No GPU needed."""
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
WORLD = os.path.join(HERE, "golden", "world")

EDGE_FASTA = [
    b">r1\nACGT\n",
    b">r1\nACGT",                                    # no trailing newline
    b">r1 desc here\nACGTACGTACGTACGTACGTACGTACGTACGTACGT\n>r2\tx\nACGTNACGTACGTACGTACGTACGTACGTACGTACGTACGTACGTA\n",
    b">empty\n>next\nACGTACGTACGTACGTACGTACGTACGTACGTAAA\n",
    b">multi\nACGTACGTAC\nGTACGTACGTACGTACG\nTACGTACGTACG\n>x\nacgtnnacgu\n",
    b">a_name_that_is_much_longer_than_thirty_nine_characters_for_sure and more\nACGTACGTACGTACGTACGTACGTACGTACGTACGTACGT\n",
    b">q\nACGT>ACGT\n>ACGT\nAC\n",                   # '>' inside a line; '>' line directly after a header
    b">\nACGT\nTTTT\n>z\nAC\n",                      # empty name swallows the next line
    b"> lead\nACGTACGTACGTACGTACGTACGTACGTACGTA\n\n\n>t\n\nACGT\n",
]
EDGE_FASTQ = [
    b"@r1\nACGT\n+\nIIII\n",
    b"@r1\nACGT\n+\nIIII",
    b"@r1 d\nACGTNNACGTACGTACGTACGTACGTACGTACGTACGTACGTAC\n+r1\nIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIIII\n@r2/1\nAC\n+\n@I\n",
    b"@r1\nACGT\n+\nIIII\n\n",
    b"@r1\nACGT\n+\nIIII\n\n\n",
    b"@\nACGT\n+\nIIII\n@r2\nACGT\n+\nIIII\n",     # empty name
    b"@r1\nACGT\n+\nIIII\n@r2\nACGT\n",              # truncated record
]


def _same_index(oracle, host, data, threads=1):
    a = oracle.IndexedReads(data)
    b = host.HostReads(data, threads)
    assert a.n == b.n
    for f in ("name_s", "name_e", "seq_s", "seq_e", "len"):
        assert np.array_equal(a.arr(f), b.arr(f)), f
    return a, b


@pytest.mark.parametrize("data", EDGE_FASTA + EDGE_FASTQ)
def test_index_pack_csv_edge_cases(oracle, data):
    from jn_cuclark_b200 import host
    a, b = _same_index(oracle, host, data)
    for k in (4, 12, 31):
        oa, ca = a.pack(k)
        ob, cb = b.pack(k)
        assert np.array_equal(oa, ob) and np.array_equal(ca, cb)
    rng = np.random.default_rng(1)
    out5 = rng.integers(0, 40, size=(a.n, 5)).astype(np.uint16)
    out5[:, 1] = rng.integers(0, 4, size=a.n)
    out5[0, :] = 0                                   # gamma 0 / confidence 0
    names = ["NA", "T0", "T1", "T2"]
    for paired in (False, True):
        p = "/tmp/ccb_csv_%d.csv" % os.getpid()
        a.write_csv(p, 31, paired, names, out5)
        assert open(p, "rb").read() == b.csv(31, paired, names, out5)
        os.remove(p)


@pytest.mark.parametrize("fname", ["reads.fa", "reads_ml.fa", "reads.fq"])
def test_world_files(oracle, fname):
    from jn_cuclark_b200 import host
    data = open(os.path.join(WORLD, fname), "rb").read()
    a, b = _same_index(oracle, host, data)
    for k in (27, 31):
        oa, ca = a.pack(k)
        ob, cb = b.pack(k)
        assert np.array_equal(oa, ob) and np.array_equal(ca, cb)
    # sub-range packing = slices of the whole
    ob2, cb2 = b.pack(31, 100, 50)
    ob, cb = b.pack(31)
    assert np.array_equal(ob2, ob[100:151] - ob[100]) and np.array_equal(cb2, cb[ob[100]:ob[150]])


def _big(kind, n, rng, odd=False):
    parts = []
    for i in range(n):
        L = int(rng.integers(20, 300))
        s = bytes(rng.choice(np.frombuffer(b"ACGTN", np.uint8), size=L, p=[.245, .245, .245, .245, .02]))
        if kind == "fa":
            if i % 7 == 0:
                s = b"\n".join(s[j:j + 60] for j in range(0, L, 60))
            parts.append(b">r%d some text\n%s\n" % (i, s))
        else:
            q = b"@" * L if i % 5 == 0 else b"I" * L       # quality lines that start with '@'
            parts.append(b"@r%d/1\n%s\n+\n%s\n" % (i, s, q))
    if odd:
        parts.insert(n // 2, b">\nACGT\n" if kind == "fa" else b"@\nACGT\n+\nIIII\n")
    return b"".join(parts)


@pytest.mark.parametrize("kind", ["fa", "fq"])
@pytest.mark.parametrize("odd", [False, True])
def test_parallel_index_equals_sequential(oracle, kind, odd):
    from jn_cuclark_b200 import host
    data = _big(kind, 12000, np.random.default_rng(5), odd)
    assert len(data) > (1 << 20)
    for threads in (3, 8):
        _same_index(oracle, host, data, threads)


def test_join_pairs(oracle, tmp_path):
    from jn_cuclark_b200 import host
    m = str(tmp_path / "m.fa")
    f1, f2 = os.path.join(WORLD, "p1.fq"), os.path.join(WORLD, "p2.fq")
    assert oracle.merge_paired(f1, f2, m) == 0
    assert open(m, "rb").read() == host.join_pairs(f1, f2)
    with pytest.raises(RuntimeError, match="FASTQ"):
        host.join_pairs(os.path.join(WORLD, "reads.fa"), os.path.join(WORLD, "reads.fa"))
    bad = tmp_path / "bad.fq"
    bad.write_bytes(b"@other/2\nACGT\n+\nIIII\n")
    with pytest.raises(RuntimeError, match="does not match"):
        host.join_pairs(f1, str(bad))


def test_extended_csv(oracle):
    from jn_cuclark_b200 import host
    data = open(os.path.join(WORLD, "reads.fa"), "rb").read()
    a, b = _same_index(oracle, host, data)
    rng = np.random.default_rng(2)
    names = ["NA"] + ["T%d" % i for i in range(6)]
    out5 = rng.integers(0, 100, size=(a.n, 5)).astype(np.uint16)
    out5[:, 1] = rng.integers(0, 7, size=a.n)
    rows = np.zeros((a.n, 32), np.uint16)
    for i in range(a.n):
        t = np.sort(rng.choice(6, size=int(rng.integers(0, 5)), replace=False))
        rows[i, 0] = t.size
        rows[i, 1:1 + 2 * t.size:2] = t
        rows[i, 2:2 + 2 * t.size:2] = rng.integers(1, 99, size=t.size)
    p = "/tmp/ccb_csvx_%d.csv" % os.getpid()
    a.write_csv(p, 31, False, names, out5, rows)
    assert open(p, "rb").read() == b.csv(31, False, names, out5, rows)
    os.remove(p)


def test_long_part_is_split_not_wrapped():
    """Parts longer than 65528 bases are split with k-1 overlap (the reference's u16 length wraps)."""
    from jn_cuclark_b200 import host
    rng = np.random.default_rng(3)
    L, k = 150000, 31
    s = bytes(rng.choice(np.frombuffer(b"ACGT", np.uint8), size=L))
    b = host.HostReads(b">long\n" + s + b"\n")
    off, cont = b.pack(k)
    lens, p, total_kmers = [], 0, 0
    while p < off[1]:
        n = int(cont[p]); lens.append(n); total_kmers += n - k + 1; p += 1 + (n + 7) // 8
    assert lens == [65528, 65528, L - 2 * (65528 - (k - 1))]
    assert total_kmers == L - k + 1


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_pack_random_texts_equal_oracle(oracle, seed):
    """Random FASTA with ragged line widths, N runs, lower case, U, stray characters and reads of every length around
    the container size: the packer's eight-bases-per-step path and its byte path write the oracle's containers."""
    from jn_cuclark_b200 import host
    rng = np.random.default_rng(seed)
    alphabet = np.frombuffer(b"ACGTACGTACGTACGTACGTACGTacgtuUN.-", dtype=np.uint8)
    recs = []
    for i in range(400):
        n = int(rng.integers(0, 260)) if i % 7 else int(rng.integers(0, 20))
        seq = alphabet[rng.integers(0, alphabet.size if i % 3 == 0 else 24, n)].tobytes()
        width = int(rng.integers(1, 90))
        lines = b"\n".join(seq[j:j + width] for j in range(0, len(seq), width)) if i % 2 else seq
        recs.append(b">r%d x\n" % i + lines + b"\n")
    data = b"".join(recs)
    a, b = _same_index(oracle, host, data)
    for k in (2, 8, 9, 21, 31, 32):
        oa, ca = a.pack(k)
        ob, cb = b.pack(k)
        assert np.array_equal(oa, ob) and np.array_equal(ca[:int(oa[-1])], cb[:int(ob[-1])]), k


def test_numa_binding_is_scoped_and_optional(monkeypatch):
    """Pinned batch buffers are allocated with the calling thread bound to the memory node of the slot's GPU
    (csrc/hostnuma.hpp).  An unknown device changes nothing; a known node's CPU list parses to the count the kernel reports;
    whatever happens while bound, the thread's affinity is what it was afterwards."""
    import os
    from jn_cuclark_b200 import host
    before = len(os.sched_getaffinity(0))
    node, during, after = host.numa_probe("ffff:ff:1f.7")             # no such device
    assert node == -1 and during == before and after == before
    p = "/sys/devices/system/node/node0/cpulist"
    if os.path.exists(p):
        want = 0
        for part in open(p).read().strip().split(","):
            a, _, b = part.partition("-")
            want += int(b or a) - int(a) + 1
        assert host.numa_node_cpus(0) == want
    assert host.numa_node_cpus(4095) == -1
    # a real PCI device with a node, if the machine has one: bound to a subset, restored afterwards
    root = "/sys/bus/pci/devices"
    for dev in sorted(os.listdir(root)) if os.path.isdir(root) else []:
        try:
            n = int(open(os.path.join(root, dev, "numa_node")).read())
        except (OSError, ValueError):
            continue
        if n < 0:
            continue
        node, during, after = host.numa_probe(dev.upper())
        assert after == before and 0 < during <= before
        assert node in (n, -1)                                         # -1: the node's CPUs are outside this process's cpuset
        monkeypatch.setenv("CCB_NUMA", "0")
        assert host.numa_probe(dev)[0] == -1
        break


def _fastq(n, rng, at_quals=False, plus_seqs=0):
    recs = []
    for i in range(n):
        L = int(rng.integers(1, 200))
        seq = bytes(rng.choice(np.frombuffer(b"ACGTN", np.uint8), size=L))
        if plus_seqs and i % plus_seqs == 1:
            seq = b"+" + seq[1:]
        q = bytes(rng.integers(33, 74, L, dtype=np.uint8))
        if at_quals:
            q = b"@" + q[1:]
        recs.append(b"@r%d some text/1\n" % i + seq + b"\n+\n" + q + b"\n")
    return b"".join(recs)


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_fastq_index_in_chunks_equals_sequential_index(seed):
    """The streaming pipeline cuts a FASTQ text at arbitrary bytes; every chunk finds its first record by itself (a line
    starting with '@' two lines above a line starting with '+') and takes every fourth line from there; the pieces are
    checked against each other.  Whatever it accepts must be exactly the sequential index (reference
    src/CuCLARK_hh.hh:1496-1524) -- for every chunk size, also when every quality line starts with '@', with CRLF-free
    ragged reads, and with a tail cut short; what tries to fool it must be refused (None), never mis-indexed."""
    from jn_cuclark_b200 import host
    rng = np.random.default_rng(seed)

    def same(data, chunk):
        want = host.HostReads(data)
        got = host.index_fastq_chunked(data, chunk)
        if got is None:
            return False
        for f in host.FIELDS:
            assert np.array_equal(got[f], want.arr(f)), (f, chunk)
        return True

    plain = _fastq(300, rng)
    atq = _fastq(300, rng, at_quals=True)
    for chunk in (64, 257, 1000, 4096, 10 ** 6):
        assert same(plain, chunk), chunk
        assert same(atq, chunk), chunk                      # quality lines that start with '@' are not headers: line + 2 tells
        for cut in (1, 7, 40):
            same(plain[:-cut], chunk)                       # a truncated tail: accepted only if equal (asserted inside)
    # sequences that start with '+' behind quality lines that start with '@': a quality line then looks exactly like a header.
    # Some chunk sizes start on the wrong line -- the check at the end must refuse those; none may return a wrong index.
    fooled = _fastq(300, rng, at_quals=True, plus_seqs=3)
    refused = 0
    for chunk in range(50, 1200, 37):
        refused += 0 if same(fooled, chunk) else 1
    assert refused > 0
    # a marker alone on its line is for the sequential indexer
    odd = plain[:len(plain) // 2]
    odd = odd[:odd.rindex(b"\n@") + 1] + b"@\nACGT\n+\nIIII\n" + plain[len(plain) // 2:][plain[len(plain) // 2:].index(b"\n@") + 1:]
    assert host.index_fastq_chunked(odd, 500) is None
    # missing line: records shifted from there on; either refused or (chunks that never see the damage) equal
    lines = plain.split(b"\n")
    broken = b"\n".join(lines[:401] + lines[402:])
    for chunk in (100, 333, 5000, 10 ** 6):
        same(broken, chunk)


def test_parallel_pair_join_equals_the_line_by_line_join(oracle, tmp_path, monkeypatch):
    """join_pairs on large files splits the records over threads (hostpipe.cpp join_pairs_texts); small files and anything
    unusual go line by line as the reference reads them (src/file.cc:205-268).  Forced onto small files here
    (CCB_JOIN_MIN_BYTES=0): the same bytes as the line-by-line join (CCB_JOIN_SEQUENTIAL=1) and as the oracle's merge, the same
    error messages, for well-formed files, files without a final newline, CRLF, ids with separators, empty sequences,
    unequal record counts, a damaged record, a line that is no header where one belongs."""
    from jn_cuclark_b200 import host
    rng = np.random.default_rng(5)

    def fq(n, tag, lens=None, crlf=False, ids=None):
        out = []
        for i in range(n):
            L = int(lens[i]) if lens is not None else int(rng.integers(0, 180))
            s = bytes(rng.choice(np.frombuffer(b"ACGTN", np.uint8), size=L))
            name = ids[i] if ids else b"read%d" % i
            eol = b"\r\n" if crlf else b"\n"
            out.append(b"@" + name + tag + eol + s + eol + b"+" + eol + b"I" * L + eol)
        return b"".join(out)

    def both(a, b):
        fa, fb = tmp_path / "a.fq", tmp_path / "b.fq"
        fa.write_bytes(a); fb.write_bytes(b)
        res = []
        for seq in (True, False):
            if seq:
                monkeypatch.setenv("CCB_JOIN_SEQUENTIAL", "1")
            else:
                monkeypatch.delenv("CCB_JOIN_SEQUENTIAL", raising=False)
            try:
                res.append(("ok", host.join_pairs(str(fa), str(fb))))
            except RuntimeError as e:
                res.append(("err", str(e)))
        assert res[0] == res[1], (res[0][0], res[1][0], res[0][1][:200], res[1][1][:200])
        return res[0], str(fa), str(fb)

    monkeypatch.setenv("CCB_JOIN_MIN_BYTES", "0")
    ids = [b"r%d" % i if i % 5 else b"weird id %d/x" % i for i in range(1000)]
    a, b = fq(1000, b"/1", ids=ids), fq(1000, b"/2", ids=ids)
    (kind, text), fa, fb = both(a, b)
    assert kind == "ok" and text.count(b"\n") == 2000
    m = str(tmp_path / "m.fa")
    assert oracle.merge_paired(fa, fb, m) == 0 and open(m, "rb").read() == text
    assert both(a[:-1], b[:-1])[0][0] == "ok"                         # no newline at the end of either file
    assert both(a[:-1], b)[0][0] == "ok"
    assert both(fq(300, b"/1", crlf=True), fq(300, b"/2", crlf=True))[0][0] == "ok"
    assert both(fq(50, b" 1"), fq(50, b" 2"))[0][0] == "ok"
    assert both(fq(3, b"/1"), fq(3, b"/2"))[0][0] == "ok"             # fewer records than threads
    # unequal counts, a record cut short, a missing line, an id that differs, a non-header where a header belongs
    assert both(a, fq(900, b"/2", ids=ids))[0][0] in ("ok", "err")
    assert both(a, b[: len(b) // 2])[0][0] in ("ok", "err")
    la = a.split(b"\n")
    assert both(b"\n".join(la[:41] + la[42:]), b)[0][0] in ("ok", "err")
    bad_id = b.replace(b"@r501/2", b"@q501/2")
    assert bad_id != b
    (kind, msg), _, _ = both(a, bad_id)
    assert kind == "err" and "does not match" in msg
    shifted = b"\n".join(la[:400] + [b"x" + la[400][1:]] + la[401:])    # line 400 is a header: make it start with 'x'
    assert both(shifted, b)[0][0] in ("ok", "err")
    (kind, msg), _, _ = both(b">r0\nACGT\n", b">r0\nACGT\n")
    assert kind == "err" and "FASTQ" in msg
    # the golden pair files through the parallel path
    f1, f2 = os.path.join(WORLD, "p1.fq"), os.path.join(WORLD, "p2.fq")
    assert oracle.merge_paired(f1, f2, m) == 0
    assert open(m, "rb").read() == host.join_pairs(f1, f2)


def test_host_suite_on_the_scalar_paths():
    """The AVX2 routines (newline events, marker scan, 2-bit packer: csrc/hostscan_avx2.cpp) are chosen once per process;
    CCB_NO_AVX2=1 takes the scalar code instead.  Every test of this file once more in a process that runs without them."""
    import subprocess
    import sys
    if os.environ.get("CCB_NO_AVX2"):
        pytest.skip("already the scalar run")
    env = dict(os.environ, CCB_NO_AVX2="1")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-q", "-x", "-p", "no:cacheprovider"],
                       env=env, capture_output=True, text=True, timeout=900, cwd=os.path.dirname(HERE))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-1000:]
