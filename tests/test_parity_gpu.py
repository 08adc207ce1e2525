"""GPU parity: CUDA path (through the C ABI) vs the CPU oracle, bit-exact u16[5] and rows."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True, params=["block", "warp"])
def kernel_flavour(request):
    """Every parity case runs twice: the default path (thread-per-read block kernel for short single-part reads,
    warp-per-read kernel for the rest) and the warp-per-read kernel for every read."""
    from jn_cuclark_b200 import lib
    lib.default_flags = lib.CCB_WARP_KERNEL if request.param == "warp" else 0
    yield
    lib.default_flags = 0

HT_SMALL = 1000003


def _reads(world, name, oracle):
    return oracle.IndexedReads(open(os.path.join(world["dir"], name), "rb").read())


@pytest.fixture(scope="module")
def full_db(small_world, oracle):
    return oracle.OracleDb.build(small_world["targets"], 31, HT_SMALL)


@pytest.fixture(scope="module")
def gpu_db(full_db):
    from jn_cuclark_b200 import CuClarkDB
    db = CuClarkDB(1, 31, 2, len(full_db.labels))
    c, l = full_db.pairs()
    db.load_pairs(c, l)
    yield db
    db.close()


@pytest.mark.parametrize("fname", ["reads.fa", "reads_ml.fa", "reads.fq"])
def test_u16x5_bit_exact(small_world, oracle, full_db, gpu_db, fname):
    from jn_cuclark_b200 import CCB_STATS
    ir = _reads(small_world, fname, oracle)
    off, cont = ir.pack(31)
    want, _, nl = full_db.query_packed(31, off, cont)
    before = gpu_db.stats().lookups
    got, _ = gpu_db.classify_packed(off, cont, flags=CCB_STATS)
    assert np.array_equal(got, want)
    assert gpu_db.stats().lookups - before == nl
    assert (want[:, 1] > 0).sum() > 1000          # the test is not vacuous


def test_extended_rows(small_world, oracle, full_db, gpu_db):
    ir = _reads(small_world, "reads.fa", oracle)
    off, cont = ir.pack(31)
    want, wrows, _ = full_db.query_packed(31, off, cont, row_stride=gpu_db.row_stride)
    got, rows = gpu_db.classify_packed(off, cont, extended=True)
    assert np.array_equal(got, want)
    n = wrows[:, 0].astype(int)
    for i in range(len(n)):
        assert rows[i, 0] == n[i]
        assert np.array_equal(rows[i, 1:1 + 2 * n[i]], wrows[i, 1:1 + 2 * n[i]])


def test_dense_path_equals_fast_path(small_world, oracle, full_db, gpu_db):
    from jn_cuclark_b200 import CCB_FORCE_DENSE
    ir = _reads(small_world, "reads.fa", oracle)
    off, cont = ir.pack(31)
    want, _, _ = full_db.query_packed(31, off, cont)
    got, _ = gpu_db.classify_packed(off, cont, flags=CCB_FORCE_DENSE)
    assert np.array_equal(got, want)


def test_paired(small_world, oracle, full_db, gpu_db, tmp_path):
    m = str(tmp_path / "merged.fa")
    assert oracle.merge_paired(os.path.join(small_world["dir"], "p1.fq"), os.path.join(small_world["dir"], "p2.fq"), m) == 0
    ir = oracle.IndexedReads(open(m, "rb").read())
    off, cont = ir.pack(31)
    want, _, _ = full_db.query_packed(31, off, cont)
    got, _ = gpu_db.classify_packed(off, cont)
    assert np.array_equal(got, want)


def test_ref_files_loader(small_world, oracle, full_db, tmp_path):
    """CuClarkDB::read equivalent: load .sz/.ky/.lb written in the reference format, incl. -s sampling."""
    from jn_cuclark_b200 import CuClarkDB
    prefix = str(tmp_path / "db")
    full_db.write(prefix)
    ir = _reads(small_world, "reads.fa", oracle)
    off, cont = ir.pack(31)
    for s in (1, 3):
        odb = oracle.OracleDb.read(prefix, HT_SMALL, 31, s)
        want, _, _ = odb.query_packed(31, off, cont)
        db = CuClarkDB(1, 31, 1, len(full_db.labels))
        n = db.read(prefix, HT_SMALL, oracle.lib().ccbo_key_bytes(31, HT_SMALL), s)
        assert n == odb.n
        got, _ = db.classify_packed(off, cont)
        db.close()
        assert np.array_equal(got, want)


@pytest.mark.parametrize("k", [12, 21, 27, 32])
def test_other_k(small_world, oracle, k):
    from jn_cuclark_b200 import CuClarkDB
    odb = oracle.OracleDb.build(small_world["targets"], k, HT_SMALL)
    ir = _reads(small_world, "reads.fa", oracle)
    off, cont = ir.pack(k)
    want, _, _ = odb.query_packed(k, off, cont)
    db = CuClarkDB(1, k, 1, len(odb.labels))
    c, l = odb.pairs()
    db.load_pairs(c, l)
    got, _ = db.classify_packed(off, cont)
    db.close()
    assert np.array_equal(got, want)


def test_long_reads_and_many_targets(oracle, tmp_path):
    """10 kbp reads with 5 % errors (tile loop) and a read touching > 32 targets (dense path)."""
    import synth
    from jn_cuclark_b200 import CuClarkDB
    tp, genomes, _ = synth.make_genomes(str(tmp_path / "g"), 48, 12000, seed=5, shared=False, quirks=False, same_label_pair=False)
    odb = oracle.OracleDb.build(tp, 31, HT_SMALL)
    rng = np.random.default_rng(9)
    reads = []
    for i in range(40):
        g = genomes[i % 48]
        s = g[500:10500].copy()
        m = rng.random(s.size) < 0.05
        s[m] = synth.BASES[rng.integers(0, 4, size=int(m.sum()))]
        if i % 5 == 0:
            s[rng.integers(0, s.size, size=7)] = ord("N")
        reads.append((f"long{i}", s))
    reads.append(("mosaic48", np.concatenate([g[100:180] for g in genomes])))     # 48 targets
    reads.append(("mosaic33", np.concatenate([g[300:360] for g in genomes[:33]])))
    synth.write_reads_fasta(str(tmp_path / "r.fa"), reads)
    ir = oracle.IndexedReads(open(str(tmp_path / "r.fa"), "rb").read())
    off, cont = ir.pack(31)
    want, _, _ = odb.query_packed(31, off, cont)
    db = CuClarkDB(1, 31, 1, len(odb.labels))
    c, l = odb.pairs()
    db.load_pairs(c, l)
    got, _ = db.classify_packed(off, cont)
    st = db.stats()
    db.close()
    assert np.array_equal(got, want)
    assert st.dense_reads == 2


def test_discriminative_rule_on_device(oracle):
    """k-mers inserted under two labels become misses (HashTableStorage_hh.hh:244)."""
    from jn_cuclark_b200 import CuClarkDB
    rng = np.random.default_rng(3)
    k = 31
    x = rng.integers(0, 1 << 62, size=20000, dtype=np.uint64)
    canon = np.array([oracle.canonical(int(v), k) for v in x], dtype=np.uint64)
    labels = rng.integers(0, 7, size=canon.size).astype(np.uint16)
    canon = np.concatenate([canon, canon[:5000], canon[5000:8000]])
    labels = np.concatenate([labels, (labels[:5000] + 1) % 7, labels[5000:8000]]).astype(np.uint16)
    odb = oracle.OracleDb.from_pairs(canon, labels, k, HT_SMALL)
    assert odb.n == 15000
    db = CuClarkDB(1, k, 1, 7)
    db.load_pairs(canon, labels)
    info = db.db_info()
    assert info.tombstones >= 1 and info.entries == 20000
    # one read per k-mer: [len=31][4 containers]
    n = 20000
    off = (np.arange(n + 1, dtype=np.uint32) * 5)
    cont = np.zeros(n * 5, np.uint16)
    cont[0::5] = 31
    v = x[:n] << np.uint64(2)
    for j in range(4):
        cont[1 + j::5] = ((v >> np.uint64(48 - 16 * j)) & np.uint64(0xFFFF)).astype(np.uint16)
    want, _, _ = odb.query_packed(k, off, cont)
    got, _ = db.classify_packed(off, cont)
    db.close()
    assert np.array_equal(got, want)
    assert (want[:5000, 0] == 0).all() and (want[5000:, 0] == 1).all()


def test_synthetic_workload_matches_oracle(oracle):
    """bench.py's device-generated DB + packed reads (include/ccb_synth.h) against the CPU oracle."""
    import ctypes as C
    import torch
    from jn_cuclark_b200 import CuClarkDB, CCB_STATS
    G, LG, NR, L, k = 5, 60000, 6000, 150, 31
    db = CuClarkDB(1, k, 1, G)
    n = db.synth_db(0xC1A2C000, G, LG, G)
    odb = oracle.OracleDb.synth(0xC1A2C000, G, LG, G, k, HT_SMALL)
    assert db.db_info().entries == odb.n and n == G * (LG - k + 1)
    d_off, d_cont, nc = db.synth_reads_packed(0x5EED0001, 0xC1A2C000, 0, NR, L, G, LG)
    off = torch.empty(NR + 1, dtype=torch.int32, device="cuda")
    cont = torch.empty(nc, dtype=torch.int16, device="cuda")
    C.cdll.LoadLibrary("libcudart.so").cudaMemcpy(C.c_void_p(off.data_ptr()), C.c_void_p(d_off), C.c_size_t((NR + 1) * 4), 3)
    C.cdll.LoadLibrary("libcudart.so").cudaMemcpy(C.c_void_p(cont.data_ptr()), C.c_void_p(d_cont), C.c_size_t(nc * 2), 3)
    off_h = off.cpu().numpy().view(np.uint32)
    cont_h = cont.cpu().numpy().view(np.uint16)
    # the packed batch equals packing the text of the same reads with the reference's rules
    txt = db.synth_reads_text(0x5EED0001, 0xC1A2C000, 0, NR, L, G, LG)
    fa = b"".join(b">r%d\n" % i + txt[i].tobytes() for i in range(NR))
    ir = oracle.IndexedReads(fa)
    off_o, cont_o = ir.pack(k)
    assert np.array_equal(off_h, off_o) and np.array_equal(cont_h, cont_o)
    want, _, nl = odb.query_packed(k, off_o, cont_o)
    out = torch.zeros(NR * 5, dtype=torch.int16, device="cuda")
    stats = torch.zeros(2, dtype=torch.int64, device="cuda")
    db.query_device(d_off, d_cont, NR, out.data_ptr(), 0, stats.data_ptr(), CCB_STATS, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    got = out.cpu().numpy().view(np.uint16).reshape(-1, 5)
    assert np.array_equal(got, want)
    assert int(stats[0]) == nl
    frac_hit = (want[:, 1] > 0).mean()
    assert 0.85 < frac_hit < 0.95                      # 90 % of reads come from a genome
    db.device_free(d_off); db.device_free(d_cont)
    db.close()


@pytest.mark.parametrize("extended", [False, True])
def test_malloc_ref_contiguous_results(small_world, oracle, full_db, extended):
    """ccb_batch_alloc_ref = CuClarkDB::malloc's own result layout (src/CuClarkDB.cu:412-416): the batches of a file
    deliver into ONE final / full result array at row indexBatches[b], in any submission order."""
    from jn_cuclark_b200 import CuClarkDB
    ir = _reads(small_world, "reads.fa", oracle)
    off, cont = ir.pack(31)
    n = off.size - 1
    want, wrows, _ = full_db.query_packed(31, off, cont, row_stride=32)
    nb = 3
    db = CuClarkDB(1, 31, nb, len(full_db.labels))
    c, l = full_db.pairs()
    db.load_pairs(c, l)
    index = [0, n // 5, n // 2]                       # uneven batches
    ends = index[1:] + [n]
    max_reads = max(e - s for s, e in zip(index, ends))
    max_cont = max(int(off[e]) - int(off[s]) for s, e in zip(index, ends))
    fin, rows = db.malloc_ref(n, max_reads, max_cont, index, extended)
    for b in (2, 0, 1):
        s, e = index[b], ends[b]
        o, cc, _, _ = db.buffers(b)
        o[: e - s + 1] = off[s:e + 1] - off[s]
        cc[: int(off[e]) - int(off[s])] = cont[int(off[s]):int(off[e])]
        db.readyBatch(b, e - s, int(off[e]) - int(off[s]))
        db.queryBatch(b, extended)
    for b in range(nb):
        db.waitForBatch(b)
    assert np.array_equal(fin, want)
    if extended:
        cnt = wrows[:, 0].astype(int)
        for i in range(n):
            assert rows[i, 0] == cnt[i] and np.array_equal(rows[i, 1:1 + 2 * cnt[i]], wrows[i, 1:1 + 2 * cnt[i]])
    db.freeBatchMemory()
    db.close()


@pytest.mark.parametrize("which", ["c3_pairs_light", "c4_long_indels"])
def test_synthetic_c3_c4_workloads_match_oracle(oracle, tmp_path, which):
    """The other bench workloads (include/ccb_synth.h: pairs joined by N on a light database = C3, long reads with 5 %
    errors incl. indels = C4): the device generator's packed batch equals the oracle's text of the same reads packed with
    the reference's rules, the database equals the oracle's scan of the same genomes, results are bit-exact, with and
    without the anchor index."""
    import ctypes as C
    import torch
    from jn_cuclark_b200 import CuClarkDB, Workload, CCB_STATS, CCB_NO_ANCHOR
    if which == "c3_pairs_light":
        G, LG, NR, k, gap = 6, 40000, 3000, 27, 4
        w = Workload(1, 150, 400, 10, 0, 0); ow = oracle.Workload(1, 150, 400, 10, 0, 0)
        db = CuClarkDB(1, k, 1, G, max_hits=23)
        n = db.synth_db_light(0xC1A2C000, G, LG, G, gap)
        odb = oracle.OracleDb.synth(0xC1A2C000, G, LG, G, k, HT_SMALL, light_gap=gap)
        assert n == G * ((LG - k) // (gap * k) + 1)
    else:
        G, LG, NR, k = 5, 60000, 300, 31
        w = Workload(2, 10000, 0, 30, 10, 10); ow = oracle.Workload(2, 10000, 0, 30, 10, 10)
        db = CuClarkDB(1, k, 1, G)
        db.synth_db(0xC1A2C000, G, LG, G)
        odb = oracle.OracleDb.synth(0xC1A2C000, G, LG, G, k, HT_SMALL)
    assert db.db_info().entries == odb.n
    d_off, d_cont, nc = db.synth2_reads_packed(w, 0x5EED0003, 0xC1A2C000, 7, NR, G, LG)
    off = torch.empty(NR + 1, dtype=torch.int32, device="cuda"); cont = torch.empty(nc, dtype=torch.int16, device="cuda")
    rt = C.cdll.LoadLibrary("libcudart.so")
    rt.cudaMemcpy(C.c_void_p(off.data_ptr()), C.c_void_p(d_off), C.c_size_t((NR + 1) * 4), 3)
    rt.cudaMemcpy(C.c_void_p(cont.data_ptr()), C.c_void_p(d_cont), C.c_size_t(nc * 2), 3)
    fa = str(tmp_path / "r.fa")
    oracle.synth2_reads_fasta(fa, ow, 0x5EED0003, 0xC1A2C000, 7, NR, G, LG)
    ir = oracle.IndexedReads(open(fa, "rb").read())
    off_o, cont_o = ir.pack(k)
    assert np.array_equal(off.cpu().numpy().view(np.uint32), off_o) and np.array_equal(cont.cpu().numpy().view(np.uint16), cont_o)
    want, _, nl = odb.query_packed(k, off_o, cont_o)
    assert (want[:, 1] > 0).mean() > 0.5
    out = torch.zeros(NR * 5, dtype=torch.int16, device="cuda"); stats = torch.zeros(2, dtype=torch.int64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    db.query_device(d_off, d_cont, NR, out.data_ptr(), 0, stats.data_ptr(), CCB_STATS, st)
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy().view(np.uint16).reshape(-1, 5), want) and int(stats[0]) == nl
    db.anchor_synth(0xC1A2C000, G, LG, G)
    out.zero_()
    db.query_device(d_off, d_cont, NR, out.data_ptr(), 0, 0, 0, st)
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy().view(np.uint16).reshape(-1, 5), want)
    db.device_free(d_off); db.device_free(d_cont); db.close()
