"""Anchor index (sequences + sampled k-mer positions in front of the table): results are identical with and
without it, and identical to the oracle; it only changes how many table sectors are read."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True, params=["split", "fused"])
def anchor_flavour(request, monkeypatch):
    """Every case runs twice: the anchor pass as a kernel of its own in front of the probe kernel (the default:
    k_anchor_scan + k_query_block<PRE>), and both passes fused in one kernel (CCB_FUSED_ANCHOR=1)."""
    if request.param == "fused":
        monkeypatch.setenv("CCB_FUSED_ANCHOR", "1")
    else:
        monkeypatch.delenv("CCB_FUSED_ANCHOR", raising=False)
    yield


def _classify(db, d_off, d_cont, n, flags):
    import torch
    out = torch.zeros(n * 5, dtype=torch.int16, device="cuda")
    st = torch.zeros(2, dtype=torch.int64, device="cuda")
    db.query_device(d_off, d_cont, n, out.data_ptr(), 0, st.data_ptr(), flags)
    torch.cuda.synchronize()
    return out.cpu().numpy().view(np.uint16).reshape(n, 5), int(st[0]), int(st[1])


@pytest.mark.parametrize("k,L,G,LG", [(31, 150, 6, 60000), (31, 100, 3, 30000), (21, 140, 5, 50000), (31, 151, 4, 40000), (31, 200, 4, 40000),
                                       (32, 192, 4, 30000), (12, 120, 3, 20000), (27, 33, 3, 20000)])
def test_anchor_path_equals_plain_path_and_oracle(oracle, k, L, G, LG):
    import torch
    from jn_cuclark_b200 import CuClarkDB, CCB_STATS, CCB_NO_ANCHOR, CCB_WARP_KERNEL
    NR = 6000
    db = CuClarkDB(1, k, 1, G)
    db.synth_db(11, G, LG, G)
    info = db.anchor_synth(11, G, LG, G)
    assert info.ready and info.entries > 0.08 * G * LG and info.dropped < 0.08 * info.entries
    d_off, d_cont, nc = db.synth_reads_packed(22, 11, 0, NR, L, G, LG)
    plain, l0, s0 = _classify(db, d_off, d_cont, NR, CCB_STATS | CCB_WARP_KERNEL)      # warp-per-read kernel, table only
    # the index in use: the block kernel takes the short reads, the warp kernel the rest
    fast, l1, s1 = _classify(db, d_off, d_cont, NR, CCB_STATS)
    assert np.array_equal(fast, plain)
    assert l1 == l0                                   # every k-mer is still accounted for
    if 100 <= L <= 192:
        assert s1 < 0.6 * s0, (s0, s1)                # and most of them no longer touch the table
    elif L < 100:
        pass                                          # a handful of windows: the index's own sectors outweigh the saving
    else:
        assert 0.99 * s0 < s1 <= s0                   # longer reads take the plain path (all but a few cut short by an N)
    # thread-per-read kernel without the index: same probes, same sectors
    lane, l2, s2_ = _classify(db, d_off, d_cont, NR, CCB_STATS | CCB_NO_ANCHOR)
    assert np.array_equal(lane, plain) and l2 == l0 and s2_ == s0
    # against the oracle
    off = torch.empty(NR + 1, dtype=torch.int32, device="cuda"); cont = torch.empty(nc, dtype=torch.int16, device="cuda")
    import ctypes as C
    rt = C.CDLL("libcudart.so")
    rt.cudaMemcpy(C.c_void_p(off.data_ptr()), C.c_void_p(d_off), C.c_size_t((NR + 1) * 4), 3)
    rt.cudaMemcpy(C.c_void_p(cont.data_ptr()), C.c_void_p(d_cont), C.c_size_t(nc * 2), 3)
    odb = oracle.OracleDb.synth(11, G, LG, G, k, 1000003)
    want, _, nl = odb.query_packed(k, off.cpu().numpy().view(np.uint32), cont.cpu().numpy().view(np.uint16))
    assert np.array_equal(fast, want) and nl == l1
    # dropping the index restores the plain path
    db.anchor_drop()
    again, _, s2 = _classify(db, d_off, d_cont, NR, CCB_STATS)
    assert np.array_equal(again, plain) and s2 == s0
    db.device_free(d_off); db.device_free(d_cont); db.close()


@pytest.mark.parametrize("fname", ["reads.fa", "reads_ml.fa", "reads.fq", "pairs"])
def test_lane_kernel_on_files_with_quirks(small_world, oracle, fname, tmp_path):
    """Short reads, N runs (several parts), reads shorter than k, lower case, paired reads joined by N: the
    thread-per-read kernel takes what it can and hands the rest back; u16[5] and sparse rows equal the oracle's."""
    import os
    from jn_cuclark_b200 import CuClarkDB, CCB_STATS
    odb = oracle.OracleDb.build(small_world["targets"], 31, 1000003)
    if fname == "pairs":
        m = str(tmp_path / "m.fa")
        assert oracle.merge_paired(os.path.join(small_world["dir"], "p1.fq"), os.path.join(small_world["dir"], "p2.fq"), m) == 0
        data = open(m, "rb").read()
    else:
        data = open(os.path.join(small_world["dir"], fname), "rb").read()
    ir = oracle.IndexedReads(data)
    off, cont = ir.pack(31)
    want, wrows, nl = odb.query_packed(31, off, cont, row_stride=32)
    db = CuClarkDB(1, 31, 2, len(odb.labels))
    c, l = odb.pairs()
    db.load_pairs(c, l)
    before = db.stats().lookups
    got, rows = db.classify_packed(off, cont, extended=True, flags=CCB_STATS)
    assert db.stats().lookups - before == nl
    db.close()
    assert np.array_equal(got, want)
    n = wrows[:, 0].astype(int)
    for i in range(len(n)):
        assert rows[i, 0] == n[i] and np.array_equal(rows[i, 1:1 + 2 * n[i]], wrows[i, 1:1 + 2 * n[i]]), i


def test_lane_kernel_many_targets_goes_back_to_warp_kernel(oracle, tmp_path):
    import synth
    from jn_cuclark_b200 import CuClarkDB
    tp, genomes, _ = synth.make_genomes(str(tmp_path / "g"), 48, 6000, seed=5, shared=False, quirks=False, same_label_pair=False)
    odb = oracle.OracleDb.build(tp, 31, 1000003)
    reads = [("five", np.concatenate([g[100:136] for g in genomes[:5]])),            # 5 targets in 180 bases
             ("four", np.concatenate([g[200:240] for g in genomes[:4]])),            # exactly 4
             ("mosaic40", np.concatenate([g[300:340] for g in genomes[:40]])),       # long and > 32 targets: dense pass
             ("plain", genomes[7][1000:1150])]
    synth.write_reads_fasta(str(tmp_path / "r.fa"), reads)
    ir = oracle.IndexedReads(open(str(tmp_path / "r.fa"), "rb").read())
    off, cont = ir.pack(31)
    want, _, _ = odb.query_packed(31, off, cont)
    db = CuClarkDB(1, 31, 1, len(odb.labels))
    c, l = odb.pairs()
    db.load_pairs(c, l)
    got, _ = db.classify_packed(off, cont)
    db.close()
    assert np.array_equal(got, want)


def _world_texts(small_world, odb):
    out = []
    for line in open(small_world["targets"]):
        f, lab = line.split()[:2]
        out.append((open(f, "rb").read(), odb.labels.index(lab)))
    return out


@pytest.mark.parametrize("light", [False, True])
@pytest.mark.parametrize("fname", ["reads.fa", "reads.fq", "pairs"])
def test_anchor_from_fasta_targets(small_world, oracle, fname, light, tmp_path):
    """The index built from the target FASTA files (N runs, lower case, several records per file, sequence
    shared between targets, two files under one label).  In a light database a text k-mer can be stored under
    ANOTHER target's label (it was sampled there, not here): those windows must still be probed."""
    import os
    from jn_cuclark_b200 import CuClarkDB, CCB_STATS, CCB_NO_ANCHOR
    k = 27 if light else 31
    odb = oracle.OracleDb.build(small_world["targets"], k, 1000003, light=light, gap=4)
    if fname == "pairs":      # two parts per read (mates joined by N): both go through the block kernel, hence the index
        m = str(tmp_path / "m.fa")
        assert oracle.merge_paired(os.path.join(small_world["dir"], "p1.fq"), os.path.join(small_world["dir"], "p2.fq"), m) == 0
        data = open(m, "rb").read()
    else:
        data = open(os.path.join(small_world["dir"], fname), "rb").read()
    ir = oracle.IndexedReads(data)
    off, cont = ir.pack(k)
    want, wrows, nl = odb.query_packed(k, off, cont, row_stride=32)
    db = CuClarkDB(1, k, 2, len(odb.labels))
    c, l = odb.pairs()
    db.load_pairs(c, l)
    info = db.anchor_from_fasta(_world_texts(small_world, odb))
    assert info.ready and info.entries > 0 and info.dropped <= 0.02 * info.entries
    b0 = db.stats()
    got, rows = db.classify_packed(off, cont, extended=True, flags=CCB_STATS)
    b1 = db.stats()
    plain, _ = db.classify_packed(off, cont, extended=True, flags=CCB_STATS | CCB_NO_ANCHOR)
    b2 = db.stats()
    db.close()
    assert np.array_equal(got, want) and np.array_equal(plain, want)
    assert b1.lookups - b0.lookups == nl == b2.lookups - b1.lookups
    assert b1.sectors - b0.sectors < 0.85 * (b2.sectors - b1.sectors)
    n = wrows[:, 0].astype(int)
    for i in range(len(n)):
        assert rows[i, 0] == n[i] and np.array_equal(rows[i, 1:1 + 2 * n[i]], wrows[i, 1:1 + 2 * n[i]]), i


def test_anchor_with_a_table_from_other_labels(oracle, tmp_path):
    """Worst case for the text's own bits: the table says label B for k-mers the text holds under label A
    (database and target files do not belong together).  'other' bits send those windows to the table."""
    import synth
    from jn_cuclark_b200 import CuClarkDB
    tp, genomes, _ = synth.make_genomes(str(tmp_path / "g"), 4, 8000, seed=9, shared=False, quirks=False, same_label_pair=False)
    odb = oracle.OracleDb.build(tp, 31, 1000003)
    c, l = odb.pairs()
    l2 = ((l.astype(np.int64) + 1) % 4).astype(np.uint16)            # every label rotated
    odb2 = oracle.OracleDb.from_pairs(c, l2, 31, 1000003)
    reads = synth.make_reads(genomes, 600, 150, seed=10, k=31)
    synth.write_reads_fasta(str(tmp_path / "r.fa"), reads)
    ir = oracle.IndexedReads(open(str(tmp_path / "r.fa"), "rb").read())
    off, cont = ir.pack(31)
    want, _, _ = odb2.query_packed(31, off, cont)
    db = CuClarkDB(1, 31, 1, 4)
    db.load_pairs(c, l2)
    texts = [(open(line.split()[0], "rb").read(), odb.labels.index(line.split()[1])) for line in open(tp)]
    db.anchor_from_fasta(texts)
    got, _ = db.classify_packed(off, cont)
    db.close()
    assert np.array_equal(got, want)
