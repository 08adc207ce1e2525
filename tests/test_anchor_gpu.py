"""Anchor index (sequences + sampled k-mer positions in front of the table): results are identical with and
without it, and identical to the oracle; it only changes how many table sectors are read."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _classify(db, d_off, d_cont, n, flags):
    import torch
    out = torch.zeros(n * 5, dtype=torch.int16, device="cuda")
    st = torch.zeros(2, dtype=torch.int64, device="cuda")
    db.query_device(d_off, d_cont, n, out.data_ptr(), 0, st.data_ptr(), flags)
    torch.cuda.synchronize()
    return out.cpu().numpy().view(np.uint16).reshape(n, 5), int(st[0]), int(st[1])


@pytest.mark.parametrize("k,L,G,LG", [(31, 150, 6, 60000), (31, 100, 3, 30000), (21, 140, 5, 50000), (31, 151, 4, 40000), (31, 200, 4, 40000)])
def test_anchor_path_equals_plain_path_and_oracle(oracle, k, L, G, LG):
    import torch
    from jn_cuclark_b200 import CuClarkDB, CCB_STATS, CCB_NO_ANCHOR
    NR = 6000
    db = CuClarkDB(1, k, 1, G)
    db.synth_db(11, G, LG, G)
    info = db.anchor_synth(11, G, LG, G)
    assert info.ready and info.entries > 0.08 * G * LG and info.dropped < 0.08 * info.entries
    d_off, d_cont, nc = db.synth_reads_packed(22, 11, 0, NR, L, G, LG)
    plain, l0, s0 = _classify(db, d_off, d_cont, NR, CCB_STATS | CCB_NO_ANCHOR)
    fast, l1, s1 = _classify(db, d_off, d_cont, NR, CCB_STATS)
    assert np.array_equal(fast, plain)
    # thread-per-read kernel, without and with the index
    from jn_cuclark_b200 import CCB_LANE_KERNEL
    lane, l2, s2_ = _classify(db, d_off, d_cont, NR, CCB_STATS | CCB_LANE_KERNEL | CCB_NO_ANCHOR)
    assert np.array_equal(lane, plain) and l2 == l0 and s2_ == s0
    lane_a, l3, s3 = _classify(db, d_off, d_cont, NR, CCB_STATS | CCB_LANE_KERNEL)
    assert np.array_equal(lane_a, plain) and l3 == l0
    if L <= 192:
        assert s3 < 0.6 * s0, (s0, s3)
    assert l1 == l0                                   # every k-mer is still accounted for
    if L - k + 1 <= 128:
        assert s1 < 0.6 * s0, (s0, s1)                # and most of them no longer touch the table
    else:
        assert s1 == s0                               # longer reads take the plain path for now
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
    from jn_cuclark_b200 import CuClarkDB, CCB_LANE_KERNEL, CCB_STATS
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
    got, rows = db.classify_packed(off, cont, extended=True, flags=CCB_LANE_KERNEL | CCB_STATS)
    assert db.stats().lookups - before == nl
    db.close()
    assert np.array_equal(got, want)
    n = wrows[:, 0].astype(int)
    for i in range(len(n)):
        assert rows[i, 0] == n[i] and np.array_equal(rows[i, 1:1 + 2 * n[i]], wrows[i, 1:1 + 2 * n[i]]), i


def test_lane_kernel_many_targets_goes_back_to_warp_kernel(oracle, tmp_path):
    import synth
    from jn_cuclark_b200 import CuClarkDB, CCB_LANE_KERNEL
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
    got, _ = db.classify_packed(off, cont, flags=CCB_LANE_KERNEL)
    db.close()
    assert np.array_equal(got, want)
