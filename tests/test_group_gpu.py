"""Device group (row E2 as a product): ccb_group_create -- one process, the table sharded over the listed devices, the
ordinary batch calls on the leader handle.  Results must equal the oracle's bit for bit.  With one GPU the shards share
device 0 (the same kernels, events and peer-pointer plumbing); with two or more they sit on different GPUs."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
HT_SMALL = 1000003


def _devices(G):
    import torch
    n = torch.cuda.device_count()
    return [g % n for g in range(G)]


def _group(odb, G, slots=3, k=31):
    from jn_cuclark_b200 import CuClarkGroup
    grp = CuClarkGroup(_devices(G), k, slots, len(odb.labels))
    c, l = odb.pairs()
    for m in grp.members:
        m.load_pairs(c, l)                      # every member keeps its own shard of the pairs
    assert sum(m.db_info().entries for m in grp.members) == odb.n == grp.db_info().entries
    return grp


@pytest.mark.parametrize("G", [1, 2, 3])
def test_group_equals_oracle(small_world, oracle, tmp_path, G):
    odb = oracle.OracleDb.build(small_world["targets"], 31, HT_SMALL)
    grp = _group(odb, G)
    try:
        m = str(tmp_path / "merged.fa")
        assert oracle.merge_paired(os.path.join(small_world["dir"], "p1.fq"), os.path.join(small_world["dir"], "p2.fq"), m) == 0
        for path, ext in ((os.path.join(small_world["dir"], "reads.fa"), True), (os.path.join(small_world["dir"], "reads.fq"), False), (m, True)):
            ir = oracle.IndexedReads(open(path, "rb").read())
            off, cont = ir.pack(31)
            want, wrows, _ = odb.query_packed(31, off, cont, row_stride=32)
            got, rows = grp.classify_packed(off, cont, extended=ext)
            assert np.array_equal(got, want), path
            if ext:
                n = wrows[:, 0].astype(int)
                for i in range(len(n)):
                    assert rows[i, 0] == n[i] and np.array_equal(rows[i, 1:1 + 2 * n[i]], wrows[i, 1:1 + 2 * n[i]])
    finally:
        grp.close()


def test_group_slots_in_flight_and_odd_batches(oracle, tmp_path):
    """Several slots in flight at once (homes on different devices when there are several), a long-read batch, an empty
    batch, and an exchange region that is too small (CCB_SHARD_CAP): the batch fails loudly in waitForBatch."""
    import synth
    from jn_cuclark_b200 import CcbError
    tp, genomes, _ = synth.make_genomes(str(tmp_path / "g"), 12, 12000, seed=5, shared=False, quirks=False, same_label_pair=False)
    odb = oracle.OracleDb.build(tp, 31, HT_SMALL)
    rng = np.random.default_rng(3)
    short = [(f"s{i}", genomes[i % 12][1000 + 5 * i:1150 + 5 * i].copy()) for i in range(900)]
    long_ = []
    for i in range(6):
        s = genomes[i][500:10500].copy()
        mm = rng.random(s.size) < 0.05
        s[mm] = synth.BASES[rng.integers(0, 4, size=int(mm.sum()))]
        long_.append((f"long{i}", s))
    packs = []
    for name, reads in (("a", short[:300]), ("b", short[300:600] + long_[:2]), ("c", long_), ("d", short[600:])):
        synth.write_reads_fasta(str(tmp_path / f"{name}.fa"), reads)
        ir = oracle.IndexedReads(open(str(tmp_path / f"{name}.fa"), "rb").read())
        off, cont = ir.pack(31)
        packs.append((off, cont, odb.query_packed(31, off, cont)[0]))
    grp = _group(odb, 2, slots=4)
    try:
        grp.malloc(max(p[0].size - 1 for p in packs), max(int(p[0][-1]) for p in packs), False)
        for s, (off, cont, _) in enumerate(packs):
            o, c, _, _ = grp.buffers(s)
            o[: off.size] = off; c[: int(off[-1])] = cont[: int(off[-1])]
            grp.readyBatch(s, off.size - 1, int(off[-1])); grp.queryBatch(s)
        for s in (3, 1, 0, 2):
            grp.waitForBatch(s)
        for s, (off, _, want) in enumerate(packs):
            assert np.array_equal(grp.buffers(s)[2][: off.size - 1], want), s
        grp.readyBatch(0, 0, 0); grp.queryBatch(0); grp.waitForBatch(0)          # empty batch
        grp.freeBatchMemory()
        os.environ["CCB_SHARD_CAP"] = "500"
        try:
            with pytest.raises(CcbError) as e:
                grp.classify_packed(packs[0][0], packs[0][1])
            assert "increase the number of batches" in str(e.value)
        finally:
            del os.environ["CCB_SHARD_CAP"]
    finally:
        grp.close()
