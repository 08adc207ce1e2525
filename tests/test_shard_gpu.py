"""Sharded-table mode (row E2): count -> route -> probe -> vote gives the oracle's u16[5] / rows bit for bit,
for 1, 2, 3 and 8 shards.  With one GPU all shards live on the same device and the exchange is a pointer
hand-over; with two or more GPUs the same kernels run one process per GPU with NCCL all-to-all in between
(jn_cuclark_b200/sharded.py)."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
HT_SMALL = 1000003
HERE = os.path.dirname(os.path.abspath(__file__))


def _classify_sharded_one_gpu(shards, off, cont, extended=False):
    """All shards on this device: shard g's handle probes segment g of the send buffer."""
    import torch
    G = len(shards)
    n = off.size - 1
    dev = torch.device("cuda", 0)
    d_off = torch.from_numpy(np.ascontiguousarray(off, dtype=np.uint32).view(np.int32)).to(dev)
    d_cont = torch.from_numpy(np.ascontiguousarray(cont, dtype=np.uint16).view(np.int16)).to(dev)
    pos = torch.empty(G * n + 1, dtype=torch.int32, device=dev)
    rq = shards[0]
    rq.shard_count(d_off.data_ptr(), d_cont.data_ptr(), n, G, pos.data_ptr())
    bounds = [int(pos[g * n].item()) & 0xFFFFFFFF for g in range(G + 1)]
    total = bounds[G]
    send = torch.full((max(total, 1),), -1, dtype=torch.int64, device=dev)
    rq.shard_route(d_off.data_ptr(), d_cont.data_ptr(), n, G, pos.data_ptr(), bounds[:G],
                   [send.data_ptr() + 8 * bounds[g] for g in range(G)])
    labels = torch.full((max(total, 1),), 0x7777, dtype=torch.int16, device=dev)
    stats = torch.zeros(2, dtype=torch.int64, device=dev)
    for g in range(G):
        shards[g].shard_probe(send.data_ptr() + 8 * bounds[g], bounds[g + 1] - bounds[g], labels.data_ptr() + 2 * bounds[g], stats.data_ptr())
    out = torch.zeros(n * 5, dtype=torch.int16, device=dev)
    rows = torch.zeros(n * rq.row_stride, dtype=torch.int16, device=dev) if extended else None
    rq.shard_vote(n, G, pos.data_ptr(), bounds[:G], [labels.data_ptr() + 2 * bounds[g] for g in range(G)],
                  out.data_ptr(), rows.data_ptr() if extended else 0)
    torch.cuda.synchronize()
    o = out.cpu().numpy().view(np.uint16).reshape(n, 5)
    r = rows.cpu().numpy().view(np.uint16).reshape(n, rq.row_stride) if extended else None
    return o, r, int(stats[0]), bounds


def _make_shards(odb, G, k=31, ext=False):
    from jn_cuclark_b200 import CuClarkDB
    c, l = odb.pairs()
    out = []
    for g in range(G):
        db = CuClarkDB(1, k, 1, len(odb.labels))
        db.shard_config(G, g)
        db.load_pairs(c, l)
        if ext:
            db.malloc(1, 1, True)
        out.append(db)
    return out


@pytest.mark.parametrize("G", [1, 2, 3, 8])
def test_sharded_equals_oracle(small_world, oracle, G):
    odb = oracle.OracleDb.build(small_world["targets"], 31, HT_SMALL)
    ir = oracle.IndexedReads(open(os.path.join(small_world["dir"], "reads.fa"), "rb").read())
    off, cont = ir.pack(31)
    want, wrows, nl = odb.query_packed(31, off, cont, row_stride=32)
    shards = _make_shards(odb, G)
    try:
        entries = [s.db_info().entries for s in shards]
        assert sum(entries) == odb.n                       # every k-mer lives on exactly one shard
        if G > 1:
            assert min(entries) > 0.5 * odb.n / G          # and the hash space is cut evenly
        got, rows, lookups, bounds = _classify_sharded_one_gpu(shards, off, cont, extended=True)
        assert lookups == nl and bounds[G] == nl
        assert np.array_equal(got, want)
        n = wrows[:, 0].astype(int)
        for i in range(len(n)):
            assert rows[i, 0] == n[i] and np.array_equal(rows[i, 1:1 + 2 * n[i]], wrows[i, 1:1 + 2 * n[i]])
        # a whole-table query on a shard must refuse, not read out of bounds
        if G > 1:
            from jn_cuclark_b200 import CcbError
            with pytest.raises(CcbError):
                shards[0].classify_packed(off, cont)
    finally:
        for s in shards:
            s.close()


def test_sharded_long_reads_many_targets_and_empty(oracle, tmp_path):
    import synth
    tp, genomes, _ = synth.make_genomes(str(tmp_path / "g"), 48, 12000, seed=5, shared=False, quirks=False, same_label_pair=False)
    odb = oracle.OracleDb.build(tp, 31, HT_SMALL)
    rng = np.random.default_rng(9)
    reads = []
    for i in range(12):
        s = genomes[i % 48][500:10500].copy()
        m = rng.random(s.size) < 0.05
        s[m] = synth.BASES[rng.integers(0, 4, size=int(m.sum()))]
        reads.append((f"long{i}", s))
    reads.append(("mosaic48", np.concatenate([g[100:180] for g in genomes])))      # > 32 targets: dense vote
    reads.append(("short", genomes[0][:20]))                                       # no k-mer at all
    reads.append(("allN", np.full(200, ord("N"), np.uint8)))
    synth.write_reads_fasta(str(tmp_path / "r.fa"), reads)
    ir = oracle.IndexedReads(open(str(tmp_path / "r.fa"), "rb").read())
    off, cont = ir.pack(31)
    want, _, nl = odb.query_packed(31, off, cont)
    shards = _make_shards(odb, 4)
    try:
        got, _, lookups, _ = _classify_sharded_one_gpu(shards, off, cont)
        assert lookups == nl and np.array_equal(got, want)
        # empty batch
        got0, _, l0, _ = _classify_sharded_one_gpu(shards, np.zeros(1, np.uint32), np.zeros(0, np.uint16))
        assert got0.shape == (0, 5) and l0 == 0
    finally:
        for s in shards:
            s.close()


def test_peer_exchange_kernels_one_gpu(small_world, oracle):
    """The peer-store exchange with every shard on this one device: route kernels store into the owners'
    inbox regions, probe kernels store labels into the requesters' label boxes, counts travel as device words."""
    import torch
    from jn_cuclark_b200.sharded import PeerShardedClassifier
    G = 3
    odb = oracle.OracleDb.build(small_world["targets"], 31, HT_SMALL)
    ir = oracle.IndexedReads(open(os.path.join(small_world["dir"], "reads.fa"), "rb").read())
    off, cont = ir.pack(31)
    n = off.size - 1
    want, _, nl = odb.query_packed(31, off, cont)
    shards = _make_shards(odb, G)
    dev = torch.device("cuda", 0)
    cap = int(nl)                                           # worst case: everything to one owner
    inbox = [torch.zeros(G * cap, dtype=torch.int64, device=dev) for _ in range(G)]
    labbox = [torch.zeros(G * cap, dtype=torch.int16, device=dev) for _ in range(G)]
    counts = [torch.zeros(G, dtype=torch.int64, device=dev) for _ in range(G)]
    peers = lambda r: {"rank": r, "inbox": [t.data_ptr() for t in inbox], "labbox": [t.data_ptr() for t in labbox],
                       "counts": [t.data_ptr() for t in counts]}
    try:
        cls = [PeerShardedClassifier(shards[r], cap, n, peers=peers(r)) for r in range(G)]
        # "rank" r classifies the r-th third of the reads
        cut = [0, n // 3, 2 * n // 3, n]
        parts = []
        for r in range(G):
            a, b = cut[r], cut[r + 1]
            o = (off[a:b + 1] - off[a]).astype(np.uint32)
            c = cont[int(off[a]):int(off[b])]
            parts.append((torch.from_numpy(o.view(np.int32)).to(dev), torch.from_numpy(np.ascontiguousarray(c).view(np.int16)).to(dev), b - a))
        stats = torch.zeros(2, dtype=torch.int64, device=dev)
        for r in range(G):
            cls[r].stage_route(parts[r][0].data_ptr(), parts[r][1].data_ptr(), parts[r][2])
        for r in range(G):
            cls[r].stage_probe(stats)
        outs = [torch.zeros(parts[r][2] * 5, dtype=torch.int16, device=dev) for r in range(G)]
        for r in range(G):
            cls[r].stage_vote(parts[r][2], outs[r])
        torch.cuda.synchronize()
        got = np.concatenate([o.cpu().numpy().view(np.uint16).reshape(-1, 5) for o in outs])
        assert int(stats[0]) == nl and not any(c.overflowed() for c in cls)
        assert np.array_equal(got, want)
        assert int(sum(int(c.sum()) for c in counts)) == nl
        for c in cls:
            c.close()
        # capacity too small: flagged, nothing written or read out of bounds -- the whole classify(), vote included
        small = [PeerShardedClassifier(shards[r], 1000, n, peers=peers(r)) for r in range(G)]
        junk = torch.zeros(parts[0][2] * 5, dtype=torch.int16, device=dev)
        small[0].classify(parts[0][0].data_ptr(), parts[0][1].data_ptr(), parts[0][2], junk)
        torch.cuda.synchronize()
        assert small[0].overflowed() and not small[0].overflowed()          # reading the flag clears it
        for c in small:
            c.close()
    finally:
        for s_ in shards:
            s_.close()


def _nccl_worker(rank, world, port, q):
    try:
        _nccl_worker_body(rank, world, port, q)
    except BaseException as e:                       # report instead of leaving the parent to time out
        import traceback
        q.put((rank, False, -1, 0, 0, traceback.format_exc()[-1500:]))
        raise


def _nccl_worker_body(rank, world, port, q):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, HERE); sys.path.insert(0, os.path.dirname(HERE))
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    from jn_cuclark_b200 import CuClarkDB
    from jn_cuclark_b200.sharded import ShardedClassifier
    G_, LG, L, NR, k = 6, 50000, 150, 4000 + 500 * rank, 31
    whole = CuClarkDB(1, k, 1, G_, device=rank)
    whole.synth_db(11, G_, LG, G_)
    d_off, d_cont, _ = whole.synth_reads_packed(22, 11, rank * 100000, NR, L, G_, LG)
    ref = torch.zeros(NR * 5, dtype=torch.int16, device="cuda")
    whole.query_device(d_off, d_cont, NR, ref.data_ptr())
    torch.cuda.synchronize()
    part = CuClarkDB(1, k, 1, G_, device=rank)
    part.shard_config(world, rank)
    part.synth_db(11, G_, LG, G_)
    cap = (4000 + 500 * world) * (L - k + 1)                                       # the same on every rank
    sc = ShardedClassifier(part, cap, 4000 + 500 * world)
    out = torch.zeros(NR * 5, dtype=torch.int16, device="cuda")
    for _ in range(2):
        out.zero_()
        sc.classify(d_off, d_cont, NR, out)
    torch.cuda.synchronize()
    same = bool(torch.equal(out, ref))
    sc.close()
    # the same through the peer-store exchange (CUDA IPC inboxes, no collective on the data path)
    from jn_cuclark_b200.sharded import PeerShardedClassifier
    pc = PeerShardedClassifier(part, cap, 4000 + 500 * world)
    for _ in range(3):                                   # buffers are reused batch after batch
        out2 = torch.zeros(NR * 5, dtype=torch.int16, device="cuda")
        pc.classify(d_off, d_cont, NR, out2)
    torch.cuda.synchronize()
    same = same and bool(torch.equal(out2, ref)) and not pc.overflowed()
    pc.close()
    classified = int((ref.view(-1, 5)[:, 1] != 0).sum())
    q.put((rank, same, classified, part.db_info().entries, whole.db_info().entries, sc.last["sent"]))
    dist.barrier()
    whole.device_free(d_off); whole.device_free(d_cont); whole.close(); part.close()
    dist.destroy_process_group()


def test_sharded_two_processes_nccl():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29700 + os.getpid() % 200
    ps = [ctx.Process(target=_nccl_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    got = sorted(q.get(timeout=300) for _ in ps)
    assert all(g[1] for g in got), got
    for p in ps:
        p.join(120)
        assert p.exitcode == 0
    assert all(g[1] for g in got), got                      # sharded == replicated, bit for bit
    assert all(g[2] > 1000 for g in got)
    assert got[0][3] + got[1][3] == got[0][4]               # shards partition the table


def _classify_one_pass_one_gpu(shards, off, cont, cap, extended=False, n_containers=-1):
    """The one-pass pipeline (ccb_shard_route2 / ccb_shard_vote2) with every shard on this device: requester = shard 0's
    handle, owner g probes region g of the requester's send buffer."""
    import torch
    G = len(shards)
    n = off.size - 1
    dev = torch.device("cuda", 0)
    d_off = torch.from_numpy(np.ascontiguousarray(off, dtype=np.uint32).view(np.int32)).to(dev)
    d_cont = torch.from_numpy(np.ascontiguousarray(cont, dtype=np.uint16).view(np.int16)).to(dev) if cont.size else torch.zeros(1, dtype=torch.int16, device=dev)
    rq = shards[0]
    ws = rq.shardws_create(max(n, 1), G, cap)
    send = torch.full((G * cap,), -1, dtype=torch.int64, device=dev)
    sent = torch.zeros(G, dtype=torch.int64, device=dev)
    labels = torch.full((G * cap,), 0x7777, dtype=torch.int16, device=dev)
    stats = torch.zeros(2, dtype=torch.int64, device=dev)
    rq.shard_route2(ws, d_off.data_ptr(), d_cont.data_ptr(), n, n_containers, [send.data_ptr() + 8 * g * cap for g in range(G)],
                    [sent.data_ptr() + 8 * g for g in range(G)])
    counts = [int(x) for x in sent.cpu().tolist()]
    for g in range(G):
        shards[g].shard_probe(send.data_ptr() + 8 * g * cap, min(counts[g], cap), labels.data_ptr() + 2 * g * cap, stats.data_ptr())
    out = torch.full((max(n, 1) * 5,), 0x5555, dtype=torch.int16, device=dev)
    rows = torch.zeros(max(n, 1) * rq.row_stride, dtype=torch.int16, device=dev) if extended else None
    rq.shard_vote2(ws, n, [labels.data_ptr() + 2 * g * cap for g in range(G)], out.data_ptr(), rows.data_ptr() if extended else 0)
    torch.cuda.synchronize()
    err = rq.shardws_error(ws)
    rq.shardws_destroy(ws)
    o = out.cpu().numpy().view(np.uint16)[: n * 5].reshape(n, 5)
    r = rows.cpu().numpy().view(np.uint16)[: n * rq.row_stride].reshape(n, rq.row_stride) if extended else None
    return o, r, int(stats[0]), counts, err


@pytest.mark.parametrize("G", [1, 2, 3, 8])
@pytest.mark.parametrize("fname", ["reads.fa", "pairs"])
def test_one_pass_pipeline_equals_oracle(small_world, oracle, tmp_path, G, fname):
    """Short reads of one part and pairs joined by N (two parts) through k_route_block / k_vote_block; the odd reads of
    the world (several parts, shorter than k) through the generic passes behind them; rows included."""
    odb = oracle.OracleDb.build(small_world["targets"], 31, HT_SMALL)
    if fname == "pairs":
        m = str(tmp_path / "merged.fa")
        assert oracle.merge_paired(os.path.join(small_world["dir"], "p1.fq"), os.path.join(small_world["dir"], "p2.fq"), m) == 0
        data = open(m, "rb").read()
    else:
        data = open(os.path.join(small_world["dir"], fname), "rb").read()
    ir = oracle.IndexedReads(data)
    off, cont = ir.pack(31)
    want, wrows, nl = odb.query_packed(31, off, cont, row_stride=32)
    shards = _make_shards(odb, G)
    try:
        got, rows, lookups, counts, err = _classify_one_pass_one_gpu(shards, off, cont, cap=int(nl) + 64, extended=True, n_containers=int(off[-1]))
        assert not err and lookups == nl and sum(counts) == nl
        assert np.array_equal(got, want)
        n = wrows[:, 0].astype(int)
        for i in range(len(n)):
            assert rows[i, 0] == n[i] and np.array_equal(rows[i, 1:1 + 2 * n[i]], wrows[i, 1:1 + 2 * n[i]])
    finally:
        for s in shards:
            s.close()


def test_one_pass_pipeline_odd_batches(oracle, tmp_path):
    """Long reads (generic path only), a short read that meets more than four targets (block path -> dense re-vote), a read
    that meets 48 (generic -> dense), reads without k-mers, an empty batch, a region that is too small."""
    import synth
    tp, genomes, _ = synth.make_genomes(str(tmp_path / "g"), 48, 12000, seed=5, shared=False, quirks=False, same_label_pair=False)
    odb = oracle.OracleDb.build(tp, 31, HT_SMALL)
    rng = np.random.default_rng(9)
    short = [(f"s{i}", genomes[i % 48][1000 + 7 * i:1150 + 7 * i].copy()) for i in range(600)]
    short.append(("mosaic6", np.concatenate([g[100:132] for g in genomes[:6]])))      # 192 bases, six targets: block path, fifth target
    short.append(("mosaic5", np.concatenate([g[200:236] for g in genomes[3:8]])))
    short.append(("tiny", genomes[0][:20])); short.append(("allN", np.full(150, ord("N"), np.uint8)))
    long_ = []
    for i in range(12):
        s = genomes[i % 48][500:10500].copy()
        m = rng.random(s.size) < 0.05
        s[m] = synth.BASES[rng.integers(0, 4, size=int(m.sum()))]
        long_.append((f"long{i}", s))
    long_.append(("mosaic48", np.concatenate([g[100:180] for g in genomes])))
    shards = _make_shards(odb, 4)
    try:
        for name, reads, hint in (("short", short, True), ("mixed", short + long_, True), ("long", long_, True), ("long_nohint", long_, False)):
            synth.write_reads_fasta(str(tmp_path / f"{name}.fa"), reads)
            ir = oracle.IndexedReads(open(str(tmp_path / f"{name}.fa"), "rb").read())
            off, cont = ir.pack(31)
            want, _, nl = odb.query_packed(31, off, cont)
            got, _, lookups, counts, err = _classify_one_pass_one_gpu(shards, off, cont, cap=int(nl) + 64, n_containers=int(off[-1]) if hint else -1)
            assert not err and lookups == nl, name
            assert np.array_equal(got, want), name
            if name == "short":
                assert (want[-4:-2, 1] > 0).all()                                        # the mosaics are classified
        got0, _, l0, _, err0 = _classify_one_pass_one_gpu(shards, np.zeros(1, np.uint32), np.zeros(0, np.uint16), cap=64)
        assert got0.shape == (0, 5) and l0 == 0 and not err0
        # a region that is too small: flagged; the vote reads nothing beyond the region
        _, _, _, counts, err = _classify_one_pass_one_gpu(shards, off, cont, cap=1000)
        assert err
    finally:
        for s in shards:
            s.close()


def test_probe_range_by_range(small_world, oracle, monkeypatch):
    """A shard larger than the TLBs reach is probed range by range (query.cu probe_passes): with the range set to 1 MB the
    small table takes several passes -- every k-mer is still probed exactly once."""
    odb = oracle.OracleDb.build(small_world["targets"], 31, HT_SMALL)
    ir = oracle.IndexedReads(open(os.path.join(small_world["dir"], "reads.fa"), "rb").read())
    off, cont = ir.pack(31)
    want, _, nl = odb.query_packed(31, off, cont)
    shards = _make_shards(odb, 2)
    try:
        assert shards[0].db_info().bytes > 3 * (1 << 20)
        monkeypatch.setenv("CCB_PROBE_RANGE_MB", "1")
        got, _, lookups, counts, err = _classify_one_pass_one_gpu(shards, off, cont, cap=int(nl) + 64)
        assert not err and lookups == nl and np.array_equal(got, want)
        got2, _, lookups2, _ = _classify_sharded_one_gpu(shards, off, cont)
        assert lookups2 == nl and np.array_equal(got2, want)
    finally:
        for s in shards:
            s.close()
