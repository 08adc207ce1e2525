"""world_size-2 CPU tests of the N>1 path (row E1 of SURVEY.md section 8): reads are independent units,
the table is replicated, every rank classifies its own contiguous slice of the read stream and no
data-path collective exists.  What N>1 adds on the host side is (1) the slice arithmetic of the
counter-based read stream, (2) the max-over-ranks timing reduction and (3) the launch contract of
bench.py under torchrun; all three run here over gloo."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def _pack_reads(first, n, length=80, genomes=3, genome_len=5000):
    """The packed slice [first, first+n) of the synthetic read stream, from the oracle's generator
    (same counter-based stream as ccb_synth.h)."""
    import oracle_py
    import tempfile
    d = tempfile.mkdtemp(prefix="ccb_mr_")
    p = os.path.join(d, "r.fa")
    oracle_py.lib().ccbo_synth_reads_fasta(p.encode(), 77, 5, first, n, length, genomes, genome_len)
    return open(p, "rb").read()


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    per = 500
    mine = _pack_reads(rank * per, per)
    # rank r's slice equals the r-th slice of the stream generated in one piece
    whole = _pack_reads(0, per * world)
    recs = whole.split(b">")[1:]
    mine_recs = mine.split(b">")[1:]
    seqs_whole = [r.split(b"\n", 1)[1] for r in recs[rank * per:(rank + 1) * per]]
    seqs_mine = [r.split(b"\n", 1)[1] for r in mine_recs]
    ok = seqs_whole == seqs_mine and len(mine_recs) == per
    # max-over-ranks timing reduction as bench.py does it
    t = torch.tensor([1.0 + rank], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    n = torch.tensor([per], dtype=torch.int64)
    dist.all_reduce(n)
    q.put((rank, ok, float(t[0]), int(n[0])))
    dist.barrier()
    dist.destroy_process_group()


def test_read_stream_slices_and_reductions_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + os.getpid() % 300
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    got = sorted(q.get(timeout=180) for _ in ps)
    for p in ps:
        p.join(60)
        assert p.exitcode == 0
    assert [g[1] for g in got] == [True, True]
    assert [g[2] for g in got] == [2.0, 2.0]          # max over ranks
    assert [g[3] for g in got] == [1000, 1000]        # units of all ranks


@pytest.mark.skipif(not os.path.exists(os.path.join(ROOT, "oracle", "build", "liboracle.so")), reason="oracle not built")
def test_bench_reference_arm_under_torchrun_prints_one_json_line():
    port = 29300 + os.getpid() % 300
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "bench.py"),
           "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0",
           "--sample-genomes", "1", "--genome-len", "200000", "--sample-reads", "2000"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, r.stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == 2 and d["value"] > 0
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["cpu_baseline"]["kind"] in ("reference", "port")
