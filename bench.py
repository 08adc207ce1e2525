#!/usr/bin/env python
"""bench.py -- reads/s and k-mer lookups/s of the cuCLARK classification hot path on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W]            # this repo's CUDA path
  python bench.py --impl reference [...]                         # reference CPU lookup (oracle/_ref)

A step = one pass of the hot path over one batch of synthetic reads
(include/ccb_synth.h).  Default workload (N=1) = BASELINE.json configs[1]:
cuCLARK full mode, k=31, 1,000 x 4 Mbp genomes (~4e9 31-mers), 10 M x 150 bp reads.
`value`  : whole-job k-mer lookups/s, batch resident in HBM, CUDA events on the launch stream.
`e2e`    : the same through the C ABI with HOST (pinned) buffers: H2D + query + D2H per step.
`roofline`: algorithmic bytes = 32 B (one table sector) per lookup vs the measured HBM peak.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

DB_SEED, READS_SEED = 0xC1A2C000 + 2000, 0x5EED0002
HTSIZE_FULL = 1610612741


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--genomes", type=int, default=1000)
    ap.add_argument("--genome-len", type=int, default=4_000_000)
    ap.add_argument("--reads", type=int, default=10_000_000, help="reads per GPU per step")
    ap.add_argument("--read-len", type=int, default=150)
    ap.add_argument("-k", type=int, default=31)
    ap.add_argument("--load-factor", type=float, default=0.5)
    ap.add_argument("--slots", type=int, default=16, help="pinned batch slots of the e2e pipeline (4: 31.0 / 15.0 ms per step, 8: 29.2 / 13.2, 16: 27.1 / 12.1 = the kernel time)")
    ap.add_argument("--sample-genomes", type=int, default=4, help="cpu baseline: genomes in the sample DB")
    ap.add_argument("--sample-reads", type=int, default=200_000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--oracle-htsize", type=int, default=HTSIZE_FULL, help="cpu baseline: HTSIZE of the oracle's table (the reference's 1610612741; tests use a small one)")
    ap.add_argument("--no-gather", action="store_true")
    ap.add_argument("--warp", action="store_true", help="warp-per-read kernel for every read (CCB_WARP_KERNEL)")
    ap.add_argument("--no-anchor", action="store_true", help="do not build the anchor index: table-only path as the headline, no second leg")
    ap.add_argument("--table-only-headline", action="store_true", help="headline = table-only path (SURVEY D3's one sector per lookup), anchor index as the second leg")
    ap.add_argument("--no-workloads", action="store_true", help="skip the C3 (light database, pairs) and C4 (long reads with indels) legs")
    ap.add_argument("--c4-reads", type=int, default=125_000, help="C4: long reads per GPU per step")
    ap.add_argument("--c4-read-len", type=int, default=10_000)
    ap.add_argument("--no-sharded-leg", action="store_true", help="N >= 2: skip the leg with the table sharded over the ranks")
    ap.add_argument("--no-cli-e2e", action="store_true", help="N = 1: skip the file -> CSV leg through bin/cuCLARK")
    ap.add_argument("--cli-genomes", type=int, default=128, help="cli_e2e: genomes of its database (x genome-len, built by bin/cuCLARK itself)")
    ap.add_argument("--cli-reads", type=int, default=10_000_000)
    ap.add_argument("--sharded", action="store_true",
                    help="row E2: the table sharded over the N GPUs (k-mer all-to-all) instead of replicated")
    ap.add_argument("--exchange", default="nccl", choices=["nccl", "peer"],
                    help="--sharded: NCCL all-to-all, or kernels storing into peer inboxes (CUDA IPC over NVLink)")
    ap.add_argument("--chunks", type=int, default=1, help="--sharded: split each rank's batch into this many exchange rounds")
    ap.add_argument("--overlap", action="store_true",
                    help="--sharded --exchange peer: rounds alternate between two streams, each with its own inboxes and barrier "
                         "group, so one round's probe (HBM-bound) runs under the next round's count/route (issue- and NVLink-bound)")
    ap.add_argument("--kernel-only", action="store_true", help="profiling runs: resident-batch loop only (prints no bench line)")
    return ap.parse_args()


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        sm = sorted(float(r[1]) for r in self.rows if len(r) > 2 and r[1].replace(".", "").isdigit())
        mx = max([float(r[2]) for r in self.rows if len(r) > 2 and r[2].replace(".", "").isdigit()] or [0])
        reasons = set()
        for r in self.rows:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


def cpu_sample_oracle(args, threads):
    """Oracle ('port') on a bounded sample: DB of the first sample-genomes genomes, reads drawn from them."""
    import oracle_py
    t0 = time.time()
    odb = oracle_py.OracleDb.synth(DB_SEED, args.sample_genomes, args.genome_len, args.sample_genomes, args.k, HTSIZE_FULL)
    t_build = time.time() - t0
    return odb, t_build


def run_reference(args):
    """Reference arm: the reference's own CPU lookup (hTable::find through oracle/_ref/ref_harness),
    all host threads, on a bounded sample of the same workload; rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import oracle_py
    cores = os.cpu_count() or 1
    import tempfile
    d = tempfile.mkdtemp(prefix="ccb_ref_")
    G, LG, NR, L, k = args.sample_genomes, args.genome_len, args.sample_reads, args.read_len, args.k
    sample = f"DB = first {G} of {args.genomes} genomes x {LG} bp ({G * (LG - k + 1)} {k}-mers, HTSIZE {HTSIZE_FULL}); {NR} x {L} bp reads drawn from them per step"
    harness = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
    odb = oracle_py.OracleDb.synth(DB_SEED, G, LG, G, k, HTSIZE_FULL)
    reads = os.path.join(d, "reads.fa")
    oracle_py.lib().ccbo_synth_reads_fasta(reads.encode(), READS_SEED, DB_SEED, 0, NR, L, G, LG)
    times, lookups, kind = [], 0, "port"
    if os.path.exists(harness) and k == 31:
        kind = "reference"
        tp = os.path.join(d, "targets.txt")
        with open(tp, "w") as f:
            for g in range(G):
                f.write(f"{d}/g{g}.fa\tT{g}\n")
        prefix = oracle_py.db_name(d + "/", k, G, HTSIZE_FULL)
        odb.write(prefix)
        del odb
        r = subprocess.run([harness, "-k", str(k), "-T", tp, "-D", d, "-O", reads, "-n", str(cores),
                            "--reps", str(args.steps + args.warmup)], capture_output=True, text=True)
        for line in r.stdout.splitlines():
            if line.startswith("REFTIME"):
                kv = dict(x.split("=") for x in line.split()[1:])
                times.append(float(kv["seconds"])); lookups = int(kv["lookups"])
        if len(times) < args.steps + args.warmup:
            emit(json.dumps({"impl": "reference", "unavailable": "ref_harness failed: " + r.stderr[-200:].replace("\n", " ")}))
            return
    else:
        ir = oracle_py.IndexedReads(open(reads, "rb").read())
        off, cont = ir.pack(k)
        for _ in range(args.steps + args.warmup):
            t0 = time.time()
            _, _, lookups = odb.query_packed(k, off, cont, threads=cores)
            times.append(time.time() - t0)
    import shutil
    shutil.rmtree(d, ignore_errors=True)
    t = sum(times[args.warmup:])
    v = lookups * args.steps / t
    emit(json.dumps({
        "impl": "reference", "metric": "kmer_lookups_per_s", "value": v, "unit": "lookups/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "reads_per_s": NR * args.steps / t,
        "config": {"workload": f"cuCLARK full k={k}, synthetic {args.genomes} x {LG} bp DB, {L} bp SE reads", "sample": sample},
        "cpu_baseline": {"value": v, "unit": "lookups/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": "lookups/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def _dev_to_host(cudart, np, torch, ptr, n, dtype_t, dtype_np):
    t = torch.empty(n, dtype=dtype_t, device="cuda")
    cudart.cudaMemcpy(C.c_void_p(t.data_ptr()), C.c_void_p(ptr), C.c_size_t(n * t.element_size()), 3)
    return t.cpu().numpy().view(dtype_np)


def run_b200(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from jn_cuclark_b200 import CuClarkDB, Workload, CCB_STATS, CCB_STATS_EX, CCB_NO_ANCHOR, CCB_WARP_KERNEL

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    cudart = C.CDLL("libcudart.so")
    k, G, LG, NR, L = args.k, args.genomes, args.genome_len, args.reads, args.read_len
    NL = min(G, 65534)
    stream = torch.cuda.current_stream().cuda_stream
    peak, peak_src = measured_peak()
    default_workload = (G, LG, NR, L, k) == (1000, 4_000_000, 10_000_000, 150, 31)      # what the ncu captures were taken on
    traffic_tab = {}
    try:
        traffic_tab = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
    except Exception:
        pass

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.int64, device="cuda"); dist.all_reduce(t); return int(t[0])

    def all_ranks(flag):
        if world == 1:
            return flag
        t = torch.tensor([1 if flag else 0], dtype=torch.int32, device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MIN); return bool(int(t[0]))

    # ---- the workload: replicated table per GPU (+ the anchor index, the product default), reads partitioned by rank ----
    db = CuClarkDB(1, k, args.slots, NL, device=local, load_factor=args.load_factor)
    t0 = time.time(); n_kmers = db.synth_db(DB_SEED, G, LG, NL); t_db = time.time() - t0
    info = db.db_info()
    ainfo, anchor_why, t_anchor = None, None, 0.0
    if not args.no_anchor:
        t0 = time.time()
        try:
            ainfo = db.anchor_synth(DB_SEED, G, LG, NL)
        except Exception as e:                                     # CcbError: text over 4 Gbp, memory
            anchor_why = str(e)[:200]
        t_anchor = time.time() - t0
        if not all_ranks(ainfo is not None):
            if ainfo is not None:
                db.anchor_drop()
            ainfo, anchor_why = None, anchor_why or "another rank could not build the index"
    d_off, d_cont, n_cont = db.synth_reads_packed(READS_SEED, DB_SEED, rank * NR, NR, L, G, LG)
    out = torch.zeros(NR * 5, dtype=torch.int16, device="cuda")
    stats = torch.zeros(4, dtype=torch.int64, device="cuda")
    WK = CCB_WARP_KERNEL if args.warp else 0

    def counted(dbx, off, cont, n, o, qf):
        stats.zero_(); o.zero_()
        dbx.query_device(off, cont, n, o.data_ptr(), 0, stats.data_ptr(), CCB_STATS | CCB_STATS_EX | qf, stream)
        torch.cuda.synchronize()
        return {"lookups": int(stats[0]), "sectors": int(stats[1]), "table_probes": int(stats[2])}

    def time_resident(dbx, off, cont, n, o, qf, with_clocks=False):
        for _ in range(args.warmup):
            dbx.query_device(off, cont, n, o.data_ptr(), flags=qf, stream=stream)
        barrier()
        smp = None
        if with_clocks:
            smp = ClockSampler(local); smp.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            dbx.query_device(off, cont, n, o.data_ptr(), flags=qf, stream=stream)
        e1.record()
        barrier()
        t_ms = max_over_ranks(e0.elapsed_time(e1))
        if smp:
            smp.stop_flag = True; smp.join()
        return t_ms / args.steps, (smp.summary() if smp else None)

    # table-only first (its result words are the reference for every other path), then the product default
    c_tab = counted(db, d_off, d_cont, NR, out, CCB_NO_ANCHOR | WK)
    out_h = out.cpu().numpy().view(np.uint16).reshape(-1, 5).copy()
    classified = int((out_h[:, 1] != 0).sum())
    lookups = c_tab["lookups"]
    total_lookups = sum_over_ranks(lookups)
    c_anc, anchor_same = None, None
    if ainfo is not None:
        c_anc = counted(db, d_off, d_cont, NR, out, WK)
        anchor_same = bool(np.array_equal(out.cpu().numpy().view(np.uint16).reshape(-1, 5), out_h))
    head_anchor = ainfo is not None and not args.warp and not args.table_only_headline
    QF_HEAD = WK if head_anchor else (CCB_NO_ANCHOR | WK)
    ms, clocks = time_resident(db, d_off, d_cont, NR, out, QF_HEAD, True)
    ms_other = None
    if ainfo is not None and not args.warp:
        ms_other, _ = time_resident(db, d_off, d_cont, NR, out, (CCB_NO_ANCHOR | WK) if head_anchor else WK)
    value = total_lookups / (ms * 1e-3)

    if args.kernel_only:
        if rank == 0:
            emit(json.dumps({"kernel_only": True, "ms_per_step": ms, "lookups_per_s": value, "anchor": head_anchor, "ms_other_leg": ms_other}))
        db.device_free(d_off); db.device_free(d_cont); db.close()
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- random 32-B gather probe over the same table (what one L2 miss per probe can reach) -------
    gather = None
    if not args.no_gather and rank == 0:
        npr = 1 << 30
        res = {}
        for nbytes in (32, 64):
            db.bench_gather(npr, nbytes, stream); torch.cuda.synchronize()
            g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            g0.record(); db.bench_gather(npr, nbytes, stream); g1.record(); torch.cuda.synchronize()
            res[nbytes] = npr * nbytes / (g0.elapsed_time(g1) * 1e-3) / 1e9
        gather = res
    barrier()

    # ---- end to end through the C ABI with host buffers ------------------------------------------
    S = args.slots
    per = (NR + S - 1) // S
    off_h = _dev_to_host(cudart, np, torch, d_off, NR + 1, torch.int32, np.uint32)
    cont_h = _dev_to_host(cudart, np, torch, d_cont, n_cont, torch.int16, np.uint16)
    bounds = [(s * per, min(NR, (s + 1) * per)) for s in range(S)]
    max_c = max(int(off_h[b]) - int(off_h[a]) for a, b in bounds)
    db.malloc(per, max_c, False)
    sizes = []
    for s, (a, b) in enumerate(bounds):
        o, c, f, _ = db.buffers(s)
        o[: b - a + 1] = off_h[a:b + 1] - off_h[a]
        nc = int(off_h[b]) - int(off_h[a])
        c[:nc] = cont_h[int(off_h[a]):int(off_h[b])]
        sizes.append((b - a, nc))
    h2d = sum((n + 1) * 4 + nc * 2 for n, nc in sizes); d2h = sum(n * 10 for n, _ in sizes)
    del off_h, cont_h

    # rolling pipeline, as the CLI drives it: a slot is waited for only when it is about to be
    # refilled, so the H2D of one batch overlaps the kernel of the previous one across steps too.
    busy = [False] * S

    def time_e2e(qf):
        def e2e_step():
            for s, (n, nc) in enumerate(sizes):
                if busy[s]:
                    db.waitForBatch(s)
                db.readyBatch(s, n, nc); db.queryBatch(s, flags=qf); busy[s] = True

        def e2e_drain():
            for s in range(S):
                if busy[s]:
                    db.waitForBatch(s); busy[s] = False

        for _ in range(args.warmup):
            e2e_step()
        e2e_drain()
        barrier()
        t0 = time.perf_counter()
        marks = []
        for _ in range(args.steps):
            e2e_step(); marks.append(time.perf_counter())
        e2e_drain()
        torch.cuda.synchronize()
        t = max_over_ranks(time.perf_counter() - t0)
        barrier()
        # host-side pace of the pipeline, step by step (a one-off stall shows up as one long step)
        pace = [round(1e3 * (b - a), 2) for a, b in zip([t0] + marks[:-1], marks)]
        fin = np.concatenate([db.buffers(s)[2][:n] for s, (n, _) in enumerate(sizes)])
        return t, bool(np.array_equal(fin, out_h)), pace

    t_e2e, same, pace = time_e2e(QF_HEAD)
    e2e_other = None
    if ms_other is not None:
        t2, same2, _ = time_e2e((CCB_NO_ANCHOR | WK) if head_anchor else WK)
        e2e_other = {"value": total_lookups * args.steps / t2, "unit": "lookups/s", "ms_per_step": 1e3 * t2 / args.steps,
                     "results_equal_resident_path": same2}
    db.freeBatchMemory()

    def roofline(cnt, ms_, kernel, traffic_key):
        ach = cnt["lookups"] * 32 / (ms_ * 1e-3) / 1e9                      # 32 B per k-mer answered
        req = cnt["sectors"] * 32 / (ms_ * 1e-3) / 1e9                      # bytes the kernel asked the memory system for
        tr = traffic_tab.get(traffic_key) if default_workload else None
        dram = (tr / (ms_ * 1e-3)) / 1e9 if tr else None
        return {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": tr,
                "bytes_per_lookup": 32, "requested_gbs": req, "requested_frac": req / peak,
                "dram_gbs": dram, "dram_frac": (dram / peak) if dram else None, "peak_source": peak_src, "kernel": kernel,
                "table_probes_per_lookup": cnt["table_probes"] / max(cnt["lookups"], 1),
                "sectors_per_lookup": cnt["sectors"] / max(cnt["lookups"], 1)}

    kname_tab = "ccb::k_query" if args.warp else "ccb::k_query_block"
    ms_tab = ms_other if head_anchor else ms
    ms_anc = ms if head_anchor else ms_other
    roof_tab = roofline(c_tab, ms_tab, kname_tab, "k_query_dram_bytes_per_launch" if args.warp else "k_query_block_dram_bytes_per_launch") if ms_tab else None
    roof_anc = roofline(c_anc, ms_anc, "ccb::k_anchor_scan + ccb::k_query_block<PRE>", "k_query_block_anchor_dram_bytes_per_launch") if (c_anc and ms_anc) else None
    if roof_tab and gather:
        roof_tab.update({"gather32_gbs": gather[32], "gather64_gbs": gather[64], "frac_of_gather32": roof_tab["achieved"] / gather[32]})

    # ---- workloads C4 (same database, long reads with indels) and C3 (light database, pairs) -----------------
    workloads = {}

    def workload_leg(dbx, w, n_reads, kname, seed):
        o, c, nc = dbx.synth2_reads_packed(w, seed, DB_SEED, rank * n_reads, n_reads, G, LG)
        res = torch.zeros(n_reads * 5, dtype=torch.int16, device="cuda")
        cnt = counted(dbx, o, c, n_reads, res, 0)
        ref = res.cpu().numpy().view(np.uint16).copy()
        cnt_t = counted(dbx, o, c, n_reads, res, CCB_NO_ANCHOR)
        same_ = bool(np.array_equal(res.cpu().numpy().view(np.uint16), ref))
        m, _ = time_resident(dbx, o, c, n_reads, res, 0)
        m_t, _ = time_resident(dbx, o, c, n_reads, res, CCB_NO_ANCHOR)
        tl = sum_over_ranks(cnt["lookups"])
        leg = {"value": tl / (m * 1e-3), "unit": "lookups/s", "ms_per_step": m, "reads_per_s": n_reads * world / (m * 1e-3),
               "reads_per_gpu": n_reads, "lookups_per_step_per_gpu": cnt["lookups"], "sectors_per_lookup": cnt["sectors"] / max(cnt["lookups"], 1),
               "table_probes_per_lookup": cnt["table_probes"] / max(cnt["lookups"], 1), "kernel": kname,
               "roofline_frac": cnt["lookups"] * 32 / (m * 1e-3) / 1e9 / peak,
               "reads_classified_frac": float((ref.reshape(-1, 5)[:, 1] != 0).mean()),
               "table_only": {"value": tl / (m_t * 1e-3), "ms_per_step": m_t, "sectors_per_lookup": cnt_t["sectors"] / max(cnt_t["lookups"], 1),
                              "results_equal": same_}}
        dbx.device_free(o); dbx.device_free(c)
        return leg

    W4 = Workload(2, args.c4_read_len, 0, 30, 10, 10)
    if not args.no_workloads:
        workloads["c4"] = dict(workload_leg(db, W4, args.c4_reads, "ccb::k_query (long reads: warp per read, table only)", 0x5EED0004),
                               workload=f"long reads: {args.c4_reads} x {args.c4_read_len} bp per GPU, 5 % errors (3 % sub, 1 % ins, 1 % del), same k={k} database")
    db.close(); del db
    torch.cuda.empty_cache()

    # ---- the same reads against the table SHARDED over the ranks (row E2) ----------------------------------------
    sharded = None
    if world > 1 and not args.no_sharded_leg:
        sharded = sharded_leg(args, rank, world, local, d_off, d_cont, n_cont, out_h, total_lookups, barrier, max_over_ranks, sum_over_ranks)

    W3 = Workload(1, L, 400, 10, 0, 0)
    if not args.no_workloads:
        db3 = CuClarkDB(1, 27, 1, NL, device=local, max_hits=23, load_factor=args.load_factor)
        t0 = time.time(); n3 = db3.synth_db_light(DB_SEED, G, LG, NL, 4)
        try:
            db3.anchor_synth(DB_SEED, G, LG, NL)
        except Exception:
            pass
        t3 = time.time() - t0
        workloads["c3"] = dict(workload_leg(db3, W3, NR, "ccb::k_query_block<anchor> (two parts per read)", 0x5EED0003),
                               workload=f"cuCLARK-l light k=27 gap 4 ({n3} sampled 27-mers), {NR} pairs 2 x {L} bp (insert 400, opposite strands, joined by N) per GPU",
                               table_gb=db3.db_info().bytes / 1e9, db_build_s=t3)
        db3.close(); del db3
        torch.cuda.empty_cache()

    # ---- file -> CSV through bin/cuCLARK (rows A1/A2/A7/F2) ---------------------------------------------------------
    cli = None
    if rank == 0 and world == 1 and not args.no_cli_e2e:
        try:
            cli = cli_e2e_leg(args, local)
        except Exception as e:
            cli = {"unavailable": str(e)[:300]}

    # ---- CPU baseline + bit-exact checks on bounded samples (rank 0, N=1 only; after every timed GPU leg) ----------
    cpu_baseline, parity, parity_w = None, None, {}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        import oracle_py
        cores = os.cpu_count() or 1
        SG, SR = args.sample_genomes, args.sample_reads
        t0 = time.time()
        odb = oracle_py.OracleDb.synth(DB_SEED, SG, LG, SG, k, args.oracle_htsize)
        t_build = time.time() - t0
        sdb = CuClarkDB(1, k, 1, SG, device=local, load_factor=args.load_factor)
        sdb.synth_db(DB_SEED, SG, LG, SG)
        sdb.anchor_synth(DB_SEED, SG, LG, SG)

        def sample(so, sc, snc, n):
            o = torch.zeros(n * 5, dtype=torch.int16, device="cuda")
            sdb.query_device(so, sc, n, o.data_ptr(), stream=stream)
            torch.cuda.synchronize()
            g = o.cpu().numpy().view(np.uint16).reshape(-1, 5)
            oh = _dev_to_host(cudart, np, torch, so, n + 1, torch.int32, np.uint32)
            ch = _dev_to_host(cudart, np, torch, sc, snc, torch.int16, np.uint16)
            sdb.device_free(so); sdb.device_free(sc)
            return g, oh, ch

        got, off_s, cont_s = sample(*sdb.synth_reads_packed(READS_SEED, DB_SEED, 0, SR, L, SG, LG), SR)
        odb.query_packed(k, off_s[:2001], cont_s, threads=cores)               # warm-up (builds bucket pointers)
        t0 = time.time(); want, _, nl = odb.query_packed(k, off_s, cont_s, threads=cores); t_q = time.time() - t0
        one = min(SR, 20000)
        t0 = time.time(); _, _, nl1 = odb.query_packed(k, off_s[:one + 1], cont_s, threads=1); t_q1 = time.time() - t0
        parity = "bit-exact" if np.array_equal(got, want) else "MISMATCH"
        smp = (f"DB = first {SG} of {G} genomes x {LG} bp ({odb.n} {k}-mers, HTSIZE {args.oracle_htsize}, built in {t_build:.1f} s on 1 core); "
               f"{SR} x {L} bp reads drawn from them; lookup+vote loop only")
        cpu_baseline = {"value": nl / t_q, "unit": "lookups/s", "cores": cores, "kind": "port", "sample": smp,
                        "reads_per_s": SR / t_q, "db_build_kmers_per_s": odb.n / t_build,
                        "value_1core": nl1 / t_q1, "sample_1core": f"the first {one} reads of the same sample on one thread"}
        if not args.no_workloads:
            n4 = max(200, min(2000, SR // 100))
            g4, o4, c4 = sample(*sdb.synth2_reads_packed(W4, 0x5EED0004, DB_SEED, 0, n4, SG, LG), n4)
            w4, _, _ = odb.query_packed(k, o4, c4, threads=cores)
            parity_w["c4"] = "bit-exact" if np.array_equal(g4, w4) else "MISMATCH"
        sdb.close(); del odb
        if not args.no_workloads:
            odb3 = oracle_py.OracleDb.synth(DB_SEED, SG, LG, SG, 27, 57777779, light_gap=4)
            sdb = CuClarkDB(1, 27, 1, SG, device=local, max_hits=23)
            sdb.synth_db_light(DB_SEED, SG, LG, SG, 4)
            sdb.anchor_synth(DB_SEED, SG, LG, SG)
            n3s = min(SR, 50000)
            g3, o3, c3 = sample(*sdb.synth2_reads_packed(W3, 0x5EED0003, DB_SEED, 0, n3s, SG, LG), n3s)
            w3, _, _ = odb3.query_packed(27, o3, c3, threads=cores)
            parity_w["c3"] = "bit-exact" if np.array_equal(g3, w3) else "MISMATCH"
            sdb.close(); del odb3
        torch.cuda.empty_cache()
    for name, p_ in parity_w.items():
        workloads[name]["parity_sample"] = p_

    if rank == 0:
        head_cnt = c_anc if head_anchor else c_tab
        head_roof = roof_anc if head_anchor else roof_tab
        other = None
        if ms_other is not None:
            o_cnt, o_roof = (c_tab, roof_tab) if head_anchor else (c_anc, roof_anc)
            other = {"what": ("table only: every k-mer probed in the k-mer table (CCB_NO_ANCHOR), same reads, same result words" if head_anchor else
                              "anchor index (target text + sampled k-mer positions) in front of the same table"),
                     "value": total_lookups / (ms_other * 1e-3), "unit": "lookups/s", "ms_per_step": ms_other,
                     "reads_per_s": NR * world / (ms_other * 1e-3), "e2e": e2e_other, "roofline": o_roof,
                     "table_probes_per_s": o_cnt["table_probes"] * world / (ms_other * 1e-3)}
        line = {
            "metric": "kmer_lookups_per_s", "value": value, "unit": "lookups/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "reads_per_s": NR * world / (ms * 1e-3),
            "table_probes_per_s": head_cnt["table_probes"] * world / (ms * 1e-3),
            "config": {"workload": f"cuCLARK full k={k}, synthetic {G} x {LG} bp DB ({n_kmers} {k}-mers), {NR} x {L} bp SE reads per GPU",
                       "path": ("product default: anchor index (target text + sampled k-mer positions) in front of the k-mer table; a lookup = one k-mer of a read "
                                "answered (from the text when the read matches it base by base, else by a table probe); results word-identical to the table-only path"
                                if head_anchor else "table only: every k-mer probed in the k-mer table"),
                       "parallelism": f"replicated table, reads partitioned x{world}",
                       "l2": "inputs larger than L2 (table %.1f GB, random access)" % (info.bytes / 1e9),
                       "table_gb": info.bytes / 1e9, "load_factor": args.load_factor, "db_build_s": t_db,
                       "anchor_index": ({"gb": ainfo.bytes / 1e9, "entries": ainfo.entries, "dropped": ainfo.dropped, "build_s": t_anchor,
                                         "results_equal_table_only_path": anchor_same} if ainfo is not None else {"unavailable": anchor_why}),
                       "max_displacement": info.max_displacement, "lookups_per_step_per_gpu": lookups,
                       "sectors_per_lookup": head_cnt["sectors"] / max(head_cnt["lookups"], 1), "reads_classified_frac": classified / NR},
            "clocks": clocks,
            "e2e": {"value": total_lookups * args.steps / t_e2e, "unit": "lookups/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "reads_per_s": NR * world * args.steps / t_e2e, "ms_per_step": 1e3 * t_e2e / args.steps,
                    "slots": S, "results_equal_resident_path": same, "host_step_ms": pace,
                    "host_feed_gbs": (h2d + d2h) * world * args.steps / t_e2e / 1e9},
            # per step of the timed region: [k_anchor_scan,] k_query_block, k_query (handed-back reads), k_query_dense
            "gpu_launches": (2 if args.warp else 4 if head_anchor else 3) * args.steps,
            "roofline": head_roof,
        }
        if other:
            line["table_only" if head_anchor else "anchor_index"] = other
        if workloads:
            line["workloads"] = workloads
        if sharded:
            line["sharded"] = sharded
        if cli:
            line["cli_e2e"] = cli
        if cpu_baseline:
            line["cpu_baseline"] = cpu_baseline
            line["parity_sample"] = parity
        emit(json.dumps(line))
    cudart.cudaFree(C.c_void_p(d_off)); cudart.cudaFree(C.c_void_p(d_cont))
    if world > 1:
        dist.destroy_process_group()


def sharded_leg(args, rank, world, local, d_off, d_cont, n_cont, out_h, total_lookups, barrier, max_over_ranks, sum_over_ranks):
    """Row E2 inside the default multi-GPU line: the SAME reads against the table cut into `world` shards (shard g on
    rank g), both exchanges (kernels storing into peer inboxes over CUDA IPC; NCCL all-to-all), result words compared with
    the replicated path's on every rank."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from jn_cuclark_b200 import CuClarkDB
    from jn_cuclark_b200.sharded import ShardedClassifier, PeerShardedClassifier, PipelinedPeerClassifier, region_size
    k, G, LG, NR, L = args.k, args.genomes, args.genome_len, args.reads, args.read_len
    cudart = C.CDLL("libcudart.so")
    sdb = CuClarkDB(1, k, 1, min(G, 65534), device=local, load_factor=args.load_factor)
    sdb.shard_config(world, rank)
    t0 = time.time(); sdb.synth_db(DB_SEED, G, LG, min(G, 65534)); t_db = time.time() - t0
    info = sdb.db_info()
    cap = region_size(NR, L - k + 1, world)
    NCH = 4                                                            # chunks of the pipelined variant, two in flight
    out = torch.zeros(NR * 5, dtype=torch.int16, device="cuda")
    stats = torch.zeros(2, dtype=torch.int64, device="cuda")
    leg = {"what": f"table sharded x{world} by hash range (shard {info.bytes / 1e9:.1f} GB per GPU), the same reads; one-pass route "
                   "(k_route_block) -> owner-side probe (k_probe_inbox / k_probe_list) -> block-wise vote (k_vote_block)",
           "shard_gb": info.bytes / 1e9, "shard_entries": info.entries, "db_build_s": t_db, "region_kmers": cap,
           "exchange_variants": {"peer": "kernels store k-mers / labels straight into peer memory (CUDA IPC over NVLink), one batch at a time",
                                 "peer_pipelined": f"the same, the batch cut into {NCH} chunks through two pipelines on two streams: one chunk's probe runs beside the next one's route",
                                 "nccl": "NCCL all-to-all of the k-mers and of the labels (one host sync per batch for the sizes)"},
           "exchanges": {}}
    per_chunk = (NR + NCH - 1) // NCH
    for exch in ("peer_pipelined", "peer", "nccl"):
        if exch == "peer_pipelined":
            sc = PipelinedPeerClassifier(sdb, region_size(per_chunk, L - k + 1, world), per_chunk, n_pipes=2)
        else:
            sc = PeerShardedClassifier(sdb, cap, NR) if exch == "peer" else ShardedClassifier(sdb, cap, NR)
        stats.zero_(); out.zero_()
        sc.classify(d_off, d_cont, NR, out, stats=stats, n_containers=n_cont)
        torch.cuda.synchronize()
        same = bool(np.array_equal(out.cpu().numpy().view(np.uint16).reshape(-1, 5), out_h))
        overflow = sc.overflowed() if exch != "nccl" else False
        ok = torch.tensor([1 if (same and not overflow) else 0], dtype=torch.int32, device="cuda")
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        probes_here = int(stats[0])
        for _ in range(args.warmup):
            sc.classify(d_off, d_cont, NR, out, n_containers=n_cont)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            sc.classify(d_off, d_cont, NR, out, n_containers=n_cont)
        e1.record(); barrier()
        ms = max_over_ranks(e0.elapsed_time(e1)) / args.steps
        res = {"value": total_lookups / (ms * 1e-3), "unit": "lookups/s", "ms_per_step": ms, "reads_per_s": NR * world / (ms * 1e-3),
               "results_equal_replicated_path": bool(int(ok[0])), "probes_served_rank0": probes_here,
               # 8 B per k-mer out and 2 B per label back, for the (world-1)/world of the k-mers another GPU owns
               "nvlink_bytes_per_step_per_gpu": int(probes_here * (world - 1) / world * 10)}
        if exch == "peer":
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
            barrier()
            ev[0].record(); sc.stage_route(d_off, d_cont, NR, n_cont); ev[1].record(); barrier()
            ev[2].record(); sc.stage_probe(); ev[3].record(); barrier()
            ev[4].record(); sc.stage_vote(NR, out); ev[5].record(); torch.cuda.synchronize(); barrier()
            res["phases_rank0_ms"] = {"route": ev[0].elapsed_time(ev[1]), "probe": ev[2].elapsed_time(ev[3]), "vote": ev[4].elapsed_time(ev[5])}
            res["roofline_probe"] = {"bound": "hbm", "achieved": probes_here * 32 / (ev[2].elapsed_time(ev[3]) * 1e-3) / 1e9, "unit": "GB/s",
                                     "kernel": "ccb::k_probe_inbox", "bytes_per_lookup": 32}
        leg["exchanges"][exch] = res
        sc.close()
        torch.cuda.empty_cache()
    best = min(leg["exchanges"].values(), key=lambda r: r["ms_per_step"])
    leg.update({"value": best["value"], "unit": "lookups/s", "ms_per_step": best["ms_per_step"],
                "results_equal_replicated_path": all(r["results_equal_replicated_path"] for r in leg["exchanges"].values())})
    sdb.close()
    return leg


def cli_e2e_leg(args, local):
    """File -> CSV through the drop-in binary (rows A1/A2/A7/F2): synthetic genomes as FASTA targets, the database built by
    bin/cuCLARK itself on the first run, reads as one FASTQ file in the page cache, then `bin/cuCLARK -n <cores>`."""
    import shutil
    import tempfile
    import re
    import numpy as np
    from jn_cuclark_b200 import CuClarkDB
    G, LG, NR, L, k = args.cli_genomes, args.genome_len, args.cli_reads, args.read_len, args.k
    cores = min(os.cpu_count() or 1, 32)
    base = "/dev/shm" if os.path.isdir("/dev/shm") and shutil.disk_usage("/dev/shm").free > 12 * (1 << 30) else None
    d = tempfile.mkdtemp(prefix="ccb_cli_", dir=base)
    try:
        t0 = time.time()
        lut = np.frombuffer(b"TGCA", dtype=np.uint8)

        def sm64(z):
            z = z + np.uint64(0x9E3779B97F4A7C15)
            z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
            z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
            return z ^ (z >> np.uint64(31))

        with np.errstate(over="ignore"), open(os.path.join(d, "targets.txt"), "w") as tf:
            nw = (LG + 31) // 32
            sh = (np.arange(32, dtype=np.uint64) * np.uint64(2))[None, :]
            for g in range(G):
                gs = sm64(np.uint64(DB_SEED) ^ (np.uint64(g + 1) * np.uint64(0xD1342543DE82EF95)))
                w = sm64(gs + np.arange(nw, dtype=np.uint64) * np.uint64(0x9E3779B97F4A7C15))
                seq = lut[((w[:, None] >> sh) & np.uint64(3)).astype(np.uint8).reshape(-1)[:LG]]      # include/ccb_synth.h ccb_genome_code
                p = os.path.join(d, f"g{g}.fa")
                rows = seq[: LG // 80 * 80].reshape(-1, 80)
                with open(p, "wb") as f:
                    f.write(b">g%d\n" % g)
                    f.write(np.concatenate([rows, np.full((rows.shape[0], 1), 10, np.uint8)], axis=1).tobytes())
                    f.write(seq[LG // 80 * 80:].tobytes() + b"\n")
                tf.write(f"{p}\tT{g}\n")
        # reads of the same generator as the headline workload, as fixed-width FASTQ records
        gen = CuClarkDB(1, k, 1, 1, device=local)
        rec = np.empty((NR, 15 + 2 * L), dtype=np.uint8)
        CH = 2_000_000
        for s0 in range(0, NR, CH):
            n = min(CH, NR - s0)
            rec[s0:s0 + n, 11:11 + L] = gen.synth_reads_text(READS_SEED, DB_SEED, s0, n, L, G, LG)[:, :L]
        gen.close()
        ids = np.arange(NR)
        rec[:, 0] = ord("@"); rec[:, 1] = ord("r")
        for j in range(8):
            rec[:, 9 - j] = 48 + (ids // 10 ** j) % 10
        rec[:, 10] = 10; rec[:, 11 + L] = 10; rec[:, 12 + L] = ord("+"); rec[:, 13 + L] = 10
        rec[:, 14 + L:14 + 2 * L] = ord("I"); rec[:, 14 + 2 * L] = 10
        fq = os.path.join(d, "reads.fq")
        rec.tofile(fq)
        fq_bytes = rec.nbytes
        del rec
        t_gen = time.time() - t0
        os.makedirs(os.path.join(d, "db"))
        exe = os.path.join(ROOT, "bin", "cuCLARK")
        runs = []
        for i in range(3):                      # the first run builds the database; the fastest of the others is reported
            t0 = time.time()
            r = subprocess.run([exe, "-T", os.path.join(d, "targets.txt"), "-D", os.path.join(d, "db") + "/", "-O", fq, "-R", os.path.join(d, f"res{i}"),
                                "-n", str(cores), "--verbose"], capture_output=True, text=True, env=dict(os.environ, CUDA_VISIBLE_DEVICES=str(local)))
            wall = time.time() - t0
            if r.returncode != 0:
                return {"unavailable": "bin/cuCLARK failed: " + r.stderr[-300:].replace("\n", " ")}
            m = re.search(r"host pipeline: index ([0-9.]+) s, slots ([0-9.]+) s, total ([0-9.]+) s", r.stderr)
            runs.append({"wall_s": wall, "classify_s": float(m.group(3)) if m else None, "index_s": float(m.group(1)) if m else None,
                         "built_db": "successfully stored" in r.stderr, "anchor": "Anchor index: " in r.stderr,
                         "line": next((l.strip() for l in r.stderr.splitlines() if "host pipeline" in l), None)})
        timed = [x for x in runs[1:] if x["classify_s"]]
        best = min(timed, key=lambda x: x["classify_s"])
        same_csv = open(os.path.join(d, "res1.csv"), "rb").read() == open(os.path.join(d, "res2.csv"), "rb").read()
        # SURVEY D4 baseline 3: the reference's own cuCLARK build for this GPU (oracle/_ref/cuCLARK, unmodified sources, sm_100)
        # on the same files; it only runs with a number of targets that is a multiple of 32 (tests/golden/ref_gpu/README.md)
        ref_gpu = None
        ref_exe = os.path.join(ROOT, "oracle", "_ref", "cuCLARK")
        if not os.path.exists(ref_exe):
            ref_gpu = {"unavailable": "oracle/_ref/cuCLARK not built"}
        elif G % 32:
            ref_gpu = {"unavailable": f"{G} targets: the reference kernels deadlock unless the number of targets is a multiple of 32"}
        else:
            t0 = time.time()
            try:
                r = subprocess.run([ref_exe, "-T", os.path.join(d, "targets.txt"), "-D", os.path.join(d, "db") + "/", "-O", fq, "-R", os.path.join(d, "ref"),
                                    "-n", str(cores), "-b", "32"], capture_output=True, text=True, timeout=600, env=dict(os.environ, CUDA_VISIBLE_DEVICES=str(local)))
                m = re.search(r"Done in ([0-9.]+)s \((\d+) reads/min", r.stderr.replace("\n", " "))
                if r.returncode == 0 and m:
                    ref_gpu = {"what": f"oracle/_ref/cuCLARK -n {cores} -b 32 on the same targets, database files and FASTQ; its own 'Done in' line",
                               "classify_s": float(m.group(1)), "reads_per_s": int(m.group(2)) / 60.0, "process_wall_s": time.time() - t0,
                               "csv_identical_to_ours": open(os.path.join(d, "ref.csv"), "rb").read() == open(os.path.join(d, "res1.csv"), "rb").read()}
                else:
                    ref_gpu = {"unavailable": "reference binary failed: " + r.stderr[-200:].replace("\n", " ")}
            except subprocess.TimeoutExpired:
                ref_gpu = {"unavailable": "reference binary did not finish in 600 s"}
        return {"reference_gpu": ref_gpu,"what": f"bin/cuCLARK -n {cores}: {NR} x {L} bp reads as one FASTQ file ({fq_bytes / 1e9:.2f} GB, page cache) -> CSV; database of "
                        f"{G} x {LG} bp genomes built by the binary itself on the first run; anchor index on (product default)",
                "reads_per_s": NR / best["classify_s"], "classify_s": best["classify_s"], "index_s": best["index_s"], "threads": cores,
                "fastq_gbs": fq_bytes / best["classify_s"] / 1e9, "process_wall_s": best["wall_s"], "first_run_with_db_build_wall_s": runs[0]["wall_s"],
                "anchor_index": best["anchor"], "csv_identical_across_runs": same_csv, "generate_s": t_gen, "pipeline": best["line"]}
    finally:
        shutil.rmtree(d, ignore_errors=True)


def run_sharded(args):
    """Row E2 on its own (`--sharded`): the table sharded over the N GPUs -- for databases that do not fit one GPU
    (`--genomes 3000` = 12 G 31-mers = 192 GB of table).  Same metric; `value` times device-resident batches, `e2e` adds
    the H2D/D2H copies."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from jn_cuclark_b200 import CuClarkDB
    from jn_cuclark_b200.sharded import ShardedClassifier, PeerShardedClassifier, region_size
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    cudart = C.CDLL("libcudart.so")
    k, G, LG, NR, L = args.k, args.genomes, args.genome_len, args.reads, args.read_len
    db = CuClarkDB(1, k, 1, min(G, 65534), device=local, load_factor=args.load_factor)
    db.shard_config(world, rank)
    t0 = time.time(); n_kmers = db.synth_db(DB_SEED, G, LG, min(G, 65534)); t_db = time.time() - t0
    info = db.db_info()
    nch = max(1, args.chunks)
    per = (NR + nch - 1) // nch
    batches = []
    for c in range(nch):
        n = min(per, NR - c * per)
        o, cc, nc = db.synth_reads_packed(READS_SEED, DB_SEED, rank * NR + c * per, n, L, G, LG)
        batches.append((o, cc, nc, n, torch.zeros(n * 5, dtype=torch.int16, device="cuda")))
    peer = args.exchange == "peer"
    cap = region_size(per, L - k + 1, world)
    sc = PeerShardedClassifier(db, cap, per) if peer else ShardedClassifier(db, cap, per)
    stats = torch.zeros(2, dtype=torch.int64, device="cuda")
    overlap = bool(args.overlap and peer and nch >= 2)
    if overlap:
        # second pipeline: own inboxes / label boxes / counts, own barrier group (its all-reduces run on another stream)
        g2 = dist.new_group(list(range(world))) if world > 1 else None
        sc2 = PeerShardedClassifier(db, cap, per, group=g2)
        scs, streams = [sc, sc2], [torch.cuda.Stream(), torch.cuda.Stream()]

    def step(st=None):
        if not overlap:
            for o, cc, nc, n, out in batches:
                sc.classify(o, cc, n, out, stats=st, n_containers=nc)
            return
        cur = torch.cuda.current_stream()
        for s_ in streams:
            s_.wait_stream(cur)
        for i, (o, cc, nc, n, out) in enumerate(batches):
            with torch.cuda.stream(streams[i & 1]):
                scs[i & 1].classify(o, cc, n, out, stats=st, n_containers=nc)
        for s_ in streams:
            cur.wait_stream(s_)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX); return float(t[0])

    step(stats); torch.cuda.synchronize()
    lookups_here = int(stats[0])                                   # probes served by this shard
    tl = torch.tensor([lookups_here], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(tl)
    total_lookups = int(tl[0])
    classified = int(sum(int((b[4].view(-1, 5)[:, 1] != 0).sum()) for b in batches))
    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local); sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record(); barrier()
    ms = allmax(e0.elapsed_time(e1))
    sampler.stop_flag = True; sampler.join()
    value = total_lookups * args.steps / (ms * 1e-3)
    phases, probe_ms, n_in = None, None, lookups_here // nch
    if peer:
        assert not sc.overflowed(), "inbox capacity exceeded"
        o, cc, nc, n, out = batches[0]
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
        barrier()
        ev[0].record(); sc.stage_route(o, cc, n, nc); ev[1].record(); barrier()
        ev[2].record(); sc.stage_probe(); ev[3].record(); barrier()
        ev[4].record(); sc.stage_vote(n, out); ev[5].record(); torch.cuda.synchronize(); barrier()
        phases = {"route_ms": ev[0].elapsed_time(ev[1]), "probe_ms": ev[2].elapsed_time(ev[3]),
                  "vote_ms": ev[4].elapsed_time(ev[5]), "reads": n}
        probe_ms = phases["probe_ms"]
    # end to end: packed batches start in pinned host memory, results end there
    host = []
    for o, cc, nc, n, out in batches:
        ho = torch.empty(n + 1, dtype=torch.int32).pin_memory(); hc = torch.empty(nc, dtype=torch.int16).pin_memory()
        cudart.cudaMemcpy(C.c_void_p(ho.data_ptr()), C.c_void_p(o), C.c_size_t((n + 1) * 4), 2)
        cudart.cudaMemcpy(C.c_void_p(hc.data_ptr()), C.c_void_p(cc), C.c_size_t(nc * 2), 2)
        host.append((ho, hc, torch.empty(n * 5, dtype=torch.int16).pin_memory(),
                     torch.empty(n + 1, dtype=torch.int32, device="cuda"), torch.empty(nc, dtype=torch.int16, device="cuda")))
    h2d = sum(h[0].numel() * 4 + h[1].numel() * 2 for h in host); d2h = sum(h[2].numel() * 2 for h in host)

    def e2e_step():
        for (o, cc, nc, n, out), (ho, hc, hr, do, dc) in zip(batches, host):
            do.copy_(ho, non_blocking=True); dc.copy_(hc, non_blocking=True)
            sc.classify(do.data_ptr(), dc.data_ptr(), n, out, n_containers=nc)
            hr.copy_(out, non_blocking=True)
        torch.cuda.synchronize()

    e2e_step(); barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    t_e2e = allmax(time.perf_counter() - t0)
    barrier()
    if rank == 0:
        peak, peak_src = measured_peak()
        roof = None
        if probe_ms:
            achieved = n_in * 32 / (probe_ms * 1e-3) / 1e9
            roof = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                    "peak_source": peak_src, "bytes_per_lookup": 32, "kernel": "ccb::k_probe_inbox", "probe_ms": probe_ms, "probes": n_in}
        emit(json.dumps({
            "metric": "kmer_lookups_per_s", "value": value, "unit": "lookups/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u64", "data": "synthetic", "reads_per_s": NR * world * args.steps / (ms * 1e-3),
            "config": {"workload": f"cuCLARK full k={k}, synthetic {G} x {LG} bp DB ({n_kmers} {k}-mers), {NR} x {L} bp SE reads per GPU",
                       "parallelism": f"table sharded x{world} by hash range, k-mer exchange = " + ("kernel stores into peer inboxes (CUDA IPC over NVLink)" if peer else "NCCL all-to-all") + f", {nch} exchange round(s) per step" + (", two rounds in flight on two streams" if overlap else ""),
                       "l2": "inputs larger than L2 (shard %.1f GB, random access)" % (info.bytes / 1e9),
                       "shard_gb": info.bytes / 1e9, "table_gb_all_shards": info.bytes * world / 1e9, "shard_entries": info.entries, "db_build_s": t_db,
                       "probes_per_step_rank0": lookups_here, "phases_rank0": phases, "reads_classified_frac_rank0": classified / NR},
            "clocks": sampler.summary(),
            "e2e": {"value": total_lookups * args.steps / t_e2e, "unit": "lookups/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "ms_per_step": 1e3 * t_e2e / args.steps},
            "gpu_launches": 10 * nch * args.steps,
            "roofline": roof}))
    if overlap:
        sc2.close()
    sc.close()
    for o, cc, nc, n, out in batches:
        db.device_free(o); db.device_free(cc)
    db.close()
    if world > 1:
        dist.destroy_process_group()


_REAL_STDOUT = None


def emit(text):
    """The JSON line goes to the real stdout; everything else a library prints (NCCL's version
    banner, for one) was moved to stderr by quiet_stdout()."""
    out = _REAL_STDOUT or sys.stdout
    out.write(text + "\n"); out.flush()


def quiet_stdout():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


if __name__ == "__main__":
    quiet_stdout()
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    elif a.sharded:
        run_sharded(a)
    else:
        run_b200(a)
