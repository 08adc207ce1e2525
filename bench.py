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
    ap.add_argument("--no-gather", action="store_true")
    ap.add_argument("--warp", action="store_true", help="warp-per-read kernel for every read (CCB_WARP_KERNEL)")
    ap.add_argument("--anchor", action="store_true",
                    help="build the anchor index (target sequences + sampled k-mer positions) in front of the table")
    ap.add_argument("--no-anchor-leg", action="store_true", help="skip the second leg (anchor index in front of the table)")
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


def run_b200(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from jn_cuclark_b200 import CuClarkDB, CCB_STATS

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    cudart = C.CDLL("libcudart.so")
    k, G, LG, NR, L = args.k, args.genomes, args.genome_len, args.reads, args.read_len
    stream = torch.cuda.current_stream().cuda_stream

    def d2d(dst, src, nbytes):
        cudart.cudaMemcpy(C.c_void_p(dst), C.c_void_p(src), C.c_size_t(nbytes), 3)

    # ---- CPU baseline + bit-exact check on a bounded sample (rank 0, N=1 only) --------------
    cpu_baseline, parity = None, None
    if rank == 0 and world == 1 and not args.no_cpu_baseline and not args.kernel_only:
        import oracle_py
        cores = os.cpu_count() or 1
        SG, SR = args.sample_genomes, args.sample_reads
        t0 = time.time()
        odb = oracle_py.OracleDb.synth(DB_SEED, SG, LG, SG, k, HTSIZE_FULL)
        t_build = time.time() - t0
        sdb = CuClarkDB(1, k, 1, SG, device=local, load_factor=args.load_factor)
        sdb.synth_db(DB_SEED, SG, LG, SG)
        so, sc, snc = sdb.synth_reads_packed(READS_SEED, DB_SEED, 0, SR, L, SG, LG)
        off = torch.empty(SR + 1, dtype=torch.int32, device="cuda"); cont = torch.empty(snc, dtype=torch.int16, device="cuda")
        d2d(off.data_ptr(), so, (SR + 1) * 4); d2d(cont.data_ptr(), sc, snc * 2)
        out = torch.zeros(SR * 5, dtype=torch.int16, device="cuda")
        sdb.query_device(so, sc, SR, out.data_ptr(), stream=stream)
        torch.cuda.synchronize()
        got = out.cpu().numpy().view(np.uint16).reshape(-1, 5)
        off_h = off.cpu().numpy().view(np.uint32); cont_h = cont.cpu().numpy().view(np.uint16)
        odb.query_packed(k, off_h[:2001], cont_h, threads=cores)               # warm-up (builds bucket pointers)
        t0 = time.time()
        want, _, nl = odb.query_packed(k, off_h, cont_h, threads=cores)
        t_q = time.time() - t0
        parity = "bit-exact" if np.array_equal(got, want) else "MISMATCH"
        sample = (f"DB = first {SG} of {G} genomes x {LG} bp ({odb.n} {k}-mers, HTSIZE {HTSIZE_FULL}, built in {t_build:.1f} s on 1 core); "
                  f"{SR} x {L} bp reads drawn from them; lookup+vote loop only")
        cpu_baseline = {"value": nl / t_q, "unit": "lookups/s", "cores": cores, "kind": "port", "sample": sample,
                        "reads_per_s": SR / t_q, "db_build_kmers_per_s": odb.n / t_build}
        sdb.device_free(so); sdb.device_free(sc); sdb.close(); del odb, out, off, cont
        torch.cuda.empty_cache()

    # ---- the workload: replicated table per GPU, reads partitioned by rank -------------------
    db = CuClarkDB(1, k, args.slots, min(G, 65534), device=local, load_factor=args.load_factor)
    t0 = time.time()
    n_kmers = db.synth_db(DB_SEED, G, LG, min(G, 65534))
    t_db = time.time() - t0
    info = db.db_info()
    anchor = None
    if args.anchor:
        t0 = time.time(); anchor = db.anchor_synth(DB_SEED, G, LG, min(G, 65534)); t_anchor = time.time() - t0
    d_off, d_cont, n_cont = db.synth_reads_packed(READS_SEED, DB_SEED, rank * NR, NR, L, G, LG)
    out = torch.zeros(NR * 5, dtype=torch.int16, device="cuda")
    stats = torch.zeros(2, dtype=torch.int64, device="cuda")
    QF = 32 if args.warp else 0                                  # CCB_WARP_KERNEL; default: block kernel + warp kernel for the rest
    db.query_device(d_off, d_cont, NR, out.data_ptr(), 0, stats.data_ptr(), CCB_STATS | QF, stream)
    torch.cuda.synchronize()
    lookups, sectors = int(stats[0]), int(stats[1])
    classified = int((out.view(-1, 5)[:, 1] != 0).sum())

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

    # ---- kernel-only: batch resident in HBM ----------------------------------------------------
    def time_resident(qf, with_clocks):
        for _ in range(args.warmup):
            db.query_device(d_off, d_cont, NR, out.data_ptr(), flags=qf, stream=stream)
        barrier()
        smp = None
        if with_clocks:
            smp = ClockSampler(local); smp.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            db.query_device(d_off, d_cont, NR, out.data_ptr(), flags=qf, stream=stream)
        e1.record()
        barrier()
        t_ms = max_over_ranks(e0.elapsed_time(e1))
        if smp:
            smp.stop_flag = True; smp.join()
        return t_ms, (smp.summary() if smp else None)

    ms, clocks = time_resident(QF, True)
    total_lookups = lookups * world
    if world > 1:
        t = torch.tensor([lookups], dtype=torch.int64, device="cuda"); dist.all_reduce(t); total_lookups = int(t[0])
    value = total_lookups * args.steps / (ms * 1e-3)

    if args.kernel_only:
        if rank == 0:
            emit(json.dumps({"kernel_only": True, "ms_per_step": ms / args.steps, "lookups_per_s": value}))
        db.device_free(d_off); db.device_free(d_cont); db.close()
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- random 32-B gather probe over the same table (practical random-access roofline) -------
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

    # ---- end to end through the C ABI with host buffers ------------------------------------------
    S = args.slots
    per = (NR + S - 1) // S
    off_all = torch.empty(NR + 1, dtype=torch.int32, device="cuda"); d2d(off_all.data_ptr(), d_off, (NR + 1) * 4)
    cont_all = torch.empty(n_cont, dtype=torch.int16, device="cuda"); d2d(cont_all.data_ptr(), d_cont, n_cont * 2)
    off_h = off_all.cpu().numpy().view(np.uint32); cont_h = cont_all.cpu().numpy().view(np.uint16)
    del off_all, cont_all
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

    # rolling pipeline, as the CLI drives it: a slot is waited for only when it is about to be
    # refilled, so the H2D of one batch overlaps the kernel of the previous one across steps too.
    busy = [False] * S
    e2e_steps_ms = []

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
        e2e_steps_ms.append([round(1e3 * (b - a), 2) for a, b in zip([t0] + marks[:-1], marks)])
        # results of the e2e path equal the resident path
        fin = np.concatenate([db.buffers(s)[2][:n] for s, (n, _) in enumerate(sizes)])
        return t, bool(np.array_equal(fin, out_h))

    out_h = out.cpu().numpy().view(np.uint16).reshape(-1, 5).copy()
    t_e2e, same = time_e2e(QF)
    e2e_value = total_lookups * args.steps / t_e2e

    # ---- second leg: the anchor index in front of the same table --------------
    # Same reads, same table, same result words; windows that match the target text base by base are
    # answered without a table probe.  Reported beside the headline, which stays the table-only path.
    anchor_leg = None
    anchor_ok, anchor_why = False, None
    if not args.anchor and not args.no_anchor_leg:
        # the index may not fit (text over 4 Gbp, memory): then every rank skips the leg, the headline stands
        t0 = time.time()
        try:
            ainfo = db.anchor_synth(DB_SEED, G, LG, min(G, 65534)); anchor_ok = True
        except Exception as e:                                     # CcbError: message of the library
            anchor_why = str(e)[:200]
        t_anchor = time.time() - t0
        if world > 1:
            f = torch.tensor([1 if anchor_ok else 0], dtype=torch.int32, device="cuda"); dist.all_reduce(f, op=dist.ReduceOp.MIN)
            if anchor_ok and int(f[0]) == 0:
                db.anchor_drop(); anchor_ok, anchor_why = False, "another rank could not build the index"
        if not anchor_ok:
            anchor_leg = {"unavailable": anchor_why}
    if anchor_ok:
        AQF = 0                                                    # default path; the anchor is used because it exists
        stats.zero_(); out.zero_()
        db.query_device(d_off, d_cont, NR, out.data_ptr(), 0, stats.data_ptr(), CCB_STATS | AQF, stream)
        torch.cuda.synchronize()
        a_lookups, a_sectors = int(stats[0]), int(stats[1])
        a_same = bool(np.array_equal(out.cpu().numpy().view(np.uint16).reshape(-1, 5), out_h))
        a_ms, _ = time_resident(AQF, False)
        a_t_e2e, a_same_e2e = time_e2e(AQF)
        anchor_leg = {
            "what": "anchor index (target text + sampled k-mer positions) in front of the same table, block-cooperative "
                    "thread-per-read kernel; results compared word for word with the table-only path",
            "value": total_lookups * args.steps / (a_ms * 1e-3), "unit": "lookups/s", "ms_per_step": a_ms / args.steps,
            "reads_per_s": NR * world * args.steps / (a_ms * 1e-3),
            "e2e": {"value": total_lookups * args.steps / a_t_e2e, "unit": "lookups/s", "ms_per_step": 1e3 * a_t_e2e / args.steps,
                    "results_equal_table_only_path": a_same_e2e},
            "results_equal_table_only_path": a_same, "lookups_per_step_per_gpu": a_lookups,
            "sectors_per_lookup": a_sectors / max(a_lookups, 1),
            "index_gb": ainfo.bytes / 1e9, "index_entries": ainfo.entries, "index_dropped": ainfo.dropped, "index_build_s": t_anchor,
            "kernel": "ccb::k_query_block<anchor>",
        }
        if rank == 0:
            a_peak, _ = measured_peak()
            a_ach = (a_lookups * 32 / (a_ms * 1e-3 / args.steps)) / 1e9
            a_traffic = None
            try:
                a_traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get("k_query_block_anchor_dram_bytes_per_launch")
            except Exception:
                pass
            # same definition as the headline (32 B per lookup answered); most lookups are answered from the target
            # text here, so `traffic` (ncu, real DRAM bytes per launch) is what shows how close to HBM the kernel runs
            anchor_leg["roofline"] = {"bound": "hbm", "achieved": a_ach, "peak": a_peak, "unit": "GB/s", "frac": a_ach / a_peak,
                                      "traffic": a_traffic, "bytes_per_lookup": 32}

    if rank == 0:
        peak, peak_src = measured_peak()
        achieved = (lookups * 32 / (ms * 1e-3 / args.steps)) / 1e9          # this rank's kernel
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        default_workload = (G, LG, NR, L, k) == (1000, 4_000_000, 10_000_000, 150, 31)      # what the ncu capture was taken on
        if os.path.exists(tp) and default_workload:
            try:
                traffic = json.load(open(tp)).get("k_query_dram_bytes_per_launch" if args.warp else "k_query_block_dram_bytes_per_launch")
            except Exception:
                pass
        # real DRAM bytes per launch (ncu) over this run's launch time: how close to HBM the kernel runs on the bytes
        # the memory system moves (a 128-byte line per 32-byte sector), next to `frac` on the algorithmic bytes
        dram_gbs = (traffic / (ms * 1e-3 / args.steps)) / 1e9 if traffic else None
        line = {
            "metric": "kmer_lookups_per_s", "value": value, "unit": "lookups/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "reads_per_s": NR * world * args.steps / (ms * 1e-3),
            "config": {"workload": f"cuCLARK full k={k}, synthetic {G} x {LG} bp DB ({n_kmers} {k}-mers), {NR} x {L} bp SE reads per GPU",
                       "parallelism": f"replicated table, reads partitioned x{world}",
                       "l2": "inputs larger than L2 (table %.1f GB, random access)" % (info.bytes / 1e9),
                       "table_gb": info.bytes / 1e9, "load_factor": args.load_factor, "db_build_s": t_db,
                       "max_displacement": info.max_displacement, "lookups_per_step_per_gpu": lookups,
                       "sectors_per_lookup": sectors / max(lookups, 1), "reads_classified_frac": classified / NR},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "lookups/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "reads_per_s": NR * world * args.steps / t_e2e, "ms_per_step": 1e3 * t_e2e / args.steps,
                    "slots": S, "results_equal_resident_path": same, "host_step_ms": e2e_steps_ms[0]},
            "gpu_launches": (2 if args.warp else 3) * args.steps,      # k_query_block, k_query (handed-back reads), k_query_dense
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "dram_gbs": dram_gbs, "dram_frac": (dram_gbs / peak) if dram_gbs else None,
                         "peak_source": peak_src, "bytes_per_lookup": 32,
                         "kernel": "ccb::k_query" if args.warp else "ccb::k_query_block", "gather32_gbs": gather and gather[32], "gather64_gbs": gather and gather[64],
                         "frac_of_gather32": (achieved / gather[32]) if gather else None},
        }
        if anchor_leg:
            line["anchor_index"] = anchor_leg
        if cpu_baseline:
            line["cpu_baseline"] = cpu_baseline
            line["parity_sample"] = parity
        emit(json.dumps(line))
    db.device_free(d_off); db.device_free(d_cont)
    db.close()
    if world > 1:
        dist.destroy_process_group()


def run_sharded(args):
    """Row E2: table sharded over the N GPUs, each step = count -> route -> all-to-all -> probe -> all-to-all -> vote
    for every rank's own reads.  Same metric; `value` times device-resident batches, `e2e` adds the H2D/D2H copies."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from jn_cuclark_b200 import CuClarkDB
    from jn_cuclark_b200.sharded import ShardedClassifier, PeerShardedClassifier
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
    if peer:
        # region size: the batch's k-mers spread evenly over the owners, with 25 % slack (mix2k is uniform)
        cap = int(per * (L - k + 1) / world * 1.25) + 4096
        sc = PeerShardedClassifier(db, cap)
    else:
        sc = ShardedClassifier(db)
    stats = torch.zeros(2, dtype=torch.int64, device="cuda")
    overlap = bool(args.overlap and peer and nch >= 2)
    if overlap:
        # second pipeline: own inboxes / label boxes / counts, own barrier group (its all-reduces run on another stream)
        g2 = dist.new_group(list(range(world))) if world > 1 else None
        sc2 = PeerShardedClassifier(db, cap, group=g2)
        scs, streams = [sc, sc2], [torch.cuda.Stream(), torch.cuda.Stream()]

    def step(st=None):
        if not overlap:
            for o, cc, nc, n, out in batches:
                sc.classify(o, cc, n, out, stats=st)
            return
        cur = torch.cuda.current_stream()
        for s_ in streams:
            s_.wait_stream(cur)
        for i, (o, cc, nc, n, out) in enumerate(batches):
            with torch.cuda.stream(streams[i & 1]):
                scs[i & 1].classify(o, cc, n, out, stats=st)
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
    # probe kernel alone (the HBM-bound one) on this rank's inbox of the last exchange
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if peer:
        assert not sc.overflowed(), "inbox capacity exceeded"
        n_in = lookups_here // nch
        barrier()
        sc.stage_probe(); torch.cuda.synchronize()
        p0.record(); sc.stage_probe(); p1.record(); torch.cuda.synchronize()
        barrier()
        sent_rank0 = lookups_here
        # one more batch, phase by phase (count+route / probe / vote), barriers outside the brackets
        o, cc, nc, n, out = batches[0]
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
        ev[0].record(); pos_ = sc.stage_route(o, cc, n); ev[1].record(); barrier()
        ev[2].record(); sc.stage_probe(); ev[3].record(); barrier()
        ev[4].record(); sc.stage_vote(pos_, n, out); ev[5].record(); torch.cuda.synchronize(); barrier()
        phases = {"count_route_ms": ev[0].elapsed_time(ev[1]), "probe_ms": ev[2].elapsed_time(ev[3]),
                  "vote_ms": ev[4].elapsed_time(ev[5]), "reads": n}
    else:
        inbox, n_in = sc._keep[2], sc.last["received"]
        lab = sc._keep[3]
        db.shard_probe(inbox.data_ptr(), n_in, lab.data_ptr()); torch.cuda.synchronize()
        p0.record(); db.shard_probe(inbox.data_ptr(), n_in, lab.data_ptr()); p1.record(); torch.cuda.synchronize()
        sent_rank0 = sc.last["sent"] * nch
    probe_ms = p0.elapsed_time(p1)
    if not peer:
        phases = None
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
            sc.classify(do.data_ptr(), dc.data_ptr(), n, out)
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
        achieved = n_in * 32 / (probe_ms * 1e-3) / 1e9
        emit(json.dumps({
            "metric": "kmer_lookups_per_s", "value": value, "unit": "lookups/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u64", "data": "synthetic", "reads_per_s": NR * world * args.steps / (ms * 1e-3),
            "config": {"workload": f"cuCLARK full k={k}, synthetic {G} x {LG} bp DB ({n_kmers} {k}-mers), {NR} x {L} bp SE reads per GPU",
                       "parallelism": f"table sharded x{world} by hash range, k-mer exchange = " + ("kernel stores into peer inboxes (CUDA IPC over NVLink)" if peer else "NCCL all-to-all") + f", {nch} exchange round(s) per step" + (", two rounds in flight on two streams" if overlap else ""),
                       "l2": "inputs larger than L2 (shard %.1f GB, random access)" % (info.bytes / 1e9),
                       "shard_gb": info.bytes / 1e9, "shard_entries": info.entries, "db_build_s": t_db,
                       "kmers_sent_per_step_rank0": sent_rank0, "probes_per_step_rank0": lookups_here, "phases_rank0": phases},
            "clocks": sampler.summary(),
            "e2e": {"value": total_lookups * args.steps / t_e2e, "unit": "lookups/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "ms_per_step": 1e3 * t_e2e / args.steps},
            "gpu_launches": 6 * nch * args.steps,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                         "peak_source": peak_src, "bytes_per_lookup": 32, "kernel": "ccb::k_probe_list", "probe_ms": probe_ms,
                         "probes": n_in}}))
    if overlap:
        sc2.close()
    if peer:
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
