"""Sharded-table classification: one process per GPU, shard g of the
k-mer table on rank g, one exchange step each way per batch.

Replaces the reference's way of handling a database that does not fit one device -- bucket-range
parts swapped through the same GPU and merged per read (src/CuClarkDB.cu:552-559, 776-815,
1211-1214, 1261-1355) -- by keeping every part resident on its own GPU:

    requester            owner                 requester
    count -> route  ==>  probe           ==>   vote
             k-mers (8 B each)    labels (2 B each)

`exchange="nccl"`  : all_to_all_single over NVLink (variable split sizes, one host sync per batch
                     for the sizes).
PyTorch is plumbing here (buffers, streams, the process group); every compute step is a kernel of
libcuclark_b200.so reached through the C ABI (include/cuclark_b200.h, ccb_shard_*).
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from .lib import CuClarkDB


class ShardedClassifier:
    """db: a CuClarkDB on this rank's GPU configured with shard_config(world, rank) BEFORE its table was
    loaded.  classify() is collective: every rank calls it with its own batch (possibly empty)."""

    def __init__(self, db: CuClarkDB, group=None):
        self.db = db
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.dev = torch.device("cuda", torch.cuda.current_device())
        self.last = {}

    def classify(self, d_offsets: int, d_containers: int, n_reads: int, out5: torch.Tensor,
                 rows: torch.Tensor | None = None, stats: torch.Tensor | None = None) -> torch.Tensor:
        """d_offsets/d_containers: device pointers of a packed batch; out5: int16[n_reads*5] on the GPU."""
        G, db = self.world, self.db
        stream = torch.cuda.current_stream().cuda_stream
        pos = torch.empty(G * n_reads + 1, dtype=torch.int32, device=self.dev)
        db.shard_count(d_offsets, d_containers, n_reads, G, pos.data_ptr(), stream)
        # segment boundaries -> host (sizes of the variable all-to-all)
        idx = torch.arange(0, G + 1, device=self.dev, dtype=torch.int64) * n_reads
        bounds = pos.view(torch.int32)[idx].to(torch.int64) & 0xFFFFFFFF
        bounds_h = bounds.cpu().tolist()
        send_counts = [bounds_h[g + 1] - bounds_h[g] for g in range(G)]
        total = bounds_h[G]
        send = torch.empty(max(total, 1), dtype=torch.int64, device=self.dev)
        seg_ptr = [send.data_ptr() + 8 * bounds_h[g] for g in range(G)]
        db.shard_route(d_offsets, d_containers, n_reads, G, pos.data_ptr(), bounds_h[:G], seg_ptr, stream)
        # sizes each owner will receive
        sc = torch.tensor(send_counts, dtype=torch.int64, device=self.dev)
        rc = torch.empty_like(sc)
        if G > 1:
            dist.all_to_all_single(rc, sc, group=self.group)
        else:
            rc.copy_(sc)
        recv_counts = rc.cpu().tolist()
        n_in = sum(recv_counts)
        inbox = torch.empty(max(n_in, 1), dtype=torch.int64, device=self.dev)
        if G > 1:
            dist.all_to_all_single(inbox[:n_in], send[:total], recv_counts, send_counts, group=self.group)
        else:
            inbox[:n_in].copy_(send[:total])
        lab_out = torch.empty(max(n_in, 1), dtype=torch.int16, device=self.dev)
        db.shard_probe(inbox.data_ptr(), n_in, lab_out.data_ptr(), stats.data_ptr() if stats is not None else 0, stream)
        lab_in = torch.empty(max(total, 1), dtype=torch.int16, device=self.dev)
        if G > 1:
            # NCCL has no 16-bit integer type: the labels travel as bytes
            dist.all_to_all_single(lab_in[:total].view(torch.uint8), lab_out[:n_in].view(torch.uint8),
                                   [2 * c for c in send_counts], [2 * c for c in recv_counts], group=self.group)
        else:
            lab_in[:total].copy_(lab_out[:n_in])
        src_ptr = [lab_in.data_ptr() + 2 * bounds_h[g] for g in range(G)]
        db.shard_vote(n_reads, G, pos.data_ptr(), bounds_h[:G], src_ptr, out5.data_ptr(),
                      rows.data_ptr() if rows is not None else 0, stream)
        self.last = {"sent": total, "received": n_in, "send_counts": send_counts}
        # keep the temporaries alive until the stream has consumed them
        self._keep = (pos, send, inbox, lab_out, lab_in)
        return out5


class PeerShardedClassifier:
    """Same phases, but the exchange is done BY the kernels: `k_query<route>` stores each k-mer straight into
    the owner GPU's inbox and `k_probe_inbox` stores each label straight into the requester GPU's label box
    (peer memory mapped with CUDA IPC, plain stores over NVLink / NVSwitch).  No collective moves data, the
    host never learns the segment sizes (they travel as device words), and nothing on the host synchronises
    per batch; two one-word all-reduces per batch act as stream-ordered barriers between the phases.

    cap = k-mers one rank may send to one owner per batch (the inbox region size); a batch that needs
    more sets `error` and must be re-run in smaller pieces."""

    def __init__(self, db: CuClarkDB, cap: int, group=None, peers=None):
        self.db, self.cap, self.group = db, int(cap), group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.dev = torch.device("cuda", torch.cuda.current_device())
        G = self.world if peers is None else len(peers["inbox"])
        self.G = G
        self._opened = []
        if peers is None:
            # own buffers (plain cudaMalloc so that they can be exported), zeroed counts
            self.inbox = db.device_alloc(G * self.cap * 8)
            self.labbox = db.device_alloc(G * self.cap * 2)
            self.counts_in = db.device_alloc(G * 8)
            self._owned = [self.inbox, self.labbox, self.counts_in]
            mine = (db.ipc_export(self.inbox), db.ipc_export(self.labbox), db.ipc_export(self.counts_in), self.cap)
            allh = [None] * G
            if G > 1:
                dist.all_gather_object(allh, mine, group=group)
            else:
                allh[0] = mine
            if any(hd[3] != self.cap for hd in allh):
                raise ValueError(f"every rank must use the same inbox region size, got {[hd[3] for hd in allh]}")
            self.peer_inbox, self.peer_labbox, self.peer_counts = [], [], []
            for g in range(G):
                if g == self.rank:
                    ptrs = (self.inbox, self.labbox, self.counts_in)
                else:
                    ptrs = tuple(db.ipc_open(hd) for hd in allh[g][:3])
                    self._opened.extend(ptrs)
                self.peer_inbox.append(ptrs[0]); self.peer_labbox.append(ptrs[1]); self.peer_counts.append(ptrs[2])
        else:
            # explicit pointer tables (tests: every shard on one device of one process)
            self._owned = []
            self.rank = peers["rank"]
            self.peer_inbox, self.peer_labbox, self.peer_counts = peers["inbox"], peers["labbox"], peers["counts"]
            self.inbox, self.labbox, self.counts_in = self.peer_inbox[self.rank], self.peer_labbox[self.rank], self.peer_counts[self.rank]
        self.error = torch.zeros(1, dtype=torch.int32, device=self.dev)
        self._bar = torch.zeros(1, dtype=torch.int32, device=self.dev)
        self._keep = None

    def barrier(self):
        if self.world > 1 and dist.is_initialized():
            dist.all_reduce(self._bar, group=self.group)

    # the three stages are separate so that a single-process test can interleave several "ranks"
    def stage_route(self, d_offsets, d_containers, n_reads):
        G, r, cap, db = self.G, self.rank, self.cap, self.db
        stream = torch.cuda.current_stream().cuda_stream
        pos = torch.empty(G * n_reads + 1, dtype=torch.int32, device=self.dev)
        db.shard_count(d_offsets, d_containers, n_reads, G, pos.data_ptr(), stream)
        db.shard_route_peer(d_offsets, d_containers, n_reads, G, pos.data_ptr(),
                            [self.peer_inbox[g] + r * cap * 8 for g in range(G)], cap,
                            [self.peer_counts[g] + r * 8 for g in range(G)], self.error.data_ptr(), stream)
        self._keep = pos
        return pos

    def stage_probe(self, stats=None):
        G, r, cap = self.G, self.rank, self.cap
        stream = torch.cuda.current_stream().cuda_stream
        self.db.shard_probe_inbox(G, self.inbox, cap, self.counts_in, [self.peer_labbox[s] + r * cap * 2 for s in range(G)],
                                  stats.data_ptr() if stats is not None else 0, stream)

    def stage_vote(self, pos, n_reads, out5, rows=None):
        G, cap = self.G, self.cap
        stream = torch.cuda.current_stream().cuda_stream
        self.db.shard_vote(n_reads, G, pos.data_ptr(), None, [self.labbox + g * cap * 2 for g in range(G)],
                           out5.data_ptr(), rows.data_ptr() if rows is not None else 0, stream)

    def classify(self, d_offsets: int, d_containers: int, n_reads: int, out5: torch.Tensor,
                 rows: torch.Tensor | None = None, stats: torch.Tensor | None = None) -> torch.Tensor:
        pos = self.stage_route(d_offsets, d_containers, n_reads)
        self.barrier()                       # every rank's k-mers and counts have landed
        self.stage_probe(stats)
        self.barrier()                       # every owner's labels have landed
        self.stage_vote(pos, n_reads, out5, rows)
        return out5

    def overflowed(self) -> bool:
        """True if some batch since the last call sent more than cap k-mers to one owner (host sync)."""
        v = bool(int(self.error.item()))
        self.error.zero_()
        return v

    def close(self):
        torch.cuda.synchronize()
        if self.world > 1 and dist.is_initialized():
            dist.barrier(group=self.group)
        for p in self._opened:
            self.db.ipc_close(p)
        for p in self._owned:
            self.db.device_free(p)
        self._opened, self._owned = [], []
