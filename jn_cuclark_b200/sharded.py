"""Sharded-table classification: one process per GPU, shard g of the
k-mer table on rank g, one exchange step each way per batch.

Replaces the reference's way of handling a database that does not fit one device -- bucket-range
parts swapped through the same GPU and merged per read (src/CuClarkDB.cu:552-559, 776-815,
1211-1214, 1261-1355) -- by keeping every part resident on its own GPU:

    requester            owner                 requester
    route           ==>  probe           ==>   vote
             k-mers (8 B each)    labels (2 B each)

The device side is the one-pass pipeline of csrc/shard_block.cu (ccb_shard_route2 / ccb_shard_vote2): a block of
256 reads shares one segment per owner, reserved with one atomic; no count pass, no scan.  Two ways to move the bytes:

`ShardedClassifier`      NCCL all-to-all over NVLink (variable sizes, one host sync per batch for them).
`PeerShardedClassifier`  the kernels store straight into peer memory (CUDA IPC); no collective on the data path.

PyTorch is plumbing here (buffers, streams, the process group); every compute step is a kernel of
libcuclark_b200.so reached through the C ABI (include/cuclark_b200.h, ccb_shard_*).  The single-process,
multi-device driver of the same kernels is C++ (csrc/group.cu, `bin/cuCLARK --sharded`).
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from .lib import CuClarkDB


def region_size(max_reads: int, windows_per_read: int, world: int, slack: float = 1.25) -> int:
    """k-mers one rank may send to one owner per batch: the batch's k-mers spread evenly (mix2k is uniform), plus slack."""
    return int(max_reads * windows_per_read / max(world, 1) * slack) + 8192


class ShardedClassifier:
    """db: a CuClarkDB on this rank's GPU configured with shard_config(world, rank) BEFORE its table was
    loaded.  classify() is collective: every rank calls it with its own batch (possibly empty).
    cap = k-mers one rank may send to one owner per batch, max_reads = reads per batch."""

    def __init__(self, db: CuClarkDB, cap: int, max_reads: int, group=None):
        self.db, self.cap, self.group = db, int(cap), group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.dev = torch.device("cuda", torch.cuda.current_device())
        G = self.world
        self.ws = db.shardws_create(max_reads, G, self.cap)
        self.send = torch.empty(G * self.cap, dtype=torch.int64, device=self.dev)
        self.lab_in = torch.empty(G * self.cap, dtype=torch.int16, device=self.dev)
        self.sent = torch.zeros(G, dtype=torch.int64, device=self.dev)
        self.last = {}

    def classify(self, d_offsets: int, d_containers: int, n_reads: int, out5: torch.Tensor,
                 rows: torch.Tensor | None = None, stats: torch.Tensor | None = None, n_containers: int = -1) -> torch.Tensor:
        """d_offsets/d_containers: device pointers of a packed batch; out5: int16[n_reads*5] on the GPU."""
        G, db, cap = self.world, self.db, self.cap
        stream = torch.cuda.current_stream().cuda_stream
        db.shard_route2(self.ws, d_offsets, d_containers, n_reads, n_containers,
                        [self.send.data_ptr() + 8 * g * cap for g in range(G)],
                        [self.sent.data_ptr() + 8 * g for g in range(G)], stream)
        # sizes each owner will receive (one host sync per batch)
        rc = torch.empty_like(self.sent)
        if G > 1:
            dist.all_to_all_single(rc, self.sent, group=self.group)
        else:
            rc.copy_(self.sent)
        send_counts = self.sent.cpu().tolist()
        recv_counts = rc.cpu().tolist()
        if max(send_counts) > cap:
            raise RuntimeError(f"a region of {cap} k-mers is too small for this batch ({max(send_counts)} to one owner)")
        n_in = sum(recv_counts)
        inbox = torch.empty(max(n_in, 1), dtype=torch.int64, device=self.dev)
        seg_in = [self.send[g * cap: g * cap + send_counts[g]] for g in range(G)]
        if G > 1:
            dist.all_to_all(list(inbox[:n_in].split(recv_counts)), seg_in, group=self.group)
        else:
            inbox[:n_in].copy_(seg_in[0])
        lab_out = torch.empty(max(n_in, 1), dtype=torch.int16, device=self.dev)
        db.shard_probe(inbox.data_ptr(), n_in, lab_out.data_ptr(), stats.data_ptr() if stats is not None else 0, stream)
        if G > 1:
            # NCCL has no 16-bit integer type: the labels travel as bytes
            dist.all_to_all([self.lab_in[g * cap: g * cap + send_counts[g]].view(torch.uint8) for g in range(G)],
                            [t.view(torch.uint8) for t in lab_out[:n_in].split(recv_counts)], group=self.group)
        else:
            self.lab_in[:n_in].copy_(lab_out[:n_in])
        db.shard_vote2(self.ws, n_reads, [self.lab_in.data_ptr() + 2 * g * cap for g in range(G)], out5.data_ptr(),
                       rows.data_ptr() if rows is not None else 0, stream)
        self.last = {"sent": sum(send_counts), "received": n_in, "send_counts": send_counts}
        self._keep = (inbox, lab_out)          # alive until the stream has consumed them
        return out5

    def close(self):
        torch.cuda.synchronize()
        if self.ws:
            self.db.shardws_destroy(self.ws); self.ws = 0


class PeerShardedClassifier:
    """Same phases, but the exchange is done BY the kernels: the route kernels store each k-mer straight into
    the owner GPU's inbox and `k_probe_inbox` stores each label straight into the requester GPU's label box
    (peer memory mapped with CUDA IPC, plain stores over NVLink / NVSwitch).  No collective moves data, the
    host never learns the segment sizes (they travel as device words), and nothing on the host synchronises
    per batch; two one-word all-reduces per batch act as stream-ordered barriers between the phases.

    cap = k-mers one rank may send to one owner per batch (the inbox region size); a batch that needs
    more sets the error word (overflowed()) and must be re-run in smaller pieces."""

    def __init__(self, db: CuClarkDB, cap: int, max_reads: int, group=None, peers=None):
        self.db, self.cap, self.group = db, int(cap), group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.dev = torch.device("cuda", torch.cuda.current_device())
        G = self.world if peers is None else len(peers["inbox"])
        self.G = G
        self._opened = []
        self.ws = db.shardws_create(max_reads, G, self.cap)
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
        self._bar = torch.zeros(1, dtype=torch.int32, device=self.dev)

    def barrier(self):
        if self.world > 1 and dist.is_initialized():
            dist.all_reduce(self._bar, group=self.group)

    # the three stages are separate so that a single-process test can interleave several "ranks"
    def stage_route(self, d_offsets, d_containers, n_reads, n_containers=-1):
        G, r, cap = self.G, self.rank, self.cap
        stream = torch.cuda.current_stream().cuda_stream
        self.db.shard_route2(self.ws, d_offsets, d_containers, n_reads, n_containers,
                             [self.peer_inbox[g] + r * cap * 8 for g in range(G)],
                             [self.peer_counts[g] + r * 8 for g in range(G)], stream)

    def stage_probe(self, stats=None):
        G, r, cap = self.G, self.rank, self.cap
        stream = torch.cuda.current_stream().cuda_stream
        self.db.shard_probe_inbox(G, self.inbox, cap, self.counts_in, [self.peer_labbox[s] + r * cap * 2 for s in range(G)],
                                  stats.data_ptr() if stats is not None else 0, stream)

    def stage_vote(self, n_reads, out5, rows=None):
        G, cap = self.G, self.cap
        stream = torch.cuda.current_stream().cuda_stream
        self.db.shard_vote2(self.ws, n_reads, [self.labbox + g * cap * 2 for g in range(G)],
                            out5.data_ptr(), rows.data_ptr() if rows is not None else 0, stream)

    def classify(self, d_offsets: int, d_containers: int, n_reads: int, out5: torch.Tensor,
                 rows: torch.Tensor | None = None, stats: torch.Tensor | None = None, n_containers: int = -1) -> torch.Tensor:
        self.stage_route(d_offsets, d_containers, n_reads, n_containers)
        self.barrier()                       # every rank's k-mers and counts have landed
        self.stage_probe(stats)
        self.barrier()                       # every owner's labels have landed
        self.stage_vote(n_reads, out5, rows)
        return out5

    def overflowed(self) -> bool:
        """True if some batch since the last call sent more than cap k-mers to one owner (host sync).  The vote of
        such a batch ran on the k-mers that fitted: its results are not the whole-table results."""
        return self.db.shardws_error(self.ws, torch.cuda.current_stream().cuda_stream)

    def close(self):
        torch.cuda.synchronize()
        if self.world > 1 and dist.is_initialized():
            dist.barrier(group=self.group)
        for p in self._opened:
            self.db.ipc_close(p)
        for p in self._owned:
            self.db.device_free(p)
        if self.ws:
            self.db.shardws_destroy(self.ws); self.ws = 0
        self._opened, self._owned = [], []


class PipelinedPeerClassifier:
    """A batch cut into chunks that travel through `n_pipes` PeerShardedClassifier pipelines (own inboxes, label boxes,
    workspaces, barrier groups and streams): while the owners probe chunk i, the requesters already route chunk i+1 and
    vote chunk i-1.  The three kernels are launched on persistent grids that fit one SM together (two blocks each), so
    the HBM-bound probe and the issue/NVLink-bound route really share the SMs instead of queueing behind each other.
    max_reads = reads per CHUNK, cap = region size per chunk."""

    def __init__(self, db: CuClarkDB, cap: int, max_reads: int, n_pipes: int = 2):
        self.db, self.n_pipes = db, n_pipes
        world = dist.get_world_size() if dist.is_initialized() else 1
        self.groups = [dist.new_group(list(range(world))) if world > 1 else None for _ in range(n_pipes)]
        self.pipes = [PeerShardedClassifier(db, cap, max_reads, group=g) for g in self.groups]
        self.streams = [torch.cuda.Stream() for _ in range(n_pipes)]
        self.max_reads = max_reads

    def classify(self, d_offsets: int, d_containers: int, n_reads: int, out5: torch.Tensor, stats=None, n_containers: int = -1):
        per = self.max_reads
        cur = torch.cuda.current_stream()
        for s_ in self.streams:
            s_.wait_stream(cur)
        c = 0
        for a in range(0, max(n_reads, 1), per):
            n = min(per, n_reads - a)
            p = c % self.n_pipes
            with torch.cuda.stream(self.streams[p]):
                # offsets are absolute container indices: a chunk is a window of the offsets array over the same containers
                self.pipes[p].classify(d_offsets + 4 * a, d_containers, n, out5[5 * a:], stats=stats,
                                       n_containers=(n_containers * n // max(n_reads, 1)) if n_containers >= 0 else -1)
            c += 1
        for s_ in self.streams:
            cur.wait_stream(s_)
        return out5

    def overflowed(self) -> bool:
        return any([p.overflowed() for p in self.pipes])

    def close(self):
        for p in self.pipes:
            p.close()
