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
