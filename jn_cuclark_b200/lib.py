"""ctypes binding of ``libcuclark_b200.so`` (include/cuclark_b200.h).

``CuClarkDB`` mirrors the public interface of the reference's
``class CuClarkDB<HKMERr>`` (reference src/CuClarkDB.cuh:99-151) method for
method -- ``read / malloc / readyBatch / queryBatch / waitForBatch /
checkBatch / sync / freeBatchMemory`` -- so the parity tests read like the
reference's host engine (src/CuCLARK_hh.hh:599-632, 1599-1605, 1734-1741,
2081-2091).  All computing happens in the CUDA library; there is no Python or
CPU implementation behind these calls.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

CCB_EXTENDED, CCB_STATS, CCB_FORCE_DENSE, CCB_NO_ANCHOR, CCB_WARP_KERNEL, CCB_STATS_EX = 1, 2, 4, 8, 32, 64
default_flags = 0          # ORed into every query (tests run the parity suite once per kernel with it)

_HERE = os.path.dirname(os.path.abspath(__file__))


def lib_path() -> str:
    # CCB_LIBRARY: a variant build of the same library (measurements: `make -C jn_cuclark_b200 VARIANT=...`)
    return os.environ.get("CCB_LIBRARY") or os.path.join(_HERE, "libcuclark_b200.so")


class CcbError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"[ccb {code}] {msg}")
        self.code = code
        self.msg = msg


class _Config(C.Structure):
    _fields_ = [("device", C.c_int), ("k", C.c_int), ("n_targets", C.c_int), ("max_hits", C.c_int),
                ("n_slots", C.c_int), ("load_factor", C.c_double), ("verbose", C.c_int)]


class Workload(C.Structure):
    """struct ccb_workload (include/ccb_synth.h): kind 1 pairs joined by N, kind 2 long reads with indels"""
    _fields_ = [("kind", C.c_uint32), ("read_len", C.c_uint32), ("insert", C.c_uint32),
                ("sub_pm", C.c_uint32), ("ins_pm", C.c_uint32), ("del_pm", C.c_uint32)]


class DbInfo(C.Structure):
    _fields_ = [("entries", C.c_uint64), ("buckets", C.c_uint64), ("bytes", C.c_uint64),
                ("tombstones", C.c_uint64), ("max_displacement", C.c_uint32)]


class BuildInfo(C.Structure):
    _fields_ = [("bases", C.c_uint64), ("occurrences", C.c_uint64), ("distinct", C.c_uint64),
                ("stored", C.c_uint64), ("max_bucket", C.c_uint32), ("reserved", C.c_uint32)]


class AnchorInfo(C.Structure):
    _fields_ = [("ready", C.c_int), ("bases", C.c_uint64), ("entries", C.c_uint64), ("dropped", C.c_uint64), ("bytes", C.c_uint64)]


class Stats(C.Structure):
    _fields_ = [("reads", C.c_uint64), ("lookups", C.c_uint64), ("sectors", C.c_uint64),
                ("dense_reads", C.c_uint64), ("table_probes", C.c_uint64)]


_lib = None

# every symbol include/cuclark_b200.h declares: (restype, argtypes)
_VP, _H = C.c_void_p, C.c_void_p
SYMBOLS = {
    "ccb_create": (C.c_int, [C.POINTER(_Config), C.POINTER(_H)]),
    "ccb_destroy": (None, [_H]),
    "ccb_last_error": (C.c_char_p, [_H]),
    "ccb_db_load_files": (C.c_int, [_H, C.c_char_p, C.c_uint64, C.c_int, C.c_uint, C.POINTER(C.c_uint64)]),
    "ccb_db_begin": (C.c_int, [_H, C.c_uint64]),
    "ccb_db_add_pairs": (C.c_int, [_H, _VP, _VP, C.c_uint64]),
    "ccb_db_add_pairs_device": (C.c_int, [_H, _VP, _VP, C.c_uint64]),
    "ccb_db_end": (C.c_int, [_H]),
    "ccb_db_get_info": (C.c_int, [_H, C.POINTER(DbInfo)]),
    "ccb_build_begin": (C.c_int, [_H, C.c_uint64, C.c_int, C.c_int, C.c_uint, C.c_uint, C.c_uint64]),
    "ccb_build_add_fasta": (C.c_int, [_H, C.c_char_p, C.c_uint64, C.c_uint16]),
    "ccb_build_end": (C.c_int, [_H, C.c_char_p, C.c_int, C.POINTER(BuildInfo)]),
    "ccb_shard_config": (C.c_int, [_H, C.c_int, C.c_int]),
    "ccb_shard_count": (C.c_int, [_H, _VP, _VP, C.c_uint32, C.c_int, _VP, _VP]),
    "ccb_shard_route": (C.c_int, [_H, _VP, _VP, C.c_uint32, C.c_int, _VP, C.POINTER(C.c_uint64), C.POINTER(_VP), _VP]),
    "ccb_shard_probe": (C.c_int, [_H, _VP, C.c_uint64, _VP, _VP, _VP]),
    "ccb_shard_vote": (C.c_int, [_H, C.c_uint32, C.c_int, _VP, C.POINTER(C.c_uint64), C.POINTER(_VP), _VP, _VP, _VP]),
    "ccb_shard_probe_inbox": (C.c_int, [_H, C.c_int, _VP, C.c_uint64, _VP, C.POINTER(_VP), _VP, _VP]),
    "ccb_shardws_create": (C.c_int, [_H, C.c_uint32, C.c_int, C.c_uint64, C.POINTER(_VP)]),
    "ccb_shardws_destroy": (None, [_H, _VP]),
    "ccb_shardws_error": (C.c_int, [_H, _VP, _VP, C.POINTER(C.c_int)]),
    "ccb_shard_route2": (C.c_int, [_H, _VP, _VP, _VP, C.c_uint32, C.c_int64, C.POINTER(_VP), C.POINTER(_VP), _VP]),
    "ccb_shard_vote2": (C.c_int, [_H, _VP, C.c_uint32, C.POINTER(_VP), _VP, _VP, _VP]),
    "ccb_group_create": (C.c_int, [C.POINTER(C.c_int), C.c_int, C.POINTER(_Config), C.POINTER(_H)]),
    "ccb_group_size": (C.c_int, [_H]),
    "ccb_group_member": (_H, [_H, C.c_int]),
    "ccb_device_alloc": (C.c_int, [_H, C.c_uint64, C.POINTER(_VP)]),
    "ccb_ipc_export": (C.c_int, [_H, _VP, C.c_char_p]),
    "ccb_ipc_open": (C.c_int, [_H, C.c_char_p, C.POINTER(_VP)]),
    "ccb_ipc_close": (C.c_int, [_H, _VP]),
    "ccb_anchor_synth": (C.c_int, [_H, C.c_uint64, C.c_uint32, C.c_uint64, C.c_uint32]),
    "ccb_anchor_begin": (C.c_int, [_H, C.c_uint64]),
    "ccb_anchor_add_fasta": (C.c_int, [_H, C.c_char_p, C.c_uint64, C.c_uint16]),
    "ccb_anchor_end": (C.c_int, [_H, C.POINTER(AnchorInfo)]),
    "ccb_anchor_drop": (C.c_int, [_H]),
    "ccb_anchor_get_info": (C.c_int, [_H, C.POINTER(AnchorInfo)]),
    "ccb_batch_alloc": (C.c_int, [_H, C.c_uint32, C.c_uint64, C.c_int]),
    "ccb_batch_alloc_ref": (C.c_int, [_H, C.c_uint64, C.c_uint32, C.c_uint64, C.POINTER(C.c_uint64), C.c_int, C.c_int,
                                      C.POINTER(_VP), C.POINTER(_VP)]),
    "ccb_batch_buffers": (C.c_int, [_H, C.c_int, C.POINTER(_VP), C.POINTER(_VP), C.POINTER(_VP), C.POINTER(_VP)]),
    "ccb_batch_submit": (C.c_int, [_H, C.c_int, C.c_uint32, C.c_uint64, C.c_int]),
    "ccb_batch_wait": (C.c_int, [_H, C.c_int]),
    "ccb_batch_poll": (C.c_int, [_H, C.c_int]),
    "ccb_batch_free": (C.c_int, [_H]),
    "ccb_sync": (C.c_int, [_H]),
    "ccb_query_device": (C.c_int, [_H, _VP, _VP, C.c_uint32, _VP, _VP, _VP, C.c_int, _VP]),
    "ccb_query_device_noalloc": (C.c_int, [_H, _VP, _VP, C.c_uint32, _VP, _VP, _VP, C.c_int, _VP]),
    "ccb_get_stats": (C.c_int, [_H, C.POINTER(Stats)]),
    "ccb_version": (C.c_char_p, []),
    "ccb_l2_fetch_granularity": (C.c_int, [_H]),
    "ccb_device_count": (C.c_int, []),
    "ccb_device_memory": (C.c_int, [C.c_int, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "ccb_batch_slots": (C.c_int, [_H]),
    "ccb_synth_db": (C.c_int, [_H, C.c_uint64, C.c_uint32, C.c_uint64, C.c_uint32, C.POINTER(C.c_uint64)]),
    "ccb_synth_reads_packed": (C.c_int, [_H, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32,
                                         C.c_uint64, C.POINTER(_VP), C.POINTER(_VP), C.POINTER(C.c_uint64)]),
    "ccb_synth_reads_text": (C.c_int, [_H, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32,
                                       C.c_uint64, _VP]),
    "ccb_synth_db_light": (C.c_int, [_H, C.c_uint64, C.c_uint32, C.c_uint64, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint64)]),
    "ccb_synth2_reads_packed": (C.c_int, [_H, _VP, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint64,
                                          C.POINTER(_VP), C.POINTER(_VP), C.POINTER(C.c_uint64)]),
    "ccb_device_free": (C.c_int, [_H, _VP]),
    "ccb_bench_gather": (C.c_int, [_H, C.c_uint64, C.c_int, _VP]),
    "ccb_bench_gather_ex": (C.c_int, [_H, C.c_uint64, C.c_int, C.c_int, C.c_uint64, _VP]),
}


def load_library() -> C.CDLL:
    """Load the CUDA library; fail loudly when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    p = lib_path()
    if not os.path.exists(p):
        raise ImportError(
            f"{p} is missing: the CUDA extension has not been built "
            "(run `make -C jn_cuclark_b200` or __graft_entry__.build()). There is no CPU fallback.")
    lib = C.CDLL(p)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)          # AttributeError if the ABI lost a symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def _np_view(ptr: int, n: int, dtype) -> np.ndarray:
    if n == 0 or not ptr:
        return np.zeros(0, dtype=dtype)
    buf = (C.c_char * (n * np.dtype(dtype).itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype, count=n)


class CuClarkDB:
    """Device database + batch scheduler; same method names as the reference class."""

    def __init__(self, numDevices: int = 1, k: int = 31, numBatches: int = 1, numTargets: int = 1,
                 verbose: bool = False, *, device: int = 0, max_hits: int = 15, load_factor: float = 0.0):
        # reference ctor: CuClarkDB(numDevices, k, numBatches, numTargets, verbose)
        # (src/CuClarkDB.cu:95); one handle drives one device here, so numDevices
        # must be 1 and `device` picks the ordinal.
        if numDevices not in (0, 1):
            raise ValueError("one CuClarkDB drives one GPU; create one per device")
        self._lib = load_library()
        self._h = _H()
        cfg = _Config(device, k, numTargets, max_hits, numBatches, load_factor, int(verbose))
        rc = self._lib.ccb_create(C.byref(cfg), C.byref(self._h))
        if rc:
            raise CcbError(rc, self._lib.ccb_last_error(None).decode())
        self.k, self.numTargets, self.numBatches, self.max_hits = k, numTargets, numBatches, max_hits
        self.row_stride = 2 * max_hits + 2
        self._max_reads = 0
        self._n = {}

    @classmethod
    def _view(cls, lib, handle, k, numTargets, max_hits):
        """Non-owning wrapper of a handle the library owns (a member of a device group)."""
        o = cls.__new__(cls)
        o._lib, o._h, o._borrowed = lib, _H(handle), True
        o.k, o.numTargets, o.numBatches, o.max_hits = k, numTargets, 1, max_hits
        o.row_stride = 2 * max_hits + 2
        o._max_reads, o._n = 0, {}
        return o

    # -- helpers ------------------------------------------------------------
    def _ck(self, rc: int):
        if rc:
            raise CcbError(rc, self._lib.ccb_last_error(self._h).decode())

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            if not getattr(self, "_borrowed", False):
                self._lib.ccb_destroy(self._h)
            self._h = _H()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- database -----------------------------------------------------------
    def read(self, filename: str, htsize: int, key_bytes: int, modCollision: int = 1) -> int:
        """CuClarkDB::read + swapDbParts (src/CuClarkDB.cu:464, 776). Returns #k-mers loaded."""
        n = C.c_uint64(0)
        self._ck(self._lib.ccb_db_load_files(self._h, filename.encode(), htsize, key_bytes, modCollision, C.byref(n)))
        return n.value

    def load_pairs(self, canon: np.ndarray, labels: np.ndarray):
        canon = np.ascontiguousarray(canon, dtype=np.uint64)
        labels = np.ascontiguousarray(labels, dtype=np.uint16)
        assert canon.shape == labels.shape
        self._ck(self._lib.ccb_db_begin(self._h, canon.size))
        self._ck(self._lib.ccb_db_add_pairs(self._h, canon.ctypes.data, labels.ctypes.data, canon.size))
        self._ck(self._lib.ccb_db_end(self._h))

    def db_begin(self, expected: int):
        self._ck(self._lib.ccb_db_begin(self._h, expected))

    def db_add_device(self, d_canon: int, d_labels: int, n: int):
        self._ck(self._lib.ccb_db_add_pairs_device(self._h, d_canon, d_labels, n))

    def db_end(self):
        self._ck(self._lib.ccb_db_end(self._h))

    def db_info(self) -> DbInfo:
        info = DbInfo()
        self._ck(self._lib.ccb_db_get_info(self._h, C.byref(info)))
        return info

    def build(self, targets, htsize: int, key_bytes: int = 4, light: bool = False, gap: int = 4,
              min_count: int = 0, prefix: str | None = None, load_table: bool = True) -> BuildInfo:
        """makeSpecificTargetSets + hTable::write (src/CuCLARK_hh.hh:691-1329, src/hashTable_hh.hh:474-546)
        on the device.  targets: iterable of (fasta bytes, 0-based label index)."""
        targets = list(targets)
        total = sum(len(t) for t, _ in targets)
        self._ck(self._lib.ccb_build_begin(self._h, htsize, key_bytes, int(light), gap, min_count, total))
        for text, label in targets:
            self._ck(self._lib.ccb_build_add_fasta(self._h, text, len(text), label))
        info = BuildInfo()
        self._ck(self._lib.ccb_build_end(self._h, prefix.encode() if prefix else None, int(load_table), C.byref(info)))
        return info

    # -- batches --------------------------------------------------------------
    def malloc(self, maxReads: int, maxContainers: int, isExtended: bool = False):
        """CuClarkDB::malloc (src/CuClarkDB.cu:322): library-owned pinned buffers."""
        self._ck(self._lib.ccb_batch_alloc(self._h, maxReads, maxContainers, int(isExtended)))
        self._max_reads, self._max_cont, self._ext = maxReads, maxContainers, isExtended

    def malloc_ref(self, numReads: int, maxReads: int, maxContainers: int, indexBatches, isExtended: bool = False):
        """CuClarkDB::malloc with the reference's result layout (src/CuClarkDB.cu:322-421): one contiguous final (and
        full) result array for all numReads reads; batch b delivers at row indexBatches[b].  Returns (final5, rows|None)."""
        idx = (C.c_uint64 * len(indexBatches))(*[int(x) for x in indexBatches])
        full, fin = _VP(), _VP()
        self._ck(self._lib.ccb_batch_alloc_ref(self._h, numReads, maxReads, maxContainers, idx, len(indexBatches), int(isExtended),
                                               C.byref(full), C.byref(fin)))
        self._max_reads, self._max_cont, self._ext = maxReads, maxContainers, isExtended
        f = _np_view(fin.value, numReads * 5, np.uint16).reshape(-1, 5)
        r = _np_view(full.value, numReads * self.row_stride, np.uint16).reshape(-1, self.row_stride) if (isExtended and full.value) else None
        return f, r

    def buffers(self, batchId: int):
        """(offsets u32[maxReads+1], containers u16[maxContainers], final5 u16[maxReads,5], rows|None)"""
        o, c, f, r = _VP(), _VP(), _VP(), _VP()
        self._ck(self._lib.ccb_batch_buffers(self._h, batchId, C.byref(o), C.byref(c), C.byref(f), C.byref(r)))
        offs = _np_view(o.value, self._max_reads + 1, np.uint32)
        cont = _np_view(c.value, self._max_cont, np.uint16)
        fin = _np_view(f.value, self._max_reads * 5, np.uint16).reshape(-1, 5)
        rows = _np_view(r.value, self._max_reads * self.row_stride, np.uint16).reshape(-1, self.row_stride) if r.value else None
        return offs, cont, fin, rows

    def readyBatch(self, batchId: int, numReads: int, containerCount: int):
        """CuClarkDB::readyBatch (src/CuClarkDB.cu:821)."""
        self._n[batchId] = (numReads, containerCount)

    def queryBatch(self, batchId: int, isExtended: bool = False, flags: int = 0) -> bool:
        """CuClarkDB::queryBatch (src/CuClarkDB.cu:836); asynchronous."""
        n, c = self._n[batchId]
        self._ck(self._lib.ccb_batch_submit(self._h, batchId, n, c, flags | default_flags | (CCB_EXTENDED if isExtended else 0)))
        return True

    def waitForBatch(self, batchId: int) -> bool:
        self._ck(self._lib.ccb_batch_wait(self._h, batchId))
        return True

    def checkBatch(self, batchId: int) -> bool:
        rc = self._lib.ccb_batch_poll(self._h, batchId)
        if rc not in (0, 1):
            self._ck(rc)
        return rc == 0

    def sync(self) -> bool:
        self._ck(self._lib.ccb_sync(self._h))
        return True

    def freeBatchMemory(self):
        self._ck(self._lib.ccb_batch_free(self._h))
        self._max_reads = 0

    def query_device(self, d_offsets: int, d_containers: int, n_reads: int, d_final5: int,
                     d_rows: int = 0, d_stats: int = 0, flags: int = 0, stream: int = 0):
        self._ck(self._lib.ccb_query_device(self._h, d_offsets, d_containers, n_reads, d_final5,
                                            d_rows or None, d_stats or None, flags | default_flags, stream or None))

    def query_device_noalloc(self, d_offsets: int, d_containers: int, n_reads: int, d_final5: int,
                             d_rows: int = 0, d_stats: int = 0, flags: int = 0, stream: int = 0):
        """ccb_query_device_noalloc: query_device on the library's fixed scratch array -- no device allocation, at most
        NOALLOC_MAX_READS reads per call and NOALLOC_MAX_TARGETS targets; calls on one GPU must not overlap."""
        self._ck(self._lib.ccb_query_device_noalloc(self._h, d_offsets, d_containers, n_reads, d_final5,
                                                    d_rows or None, d_stats or None, flags | default_flags, stream or None))

    # -- sharded table (row E2): raw device pointers, the caller owns buffers and the exchange ----
    def shard_config(self, n_shards: int, shard: int):
        self._ck(self._lib.ccb_shard_config(self._h, n_shards, shard))

    def shard_count(self, d_offsets: int, d_containers: int, n_reads: int, n_shards: int, d_pos: int, stream: int = 0):
        self._ck(self._lib.ccb_shard_count(self._h, d_offsets, d_containers, n_reads, n_shards, d_pos, stream or None))

    def shard_route(self, d_offsets: int, d_containers: int, n_reads: int, n_shards: int, d_pos: int,
                    seg_base, dst, stream: int = 0):
        sb = (C.c_uint64 * n_shards)(*[int(x) for x in seg_base]) if seg_base is not None else None
        dp = (_VP * n_shards)(*[int(x) for x in dst])
        self._ck(self._lib.ccb_shard_route(self._h, d_offsets, d_containers, n_reads, n_shards, d_pos, sb, dp, stream or None))

    def shard_probe(self, d_hashes: int, n: int, d_labels: int, d_stats: int = 0, stream: int = 0):
        self._ck(self._lib.ccb_shard_probe(self._h, d_hashes or None, n, d_labels or None, d_stats or None, stream or None))

    def shard_vote(self, n_reads: int, n_shards: int, d_pos: int, seg_base, src, d_final5: int, d_rows: int = 0, stream: int = 0):
        sb = (C.c_uint64 * n_shards)(*[int(x) for x in seg_base]) if seg_base is not None else None
        sp = (_VP * n_shards)(*[int(x) for x in src])
        self._ck(self._lib.ccb_shard_vote(self._h, n_reads, n_shards, d_pos, sb, sp, d_final5, d_rows or None, stream or None))

    def shard_probe_inbox(self, n_shards: int, d_inbox: int, cap: int, d_counts_in: int, lab_dst, d_stats: int = 0, stream: int = 0):
        lp = (_VP * n_shards)(*[int(x) for x in lab_dst])
        self._ck(self._lib.ccb_shard_probe_inbox(self._h, n_shards, d_inbox, cap, d_counts_in, lp, d_stats or None, stream or None))

    # one-pass pipeline (shard_block.cu): a workspace per batch in flight
    def shardws_create(self, max_reads: int, n_shards: int, cap: int) -> int:
        p = _VP()
        self._ck(self._lib.ccb_shardws_create(self._h, max_reads, n_shards, cap, C.byref(p)))
        return p.value

    def shardws_destroy(self, ws: int):
        self._lib.ccb_shardws_destroy(self._h, ws)

    def shardws_error(self, ws: int, stream: int = 0) -> bool:
        v = C.c_int(0)
        self._ck(self._lib.ccb_shardws_error(self._h, ws, stream or None, C.byref(v)))
        return bool(v.value)

    def shard_route2(self, ws: int, d_offsets: int, d_containers: int, n_reads: int, n_containers: int, dst, count_dst, stream: int = 0):
        dp = (_VP * len(dst))(*[int(x) for x in dst])
        cp = (_VP * len(count_dst))(*[int(x) for x in count_dst])
        self._ck(self._lib.ccb_shard_route2(self._h, ws, d_offsets, d_containers, n_reads, n_containers, dp, cp, stream or None))

    def shard_vote2(self, ws: int, n_reads: int, src, d_final5: int, d_rows: int = 0, stream: int = 0):
        sp = (_VP * len(src))(*[int(x) for x in src])
        self._ck(self._lib.ccb_shard_vote2(self._h, ws, n_reads, sp, d_final5, d_rows or None, stream or None))

    def device_alloc(self, nbytes: int) -> int:
        p = _VP()
        self._ck(self._lib.ccb_device_alloc(self._h, nbytes, C.byref(p)))
        return p.value

    def ipc_export(self, ptr: int) -> bytes:
        buf = C.create_string_buffer(64)
        self._ck(self._lib.ccb_ipc_export(self._h, ptr, buf))
        return buf.raw

    def ipc_open(self, handle: bytes) -> int:
        p = _VP()
        self._ck(self._lib.ccb_ipc_open(self._h, handle, C.byref(p)))
        return p.value

    def ipc_close(self, ptr: int):
        self._ck(self._lib.ccb_ipc_close(self._h, ptr))

    # -- anchor index ---------------------------------------------------------------------------
    def anchor_synth(self, db_seed: int, n_genomes: int, genome_len: int, n_labels: int) -> AnchorInfo:
        self._ck(self._lib.ccb_anchor_synth(self._h, db_seed, n_genomes, genome_len, n_labels))
        return self.anchor_info()

    def anchor_from_fasta(self, texts) -> AnchorInfo:
        """texts: iterable of (FASTA bytes, 0-based label) -- the target files the database was built from."""
        texts = list(texts)
        self._ck(self._lib.ccb_anchor_begin(self._h, sum(len(t) for t, _ in texts)))
        for t, label in texts:
            self._ck(self._lib.ccb_anchor_add_fasta(self._h, t, len(t), label))
        info = AnchorInfo()
        self._ck(self._lib.ccb_anchor_end(self._h, C.byref(info)))
        return info

    def anchor_drop(self):
        self._ck(self._lib.ccb_anchor_drop(self._h))

    def anchor_info(self) -> AnchorInfo:
        info = AnchorInfo()
        self._ck(self._lib.ccb_anchor_get_info(self._h, C.byref(info)))
        return info

    def l2_fetch_granularity(self) -> int:
        return self._lib.ccb_l2_fetch_granularity(self._h)

    def stats(self) -> Stats:
        s = Stats()
        self._ck(self._lib.ccb_get_stats(self._h, C.byref(s)))
        return s

    # -- synthetic workload (include/ccb_synth.h) ------------------------------
    def synth_db(self, db_seed: int, n_genomes: int, genome_len: int, n_labels: int) -> int:
        n = C.c_uint64(0)
        self._ck(self._lib.ccb_synth_db(self._h, db_seed, n_genomes, genome_len, n_labels, C.byref(n)))
        return n.value

    def synth_reads_packed(self, reads_seed: int, db_seed: int, first: int, n_reads: int, read_len: int,
                           n_genomes: int, genome_len: int):
        """-> (d_offsets, d_containers, n_containers): raw device pointers owned by the caller (device_free)."""
        o, c, n = _VP(), _VP(), C.c_uint64(0)
        self._ck(self._lib.ccb_synth_reads_packed(self._h, reads_seed, db_seed, first, n_reads, read_len, n_genomes,
                                                  genome_len, C.byref(o), C.byref(c), C.byref(n)))
        return o.value, c.value, n.value

    def synth_db_light(self, db_seed: int, n_genomes: int, genome_len: int, n_labels: int, gap: int = 4) -> int:
        """Light database of the synthetic genomes: every gap-th non-overlapping k-mer (src/CuCLARK_hh.hh:711-767)."""
        n = C.c_uint64(0)
        self._ck(self._lib.ccb_synth_db_light(self._h, db_seed, n_genomes, genome_len, n_labels, gap, C.byref(n)))
        return n.value

    def synth2_reads_packed(self, w: "Workload", reads_seed: int, db_seed: int, first: int, n_reads: int, n_genomes: int, genome_len: int):
        """Reads of a ccb_workload (pairs joined by N / long reads with indels), packed on the device.
        -> (d_offsets, d_containers, n_containers): raw device pointers owned by the caller (device_free)."""
        o, c, n = _VP(), _VP(), C.c_uint64(0)
        self._ck(self._lib.ccb_synth2_reads_packed(self._h, C.byref(w), reads_seed, db_seed, first, n_reads, n_genomes, genome_len,
                                                   C.byref(o), C.byref(c), C.byref(n)))
        return o.value, c.value, n.value

    def synth_reads_text(self, reads_seed: int, db_seed: int, first: int, n_reads: int, read_len: int,
                         n_genomes: int, genome_len: int) -> np.ndarray:
        out = np.empty((n_reads, read_len + 1), dtype=np.uint8)
        self._ck(self._lib.ccb_synth_reads_text(self._h, reads_seed, db_seed, first, n_reads, read_len, n_genomes,
                                                genome_len, out.ctypes.data))
        return out

    def device_free(self, ptr: int):
        self._lib.ccb_device_free(self._h, ptr)

    def bench_gather(self, n_probes: int, nbytes: int = 32, stream: int = 0, mode: int = 0, footprint: int = 0):
        self._ck(self._lib.ccb_bench_gather_ex(self._h, n_probes, nbytes, mode, footprint, stream or None))

    # -- convenience used by tests/bench: one whole batch, synchronously ------
    def classify_packed(self, offsets: np.ndarray, containers: np.ndarray, extended: bool = False,
                        flags: int = 0, slot: int = 0):
        n = offsets.size - 1
        nc = int(offsets[-1])
        if n > self._max_reads or nc > getattr(self, "_max_cont", 0) or extended != getattr(self, "_ext", None):
            self.malloc(max(n, 1), max(nc, 1), extended)
        o, c, f, r = self.buffers(slot)
        o[: n + 1] = offsets
        c[:nc] = containers[:nc]
        self.readyBatch(slot, n, nc)
        self.queryBatch(slot, extended, flags)
        self.waitForBatch(slot)
        return f[:n].copy(), (r[:n].copy() if extended else None)


class CuClarkGroup(CuClarkDB):
    """One process, several GPUs, the table sharded over them (ccb_group_create): the leader handle takes the same
    batch calls as a one-device CuClarkDB (malloc / buffers / readyBatch / queryBatch / waitForBatch, classify_packed);
    `members[g]` is the handle that holds shard g (load it with load_pairs / synth_db, or read() on the leader loads all)."""

    def __init__(self, devices, k: int = 31, numBatches: int = 1, numTargets: int = 1, verbose: bool = False, *,
                 max_hits: int = 15, load_factor: float = 0.0):
        self._lib = load_library()
        self._h = _H()
        cfg = _Config(int(devices[0]), k, numTargets, max_hits, numBatches, load_factor, int(verbose))
        dv = (C.c_int * len(devices))(*[int(d) for d in devices])
        rc = self._lib.ccb_group_create(dv, len(devices), C.byref(cfg), C.byref(self._h))
        if rc:
            raise CcbError(rc, self._lib.ccb_last_error(None).decode())
        self.k, self.numTargets, self.numBatches, self.max_hits = k, numTargets, numBatches, max_hits
        self.row_stride = 2 * max_hits + 2
        self._max_reads = 0
        self._n = {}
        self.members = [CuClarkDB._view(self._lib, self._lib.ccb_group_member(self._h, g), k, numTargets, max_hits)
                        for g in range(self._lib.ccb_group_size(self._h))]
