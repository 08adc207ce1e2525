"""ctypes binding of the host-side C ABI (include/ccb_host.h): read indexing, packing, paired-end
join and CSV formatting exactly as bin/cuCLARK[-l] use them.  Compiled C++ inside
libcuclark_b200.so; nothing here computes in Python."""
from __future__ import annotations

import ctypes as C

import numpy as np

from .lib import load_library

_VP = C.c_void_p
HOST_SYMBOLS = {
    "ccb_host_index": (_VP, [_VP, C.c_size_t, C.c_int]),
    "ccb_host_index_fastq_chunked": (C.c_int, [_VP, C.c_size_t, C.c_size_t, C.POINTER(_VP)]),
    "ccb_host_reads_free": (None, [_VP]),
    "ccb_host_reads_count": (C.c_size_t, [_VP]),
    "ccb_host_reads_is_fasta": (C.c_int, [_VP]),
    "ccb_host_reads_field": (C.POINTER(C.c_uint64), [_VP, C.c_int]),
    "ccb_host_pack_bound": (C.c_size_t, [_VP, C.c_size_t, C.c_size_t, C.c_int]),
    "ccb_host_pack": (C.c_int64, [_VP, C.c_size_t, C.c_size_t, C.c_int, _VP, _VP, C.c_size_t]),
    "ccb_host_join_pairs": (C.c_int, [C.c_char_p, C.c_char_p, C.POINTER(_VP), C.POINTER(C.c_size_t), C.c_char_p, C.c_size_t]),
    "ccb_host_csv": (C.c_int, [_VP, C.c_size_t, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_char_p), C.c_int,
                               _VP, _VP, C.c_int, C.POINTER(_VP), C.POINTER(C.c_size_t)]),
    "ccb_host_free": (None, [_VP]),
    "ccb_host_numa_probe": (C.c_int, [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "ccb_host_numa_node_cpus": (C.c_int, [C.c_int]),
}
_bound = False


def _lib():
    global _bound
    L = load_library()
    if not _bound:
        for name, (res, args) in HOST_SYMBOLS.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _bound = True
    return L


FIELDS = ("name_s", "name_e", "seq_s", "seq_e", "len")


class HostReads:
    """Index of one FASTA/FASTQ text (reference src/CuCLARK_hh.hh:1340-1534)."""

    def __init__(self, data: bytes, threads: int = 1):
        self.data = np.frombuffer(data, dtype=np.uint8)
        self._L = _lib()
        self._r = self._L.ccb_host_index(self.data.ctypes.data, self.data.size, threads)
        if not self._r:
            raise ValueError("Failed to recognize the format of the file.")
        self.n = self._L.ccb_host_reads_count(self._r)
        self.is_fasta = bool(self._L.ccb_host_reads_is_fasta(self._r))

    def arr(self, name: str) -> np.ndarray:
        p = self._L.ccb_host_reads_field(self._r, FIELDS.index(name))
        return np.ctypeslib.as_array(p, shape=(self.n,)).copy() if self.n else np.zeros(0, np.uint64)

    def pack(self, k: int, first: int = 0, count: int | None = None):
        count = self.n - first if count is None else count
        cap = self._L.ccb_host_pack_bound(self._r, first, count, k)
        off = np.zeros(count + 1, np.uint32)
        cont = np.zeros(cap, np.uint16)
        nc = self._L.ccb_host_pack(self._r, first, count, k, off.ctypes.data, cont.ctypes.data, cap)
        if nc < 0:
            raise RuntimeError(f"ccb_host_pack failed ({nc})")
        return off, cont[:nc].copy()

    def csv(self, k: int, paired: bool, names: list[str], out5: np.ndarray, rows: np.ndarray | None = None,
            header: bool = True) -> bytes:
        arr = (C.c_char_p * len(names))(*[s.encode() for s in names])
        out5 = np.ascontiguousarray(out5, np.uint16)
        ext = rows is not None
        if ext:
            rows = np.ascontiguousarray(rows, np.uint16)
        out, n = _VP(), C.c_size_t(0)
        rc = self._L.ccb_host_csv(self._r, 0, self.n, k, int(paired), int(ext), int(header), arr, len(names),
                                  out5.ctypes.data, rows.ctypes.data if ext else None, rows.shape[1] if ext else 0,
                                  C.byref(out), C.byref(n))
        if rc:
            raise RuntimeError(f"ccb_host_csv failed ({rc})")
        s = C.string_at(out.value, n.value)
        self._L.ccb_host_free(out)
        return s

    def __del__(self):
        try:
            if self._r:
                self._L.ccb_host_reads_free(self._r)
                self._r = None
        except Exception:
            pass


def join_pairs(f1: str, f2: str) -> bytes:
    """src/file.cc:205-268 in memory: b'>id\\nR1NR2\\n' per pair."""
    L = _lib()
    out, n = _VP(), C.c_size_t(0)
    err = C.create_string_buffer(512)
    rc = L.ccb_host_join_pairs(f1.encode(), f2.encode(), C.byref(out), C.byref(n), err, 512)
    if rc:
        raise RuntimeError(err.value.decode())
    s = C.string_at(out.value, n.value)
    L.ccb_host_free(out)
    return s


def numa_probe(pci_bus_id: str):
    """-> (node, cpus while bound, cpus after): what ccb_batch_alloc's binding to the GPU's memory node does to this thread."""
    L = _lib()
    a, b = C.c_int(0), C.c_int(0)
    node = L.ccb_host_numa_probe(pci_bus_id.encode(), C.byref(a), C.byref(b))
    return node, a.value, b.value


def numa_node_cpus(node: int) -> int:
    return _lib().ccb_host_numa_node_cpus(node)


def index_fastq_chunked(data: bytes, chunk: int):
    """The FASTQ index as the streaming pipeline finds it with the text cut into chunks of `chunk` bytes
    (ccb_host_index_fastq_chunked): -> dict of the five field arrays, or None when the pieces do not fit together (such a
    file goes through the two-phase path)."""
    L = _lib()
    buf = np.frombuffer(data, dtype=np.uint8)
    r = _VP()
    rc = L.ccb_host_index_fastq_chunked(buf.ctypes.data, buf.size, chunk, C.byref(r))
    if rc == 1:
        return None
    if rc:
        raise ValueError("not a FASTQ text")
    try:
        n = L.ccb_host_reads_count(r)
        out = {}
        for i, name in enumerate(FIELDS):
            p = L.ccb_host_reads_field(r, i)
            out[name] = np.ctypeslib.as_array(p, shape=(n,)).copy() if n else np.zeros(0, np.uint64)
        return out
    finally:
        L.ccb_host_reads_free(r)
