"""ctypes binding of the CPU oracle (oracle/build/liboracle.so) -- test infrastructure only."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
HTSIZE_FULL, HTSIZE_LIGHT = 1610612741, 57777779


class Db(C.Structure):
    _fields_ = [("htsize", C.c_uint64), ("key_bytes", C.c_int), ("k", C.c_int), ("n", C.c_uint64),
                ("sz", C.POINTER(C.c_uint8)), ("keys", C.POINTER(C.c_uint64)),
                ("labels", C.POINTER(C.c_uint16)), ("ptr", C.POINTER(C.c_uint32))]


class Targets(C.Structure):
    _fields_ = [("n_files", C.c_int), ("files", C.POINTER(C.c_char_p)), ("file_label", C.POINTER(C.c_uint16)),
                ("n_labels", C.c_int), ("labels", C.POINTER(C.c_char_p))]


class Reads(C.Structure):
    _fields_ = [("n", C.c_size_t), ("name_s", C.POINTER(C.c_uint64)), ("name_e", C.POINTER(C.c_uint64)),
                ("seq_s", C.POINTER(C.c_uint64)), ("seq_e", C.POINTER(C.c_uint64)),
                ("len", C.POINTER(C.c_uint64)), ("is_fasta", C.c_int)]


class OccVec(C.Structure):
    _fields_ = [("v", C.c_void_p), ("n", C.c_size_t), ("cap", C.c_size_t)]


_lib = None


def build():
    subprocess.run(["make", "-C", ORACLE_DIR, "all"], check=True, capture_output=True)


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        p = os.path.join(ORACLE_DIR, "build", "liboracle.so")
        if not os.path.exists(p):
            build()
        L = C.CDLL(p)
        L.ccbo_revcomp.restype = C.c_uint64
        L.ccbo_revcomp.argtypes = [C.c_uint64, C.c_int]
        L.ccbo_canonical.restype = C.c_uint64
        L.ccbo_canonical.argtypes = [C.c_uint64, C.c_int]
        L.ccbo_rcode.argtypes = [C.c_ubyte]
        L.ccbo_key_bytes.argtypes = [C.c_int, C.c_uint64]
        L.ccbo_targets_read.argtypes = [C.c_char_p, C.POINTER(Targets)]
        L.ccbo_db_build.argtypes = [C.POINTER(Targets), C.c_int, C.c_uint64, C.c_int, C.c_uint, C.c_uint, C.POINTER(Db)]
        L.ccbo_db_write.argtypes = [C.POINTER(Db), C.c_char_p]
        L.ccbo_db_read.argtypes = [C.c_char_p, C.c_uint64, C.c_int, C.c_int, C.c_uint, C.POINTER(Db)]
        L.ccbo_db_free.argtypes = [C.POINTER(Db)]
        L.ccbo_db_find.argtypes = [C.POINTER(Db), C.c_uint64, C.POINTER(C.c_uint16)]
        L.ccbo_db_name.argtypes = [C.c_char_p, C.c_size_t, C.c_char_p, C.c_int, C.c_int, C.c_uint64, C.c_uint, C.c_int, C.c_uint]
        L.ccbo_index_reads.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(Reads)]
        L.ccbo_reads_free.argtypes = [C.POINTER(Reads)]
        L.ccbo_pack.restype = C.c_int64
        L.ccbo_pack.argtypes = [C.c_void_p, C.POINTER(Reads), C.c_size_t, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p, C.c_size_t]
        L.ccbo_query_packed.argtypes = [C.POINTER(Db), C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int,
                                        C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_uint64), C.c_int]
        L.ccbo_write_csv.argtypes = [C.c_char_p, C.c_void_p, C.POINTER(Reads), C.c_int, C.c_int, C.c_int, C.c_int,
                                     C.POINTER(C.c_char_p), C.c_void_p, C.c_void_p, C.c_int]
        L.ccbo_merge_paired.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p]
        L.ccbo_extract_fasta.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_uint, C.c_uint16, C.POINTER(OccVec)]
        L.ccbo_db_from_occ.argtypes = [C.POINTER(OccVec), C.c_int, C.c_uint64, C.c_uint, C.POINTER(Db)]
        L.ccbo_synth_db.argtypes = [C.c_uint64, C.c_uint32, C.c_uint64, C.c_uint32, C.c_int, C.c_uint64, C.POINTER(Db)]
        L.ccbo_synth_reads_fasta.argtypes = [C.c_char_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint64]
        L.ccbo_synth_db_ex.argtypes = [C.c_uint64, C.c_uint32, C.c_uint64, C.c_uint32, C.c_int, C.c_uint64, C.c_int, C.c_uint, C.POINTER(Db)]
        L.ccbo_synth2_reads_fasta.argtypes = [C.c_char_p, C.POINTER(Workload), C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint64]
        _lib = L
    return _lib


class Workload(C.Structure):
    """struct ccb_workload (include/ccb_synth.h): kind 1 pairs joined by N, kind 2 long reads with indels"""
    _fields_ = [("kind", C.c_uint32), ("read_len", C.c_uint32), ("insert", C.c_uint32),
                ("sub_pm", C.c_uint32), ("ins_pm", C.c_uint32), ("del_pm", C.c_uint32)]


def synth2_reads_fasta(path: str, w: "Workload", reads_seed: int, db_seed: int, first: int, n: int, n_genomes: int, genome_len: int):
    rc = lib().ccbo_synth2_reads_fasta(path.encode(), C.byref(w), reads_seed, db_seed, first, n, n_genomes, genome_len)
    assert rc == 0, rc


class OracleDb:
    """Owns a ccbo_db; exposes the canonical (k-mer, label) pairs as numpy arrays."""

    def __init__(self):
        self.db = Db()
        self.labels: list[str] = []

    @classmethod
    def build(cls, targets_path: str, k: int, htsize: int, light: bool = False, gap: int = 4, min_t: int = 0):
        self = cls()
        t = Targets()
        rc = lib().ccbo_targets_read(targets_path.encode(), C.byref(t))
        assert rc == 0, rc
        self.labels = [t.labels[i].decode() for i in range(t.n_labels)]
        rc = lib().ccbo_db_build(C.byref(t), k, htsize, int(light), gap if light else 0, min_t, C.byref(self.db))
        assert rc == 0, rc
        return self

    @classmethod
    def from_pairs(cls, canon: np.ndarray, labels: np.ndarray, k: int, htsize: int, min_t: int = 0):
        """Database from explicit (canonical, label) occurrences (duplicates allowed)."""
        self = cls()
        n = canon.size
        occ = np.zeros(n, dtype=[("c", np.uint64), ("label", np.uint16), ("pad", np.uint8, 6)])
        occ["c"] = canon
        occ["label"] = labels
        libc = C.CDLL(None)
        libc.malloc.restype = C.c_void_p
        libc.malloc.argtypes = [C.c_size_t]
        p = libc.malloc(max(occ.nbytes, 16))
        C.memmove(p, occ.ctypes.data, occ.nbytes)
        v = OccVec(p, n, n)
        rc = lib().ccbo_db_from_occ(C.byref(v), k, htsize, min_t, C.byref(self.db))
        assert rc == 0, rc
        return self

    @classmethod
    def synth(cls, db_seed: int, n_genomes: int, genome_len: int, n_labels: int, k: int, htsize: int, light_gap: int = 0):
        self = cls()
        rc = lib().ccbo_synth_db_ex(db_seed, n_genomes, genome_len, n_labels, k, htsize, 1 if light_gap else 0, light_gap, C.byref(self.db))
        assert rc == 0, rc
        return self

    @classmethod
    def read(cls, prefix: str, htsize: int, k: int, sampling: int = 1):
        self = cls()
        kb = lib().ccbo_key_bytes(k, htsize)
        rc = lib().ccbo_db_read(prefix.encode(), htsize, kb, k, sampling, C.byref(self.db))
        assert rc == 0, rc
        return self

    def write(self, prefix: str):
        assert lib().ccbo_db_write(C.byref(self.db), prefix.encode()) == 0

    @property
    def n(self) -> int:
        return int(self.db.n)

    def pairs(self):
        """(canonical u64[n], label u16[n]): canonical = key*htsize + bucket (A9)."""
        n, ht = self.n, int(self.db.htsize)
        if n == 0:
            return np.zeros(0, np.uint64), np.zeros(0, np.uint16)
        sz = np.ctypeslib.as_array(self.db.sz, shape=(ht,))
        keys = np.ctypeslib.as_array(self.db.keys, shape=(n,))
        labels = np.ctypeslib.as_array(self.db.labels, shape=(n,)).copy()
        nz = np.flatnonzero(sz)
        buckets = np.repeat(nz.astype(np.uint64), sz[nz])
        return keys * np.uint64(ht) + buckets, labels

    def find(self, x: int):
        lb = C.c_uint16(0)
        hit = lib().ccbo_db_find(C.byref(self.db), x, C.byref(lb))
        return (lb.value if hit else None)

    def query_packed(self, k: int, offsets: np.ndarray, containers: np.ndarray, row_stride: int = 0, threads: int = 1):
        n = offsets.size - 1
        out5 = np.zeros((n, 5), np.uint16)
        rows = np.zeros((n, row_stride), np.uint16) if row_stride else None
        nl = C.c_uint64(0)
        offsets = np.ascontiguousarray(offsets, np.uint32)
        containers = np.ascontiguousarray(containers, np.uint16)
        lib().ccbo_query_packed(C.byref(self.db), k, offsets.ctypes.data, containers.ctypes.data, n, 0,
                                out5.ctypes.data, rows.ctypes.data if rows is not None else None, row_stride,
                                C.byref(nl), threads)
        return out5, rows, nl.value

    def __del__(self):
        try:
            lib().ccbo_db_free(C.byref(self.db))
        except Exception:
            pass


def key_bytes(k: int, htsize: int) -> int:
    """src/main.cc:251-286: width of the stored keys (T16/T32/T64)."""
    return int(lib().ccbo_key_bytes(k, htsize))


def db_name(folder: str, k: int, n_labels: int, htsize: int, min_t: int = 0, light: bool = False, gap: int = 4) -> str:
    buf = C.create_string_buffer(4400)
    lib().ccbo_db_name(buf, 4400, folder.encode(), k, n_labels, htsize, min_t, int(light), gap)
    return buf.value.decode()


class IndexedReads:
    def __init__(self, data: bytes):
        self.data = np.frombuffer(data, dtype=np.uint8)
        self.r = Reads()
        rc = lib().ccbo_index_reads(self.data.ctypes.data, self.data.size, C.byref(self.r))
        if rc:
            raise ValueError("Failed to recognize the format of the file.")
        self.n = int(self.r.n)

    def arr(self, name):
        return np.ctypeslib.as_array(getattr(self.r, name), shape=(self.n,)).copy() if self.n else np.zeros(0, np.uint64)

    def pack(self, k: int, first: int = 0, count: int | None = None):
        count = self.n - first if count is None else count
        cap = self.data.size // 4 + 16 * count + 64
        off = np.zeros(count + 1, np.uint32)
        cont = np.zeros(cap, np.uint16)
        nc = lib().ccbo_pack(self.data.ctypes.data, C.byref(self.r), first, count, k, off.ctypes.data, cont.ctypes.data, cap)
        assert nc >= 0, nc
        return off, cont[:nc].copy()

    def write_csv(self, path: str, k: int, paired: bool, names: list[str], out5: np.ndarray,
                  rows: np.ndarray | None = None):
        arr = (C.c_char_p * len(names))(*[s.encode() for s in names])
        out5 = np.ascontiguousarray(out5, np.uint16)
        ext = rows is not None
        if ext:
            rows = np.ascontiguousarray(rows, np.uint16)
        rc = lib().ccbo_write_csv(path.encode(), self.data.ctypes.data, C.byref(self.r), k, int(paired), int(ext),
                                  len(names) - 1, arr, out5.ctypes.data, rows.ctypes.data if ext else None,
                                  rows.shape[1] if ext else 0)
        assert rc == 0

    def __del__(self):
        try:
            lib().ccbo_reads_free(C.byref(self.r))
        except Exception:
            pass


def merge_paired(f1: str, f2: str, out: str) -> int:
    return lib().ccbo_merge_paired(f1.encode(), f2.encode(), out.encode())


def canonical(x: int, k: int) -> int:
    return lib().ccbo_canonical(x, k)


def revcomp(x: int, k: int) -> int:
    return lib().ccbo_revcomp(x, k)
