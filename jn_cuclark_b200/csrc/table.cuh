// table.cuh -- device-resident table of discriminative k-mers (layout v2).
//
// Replaces the reference's three-array layout (bucket pointers + keys + labels,
// src/CuClarkDB.cu:567-633,671-741) whose probe needs 2-4 dependent global
// loads (src/CuClarkDB.cu:1216-1247).
//
// Measured on B200: a load that misses L2 costs one
// 128-byte line request whatever part of the line is wanted, and ~45 G such
// requests/s is all the part delivers.  The table is therefore organised in
// 128-byte LINES addressed by a MINIMIZER of the k-mer, so that the 6-12
// consecutive k-mers of a read that share a minimizer are answered by one line:
//
//   line    = 4 sectors x 32 B, 128-byte aligned
//   sector  = 3 keys (u64 canonical k-mer, all ones = empty)
//             + 3 labels (u16; 0xFFFF unset, 0xFFFE = seen under two labels) + u16 meta
//   home    = 4 * line_of(min over the k-mer's m-mers of phi64(canonical m-mer)) + sector_of(c)
//
// Keys are stored whole, so membership is exact.  A k-mer lives in any of the 12
// slots of its home line (insertion starts at sector_of(c) to spread contention);
// when the line is full it goes to hop_line(home, 1), hop_line(home, 2), ... (all
// k-mers of a line hop together, so a run still needs one line per hop).  The
// 4 x 16 meta bits of a line are a Bloom filter of the keys that had to leave it
// (bit leave_bit(c)): a lookup that does not find its key in a line goes on only
// if its bit is set, so a miss usually costs one line.
#pragma once
#include <cuda_runtime.h>
#include "kmer.cuh"

namespace ccb {

struct TableView {
    const ulonglong4* sectors;   // 4 * n_lines
    HashParams        hp;
};

struct alignas(32) Sector { uint64_t key[3]; uint64_t lm; };   // lm = lab0 | lab1<<16 | lab2<<32 | meta<<48
constexpr uint64_t kLmInit = 0x0000FFFFFFFFFFFFull;
// bit (0..63) of key c in the line's filter of leavers: sector bit>>4, meta bit bit&15
__device__ __forceinline__ uint32_t leave_bit(uint64_t c) { return (uint32_t)((c * 0xD6E8FEB86659FD93ull) >> 58); }

__device__ __forceinline__ void load_bucket(const ulonglong4* p, uint64_t (&e)[4])
{
    // one 256-bit read-only load = one 32-byte sector
    asm volatile("ld.global.nc.L1::no_allocate.v4.u64 {%0,%1,%2,%3}, [%4];"
                 : "=l"(e[0]), "=l"(e[1]), "=l"(e[2]), "=l"(e[3]) : "l"(p));
}

// keys e[0..2], label word e[3]: label of c, or kLabelNone
__device__ __forceinline__ uint32_t match_sector(const uint64_t (&e)[4], uint64_t c)
{
    uint32_t lab = kLabelNone;
    if (e[0] == c) lab = (uint32_t)e[3] & 0xFFFFu;
    if (e[1] == c) lab = (uint32_t)(e[3] >> 16) & 0xFFFFu;
    if (e[2] == c) lab = (uint32_t)(e[3] >> 32) & 0xFFFFu;
    return lab;
}

__device__ __forceinline__ uint32_t probe_line(const HashParams& hp, uint32_t home, int h)
{
    return h == 0 ? home : hop_line(home, h, hp.n_lines);
}

// Walk the probe sequence of c in global memory from hop h0.  Returns the raw label
// (target, kLabelTomb) or kLabelNone.  probes (optional) counts sectors read.
static __device__ __noinline__ uint32_t probe_chain(const TableView& t, uint32_t home, uint64_t c, int h0, uint32_t* probes)
{
    const uint32_t lb = leave_bit(c);
    for (int h = h0; h < kMaxHops; h++) {
        const ulonglong4* L = t.sectors + 4ull * probe_line(t.hp, home, h);
        uint64_t e[4][4];
#pragma unroll
        for (int q = 0; q < 4; q++) load_bucket(L + q, e[q]);
        if (probes) (*probes) += 4;
        uint32_t lab = kLabelNone; uint64_t meta = 0;
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const uint32_t l = match_sector(e[q], c);
            if (l != kLabelNone) lab = l;
            if ((lb >> 4) == (uint32_t)q) meta = e[q][3] >> 48;
        }
        if (lab != kLabelNone) return lab;
        if (!((meta >> (lb & 15u)) & 1ull)) return kLabelNone;
    }
    return kLabelNone;
}

__device__ __forceinline__ uint32_t home_line(const HashParams& hp, uint64_t canon)
{
    return line_of(minimizer_phi(canon, hp.k, hp.m), hp.n_lines);
}

// generic lookup (dense path, tests): target label or kLabelNone
__device__ __forceinline__ uint32_t lookup(const TableView& t, uint64_t canon)
{
    uint32_t lab = probe_chain(t, home_line(t.hp, canon), canon, 0, nullptr);
    return lab >= kLabelTomb ? kLabelNone : lab;
}

// ---- insertion (used by table.cu and by the synthetic-workload builder) -------
__device__ __forceinline__ uint64_t ld_volatile(const uint64_t* p)
{
    uint64_t v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p));
    return v;
}

// Label of slot j of a sector: unset -> mine; equal -> keep; different -> tombstone.  Every inserter
// of the key runs this, so the outcome does not depend on their order (reference rule:
// HashTableStorage_hh.hh:244, a k-mer is kept iff it was seen under exactly one label).
__device__ __forceinline__ void merge_label(uint64_t* lm, int j, uint32_t label, unsigned long long* counts)
{
    const int sh = 16 * j;
    uint64_t cur = ld_volatile(lm);
    for (;;) {
        const uint32_t have = (uint32_t)(cur >> sh) & 0xFFFFu;
        if (have == label || have == kLabelTomb) { atomicAdd(counts + 1, 1ull); return; }
        const uint32_t want = have == kLabelNone ? label : kLabelTomb;
        const uint64_t nxt = (cur & ~(0xFFFFull << sh)) | ((uint64_t)want << sh);
        const uint64_t old = atomicCAS((unsigned long long*)lm, (unsigned long long)cur, (unsigned long long)nxt);
        if (old == cur) { if (want == kLabelTomb) atomicAdd(counts + 2, 1ull); return; }
        cur = old;
    }
}

// Insert (canonical, label).  Key slots of a sector fill front to back and never change afterwards,
// so every inserter of the same key meets either the key itself or the same first empty slot: no
// duplicates under concurrency.
static __device__ __noinline__ void insert_one(uint64_t* sectors, uint32_t* status, unsigned long long* counts,
                                               const HashParams hp, uint64_t canon, uint32_t label)
{
    const uint32_t home = home_line(hp, canon), s0 = sector_of(canon), lb = leave_bit(canon);
    for (int h = 0; h < kMaxHops; h++) {
        uint64_t* line = sectors + 16ull * probe_line(hp, home, h);
        for (int q = 0; q < 4; q++) {
            uint64_t* sec = line + 4 * ((s0 + q) & 3u);
#pragma unroll 1
            for (int j = 0; j < kSectorKeys; j++) {
                uint64_t e = ld_volatile(sec + j);
                if (e == kEmpty) {
                    e = atomicCAS((unsigned long long*)(sec + j), (unsigned long long)kEmpty, (unsigned long long)canon);
                    if (e == kEmpty) {
                        if (h) { atomicMax(status + 1, (uint32_t)h); atomicAdd(counts + 3, 1ull); }   // lives outside its home line
                        atomicAdd(counts + 0, 1ull);
                        e = canon;
                    }
                }
                if (e == canon) { merge_label(sec + 3, j, label, counts); return; }
            }
        }
        // line full and key absent: later lookups of this key have to go on from here
        uint64_t* lm = line + 4 * (lb >> 4) + 3;
        const uint64_t bit = 1ull << (48 + (lb & 15u));
        if (!(ld_volatile(lm) & bit)) atomicOr((unsigned long long*)lm, (unsigned long long)bit);
    }
    atomicExch(status + 0, 1u);                            // probe sequence exhausted
}

// ---- host-side owner --------------------------------------------------------
struct Table {
    Sector*   d_sectors = nullptr;
    uint32_t* d_status  = nullptr;   // [0] error (chain exhausted) [1] max displacement seen
    unsigned long long* d_counts = nullptr; // [0] inserted new [1] repeats (same label / already tomb) [2] tombstoned [3] outside home line
    uint64_t  n_home = 0;            // home sectors = 4 * n_lines
    uint64_t  n_phys = 0;            // = n_home
    uint64_t  n_hopped = 0;          // keys stored outside their home line
    HashParams hp{};
    bool      finalized = false;
    uint64_t  n_inserted = 0;        // keys occupying a slot (incl. tombstones)

    TableView view() const { return TableView{reinterpret_cast<const ulonglong4*>(d_sectors), hp}; }
    size_t bytes() const { return n_phys * sizeof(Sector); }
};

// All return cudaError_t-compatible ints (0 = ok) or negative library codes.
int table_create(Table& t, int k, uint64_t expected_entries, double load_factor, cudaStream_t s);
int table_insert_device(Table& t, const uint64_t* d_canon, const uint16_t* d_labels, uint64_t n, cudaStream_t s);
// reference on-disk layout chunk (A9): buckets [r0, r0+nbk) with per-bucket
// size (0 when skipped by -s sampling) and key offset relative to d_keys.
int table_insert_ref_chunk(Table& t, uint64_t htsize, uint64_t r0, uint32_t nbk,
                           const uint8_t* d_sz, const uint32_t* d_off,
                           const void* d_keys, int key_bytes, const uint16_t* d_labels, cudaStream_t s);
int table_finalize(Table& t, cudaStream_t s);   // reads back status and counts
void table_destroy(Table& t);

}  // namespace ccb
