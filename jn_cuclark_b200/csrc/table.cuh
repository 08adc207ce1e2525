// table.cuh -- device-resident table of discriminative k-mers.
//
// Replaces the reference's three-array layout (bucket pointers + keys + labels,
// src/CuClarkDB.cu:567-633,671-741) whose probe needs 2-4 dependent global
// loads (src/CuClarkDB.cu:1216-1247) by a bucketed open-addressing table in
// which a probe is ONE aligned 32-byte sector:
//
//   bucket  = 4 entries x 8 B, 32-byte aligned
//   entry   = key48 << 16 | label16         (kEmpty = all ones)
//   h       = mix2k(canonical)              (bijection on 2k bits)
//   home    = floor(h * nb / 2^2k), key48 = h & (2^48-1)
//
// Overflow: an entry whose home bucket is full goes to the next bucket that has
// room (at most kMaxDisp further; the table carries kPadBuckets extra buckets,
// there is no wrap-around).  Two flags per FULL bucket are encoded in the ORDER
// of its (pairwise distinct) entries, so they cost no bits:
//   A = e0 > e1 : some entry whose home is this bucket lives in a later bucket
//   B = e2 > e3 : some entry passes through this bucket (home < b < stored)
// A miss therefore stops after one sector unless the home bucket is full AND
// has A set.
#pragma once
#include <cuda_runtime.h>
#include "kmer.cuh"

namespace ccb {

struct TableView {
    const ulonglong4* buckets;   // indexed by GLOBAL home bucket: nb + kPadBuckets of them, or, for a
                                 // shard, a pointer moved back by the shard's first bucket
    HashParams        hp;
};

struct alignas(32) Bucket { uint64_t e[4]; };

// One 256-bit read-only load = one 32-byte sector.  Without a size hint an L2 miss fills the whole 128-byte line
// from HBM (126 B of DRAM reads per probe, ncu); with .L2::64B it fills 64 bytes (63 B per probe) at the same
// request rate -- profiles/r02_probe_fill_ncu.csv, DESIGN.md section 3.  What the rest of the fill is good for is the
// overflow bucket (the next one, 5.7 % of the lookups read it): a 64-byte fill holds it for even buckets, a 128-byte
// fill also for the second bucket of a line, nothing helps the fourth.  CCB_LD_HINT: 64 (default) = always 64 bytes
// (1.2 % slower than whole lines, half the DRAM bytes: profiles/r02_ld_hint_ab.json), 0 = no hint, 1 = the smallest fill
// that still covers the next bucket, chosen per load from the address (82 B per probe; the branch around two load
// instructions costs the latency-bound kernels 8-20 %: 13.9 instead of 12.5 ms with the anchor index -- rejected).
#ifndef CCB_LD_HINT
#define CCB_LD_HINT 64
#endif
__device__ __forceinline__ void load_bucket64(const ulonglong4* p, uint64_t (&e)[4])
{
    asm volatile("ld.global.nc.L1::no_allocate.L2::64B.v4.u64 {%0,%1,%2,%3}, [%4];"
                 : "=l"(e[0]), "=l"(e[1]), "=l"(e[2]), "=l"(e[3]) : "l"(p));
}
__device__ __forceinline__ void load_bucket128(const ulonglong4* p, uint64_t (&e)[4])
{
    asm volatile("ld.global.nc.L1::no_allocate.v4.u64 {%0,%1,%2,%3}, [%4];"
                 : "=l"(e[0]), "=l"(e[1]), "=l"(e[2]), "=l"(e[3]) : "l"(p));
}
__device__ __forceinline__ void load_bucket(const ulonglong4* p, uint64_t (&e)[4])
{
#if CCB_LD_HINT == 64
    load_bucket64(p, e);
#elif CCB_LD_HINT == 0
    load_bucket128(p, e);
#else
    if (((unsigned)(uintptr_t)p & 0x60u) == 0x20u) load_bucket128(p, e);     // second bucket of a line: the third comes along
    else load_bucket64(p, e);
#endif
}

// Match the 4 entries of a bucket against q = key48 << 16.
// returns label (0..0xFFFD), kLabelTomb, or kLabelNone when absent.
__device__ __forceinline__ uint32_t match_bucket(const uint64_t (&e)[4], uint64_t q)
{
    uint32_t lab = kLabelNone;
#pragma unroll
    for (int s = 0; s < 4; s++) {
        uint64_t d = e[s] ^ q;
        // key bits agree <=> d < 2^16, and then d is the label; an empty slot (all ones) can only agree with an
        // all-ones key, where d = 0xFFFF, the one label no entry carries
        if (d < 0xFFFFull) lab = (uint32_t)d;
    }
    return lab;
}

// Resolve a probe whose home bucket is already in registers (e); follows the
// overflow chain only when the flags demand it.  probes (optional) counts sectors.
__device__ __forceinline__ uint32_t resolve(const TableView& t, uint64_t home, uint64_t q,
                                            uint64_t (&e)[4], uint32_t* probes = nullptr)
{
    uint32_t lab = match_bucket(e, q);
    if (lab != kLabelNone) return lab;
    if (e[3] == kEmpty || !(e[0] > e[1])) return kLabelNone;
    for (int d = 1; d <= kMaxDisp; d++) {
        load_bucket(t.buckets + home + d, e);
        if (probes) (*probes)++;
        lab = match_bucket(e, q);
        if (lab != kLabelNone) return lab;
        if (e[3] == kEmpty || !(e[2] > e[3])) return kLabelNone;
    }
    return kLabelNone;
}

__device__ __forceinline__ uint32_t lookup(const TableView& t, uint64_t canon)
{
    uint64_t h = mix2k(canon, t.hp.k, t.hp.mask);
    uint64_t home = home_of(h, t.hp);
    uint64_t e[4];
    load_bucket(t.buckets + home, e);
    uint32_t lab = resolve(t, home, key48_of(h) << 16, e);
    return lab >= kLabelTomb ? kLabelNone : lab;
}

// ---- insertion (used by table.cu and by the synthetic-workload builder) -------
__device__ __forceinline__ uint64_t ld_volatile(const uint64_t* p)
{
    uint64_t v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p));
    return v;
}

// Insert (canonical, label).  Slots of a bucket fill front to back and never
// change key afterwards, so every inserter of the same key meets either the
// key itself or the same first empty slot: no duplicates under concurrency.
static __device__ __noinline__ void insert_one(uint64_t* buckets, uint32_t* flags, uint32_t* status,
                           unsigned long long* counts, const HashParams hp,
                           uint64_t canon, uint32_t label)
{
    const uint64_t h = mix2k(canon, hp.k, hp.mask);
    if (hp.n_shards > 1 && owner_of(h, hp.shift, hp.n_shards) != hp.shard) return;   // another shard's k-mer
    const uint64_t home = home_of(h, hp);
    const uint64_t q = key48_of(h) << 16;
    const uint64_t mine = q | label;
    for (int d = 0; d <= kMaxDisp; d++) {
        uint64_t* b = buckets + (home + d) * kSlots;
#pragma unroll 1
        for (int s = 0; s < kSlots; s++) {
            uint64_t e = ld_volatile(b + s);
            if (e == kEmpty) {
                e = atomicCAS((unsigned long long*)(b + s), (unsigned long long)kEmpty, (unsigned long long)mine);
                if (e == kEmpty) {
                    if (d) atomicMax(status + 1, (uint32_t)d);
                    atomicAdd(counts + 0, 1ull);
                    return;
                }
            }
            if ((e ^ q) < 0x10000ull) {                    // same k-mer already present
                uint32_t have = (uint32_t)(e & 0xFFFF);
                if (have == label || have == kLabelTomb) { atomicAdd(counts + 1, 1ull); return; }
                // seen under two labels: not discriminative (reference
                // HashTableStorage_hh.hh:244, Multiplicity != 1)
                uint64_t tomb = q | kLabelTomb;
                for (;;) {
                    uint64_t old = atomicCAS((unsigned long long*)(b + s), (unsigned long long)e, (unsigned long long)tomb);
                    if (old == e || (old & 0xFFFF) == kLabelTomb) break;
                    e = old;
                }
                atomicAdd(counts + 2, 1ull);
                return;
            }
        }
        // bucket full and key absent: leave a breadcrumb for the finalize pass
        uint64_t bi = home + d;
        atomicOr(flags + (bi >> 4), (d == 0 ? 1u : 2u) << (2 * (bi & 15)));
    }
    atomicExch(status + 0, 1u);                            // displacement limit exceeded
}

// ---- host-side owner --------------------------------------------------------
struct Table {
    Bucket*   d_buckets = nullptr;
    uint32_t* d_flags   = nullptr;   // 2 bits per bucket while building
    uint32_t* d_status  = nullptr;   // [0] error (displacement overflow) [1] max displacement seen
    unsigned long long* d_counts = nullptr; // [0] inserted new [1] duplicates same label [2] tombstoned
    uint64_t  nb = 0;                // home buckets of the whole table
    uint64_t  lo = 0;                // first home bucket held here (0 unless sharded)
    uint64_t  n_phys = 0;            // buckets held: nb (or nb / n_shards) + kPadBuckets
    HashParams hp{};
    bool      finalized = false;
    uint64_t  n_inserted = 0;        // entries occupying a slot (incl. tombstones)

    TableView view() const { return TableView{reinterpret_cast<const ulonglong4*>(d_buckets) - lo, hp}; }
    size_t bytes() const { return n_phys * sizeof(Bucket); }
};

// All return cudaError_t-compatible ints (0 = ok) or negative library codes.
// expected_entries counts the WHOLE table; shard g of n_shards allocates 1/n_shards of the buckets
int table_create(Table& t, int k, uint64_t expected_entries, double load_factor, cudaStream_t s,
                 uint32_t n_shards = 1, uint32_t shard = 0);
int table_insert_device(Table& t, const uint64_t* d_canon, const uint16_t* d_labels, uint64_t n, cudaStream_t s);
// reference on-disk layout chunk (A9): buckets [r0, r0+nbk) with their sizes as in the .sz file; keys and
// labels of the chunk in file order.  d_off / d_rank: nbk u32 each, d_scratch: table_ref_chunk_scratch().
// rank_base = non-empty buckets before r0 in the file; sampling = -s factor (<= 1: every bucket).
size_t table_ref_chunk_scratch(uint32_t max_buckets);
int table_insert_ref_chunk(Table& t, uint64_t htsize, uint64_t r0, uint32_t nbk, const uint8_t* d_sz,
                           uint32_t* d_off, uint32_t* d_rank, void* d_scratch, size_t scratch_bytes,
                           uint64_t rank_base, uint32_t sampling,
                           const void* d_keys, int key_bytes, const uint16_t* d_labels, cudaStream_t s);
int table_finalize(Table& t, cudaStream_t s);   // orders entries (flags), frees build scratch, checks status
void table_destroy(Table& t);

}  // namespace ccb
