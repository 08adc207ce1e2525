// kmer.cuh -- k-mer arithmetic shared by host and device code.
//
// Encoding follows the reference so that canonical values are identical to the
// ones stored in its database files:
//   * x-encoding A=3 C=2 G=1 T/U=0, first base most significant
//     (reference src/CuCLARK_hh.hh:294-297, src/CuClarkDB.cu:1063-1083)
//   * reverse complement R(x) and canonical c = min(x, R(x))
//     (reference src/CuClarkDB.cu:1195-1206)
// Everything after the canonical value (hash, bucket, stored key) is this
// implementation's own table layout, see DESIGN.md "Table".
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define CCB_HD __host__ __device__ __forceinline__
#else
#define CCB_HD inline
#endif

namespace ccb {

// 64-bit entry = key48 << 16 | label.  Labels 0..0xFFFD are targets.
constexpr uint64_t kEmpty     = ~0ull;     // never a valid entry: label 0xFFFF is reserved
constexpr uint32_t kLabelTomb = 0xFFFEu;   // k-mer seen under >1 label (not discriminative)
constexpr uint32_t kLabelNone = 0xFFFFu;
constexpr int      kSlots     = 4;         // 4 x 8 B = one 32-byte DRAM sector
constexpr int      kMaxDisp   = 62;        // max buckets an entry may sit past its home
constexpr int      kPadBuckets = 64;       // physical buckets after the last home bucket (no wrap)

CCB_HD int code_of(unsigned char c)
{   // >=0: x-encoding code; -1: anything else
    switch (c) {
    case 'A': case 'a': return 3;
    case 'C': case 'c': return 2;
    case 'G': case 'g': return 1;
    case 'T': case 't': case 'U': case 'u': return 0;
    default: return -1;
    }
}

CCB_HD uint64_t kmask(int k) { return k >= 32 ? ~0ull : ((1ull << (2 * k)) - 1); }

CCB_HD uint64_t revcomp(uint64_t x, int k)
{
    uint64_t r = x;
    r = ((r >> 2)  & 0x3333333333333333ull) | ((r & 0x3333333333333333ull) << 2);
    r = ((r >> 4)  & 0x0F0F0F0F0F0F0F0Full) | ((r & 0x0F0F0F0F0F0F0F0Full) << 4);
    r = ((r >> 8)  & 0x00FF00FF00FF00FFull) | ((r & 0x00FF00FF00FF00FFull) << 8);
    r = ((r >> 16) & 0x0000FFFF0000FFFFull) | ((r & 0x0000FFFF0000FFFFull) << 16);
    r = (r >> 32) | (r << 32);
    return (~r) >> (64 - 2 * k);
}

CCB_HD uint64_t canonical(uint64_t x, int k)
{
    uint64_t r = revcomp(x, k);
    return x < r ? x : r;
}

// Bijection on [0, 2^(2k)): xorshift and odd multiplies, all reduced mod 2^(2k).
// Being invertible is what lets the table drop the bits implied by the bucket.
CCB_HD uint64_t mix2k(uint64_t c, int k, uint64_t mask)
{
    uint64_t h = c;
    h ^= h >> k;
    h = (h * 0xff51afd7ed558ccdull) & mask;
    h ^= h >> k;
    h = (h * 0xc4ceb9fe1a85ec53ull) & mask;
    h ^= h >> k;
    return h;
}

struct HashParams {
    uint64_t nb;      // number of home buckets (of the WHOLE table, all shards)
    uint64_t mask;    // 2^(2k)-1
    int      k;
    int      shift;   // 64 - 2k
    // sharded table (DESIGN.md "Sharded mode"): the hash space is cut into n_shards equal ranges,
    // nb is a multiple of n_shards, so shard g holds the home buckets [g*nb/n_shards, (g+1)*nb/n_shards)
    uint32_t n_shards;   // 0 or 1: not sharded
    uint32_t shard;
};

#if defined(__CUDACC__)
__device__ __forceinline__ uint64_t mulhi64(uint64_t a, uint64_t b) { return __umul64hi(a, b); }
// revcomp() with the five swap stages replaced by two bit reversals and one swap inside the pairs (14 instructions)
__device__ __forceinline__ uint64_t revcomp_brev(uint64_t x, int k)
{
    uint32_t lo = __brev((uint32_t)(x >> 32)), hi = __brev((uint32_t)x);
    lo = ((lo >> 1) & 0x55555555u) | ((lo & 0x55555555u) << 1);
    hi = ((hi >> 1) & 0x55555555u) | ((hi & 0x55555555u) << 1);
    return (~(((uint64_t)hi << 32) | lo)) >> (64 - 2 * k);
}
#endif
inline uint64_t mulhi64_host(uint64_t a, uint64_t b) { return (uint64_t)(((unsigned __int128)a * b) >> 64); }

// home bucket = floor(h * nb / 2^(2k)); stored key = low 48 bits of h.
// (home, key48) determines h as long as nb >= 64 * 2^max(0,2k-48) and no entry
// sits more than kMaxDisp buckets past its home (DESIGN.md "Table").
CCB_HD uint64_t home_of(uint64_t h, const HashParams& p)
{
#if defined(__CUDA_ARCH__)
    return mulhi64(h << p.shift, p.nb);
#else
    return mulhi64_host(h << p.shift, p.nb);
#endif
}
// owner shard of a mixed k-mer: floor(h * n_shards / 2^(2k)) == home / (nb / n_shards)
CCB_HD uint32_t owner_of(uint64_t h, int shift, uint32_t n_shards)
{
#if defined(__CUDA_ARCH__)
    return (uint32_t)mulhi64(h << shift, (uint64_t)n_shards);
#else
    return (uint32_t)mulhi64_host(h << shift, (uint64_t)n_shards);
#endif
}
CCB_HD uint64_t key48_of(uint64_t h) { return h & 0xFFFFFFFFFFFFull; }

inline uint64_t min_buckets_for_k(int k)
{
    int extra = 2 * k - 48;
    return extra > 0 ? (64ull << extra) : 64ull;
}

}  // namespace ccb
