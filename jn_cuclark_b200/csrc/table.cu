// table.cu -- building the bucketed table on the device (see table.cuh).
#include "table.cuh"
#include <cub/device/device_scan.cuh>
#include <thrust/iterator/transform_iterator.h>
#include <stdio.h>

namespace ccb {

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return (int)e_; } while (0)

__global__ void k_fill_empty(ulonglong4* b, uint64_t n)
{
    uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    uint64_t stride = gridDim.x * (uint64_t)blockDim.x;
    const ulonglong4 v = make_ulonglong4(kEmpty, kEmpty, kEmpty, kEmpty);
    for (; i < n; i += stride) b[i] = v;
}

__global__ void k_insert_pairs(uint64_t* buckets, uint32_t* flags, uint32_t* status,
                               unsigned long long* counts, HashParams hp,
                               const uint64_t* canon, const uint16_t* labels, uint64_t n)
{
    uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    uint64_t stride = gridDim.x * (uint64_t)blockDim.x;
    for (; i < n; i += stride) insert_one(buckets, flags, status, counts, hp, canon[i], labels[i]);
}

// Reference layout (A9): canonical = key * HTSIZE + bucket (src/hashTable_hh.hh:469).  off = exclusive
// scan of the chunk's bucket sizes, nzrank = inclusive scan of (size != 0); with -s sampling only every
// sampling-th non-empty bucket of the whole file is kept (src/CuClarkDB.cu:503-513).
template <typename KEY>
__global__ void k_insert_ref_chunk(uint64_t* buckets, uint32_t* flags, uint32_t* status,
                                   unsigned long long* counts, HashParams hp,
                                   uint64_t htsize, uint64_t r0, uint32_t nbk,
                                   const uint8_t* sz, const uint32_t* off, const uint32_t* nzrank,
                                   uint64_t rank_base, uint32_t sampling,
                                   const KEY* keys, const uint16_t* labels)
{
    uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    uint64_t stride = gridDim.x * (uint64_t)blockDim.x;
    for (; i < nbk; i += stride) {
        uint32_t n = sz[i];
        if (n == 0) continue;
        if (sampling > 1 && (rank_base + nzrank[i]) % sampling != 0) continue;
        uint32_t o = off[i];
        for (uint32_t j = 0; j < n; j++) {
            uint64_t c = (uint64_t)keys[o + j] * htsize + (r0 + i);
            insert_one(buckets, flags, status, counts, hp, c, labels[o + j]);
        }
    }
}

struct NonZero { __host__ __device__ uint32_t operator()(uint8_t v) const { return v ? 1u : 0u; } };
struct Widen   { __host__ __device__ uint32_t operator()(uint8_t v) const { return v; } };

// Order the entries of every full bucket so that (e0 > e1) == A, (e2 > e3) == B.
__global__ void k_finalize(ulonglong4* b, const uint32_t* flags, uint64_t n)
{
    uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    uint64_t stride = gridDim.x * (uint64_t)blockDim.x;
    for (; i < n; i += stride) {
        ulonglong4 v = b[i];
        if (v.w == kEmpty) continue;                         // not full: flags are never set
        uint32_t f = (flags[i >> 4] >> (2 * (i & 15))) & 3u;
        bool A = f & 1u, B = f & 2u;
        if ((v.x > v.y) != A) { unsigned long long t = v.x; v.x = v.y; v.y = t; }
        if ((v.z > v.w) != B) { unsigned long long t = v.z; v.z = v.w; v.w = t; }
        b[i] = v;
    }
}

static int grid_for(uint64_t n, int block)
{
    uint64_t g = (n + block - 1) / block;
    uint64_t cap = 148ull * 32;
    if (g > cap) g = cap;
    if (g == 0) g = 1;
    return (int)g;
}

int table_create(Table& t, int k, uint64_t expected_entries, double load_factor, cudaStream_t s,
                 uint32_t n_shards, uint32_t shard)
{
    if (n_shards == 0) n_shards = 1;
    if (shard >= n_shards) return -10;
    if (k < 2 || k > 32) return -10;
    if (!(load_factor > 0.0 && load_factor <= 0.95)) load_factor = 0.5;
    uint64_t nb = (uint64_t)((double)expected_entries / (kSlots * load_factor)) + 1;
    uint64_t mn = min_buckets_for_k(k);
    if (nb < mn) nb = mn;
    const uint64_t unit = 16ull * n_shards;               // shard boundaries on multiples of 16 buckets (flag words)
    nb = (nb + unit - 1) / unit * unit;
    const uint64_t per = nb / n_shards;
    t.nb = nb; t.lo = per * shard;
    t.n_phys = per + kPadBuckets;
    t.hp.nb = nb; t.hp.k = k; t.hp.shift = 64 - 2 * k; t.hp.mask = kmask(k);
    t.hp.n_shards = n_shards; t.hp.shard = shard;
    t.finalized = false; t.n_inserted = 0;
    CK(cudaMalloc(&t.d_buckets, t.n_phys * sizeof(Bucket)));
    size_t fw = (t.n_phys + 15) / 16;
    CK(cudaMalloc(&t.d_flags, fw * sizeof(uint32_t)));
    CK(cudaMalloc(&t.d_status, 2 * sizeof(uint32_t)));
    CK(cudaMalloc(&t.d_counts, 3 * sizeof(unsigned long long)));
    CK(cudaMemsetAsync(t.d_flags, 0, fw * sizeof(uint32_t), s));
    CK(cudaMemsetAsync(t.d_status, 0, 2 * sizeof(uint32_t), s));
    CK(cudaMemsetAsync(t.d_counts, 0, 3 * sizeof(unsigned long long), s));
    k_fill_empty<<<grid_for(t.n_phys, 256), 256, 0, s>>>((ulonglong4*)t.d_buckets, t.n_phys);
    CK(cudaGetLastError());
    return 0;
}

int table_insert_device(Table& t, const uint64_t* d_canon, const uint16_t* d_labels, uint64_t n, cudaStream_t s)
{
    if (t.finalized || !t.d_buckets) return -11;
    if (n == 0) return 0;
    k_insert_pairs<<<grid_for(n, 256), 256, 0, s>>>((uint64_t*)t.d_buckets - t.lo * kSlots, t.d_flags - t.lo / 16, t.d_status, t.d_counts,
                                                    t.hp, d_canon, d_labels, n);
    CK(cudaGetLastError());
    return 0;
}

size_t table_ref_chunk_scratch(uint32_t max_buckets)
{
    size_t a = 0, b = 0;
    auto in1 = thrust::make_transform_iterator((const uint8_t*)nullptr, Widen());
    auto in2 = thrust::make_transform_iterator((const uint8_t*)nullptr, NonZero());
    cub::DeviceScan::ExclusiveSum(nullptr, a, in1, (uint32_t*)nullptr, max_buckets);
    cub::DeviceScan::InclusiveSum(nullptr, b, in2, (uint32_t*)nullptr, max_buckets);
    return (a > b ? a : b) + 256;
}

int table_insert_ref_chunk(Table& t, uint64_t htsize, uint64_t r0, uint32_t nbk, const uint8_t* d_sz,
                           uint32_t* d_off, uint32_t* d_rank, void* d_scratch, size_t scratch_bytes,
                           uint64_t rank_base, uint32_t sampling,
                           const void* d_keys, int key_bytes, const uint16_t* d_labels, cudaStream_t s)
{
    if (t.finalized || !t.d_buckets) return -11;
    if (nbk == 0) return 0;
    auto in1 = thrust::make_transform_iterator(d_sz, Widen());
    auto in2 = thrust::make_transform_iterator(d_sz, NonZero());
    CK(cub::DeviceScan::ExclusiveSum(d_scratch, scratch_bytes, in1, d_off, nbk, s));
    if (sampling > 1) CK(cub::DeviceScan::InclusiveSum(d_scratch, scratch_bytes, in2, d_rank, nbk, s));
    int g = grid_for(nbk, 256);
    uint64_t* b = (uint64_t*)t.d_buckets - t.lo * kSlots;   // indexed by global bucket
    uint32_t* fl = t.d_flags - t.lo / 16;
    if (key_bytes == 2)
        k_insert_ref_chunk<uint16_t><<<g, 256, 0, s>>>(b, fl, t.d_status, t.d_counts, t.hp, htsize, r0, nbk, d_sz, d_off, d_rank, rank_base, sampling, (const uint16_t*)d_keys, d_labels);
    else if (key_bytes == 4)
        k_insert_ref_chunk<uint32_t><<<g, 256, 0, s>>>(b, fl, t.d_status, t.d_counts, t.hp, htsize, r0, nbk, d_sz, d_off, d_rank, rank_base, sampling, (const uint32_t*)d_keys, d_labels);
    else if (key_bytes == 8)
        k_insert_ref_chunk<uint64_t><<<g, 256, 0, s>>>(b, fl, t.d_status, t.d_counts, t.hp, htsize, r0, nbk, d_sz, d_off, d_rank, rank_base, sampling, (const uint64_t*)d_keys, d_labels);
    else return -12;
    CK(cudaGetLastError());
    return 0;
}

int table_finalize(Table& t, cudaStream_t s)
{
    if (t.finalized || !t.d_buckets) return -11;
    k_finalize<<<grid_for(t.n_phys, 256), 256, 0, s>>>((ulonglong4*)t.d_buckets, t.d_flags, t.n_phys);
    CK(cudaGetLastError());
    uint32_t st[2]; unsigned long long cnt[3];
    CK(cudaMemcpyAsync(st, t.d_status, sizeof st, cudaMemcpyDeviceToHost, s));
    CK(cudaMemcpyAsync(cnt, t.d_counts, sizeof cnt, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    CK(cudaFree(t.d_flags)); t.d_flags = nullptr;
    t.n_inserted = cnt[0];
    t.finalized = true;
    if (st[0]) return -13;                                  // table too full: displacement > kMaxDisp
    return 0;
}

void table_destroy(Table& t)
{
    if (t.d_buckets) cudaFree(t.d_buckets);
    if (t.d_flags) cudaFree(t.d_flags);
    if (t.d_status) cudaFree(t.d_status);
    if (t.d_counts) cudaFree(t.d_counts);
    t = Table{};
}

}  // namespace ccb
