// shard.cu -- C ABI of the sharded-table mode.
//
// The reference splits a database that does not fit one device by bucket range and
// time-slices the parts through the same GPU (src/CuClarkDB.cu:552-559, 776-815).
// Here the hash space is cut into n_shards equal ranges, one per GPU, and a batch is
// classified in four device phases around one exchange step each way:
//
//   requester  ccb_shard_count   per read, per owner shard: how many k-mers      (k_query<kModeCount>)
//                                + exclusive scan -> pos[owner][read]
//              ccb_shard_route   mixed k-mers h written to dst[owner] + pos...  (k_query<kModeRoute>)
//                                dst[g] is a local send segment (NCCL all-to-all follows) or a
//                                peer's inbox mapped into this process (stores travel over NVLink)
//   owner      ccb_shard_probe   labels for a list of mixed k-mers               (k_probe_list)
//   requester  ccb_shard_vote    votes from the label lists, same indices        (k_vote_lists)
//
// The caller (jn_cuclark_b200/sharded.py, one process per GPU) owns the buffers and the exchange.
#include "capi_internal.cuh"
#include <string.h>

using namespace ccb;

#define CU(h, x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) \
    return ccb_fail(h, (int)e_, "%s failed: %s (%s:%d)", #x, cudaGetErrorString(e_), __FILE__, __LINE__); } while (0)
#define USE_DEV(h) CU(h, cudaSetDevice((h)->cfg.device))

extern "C" int ccb_shard_config(ccb_handle* h, int n_shards, int shard)
{
    if (!h) return -1;
    if (n_shards < 1 || n_shards > kMaxShards || shard < 0 || shard >= n_shards)
        return ccb_fail(h, -60, "ccb_shard_config: shard %d of %d (at most %d shards)", shard, n_shards, kMaxShards);
    if (h->db_ready || h->db_building) return ccb_fail(h, -61, "ccb_shard_config: call before the database is loaded");
    h->n_shards = (uint32_t)n_shards; h->shard = (uint32_t)shard;
    return 0;
}

static int hp_for_routing(ccb_handle* h, int n_shards, TableView& tv)
{
    if (n_shards < 1 || n_shards > kMaxShards) return ccb_fail(h, -60, "at most %d shards", kMaxShards);
    tv.buckets = nullptr;
    tv.hp = HashParams{};
    tv.hp.k = h->cfg.k; tv.hp.mask = kmask(h->cfg.k); tv.hp.shift = 64 - 2 * h->cfg.k;
    return 0;
}

// ---------------------------------------------------------------------------
// exclusive prefix sum of u32, in place, three small kernels (tile sums -> scan of the sums -> apply); replaces
// cub::DeviceScan so that no library kernel sits on the classification path.  tmp: scan_tmp_words(n) u32.
// ---------------------------------------------------------------------------
namespace {
constexpr int kScanThreads = 256, kScanPer = 16, kScanTile = kScanThreads * kScanPer;

__device__ __forceinline__ uint32_t block_exclusive(uint32_t v, uint32_t* warp_sums, uint32_t& total)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, o); if (lane >= o) x += y; }
    if (lane == 31) warp_sums[w] = x;
    __syncthreads();
    if (w == 0) {
        uint32_t s = lane < (int)(blockDim.x >> 5) ? warp_sums[lane] : 0u;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, s, o); if (lane >= o) s += y; }
        warp_sums[lane] = s;                      // inclusive over the warps
    }
    __syncthreads();
    total = warp_sums[(blockDim.x >> 5) - 1];
    const uint32_t before = w ? warp_sums[w - 1] : 0u;
    __syncthreads();
    return before + x - v;
}

__global__ void __launch_bounds__(kScanThreads) k_scan_tile_sums(const uint32_t* __restrict__ d, size_t n, uint32_t* __restrict__ sums)
{
    __shared__ uint32_t ws[32];
    const size_t base = (size_t)blockIdx.x * kScanTile;
    uint32_t v = 0;
#pragma unroll
    for (int j = 0; j < kScanPer; j++) { const size_t i = base + (size_t)j * kScanThreads + threadIdx.x; if (i < n) v += d[i]; }
    uint32_t total;
    block_exclusive(v, ws, total);
    if (threadIdx.x == 0) sums[blockIdx.x] = total;
}

__global__ void __launch_bounds__(1024) k_scan_sums(uint32_t* sums, uint32_t nb)
{
    __shared__ uint32_t ws[32];
    uint32_t carry = 0;
    for (uint32_t b0 = 0; b0 < nb; b0 += 1024) {
        const uint32_t i = b0 + threadIdx.x;
        const uint32_t v = i < nb ? sums[i] : 0u;
        uint32_t total;
        const uint32_t ex = block_exclusive(v, ws, total);
        if (i < nb) sums[i] = carry + ex;
        carry += total;
    }
}

__global__ void __launch_bounds__(kScanThreads) k_scan_apply(uint32_t* __restrict__ d, size_t n, const uint32_t* __restrict__ sums)
{
    __shared__ uint32_t ws[32];
    const size_t base = (size_t)blockIdx.x * kScanTile + (size_t)threadIdx.x * kScanPer;     // each thread owns 16 consecutive elements
    uint32_t v[kScanPer], mine = 0;
#pragma unroll
    for (int j = 0; j < kScanPer; j++) { v[j] = base + j < n ? d[base + j] : 0u; mine += v[j]; }
    uint32_t total;
    uint32_t run = sums[blockIdx.x] + block_exclusive(mine, ws, total);
#pragma unroll
    for (int j = 0; j < kScanPer; j++) { if (base + j < n) d[base + j] = run; run += v[j]; }
}

}  // namespace

namespace ccb {
size_t scan_tmp_words(size_t n) { return (n + kScanTile - 1) / kScanTile + 1; }

int exclusive_scan_u32(uint32_t* d, size_t n, uint32_t* tmp, cudaStream_t s)
{
    if (n == 0) return 0;
    const uint32_t nb = (uint32_t)((n + kScanTile - 1) / kScanTile);
    k_scan_tile_sums<<<nb, kScanThreads, 0, s>>>(d, n, tmp);
    k_scan_sums<<<1, 1024, 0, s>>>(tmp, nb);
    k_scan_apply<<<nb, kScanThreads, 0, s>>>(d, n, tmp);
    return (int)cudaGetLastError();
}
}  // namespace ccb

// the scratch set of a stream (first use claims one of four)
static ccb_handle::ShardScratch* shard_scratch_for(ccb_handle* h, void* stream)
{
    std::lock_guard<std::mutex> g(h->mu);
    for (auto& x : h->shard_scratch) if (x.used && x.stream == stream) return &x;
    for (auto& x : h->shard_scratch) if (!x.used) { x.used = true; x.stream = stream; return &x; }
    return nullptr;
}

extern "C" int ccb_shard_count(ccb_handle* h, const void* d_offsets, const void* d_containers, uint32_t n_reads,
                               int n_shards, void* d_pos, void* stream)
{
    if (!h) return -1;
    USE_DEV(h);
    TableView tv; int rc = hp_for_routing(h, n_shards, tv); if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    const size_t n = (size_t)n_shards * n_reads + 1;
    if (n > 0xFFFFFFFFull) return ccb_fail(h, -62, "ccb_shard_count: batch too large");
    QueryArgs a{}; a.offsets = (const uint32_t*)d_offsets; a.containers = (const uint16_t*)d_containers; a.n_reads = n_reads;
    RouteArgs ra{}; ra.n_shards = (uint32_t)n_shards; ra.counts = (uint32_t*)d_pos;
    CU(h, cudaMemsetAsync((uint32_t*)d_pos + (n - 1), 0, 4, s));
    rc = launch_route(tv, a, ra, false, s);
    if (rc) return ccb_fail(h, rc, "count launch failed");
    const size_t need = scan_tmp_words(n) * 4;
    ccb_handle::ShardScratch* x = shard_scratch_for(h, stream);
    if (!x) return ccb_fail(h, -64, "ccb_shard_*: more than 4 streams in use on one handle");
    if (need > x->scan_bytes) {
        cudaFree(x->scan_tmp); x->scan_tmp = nullptr; x->scan_bytes = 0;
        CU(h, cudaMalloc(&x->scan_tmp, need)); x->scan_bytes = need;
    }
    rc = exclusive_scan_u32((uint32_t*)d_pos, n, (uint32_t*)x->scan_tmp, s);
    if (rc) return ccb_fail(h, rc, "scan launch failed");
    return 0;
}

extern "C" int ccb_shard_route(ccb_handle* h, const void* d_offsets, const void* d_containers, uint32_t n_reads,
                               int n_shards, const void* d_pos, const uint64_t* seg_base, void* const* dst, void* stream)
{
    if (!h || !dst) return -1;
    USE_DEV(h);
    TableView tv; int rc = hp_for_routing(h, n_shards, tv); if (rc) return rc;
    QueryArgs a{}; a.offsets = (const uint32_t*)d_offsets; a.containers = (const uint16_t*)d_containers; a.n_reads = n_reads;
    RouteArgs ra{}; ra.n_shards = (uint32_t)n_shards; ra.pos = (const uint32_t*)d_pos;
    ra.derive_base = seg_base ? 0u : 1u;
    for (int g = 0; g < n_shards; g++) { ra.seg_base[g] = seg_base ? seg_base[g] : 0; ra.dst[g] = (uint64_t*)dst[g]; }
    rc = launch_route(tv, a, ra, true, (cudaStream_t)stream);
    if (rc) return ccb_fail(h, rc, "route launch failed");
    return 0;
}

extern "C" int ccb_shard_probe(ccb_handle* h, const void* d_hashes, uint64_t n, void* d_labels, void* d_stats, void* stream)
{
    if (!h) return -1;
    if (!h->db_ready) return ccb_fail(h, -32, "Failed to find the database.");
    USE_DEV(h);
    int rc = launch_probe_list(h->table.view(), (const uint64_t*)d_hashes, n, (uint16_t*)d_labels, (unsigned long long*)d_stats, (cudaStream_t)stream);
    if (rc) return ccb_fail(h, rc, "probe launch failed");
    return 0;
}

extern "C" int ccb_shard_vote(ccb_handle* h, uint32_t n_reads, int n_shards, const void* d_pos, const uint64_t* seg_base,
                              const void* const* src, void* d_final5, void* d_rows, void* stream)
{
    if (!h || !src) return -1;
    if (n_shards < 1 || n_shards > kMaxShards) return ccb_fail(h, -60, "at most %d shards", kMaxShards);
    USE_DEV(h);
    cudaStream_t s = (cudaStream_t)stream;
    ccb_handle::ShardScratch* x = shard_scratch_for(h, stream);
    if (!x) return ccb_fail(h, -64, "ccb_shard_*: more than 4 streams in use on one handle");
    if (x->ovf_cap < n_reads || !x->ovf_list) {
        cudaFree(x->ovf_list); x->ovf_list = nullptr;
        CU(h, cudaMalloc(&x->ovf_list, (size_t)(n_reads ? n_reads : 1) * 4));
        x->ovf_cap = n_reads;
    }
    if (!x->ovf_count) CU(h, cudaMalloc(&x->ovf_count, 8));
    if (!x->dense) CU(h, cudaMalloc(&x->dense, (size_t)kDenseBlocks * h->cfg.n_targets * 4));
    QueryArgs a{}; a.n_reads = n_reads; a.out5 = (uint16_t*)d_final5; a.rows = (uint16_t*)d_rows; a.row_stride = h->row_stride;
    a.ovf_list = x->ovf_list; a.ovf_count = x->ovf_count;
    VoteArgs va{}; va.n_shards = (uint32_t)n_shards; va.pos = (const uint32_t*)d_pos;
    va.derive_base = seg_base ? 0u : 1u;
    for (int g = 0; g < n_shards; g++) { va.seg_base[g] = seg_base ? seg_base[g] : 0; va.src[g] = (const uint16_t*)src[g]; }
    CU(h, cudaMemsetAsync(x->ovf_count, 0, 4, s));
    int rc = launch_vote_lists(a, va, s);
    if (!rc) rc = launch_vote_lists_dense(a, va, x->ovf_list, x->ovf_count, n_reads, (uint32_t)h->cfg.n_targets, x->dense, kDenseBlocks, s);
    if (rc) return ccb_fail(h, rc, "vote launch failed");
    return 0;
}

// ---------------------------------------------------------------------------
// peer exchange: the route kernel stores into the owners' inboxes, the probe kernel stores the labels
// into the requesters' label boxes -- NVLink carries plain stores, no collective on the data path.
// ---------------------------------------------------------------------------
extern "C" int ccb_shard_probe_inbox(ccb_handle* h, int n_shards, const void* d_inbox, uint64_t cap, const void* d_counts_in,
                                     void* const* lab_dst, void* d_stats, void* stream)
{
    if (!h || !lab_dst) return -1;
    if (!h->db_ready) return ccb_fail(h, -32, "Failed to find the database.");
    if (n_shards < 1 || n_shards > kMaxShards) return ccb_fail(h, -60, "at most %d shards", kMaxShards);
    USE_DEV(h);
    InboxArgs ia{}; ia.n_shards = (uint32_t)n_shards; ia.cap = cap; ia.inbox = (const uint64_t*)d_inbox;
    ia.counts_in = (const unsigned long long*)d_counts_in;
    for (int g = 0; g < n_shards; g++) ia.lab_dst[g] = (uint16_t*)lab_dst[g];
    int rc = launch_probe_inbox(h->table.view(), ia, (unsigned long long*)d_stats, (cudaStream_t)stream);
    if (rc) return ccb_fail(h, rc, "probe launch failed");
    return 0;
}

extern "C" int ccb_device_alloc(ccb_handle* h, uint64_t bytes, void** ptr)
{
    if (!h || !ptr) return -1;
    USE_DEV(h);
    CU(h, cudaMalloc(ptr, bytes ? bytes : 1));
    return 0;
}

// CUDA IPC: 64-byte handle of a ccb_device_alloc'ed buffer, opened by the process that drives a peer GPU
extern "C" int ccb_ipc_export(ccb_handle* h, void* ptr, unsigned char out[64])
{
    if (!h || !ptr || !out) return -1;
    USE_DEV(h);
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    cudaIpcMemHandle_t mh;
    CU(h, cudaIpcGetMemHandle(&mh, ptr));
    memcpy(out, &mh, 64);
    return 0;
}

extern "C" int ccb_ipc_open(ccb_handle* h, const unsigned char in[64], void** ptr)
{
    if (!h || !in || !ptr) return -1;
    USE_DEV(h);
    cudaIpcMemHandle_t mh;
    memcpy(&mh, in, 64);
    CU(h, cudaIpcOpenMemHandle(ptr, mh, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}

extern "C" int ccb_ipc_close(ccb_handle* h, void* ptr)
{
    if (!h || !ptr) return -1;
    USE_DEV(h);
    CU(h, cudaIpcCloseMemHandle(ptr));
    return 0;
}

// ---------------------------------------------------------------------------
// One-pass pipeline (shard_block.cu): ccb_shardws_create / ccb_shard_route2 / ccb_shard_vote2
// ---------------------------------------------------------------------------
struct ccb_shardws {
    uint32_t max_reads = 0; int n_shards = 0; uint64_t cap = 0;
    // control words, cleared per batch in one memset: cursor[8] u64 | redo_count, ovf_count (generic), ovf_count (block) u32 | error u32 (not cleared)
    unsigned long long* ctrl = nullptr;
    uint32_t* blkseg = nullptr; uint8_t* rstate = nullptr; uint8_t* rid = nullptr;     // rid: n_shards regions of cap bytes
    uint32_t* redo_list = nullptr; uint32_t* pos = nullptr; uint32_t* scan_tmp = nullptr;
    uint32_t* ovf_list = nullptr; uint32_t* ovf_list_b = nullptr; uint32_t* dense = nullptr;
    bool long_reads = false;             // the last routed batch went through the generic path only
    unsigned long long* cursor() const { return ctrl; }
    uint32_t* words() const { return reinterpret_cast<uint32_t*>(ctrl + kMaxShards); }   // [0] redo_count [1] ovf generic [2] ovf block [3] error (sticky until read)
};

extern "C" void ccb_shardws_destroy(ccb_handle* h, ccb_shardws* w)
{
    if (!w) return;
    if (h) cudaSetDevice(h->cfg.device);
    cudaFree(w->ctrl); cudaFree(w->blkseg); cudaFree(w->rstate); cudaFree(w->rid); cudaFree(w->redo_list); cudaFree(w->pos);
    cudaFree(w->scan_tmp); cudaFree(w->ovf_list); cudaFree(w->ovf_list_b); cudaFree(w->dense);
    cudaGetLastError();
    delete w;
}

extern "C" int ccb_shardws_create(ccb_handle* h, uint32_t max_reads, int n_shards, uint64_t cap, ccb_shardws** out)
{
    if (!h || !out) return -1;
    if (n_shards < 1 || n_shards > kMaxShards) return ccb_fail(h, -60, "at most %d shards", kMaxShards);
    if (cap == 0 || cap > 0xFFFFFFF0ull) return ccb_fail(h, -62, "ccb_shardws_create: region size %llu out of range", (unsigned long long)cap);
    const size_t npos = (size_t)n_shards * max_reads + 1;
    if (npos > 0xFFFFFFFFull) return ccb_fail(h, -62, "ccb_shardws_create: batch too large");
    USE_DEV(h);
    ccb_shardws* w = new ccb_shardws();
    w->max_reads = max_reads; w->n_shards = n_shards; w->cap = cap;
    const size_t nblk = ((size_t)max_reads + kRouteBlk - 1) / kRouteBlk + 1;
    const size_t mr = max_reads ? max_reads : 1;
    cudaError_t e = cudaSuccess;
    auto ok = [&](cudaError_t x) { if (e == cudaSuccess) e = x; };
    ok(cudaMalloc(&w->ctrl, kMaxShards * 8 + 16));
    ok(cudaMalloc(&w->blkseg, nblk * 2 * kMaxShards * 2 * 4));
    ok(cudaMalloc(&w->rstate, mr));
    ok(cudaMalloc(&w->rid, (size_t)n_shards * cap));
    ok(cudaMalloc(&w->redo_list, mr * 4));
    ok(cudaMalloc(&w->pos, npos * 4));
    ok(cudaMalloc(&w->scan_tmp, scan_tmp_words(npos) * 4));
    ok(cudaMalloc(&w->ovf_list, mr * 4)); ok(cudaMalloc(&w->ovf_list_b, mr * 4));
    ok(cudaMalloc(&w->dense, (size_t)kDenseBlocks * h->cfg.n_targets * 4));
    if (e == cudaSuccess) e = cudaMemset(w->ctrl, 0, kMaxShards * 8 + 16);
    if (e != cudaSuccess) {
        ccb_shardws_destroy(h, w);
        return ccb_fail(h, (int)e, "ccb_shardws_create: %s", cudaGetErrorString(e));
    }
    *out = w;
    return 0;
}

extern "C" int ccb_shard_route2(ccb_handle* h, ccb_shardws* w, const void* d_offsets, const void* d_containers, uint32_t n_reads,
                                int64_t n_containers, void* const* dst, void* const* count_dst, void* stream)
{
    if (!h || !w || !dst || !count_dst) return -1;
    if (n_reads > w->max_reads) return ccb_fail(h, -33, "ccb_shard_route2: %u reads, workspace made for %u", n_reads, w->max_reads);
    USE_DEV(h);
    const int G = w->n_shards;
    TableView tv; int rc = hp_for_routing(h, G, tv); if (rc) return rc;
    tv.hp.n_shards = (uint32_t)G;
    cudaStream_t s = (cudaStream_t)stream;
    QueryArgs a{}; a.offsets = (const uint32_t*)d_offsets; a.containers = (const uint16_t*)d_containers; a.n_reads = n_reads;
    CU(h, cudaMemsetAsync(w->ctrl, 0, kMaxShards * 8 + 12, s));
    // a batch that averages more than 50 containers per read holds long reads only: generic path for all of them
    w->long_reads = n_containers >= 0 && n_containers > 50ll * (int64_t)n_reads;
    if (!w->long_reads) {
        BlockRouteArgs ra{}; ra.n_shards = (uint32_t)G; ra.cap = w->cap; ra.cursor = w->cursor(); ra.blkseg = w->blkseg;
        ra.rstate = w->rstate; ra.redo_list = w->redo_list; ra.redo_count = w->words() + 0; ra.error = w->words() + 3;
        for (int g = 0; g < G; g++) { ra.dst[g] = (uint64_t*)dst[g]; ra.rid[g] = w->rid + (size_t)g * w->cap; }
        rc = launch_route_block(tv.hp, a, ra, s);
        if (rc) return ccb_fail(h, rc, "route launch failed");
        a.read_list = w->redo_list; a.n_list = w->words() + 0;
    }
    // generic path for the listed reads (or all): count -> scan -> route behind the block path's k-mers
    const size_t npos = (size_t)G * n_reads + 1;
    CU(h, cudaMemsetAsync(w->pos, 0, npos * 4, s));
    RouteArgs ga{}; ga.n_shards = (uint32_t)G; ga.counts = w->pos;
    const uint32_t small = w->long_reads ? 0u : 148u * 16u;          // the list is short when the block kernel took the batch
    rc = launch_route(tv, a, ga, false, s, small);
    if (!rc) rc = exclusive_scan_u32(w->pos, npos, w->scan_tmp, s);
    if (rc) return ccb_fail(h, rc, "count launch failed");
    ga.counts = nullptr; ga.pos = w->pos; ga.derive_base = 1; ga.cap = w->cap; ga.dst_off = w->cursor();
    for (int g = 0; g < G; g++) ga.dst[g] = (uint64_t*)dst[g];
    rc = launch_route(tv, a, ga, true, s, small);
    if (!rc) rc = launch_post_counts(w->pos, n_reads, (uint32_t)G, w->cap, (unsigned long long* const*)count_dst, w->words() + 3, s, w->cursor());
    if (rc) return ccb_fail(h, rc, "route launch failed");
    return 0;
}

extern "C" int ccb_shard_vote2(ccb_handle* h, ccb_shardws* w, uint32_t n_reads, const void* const* src, void* d_final5, void* d_rows,
                               void* stream)
{
    if (!h || !w || !src) return -1;
    if (n_reads > w->max_reads) return ccb_fail(h, -33, "ccb_shard_vote2: %u reads, workspace made for %u", n_reads, w->max_reads);
    USE_DEV(h);
    const int G = w->n_shards;
    cudaStream_t s = (cudaStream_t)stream;
    QueryArgs a{}; a.n_reads = n_reads; a.out5 = (uint16_t*)d_final5; a.rows = (uint16_t*)d_rows; a.row_stride = h->row_stride;
    int rc = 0;
    if (!w->long_reads) {
        BlockVoteArgs bv{}; bv.n_shards = (uint32_t)G; bv.cap = w->cap; bv.blkseg = w->blkseg; bv.rstate = w->rstate;
        for (int g = 0; g < G; g++) { bv.src[g] = (const uint16_t*)src[g]; bv.rid[g] = w->rid + (size_t)g * w->cap; }
        a.ovf_list = w->ovf_list_b; a.ovf_count = w->words() + 2;
        rc = launch_vote_block(a, bv, s);
        if (!rc) rc = launch_vote_block_dense(a, bv, w->ovf_list_b, w->words() + 2, n_reads, (uint32_t)h->cfg.n_targets, w->dense, kDenseBlocks, s);
        if (rc) return ccb_fail(h, rc, "vote launch failed");
        a.read_list = w->redo_list; a.n_list = w->words() + 0;
    }
    a.ovf_list = w->ovf_list; a.ovf_count = w->words() + 1;
    VoteArgs va{}; va.n_shards = (uint32_t)G; va.pos = w->pos; va.derive_base = 1; va.src_off = w->cursor(); va.cap = w->cap;
    for (int g = 0; g < G; g++) va.src[g] = (const uint16_t*)src[g];
    rc = launch_vote_lists(a, va, s);
    if (!rc) rc = launch_vote_lists_dense(a, va, w->ovf_list, w->words() + 1, n_reads, (uint32_t)h->cfg.n_targets, w->dense, kDenseBlocks, s);
    if (rc) return ccb_fail(h, rc, "vote launch failed");
    return 0;
}

// *error of the last batches routed through this workspace (host sync on the stream): 1 = a region overflowed
extern "C" int ccb_shardws_error(ccb_handle* h, ccb_shardws* w, void* stream, int* out)
{
    if (!h || !w || !out) return -1;
    USE_DEV(h);
    uint32_t v = 0;
    CU(h, cudaMemcpyAsync(&v, w->words() + 3, 4, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    CU(h, cudaMemsetAsync(w->words() + 3, 0, 4, (cudaStream_t)stream));
    CU(h, cudaStreamSynchronize((cudaStream_t)stream));
    *out = (int)v;
    return 0;
}
