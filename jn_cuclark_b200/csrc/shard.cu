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
#include <cub/device/device_scan.cuh>
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
    size_t need = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, need, (uint32_t*)d_pos, (uint32_t*)d_pos, n, s);
    ccb_handle::ShardScratch* x = shard_scratch_for(h, stream);
    if (!x) return ccb_fail(h, -64, "ccb_shard_*: more than 4 streams in use on one handle");
    if (need > x->scan_bytes) {
        cudaFree(x->scan_tmp); x->scan_tmp = nullptr; x->scan_bytes = 0;
        CU(h, cudaMalloc(&x->scan_tmp, need)); x->scan_bytes = need;
    }
    CU(h, cub::DeviceScan::ExclusiveSum(x->scan_tmp, need, (uint32_t*)d_pos, (uint32_t*)d_pos, n, s));
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
extern "C" int ccb_shard_route_peer(ccb_handle* h, const void* d_offsets, const void* d_containers, uint32_t n_reads,
                                    int n_shards, const void* d_pos, void* const* dst, uint64_t cap,
                                    void* const* count_dst, void* d_error, void* stream)
{
    if (!h || !dst || !count_dst) return -1;
    USE_DEV(h);
    TableView tv; int rc = hp_for_routing(h, n_shards, tv); if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    QueryArgs a{}; a.offsets = (const uint32_t*)d_offsets; a.containers = (const uint16_t*)d_containers; a.n_reads = n_reads;
    RouteArgs ra{}; ra.n_shards = (uint32_t)n_shards; ra.pos = (const uint32_t*)d_pos; ra.derive_base = 1; ra.cap = cap;
    for (int g = 0; g < n_shards; g++) ra.dst[g] = (uint64_t*)dst[g];
    rc = launch_route(tv, a, ra, true, s);
    if (!rc) rc = launch_post_counts((const uint32_t*)d_pos, n_reads, (uint32_t)n_shards, cap,
                                     (unsigned long long* const*)count_dst, (uint32_t*)d_error, s);
    if (rc) return ccb_fail(h, rc, "route launch failed");
    return 0;
}

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
