// capi.cu -- C ABI (include/cuclark_b200.h) over the table + query kernels.
#include "../../include/cuclark_b200.h"
#include "capi_internal.cuh"
#include "dbbuild.cuh"
#include <algorithm>
#include <atomic>
#include <memory>
#include <thread>
#include <fcntl.h>
#include <unistd.h>

#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <mutex>
#include <string>
#include <vector>

using namespace ccb;

namespace { thread_local std::string g_create_error; }

int ccb_fail(ccb_handle* h, int code, const char* fmt, ...)
{
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    if (h) { std::lock_guard<std::mutex> g(h->mu); h->err = buf; } else g_create_error = buf;
    return code ? code : -1;
}
#define CU(h, x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) \
    return ccb_fail(h, (int)e_, "%s failed: %s (%s:%d)", #x, cudaGetErrorString(e_), __FILE__, __LINE__); } while (0)
#define USE_DEV(h) CU(h, cudaSetDevice((h)->cfg.device))

extern "C" const char* ccb_version(void) { return "cuclark_b200 0.1 sm_100a"; }

extern "C" const char* ccb_last_error(const ccb_handle* h)
{
    return h ? h->err.c_str() : g_create_error.c_str();
}

extern "C" int ccb_create(const ccb_config* cfg, ccb_handle** out)
{
    if (!cfg || !out) return ccb_fail(nullptr, -1, "ccb_create: null argument");
    if (cfg->k < 2 || cfg->k > 32) return ccb_fail(nullptr, -2, "The k-mer length should be in [2,32].");
    if (cfg->n_targets < 1 || cfg->n_targets > 65534) return ccb_fail(nullptr, -3, "n_targets must be in [1,65534]");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev <= 0)
        return ccb_fail(nullptr, -4, "No CUDA devices found. Abort. (%s)", cudaGetErrorString(e));   // src/CuClarkDB.cu:128
    if (cfg->device < 0 || cfg->device >= ndev)
        return ccb_fail(nullptr, -5, "%d CUDA devices requested. Insufficient devices found. Abort.", cfg->device + 1);
    ccb_handle* h = new ccb_handle();
    h->cfg = *cfg;
    if (h->cfg.n_slots < 1) h->cfg.n_slots = 1;
    if (h->cfg.max_hits < 1) h->cfg.max_hits = 15;
    if (h->cfg.load_factor <= 0) h->cfg.load_factor = 0.5;
    h->row_stride = 2 * h->cfg.max_hits + 2;
    if (cudaSetDevice(cfg->device) != cudaSuccess || cudaStreamCreateWithFlags(&h->db_stream, cudaStreamNonBlocking) != cudaSuccess) {
        ccb_fail(nullptr, -6, "cannot initialise CUDA device %d: %s", cfg->device, cudaGetErrorString(cudaGetLastError()));
        delete h; return -6;
    }
    // cudaLimitMaxL2FetchGranularity is left alone: it was measured to change nothing on this part (every L2 miss
    // fills the whole 128-byte line whatever the limit says).  CCB_L2_FETCH=32|64|128 still sets
    // it for measurements; the limit is device-wide, so a library should not touch it unasked.
    {
        if (const char* e = getenv("CCB_L2_FETCH")) { const size_t want = (size_t)atoi(e); if (want) cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, want); }
        size_t got = 0; cudaDeviceGetLimit(&got, cudaLimitMaxL2FetchGranularity);
        h->l2_fetch = (int)got;
        cudaGetLastError();
    }
    *out = h;
    return 0;
}

extern "C" int ccb_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

extern "C" int ccb_device_memory(int device, uint64_t* free_bytes, uint64_t* total_bytes)
{
    size_t f = 0, t = 0;
    if (cudaSetDevice(device) != cudaSuccess || cudaMemGetInfo(&f, &t) != cudaSuccess) { cudaGetLastError(); return -1; }
    if (free_bytes) *free_bytes = f;
    if (total_bytes) *total_bytes = t;
    return 0;
}

extern "C" int ccb_batch_slots(const ccb_handle* h) { return h ? h->cfg.n_slots : 0; }

extern "C" int ccb_l2_fetch_granularity(const ccb_handle* h) { return h ? h->l2_fetch : -1; }

static void free_slots(ccb_handle* h)
{
    for (Slot& s : h->slots) {
        if (s.stream) cudaStreamSynchronize(s.stream);
        cudaFreeHost(s.h_offsets); cudaFreeHost(s.h_cont);
        if (!s.shared_results) { cudaFreeHost(s.h_final5); cudaFreeHost(s.h_rows); }
        cudaFreeHost(s.h_ovf_count); cudaFreeHost(s.h_stats);
        cudaFree(s.d_offsets); cudaFree(s.d_cont); cudaFree(s.d_final5); cudaFree(s.d_rows);
        cudaFree(s.d_ovf_list); cudaFree(s.d_ovf_count); cudaFree(s.d_dense_scratch); cudaFree(s.d_stats); cudaFree(s.d_anchor_pre);
        if (s.done) cudaEventDestroy(s.done);
        if (s.stream) cudaStreamDestroy(s.stream);
    }
    h->slots.clear();
    cudaFreeHost(h->ref_final); cudaFreeHost(h->ref_full); h->ref_final = h->ref_full = nullptr;
    cudaGetLastError();
}

extern "C" void ccb_destroy(ccb_handle* h)
{
    if (!h) return;
    if (h->group) group_destroy(h);
    cudaSetDevice(h->cfg.device);
    cudaDeviceSynchronize();
    free_slots(h);
    build_free(h);
    anchor_build_free(h);
    anchor_free(h->anchor);
    table_destroy(h->table);
    cudaFree(h->q_ovf_list); cudaFree(h->q_ovf_count); cudaFree(h->q_dense); cudaFree(h->q_anchor_pre);
    for (auto& x : h->shard_scratch) { cudaFree(x.scan_tmp); cudaFree(x.ovf_list); cudaFree(x.ovf_count); cudaFree(x.dense); }
    if (h->db_stream) cudaStreamDestroy(h->db_stream);
    delete h;
}

// ---------------------------------------------------------------------------
// database
// ---------------------------------------------------------------------------
extern "C" int ccb_db_begin(ccb_handle* h, uint64_t expected_entries)
{
    if (!h) return -1;
    USE_DEV(h);
    if (h->db_building || h->db_ready) { table_destroy(h->table); h->db_ready = h->db_building = false; }
    anchor_free(h->anchor);                      // its "stored in the table" bits belong to the old table
    int rc = table_create(h->table, h->cfg.k, expected_entries, h->cfg.load_factor, h->db_stream, h->n_shards, h->shard);
    if (rc) {
        table_destroy(h->table);
        if (rc == (int)cudaErrorMemoryAllocation)
            return ccb_fail(h, rc, "ERROR: Out of GPU memory for the k-mer table (%llu entries).", (unsigned long long)expected_entries);
        return ccb_fail(h, rc, "table_create failed (%d)", rc);
    }
    h->db_building = true;
    return 0;
}

extern "C" int ccb_db_add_pairs_device(ccb_handle* h, const void* d_canon, const void* d_labels, uint64_t n)
{
    if (!h || !h->db_building) return ccb_fail(h, -1, "ccb_db_add_pairs_device: call ccb_db_begin first");
    USE_DEV(h);
    int rc = table_insert_device(h->table, (const uint64_t*)d_canon, (const uint16_t*)d_labels, n, h->db_stream);
    if (rc) return ccb_fail(h, rc, "table insert failed (%d)", rc);
    return 0;
}

extern "C" int ccb_db_add_pairs(ccb_handle* h, const uint64_t* canon, const uint16_t* labels, uint64_t n)
{
    if (!h || !h->db_building) return ccb_fail(h, -1, "ccb_db_add_pairs: call ccb_db_begin first");
    USE_DEV(h);
    const uint64_t chunk = 1ull << 24;
    uint64_t* dc = nullptr; uint16_t* dl = nullptr;
    uint64_t cap = n < chunk ? n : chunk;
    if (cap == 0) return 0;
    CU(h, cudaMalloc(&dc, cap * 8)); CU(h, cudaMalloc(&dl, cap * 2));
    int rc = 0;
    for (uint64_t o = 0; o < n && !rc; o += chunk) {
        uint64_t m = n - o < chunk ? n - o : chunk;
        cudaMemcpyAsync(dc, canon + o, m * 8, cudaMemcpyHostToDevice, h->db_stream);
        cudaMemcpyAsync(dl, labels + o, m * 2, cudaMemcpyHostToDevice, h->db_stream);
        rc = table_insert_device(h->table, dc, dl, m, h->db_stream);
        cudaStreamSynchronize(h->db_stream);
    }
    cudaFree(dc); cudaFree(dl);
    if (rc) return ccb_fail(h, rc, "table insert failed (%d)", rc);
    CU(h, cudaGetLastError());
    return 0;
}

extern "C" int ccb_db_end(ccb_handle* h)
{
    if (!h || !h->db_building) return ccb_fail(h, -1, "ccb_db_end: no table under construction");
    USE_DEV(h);
    uint32_t st[2] = {0, 0}; unsigned long long cnt[3] = {0, 0, 0};
    cudaMemcpyAsync(st, h->table.d_status, sizeof st, cudaMemcpyDeviceToHost, h->db_stream);
    cudaMemcpyAsync(cnt, h->table.d_counts, sizeof cnt, cudaMemcpyDeviceToHost, h->db_stream);
    int rc = table_finalize(h->table, h->db_stream);
    h->db_building = false;
    h->max_disp = st[1]; h->tombstones = cnt[2];
    if (rc == -13) { table_destroy(h->table); return ccb_fail(h, rc, "k-mer table too full: an entry moved more than %d buckets; lower load_factor", kMaxDisp); }
    if (rc) { table_destroy(h->table); return ccb_fail(h, rc, "table finalize failed (%d): %s", rc, rc > 0 ? cudaGetErrorString((cudaError_t)rc) : ""); }
    h->db_ready = true;
    return 0;
}

extern "C" int ccb_db_get_info(const ccb_handle* h, ccb_db_info* info)
{
    if (h && h->group && info) return group_db_get_info(h, info);
    if (!h || !info || !h->db_ready) return -1;
    info->entries = h->table.n_inserted; info->buckets = h->table.nb; info->bytes = h->table.bytes();
    info->tombstones = h->tombstones; info->max_displacement = h->max_disp;
    return 0;
}

// Reference files (A9): .sz = HTSIZE bucket sizes, .ky/.lb = keys/labels in
// bucket order.  Streamed in chunks; the device rebuilds canonical = key*HTSIZE
// + bucket and inserts (table.cu:k_insert_ref_chunk).
extern "C" int ccb_db_load_files(ccb_handle* h, const char* prefix, uint64_t htsize, int key_bytes,
                                 unsigned sampling, uint64_t* n_loaded)
{
    if (!h || !prefix) return -1;
    if (h->group) return group_db_load_files(h, prefix, htsize, key_bytes, sampling, n_loaded);
    USE_DEV(h);
    if (key_bytes != 2 && key_bytes != 4 && key_bytes != 8) return ccb_fail(h, -2, "key_bytes must be 2, 4 or 8");
    std::string p = prefix;
    // the three files are read with pread() from several threads: one thread copies out of the page cache at
    // 2-3 GB/s, the .sz file alone is HTSIZE = 1.6 GB
    const unsigned nthr = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
    auto pread_all = [&](int fd, void* dst, uint64_t bytes, uint64_t off) -> bool {
        const uint64_t grain = 4ull << 20;
        const unsigned T = (unsigned)std::max<uint64_t>(1, std::min<uint64_t>(nthr, (bytes + grain - 1) / grain));
        std::atomic<bool> ok(true);
        auto part = [&](unsigned t) {
            uint64_t lo = bytes / T * t, hi = t + 1 == T ? bytes : bytes / T * (t + 1);
            while (lo < hi) {
                const ssize_t r = pread(fd, (char*)dst + lo, (size_t)std::min<uint64_t>(hi - lo, 64ull << 20), (off_t)(off + lo));
                if (r <= 0) { ok = false; return; }
                lo += (uint64_t)r;
            }
        };
        if (T == 1) { part(0); return ok; }
        std::vector<std::thread> th;
        for (unsigned t = 0; t < T; t++) th.emplace_back(part, t);
        for (auto& x : th) x.join();
        return ok;
    };
    const int fs = open((p + ".sz").c_str(), O_RDONLY);
    if (fs < 0) return ccb_fail(h, -20, "Failed to open %s.sz", prefix);          // src/CuClarkDB.cu:483-488
    const int fk = open((p + ".ky").c_str(), O_RDONLY);
    if (fk < 0) { close(fs); return ccb_fail(h, -20, "Failed to open %s.ky", prefix); }
    const int fl = open((p + ".lb").c_str(), O_RDONLY);
    if (fl < 0) { close(fs); close(fk); return ccb_fail(h, -20, "Failed to open %s.lb", prefix); }

    std::unique_ptr<uint8_t[]> sz(new uint8_t[htsize]);
    if (!pread_all(fs, sz.get(), htsize, 0)) { close(fs); close(fk); close(fl); return ccb_fail(h, -21, "%s.sz is shorter than HTSIZE=%llu", prefix, (unsigned long long)htsize); }
    close(fs);
    // per block of 64 Ki buckets: keys and non-empty buckets (all host threads), then the blocks are
    // grouped into chunks the staging buffers can hold
    const uint64_t BS = 1ull << 16;
    const uint64_t nblk = (htsize + BS - 1) / BS;
    std::vector<uint64_t> blk_keys(nblk), blk_nz(nblk), blk_rank(nblk + 1, 0), blk_kept(nblk, 0);
    auto par = [&](auto&& fn) {
        std::vector<std::thread> th;
        for (unsigned t = 0; t < nthr; t++) th.emplace_back([&, t] { for (uint64_t b = t; b < nblk; b += nthr) fn(b); });
        for (auto& x : th) x.join();
    };
    par([&](uint64_t b) {
        const uint64_t lo = b * BS, hi = std::min(htsize, lo + BS);
        uint64_t k = 0, z = 0;
        for (uint64_t i = lo; i < hi; i++) { k += sz[i]; z += sz[i] != 0; }
        blk_keys[b] = k; blk_nz[b] = z;
    });
    for (uint64_t b = 0; b < nblk; b++) blk_rank[b + 1] = blk_rank[b] + blk_nz[b];
    // choose buckets (src/CuClarkDB.cu:503-513): keep every sampling-th non-empty bucket
    const bool all = sampling <= 1;
    uint64_t n_keep = 0;
    if (all) { for (uint64_t b = 0; b < nblk; b++) n_keep += blk_keys[b]; }
    else {
        par([&](uint64_t b) {
            const uint64_t lo = b * BS, hi = std::min(htsize, lo + BS);
            uint64_t rank = blk_rank[b], k = 0;
            for (uint64_t i = lo; i < hi; i++) if (sz[i]) { rank++; if (rank % sampling == 0) k += sz[i]; }
            blk_kept[b] = k;
        });
        for (uint64_t b = 0; b < nblk; b++) n_keep += blk_kept[b];
    }
    int rc = ccb_db_begin(h, n_keep);
    if (rc) { close(fk); close(fl); return rc; }

    const uint32_t CB = 1u << 24;               // max buckets per chunk
    const uint64_t CK = 1ull << 24;             // max keys per chunk (a block holds at most 255 * 64 Ki < CK); pinned staging is 2 x 112 MB
    const size_t scratch_bytes = table_ref_chunk_scratch(CB);
    uint32_t* d_off = nullptr; uint32_t* d_rank = nullptr; void* d_scratch = nullptr;
    CU(h, cudaMalloc(&d_off, CB * 4ull)); CU(h, cudaMalloc(&d_rank, CB * 4ull)); CU(h, cudaMalloc(&d_scratch, scratch_bytes));
    struct Stage { uint8_t* d_sz = nullptr; void* d_keys = nullptr; uint16_t* d_lab = nullptr;
                   uint8_t* h_sz = nullptr; void* h_keys = nullptr; uint16_t* h_lab = nullptr; cudaEvent_t done = nullptr; bool busy = false; } st[2];
    for (auto& x : st) {
        CU(h, cudaMalloc(&x.d_sz, CB)); CU(h, cudaMalloc(&x.d_keys, CK * key_bytes)); CU(h, cudaMalloc(&x.d_lab, CK * 2));
        CU(h, cudaMallocHost(&x.h_sz, CB)); CU(h, cudaMallocHost(&x.h_keys, CK * key_bytes)); CU(h, cudaMallocHost(&x.h_lab, CK * 2));
        CU(h, cudaEventCreateWithFlags(&x.done, cudaEventDisableTiming));
    }
    uint64_t blk = 0, keys_read = 0; rc = 0; int cur = 0;
    while (blk < nblk && !rc) {
        uint64_t nk = 0, e = blk;
        while (e < nblk && (e - blk) * BS < CB && nk + blk_keys[e] <= CK) { nk += blk_keys[e]; e++; }
        const uint64_t b0 = blk * BS, b1 = std::min(htsize, e * BS);
        const uint32_t nb = (uint32_t)(b1 - b0);
        Stage& x = st[cur];
        if (x.busy) { cudaEventSynchronize(x.done); x.busy = false; }      // the copy engines are done with this stage
        if (nk) {
            if (!pread_all(fk, x.h_keys, nk * (uint64_t)key_bytes, keys_read * (uint64_t)key_bytes) ||
                !pread_all(fl, x.h_lab, nk * 2, keys_read * 2)) { rc = -22; break; }
            keys_read += nk;
            memcpy(x.h_sz, sz.get() + b0, nb);
            cudaMemcpyAsync(x.d_keys, x.h_keys, nk * key_bytes, cudaMemcpyHostToDevice, h->db_stream);
            cudaMemcpyAsync(x.d_lab, x.h_lab, nk * 2, cudaMemcpyHostToDevice, h->db_stream);
            cudaMemcpyAsync(x.d_sz, x.h_sz, nb, cudaMemcpyHostToDevice, h->db_stream);
            rc = table_insert_ref_chunk(h->table, htsize, b0, nb, x.d_sz, d_off, d_rank, d_scratch, scratch_bytes,
                                        blk_rank[blk], all ? 1u : sampling, x.d_keys, key_bytes, x.d_lab, h->db_stream);
            cudaEventRecord(x.done, h->db_stream); x.busy = true;
            cur ^= 1;                                                      // read the next chunk while this one is inserted
        }
        blk = e;
    }
    cudaStreamSynchronize(h->db_stream);
    close(fk); close(fl);
    cudaFree(d_off); cudaFree(d_rank); cudaFree(d_scratch);
    for (auto& x : st) { cudaFree(x.d_sz); cudaFree(x.d_keys); cudaFree(x.d_lab); cudaFreeHost(x.h_sz); cudaFreeHost(x.h_keys); cudaFreeHost(x.h_lab); if (x.done) cudaEventDestroy(x.done); }
    if (rc == -22) { table_destroy(h->table); h->db_building = false; return ccb_fail(h, rc, "%s.ky/.lb are shorter than the bucket sizes in .sz announce", prefix); }
    if (rc) { table_destroy(h->table); h->db_building = false; return ccb_fail(h, rc, "table insert failed (%d)", rc); }
    rc = ccb_db_end(h);
    if (rc) return rc;
    if (n_loaded) *n_loaded = n_keep;
    if (h->cfg.verbose) fprintf(stderr, "DB loaded on device %d: %llu k-mers, %.3f GB table\n", h->cfg.device,
                                (unsigned long long)n_keep, h->table.bytes() / 1e9);
    return 0;
}

// ---------------------------------------------------------------------------
// batches
// ---------------------------------------------------------------------------
extern "C" int ccb_batch_free(ccb_handle* h)
{
    if (!h) return -1;
    if (h->group) return group_batch_free(h);
    cudaSetDevice(h->cfg.device);
    free_slots(h);
    h->max_reads = 0; h->max_cont = 0;
    return 0;
}

static int batch_alloc(ccb_handle* h, uint32_t max_reads, uint64_t max_containers, int extended,
                       uint64_t num_reads, const uint64_t* index_batches);

extern "C" int ccb_batch_alloc(ccb_handle* h, uint32_t max_reads, uint64_t max_containers, int extended)
{
    return batch_alloc(h, max_reads, max_containers, extended, 0, nullptr);
}

extern "C" int ccb_batch_alloc_ref(ccb_handle* h, uint64_t num_reads, uint32_t max_reads, uint64_t max_containers,
                                   const uint64_t* index_batches, int n_batches, int extended,
                                   uint16_t** full_results, uint16_t** final_results)
{
    if (!h || !index_batches || !final_results) return ccb_fail(h, -1, "ccb_batch_alloc_ref: null argument");
    if (h->group) return ccb_fail(h, -37, "ccb_batch_alloc_ref: not on a device group");
    if (n_batches != h->cfg.n_slots) return ccb_fail(h, -35, "ccb_batch_alloc_ref: %d batches but the handle was created with %d slots", n_batches, h->cfg.n_slots);
    for (int b = 0; b < n_batches; b++)
        if (index_batches[b] > num_reads) return ccb_fail(h, -36, "ccb_batch_alloc_ref: batch %d starts at read %llu of %llu", b, (unsigned long long)index_batches[b], (unsigned long long)num_reads);
    const int rc = batch_alloc(h, max_reads, max_containers, extended, num_reads ? num_reads : 1, index_batches);
    if (rc) return rc;
    *final_results = h->ref_final;
    if (full_results) *full_results = h->ref_full;
    return 0;
}

// index_batches == NULL: every slot owns its result buffers (ccb_batch_alloc).  Otherwise (ccb_batch_alloc_ref) slot b's
// results land at row index_batches[b] of one contiguous array of num_reads rows, the way CuClarkDB::malloc lays
// h_resultsFinal[i] = _finalResults + rowSize * _indexBatches[i] (src/CuClarkDB.cu:412-416).
static int batch_alloc(ccb_handle* h, uint32_t max_reads, uint64_t max_containers, int extended,
                       uint64_t num_reads, const uint64_t* index_batches)
{
    if (!h) return -1;
    if (h->group) return group_batch_alloc(h, max_reads, max_containers, extended);
    USE_DEV(h);
    if (max_containers > 0xFFFFFFFFull)
        return ccb_fail(h, -30, "ERROR: Batch overflow. Please increase the number of batches (-b <numberofbatches>).");  // CuCLARK_hh.hh:1709-1713
    // slots that are large enough are kept (the next file of a list; pinning and unpinning 250 MB costs ~0.1 s each)
    if (!index_batches && !h->ref_final && !h->slots.empty() && (int)h->slots.size() == h->cfg.n_slots && h->max_reads >= max_reads &&
        h->max_cont >= max_containers && h->batch_extended == (extended != 0)) {
        bool idle = true;
        for (const Slot& s : h->slots) idle = idle && !s.in_flight;
        if (idle) return 0;
    }
    free_slots(h);
    h->max_reads = max_reads; h->max_cont = max_containers; h->batch_extended = extended != 0;
    h->slots.resize((size_t)h->cfg.n_slots);
    DeviceNodeBind near_gpu(h->cfg.device);                     // pinned buffers on the memory node the GPU hangs off
    const size_t so = ((size_t)max_reads + 1) * 4, sc = (size_t)(max_containers ? max_containers : 1) * 2,
                 sf = (size_t)(max_reads ? max_reads : 1) * 10, sr = (size_t)(max_reads ? max_reads : 1) * h->row_stride * 2;
    if (index_batches) {
        cudaError_t e = cudaMallocHost(&h->ref_final, (size_t)num_reads * 10);
        if (e == cudaSuccess && extended) e = cudaMallocHost(&h->ref_full, (size_t)num_reads * h->row_stride * 2);
        if (e != cudaSuccess) { free_slots(h); return ccb_fail(h, (int)e, "result allocation failed: %s", cudaGetErrorString(e)); }
    }
    size_t slot_no = 0;
    for (Slot& s : h->slots) {
        cudaError_t e = cudaSuccess;
        auto ok = [&](cudaError_t x) { if (e == cudaSuccess) e = x; };
        ok(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
        ok(cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming | cudaEventBlockingSync));
        ok(cudaMallocHost(&s.h_offsets, so)); ok(cudaMallocHost(&s.h_cont, sc));
        if (index_batches) {
            s.shared_results = true;
            s.h_final5 = h->ref_final + 5 * index_batches[slot_no];
            if (extended) s.h_rows = h->ref_full + (size_t)h->row_stride * index_batches[slot_no];
        } else {
            ok(cudaMallocHost(&s.h_final5, sf));
        }
        slot_no++;
        ok(cudaMallocHost(&s.h_ovf_count, 4)); ok(cudaMallocHost(&s.h_stats, 32));
        ok(cudaMalloc(&s.d_offsets, so)); ok(cudaMalloc(&s.d_cont, sc)); ok(cudaMalloc(&s.d_final5, sf));
        ok(cudaMalloc(&s.d_ovf_list, (size_t)(max_reads ? max_reads : 1) * 8)); ok(cudaMalloc(&s.d_ovf_count, 8));   // [overflow list | redo list]
        ok(cudaMalloc(&s.d_dense_scratch, (size_t)kDenseBlocks * h->cfg.n_targets * 4));
        ok(cudaMalloc(&s.d_stats, 32));
        ok(cudaMalloc(&s.d_anchor_pre, (size_t)(max_reads ? max_reads : 1) * kAnchorPreBytes));
        if (extended) { if (!index_batches) ok(cudaMallocHost(&s.h_rows, sr)); ok(cudaMalloc(&s.d_rows, sr)); }
        if (e != cudaSuccess) {
            free_slots(h);
            if (e == cudaErrorMemoryAllocation)   // src/CuClarkDB.cu:65-73
                return ccb_fail(h, (int)e, "ERROR: Out of GPU memory.\nPlease increase the number of batches (-b <numberofbatches>).");
            return ccb_fail(h, (int)e, "batch allocation failed: %s", cudaGetErrorString(e));
        }
        *s.h_ovf_count = 0; s.h_stats[0] = s.h_stats[1] = s.h_stats[2] = s.h_stats[3] = 0;
    }
    return 0;
}

extern "C" int ccb_batch_buffers(ccb_handle* h, int slot, uint32_t** offsets, uint16_t** containers,
                                 uint16_t** final5, uint16_t** rows)
{
    if (h && h->group) return group_batch_buffers(h, slot, offsets, containers, final5, rows);
    if (!h || slot < 0 || slot >= (int)h->slots.size()) return ccb_fail(h, -31, "bad slot %d", slot);
    Slot& s = h->slots[(size_t)slot];
    if (offsets) *offsets = s.h_offsets;
    if (containers) *containers = s.h_cont;
    if (final5) *final5 = s.h_final5;
    if (rows) *rows = s.h_rows;
    return 0;
}

// ovf_list holds 2 * list_cap entries and ovf_count 2 words: [0] reads for the dense pass, [1] reads the
// thread-per-read kernel handed back to the warp-per-read kernel
// n_containers: u16 units of the batch when the caller knows them (-1 otherwise): tells long-read batches apart
static int run_query(ccb_handle* h, const QueryArgs& a0, uint32_t* ovf_list, uint32_t* ovf_count, uint32_t list_cap,
                     uint32_t* dense_scratch, int flags, cudaStream_t st, int64_t n_containers = -1)
{
    QueryArgs a = a0;
    a.ovf_list = ovf_list; a.ovf_count = ovf_count;
    TableView tv = h->table.view();
    cudaMemsetAsync(ovf_count, 0, 8, st);
    int rc;
    if (flags & CCB_FORCE_DENSE) {
        // every read through the dense exact path (cross-check of the fast path)
        QueryArgs d = a; d.stats = nullptr;
        rc = launch_query_dense(tv, d, nullptr, nullptr, a.n_reads, (uint32_t)h->cfg.n_targets, dense_scratch, kDenseBlocks, st);
        return rc;
    }
    const bool use_anchor = h->anchor.ready && !(flags & CCB_NO_ANCHOR);
    const AnchorView av = h->anchor.view();
    // short single-part reads go through the thread-per-read block kernel (query_lane.cu), everything it hands
    // back (several parts, long reads, many targets) through the warp-per-read kernel
    // (a batch that averages more than 50 containers per read holds long reads only: straight to the warp kernel)
    const bool long_reads = n_containers >= 0 && n_containers > 50ll * (int64_t)a.n_reads;
    uint32_t redo_grid = 0;
    if (!(flags & CCB_WARP_KERNEL) && !long_reads && list_cap >= a.n_reads) {
        uint32_t* redo_list = ovf_list + list_cap; uint32_t* redo_count = ovf_count + 1;
        rc = launch_query_lane(tv, a, use_anchor ? &av : nullptr, redo_list, redo_count, st);
        if (rc) return rc;
        a.read_list = redo_list; a.n_list = redo_count;
        if (n_containers >= 0) redo_grid = 148u * 16u;      // short reads on average: the list is short
    }
    rc = launch_query(tv, a, st, redo_grid);
    if (rc) return rc;
    QueryArgs d = a; d.stats = nullptr; d.read_list = nullptr; d.n_list = nullptr;
    return launch_query_dense(tv, d, ovf_list, ovf_count, a.n_reads, (uint32_t)h->cfg.n_targets, dense_scratch, kDenseBlocks, st);
}

extern "C" int ccb_batch_submit(ccb_handle* h, int slot, uint32_t n_reads, uint64_t n_containers, int flags)
{
    if (h && h->group) return group_batch_submit(h, slot, n_reads, n_containers, flags);
    if (!h || slot < 0 || slot >= (int)h->slots.size()) return ccb_fail(h, -31, "bad slot %d", slot);
    if (!h->db_ready) return ccb_fail(h, -32, "Failed to find the database.");
    if (h->n_shards > 1) return ccb_fail(h, -63, "this handle holds one shard of the table: use the ccb_shard_* calls");
    if (n_reads > h->max_reads || n_containers > h->max_cont) return ccb_fail(h, -33, "batch larger than allocated (%u reads, %llu containers)", n_reads, (unsigned long long)n_containers);
    if ((flags & CCB_EXTENDED) && !h->batch_extended) return ccb_fail(h, -34, "extended results requested but ccb_batch_alloc(extended=0)");
    USE_DEV(h);
    Slot& s = h->slots[(size_t)slot];
    if (s.in_flight) { CU(h, cudaEventSynchronize(s.done)); s.in_flight = false; }
    s.n_reads = n_reads; s.flags = flags;
    if (n_reads == 0) { CU(h, cudaEventRecord(s.done, s.stream)); s.in_flight = true; return 0; }
    CU(h, cudaMemcpyAsync(s.d_offsets, s.h_offsets, ((size_t)n_reads + 1) * 4, cudaMemcpyHostToDevice, s.stream));
    if (n_containers) CU(h, cudaMemcpyAsync(s.d_cont, s.h_cont, (size_t)n_containers * 2, cudaMemcpyHostToDevice, s.stream));
    QueryArgs a{};
    a.offsets = s.d_offsets; a.containers = s.d_cont; a.n_reads = n_reads;
    a.out5 = s.d_final5; a.rows = (flags & CCB_EXTENDED) ? s.d_rows : nullptr; a.row_stride = h->row_stride;
    a.stats = (flags & CCB_STATS) ? s.d_stats : nullptr; a.stats_words = 4;
    a.anchor_pre = s.d_anchor_pre;
    if (flags & CCB_STATS) CU(h, cudaMemsetAsync(s.d_stats, 0, 32, s.stream));
    int rc = run_query(h, a, s.d_ovf_list, s.d_ovf_count, h->max_reads, s.d_dense_scratch, flags, s.stream, (int64_t)n_containers);
    if (rc) return ccb_fail(h, rc, "query launch failed: %s", cudaGetErrorString((cudaError_t)rc));
    CU(h, cudaMemcpyAsync(s.h_final5, s.d_final5, (size_t)n_reads * 10, cudaMemcpyDeviceToHost, s.stream));
    if (flags & CCB_EXTENDED) CU(h, cudaMemcpyAsync(s.h_rows, s.d_rows, (size_t)n_reads * h->row_stride * 2, cudaMemcpyDeviceToHost, s.stream));
    CU(h, cudaMemcpyAsync(s.h_ovf_count, s.d_ovf_count, 4, cudaMemcpyDeviceToHost, s.stream));
    if (flags & CCB_STATS) CU(h, cudaMemcpyAsync(s.h_stats, s.d_stats, 32, cudaMemcpyDeviceToHost, s.stream));
    CU(h, cudaEventRecord(s.done, s.stream));
    s.in_flight = true;
    return 0;
}

static void harvest(ccb_handle* h, Slot& s)
{
    std::lock_guard<std::mutex> g(h->mu);          // slots may be waited on from several host threads
    h->stats.reads += s.n_reads;
    h->stats.dense_reads += (s.flags & CCB_FORCE_DENSE) ? s.n_reads : *s.h_ovf_count;
    if (s.flags & CCB_STATS) { h->stats.lookups += s.h_stats[0]; h->stats.sectors += s.h_stats[1]; h->stats.table_probes += s.h_stats[2]; }
    s.in_flight = false;
}

extern "C" int ccb_batch_wait(ccb_handle* h, int slot)
{
    if (h && h->group) return group_batch_wait(h, slot);
    if (!h || slot < 0 || slot >= (int)h->slots.size()) return ccb_fail(h, -31, "bad slot %d", slot);
    Slot& s = h->slots[(size_t)slot];
    if (!s.in_flight) return 0;
    USE_DEV(h);
    CU(h, cudaEventSynchronize(s.done));
    harvest(h, s);
    return 0;
}

extern "C" int ccb_batch_poll(ccb_handle* h, int slot)
{
    if (h && h->group) return group_batch_poll(h, slot);
    if (!h || slot < 0 || slot >= (int)h->slots.size()) return ccb_fail(h, -31, "bad slot %d", slot);
    Slot& s = h->slots[(size_t)slot];
    if (!s.in_flight) return 0;
    cudaError_t e = cudaEventQuery(s.done);
    if (e == cudaErrorNotReady) return 1;
    if (e != cudaSuccess) return ccb_fail(h, (int)e, "cudaEventQuery: %s", cudaGetErrorString(e));
    harvest(h, s);
    return 0;
}

extern "C" int ccb_sync(ccb_handle* h)
{
    if (!h) return -1;
    if (h->group) return group_sync(h);
    USE_DEV(h);
    CU(h, cudaDeviceSynchronize());
    return 0;
}

extern "C" int ccb_query_device(ccb_handle* h, const void* d_offsets, const void* d_containers,
                                uint32_t n_reads, void* d_final5, void* d_rows, void* d_stats,
                                int flags, void* stream)
{
    if (!h) return -1;
    if (!h->db_ready) return ccb_fail(h, -32, "Failed to find the database.");
    if (h->group) return ccb_fail(h, -63, "this handle leads a device group: use the ccb_batch_* calls");
    if (h->n_shards > 1) return ccb_fail(h, -63, "this handle holds one shard of the table: use the ccb_shard_* calls");
    USE_DEV(h);
    if (h->q_ovf_cap < n_reads || !h->q_ovf_list) {
        cudaFree(h->q_ovf_list); h->q_ovf_list = nullptr;
        CU(h, cudaMalloc(&h->q_ovf_list, (size_t)(n_reads ? n_reads : 1) * 8));
        h->q_ovf_cap = n_reads;
    }
    if (!h->q_ovf_count) CU(h, cudaMalloc(&h->q_ovf_count, 8));
    if (h->anchor.ready && !(flags & CCB_NO_ANCHOR) && (h->q_pre_cap < n_reads || !h->q_anchor_pre)) {
        cudaFree(h->q_anchor_pre); h->q_anchor_pre = nullptr; h->q_pre_cap = 0;
        CU(h, cudaMalloc(&h->q_anchor_pre, (size_t)(n_reads ? n_reads : 1) * kAnchorPreBytes));
        h->q_pre_cap = n_reads;
    }
    if (!h->q_dense) CU(h, cudaMalloc(&h->q_dense, (size_t)kDenseBlocks * h->cfg.n_targets * 4));
    QueryArgs a{};
    a.offsets = (const uint32_t*)d_offsets; a.containers = (const uint16_t*)d_containers; a.n_reads = n_reads;
    a.out5 = (uint16_t*)d_final5; a.rows = (uint16_t*)d_rows; a.row_stride = h->row_stride;
    a.stats = (flags & CCB_STATS) ? (unsigned long long*)d_stats : nullptr; a.stats_words = (flags & CCB_STATS_EX) ? 4 : 2;
    a.anchor_pre = h->q_anchor_pre;
    int rc = run_query(h, a, h->q_ovf_list, h->q_ovf_count, h->q_ovf_cap, h->q_dense, flags, (cudaStream_t)stream);
    if (rc) return ccb_fail(h, rc, "query launch failed: %s", cudaGetErrorString((cudaError_t)rc));
    return 0;
}

extern "C" int ccb_get_stats(const ccb_handle* h, ccb_stats* out)
{
    if (!h || !out) return -1;
    *out = h->stats;
    return 0;
}

// ---------------------------------------------------------------------------
// ccb_query_device_noalloc: ccb_query_device for callers that must not allocate (CUDA graph capture, serving loops,
// harnesses that forbid cudaMalloc).  No counterpart in the reference.
//
// ccb_query_device keeps its scratch (overflow + redo lists, their counters, the anchor pass's 64 bytes per read, the
// dense path's per-block target counters) in buffers it cudaMallocs on first use and regrows.  Here the same scratch is
// ONE file-scope device array of a fixed size; every GPU that loads the library holds it (6.6 MB) until exit.  In host
// code `g_noalloc_scratch` is only a stand-in for the symbol: its device address is asked for once per device with
// cudaGetSymbolAddress (that device current) and kept.
// ---------------------------------------------------------------------------
namespace {

constexpr uint32_t kNoallocMaxReads = CCB_NOALLOC_MAX_READS;
constexpr uint32_t kNoallocMaxTargets = CCB_NOALLOC_MAX_TARGETS;
constexpr size_t kNoallocLists = 0;                                               // u32[2 * max reads]: overflow | redo
constexpr size_t kNoallocCount = kNoallocLists + (size_t)kNoallocMaxReads * 8;    // 2 x u32 counters (own 256 bytes)
constexpr size_t kNoallocPre = kNoallocCount + 256;                               // AnchorPre[max reads]
constexpr size_t kNoallocDense = kNoallocPre + (size_t)kNoallocMaxReads * kAnchorPreBytes;
constexpr size_t kNoallocBytes = kNoallocDense + (size_t)kDenseBlocks * kNoallocMaxTargets * 4;
constexpr int kNoallocMaxDevices = 64;

}  // namespace

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoallocBytes];

static std::mutex g_noalloc_mu;
static unsigned char* g_noalloc_ptr[kNoallocMaxDevices];                          // per device: the array's device address

extern "C" int ccb_query_device_noalloc(ccb_handle* h, const void* d_offsets, const void* d_containers,
                                        uint32_t n_reads, void* d_final5, void* d_rows, void* d_stats,
                                        int flags, void* stream)
{
    if (!h) return -1;
    if (n_reads > kNoallocMaxReads)
        return ccb_fail(h, -70, "ccb_query_device_noalloc: %u reads in one call, the fixed scratch holds %u (split the batch, or use ccb_query_device)",
                        n_reads, kNoallocMaxReads);
    if ((uint32_t)h->cfg.n_targets > kNoallocMaxTargets)
        return ccb_fail(h, -71, "ccb_query_device_noalloc: %d targets, the fixed scratch holds %u (use ccb_query_device)",
                        h->cfg.n_targets, kNoallocMaxTargets);
    if (!h->db_ready) return ccb_fail(h, -32, "Failed to find the database.");
    if (h->group) return ccb_fail(h, -63, "this handle leads a device group: use the ccb_batch_* calls");
    if (h->n_shards > 1) return ccb_fail(h, -63, "this handle holds one shard of the table: use the ccb_shard_* calls");
    const int dev = h->cfg.device;
    if (dev < 0 || dev >= kNoallocMaxDevices) return ccb_fail(h, -72, "ccb_query_device_noalloc: device %d out of range", dev);
    USE_DEV(h);
    unsigned char* base;
    {   // set-up, once per device, that device current
        std::lock_guard<std::mutex> g(g_noalloc_mu);
        if (!g_noalloc_ptr[dev]) {
            void* p = nullptr;
            CU(h, cudaGetSymbolAddress(&p, g_noalloc_scratch));
            g_noalloc_ptr[dev] = (unsigned char*)p;
        }
        base = g_noalloc_ptr[dev];
    }
    if (n_reads == 0) return 0;                                                   // (a call that only does the set-up)
    QueryArgs a{};
    a.offsets = (const uint32_t*)d_offsets; a.containers = (const uint16_t*)d_containers; a.n_reads = n_reads;
    a.out5 = (uint16_t*)d_final5; a.rows = (uint16_t*)d_rows; a.row_stride = h->row_stride;
    a.stats = (flags & CCB_STATS) ? (unsigned long long*)d_stats : nullptr; a.stats_words = (flags & CCB_STATS_EX) ? 4 : 2;
    a.anchor_pre = base + kNoallocPre;
    // the array keeps the last call's contents: the two counters are cleared by run_query, the dense counters by the
    // dense kernel itself, the lists and the anchor records are written before they are read
    int rc = run_query(h, a, (uint32_t*)(base + kNoallocLists), (uint32_t*)(base + kNoallocCount), kNoallocMaxReads,
                       (uint32_t*)(base + kNoallocDense), flags, (cudaStream_t)stream);
    if (rc) return ccb_fail(h, rc, "query launch failed: %s", cudaGetErrorString((cudaError_t)rc));
    return 0;
}
