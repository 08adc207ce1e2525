// dbbuild.cu -- database construction on the GPU.
//
// The reference builds its database on one CPU thread with a chained hash table
// (src/CuCLARK_hh.hh:691-1329, src/HashTableStorage_hh.hh:422-461,230-280) and
// writes it bucket by bucket (src/hashTable_hh.hh:474-546).  Here:
//
//   host   : one pass over each FASTA text turns it into a code stream
//            (0..3 = base in x-encoding, 4 = window reset) -- newlines dropped,
//            header lines and non-ACGTU bytes become resets
//   device : k_emit_full   every window of k codes without a reset -> canonical
//                          k-mer -> (bucket r = c % HTSIZE, key q = c / HTSIZE)
//            radix sort    by (r, q)  [cub::DeviceRadixSort -- a library sort,
//                          construction is not the classification hot path]
//            k_flag_runs   a run of equal (r,q) survives iff one label and
//                          min(count,254) > min_count
//            k_tile_*      order-preserving compaction of the survivors
//            k_bucket_sizes, k_emit_keys -> .sz / .ky / .lb exactly as the
//                          reference lays them out
//   light mode scans on the host (its sampling is sequential per file and emits
//   one k-mer per ~4k bases) and joins the same device pipeline at the sort.
#include "capi_internal.cuh"
#include "dbbuild.cuh"
#include <cub/device/device_radix_sort.cuh>
#include <algorithm>
#include <chrono>
#include <memory>
#include <stdio.h>
#include <string.h>

using namespace ccb;

#define CU(h, x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) \
    return ccb_fail(h, (int)e_, "%s failed: %s (%s:%d)", #x, cudaGetErrorString(e_), __FILE__, __LINE__); } while (0)
#define USE_DEV(h) CU(h, cudaSetDevice((h)->cfg.device))

namespace ccb {

struct Seg { uint64_t start; uint32_t label; uint32_t pad; };

struct BuildState {
    uint64_t htsize = 0; int key_bytes = 4; bool light = false; unsigned gap = 4, min_t = 0;
    int k = 31;
    bool wide = false;              // q = c / HTSIZE needs more than 32 bits: two stable sorts instead of one
    int  r_bits = 0, q_bits = 0;
    uint64_t cap = 0;               // occurrences the device arrays hold
    uint64_t* d_key[2] = {nullptr, nullptr};
    void*     d_val[2] = {nullptr, nullptr};    // u16 labels, or u64 (r<<16|label) when wide
    unsigned long long* d_count = nullptr;      // occurrences emitted so far
    uint64_t  n_occ = 0, bases = 0;
    // staging of code streams
    uint8_t*  h_codes = nullptr; uint8_t* d_codes = nullptr; uint64_t stage_cap = 0, stage_n = 0;
    Seg*      d_segs = nullptr; std::vector<Seg> segs; uint32_t seg_cap = 0;
    // light mode: (canonical, label) pairs found on the host
    std::vector<uint64_t> l_c; std::vector<uint16_t> l_lab;
    int       n_carry = 0;          // codes at the head of the staging buffer that were emitted by the previous flush
};

void build_free(ccb_handle* h)
{
    BuildState* b = h->build;
    if (!b) return;
    cudaSetDevice(h->cfg.device);
    for (int i = 0; i < 2; i++) { cudaFree(b->d_key[i]); cudaFree(b->d_val[i]); }
    cudaFree(b->d_count); cudaFree(b->d_codes); cudaFree(b->d_segs);
    if (b->h_codes) cudaFreeHost(b->h_codes);
    delete b;
    h->build = nullptr;
}

static int bits_for(uint64_t max_value)
{
    int b = 0;
    while (b < 64 && (max_value >> b)) b++;
    return b ? b : 1;
}

// ---------------------------------------------------------------------------
// kernels
// ---------------------------------------------------------------------------
namespace {

constexpr int kRun = 16;            // consecutive stream positions per thread

template <bool WIDE>
__device__ __forceinline__ void store_occ(uint64_t* keys, void* vals, uint64_t i, uint64_t c, uint32_t label, uint64_t htsize)
{
    const uint64_t q = c / htsize, r = c - q * htsize;
    if (WIDE) { keys[i] = q; ((uint64_t*)vals)[i] = (r << 16) | label; }
    else      { keys[i] = (r << 32) | q; ((uint16_t*)vals)[i] = (uint16_t)label; }
}

// Every window of k codes with no reset in it -> one occurrence.  A thread owns kRun
// consecutive positions (window END positions), rebuilds the window state from the k-1 codes
// before them, and emits for positions >= first_emit (codes before that are the carry-over of
// the previous flush, already emitted).
template <bool WIDE>
__global__ void __launch_bounds__(256) k_emit_full(const uint8_t* __restrict__ codes, uint64_t n, uint64_t first_emit,
                                                   const Seg* __restrict__ segs, int nseg, int k, uint64_t htsize,
                                                   uint64_t* keys, void* vals, unsigned long long* counter, uint64_t cap)
{
    const uint64_t s = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) * kRun;
    const uint64_t mask = kmask(k);
    uint64_t out[kRun]; uint32_t out_pos = 0;      // bit j of out_pos: position s+j emitted
    int m = 0;
    if (s < n) {
        uint64_t x = 0; int cpt = 0;
        const uint64_t lb = s >= (uint64_t)(k - 1) ? s - (k - 1) : 0;
        for (uint64_t j = lb; j < s; j++) {
            const uint32_t c = codes[j];
            if (c < 4) { x = ((x << 2) | c) & mask; cpt++; } else { x = 0; cpt = 0; }
        }
#pragma unroll
        for (int j = 0; j < kRun; j++) {
            const uint64_t i = s + j;
            if (i < n) {
                const uint32_t c = codes[i];
                if (c < 4) {
                    x = ((x << 2) | c) & mask;
                    if (cpt < k) cpt++;
                    if (cpt == k && i >= first_emit) { out[m++] = canonical(x, k); out_pos |= 1u << j; }
                } else { x = 0; cpt = 0; }
            }
        }
    }
    // warp-aggregated reservation (order is irrelevant: everything is sorted afterwards)
    const int lane = threadIdx.x & 31;
    int incl = m;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { int v = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += v; }
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    unsigned long long base = 0;
    if (lane == 31 && total) base = atomicAdd(counter, (unsigned long long)total);
    base = __shfl_sync(0xffffffffu, base, 31);
    if (m == 0) return;
    if (base + total > cap) return;                 // host checks the counter against cap afterwards
    // label of the segment (file) holding position s; windows never span a reset, files end with one
    int lo = 0, hi = nseg - 1;
    while (lo < hi) { int mid = (lo + hi + 1) >> 1; if (segs[mid].start <= s) lo = mid; else hi = mid - 1; }
    uint64_t w = base + (incl - m);
    int o = 0;
#pragma unroll
    for (int j = 0; j < kRun; j++) {
        if (out_pos & (1u << j)) {
            while (lo + 1 < nseg && segs[lo + 1].start <= s + j) lo++;
            store_occ<WIDE>(keys, vals, w++, out[o++], segs[lo].label, htsize);
        }
    }
}

template <bool WIDE>
__global__ void k_store_pairs(const uint64_t* __restrict__ c, const uint16_t* __restrict__ lab, uint64_t n, uint64_t base,
                              uint64_t htsize, uint64_t* keys, void* vals)
{
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += gridDim.x * (uint64_t)blockDim.x)
        store_occ<WIDE>(keys, vals, base + i, c[i], lab[i], htsize);
}

// sorted occurrence i as (r, q, label)
template <bool WIDE>
struct Occ {
    const uint64_t* a; const void* b;     // !WIDE: a = r<<32|q, b = u16 labels.  WIDE: a = r<<16|label, b = u64 q
    __device__ __forceinline__ bool same(uint64_t i, uint64_t j) const
    {
        if (WIDE) return (a[i] >> 16) == (a[j] >> 16) && ((const uint64_t*)b)[i] == ((const uint64_t*)b)[j];
        return a[i] == a[j];
    }
    __device__ __forceinline__ uint32_t label(uint64_t i) const { return WIDE ? (uint32_t)(a[i] & 0xFFFF) : ((const uint16_t*)b)[i]; }
    __device__ __forceinline__ uint64_t r(uint64_t i) const { return WIDE ? a[i] >> 16 : a[i] >> 32; }
    __device__ __forceinline__ uint64_t q(uint64_t i) const { return WIDE ? ((const uint64_t*)b)[i] : (a[i] & 0xFFFFFFFFull); }
};

constexpr int kTile = 2048;           // elements per compaction tile (256 threads x 8)

// flag[i] = 1 iff i starts a run of equal k-mers that is kept (addElement/RemoveCommon rule);
// tile_kept[t] = kept runs in tile t, counters[0] += distinct k-mers.
template <bool WIDE>
__global__ void __launch_bounds__(256) k_flag_runs(Occ<WIDE> occ, uint64_t n, unsigned min_t, uint8_t* flag,
                                                   uint32_t* tile_kept, unsigned long long* counters)
{
    __shared__ uint32_t s_kept, s_heads;
    if (threadIdx.x == 0) { s_kept = 0; s_heads = 0; }
    __syncthreads();
    const uint64_t t0 = blockIdx.x * (uint64_t)kTile;
    uint32_t kept = 0, heads = 0;
    for (int u = 0; u < kTile / 256; u++) {
        const uint64_t i = t0 + u * 256 + threadIdx.x;
        if (i >= n) break;
        uint8_t f = 0;
        if (i == 0 || !occ.same(i, i - 1)) {
            heads++;
            const uint32_t lab = occ.label(i);
            bool one = true; unsigned count = 1;
            for (uint64_t j = i + 1; j < n && occ.same(i, j); j++) {
                if (occ.label(j) != lab) one = false;
                if (count < 254) count++;
            }
            f = (one && count > min_t) ? 1 : 0;
        }
        flag[i] = f; kept += f;
    }
    atomicAdd(&s_kept, kept); atomicAdd(&s_heads, heads);
    __syncthreads();
    if (threadIdx.x == 0) { tile_kept[blockIdx.x] = s_kept; atomicAdd(counters, (unsigned long long)s_heads); }
}

// exclusive scan of the tile counts (one block; n_tiles is n/2048)
__global__ void __launch_bounds__(1024) k_tile_scan(const uint32_t* __restrict__ tile_kept, uint64_t n_tiles, uint64_t* tile_base,
                                                    unsigned long long* total)
{
    __shared__ uint64_t s_part[1024];
    const uint64_t per = (n_tiles + 1023) / 1024;
    const uint64_t a = threadIdx.x * per, b = a + per < n_tiles ? a + per : n_tiles;
    uint64_t sum = 0;
    for (uint64_t i = a; i < b; i++) sum += tile_kept[i];
    s_part[threadIdx.x] = sum;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint64_t acc = 0;
        for (int i = 0; i < 1024; i++) { uint64_t v = s_part[i]; s_part[i] = acc; acc += v; }
        *total = acc;
    }
    __syncthreads();
    uint64_t acc = s_part[threadIdx.x];
    for (uint64_t i = a; i < b; i++) { tile_base[i] = acc; acc += tile_kept[i]; }
}

// survivors, in sorted order: canonical k-mer c = q * HTSIZE + r and label
template <bool WIDE>
__global__ void __launch_bounds__(256) k_tile_scatter(Occ<WIDE> occ, uint64_t n, const uint8_t* __restrict__ flag,
                                                      const uint64_t* __restrict__ tile_base, uint64_t htsize,
                                                      uint64_t* out_c, uint16_t* out_lab)
{
    __shared__ uint32_t s_warp[8];
    const uint64_t t0 = blockIdx.x * (uint64_t)kTile;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint64_t base = tile_base[blockIdx.x];
    for (int u = 0; u < kTile / 256; u++) {
        const uint64_t i = t0 + u * 256 + threadIdx.x;
        const bool f = i < n && flag[i];
        const uint32_t bal = __ballot_sync(0xffffffffu, f);
        if (lane == 0) s_warp[wid] = __popc(bal);
        __syncthreads();
        uint32_t before = 0, all = 0;
#pragma unroll
        for (int w = 0; w < 8; w++) { const uint32_t v = s_warp[w]; if (w < wid) before += v; all += v; }
        if (f) {
            const uint64_t o = base + before + __popc(bal & ((1u << lane) - 1));
            out_c[o] = occ.q(i) * htsize + occ.r(i);
            out_lab[o] = (uint16_t)occ.label(i);
        }
        base += all;
        __syncthreads();
    }
}

// .sz: survivors are ordered by bucket; the first survivor of a bucket counts its bucket
__global__ void k_bucket_sizes(const uint64_t* __restrict__ c, uint64_t n, uint64_t htsize, uint8_t* sz, uint32_t* max_bucket)
{
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += gridDim.x * (uint64_t)blockDim.x) {
        const uint64_t r = c[i] % htsize;
        if (i && c[i - 1] % htsize == r) continue;
        uint32_t cnt = 1;
        for (uint64_t j = i + 1; j < n && cnt < 100000 && c[j] % htsize == r; j++) cnt++;
        sz[r] = (uint8_t)(cnt < 255 ? cnt : 255);
        atomicMax(max_bucket, cnt);
    }
}

// .ky: key = c / HTSIZE in key_bytes bytes, survivor order
__global__ void k_emit_keys(const uint64_t* __restrict__ c, uint64_t n, uint64_t htsize, int key_bytes, void* out)
{
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += gridDim.x * (uint64_t)blockDim.x) {
        const uint64_t q = c[i] / htsize;
        if (key_bytes == 2) ((uint16_t*)out)[i] = (uint16_t)q;
        else if (key_bytes == 4) ((uint32_t*)out)[i] = (uint32_t)q;
        else ((uint64_t*)out)[i] = q;
    }
}

}  // namespace
}  // namespace ccb

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
namespace {

constexpr uint64_t kStageBytes = 256ull << 20;

// FASTA text -> code stream (full mode).  Rules of src/CuCLARK_hh.hh:1129-1198: '\n' is skipped
// (:1170), any other byte outside ACGTU/acgtu resets the window (:1173), '>' resets and the rest
// of that line is skipped (:1176-1188).  A reset closes the file.
struct CodeBuf {                       // uninitialised byte buffer (a std::vector would zero-fill it first)
    std::unique_ptr<uint8_t[]> p; size_t n = 0;
    const uint8_t* data() const { return p.get(); }
    size_t size() const { return n; }
};

void compact_fasta(const unsigned char* t, uint64_t len, CodeBuf& out, uint64_t& bases)
{
    out.p.reset(new uint8_t[len + 2]);
    uint8_t* o = out.p.get();
    uint64_t i = 0, resets = 0;
    bool last_reset = true;
    while (i < len) {
        if (t[i] == '>') {                                     // header: reset, skip the line
            if (!last_reset) { *o++ = 4; resets++; last_reset = true; }
            const void* e = memchr(t + i, '\n', len - i);
            i = e ? (uint64_t)((const unsigned char*)e - t) + 1 : len;
            continue;
        }
        const void* e = memchr(t + i, '\n', len - i);
        const uint64_t end = e ? (uint64_t)((const unsigned char*)e - t) : len;
        // fast path: a line of nothing but ACGTU/acgtu, translated without branches
        const uint64_t m = end - i;
        unsigned bad = 0;
        for (uint64_t j = 0; j < m; j++) {
            const unsigned ch = t[i + j], up = ch & 0xDFu, v = (ch >> 1) & 3u;
            bad |= (unsigned)!(up == 'A' || up == 'C' || up == 'G' || up == 'T' || up == 'U');
            o[j] = (uint8_t)(3u - (v ^ (v >> 1)));             // A=3 C=2 G=1 T/U=0
        }
        if (!bad) { if (m) last_reset = false; o += m; i = end + 1; continue; }
        // slow path: byte by byte up to the end of this line (a '>' inside skips the rest of it)
        for (; i < end; i++) {
            const int c = code_of(t[i]);
            if (c >= 0) { *o++ = (uint8_t)c; last_reset = false; continue; }
            if (!last_reset) { *o++ = 4; resets++; last_reset = true; }
            if (t[i] == '>') { i = end; break; }
        }
        i = end + 1;
    }
    if (!last_reset) { *o++ = 4; resets++; }
    out.n = (size_t)(o - out.p.get());
    bases += out.n - resets;
}

// light mode (src/CuCLARK_hh.hh:711-767): non-overlapping windows, the window is cleared after each
// complete k-mer (:727); of those only every gap-th per file is kept (iter % gap == 0, :725)
void scan_light(const unsigned char* t, uint64_t len, int k, unsigned gap, uint16_t label,
                std::vector<uint64_t>& c_out, std::vector<uint16_t>& l_out, uint64_t& bases)
{
    const uint64_t mask = kmask(k);
    uint64_t x = 0, iter = 0, i = 0, nb = 0; int cpt = 0;
    while (i < len) {
        const unsigned char ch = t[i];
        const int c = code_of(ch);
        if (c >= 0) {
            nb++;
            x = ((x << 2) | (uint64_t)c) & mask;
            if (cpt == k - 1) {
                if (iter % gap == 0) { c_out.push_back(canonical(x, k)); l_out.push_back(label); }
                x = 0; cpt = 0; iter++;
            } else cpt++;
            i++; continue;
        }
        if (ch == '\n') { i++; continue; }
        x = 0; cpt = 0;
        if (ch == '>') {
            const void* e = memchr(t + i, '\n', len - i);
            i = e ? (uint64_t)((const unsigned char*)e - t) + 1 : len;
        } else i++;
    }
    bases += nb;
}

template <bool WIDE>
int flush_codes_t(ccb_handle* h, BuildState* b)
{
    if (b->stage_n <= (uint64_t)b->n_carry) return 0;
    cudaStream_t s = h->db_stream;
    if (b->segs.size() > b->seg_cap) {
        cudaFree(b->d_segs); b->seg_cap = (uint32_t)b->segs.size() * 2 + 1024;
        CU(h, cudaMalloc(&b->d_segs, b->seg_cap * sizeof(Seg)));
    }
    CU(h, cudaMemcpyAsync(b->d_codes, b->h_codes, b->stage_n, cudaMemcpyHostToDevice, s));
    CU(h, cudaMemcpyAsync(b->d_segs, b->segs.data(), b->segs.size() * sizeof(Seg), cudaMemcpyHostToDevice, s));
    const uint64_t threads = (b->stage_n + kRun - 1) / kRun;
    const uint32_t grid = (uint32_t)((threads + 255) / 256);
    k_emit_full<WIDE><<<grid, 256, 0, s>>>(b->d_codes, b->stage_n, (uint64_t)b->n_carry, b->d_segs, (int)b->segs.size(), b->k,
                                          b->htsize, b->d_key[0], b->d_val[0], b->d_count, b->cap);
    CU(h, cudaGetLastError());
    CU(h, cudaStreamSynchronize(s));
    // the last k-1 codes ride along so that windows spanning the flush are seen whole
    const int keep = (int)std::min<uint64_t>(b->stage_n, (uint64_t)(b->k - 1));
    memmove(b->h_codes, b->h_codes + b->stage_n - keep, (size_t)keep);
    b->n_carry = keep; b->stage_n = (uint64_t)keep;
    const uint32_t cur = b->segs.empty() ? 0 : b->segs.back().label;
    b->segs.clear(); b->segs.push_back(Seg{0, cur, 0});
    return 0;
}

int flush_codes(ccb_handle* h, BuildState* b) { return b->wide ? flush_codes_t<true>(h, b) : flush_codes_t<false>(h, b); }

}  // namespace

namespace ccb {
size_t fasta_to_codes(const unsigned char* text, uint64_t len, std::unique_ptr<uint8_t[]>& out, uint64_t& bases)
{
    CodeBuf b;
    compact_fasta(text, len, b, bases);
    out = std::move(b.p);
    return b.n;
}
}  // namespace ccb

extern "C" int ccb_build_begin(ccb_handle* h, uint64_t htsize, int key_bytes, int light, unsigned gap,
                               unsigned min_count, uint64_t max_bases)
{
    if (!h) return -1;
    if (htsize < 2 || htsize >= (1ull << 32)) return ccb_fail(h, -50, "ccb_build_begin: HTSIZE %llu out of range", (unsigned long long)htsize);
    if (key_bytes != 2 && key_bytes != 4 && key_bytes != 8) return ccb_fail(h, -50, "ccb_build_begin: key_bytes must be 2, 4 or 8");
    if (light && gap == 0) return ccb_fail(h, -50, "ccb_build_begin: gap must be >= 1");
    USE_DEV(h);
    std::lock_guard<std::mutex> g(h->build_mu);
    build_free(h);
    BuildState* b = new BuildState();
    h->build = b;
    b->htsize = htsize; b->key_bytes = key_bytes; b->light = light != 0; b->gap = gap; b->min_t = min_count; b->k = h->cfg.k;
    const uint64_t max_c = kmask(b->k);
    b->r_bits = bits_for(htsize - 1); b->q_bits = bits_for(max_c / htsize);
    b->wide = b->q_bits > 32 || getenv("CCB_BUILD_WIDE") != nullptr;
    if (key_bytes < 8 && (max_c / htsize) >> (8 * key_bytes))
        { build_free(h); return ccb_fail(h, -50, "ccb_build_begin: %d-byte keys cannot hold k=%d with HTSIZE %llu", key_bytes, b->k, (unsigned long long)htsize); }
    b->cap = max_bases ? max_bases : 1;
    const size_t vbytes = b->wide ? 8 : 2;
    cudaError_t e = cudaMalloc(&b->d_key[0], b->cap * 8);
    if (e == cudaSuccess) e = cudaMalloc(&b->d_val[0], b->cap * vbytes);
    if (e == cudaSuccess) e = cudaMalloc(&b->d_count, 8);
    if (e == cudaSuccess && !b->light) {
        uint64_t stage = kStageBytes;                       // CCB_BUILD_STAGE: smaller flushes (tests of the carry-over)
        if (const char* e = getenv("CCB_BUILD_STAGE")) stage = std::max<uint64_t>(64, strtoull(e, nullptr, 10));
        b->stage_cap = std::min<uint64_t>(stage, b->cap + 64);
        e = cudaMalloc(&b->d_codes, b->stage_cap);
        if (e == cudaSuccess) e = cudaMallocHost(&b->h_codes, b->stage_cap);
    }
    if (e != cudaSuccess) { build_free(h); return ccb_fail(h, (int)e, "ERROR: Out of GPU memory while preparing the database build (%llu bases): %s", (unsigned long long)max_bases, cudaGetErrorString(e)); }
    CU(h, cudaMemsetAsync(b->d_count, 0, 8, h->db_stream));
    return 0;
}

extern "C" int ccb_build_add_fasta(ccb_handle* h, const unsigned char* text, uint64_t len, uint16_t label)
{
    if (!h || !h->build) return ccb_fail(h, -51, "ccb_build_add_fasta: call ccb_build_begin first");
    if (!text || len == 0 || text[0] != '>') return ccb_fail(h, -52, "ccb_build_add_fasta: target is not FASTA (first byte must be '>')");
    BuildState* b = h->build;
    if (b->light) {
        std::vector<uint64_t> c; std::vector<uint16_t> l; uint64_t nb = 0;
        scan_light(text, len, b->k, b->gap, label, c, l, nb);
        std::lock_guard<std::mutex> g(h->build_mu);
        b->l_c.insert(b->l_c.end(), c.begin(), c.end()); b->l_lab.insert(b->l_lab.end(), l.begin(), l.end());
        b->bases += nb;
        return 0;
    }
    CodeBuf codes; uint64_t nb = 0;
    compact_fasta(text, len, codes, nb);
    std::lock_guard<std::mutex> g(h->build_mu);
    USE_DEV(h);
    b->bases += nb;
    if (!b->segs.empty() && b->segs.back().start == b->stage_n) b->segs.pop_back();   // previous file left no codes
    b->segs.push_back(Seg{b->stage_n, label, 0});
    uint64_t o = 0;
    while (o < codes.size()) {
        const uint64_t room = b->stage_cap - b->stage_n;
        const uint64_t m = std::min<uint64_t>(room, codes.size() - o);
        memcpy(b->h_codes + b->stage_n, codes.data() + o, (size_t)m);
        b->stage_n += m; o += m;
        if (b->stage_n == b->stage_cap) { int rc = flush_codes(h, b); if (rc) return rc; }
    }
    return 0;
}

namespace {

template <bool WIDE>
int build_end_t(ccb_handle* h, BuildState* b, const char* prefix, int load_table, ccb_build_info* info)
{
    cudaStream_t s = h->db_stream;
    const uint64_t htsize = b->htsize;
    auto t_last = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {          // --verbose: wall time of each phase (every phase ends synchronised)
        auto now = std::chrono::steady_clock::now();
        if (h->cfg.verbose) fprintf(stderr, "  build: %-28s %8.3f s\n", what, std::chrono::duration<double>(now - t_last).count());
        t_last = now;
    };
    // ---- remaining input -------------------------------------------------------------------
    if (b->light) {
        const uint64_t n = b->l_c.size();
        if (n > b->cap) return ccb_fail(h, -53, "database build: more k-mers (%llu) than max_bases announced (%llu)", (unsigned long long)n, (unsigned long long)b->cap);
        const uint64_t chunk = 1ull << 24;
        uint64_t* dc = nullptr; uint16_t* dl = nullptr;
        if (n) { CU(h, cudaMalloc(&dc, std::min(n, chunk) * 8)); CU(h, cudaMalloc(&dl, std::min(n, chunk) * 2)); }
        for (uint64_t o = 0; o < n; o += chunk) {
            const uint64_t m = std::min(chunk, n - o);
            CU(h, cudaMemcpyAsync(dc, b->l_c.data() + o, m * 8, cudaMemcpyHostToDevice, s));
            CU(h, cudaMemcpyAsync(dl, b->l_lab.data() + o, m * 2, cudaMemcpyHostToDevice, s));
            k_store_pairs<WIDE><<<(uint32_t)std::min<uint64_t>((m + 255) / 256, 148 * 8), 256, 0, s>>>(dc, dl, m, o, htsize, b->d_key[0], b->d_val[0]);
            CU(h, cudaStreamSynchronize(s));
        }
        cudaFree(dc); cudaFree(dl);
        b->n_occ = n;
    } else {
        int rc = flush_codes(h, b);
        if (rc) return rc;
        unsigned long long cnt = 0;
        CU(h, cudaMemcpyAsync(&cnt, b->d_count, 8, cudaMemcpyDeviceToHost, s));
        CU(h, cudaStreamSynchronize(s));
        if (cnt > b->cap) return ccb_fail(h, -53, "database build: more k-mers (%llu) than max_bases announced (%llu)", cnt, (unsigned long long)b->cap);
        b->n_occ = cnt;
    }
    cudaFree(b->d_codes); b->d_codes = nullptr;
    if (b->h_codes) { cudaFreeHost(b->h_codes); b->h_codes = nullptr; }
    const uint64_t n = b->n_occ;
    lap("last flush");

    // ---- sort by (bucket, key) -------------------------------------------------------------
    const size_t vbytes = WIDE ? 8 : 2;
    cudaError_t e = cudaMalloc(&b->d_key[1], std::max<uint64_t>(n, 1) * 8);
    if (e == cudaSuccess) e = cudaMalloc(&b->d_val[1], std::max<uint64_t>(n, 1) * vbytes);
    if (e != cudaSuccess) return ccb_fail(h, (int)e, "ERROR: Out of GPU memory while sorting %llu k-mers: %s", (unsigned long long)n, cudaGetErrorString(e));
    cub::DoubleBuffer<uint64_t> dk(b->d_key[0], b->d_key[1]);
    const uint64_t* occ_a = nullptr; const void* occ_b = nullptr;
    uint64_t* out_c = nullptr; uint16_t* out_lab = nullptr;
    int cur_key = 0, cur_val = 0;
    if (n) {
        void* tmp = nullptr; size_t tmp_bytes = 0;
        if (WIDE) {
            cub::DoubleBuffer<uint64_t> dv((uint64_t*)b->d_val[0], (uint64_t*)b->d_val[1]);
            size_t t1 = 0, t2 = 0;
            cub::DeviceRadixSort::SortPairs(nullptr, t1, dk, dv, n, 0, b->q_bits, s);
            cub::DeviceRadixSort::SortPairs(nullptr, t2, dv, dk, n, 16, 16 + b->r_bits, s);
            tmp_bytes = std::max(t1, t2);
            CU(h, cudaMalloc(&tmp, tmp_bytes));
            CU(h, cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, dk, dv, n, 0, b->q_bits, s));               // by key q
            CU(h, cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, dv, dk, n, 16, 16 + b->r_bits, s));          // stable, by bucket r
            occ_a = dv.Current(); occ_b = dk.Current();
            out_c = dk.Alternate(); out_lab = (uint16_t*)dv.Alternate();
            cur_key = dk.selector; cur_val = dv.selector;
        } else {
            cub::DoubleBuffer<uint16_t> dv((uint16_t*)b->d_val[0], (uint16_t*)b->d_val[1]);
            cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, dk, dv, n, 0, 32 + b->r_bits, s);
            CU(h, cudaMalloc(&tmp, tmp_bytes));
            CU(h, cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, dk, dv, n, 0, 32 + b->r_bits, s));
            occ_a = dk.Current(); occ_b = dv.Current();
            out_c = dk.Alternate(); out_lab = dv.Alternate();
            cur_key = dk.selector; cur_val = dv.selector;
        }
        CU(h, cudaStreamSynchronize(s));
        cudaFree(tmp);
    }

    lap("sort by (bucket, key)");
    // ---- keep the discriminative ones, in order ------------------------------------------------
    uint64_t n_keep = 0, n_distinct = 0;
    if (n) {
        const uint64_t n_tiles = (n + kTile - 1) / kTile;
        uint8_t* flag = nullptr; uint32_t* tile_kept = nullptr; uint64_t* tile_base = nullptr; unsigned long long* counters = nullptr;
        CU(h, cudaMalloc(&flag, n)); CU(h, cudaMalloc(&tile_kept, n_tiles * 4)); CU(h, cudaMalloc(&tile_base, n_tiles * 8));
        CU(h, cudaMalloc(&counters, 16)); CU(h, cudaMemsetAsync(counters, 0, 16, s));
        Occ<WIDE> occ{occ_a, occ_b};
        k_flag_runs<WIDE><<<(uint32_t)n_tiles, 256, 0, s>>>(occ, n, b->min_t, flag, tile_kept, counters);
        k_tile_scan<<<1, 1024, 0, s>>>(tile_kept, n_tiles, tile_base, counters + 1);
        k_tile_scatter<WIDE><<<(uint32_t)n_tiles, 256, 0, s>>>(occ, n, flag, tile_base, htsize, out_c, out_lab);
        unsigned long long hc[2] = {0, 0};
        CU(h, cudaMemcpyAsync(hc, counters, 16, cudaMemcpyDeviceToHost, s));
        CU(h, cudaStreamSynchronize(s));
        CU(h, cudaGetLastError());
        n_distinct = hc[0]; n_keep = hc[1];
        cudaFree(flag); cudaFree(tile_kept); cudaFree(tile_base); cudaFree(counters);
    }
    lap("filter + compaction");
    // the sorted occurrences are no longer needed
    cudaFree(b->d_key[cur_key]); b->d_key[cur_key] = nullptr;
    cudaFree(b->d_val[cur_val]); b->d_val[cur_val] = nullptr;
    if (n == 0) { out_c = b->d_key[1]; out_lab = (uint16_t*)b->d_val[1]; }

    // ---- reference layout: bucket sizes ------------------------------------------------------------
    uint8_t* d_sz = nullptr; uint32_t* d_max = nullptr; uint32_t max_bucket = 0;
    CU(h, cudaMalloc(&d_sz, htsize)); CU(h, cudaMalloc(&d_max, 4));
    CU(h, cudaMemsetAsync(d_sz, 0, htsize, s)); CU(h, cudaMemsetAsync(d_max, 0, 4, s));
    if (n_keep) k_bucket_sizes<<<(uint32_t)std::min<uint64_t>((n_keep + 255) / 256, 148 * 16), 256, 0, s>>>(out_c, n_keep, htsize, d_sz, d_max);
    CU(h, cudaMemcpyAsync(&max_bucket, d_max, 4, cudaMemcpyDeviceToHost, s));
    CU(h, cudaStreamSynchronize(s));
    cudaFree(d_max);
    if (info) { info->bases = b->bases; info->occurrences = n; info->distinct = n_distinct; info->stored = n_keep; info->max_bucket = max_bucket; info->reserved = 0; }
    if (max_bucket > 255) {   // src/hashTable_hh.hh:499-507: the reference aborts on a bucket it cannot size in a byte
        cudaFree(d_sz);
        return ccb_fail(h, -54, "database build: a bucket holds %u k-mers, the .sz format allows 255", max_bucket);
    }

    lap("bucket sizes");
    // ---- files ---------------------------------------------------------------------------------
    int rc = 0;
    if (prefix) {
        const std::string p(prefix);
        FILE* fs = fopen((p + ".sz").c_str(), "wb"); FILE* fk = fopen((p + ".ky").c_str(), "wb"); FILE* fl = fopen((p + ".lb").c_str(), "wb");
        const uint64_t chunk = 1ull << 25;
        void* hbuf = nullptr; void* dbuf = nullptr;
        if (!fs || !fk || !fl) rc = ccb_fail(h, -55, "Failed to create the database files %s.{sz,ky,lb}", prefix);
        if (!rc && cudaMallocHost(&hbuf, chunk * 8) != cudaSuccess) rc = ccb_fail(h, -56, "pinned buffer for the database files: out of memory");
        if (!rc && cudaMalloc(&dbuf, chunk * 8) != cudaSuccess) rc = ccb_fail(h, -56, "device buffer for the database files: out of memory");
        for (uint64_t o = 0; !rc && o < htsize; o += chunk * 8) {
            const uint64_t m = std::min<uint64_t>(chunk * 8, htsize - o);
            cudaMemcpyAsync(hbuf, d_sz + o, m, cudaMemcpyDeviceToHost, s); cudaStreamSynchronize(s);
            if (fwrite(hbuf, 1, m, fs) != m) rc = ccb_fail(h, -57, "short write on %s.sz", prefix);
        }
        for (uint64_t o = 0; !rc && o < n_keep; o += chunk) {
            const uint64_t m = std::min<uint64_t>(chunk, n_keep - o);
            k_emit_keys<<<(uint32_t)std::min<uint64_t>((m + 255) / 256, 148 * 16), 256, 0, s>>>(out_c + o, m, htsize, b->key_bytes, dbuf);
            cudaMemcpyAsync(hbuf, dbuf, m * b->key_bytes, cudaMemcpyDeviceToHost, s); cudaStreamSynchronize(s);
            if (fwrite(hbuf, (size_t)b->key_bytes, m, fk) != m) rc = ccb_fail(h, -57, "short write on %s.ky", prefix);
            cudaMemcpyAsync(hbuf, out_lab + o, m * 2, cudaMemcpyDeviceToHost, s); cudaStreamSynchronize(s);
            if (!rc && fwrite(hbuf, 2, m, fl) != m) rc = ccb_fail(h, -57, "short write on %s.lb", prefix);
        }
        if (fs) fclose(fs); if (fk) fclose(fk); if (fl) fclose(fl);
        if (hbuf) cudaFreeHost(hbuf);
        cudaFree(dbuf);
        if (!rc && cudaGetLastError() != cudaSuccess) rc = ccb_fail(h, -58, "database build: CUDA error while writing files");
    }
    cudaFree(d_sz);
    if (rc) return rc;
    if (prefix) lap("write .sz/.ky/.lb");

    // ---- table for queries -----------------------------------------------------------------------
    if (load_table) {
        rc = ccb_db_begin(h, n_keep);
        if (!rc && n_keep) rc = ccb_db_add_pairs_device(h, out_c, out_lab, n_keep);
        if (!rc) rc = ccb_db_end(h);
        lap("k-mer table");
    }
    return rc;
}

}  // namespace

extern "C" int ccb_build_end(ccb_handle* h, const char* prefix, int load_table, ccb_build_info* info)
{
    if (!h || !h->build) return ccb_fail(h, -51, "ccb_build_end: call ccb_build_begin first");
    USE_DEV(h);
    std::lock_guard<std::mutex> g(h->build_mu);
    BuildState* b = h->build;
    int rc = b->wide ? build_end_t<true>(h, b, prefix, load_table, info) : build_end_t<false>(h, b, prefix, load_table, info);
    build_free(h);
    return rc;
}
