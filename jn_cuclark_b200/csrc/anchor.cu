// anchor.cu -- building the anchor index (query.cuh AnchorView; DESIGN.md "Anchor and verify").
//
// Input: the targets as one code stream F (0..3 = base in x-encoding, 4 = not a base: N, record
// boundary), the same stream the database builder scans, and the label of every record.
//   text   T = F . revcomp(F), 2 bits per base + a validity bit per base
//   disc   bit p = the k-mer starting at T[p] is stored in the k-mer table (one table lookup per
//          position, once, at build time)
//   table  every 8th k-mer of F by hash (top three bits of mix2k zero) -> [fp15 | fwd | label16 | pos32];
//          first come first served, a full bucket drops the newcomer: the index only ever SAVES probes,
//          exactness never depends on it (k_query<kModeAnchor> verifies base by base).
#include "capi_internal.cuh"
#include "../../include/ccb_synth.h"
#include <algorithm>
#include <vector>

using namespace ccb;

#define CU(h, x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) \
    return ccb_fail(h, (int)e_, "%s failed: %s (%s:%d)", #x, cudaGetErrorString(e_), __FILE__, __LINE__); } while (0)
#define USE_DEV(h) CU(h, cudaSetDevice((h)->cfg.device))

namespace ccb {

void anchor_free(Anchor& a)
{
    cudaFree(a.d_seq); cudaFree(a.d_valid); cudaFree(a.d_disc); cudaFree(a.d_buckets);
    a = Anchor{};
}

namespace {

struct ASeg { uint64_t start; uint32_t label; uint32_t pad; };

// 32 bases of T per thread -> 2 sequence words + 1 validity word
__global__ void k_pack_text(const uint8_t* __restrict__ codes, uint64_t n_codes, uint64_t nF,
                            uint32_t* seq, uint32_t* valid)
{
    const uint64_t nblk = 2 * nF / 32;
    for (uint64_t b = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; b < nblk; b += gridDim.x * (uint64_t)blockDim.x) {
        uint32_t w0 = 0, w1 = 0, v = 0;
        for (int i = 0; i < 32; i++) {
            const uint64_t P = 32 * b + i;
            uint32_t c;
            if (P < nF) c = P < n_codes ? codes[P] : 4u;
            else { const uint64_t src = nF - 1 - (P - nF); c = src < n_codes ? codes[src] : 4u; if (c < 4u) c = 3u - c; }
            const uint32_t base = c < 4u ? c : 0u;
            if (c < 4u) v |= 1u << i;
            if (i < 16) w0 |= base << (30 - 2 * i); else w1 |= base << (30 - 2 * (i - 16));
        }
        seq[2 * b] = w0; seq[2 * b + 1] = w1; valid[b] = v;
    }
}

__device__ __forceinline__ uint32_t text_code(const uint32_t* seq, uint64_t p) { return (seq[p >> 4] >> (30 - 2 * (int)(p & 15))) & 3u; }
__device__ __forceinline__ bool text_valid(const uint32_t* valid, uint64_t p) { return (valid[p >> 5] >> (p & 31)) & 1u; }

// For the 32 k-mer start positions of block b: roll over the text, call f(p, forward k-mer, reverse complement).
template <typename F>
__device__ __forceinline__ void for_kmers_of_block(const uint32_t* seq, const uint32_t* valid, uint64_t n_text, uint64_t b,
                                                   int k, uint64_t mask, F&& f)
{
    const uint64_t p_first = 32 * b;
    uint64_t x = 0, rc = 0; int run = 0;
    const int s2 = 2 * (k - 1);
    for (uint64_t e = p_first; e < p_first + 32 + (uint64_t)k - 1 && e < n_text; e++) {
        if (!text_valid(valid, e)) { run = 0; x = 0; rc = 0; continue; }
        const uint64_t c = text_code(seq, e);
        x = ((x << 2) | c) & mask;
        rc = (rc >> 2) | ((3ull - c) << s2);
        if (run < k) run++;
        if (run == k) {
            const uint64_t p = e - (uint64_t)k + 1;
            if (p >= p_first) f(p, x, rc);
        }
    }
}

__global__ void k_disc_bits(const TableView t, const uint32_t* __restrict__ seq, const uint32_t* __restrict__ valid,
                            uint64_t n_text, uint32_t* disc)
{
    const uint64_t nblk = n_text / 32;
    for (uint64_t b = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; b < nblk; b += gridDim.x * (uint64_t)blockDim.x) {
        uint32_t bits = 0;
        for_kmers_of_block(seq, valid, n_text, b, t.hp.k, t.hp.mask, [&](uint64_t p, uint64_t x, uint64_t rc) {
            if (lookup(t, x < rc ? x : rc) != kLabelNone) bits |= 1u << (p - 32 * b);
        });
        disc[b] = bits;
    }
}

__global__ void k_anchor_insert(const HashParams hp, const uint32_t* __restrict__ seq, const uint32_t* __restrict__ valid,
                                uint64_t nF, const ASeg* __restrict__ segs, int nseg,
                                unsigned long long* buckets, uint64_t nb, unsigned long long* counts)
{
    const uint64_t nblk = nF / 32;
    const int k = hp.k;
    for (uint64_t b = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; b < nblk; b += gridDim.x * (uint64_t)blockDim.x) {
        int lo = 0, hi = nseg - 1;
        while (lo < hi) { int mid = (lo + hi + 1) >> 1; if (segs[mid].start <= 32 * b) lo = mid; else hi = mid - 1; }
        for_kmers_of_block(seq, valid, nF, b, k, hp.mask, [&](uint64_t p, uint64_t x, uint64_t rc) {
            const uint64_t h = mix2k(x < rc ? x : rc, k, hp.mask);
            if ((h >> (2 * k - 3)) != 0) return;
            while (lo + 1 < nseg && segs[lo + 1].start <= p) lo++;
            const uint64_t ent = ((uint64_t)((uint32_t)(h >> 3) & 0x7FFFu) << 49) | ((uint64_t)(x <= rc ? 1u : 0u) << 48) |
                                 ((uint64_t)(segs[lo].label & 0xFFFFu) << 32) | p;
            const uint64_t bi = mulhi64(((h << 3) & hp.mask) << hp.shift, nb);
            unsigned long long* s = buckets + bi * 4;
            for (int i = 0; i < 4; i++) {
                if (s[i] == ~0ull && atomicCAS(s + i, ~0ull, (unsigned long long)ent) == ~0ull) { atomicAdd(counts, 1ull); return; }
            }
            atomicAdd(counts + 1, 1ull);          // bucket full: dropped
        });
    }
}

__global__ void k_synth_codes(uint64_t seed, uint32_t n_genomes, uint64_t genome_len, uint8_t* codes)
{
    const uint64_t per = genome_len + 1, total = per * n_genomes;
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < total; i += gridDim.x * (uint64_t)blockDim.x) {
        const uint32_t g = (uint32_t)(i / per); const uint64_t j = i % per;
        codes[i] = j < genome_len ? (uint8_t)ccb_genome_code(ccb_genome_seed(seed, g), j) : (uint8_t)4;
    }
}

}  // namespace

// d_codes: n_codes code bytes on the device; segs: (first code, label) of every record, ascending
int anchor_build(ccb_handle* h, const uint8_t* d_codes, uint64_t n_codes, const std::vector<std::pair<uint64_t, uint32_t>>& segs)
{
    if (!h->db_ready) return ccb_fail(h, -70, "anchor index: the k-mer table must be loaded first");
    if (h->n_shards > 1) return ccb_fail(h, -70, "anchor index: not with a sharded table");
    anchor_free(h->anchor);
    const uint64_t nF = (n_codes + 64 + 31) / 32 * 32;
    if (2 * nF >= (1ull << 32) * 2 || nF >= (1ull << 32)) return ccb_fail(h, -71, "anchor index: %llu bases do not fit 32-bit positions", (unsigned long long)n_codes);
    cudaStream_t s = h->db_stream;
    Anchor& a = h->anchor;
    a.nF = nF;
    const uint64_t seq_words = 2 * nF / 16 + 64, bit_words = 2 * nF / 32 + 64;
    CU(h, cudaMalloc(&a.d_seq, seq_words * 4)); CU(h, cudaMalloc(&a.d_valid, bit_words * 4)); CU(h, cudaMalloc(&a.d_disc, bit_words * 4));
    CU(h, cudaMemsetAsync(a.d_seq, 0, seq_words * 4, s)); CU(h, cudaMemsetAsync(a.d_valid, 0, bit_words * 4, s));
    CU(h, cudaMemsetAsync(a.d_disc, 0, bit_words * 4, s));
    k_pack_text<<<148 * 16, 256, 0, s>>>(d_codes, n_codes, nF, a.d_seq, a.d_valid);
    k_disc_bits<<<148 * 32, 256, 0, s>>>(h->table.view(), a.d_seq, a.d_valid, 2 * nF, a.d_disc);
    a.nb = std::max<uint64_t>(1024, nF / 16);                  // nF/8 sampled k-mers at most, 4 per bucket, half full
    CU(h, cudaMalloc(&a.d_buckets, a.nb * 32));
    CU(h, cudaMemsetAsync(a.d_buckets, 0xFF, a.nb * 32, s));
    std::vector<ASeg> hs(segs.size());
    for (size_t i = 0; i < segs.size(); i++) hs[i] = ASeg{segs[i].first, segs[i].second, 0};
    ASeg* d_segs = nullptr; unsigned long long* d_cnt = nullptr;
    CU(h, cudaMalloc(&d_segs, std::max<size_t>(1, hs.size()) * sizeof(ASeg))); CU(h, cudaMalloc(&d_cnt, 16));
    CU(h, cudaMemcpyAsync(d_segs, hs.data(), hs.size() * sizeof(ASeg), cudaMemcpyHostToDevice, s));
    CU(h, cudaMemsetAsync(d_cnt, 0, 16, s));
    k_anchor_insert<<<148 * 32, 256, 0, s>>>(h->table.hp, a.d_seq, a.d_valid, nF, d_segs, (int)hs.size(),
                                            (unsigned long long*)a.d_buckets, a.nb, d_cnt);
    unsigned long long cnt[2] = {0, 0};
    CU(h, cudaMemcpyAsync(cnt, d_cnt, 16, cudaMemcpyDeviceToHost, s));
    CU(h, cudaStreamSynchronize(s));
    CU(h, cudaGetLastError());
    cudaFree(d_segs); cudaFree(d_cnt);
    a.entries = cnt[0]; a.dropped = cnt[1];
    a.bytes = seq_words * 4 + 2 * bit_words * 4 + a.nb * 32;
    a.ready = true;
    if (h->cfg.verbose) fprintf(stderr, "anchor index: %llu bases x 2 strands, %llu sampled k-mers (%llu dropped), %.2f GB\n",
                                (unsigned long long)nF, cnt[0], cnt[1], a.bytes / 1e9);
    return 0;
}

}  // namespace ccb

extern "C" int ccb_anchor_synth(ccb_handle* h, uint64_t db_seed, uint32_t n_genomes, uint64_t genome_len, uint32_t n_labels)
{
    if (!h || n_genomes == 0 || n_labels == 0) return -1;
    USE_DEV(h);
    const uint64_t n = (uint64_t)n_genomes * (genome_len + 1);
    uint8_t* d_codes = nullptr;
    CU(h, cudaMalloc(&d_codes, n));
    k_synth_codes<<<148 * 16, 256, 0, h->db_stream>>>(db_seed, n_genomes, genome_len, d_codes);
    std::vector<std::pair<uint64_t, uint32_t>> segs(n_genomes);
    for (uint32_t g = 0; g < n_genomes; g++) segs[g] = {(uint64_t)g * (genome_len + 1), g % n_labels};
    int rc = anchor_build(h, d_codes, n, segs);
    cudaFree(d_codes);
    return rc;
}

extern "C" int ccb_anchor_drop(ccb_handle* h)
{
    if (!h) return -1;
    USE_DEV(h);
    anchor_free(h->anchor);
    return 0;
}

extern "C" int ccb_anchor_get_info(const ccb_handle* h, ccb_anchor_info* info)
{
    if (!h || !info) return -1;
    info->ready = h->anchor.ready ? 1 : 0; info->bases = h->anchor.nF; info->entries = h->anchor.entries;
    info->dropped = h->anchor.dropped; info->bytes = h->anchor.bytes;
    return 0;
}
