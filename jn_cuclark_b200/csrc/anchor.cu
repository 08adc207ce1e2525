// anchor.cu -- building the anchor index (query.cuh AnchorView; DESIGN.md "Anchor and verify").
//
// Input: the targets as one code stream F (0..3 = base in x-encoding, 4 = not a base: N, record
// boundary), the same stream the database builder scans, and the label of every record.
//   text    T = F . revcomp(F), per 32 positions: 2 words of bases, validity bits, and two bits per position
//           from ONE table lookup at build time: "the k-mer starting here is stored under this record's
//           label" / "... under another label" (light databases, databases built from other files)
//   sampled every canonical k-mer of F with anchor_sampled() -> [fp15 | fwd | label16 | pos32] in a sector of
//           four, overflowing into the neighbouring sector of the same line; first come first served, a
//           k-mer that finds both full is dropped: the index only ever SAVES probes, exactness never depends
//           on it (k_query_block verifies base by base).
#include "capi_internal.cuh"
#include "dbbuild.cuh"
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
    cudaFree(a.d_txt); cudaFree(a.d_buckets);
    a = Anchor{};
}

namespace {

struct ASeg { uint64_t start; uint32_t label; uint32_t pad; };

// 32 positions of T per thread -> 2 words of bases + 1 validity word
__global__ void k_pack_text(const uint8_t* __restrict__ codes, uint64_t n_codes, uint64_t nF, uint32_t* txt)
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
        uint32_t* g = txt + kTextWords * b;
        g[0] = w0; g[1] = w1; g[2] = v; g[3] = 0; g[4] = 0;
    }
}

__device__ __forceinline__ uint32_t text_code(const uint32_t* txt, uint64_t p)
{
    return (txt[kTextWords * (p >> 5) + ((p >> 4) & 1)] >> (30 - 2 * (int)(p & 15))) & 3u;
}
__device__ __forceinline__ bool text_valid(const uint32_t* txt, uint64_t p) { return (txt[kTextWords * (p >> 5) + 2] >> (p & 31)) & 1u; }

// For the 32 k-mer start positions of block b: roll over the text, call f(p, forward k-mer, reverse complement).
template <typename F>
__device__ __forceinline__ void for_kmers_of_block(const uint32_t* txt, uint64_t n_text, uint64_t b, int k, uint64_t mask, F&& f)
{
    const uint64_t p_first = 32 * b;
    uint64_t x = 0, rc = 0; int run = 0;
    const int s2 = 2 * (k - 1);
    for (uint64_t e = p_first; e < p_first + 32 + (uint64_t)k - 1 && e < n_text; e++) {
        if (!text_valid(txt, e)) { run = 0; x = 0; rc = 0; continue; }
        const uint64_t c = text_code(txt, e);
        x = ((x << 2) | c) & mask;
        rc = (rc >> 2) | ((3ull - c) << s2);
        if (run < k) run++;
        if (run == k) {
            const uint64_t p = e - (uint64_t)k + 1;
            if (p >= p_first) f(p, x, rc);
        }
    }
}

// label of the record that holds F position p (segs ascending by start)
__device__ __forceinline__ uint32_t seg_label(const ASeg* __restrict__ segs, int nseg, uint64_t p)
{
    int lo = 0, hi = nseg - 1;
    while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (segs[mid].start <= p) lo = mid; else hi = mid - 1; }
    return segs[lo].label & 0xFFFFu;
}

__global__ void k_text_bits(const TableView t, uint32_t* txt, uint64_t nF, const ASeg* __restrict__ segs, int nseg)
{
    const uint64_t nblk = 2 * nF / 32;
    for (uint64_t b = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; b < nblk; b += gridDim.x * (uint64_t)blockDim.x) {
        uint32_t same = 0, other = 0;
        for_kmers_of_block(txt, 2 * nF, b, t.hp.k, t.hp.mask, [&](uint64_t p, uint64_t x, uint64_t rc) {
            const uint32_t l = lookup(t, x < rc ? x : rc);
            if (l >= kLabelTomb) return;                                        // absent, or stored as not discriminative
            const uint64_t pf = p < nF ? p : nF - 1 - (p - nF);                 // an F position inside the k-mer
            if (l == seg_label(segs, nseg, pf)) same |= 1u << (p - 32 * b); else other |= 1u << (p - 32 * b);
        });
        txt[kTextWords * b + 3] = same; txt[kTextWords * b + 4] = other;
    }
}

__global__ void k_anchor_insert(const HashParams hp, const uint32_t* __restrict__ txt, uint64_t nF,
                                const ASeg* __restrict__ segs, int nseg,
                                unsigned long long* buckets, uint64_t nb, unsigned long long* counts)
{
    const uint64_t nblk = nF / 32;
    const int k = hp.k;
    for (uint64_t b = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; b < nblk; b += gridDim.x * (uint64_t)blockDim.x) {
        for_kmers_of_block(txt, nF, b, k, hp.mask, [&](uint64_t p, uint64_t x, uint64_t rc) {
            const uint64_t c = x < rc ? x : rc;
            if (!anchor_sampled(c)) return;
            const uint64_t h = mix2k(c, k, hp.mask);
            const uint64_t ent = ((uint64_t)((uint32_t)h & 0x7FFFu) << 49) | ((uint64_t)(x <= rc ? 1u : 0u) << 48) |
                                 ((uint64_t)seg_label(segs, nseg, p) << 32) | p;
            const uint64_t bi = mulhi64(h << hp.shift, nb);
#pragma unroll
            for (int alt = 0; alt < 2; alt++) {
                unsigned long long* s = buckets + (bi ^ (uint64_t)alt) * 4;
                for (int i = 0; i < 4; i++) {
                    if (s[i] == ~0ull && atomicCAS(s + i, ~0ull, (unsigned long long)ent) == ~0ull) { atomicAdd(counts, 1ull); return; }
                }
            }
            atomicAdd(counts + 1, 1ull);          // both sectors full: dropped
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
    if (segs.empty()) return ccb_fail(h, -70, "anchor index: no records");
    anchor_free(h->anchor);
    const uint64_t nF = (n_codes + 64 + 31) / 32 * 32;
    if (nF >= (1ull << 32)) return ccb_fail(h, -71, "anchor index: %llu bases do not fit 32-bit positions", (unsigned long long)n_codes);
    cudaStream_t s = h->db_stream;
    Anchor& a = h->anchor;
    a.nF = nF;
    const uint64_t txt_words = (2 * nF / 32 + 16) * kTextWords;              // 16 groups of slack: the kernel reads ahead
    CU(h, cudaMalloc(&a.d_txt, txt_words * 4));
    CU(h, cudaMemsetAsync(a.d_txt, 0, txt_words * 4, s));
    std::vector<ASeg> hs(segs.size());
    for (size_t i = 0; i < segs.size(); i++) hs[i] = ASeg{segs[i].first, segs[i].second, 0};
    ASeg* d_segs = nullptr; unsigned long long* d_cnt = nullptr;
    CU(h, cudaMalloc(&d_segs, hs.size() * sizeof(ASeg))); CU(h, cudaMalloc(&d_cnt, 16));
    CU(h, cudaMemcpyAsync(d_segs, hs.data(), hs.size() * sizeof(ASeg), cudaMemcpyHostToDevice, s));
    CU(h, cudaMemsetAsync(d_cnt, 0, 16, s));
    k_pack_text<<<148 * 16, 256, 0, s>>>(d_codes, n_codes, nF, a.d_txt);
    k_text_bits<<<148 * 32, 256, 0, s>>>(h->table.view(), a.d_txt, nF, d_segs, (int)hs.size());
    a.nb = std::max<uint64_t>(1024, nF / 16) & ~1ull;          // nF/8 sampled k-mers expected, 4 per sector, half full
    CU(h, cudaMalloc(&a.d_buckets, a.nb * 32));
    CU(h, cudaMemsetAsync(a.d_buckets, 0xFF, a.nb * 32, s));
    k_anchor_insert<<<148 * 32, 256, 0, s>>>(h->table.hp, a.d_txt, nF, d_segs, (int)hs.size(),
                                            (unsigned long long*)a.d_buckets, a.nb, d_cnt);
    unsigned long long cnt[2] = {0, 0};
    CU(h, cudaMemcpyAsync(cnt, d_cnt, 16, cudaMemcpyDeviceToHost, s));
    CU(h, cudaStreamSynchronize(s));
    CU(h, cudaGetLastError());
    cudaFree(d_segs); cudaFree(d_cnt);
    a.entries = cnt[0]; a.dropped = cnt[1];
    a.bytes = txt_words * 4 + a.nb * 32;
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

// ---- from the target FASTA files (the texts the database was built from) -------------------------------
namespace ccb {
struct AnchorBuild {
    uint8_t* d_codes = nullptr; uint64_t cap = 0, n = 0;
    std::vector<std::pair<uint64_t, uint32_t>> segs;
};
void anchor_build_free(ccb_handle* h)
{
    if (!h->anchor_build) return;
    cudaFree(h->anchor_build->d_codes);
    delete h->anchor_build; h->anchor_build = nullptr;
}
}  // namespace ccb

extern "C" int ccb_anchor_begin(ccb_handle* h, uint64_t max_bases)
{
    if (!h) return -1;
    USE_DEV(h);
    if (!h->db_ready) return ccb_fail(h, -70, "anchor index: the k-mer table must be loaded first");
    if (h->n_shards > 1) return ccb_fail(h, -70, "anchor index: not with a sharded table");
    if (max_bases + 64 >= (1ull << 32)) return ccb_fail(h, -71, "anchor index: %llu bases do not fit 32-bit positions", (unsigned long long)max_bases);
    std::lock_guard<std::mutex> g(h->build_mu);
    anchor_build_free(h);
    AnchorBuild* b = new AnchorBuild;
    b->cap = max_bases + 64;
    cudaError_t e = cudaMalloc(&b->d_codes, b->cap);
    if (e != cudaSuccess) { delete b; return ccb_fail(h, (int)e, "anchor index: out of GPU memory (%llu bases): %s", (unsigned long long)max_bases, cudaGetErrorString(e)); }
    h->anchor_build = b;
    return 0;
}

extern "C" int ccb_anchor_add_fasta(ccb_handle* h, const unsigned char* text, uint64_t len, uint16_t label)
{
    if (!h || !h->anchor_build) return ccb_fail(h, -72, "ccb_anchor_add_fasta: call ccb_anchor_begin first");
    if (!text || len == 0) return 0;
    std::unique_ptr<uint8_t[]> codes; uint64_t bases = 0;
    const size_t n = fasta_to_codes(text, len, codes, bases);      // outside the lock: several host threads scan at once
    if (n == 0) return 0;
    std::lock_guard<std::mutex> g(h->build_mu);
    USE_DEV(h);
    AnchorBuild* b = h->anchor_build;
    if (b->n + n > b->cap) return ccb_fail(h, -73, "anchor index: more bases than announced in ccb_anchor_begin");
    CU(h, cudaMemcpy(b->d_codes + b->n, codes.get(), n, cudaMemcpyHostToDevice));
    b->segs.push_back({b->n, (uint32_t)label});
    b->n += n;
    return 0;
}

extern "C" int ccb_anchor_end(ccb_handle* h, ccb_anchor_info* info)
{
    if (!h || !h->anchor_build) return ccb_fail(h, -72, "ccb_anchor_end: call ccb_anchor_begin first");
    USE_DEV(h);
    std::lock_guard<std::mutex> g(h->build_mu);
    AnchorBuild* b = h->anchor_build;
    int rc = b->n ? anchor_build(h, b->d_codes, b->n, b->segs) : 0;
    anchor_build_free(h);
    if (rc == 0 && info) ccb_anchor_get_info(h, info);
    return rc;
}

extern "C" int ccb_anchor_drop(ccb_handle* h)
{
    if (!h) return -1;
    USE_DEV(h);
    anchor_build_free(h);                        // texts collected so far, if a build was under way
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
