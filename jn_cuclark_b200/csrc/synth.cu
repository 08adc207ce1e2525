// synth.cu -- synthetic benchmark workload generated on the device
// (definition: include/ccb_synth.h).  Used by bench.py
// and the GPU tests to build configs whose inputs would take hours to produce
// through files (C2: 1,000 x 4 Mbp genomes = 4e9 31-mers).
#include "../../include/cuclark_b200.h"
#include "../../include/ccb_synth.h"
#include "capi_internal.cuh"

using namespace ccb;

namespace {

constexpr int kSeg = 256;   // k-mer start positions per thread

// Every k-mer of every genome -> table (label = genome index % n_labels).
__global__ void k_synth_db(uint64_t* buckets, uint32_t* flags, uint32_t* status, unsigned long long* counts,
                           HashParams hp, uint64_t seed, uint32_t n_genomes, uint64_t genome_len, uint32_t n_labels)
{
    const int k = hp.k;
    const uint64_t per_genome = genome_len - k + 1;
    const uint64_t segs_per_genome = (per_genome + kSeg - 1) / kSeg;
    const uint64_t total = segs_per_genome * n_genomes;
    const uint64_t stride = gridDim.x * (uint64_t)blockDim.x;
    for (uint64_t s = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; s < total; s += stride) {
        const uint32_t g = (uint32_t)(s / segs_per_genome);
        const uint64_t p0 = (s % segs_per_genome) * kSeg;
        const uint64_t p1 = min(p0 + (uint64_t)kSeg, per_genome);
        const uint64_t gs = ccb_genome_seed(seed, g);
        uint64_t x = 0, rc = 0;
        for (int j = 0; j < k - 1; j++) x = (x << 2) | ccb_genome_code(gs, p0 + j);
        rc = revcomp(x, k);
        for (uint64_t p = p0; p < p1; p++) {
            uint64_t code = ccb_genome_code(gs, p + k - 1);
            x = ((x << 2) | code) & hp.mask;
            rc = (rc >> 2) | ((3ull - code) << (2 * (k - 1)));
            insert_one(buckets, flags, status, counts, hp, x < rc ? x : rc, g % n_labels);
        }
    }
}

__device__ __forceinline__ uint32_t slots_of_part(uint32_t len) { return 1u + ((len + 7u) >> 3); }

// parts of read i under the reference's packing rules (src/CuCLARK_hh.hh:1629-1707):
// maximal ACGT runs, runs shorter than k dropped, reads shorter than k empty.
__device__ __forceinline__ void read_parts(const ccb_read_plan& p, uint32_t L, int k, uint32_t (&beg)[2], uint32_t (&len)[2], int& n)
{
    n = 0;
    if (L < (uint32_t)k) return;
    if (p.kind == 2u) {
        if (p.n_pos >= (uint32_t)k) { beg[n] = 0; len[n] = p.n_pos; n++; }
        uint32_t r = L - 1 - p.n_pos;
        if (r >= (uint32_t)k) { beg[n] = p.n_pos + 1; len[n] = r; n++; }
    } else { beg[0] = 0; len[0] = L; n = 1; }
}

__global__ void k_synth_count(uint64_t rseed, uint64_t first, uint32_t n_reads, uint32_t L, int k,
                              uint32_t n_genomes, uint64_t genome_len, uint32_t* slots)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_reads) return;
    ccb_read_plan p = ccb_plan_read(rseed, first + i, n_genomes, genome_len, L);
    uint32_t b[2], l[2]; int n;
    read_parts(p, L, k, b, l, n);
    uint32_t s = 0;
    for (int j = 0; j < n; j++) s += slots_of_part(l[j]);
    slots[i] = s;
}

__global__ void k_synth_fill(uint64_t rseed, uint64_t dseed, uint64_t first, uint32_t n_reads, uint32_t L, int k,
                             uint32_t n_genomes, uint64_t genome_len, const uint32_t* offsets, uint16_t* cont)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_reads) return;
    ccb_read_plan p = ccb_plan_read(rseed, first + i, n_genomes, genome_len, L);
    uint32_t b[2], l[2]; int n;
    read_parts(p, L, k, b, l, n);
    uint16_t* o = cont + offsets[i];
    for (int j = 0; j < n; j++) {
        *o++ = (uint16_t)l[j];
        uint32_t acc = 0, cnt = 0;
        for (uint32_t q = 0; q < l[j]; q++) {
            acc = (acc << 2) | ccb_read_code(&p, dseed, b[j] + q, L);
            if (++cnt == 8) { *o++ = (uint16_t)acc; acc = 0; cnt = 0; }
        }
        if (cnt) *o++ = (uint16_t)(acc << (2 * (8 - cnt)));
    }
}

// ASCII text of the same reads (one line per read, '\n' terminated) for file-level tests
__global__ void k_synth_text(uint64_t rseed, uint64_t dseed, uint64_t first, uint32_t n_reads, uint32_t L,
                             uint32_t n_genomes, uint64_t genome_len, char* out)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_reads) return;
    ccb_read_plan p = ccb_plan_read(rseed, first + i, n_genomes, genome_len, L);
    char* o = out + (size_t)i * (L + 1);
    for (uint32_t q = 0; q < L; q++) o[q] = ccb_code_char(ccb_read_code(&p, dseed, q, L));
    o[L] = '\n';
}

}  // namespace

extern "C" int ccb_synth_db(ccb_handle* h, uint64_t db_seed, uint32_t n_genomes, uint64_t genome_len,
                            uint32_t n_labels, uint64_t* n_kmers)
{
    if (!h || n_genomes == 0 || genome_len < (uint64_t)h->cfg.k || n_labels == 0) return ccb_fail(h, -40, "ccb_synth_db: bad arguments");
    uint64_t total = (uint64_t)n_genomes * (genome_len - h->cfg.k + 1);
    int rc = ccb_db_begin(h, total);
    if (rc) return rc;
    Table& t = h->table;
    k_synth_db<<<148 * 16, 256, 0, h->db_stream>>>((uint64_t*)t.d_buckets - t.lo * kSlots, t.d_flags - t.lo / 16, t.d_status, t.d_counts, t.hp,
                                                    db_seed, n_genomes, genome_len, n_labels);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return ccb_fail(h, (int)e, "k_synth_db launch: %s", cudaGetErrorString(e));
    rc = ccb_db_end(h);
    if (rc) return rc;
    if (n_kmers) *n_kmers = total;
    return 0;
}

extern "C" int ccb_synth_reads_packed(ccb_handle* h, uint64_t reads_seed, uint64_t db_seed, uint64_t first_read,
                                      uint32_t n_reads, uint32_t read_len, uint32_t n_genomes, uint64_t genome_len,
                                      void** d_offsets, void** d_containers, uint64_t* n_containers)
{
    if (!h || !d_offsets || !d_containers || n_reads == 0 || genome_len < read_len) return ccb_fail(h, -41, "ccb_synth_reads_packed: bad arguments");
    cudaSetDevice(h->cfg.device);
    uint32_t *slots = nullptr, *off = nullptr; uint16_t* cont = nullptr; void* tmp = nullptr; size_t tmp_bytes = 0;
    cudaStream_t s = h->db_stream;
    cudaError_t e;
    if ((e = cudaMalloc(&slots, ((size_t)n_reads + 1) * 4)) != cudaSuccess || (e = cudaMalloc(&off, ((size_t)n_reads + 1) * 4)) != cudaSuccess)
        return ccb_fail(h, (int)e, "cudaMalloc: %s", cudaGetErrorString(e));
    cudaMemsetAsync(slots + n_reads, 0, 4, s);
    uint32_t g = (n_reads + 255) / 256;
    k_synth_count<<<g, 256, 0, s>>>(reads_seed, first_read, n_reads, read_len, h->cfg.k, n_genomes, genome_len, slots);
    tmp_bytes = scan_tmp_words((size_t)n_reads + 1) * 4;
    cudaMalloc(&tmp, tmp_bytes);
    cudaMemcpyAsync(off, slots, ((size_t)n_reads + 1) * 4, cudaMemcpyDeviceToDevice, s);
    exclusive_scan_u32(off, (size_t)n_reads + 1, (uint32_t*)tmp, s);
    uint32_t total = 0;
    cudaMemcpyAsync(&total, off + n_reads, 4, cudaMemcpyDeviceToHost, s);
    cudaStreamSynchronize(s);
    cudaFree(tmp); cudaFree(slots);
    if ((e = cudaMalloc(&cont, ((size_t)total + 8) * 2)) != cudaSuccess) { cudaFree(off); return ccb_fail(h, (int)e, "cudaMalloc: %s", cudaGetErrorString(e)); }
    k_synth_fill<<<g, 256, 0, s>>>(reads_seed, db_seed, first_read, n_reads, read_len, h->cfg.k, n_genomes, genome_len, off, cont);
    e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) { cudaFree(off); cudaFree(cont); return ccb_fail(h, (int)e, "synth reads: %s", cudaGetErrorString(e)); }
    *d_offsets = off; *d_containers = cont;
    if (n_containers) *n_containers = total;
    return 0;
}

extern "C" int ccb_synth_reads_text(ccb_handle* h, uint64_t reads_seed, uint64_t db_seed, uint64_t first_read,
                                    uint32_t n_reads, uint32_t read_len, uint32_t n_genomes, uint64_t genome_len,
                                    char* host_out)
{
    if (!h || !host_out || n_reads == 0) return ccb_fail(h, -42, "ccb_synth_reads_text: bad arguments");
    cudaSetDevice(h->cfg.device);
    char* d = nullptr; size_t bytes = (size_t)n_reads * (read_len + 1);
    cudaError_t e = cudaMalloc(&d, bytes);
    if (e != cudaSuccess) return ccb_fail(h, (int)e, "cudaMalloc: %s", cudaGetErrorString(e));
    k_synth_text<<<(n_reads + 255) / 256, 256, 0, h->db_stream>>>(reads_seed, db_seed, first_read, n_reads, read_len, n_genomes, genome_len, d);
    e = cudaMemcpyAsync(host_out, d, bytes, cudaMemcpyDeviceToHost, h->db_stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->db_stream);
    cudaFree(d);
    if (e != cudaSuccess) return ccb_fail(h, (int)e, "synth text: %s", cudaGetErrorString(e));
    return 0;
}

// ---------------------------------------------------------------------------
// The other configurations (include/ccb_synth.h: C3 pairs on a light database, C4 long reads with indels)
// ---------------------------------------------------------------------------
namespace {

// light database: of each genome the non-overlapping k-mers at 0, k, 2k, ... every gap-th (src/CuCLARK_hh.hh:711-767)
__global__ void k_synth_db_light(uint64_t* buckets, uint32_t* flags, uint32_t* status, unsigned long long* counts,
                                 HashParams hp, uint64_t seed, uint32_t n_genomes, uint64_t genome_len, uint32_t n_labels, uint32_t gap)
{
    const int k = hp.k;
    const uint64_t step = (uint64_t)gap * k;
    const uint64_t per_genome = (genome_len - k) / step + 1;
    const uint64_t total = per_genome * n_genomes;
    const uint64_t stride = gridDim.x * (uint64_t)blockDim.x;
    for (uint64_t s = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; s < total; s += stride) {
        const uint32_t g = (uint32_t)(s / per_genome);
        const uint64_t p0 = (s % per_genome) * step;
        const uint64_t gs = ccb_genome_seed(seed, g);
        uint64_t x = 0;
        for (int j = 0; j < k; j++) x = (x << 2) | ccb_genome_code(gs, p0 + j);
        insert_one(buckets, flags, status, counts, hp, canonical(x & hp.mask, k), g % n_labels);
    }
}

__global__ void k_synth2_text(ccb_workload w, uint64_t rseed, uint64_t dseed, uint64_t first, uint32_t n_reads,
                              uint32_t n_genomes, uint64_t genome_len, uint32_t stride, char* text, uint32_t* len)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_reads) return;
    len[i] = ccb_synth2_text(&w, rseed, dseed, first + i, n_genomes, genome_len, text + (size_t)i * stride);
}

// the reference's packing rules on a text (src/CuCLARK_hh.hh:1629-1707): maximal ACGT runs, runs shorter than k dropped
template <bool FILL>
__global__ void k_pack_text(const char* __restrict__ text, const uint32_t* __restrict__ len, uint32_t stride, uint32_t n_reads, int k,
                            uint32_t* slots, const uint32_t* offsets, uint16_t* cont)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_reads) return;
    const char* t = text + (size_t)i * stride;
    const uint32_t L = len[i];
    uint32_t total = 0;
    uint16_t* o = FILL ? cont + offsets[i] : nullptr;
    uint32_t q = 0;
    while (q < L) {
        while (q < L && t[q] == 'N') q++;
        uint32_t e = q;
        while (e < L && t[e] != 'N') e++;
        const uint32_t run = e - q;
        if (run >= (uint32_t)k) {
            total += slots_of_part(run);
            if (FILL) {
                *o++ = (uint16_t)run;
                uint32_t acc = 0, cnt = 0;
                for (uint32_t x = q; x < e; x++) {
                    const char c = t[x];
                    acc = (acc << 2) | (c == 'A' ? 3u : c == 'C' ? 2u : c == 'G' ? 1u : 0u);
                    if (++cnt == 8) { *o++ = (uint16_t)acc; acc = 0; cnt = 0; }
                }
                if (cnt) *o++ = (uint16_t)(acc << (2 * (8 - cnt)));
            }
        }
        q = e;
    }
    if (!FILL) slots[i] = total;
}

}  // namespace

extern "C" int ccb_synth_db_light(ccb_handle* h, uint64_t db_seed, uint32_t n_genomes, uint64_t genome_len,
                                  uint32_t n_labels, uint32_t gap, uint64_t* n_kmers)
{
    if (!h || n_genomes == 0 || genome_len < (uint64_t)h->cfg.k || n_labels == 0 || gap == 0) return ccb_fail(h, -40, "ccb_synth_db_light: bad arguments");
    const uint64_t total = (uint64_t)n_genomes * ((genome_len - h->cfg.k) / ((uint64_t)gap * h->cfg.k) + 1);
    int rc = ccb_db_begin(h, total);
    if (rc) return rc;
    Table& t = h->table;
    k_synth_db_light<<<148 * 16, 256, 0, h->db_stream>>>((uint64_t*)t.d_buckets - t.lo * kSlots, t.d_flags - t.lo / 16, t.d_status, t.d_counts, t.hp,
                                                          db_seed, n_genomes, genome_len, n_labels, gap);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return ccb_fail(h, (int)e, "k_synth_db_light launch: %s", cudaGetErrorString(e));
    rc = ccb_db_end(h);
    if (rc) return rc;
    if (n_kmers) *n_kmers = total;
    return 0;
}

extern "C" int ccb_synth2_reads_packed(ccb_handle* h, const ccb_workload* w, uint64_t reads_seed, uint64_t db_seed, uint64_t first_read,
                                       uint32_t n_reads, uint32_t n_genomes, uint64_t genome_len,
                                       void** d_offsets, void** d_containers, uint64_t* n_containers)
{
    if (!h || !w || !d_offsets || !d_containers || n_reads == 0 || (w->kind != 1u && w->kind != 2u) || w->read_len == 0 ||
        (w->kind == 1u && w->insert < w->read_len) || genome_len < ccb_synth2_span(w))
        return ccb_fail(h, -41, "ccb_synth2_reads_packed: bad arguments");
    cudaSetDevice(h->cfg.device);
    cudaStream_t s = h->db_stream;
    const uint32_t stride = ccb_synth2_max_len(w);
    char* text = nullptr; uint32_t *len = nullptr, *off = nullptr, *tmp = nullptr; uint16_t* cont = nullptr;
    cudaError_t e = cudaMalloc(&text, (size_t)n_reads * stride);
    if (e == cudaSuccess) e = cudaMalloc(&len, (size_t)n_reads * 4);
    if (e == cudaSuccess) e = cudaMalloc(&off, ((size_t)n_reads + 1) * 4);
    if (e == cudaSuccess) e = cudaMalloc(&tmp, scan_tmp_words((size_t)n_reads + 1) * 4);
    if (e != cudaSuccess) { cudaFree(text); cudaFree(len); cudaFree(off); cudaFree(tmp); return ccb_fail(h, (int)e, "cudaMalloc: %s", cudaGetErrorString(e)); }
    const uint32_t g = (n_reads + 127) / 128;
    k_synth2_text<<<g, 128, 0, s>>>(*w, reads_seed, db_seed, first_read, n_reads, n_genomes, genome_len, stride, text, len);
    cudaMemsetAsync(off + n_reads, 0, 4, s);
    k_pack_text<false><<<g, 128, 0, s>>>(text, len, stride, n_reads, h->cfg.k, off, nullptr, nullptr);
    exclusive_scan_u32(off, (size_t)n_reads + 1, tmp, s);
    uint32_t total = 0;
    cudaMemcpyAsync(&total, off + n_reads, 4, cudaMemcpyDeviceToHost, s);
    e = cudaStreamSynchronize(s);
    if (e == cudaSuccess) e = cudaMalloc(&cont, ((size_t)total + 8) * 2);
    if (e == cudaSuccess) {
        k_pack_text<true><<<g, 128, 0, s>>>(text, len, stride, n_reads, h->cfg.k, nullptr, off, cont);
        e = cudaStreamSynchronize(s);
    }
    cudaFree(text); cudaFree(len); cudaFree(tmp);
    if (e != cudaSuccess) { cudaFree(off); cudaFree(cont); return ccb_fail(h, (int)e, "synth reads: %s", cudaGetErrorString(e)); }
    *d_offsets = off; *d_containers = cont;
    if (n_containers) *n_containers = total;
    return 0;
}

extern "C" int ccb_device_free(ccb_handle* h, void* d_ptr)
{
    if (h) cudaSetDevice(h->cfg.device);
    cudaFree(d_ptr);
    return 0;
}

// ---------------------------------------------------------------------------
// Random-access roofline probe: n_probes independent aligned 32-byte loads at
// hashed positions of the resident table, 8 in flight per thread, nothing
// else.  The query kernel cannot beat the mode-0 number at the same footprint.
//   bytes   : 32 | 64 | 128 -- consecutive sectors fetched per probe by ONE thread
//   mode 0  : uniform over the first `footprint` bytes of the table (0 = all)
//   mode 1  : page-local -- the 32 lanes of a warp draw their sectors from one
//             random 2 MiB page per load instruction (isolates translation cost)
//   mode 2  : lane pairs read the two halves of a random 64-byte block
//   mode 3  : lane quads read the four sectors of a random 128-byte line
//   mode 4-7: mode 0 with other load flavours (plain / .cs / 2 x 16 B / L2 evict-first hint)
// ---------------------------------------------------------------------------
namespace {
template <int BYTES, int MODE>
__global__ void __launch_bounds__(256) k_gather(const ulonglong4* tbl, uint64_t n_sectors, uint64_t n_probes,
                                                 uint64_t salt, unsigned long long* sink)
{
    const uint64_t stride = gridDim.x * (uint64_t)blockDim.x;
    const uint64_t n_units = n_sectors / (BYTES / 32);
    unsigned long long acc = 0;
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n_probes; i += 8 * stride) {
        uint64_t e[8][4];
#pragma unroll
        for (int u = 0; u < 8; u++) {
            uint64_t idx = i + u * stride;
            uint64_t sector;
            if (MODE == 0 || MODE >= 4) {
                sector = __umul64hi(ccb_sm64(idx ^ salt), n_units) * (BYTES / 32);
            } else if (MODE == 1) {
                const uint64_t pages = n_sectors >> 16;                         // 2 MiB = 65536 sectors
                uint64_t pg = __umul64hi(ccb_sm64((idx >> 5) ^ salt), pages);   // same for the whole warp
                sector = (pg << 16) | (ccb_sm64(idx ^ salt ^ 0x77ull) & 0xFFFFull);
            } else if (MODE == 2) {
                sector = __umul64hi(ccb_sm64((idx >> 1) ^ salt), n_sectors >> 1) * 2 + (idx & 1);
            } else {
                sector = __umul64hi(ccb_sm64((idx >> 2) ^ salt), n_sectors >> 2) * 4 + (idx & 3);
            }
            const ulonglong4* p = tbl + sector;
            if (idx < n_probes && MODE == 4) {          // plain ld.global (L1-allocating, coherent path)
                asm volatile("ld.global.v4.u64 {%0,%1,%2,%3}, [%4];"
                             : "=l"(e[u][0]), "=l"(e[u][1]), "=l"(e[u][2]), "=l"(e[u][3]) : "l"(p));
            } else if (idx < n_probes && MODE == 5) {   // streaming (evict-first) load
                asm volatile("ld.global.cs.v4.u64 {%0,%1,%2,%3}, [%4];"
                             : "=l"(e[u][0]), "=l"(e[u][1]), "=l"(e[u][2]), "=l"(e[u][3]) : "l"(p));
            } else if (idx < n_probes && MODE == 6) {   // two 16-byte loads of the sector
                const ulonglong2* p2 = reinterpret_cast<const ulonglong2*>(p);
                asm volatile("ld.global.nc.v2.u64 {%0,%1}, [%2];" : "=l"(e[u][0]), "=l"(e[u][1]) : "l"(p2));
                asm volatile("ld.global.nc.v2.u64 {%0,%1}, [%2];" : "=l"(e[u][2]), "=l"(e[u][3]) : "l"(p2 + 1));
            } else if (idx < n_probes && MODE == 7) {   // L2 evict-first policy, no L1 allocation
                uint64_t pol;
                asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
                asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u64 {%0,%1,%2,%3}, [%4], %5;"
                             : "=l"(e[u][0]), "=l"(e[u][1]), "=l"(e[u][2]), "=l"(e[u][3]) : "l"(p), "l"(pol));
            } else if (idx < n_probes) {
                load_bucket(p, e[u]);
                if (BYTES >= 64) { uint64_t f[4]; load_bucket(p + 1, f); e[u][0] ^= f[0]; }
                if (BYTES >= 128) { uint64_t f[4]; load_bucket(p + 2, f); e[u][1] ^= f[1]; load_bucket(p + 3, f); e[u][2] ^= f[2]; }
            } else { e[u][0] = e[u][1] = e[u][2] = e[u][3] = 0; }
        }
#pragma unroll
        for (int u = 0; u < 8; u++) acc += e[u][0] ^ e[u][1] ^ e[u][2] ^ e[u][3];
    }
    if (acc == 0x123456789abcdefull) *sink = acc;                   // keep the loads alive
}
}  // namespace

// ---------------------------------------------------------------------------
// Line-reuse probes (what a minimizer-bucketed table would do): all four sectors of a random
// 128-byte line are wanted by ONE thread.
//   mode 8 : sector 0 first, sectors 1-3 only after it has arrived (dependent address)
//   mode 9 : prefetch.global.L2 of the lines of step j+1, then the 4 sectors of the lines of step j
//   mode 10: like 9 without the prefetch (all four sectors at once = bytes=128 mode 0, control)
// n_probes counts LINES.  LPT lines per thread per step.
// ---------------------------------------------------------------------------
namespace {
template <int MODE>
__global__ void __launch_bounds__(256) k_line_reuse(const ulonglong4* tbl, uint64_t n_lines, uint64_t n_probes,
                                                     uint64_t salt, unsigned long long* sink)
{
    constexpr int LPT = 4;
    const uint64_t stride = gridDim.x * (uint64_t)blockDim.x;
    const uint64_t tid = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    const uint64_t steps = n_probes / (stride * LPT);
    unsigned long long acc = 0;
    auto line_of = [&](uint64_t step, int u) { return __umul64hi(ccb_sm64(((step * LPT + u) * stride + tid) ^ salt), n_lines) * 4; };
    if (MODE == 9) {
#pragma unroll
        for (int u = 0; u < LPT; u++) asm volatile("prefetch.global.L2 [%0];" :: "l"(tbl + line_of(0, u)));
    }
    for (uint64_t st = 0; st < steps; st++) {
        if (MODE == 9 && st + 1 < steps) {
#pragma unroll
            for (int u = 0; u < LPT; u++) asm volatile("prefetch.global.L2 [%0];" :: "l"(tbl + line_of(st + 1, u)));
        }
        uint64_t e[LPT][4];
        if (MODE == 8) {
#pragma unroll
            for (int u = 0; u < LPT; u++) load_bucket(tbl + line_of(st, u), e[u]);
#pragma unroll
            for (int u = 0; u < LPT; u++) {
                const ulonglong4* p = tbl + line_of(st, u) + (e[u][0] == 0x1234567ull ? 1 : 0);   // depends on sector 0
                uint64_t f[4];
                load_bucket(p + 1, f); acc += f[0];
                load_bucket(p + 2, f); acc += f[1];
                load_bucket(p + 3, f); acc += f[2];
                acc += e[u][3];
            }
        } else {
#pragma unroll
            for (int q = 0; q < 4; q++) {
#pragma unroll
                for (int u = 0; u < LPT; u++) load_bucket(tbl + line_of(st, u) + q, e[u]);
#pragma unroll
                for (int u = 0; u < LPT; u++) acc += e[u][q];
            }
        }
    }
    if (acc == 0x123456789abcdefull) *sink = acc;
}
}  // namespace

// ---------------------------------------------------------------------------
// mode 11: every lane fetches whole random 128-byte lines with ONE bulk async copy each
// (cp.async.bulk global -> shared, completion on a per-warp mbarrier), LPT lines per lane per step.
// mode 12: the same lines with cooperative cp.async (8 lanes x 16 B per line, request list in smem).
// ---------------------------------------------------------------------------
namespace {
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int MODE>
__global__ void __launch_bounds__(128) k_line_bulk(const ulonglong4* tbl, uint64_t n_lines, uint64_t n_probes,
                                                    uint64_t salt, unsigned long long* sink)
{
    constexpr int LPT = 2, WARPS = 4;
    __shared__ alignas(128) unsigned char buf[WARPS][32 * LPT][128];
    __shared__ alignas(8) unsigned long long bar[WARPS];
    __shared__ uint32_t req[WARPS][32 * LPT];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint64_t stride = gridDim.x * (uint64_t)blockDim.x;
    const uint64_t tid = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    const uint64_t steps = n_probes / (stride * LPT);
    const uint32_t bar_a = smem_u32(&bar[w]);
    if (MODE == 11) {
        if (lane == 0) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(bar_a));
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
    }
    unsigned long long acc = 0;
    uint32_t parity = 0;
    for (uint64_t st = 0; st < steps; st++) {
        uint64_t line[LPT];
#pragma unroll
        for (int u = 0; u < LPT; u++) line[u] = __umul64hi(ccb_sm64(((st * LPT + u) * stride + tid) ^ salt), n_lines);
        if (MODE == 11) {
            if (lane == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar_a), "r"(32 * LPT * 128) : "memory");
            __syncwarp();
#pragma unroll
            for (int u = 0; u < LPT; u++)
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], 128, [%2];"
                             :: "r"(smem_u32(&buf[w][lane * LPT + u][0])), "l"(tbl + line[u] * 4), "r"(bar_a) : "memory");
            uint32_t done = 0;
            while (!done)
                asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                             : "=r"(done) : "r"(bar_a), "r"(parity) : "memory");
            parity ^= 1;
        } else {
#pragma unroll
            for (int u = 0; u < LPT; u++) req[w][lane * LPT + u] = (uint32_t)line[u];
            __syncwarp();
#pragma unroll 4
            for (int j = 0; j < 32 * LPT; j += 4) {
                const int li = j + (lane >> 3);
                const unsigned char* src = (const unsigned char*)(tbl + (uint64_t)req[w][li] * 4) + (lane & 7) * 16;
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(smem_u32(&buf[w][li][(lane & 7) * 16])), "l"(src) : "memory");
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
            asm volatile("cp.async.wait_group 0;" ::: "memory");
            __syncwarp();
        }
#pragma unroll
        for (int u = 0; u < LPT; u++) {
            const ulonglong2* q = (const ulonglong2*)&buf[w][lane * LPT + u][0];
            acc += q[0].x ^ q[3].y ^ q[5].x ^ q[7].y;
        }
        __syncwarp();
    }
    if (acc == 0x123456789abcdefull) *sink = acc;
}
}  // namespace

extern "C" int ccb_bench_gather_ex(ccb_handle* h, uint64_t n_probes, int bytes, int mode, uint64_t footprint, void* stream)
{
    if (!h || !h->db_ready) return ccb_fail(h, -32, "Failed to find the database.");
    cudaSetDevice(h->cfg.device);
    static unsigned long long* sink = nullptr;
    if (!sink) cudaMalloc(&sink, 8);
    const ulonglong4* tbl = (const ulonglong4*)h->table.d_buckets;
    uint64_t n32 = h->table.nb;
    if (footprint && footprint / 32 < n32) n32 = footprint / 32;
    if (n32 < (1u << 16)) return ccb_fail(h, -43, "ccb_bench_gather: footprint below 2 MiB");
    const int grid = 148 * 8;
    const uint64_t salt = 0x5bd1e995u;
    cudaStream_t s = (cudaStream_t)stream;
    if (mode == 0 && bytes == 32)       k_gather<32, 0><<<grid, 256, 0, s>>>(tbl, n32, n_probes, salt, sink);
    else if (mode == 0 && bytes == 64)  k_gather<64, 0><<<grid, 256, 0, s>>>(tbl, n32, n_probes, salt, sink);
    else if (mode == 0 && bytes == 128) k_gather<128, 0><<<grid, 256, 0, s>>>(tbl, n32, n_probes, salt, sink);
    else if (mode == 1 && bytes == 32)  k_gather<32, 1><<<grid, 256, 0, s>>>(tbl, n32, n_probes, salt, sink);
    else if (mode == 2 && bytes == 32)  k_gather<32, 2><<<grid, 256, 0, s>>>(tbl, n32, n_probes, salt, sink);
    else if (mode == 3 && bytes == 32)  k_gather<32, 3><<<grid, 256, 0, s>>>(tbl, n32, n_probes, salt, sink);
    else if (mode == 4 && bytes == 32)  k_gather<32, 4><<<grid, 256, 0, s>>>(tbl, n32, n_probes, salt, sink);
    else if (mode == 5 && bytes == 32)  k_gather<32, 5><<<grid, 256, 0, s>>>(tbl, n32, n_probes, salt, sink);
    else if (mode == 6 && bytes == 32)  k_gather<32, 6><<<grid, 256, 0, s>>>(tbl, n32, n_probes, salt, sink);
    else if (mode == 7 && bytes == 32)  k_gather<32, 7><<<grid, 256, 0, s>>>(tbl, n32, n_probes, salt, sink);
    else if (mode == 8)  k_line_reuse<8><<<grid, 256, 0, s>>>(tbl, n32 / 4, n_probes, salt, sink);
    else if (mode == 9)  k_line_reuse<9><<<grid, 256, 0, s>>>(tbl, n32 / 4, n_probes, salt, sink);
    else if (mode == 10) k_line_reuse<10><<<grid, 256, 0, s>>>(tbl, n32 / 4, n_probes, salt, sink);
    else if (mode == 11) k_line_bulk<11><<<148 * 3, 128, 0, s>>>(tbl, n32 / 4, n_probes, salt, sink);
    else if (mode == 12) k_line_bulk<12><<<148 * 3, 128, 0, s>>>(tbl, n32 / 4, n_probes, salt, sink);
    else return ccb_fail(h, -43, "ccb_bench_gather: unsupported bytes/mode (%d/%d)", bytes, mode);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return ccb_fail(h, (int)e, "k_gather launch: %s", cudaGetErrorString(e));
    return 0;
}

extern "C" int ccb_bench_gather(ccb_handle* h, uint64_t n_probes, int bytes, void* stream)
{
    return ccb_bench_gather_ex(h, n_probes, bytes, 0, 0, stream);
}
