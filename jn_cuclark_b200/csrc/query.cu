// query.cu -- fused k-mer extraction -> canonicalise -> probe -> vote.
//
// Replaces the reference's queryKernel + mergeKernel + resultKernel
// (src/CuClarkDB.cu:1000-1183, 1261-1355, 1361-1411) by one kernel.  k_query is the general kernel: reads of
// any length and any number of parts.  Short reads of one or two parts are taken by k_query_block
// (query_lane.cu) first, which hands everything else to k_query on a list (QueryArgs::read_list).
//
//  * one warp per read, the read is walked in tiles of 512 k-mer start
//    positions; each lane owns a CONTIGUOUS chunk of positions, so forward and
//    reverse-complement words roll (4 ALU ops per k-mer instead of the
//    reference's per-k-mer assembly + 5-stage bit reversal);
//  * the reference's packed batch layout is consumed as is: the container
//    stream IS the 2-bit stream, part headers / tail padding / part breaks
//    become an invalid-position bitmask, so N handling costs one AND per k-mer;
//  * 4 independent 32-byte probes are in flight per lane before any is
//    consumed (memory-level parallelism for the HBM random-access bound);
//  * votes never touch shared or global memory: labels are reduced with
//    warp shuffles into <=32 register accumulators (one (target,count) per
//    lane), then best / second best are two warp max-reductions on the key
//    (count desc, target asc) == the reference's ascending scan
//    (src/CuClarkDB.cu:1380-1399).
#include "query.cuh"
#include "vote.cuh"
#include <stdlib.h>

namespace ccb {

constexpr int kWarpsPerCta = 8;
constexpr int kTilePos     = 512;          // k-mer start positions per tile
constexpr int kTileSlots   = kTilePos / 8; // containers per tile
constexpr int kLoadSlots   = 72;           // 64 + halo (k-1 <= 31 bases -> 4 slots) rounded up
constexpr int kBaseWords   = kLoadSlots / 2;       // 36 x 16 bases
constexpr int kMaskWords   = 20;                   // 18 used + pad

struct WarpSmem {
    uint32_t base[kBaseWords];   // MSB-first: base p at word p>>4, bits 31-2(p&15)..
    uint32_t mask[kMaskWords];   // LSB-first: position p valid <=> bit p&31 of word p>>5
};

__device__ __forceinline__ uint32_t slot_of(const WarpSmem& w, int j)
{   // u16 container j of the tile
    return (w.base[j >> 1] >> (16 * (1 - (j & 1)))) & 0xFFFFu;
}

// Fold up to four labels per lane into the warp's register accumulators.
__device__ __forceinline__ void vote4(const uint32_t (&lab)[4], uint32_t& acc_lab, uint32_t& acc_cnt,
                                      int& nacc, bool& overflow, int lane)
{
    uint32_t pend = (lab[0] != kLabelNone ? 1u : 0u) | (lab[1] != kLabelNone ? 2u : 0u) |
                    (lab[2] != kLabelNone ? 4u : 0u) | (lab[3] != kLabelNone ? 8u : 0u);
    for (;;) {
        uint32_t b = __ballot_sync(0xFFFFFFFFu, pend != 0);
        if (b == 0) break;
        int src = __ffs(b) - 1;
        uint32_t mine = (pend & 1u) ? lab[0] : (pend & 2u) ? lab[1] : (pend & 4u) ? lab[2] : lab[3];
        uint32_t L = __shfl_sync(0xFFFFFFFFu, mine, src);
        uint32_t c = 0;
#pragma unroll
        for (int j = 0; j < 4; j++)
            if (((pend >> j) & 1u) && lab[j] == L) { c++; pend &= ~(1u << j); }
        uint32_t total = __reduce_add_sync(0xFFFFFFFFu, c);
        uint32_t hit = __ballot_sync(0xFFFFFFFFu, lane < nacc && acc_lab == L);
        if (hit) {
            if (lane == __ffs(hit) - 1) acc_cnt += total;
        } else if (nacc < 32) {
            if (lane == nacc) { acc_lab = L; acc_cnt = total; }
            nacc++;
        } else {
            overflow = true;
        }
    }
}

// Writes a read's result {sumN, idxBest, best, idxSecond, second} (src/CuClarkDB.cu:1370-1405), its
// sparse row when asked for, and lists it for the dense pass when it overflowed the accumulators.
__device__ __forceinline__ void finish_read(const QueryArgs& a, uint32_t r, uint32_t acc_lab, uint32_t acc_cnt,
                                            int nacc, bool overflow, int lane)
{
    const uint32_t cnt16 = lane < nacc ? (acc_cnt & 0xFFFFu) : 0u;   // reference counters are u16
    const uint32_t sum = __reduce_add_sync(0xFFFFFFFFu, cnt16);
    const uint32_t key = cnt16 ? ((cnt16 << 16) | (0xFFFFu - acc_lab)) : 0u;
    const uint32_t k1 = __reduce_max_sync(0xFFFFFFFFu, key);
    const uint32_t k2 = __reduce_max_sync(0xFFFFFFFFu, key == k1 ? 0u : key);
    if (lane < 5) {
        uint32_t v;
        if (lane == 0) v = sum;
        else if (lane == 1) v = k1 ? (0xFFFFu - (k1 & 0xFFFFu)) + 1u : 0u;
        else if (lane == 2) v = k1 >> 16;
        else if (lane == 3) v = k2 ? (0xFFFFu - (k2 & 0xFFFFu)) + 1u : 0u;
        else v = k2 >> 16;
        a.out5[(size_t)r * 5 + lane] = (uint16_t)v;
    }
    if (a.rows) {
        uint16_t* row = a.rows + (size_t)r * a.row_stride;
        int rank = 0;
        for (int j = 0; j < nacc; j++) {
            uint32_t l = __shfl_sync(0xFFFFFFFFu, acc_lab, j);
            rank += (l < acc_lab) ? 1 : 0;
        }
        if (lane == 0) row[0] = (uint16_t)min(nacc, (a.row_stride - 1) / 2);
        if (lane < nacc && 2 * rank + 2 < a.row_stride) {
            row[1 + 2 * rank] = (uint16_t)acc_lab;
            row[2 + 2 * rank] = (uint16_t)acc_cnt;
        }
    }
    if (overflow && lane == 0 && a.ovf_list) {
        uint32_t i = atomicAdd(a.ovf_count, 1u);
        a.ovf_list[i] = r;
    }
}

// MODE 0: classify against a table held whole (the product path).
// MODE 1/2: sharded table -- count, then write, each read's mixed k-mers per owner shard
//           (same walk, so both passes see the k-mers in the same order).
enum { kModeQuery = 0, kModeCount = 1, kModeRoute = 2 };

template <int MODE, bool STATS>
__global__ void __launch_bounds__(kWarpsPerCta * 32, 2)
k_query(const TableView t, const QueryArgs a, const RouteArgs ra)
{
    constexpr bool kProbe = MODE == kModeQuery;     // the mode that probes the table and votes
    __shared__ WarpSmem smem[kWarpsPerCta];
    const int lane = threadIdx.x & 31;
    const int wid  = threadIdx.x >> 5;
    WarpSmem& sw = smem[wid];
    const int k = t.hp.k;
    const uint64_t kmsk = t.hp.mask;
    const uint64_t validk = (k >= 64) ? ~0ull : ((1ull << k) - 1);
    const int s2 = 2 * (k - 1);
    const uint32_t total_warps = gridDim.x * kWarpsPerCta;

    // Software pipeline over the warp's reads: offsets are loaded two reads ahead, the first tile's
    // containers and the first part header one read ahead, so a read's three dependent loads
    // (offsets -> containers -> table) do not add up.
    const uint32_t n_items = a.read_list ? min(__ldg(a.n_list), a.n_reads) : a.n_reads;     // reads to do (all, or the listed ones)
    auto load_off = [&](uint32_t idx, uint32_t& b, int& n) {
        b = 0; n = 0;
        if (idx < n_items) {
            const uint32_t rr = a.read_list ? __ldg(a.read_list + idx) : idx;
            b = __ldg(a.offsets + rr); n = (int)(__ldg(a.offsets + rr + 1) - b);
        }
    };
    auto load_words = [&](uint32_t b, int n, uint32_t& xa, uint32_t& xb, uint32_t& xh) {
        const uint16_t* cc = a.containers + b;
        int s = 1 + 2 * lane;
        uint32_t v0 = s < n ? (uint32_t)__ldg(cc + s) : 0u, v1 = s + 1 < n ? (uint32_t)__ldg(cc + s + 1) : 0u;
        xa = (v0 << 16) | v1;
        xb = 0;
        if (lane < kBaseWords - 32) {
            s = 65 + 2 * lane;
            v0 = s < n ? (uint32_t)__ldg(cc + s) : 0u; v1 = s + 1 < n ? (uint32_t)__ldg(cc + s + 1) : 0u;
            xb = (v0 << 16) | v1;
        }
        xh = n > 0 ? (uint32_t)__ldg(cc) : 0u;
    };
    const uint32_t r_first = blockIdx.x * kWarpsPerCta + wid;
    uint32_t cur_beg, nxt_beg, pa, pb, ph; int cur_n, nxt_n;
    load_off(r_first, cur_beg, cur_n);
    load_off(r_first + total_warps, nxt_beg, nxt_n);
    load_words(cur_beg, cur_n, pa, pb, ph);

    for (uint32_t idx = r_first; idx < n_items; idx += total_warps) {
        const uint32_t r = a.read_list ? __ldg(a.read_list + idx) : idx;
        const uint32_t beg = cur_beg;
        const int nslots = cur_n;
        const uint32_t wa = pa, wb = pb, wh = ph;
        {
            uint32_t n2_beg; int n2_n;
            load_off(idx + 2 * total_warps, n2_beg, n2_n);
            load_words(nxt_beg, nxt_n, pa, pb, ph);
            cur_beg = nxt_beg; cur_n = nxt_n; nxt_beg = n2_beg; nxt_n = n2_n;
        }
        const uint16_t* c = a.containers + beg;
        uint32_t acc_lab = kLabelNone, acc_cnt = 0;
        int nacc = 0; bool overflow = false;
        unsigned long long st_lookups = 0; uint32_t st_sectors = 0;
        uint32_t sent[kMaxShards];                 // MODE 1/2: k-mers of this read per owner so far
        uint64_t* out_g[kMaxShards];               // MODE 2: where this read's k-mers for owner g start
        uint32_t room[kMaxShards];                 // MODE 2: how many of them fit (capacity-bounded peer inbox)
#pragma unroll
        for (int g = 0; g < kMaxShards; g++) {
            sent[g] = 0; out_g[g] = nullptr; room[g] = 0xFFFFFFFFu;
            if (MODE == kModeRoute && g < (int)ra.n_shards) {
                const uint64_t base = ra.derive_base ? (uint64_t)__ldg(ra.pos + (size_t)g * a.n_reads) : ra.seg_base[g];
                out_g[g] = ra.dst[g] + (ra.dst_off ? ra.dst_off[g] : 0ull) + (__ldg(ra.pos + (size_t)g * a.n_reads + r) - base);
                if (ra.cap) {                            // elements of this read that still fit owner g's segment
                    const uint64_t first = (uint64_t)(out_g[g] - ra.dst[g]);
                    room[g] = first < ra.cap ? (uint32_t)min(ra.cap - first, (uint64_t)0xFFFFFFFFu) : 0u;
                }
            }
        }

        const int total_pos = 8 * (nslots - 1) - (k - 1);
        if (total_pos > 0) {
            int ps = 0, pl = (int)wh;                // current part: header slot, length
            const int ntiles = (total_pos + kTilePos - 1) / kTilePos;
            for (int tile = 0; tile < ntiles; tile++) {
                const int slot0 = 1 + tile * kTileSlots;        // first container of the tile
                // ---- stage the tile's containers: 2 x u16 -> one MSB-first word ----
                __syncwarp();
                if (tile == 0) {
                    sw.base[lane] = wa;
                    if (lane < kBaseWords - 32) sw.base[lane + 32] = wb;
                } else {
                    for (int w = lane; w < kBaseWords; w += 32) {
                        int s = slot0 + 2 * w;
                        uint32_t v0 = s < nslots ? (uint32_t)c[s] : 0u;
                        uint32_t v1 = s + 1 < nslots ? (uint32_t)c[s + 1] : 0u;
                        sw.base[w] = (v0 << 16) | v1;
                    }
                }
                __syncwarp();
                // ---- valid-position mask from the part walk (src/CuClarkDB.cu:1042-1047) ----
                {
                    uint32_t m = 0;
                    const int next_tile_slot0 = slot0 + kTileSlots;
                    const int w0 = 32 * lane;            // tile-relative position of my mask word
                    int cs = ps, cl = pl; bool have_c = false;
                    for (;;) {
                        int lo = 8 * ps - kTilePos * tile, hi = lo + pl;   // valid range, tile-relative
                        int x0 = max(lo, w0) - w0, x1 = min(hi, w0 + 32) - w0;
                        if (x0 < x1) m |= (x1 - x0 == 32) ? 0xFFFFFFFFu : (((1u << (x1 - x0)) - 1u) << x0);
                        int next = ps + 1 + ((pl + 7) >> 3);
                        if (!have_c && next > next_tile_slot0) { cs = ps; cl = pl; have_c = true; }
                        if (next >= slot0 + kLoadSlots || next >= nslots) break;
                        ps = next; pl = (int)slot_of(sw, next - slot0);
                    }
                    if (have_c) { ps = cs; pl = cl; }
                    if (lane < kMaskWords) sw.mask[lane] = m;
                }
                __syncwarp();
                // ---- my chunk of positions ----
                const int npos = min(kTilePos, total_pos - tile * kTilePos);
                const int chunk = 4 * ((npos + 127) >> 7);       // multiple of 4, <= 16
                const int p0 = lane * chunk;
                uint64_t hi, lo, vm;
                {
                    int j = p0 >> 4, sh = 2 * (p0 & 15);
                    uint32_t a0 = sw.base[j], a1 = sw.base[j + 1], a2 = sw.base[j + 2],
                             a3 = sw.base[j + 3], a4 = sw.base[min(j + 4, kBaseWords - 1)];
                    uint32_t b0 = __funnelshift_l(a1, a0, sh), b1 = __funnelshift_l(a2, a1, sh),
                             b2 = __funnelshift_l(a3, a2, sh), b3 = __funnelshift_l(a4, a3, sh);
                    hi = ((uint64_t)b0 << 32) | b1; lo = ((uint64_t)b2 << 32) | b3;
                    int jm = p0 >> 5, shm = p0 & 31;
                    uint32_t m0 = sw.mask[jm], m1 = sw.mask[jm + 1], m2 = sw.mask[jm + 2];
                    vm = ((uint64_t)__funnelshift_r(m1, m2, shm) << 32) | __funnelshift_r(m0, m1, shm);
                }
                uint64_t x = hi >> (64 - s2);                       // first k-1 bases
                uint64_t rc = revcomp(x, k);
                uint64_t rest_hi = (hi << s2) | (lo >> (64 - s2)), rest_lo = lo << s2;
                const int ngroups = chunk >> 2;
                for (int g = 0; g < ngroups; g++) {
                    uint64_t e[4][4]; uint64_t home[4], q[4]; bool ok[4];
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        const int i = 4 * g + j;
                        uint64_t code = rest_hi >> 62;
                        rest_hi = (rest_hi << 2) | (rest_lo >> 62); rest_lo <<= 2;
                        x = ((x << 2) | code) & kmsk;
                        rc = (rc >> 2) | ((3ull - code) << s2);
                        ok[j] = (p0 + i < npos) && (((vm >> i) & validk) == validk);
                        uint64_t cn = x < rc ? x : rc;
                        uint64_t h = mix2k(cn, k, kmsk);
                        if (kProbe) {
                            home[j] = home_of(h, t.hp);
                            q[j] = key48_of(h) << 16;
                            if (ok[j]) load_bucket(t.buckets + home[j], e[j]);
                        } else {
                            q[j] = h;
                            home[j] = ok[j] ? owner_of(h, t.hp.shift, ra.n_shards) : 0xFFFFFFFFull;
                        }
                    }
                    if (kProbe) {
                        uint32_t lab[4];
#pragma unroll
                        for (int j = 0; j < 4; j++) {
                            lab[j] = kLabelNone;
                            if (ok[j]) {
                                uint32_t l = resolve(t, home[j], q[j], e[j], STATS ? &st_sectors : nullptr);
                                lab[j] = l >= kLabelTomb ? kLabelNone : l;
                                if (STATS) { st_lookups++; st_sectors++; }
                            }
                        }
                        vote4(lab, acc_lab, acc_cnt, nacc, overflow, lane);
                    } else {
                        const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
                        for (int j = 0; j < 4; j++) {
#pragma unroll
                            for (int og = 0; og < kMaxShards; og++) {
                                if (og < (int)ra.n_shards) {
                                    const bool mine = home[j] == (uint64_t)og;
                                    const uint32_t b = __ballot_sync(0xFFFFFFFFu, mine);
                                    if (MODE == kModeRoute && mine) {
                                        const uint32_t at = sent[og] + __popc(b & lt);
                                        if (at < room[og]) out_g[og][at] = q[j];
                                    }
                                    sent[og] += __popc(b);
                                }
                            }
                        }
                    }
                }
            }
        }
        if (kProbe) {
            finish_read(a, r, acc_lab, acc_cnt, nacc, overflow, lane);
        } else if (MODE == kModeCount) {
            uint32_t mine = 0;
#pragma unroll
            for (int g = 0; g < kMaxShards; g++) if (lane == g) mine = sent[g];
            if (lane < (int)ra.n_shards) ra.counts[(size_t)lane * a.n_reads + r] = mine;
        }
        if (STATS) {
            unsigned long long l = st_lookups, s = st_sectors;
#pragma unroll
            for (int o = 16; o; o >>= 1) {
                l += __shfl_xor_sync(0xFFFFFFFFu, l, o);
                s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
            }
            if (lane == 0) { atomicAdd(a.stats + 0, l); atomicAdd(a.stats + 1, s); if (a.stats_words > 2) atomicAdd(a.stats + 2, l); }
        }
    }
    if (MODE == kModeRoute) __threadfence_system();      // the stores may have gone to a peer GPU
}

int launch_query(const TableView& t, const QueryArgs& a, cudaStream_t s, uint32_t max_ctas)
{
    if (a.n_reads == 0) return 0;
    uint32_t ctas = (a.n_reads + kWarpsPerCta - 1) / kWarpsPerCta;
    // 2 CTAs are resident per SM; the grid holds 64x that many so the block scheduler
    // rebalances between fast and slow SMs (measured: 2/SM 36.1 ms, 8/SM 31.9, 128/SM 29.2
    // per 10 M reads -- a static split leaves the slower SMs as the tail).
    uint32_t cap = max_ctas ? max_ctas : 148u * 128u;
    if (ctas > cap) ctas = cap;
    const RouteArgs none{};
    if (a.stats) k_query<kModeQuery, true><<<ctas, kWarpsPerCta * 32, 0, s>>>(t, a, none);
    else         k_query<kModeQuery, false><<<ctas, kWarpsPerCta * 32, 0, s>>>(t, a, none);
    return (int)cudaGetLastError();
}


int launch_route(const TableView& t, const QueryArgs& a, const RouteArgs& ra, bool write, cudaStream_t s, uint32_t max_ctas)
{
    if (a.n_reads == 0) return 0;
    if (ra.n_shards < 1 || ra.n_shards > (uint32_t)kMaxShards) return -60;
    uint32_t ctas = (a.n_reads + kWarpsPerCta - 1) / kWarpsPerCta;
    uint32_t cap = max_ctas ? max_ctas : 148u * 128u;
    if (ctas > cap) ctas = cap;
    if (write) k_query<kModeRoute, false><<<ctas, kWarpsPerCta * 32, 0, s>>>(t, a, ra);
    else       k_query<kModeCount, false><<<ctas, kWarpsPerCta * 32, 0, s>>>(t, a, ra);
    return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------
// dense exact path (rare): the reference's own counting scheme
// (one counter per target, src/CuClarkDB.cu:1017-1027,1104-1108) but with the
// counters in global scratch, so any number of distinct targets is exact.
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint64_t kmer_at(const uint16_t* data, int j, int k, uint64_t mask)
{   // bases j..j+k-1 of a part whose containers start at data (src/CuClarkDB.cu:1063-1083)
    uint64_t x = 0;
    for (int b = 0; b < k; b++) {
        int p = j + b;
        x = (x << 2) | ((data[p >> 3] >> (2 * (7 - (p & 7)))) & 3u);
    }
    return x & mask;
}

__global__ void __launch_bounds__(256)
k_query_dense(const TableView t, const QueryArgs a, const uint32_t* read_list, const uint32_t* d_n_list,
              uint32_t max_list, uint32_t n_targets, uint32_t* scratch)
{
    __shared__ uint32_t s_key[2][256];
    __shared__ uint32_t s_sum[256];
    uint32_t n_list = d_n_list ? min(*d_n_list, max_list) : max_list;
    uint32_t* cnt = scratch + (size_t)blockIdx.x * n_targets;
    const int k = t.hp.k;
    for (uint32_t li = blockIdx.x; li < n_list; li += gridDim.x) {
        const uint32_t r = read_list ? read_list[li] : li;
        for (uint32_t i = threadIdx.x; i < n_targets; i += blockDim.x) cnt[i] = 0;
        __syncthreads();
        const uint32_t beg = a.offsets[r], end = a.offsets[r + 1];
        uint32_t p = beg;
        while (p < end) {
            int len = a.containers[p];
            const uint16_t* data = a.containers + p + 1;
            p += 1 + ((len + 7) >> 3);
            for (int j = threadIdx.x; j + k <= len; j += blockDim.x) {
                uint64_t x = kmer_at(data, j, k, t.hp.mask);
                uint32_t lab = lookup(t, canonical(x, k));
                if (lab != kLabelNone && lab < n_targets) atomicAdd(cnt + lab, 1u);   // a label the handle was not sized for is ignored
            }
        }
        __syncthreads();
        dense_result(a, r, cnt, n_targets, s_key, s_sum);
        __syncthreads();
    }
}

int launch_query_dense(const TableView& t, const QueryArgs& a, const uint32_t* read_list,
                       const uint32_t* d_n_list, uint32_t max_list, uint32_t n_targets,
                       uint32_t* scratch, uint32_t n_scratch_blocks, cudaStream_t s)
{
    if (max_list == 0 || n_scratch_blocks == 0) return 0;
    uint32_t g = max_list < n_scratch_blocks ? max_list : n_scratch_blocks;
    k_query_dense<<<g, 256, 0, s>>>(t, a, read_list, d_n_list, max_list, n_targets, scratch);
    return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------
// sharded table: owner-side probe of a k-mer list, requester-side vote from label lists
// ---------------------------------------------------------------------------
// A table (or shard) larger than what the TLBs reach is probed range by range: measured with the table-only kernel,
// 1.2 G random probes take 27.3 ms on a 64 GB table, 60.5 ms on 96 GB and 96.0 ms on 128 GB (profiles/r02_table_size.json)
// -- beyond 64 GB every probe also misses the translation caches (70 GB: 28.0 ms, 77 GB: 31.5 ms: the cliff starts near
// 72 GB, profiles/r02_probe_ranges.json).  A list of k-mers can be walked once per range instead; a pass costs 5-7 ms per
// 1.2 G k-mers whatever it probes, so the ranges are as large as the caches reach: <= 68 GB (a 134 GB table in two
// ranges of 67 GB: 36.1 ms, in three of 45 GB: 41.1 ms, in one: 105 ms).
static void probe_passes(const TableView& t, uint64_t& lo, uint64_t& per, int& passes)
{
    const uint64_t G = t.hp.n_shards > 1 ? t.hp.n_shards : 1;
    const uint64_t held = t.hp.nb / G;                                   // home buckets held here
    lo = held * (t.hp.n_shards > 1 ? t.hp.shard : 0);
    const char* env = getenv("CCB_PROBE_RANGE_MB");                      // (tests: ranges of a few MB on a small table)
    const uint64_t reach = env && atoll(env) > 0 ? (uint64_t)atoll(env) << 20 : 68ull << 30;
    passes = (int)((held * 32 + reach - 1) / reach);
    if (passes < 1) passes = 1;
    per = (held + (uint64_t)passes - 1) / (uint64_t)passes;
}

// 4 independent probes in flight per thread, as in k_query.
// home_lo / home_hi: only k-mers whose home bucket lies in [home_lo, home_hi) are probed by this launch (a shard larger
// than the TLB reaches is probed range by range, see launch_probe_list)
template <bool STATS>
__global__ void __launch_bounds__(256) k_probe_list(const TableView t, const uint64_t* __restrict__ hashes, uint64_t n,
                                                    uint16_t* __restrict__ labels, unsigned long long* stats,
                                                    const uint64_t home_lo, const uint64_t home_hi)
{
    const uint64_t stride = gridDim.x * (uint64_t)blockDim.x;
    uint32_t sectors = 0; unsigned long long lookups = 0;
    for (uint64_t i0 = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i0 < n; i0 += 4 * stride) {
        uint64_t e[4][4], home[4], q[4]; bool ok[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const uint64_t i = i0 + j * stride;
            ok[j] = i < n;
            const uint64_t h = ok[j] ? __ldg(hashes + i) : 0;
            home[j] = home_of(h, t.hp); q[j] = key48_of(h) << 16;
            ok[j] = ok[j] && home[j] >= home_lo && home[j] < home_hi;
            if (ok[j]) load_bucket(t.buckets + home[j], e[j]);
        }
#pragma unroll
        for (int j = 0; j < 4; j++) {
            if (ok[j]) {
                const uint32_t l = resolve(t, home[j], q[j], e[j], STATS ? &sectors : nullptr);
                labels[i0 + j * stride] = (uint16_t)(l >= kLabelTomb ? kLabelNone : l);
                if (STATS) { lookups++; sectors++; }
            }
        }
    }
    if (STATS) {
        unsigned long long l = lookups, sc = sectors;
#pragma unroll
        for (int o = 16; o; o >>= 1) { l += __shfl_xor_sync(0xFFFFFFFFu, l, o); sc += __shfl_xor_sync(0xFFFFFFFFu, sc, o); }
        if ((threadIdx.x & 31) == 0) { atomicAdd(stats + 0, l); atomicAdd(stats + 1, sc); }
    }
}

int launch_probe_list(const TableView& t, const uint64_t* hashes, uint64_t n, uint16_t* labels,
                      unsigned long long* stats, cudaStream_t s)
{
    if (n == 0) return 0;
    uint64_t g = (n + 1023) / 1024;
    if (g > 148ull * 64) g = 148ull * 64;
    uint64_t lo, per; int passes;
    probe_passes(t, lo, per, passes);
    for (int v = 0; v < passes; v++) {
        const uint64_t a = lo + per * (uint64_t)v, b = v + 1 == passes ? ~0ull : a + per;
        if (stats) k_probe_list<true><<<(uint32_t)g, 256, 0, s>>>(t, hashes, n, labels, stats, v ? a : 0ull, b);
        else       k_probe_list<false><<<(uint32_t)g, 256, 0, s>>>(t, hashes, n, labels, stats, v ? a : 0ull, b);
    }
    return (int)cudaGetLastError();
}

// one warp per read: the read's labels are n_shards contiguous runs
__global__ void __launch_bounds__(kWarpsPerCta * 32) k_vote_lists(const QueryArgs a, const VoteArgs va)
{
    const int lane = threadIdx.x & 31;
    const uint32_t total_warps = gridDim.x * kWarpsPerCta;
    const uint32_t n_items = a.read_list ? min(__ldg(a.n_list), a.n_reads) : a.n_reads;
    for (uint32_t idx = blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5); idx < n_items; idx += total_warps) {
        const uint32_t r = a.read_list ? __ldg(a.read_list + idx) : idx;
        uint32_t acc_lab = kLabelNone, acc_cnt = 0;
        int nacc = 0; bool overflow = false;
        for (uint32_t g = 0; g < va.n_shards; g++) {
            const size_t pi = (size_t)g * a.n_reads + r;
            const uint32_t b = __ldg(va.pos + pi), e = __ldg(va.pos + pi + 1);
            const uint64_t base = va.derive_base ? (uint64_t)__ldg(va.pos + (size_t)g * a.n_reads) : va.seg_base[g];
            const uint64_t first = (va.src_off ? va.src_off[g] : 0ull) + ((uint64_t)b - base);
            const uint16_t* src = va.src[g] + first;
            uint32_t cnt = e - b;
            if (va.cap) cnt = first < va.cap ? (uint32_t)min((uint64_t)cnt, va.cap - first) : 0u;
            for (uint32_t o = 0; o < cnt; o += 128) {
                uint32_t lab[4];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const uint32_t i = o + 32 * j + lane;
                    lab[j] = i < cnt ? (uint32_t)__ldg(src + i) : kLabelNone;
                    if (lab[j] >= kLabelTomb) lab[j] = kLabelNone;
                }
                vote4(lab, acc_lab, acc_cnt, nacc, overflow, lane);
            }
        }
        finish_read(a, r, acc_lab, acc_cnt, nacc, overflow, lane);
    }
}

int launch_vote_lists(const QueryArgs& a, const VoteArgs& va, cudaStream_t s)
{
    if (a.n_reads == 0) return 0;
    uint32_t ctas = (a.n_reads + kWarpsPerCta - 1) / kWarpsPerCta;
    if (ctas > 148u * 64u) ctas = 148u * 64u;
    k_vote_lists<<<ctas, kWarpsPerCta * 32, 0, s>>>(a, va);
    return (int)cudaGetLastError();
}

// exact path for reads that hit more than 32 distinct targets: dense counters in global scratch
__global__ void __launch_bounds__(256)
k_vote_lists_dense(const QueryArgs a, const VoteArgs va, const uint32_t* read_list, const uint32_t* d_n_list,
                   uint32_t max_list, uint32_t n_targets, uint32_t* scratch)
{
    __shared__ uint32_t s_key[2][256];
    __shared__ uint32_t s_sum[256];
    uint32_t n_list = d_n_list ? min(*d_n_list, max_list) : max_list;
    uint32_t* cnt = scratch + (size_t)blockIdx.x * n_targets;
    for (uint32_t li = blockIdx.x; li < n_list; li += gridDim.x) {
        const uint32_t r = read_list ? read_list[li] : li;
        for (uint32_t i = threadIdx.x; i < n_targets; i += blockDim.x) cnt[i] = 0;
        __syncthreads();
        for (uint32_t g = 0; g < va.n_shards; g++) {
            const size_t pi = (size_t)g * a.n_reads + r;
            const uint32_t b = va.pos[pi], e = va.pos[pi + 1];
            const uint64_t base = va.derive_base ? (uint64_t)va.pos[(size_t)g * a.n_reads] : va.seg_base[g];
            const uint64_t first = (va.src_off ? va.src_off[g] : 0ull) + ((uint64_t)b - base);
            const uint16_t* src = va.src[g] + first;
            uint32_t n_lab = e - b;
            if (va.cap) n_lab = first < va.cap ? (uint32_t)min((uint64_t)n_lab, va.cap - first) : 0u;
            for (uint32_t i = threadIdx.x; i < n_lab; i += blockDim.x) {
                const uint32_t lab = src[i];
                if (lab < kLabelTomb && lab < n_targets) atomicAdd(cnt + lab, 1u);
            }
        }
        __syncthreads();
        dense_result(a, r, cnt, n_targets, s_key, s_sum);
        __syncthreads();
    }
}

int launch_vote_lists_dense(const QueryArgs& a, const VoteArgs& va, const uint32_t* read_list, const uint32_t* d_n_list,
                            uint32_t max_list, uint32_t n_targets, uint32_t* scratch, uint32_t n_scratch_blocks, cudaStream_t s)
{
    if (max_list == 0 || n_scratch_blocks == 0) return 0;
    uint32_t g = max_list < n_scratch_blocks ? max_list : n_scratch_blocks;
    k_vote_lists_dense<<<g, 256, 0, s>>>(a, va, read_list, d_n_list, max_list, n_targets, scratch);
    return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------
// peer exchange: counts to the owners, owner-side probe straight out of the inbox
// ---------------------------------------------------------------------------
struct CountDst { unsigned long long* p[kMaxShards]; };

__global__ void k_post_counts(const uint32_t* __restrict__ pos, uint32_t n_reads, uint32_t n_shards, uint64_t cap,
                              CountDst dst, uint32_t* error, const unsigned long long* first)
{
    const uint32_t g = threadIdx.x;
    if (g >= n_shards) return;
    const uint64_t n = (first ? first[g] : 0ull) + ((uint64_t)pos[(size_t)(g + 1) * n_reads] - pos[(size_t)g * n_reads]);
    if (cap && n > cap) { if (error) atomicExch(error, 1u); *dst.p[g] = cap; }
    else *dst.p[g] = n;
    __threadfence_system();
}

int launch_post_counts(const uint32_t* pos, uint32_t n_reads, uint32_t n_shards, uint64_t cap,
                       unsigned long long* const* count_dst, uint32_t* error, cudaStream_t s, const unsigned long long* first)
{
    CountDst d{};
    for (uint32_t g = 0; g < n_shards; g++) d.p[g] = count_dst[g];
    k_post_counts<<<1, 32, 0, s>>>(pos, n_reads, n_shards, cap, d, error, first);
    return (int)cudaGetLastError();
}

template <bool STATS>
__global__ void __launch_bounds__(256) k_probe_inbox(const TableView t, const InboxArgs ia, unsigned long long* stats,
                                                     const uint64_t home_lo, const uint64_t home_hi)
{
    const uint64_t stride = gridDim.x * (uint64_t)blockDim.x;
    uint32_t sectors = 0; unsigned long long lookups = 0;
    for (uint32_t sidx = 0; sidx < ia.n_shards; sidx++) {
        const uint64_t n = ia.counts_in[sidx];
        const uint64_t* __restrict__ hashes = ia.inbox + (size_t)sidx * ia.cap;
        uint16_t* __restrict__ labels = ia.lab_dst[sidx];
        for (uint64_t i0 = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i0 < n; i0 += 4 * stride) {
            uint64_t e[4][4], home[4], q[4]; bool ok[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const uint64_t i = i0 + j * stride;
                ok[j] = i < n;
                const uint64_t h = ok[j] ? hashes[i] : 0;
                home[j] = home_of(h, t.hp); q[j] = key48_of(h) << 16;
                ok[j] = ok[j] && home[j] >= home_lo && home[j] < home_hi;
                if (ok[j]) load_bucket(t.buckets + home[j], e[j]);
            }
#pragma unroll
            for (int j = 0; j < 4; j++) {
                if (ok[j]) {
                    const uint32_t l = resolve(t, home[j], q[j], e[j], STATS ? &sectors : nullptr);
                    labels[i0 + j * stride] = (uint16_t)(l >= kLabelTomb ? kLabelNone : l);
                    if (STATS) { lookups++; sectors++; }
                }
            }
        }
    }
    __threadfence_system();
    if (STATS) {
        unsigned long long l = lookups, sc = sectors;
#pragma unroll
        for (int o = 16; o; o >>= 1) { l += __shfl_xor_sync(0xFFFFFFFFu, l, o); sc += __shfl_xor_sync(0xFFFFFFFFu, sc, o); }
        if ((threadIdx.x & 31) == 0) { atomicAdd(stats + 0, l); atomicAdd(stats + 1, sc); }
    }
}

int launch_probe_inbox(const TableView& t, const InboxArgs& ia, unsigned long long* stats, cudaStream_t s)
{
    // grid: 32 blocks per SM's worth (measured on one B200, 10 M reads: 2 per SM 36.8 ms, 4: 33.8, 8: 31.1, 32: 30.1 --
    // small persistent grids that would leave room for another batch's route kernel cost more than the overlap returns,
    // profiles/r02_sharded_grid_sweep.json)
    static const int per_sm = getenv("CCB_PROBE_CTAS_PER_SM") ? atoi(getenv("CCB_PROBE_CTAS_PER_SM")) : 32;
    uint64_t lo, per; int passes;
    probe_passes(t, lo, per, passes);
    for (int v = 0; v < passes; v++) {
        const uint64_t a = lo + per * (uint64_t)v, b = v + 1 == passes ? ~0ull : a + per;
        if (stats) k_probe_inbox<true><<<148 * per_sm, 256, 0, s>>>(t, ia, stats, v ? a : 0ull, b);
        else       k_probe_inbox<false><<<148 * per_sm, 256, 0, s>>>(t, ia, stats, v ? a : 0ull, b);
    }
    return (int)cudaGetLastError();
}

}  // namespace ccb
