// query_lane.cu -- thread-per-read classification of short reads.
//
// k_query (query.cu) gives a whole warp to one read; most of its instructions coordinate the 32 lanes
// (tile staging, validity masks, shuffle-merged votes).  For the reads that dominate a metagenomic
// sample -- one part, <= 192 bases, a handful of targets hit -- none of that is needed: here every lane
// walks its own read with rolling forward / reverse-complement words, keeps 4 probes in flight, and votes
// into four (label,count) register pairs.  A warp holds 32 reads, so there are 32x more independent
// requests to hide the DRAM latency behind, and with the anchor index (query.cuh AnchorView) the windows
// that match the target sequence base for base are answered without touching the table at all.
// Anything else -- several parts, longer reads, a fifth distinct target -- is handed back on a list and
// classified by k_query.  Results are the same u16[5] / sparse rows either way.
#include "query.cuh"
#include <stdlib.h>

namespace ccb {

constexpr int kLaneMaxBases = 192;
constexpr int kLaneAcc = 4;

namespace {

// 16 bases compared: diff = text word ^ read word (2 bits per base, first base in the top bits)
// -> 16 bits, bit b set iff base b differs
__device__ __forceinline__ uint32_t mismatch16(uint32_t diff)
{
    uint32_t v = __brev((diff | (diff >> 1)) & 0x55555555u) >> 1;     // base b -> bit 2b
    v &= 0x55555555u;
    v = (v | (v >> 1)) & 0x33333333u;
    v = (v | (v >> 2)) & 0x0F0F0F0Fu;
    v = (v | (v >> 4)) & 0x00FF00FFu;
    v = (v | (v >> 8)) & 0x0000FFFFu;
    return v;
}

struct Bits192 {                       // bit i <-> position i, LSB first
    uint64_t w[3];
    __device__ __forceinline__ Bits192 shr(int s) const          // 0 < s < 64
    {
        Bits192 r;
        r.w[0] = (w[0] >> s) | (w[1] << (64 - s));
        r.w[1] = (w[1] >> s) | (w[2] << (64 - s));
        r.w[2] = w[2] >> s;
        return r;
    }
    __device__ __forceinline__ void and_with(const Bits192& o) { w[0] &= o.w[0]; w[1] &= o.w[1]; w[2] &= o.w[2]; }
    __device__ __forceinline__ bool get(int i) const
    {
        const uint64_t v = i < 64 ? w[0] : (i < 128 ? w[1] : w[2]);
        return (v >> (i & 63)) & 1ull;
    }
};

struct LaneVotes {
    uint32_t lab[kLaneAcc], cnt[kLaneAcc];
    int n; bool overflow;
    __device__ __forceinline__ void init() { n = 0; overflow = false;
#pragma unroll
        for (int i = 0; i < kLaneAcc; i++) { lab[i] = kLabelNone; cnt[i] = 0; } }
    __device__ __forceinline__ void add(uint32_t L, uint32_t c)
    {
        bool done = false;
#pragma unroll
        for (int i = 0; i < kLaneAcc; i++) if (!done && i < n && lab[i] == L) { cnt[i] += c; done = true; }
        if (!done) {
            if (n < kLaneAcc) {
#pragma unroll
                for (int i = 0; i < kLaneAcc; i++) if (i == n) { lab[i] = L; cnt[i] = c; }
                n++;
            } else overflow = true;
        }
    }
};

}  // namespace

template <bool ANCHOR, bool STATS>
__global__ void __launch_bounds__(256) k_query_lane(const TableView t, const QueryArgs a, const AnchorView av,
                                                    uint32_t* redo_list, uint32_t* redo_count)
{
    const int k = t.hp.k;
    const uint64_t kmsk = t.hp.mask;
    const int s2 = 2 * (k - 1);
    unsigned long long st_lookups = 0, st_sectors = 0;
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < a.n_reads; r += stride) {
        const uint32_t beg = __ldg(a.offsets + r);
        const int nslots = (int)(__ldg(a.offsets + r + 1) - beg);
        const uint16_t* c = a.containers + beg;
        const int len = nslots > 0 ? (int)__ldg(c) : 0;
        if (nslots == 0) {                                     // no k-mer at all (read shorter than k, or all N)
            uint16_t* o5 = a.out5 + (size_t)r * 5;
            o5[0] = o5[1] = o5[2] = o5[3] = o5[4] = 0;
            if (a.rows) a.rows[(size_t)r * a.row_stride] = 0;
            continue;
        }
        if (1 + ((len + 7) >> 3) != nslots || len > kLaneMaxBases || len < k) {       // not for this kernel
            redo_list[atomicAdd(redo_count, 1u)] = r;
            continue;
        }
        const int npos = len - k + 1;
        LaneVotes votes; votes.init();
        Bits192 W; W.w[0] = W.w[1] = W.w[2] = 0;              // windows answered by the anchor index

        if (ANCHOR) {
            // ---- the first two sampled k-mers of the read are tried as anchors ----
            uint64_t x = 0, rc = 0;
            uint32_t cw = 0;
            int tries = 0;
            uint64_t ae = ~0ull; int apos = 0; uint32_t afwd = 0;
            for (int i = 0; i < len && tries < 2 && ae == ~0ull; i++) {
                if ((i & 7) == 0) cw = __ldg(c + 1 + (i >> 3));
                const uint64_t code = (cw >> (14 - 2 * (i & 7))) & 3u;
                x = ((x << 2) | code) & kmsk;
                rc = (rc >> 2) | ((3ull - code) << s2);
                if (i < k - 1) continue;
                const uint64_t h = mix2k(x < rc ? x : rc, k, kmsk);
                if ((h >> (2 * k - 3)) != 0) continue;
                tries++;
                const uint64_t bi = mulhi64(((h << 3) & kmsk) << t.hp.shift, av.nb);
                const uint32_t fp = (uint32_t)(h >> 3) & 0x7FFFu;
                uint64_t e[4];
                load_bucket(av.buckets + bi, e);
                if (STATS) st_sectors++;
#pragma unroll
                for (int sl = 3; sl >= 0; sl--) if ((uint32_t)(e[sl] >> 49) == fp && e[sl] != ~0ull) ae = e[sl];
                apos = i - (k - 1); afwd = x <= rc ? 1u : 0u;
            }
            if (ae != ~0ull) {
                const uint32_t alab = (uint32_t)(ae >> 32) & 0xFFFFu;
                const uint64_t tpos = ae & 0xFFFFFFFFull;
                const bool same = afwd == ((uint32_t)(ae >> 48) & 1u);
                const int64_t g0 = same ? (int64_t)tpos - apos : (int64_t)(2 * av.nF - tpos - (uint64_t)k) - apos;
                if (g0 >= 0 && (uint64_t)g0 + kLaneMaxBases + 64 <= 2 * av.nF) {
                    // ---- read vs text, 16 bases at a time; validity and "stored" bits of the span ----
                    const uint64_t w0 = (uint64_t)g0 >> 4; const int sh = 2 * (int)(g0 & 15);
                    const uint64_t v0 = (uint64_t)g0 >> 5; const int vs = (int)(g0 & 31);
                    uint32_t tprev = __ldg(av.seq + w0);
                    uint32_t vprev = __ldg(av.valid + v0), dprev = __ldg(av.disc + v0);
                    Bits192 good, D;
                    bool span_ok = true;
#pragma unroll
                    for (int u = 0; u < 6; u++) {                  // 32 positions per step
                        uint32_t mis = 0;
#pragma unroll
                        for (int hlf = 0; hlf < 2; hlf++) {
                            const int w = 2 * u + hlf;             // 16-base unit
                            const uint32_t tnext = __ldg(av.seq + w0 + w + 1);
                            const uint32_t tw = __funnelshift_l(tnext, tprev, sh);
                            tprev = tnext;
                            const int s0 = 1 + 2 * w;
                            const uint32_t r0 = s0 < nslots ? (uint32_t)__ldg(c + s0) : 0u, r1 = s0 + 1 < nslots ? (uint32_t)__ldg(c + s0 + 1) : 0u;
                            mis |= mismatch16(tw ^ ((r0 << 16) | r1)) << (16 * hlf);
                        }
                        const uint32_t vnext = __ldg(av.valid + v0 + u + 1), dnext = __ldg(av.disc + v0 + u + 1);
                        const uint32_t gv = __funnelshift_r(vprev, vnext, vs), gd = __funnelshift_r(dprev, dnext, vs);
                        vprev = vnext; dprev = dnext;
                        const int here = len - 32 * u;             // read bases in this word
                        const uint32_t need = here >= 32 ? 0xFFFFFFFFu : (here > 0 ? (1u << here) - 1u : 0u);
                        if ((gv & need) != need) span_ok = false;  // a separator / N in the span: other records may follow
                        const uint32_t g32 = ~mis & gv & need;
                        if (u & 1) { good.w[u >> 1] |= (uint64_t)g32 << 32; D.w[u >> 1] |= (uint64_t)gd << 32; }
                        else       { good.w[u >> 1] = g32; D.w[u >> 1] = gd; }
                    }
                    if (STATS) st_sectors += 3;
                    if (span_ok) {
                        // windows of k good bases: runs of ones by doubling
                        Bits192 run = good; int have = 1;
                        while (2 * have <= k) { run.and_with(run.shr(have)); have *= 2; }
                        if (have < k) run.and_with(run.shr(k - have));
                        // only window starts < npos count (bits beyond are already 0: good is masked to len)
                        W = run;
                        const uint32_t hits = __popcll(W.w[0] & D.w[0]) + __popcll(W.w[1] & D.w[1]) + __popcll(W.w[2] & D.w[2]);
                        if (hits) votes.add(alab, hits);
                        if (STATS) st_lookups += __popcll(W.w[0]) + __popcll(W.w[1]) + __popcll(W.w[2]);
                    }
                }
            }
        }

        // ---- walk the read: every window not answered above is probed, 4 in flight ----
        {
            uint64_t x = 0, rc = 0;
            uint32_t cw = 0;
            for (int i = 0; i < k - 1; i++) {
                if ((i & 7) == 0) cw = __ldg(c + 1 + (i >> 3));
                const uint64_t code = (cw >> (14 - 2 * (i & 7))) & 3u;
                x = (x << 2) | code;
            }
            x &= kmsk;
            rc = revcomp(x, k);                                  // of the k-1 bases, as k_query does
            for (int p0 = 0; p0 < npos && !votes.overflow; p0 += 4) {
                uint64_t e[4][4], home[4], q[4]; bool need[4];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int p = p0 + j, i = p + k - 1;           // window start, its last base
                    need[j] = false;
                    if (p < npos) {
                        if ((i & 7) == 0) cw = __ldg(c + 1 + (i >> 3));
                        const uint64_t code = (cw >> (14 - 2 * (i & 7))) & 3u;
                        x = ((x << 2) | code) & kmsk;
                        rc = (rc >> 2) | ((3ull - code) << s2);
                        need[j] = !(ANCHOR && W.get(p));
                        if (need[j]) {
                            const uint64_t h = mix2k(x < rc ? x : rc, k, kmsk);
                            home[j] = home_of(h, t.hp); q[j] = key48_of(h) << 16;
                            load_bucket(t.buckets + home[j], e[j]);
                        }
                    }
                }
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    if (need[j]) {
                        uint32_t probes = 1;
                        const uint32_t l = resolve(t, home[j], q[j], e[j], STATS ? &probes : nullptr);
                        if (STATS) { st_lookups++; st_sectors += probes; }
                        if (l < kLabelTomb) votes.add(l, 1u);
                    }
                }
            }
        }
        if (votes.overflow) {                                   // a fifth target: the warp-per-read kernel redoes it
            redo_list[atomicAdd(redo_count, 1u)] = r;
            continue;
        }
        // ---- result {sumN, idxBest, best, idxSecond, second}: top-2 under (count desc, target asc) ----
        uint32_t sum = 0, k1 = 0, k2 = 0;
#pragma unroll
        for (int i = 0; i < kLaneAcc; i++) {
            const uint32_t c16 = i < votes.n ? (votes.cnt[i] & 0xFFFFu) : 0u;     // reference counters are u16
            if (!c16) continue;
            sum += c16;
            const uint32_t key = (c16 << 16) | (0xFFFFu - votes.lab[i]);
            if (key > k1) { k2 = k1; k1 = key; } else if (key > k2) k2 = key;
        }
        uint16_t* o5 = a.out5 + (size_t)r * 5;
        o5[0] = (uint16_t)sum;
        o5[1] = (uint16_t)(k1 ? (0xFFFFu - (k1 & 0xFFFFu)) + 1u : 0u);
        o5[2] = (uint16_t)(k1 >> 16);
        o5[3] = (uint16_t)(k2 ? (0xFFFFu - (k2 & 0xFFFFu)) + 1u : 0u);
        o5[4] = (uint16_t)(k2 >> 16);
        if (a.rows) {                                           // sparse row {n,(target,count)*}, targets ascending
            uint16_t* row = a.rows + (size_t)r * a.row_stride;
            const int fit = (a.row_stride - 1) / 2;
            row[0] = (uint16_t)min(votes.n, fit);
#pragma unroll
            for (int i = 0; i < kLaneAcc; i++) {
                if (i < votes.n) {
                    int rank = 0;
#pragma unroll
                    for (int j = 0; j < kLaneAcc; j++) if (j < votes.n && votes.lab[j] < votes.lab[i]) rank++;
                    if (2 * rank + 2 < a.row_stride) { row[1 + 2 * rank] = (uint16_t)votes.lab[i]; row[2 + 2 * rank] = (uint16_t)votes.cnt[i]; }
                }
            }
        }
    }
    if (STATS) {
#pragma unroll
        for (int o = 16; o; o >>= 1) {
            st_lookups += __shfl_xor_sync(0xFFFFFFFFu, st_lookups, o);
            st_sectors += __shfl_xor_sync(0xFFFFFFFFu, st_sectors, o);
        }
        if ((threadIdx.x & 31) == 0) { atomicAdd(a.stats + 0, st_lookups); atomicAdd(a.stats + 1, st_sectors); }
    }
}

// ---------------------------------------------------------------------------
// Block-cooperative variant: the same per-read work split into three uniform phases per block of 256
// reads, so that lanes never wait for a neighbour whose read needs more probes.
//   A  thread = read   : shape check, anchor lookup, read-vs-text comparison, bulk vote; the windows still
//                        to be probed are cut into 16-aligned chunks and queued in shared memory
//   B  thread = chunk  : k-mer of the chunk's first window extracted from the containers, then rolling;
//                        4 probes in flight; votes into per-read accumulators in shared memory
//   C  thread = read   : result from the accumulators (or back to k_query when a fifth target showed up)
// ---------------------------------------------------------------------------
constexpr int kBlkReads = 256;
constexpr int kBlkQueue = kBlkReads * 12;        // <= 162 windows per read -> <= 11 chunks of 16

struct BlockSmem {
    uint32_t beg[kBlkReads];
    uint32_t acc_lab[kBlkReads][kLaneAcc];
    uint32_t acc_cnt[kBlkReads][kLaneAcc];
    uint8_t  state[kBlkReads];                   // 0 nothing to do, 1 handled here, 2 overflow
    uint16_t len[kBlkReads];
    uint32_t queue[kBlkQueue];                   // read << 20 | chunk << 16 | need mask
    uint32_t qn;
};

__device__ __forceinline__ void vote_smem(BlockSmem& sm, uint32_t lr, uint32_t label, uint32_t count)
{
#pragma unroll
    for (int s = 0; s < kLaneAcc; s++) {
        const uint32_t old = atomicCAS(&sm.acc_lab[lr][s], kLabelNone, label);
        if (old == kLabelNone || old == label) { atomicAdd(&sm.acc_cnt[lr][s], count); return; }
    }
    sm.state[lr] = 2;
}

template <bool ANCHOR, bool STATS>
__global__ void __launch_bounds__(kBlkReads, 2) k_query_block(const TableView t, const QueryArgs a, const AnchorView av,
                                                               uint32_t* redo_list, uint32_t* redo_count)
{
    __shared__ BlockSmem sm;
    const int k = t.hp.k;
    const uint64_t kmsk = t.hp.mask;
    const int s2 = 2 * (k - 1);
    const int tid = threadIdx.x;
    unsigned long long st_lookups = 0, st_sectors = 0;
    const uint32_t nblocks = (a.n_reads + kBlkReads - 1) / kBlkReads;
    for (uint32_t blk = blockIdx.x; blk < nblocks; blk += gridDim.x) {
        const uint32_t r = blk * kBlkReads + tid;
        // ---------------- phase A ----------------
        if (tid == 0) sm.qn = 0;
#pragma unroll
        for (int s = 0; s < kLaneAcc; s++) { sm.acc_lab[tid][s] = kLabelNone; sm.acc_cnt[tid][s] = 0; }
        sm.state[tid] = 0;
        __syncthreads();
        if (r < a.n_reads) {
            const uint32_t beg = __ldg(a.offsets + r);
            const int nslots = (int)(__ldg(a.offsets + r + 1) - beg);
            const uint16_t* c = a.containers + beg;
            const int len = nslots > 0 ? (int)__ldg(c) : 0;
            if (nslots == 0) {
                uint16_t* o5 = a.out5 + (size_t)r * 5;
                o5[0] = o5[1] = o5[2] = o5[3] = o5[4] = 0;
                if (a.rows) a.rows[(size_t)r * a.row_stride] = 0;
            } else if (1 + ((len + 7) >> 3) != nslots || len > kLaneMaxBases || len < k) {
                redo_list[atomicAdd(redo_count, 1u)] = r;
            } else {
                sm.state[tid] = 1; sm.beg[tid] = beg; sm.len[tid] = (uint16_t)len;
                const int npos = len - k + 1;
                Bits192 W; W.w[0] = W.w[1] = W.w[2] = 0;
                if (ANCHOR) {
                    uint64_t x = 0, rc = 0;
                    uint32_t cw = 0;
                    int tries = 0;
                    uint64_t ae = ~0ull; int apos = 0; uint32_t afwd = 0;
                    for (int i = 0; i < len && tries < 2 && ae == ~0ull; i++) {
                        if ((i & 7) == 0) cw = __ldg(c + 1 + (i >> 3));
                        const uint64_t code = (cw >> (14 - 2 * (i & 7))) & 3u;
                        x = ((x << 2) | code) & kmsk;
                        rc = (rc >> 2) | ((3ull - code) << s2);
                        if (i < k - 1) continue;
                        const uint64_t h = mix2k(x < rc ? x : rc, k, kmsk);
                        if ((h >> (2 * k - 3)) != 0) continue;
                        tries++;
                        const uint64_t bi = mulhi64(((h << 3) & kmsk) << t.hp.shift, av.nb);
                        const uint32_t fp = (uint32_t)(h >> 3) & 0x7FFFu;
                        uint64_t e[4];
                        load_bucket(av.buckets + bi, e);
                        if (STATS) st_sectors++;
#pragma unroll
                        for (int sl = 3; sl >= 0; sl--) if ((uint32_t)(e[sl] >> 49) == fp && e[sl] != ~0ull) ae = e[sl];
                        apos = i - (k - 1); afwd = x <= rc ? 1u : 0u;
                    }
                    if (ae != ~0ull) {
                        const uint32_t alab = (uint32_t)(ae >> 32) & 0xFFFFu;
                        const uint64_t tpos = ae & 0xFFFFFFFFull;
                        const bool same = afwd == ((uint32_t)(ae >> 48) & 1u);
                        const int64_t g0 = same ? (int64_t)tpos - apos : (int64_t)(2 * av.nF - tpos - (uint64_t)k) - apos;
                        if (g0 >= 0 && (uint64_t)g0 + kLaneMaxBases + 64 <= 2 * av.nF) {
                            const uint64_t w0 = (uint64_t)g0 >> 4; const int sh = 2 * (int)(g0 & 15);
                            const uint64_t v0 = (uint64_t)g0 >> 5; const int vs = (int)(g0 & 31);
                            uint32_t tprev = __ldg(av.seq + w0);
                            uint32_t vprev = __ldg(av.valid + v0), dprev = __ldg(av.disc + v0);
                            Bits192 good, D;
                            bool span_ok = true;
#pragma unroll
                            for (int u = 0; u < 6; u++) {
                                uint32_t mis = 0;
#pragma unroll
                                for (int hlf = 0; hlf < 2; hlf++) {
                                    const int w = 2 * u + hlf;
                                    const uint32_t tnext = __ldg(av.seq + w0 + w + 1);
                                    const uint32_t tw = __funnelshift_l(tnext, tprev, sh);
                                    tprev = tnext;
                                    const int s0 = 1 + 2 * w;
                                    const uint32_t r0 = s0 < nslots ? (uint32_t)__ldg(c + s0) : 0u, r1 = s0 + 1 < nslots ? (uint32_t)__ldg(c + s0 + 1) : 0u;
                                    mis |= mismatch16(tw ^ ((r0 << 16) | r1)) << (16 * hlf);
                                }
                                const uint32_t vnext = __ldg(av.valid + v0 + u + 1), dnext = __ldg(av.disc + v0 + u + 1);
                                const uint32_t gv = __funnelshift_r(vprev, vnext, vs), gd = __funnelshift_r(dprev, dnext, vs);
                                vprev = vnext; dprev = dnext;
                                const int here = len - 32 * u;
                                const uint32_t need = here >= 32 ? 0xFFFFFFFFu : (here > 0 ? (1u << here) - 1u : 0u);
                                if ((gv & need) != need) span_ok = false;
                                const uint32_t g32 = ~mis & gv & need;
                                if (u & 1) { good.w[u >> 1] |= (uint64_t)g32 << 32; D.w[u >> 1] |= (uint64_t)gd << 32; }
                                else       { good.w[u >> 1] = g32; D.w[u >> 1] = gd; }
                            }
                            if (STATS) st_sectors += 3;
                            if (span_ok) {
                                Bits192 run = good; int have = 1;
                                while (2 * have <= k) { run.and_with(run.shr(have)); have *= 2; }
                                if (have < k) run.and_with(run.shr(k - have));
                                W = run;
                                const uint32_t hits = __popcll(W.w[0] & D.w[0]) + __popcll(W.w[1] & D.w[1]) + __popcll(W.w[2] & D.w[2]);
                                if (hits) { sm.acc_lab[tid][0] = alab; sm.acc_cnt[tid][0] = hits; }
                                if (STATS) st_lookups += __popcll(W.w[0]) + __popcll(W.w[1]) + __popcll(W.w[2]);
                            }
                        }
                    }
                }
                // ---- the windows still to probe, as 16-aligned chunks ----
                uint32_t items[12]; int ni = 0;
#pragma unroll
                for (int ch = 0; ch < 12; ch++) {
                    const int p = 16 * ch;
                    if (p < npos) {
                        const uint64_t wv = ch < 4 ? W.w[0] : (ch < 8 ? W.w[1] : W.w[2]);
                        const int left = npos - p;
                        const uint32_t in = left >= 16 ? 0xFFFFu : ((1u << left) - 1u);
                        const uint32_t needm = ~(uint32_t)(wv >> (16 * (ch & 3))) & in;
                        if (needm) items[ni++] = ((uint32_t)tid << 20) | ((uint32_t)ch << 16) | needm;
                    }
                }
                if (ni) {
                    const uint32_t at = atomicAdd(&sm.qn, (uint32_t)ni);
#pragma unroll
                    for (int j = 0; j < 12; j++) if (j < ni) sm.queue[at + j] = items[j];
                }
            }
        }
        __syncthreads();
        // ---------------- phase B ----------------
        const uint32_t qn = sm.qn;
        for (uint32_t it = tid; it < qn; it += kBlkReads) {
            const uint32_t item = sm.queue[it];
            const uint32_t lr = item >> 20, ch = (item >> 16) & 15u, needm = item & 0xFFFFu;
            const uint16_t* c = a.containers + sm.beg[lr];
            const int first = 16 * (int)ch + (__ffs(needm) - 1), last = 16 * (int)ch + (31 - __clz(needm));
            // k-1 bases starting at `first`, taken from five containers
            uint64_t x, rc;
            {
                const int s0 = 1 + (first >> 3);
                const int nsl = 1 + ((sm.len[lr] + 7) >> 3);
                uint64_t v = 0;
#pragma unroll
                for (int j = 0; j < 4; j++) v = (v << 16) | (s0 + j < nsl ? (uint64_t)__ldg(c + s0 + j) : 0ull);
                const uint64_t c4 = s0 + 4 < nsl ? (uint64_t)__ldg(c + s0 + 4) : 0ull;
                const int sh = 2 * (first & 7);
                const uint64_t full = sh ? ((v << sh) | (c4 >> (16 - sh))) : v;       // 32 bases from `first`
                x = k > 1 ? full >> (64 - s2) : 0ull;
                rc = revcomp(x, k);
            }
            uint32_t run_lab = kLabelNone, run_cnt = 0;
            uint32_t cw = 0; bool have_cw = false;
            for (int p0 = first; p0 <= last; p0 += 4) {
                uint64_t e[4][4], home[4], q[4]; bool need[4];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int p = p0 + j, i = p + k - 1;
                    need[j] = false;
                    if (p <= last) {
                        if (!have_cw || (i & 7) == 0) { cw = __ldg(c + 1 + (i >> 3)); have_cw = true; }
                        const uint64_t code = (cw >> (14 - 2 * (i & 7))) & 3u;
                        x = ((x << 2) | code) & kmsk;
                        rc = (rc >> 2) | ((3ull - code) << s2);
                        need[j] = (needm >> (p - 16 * (int)ch)) & 1u;
                        if (need[j]) {
                            const uint64_t h = mix2k(x < rc ? x : rc, k, kmsk);
                            home[j] = home_of(h, t.hp); q[j] = key48_of(h) << 16;
                            load_bucket(t.buckets + home[j], e[j]);
                        }
                    }
                }
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    if (need[j]) {
                        uint32_t probes = 1;
                        const uint32_t l = resolve(t, home[j], q[j], e[j], STATS ? &probes : nullptr);
                        if (STATS) { st_lookups++; st_sectors += probes; }
                        if (l < kLabelTomb) {
                            if (l == run_lab) run_cnt++;
                            else { if (run_cnt) vote_smem(sm, lr, run_lab, run_cnt); run_lab = l; run_cnt = 1; }
                        }
                    }
                }
            }
            if (run_cnt) vote_smem(sm, lr, run_lab, run_cnt);
        }
        __syncthreads();
        // ---------------- phase C ----------------
        if (r < a.n_reads && sm.state[tid]) {
            if (sm.state[tid] == 2) {
                redo_list[atomicAdd(redo_count, 1u)] = r;
            } else {
                uint32_t sum = 0, k1 = 0, k2 = 0; int n = 0;
#pragma unroll
                for (int i = 0; i < kLaneAcc; i++) {
                    const uint32_t lab = sm.acc_lab[tid][i];
                    if (lab == kLabelNone) continue;
                    n++;
                    const uint32_t c16 = sm.acc_cnt[tid][i] & 0xFFFFu;
                    if (!c16) continue;
                    sum += c16;
                    const uint32_t key = (c16 << 16) | (0xFFFFu - lab);
                    if (key > k1) { k2 = k1; k1 = key; } else if (key > k2) k2 = key;
                }
                uint16_t* o5 = a.out5 + (size_t)r * 5;
                o5[0] = (uint16_t)sum;
                o5[1] = (uint16_t)(k1 ? (0xFFFFu - (k1 & 0xFFFFu)) + 1u : 0u);
                o5[2] = (uint16_t)(k1 >> 16);
                o5[3] = (uint16_t)(k2 ? (0xFFFFu - (k2 & 0xFFFFu)) + 1u : 0u);
                o5[4] = (uint16_t)(k2 >> 16);
                if (a.rows) {
                    uint16_t* row = a.rows + (size_t)r * a.row_stride;
                    const int fit = (a.row_stride - 1) / 2;
                    row[0] = (uint16_t)min(n, fit);
#pragma unroll
                    for (int i = 0; i < kLaneAcc; i++) {
                        const uint32_t lab = sm.acc_lab[tid][i];
                        if (lab == kLabelNone) continue;
                        int rank = 0;
#pragma unroll
                        for (int j = 0; j < kLaneAcc; j++) { const uint32_t o = sm.acc_lab[tid][j]; if (o != kLabelNone && o < lab) rank++; }
                        if (2 * rank + 2 < a.row_stride) { row[1 + 2 * rank] = (uint16_t)lab; row[2 + 2 * rank] = (uint16_t)sm.acc_cnt[tid][i]; }
                    }
                }
            }
        }
        __syncthreads();
    }
    if (STATS) {
#pragma unroll
        for (int o = 16; o; o >>= 1) {
            st_lookups += __shfl_xor_sync(0xFFFFFFFFu, st_lookups, o);
            st_sectors += __shfl_xor_sync(0xFFFFFFFFu, st_sectors, o);
        }
        if ((threadIdx.x & 31) == 0) { atomicAdd(a.stats + 0, st_lookups); atomicAdd(a.stats + 1, st_sectors); }
    }
}

int launch_query_lane(const TableView& t, const QueryArgs& a, const AnchorView* av, uint32_t* redo_list,
                      uint32_t* redo_count, cudaStream_t s)
{
    if (a.n_reads == 0) return 0;
    uint32_t ctas = (a.n_reads + 255) / 256;
    if (ctas > 148u * 64u) ctas = 148u * 64u;
    const AnchorView none{};
    static const bool simple = getenv("CCB_LANE_SIMPLE") != nullptr;      // the one-phase kernel, kept for comparison
    if (!simple) {
        if (av) {
            if (a.stats) k_query_block<true, true><<<ctas, kBlkReads, 0, s>>>(t, a, *av, redo_list, redo_count);
            else         k_query_block<true, false><<<ctas, kBlkReads, 0, s>>>(t, a, *av, redo_list, redo_count);
        } else {
            if (a.stats) k_query_block<false, true><<<ctas, kBlkReads, 0, s>>>(t, a, none, redo_list, redo_count);
            else         k_query_block<false, false><<<ctas, kBlkReads, 0, s>>>(t, a, none, redo_list, redo_count);
        }
        return (int)cudaGetLastError();
    }
    if (av) {
        if (a.stats) k_query_lane<true, true><<<ctas, 256, 0, s>>>(t, a, *av, redo_list, redo_count);
        else         k_query_lane<true, false><<<ctas, 256, 0, s>>>(t, a, *av, redo_list, redo_count);
    } else {
        if (a.stats) k_query_lane<false, true><<<ctas, 256, 0, s>>>(t, a, none, redo_list, redo_count);
        else         k_query_lane<false, false><<<ctas, 256, 0, s>>>(t, a, none, redo_list, redo_count);
    }
    return (int)cudaGetLastError();
}

}  // namespace ccb
