// shard_block.cu -- sharded table (row E2): one-pass routing and block-wise voting for short reads.
//
// k_query<count> + scan + k_query<route> (query.cu) walk every read twice with eight ballots per k-mer and need a
// global scan in between, because every (read, owner) pair gets its own contiguous segment.  The vote does not need
// that: counts are order-independent.  Here a BLOCK of 256 reads owns one contiguous segment per owner:
//
//   k_route_block   thread per read for the headers, then the block's k-mer windows are dealt out four at a time
//                   (the item queue of k_query_block, query_lane.cu).  Pass 1 counts the block's k-mers per owner
//                   (packed counters in registers, one shared-memory add per warp and owner); ONE atomicAdd per owner
//                   on the requester's cursor reserves the block's segment; pass 2 recomputes the k-mers (cheaper than
//                   keeping 30 k of them) and stores each at  segment base + rank, the rank coming from
//                   __match_any_sync (lanes bound for the same owner take consecutive slots: stores of one warp
//                   instruction to one owner are contiguous) + one shared-memory add per warp and owner.  Beside every
//                   k-mer goes ONE local byte: which read of the block it belongs to.
//   owner side      k_probe_inbox (query.cu), unchanged: labels come back at the same indices.
//   k_vote_block    the block's segments are read back coalesced (label u16 + read byte), votes go into the read's
//                   four (label,count) pairs in shared memory, results as in k_query_block.
// Reads the block kernel does not take (three or more parts, parts over 192 bases) go through the generic pair of
// k_query<count/route> passes on a list; a read that meets a fifth target is re-voted by k_vote_block_dense.
#include "query.cuh"
#include "vote.cuh"
#include <stdlib.h>

namespace ccb {

namespace {

constexpr int kMaxBases = 192;             // as k_query_block
constexpr int kAcc = 4;
constexpr int kItems = 48;                 // <= 191 windows -> 12 chunks of 16 -> <= 4 items each

template <int BLK>
struct RouteSmem {
    uint32_t beg[BLK];
    uint16_t len[BLK];
    uint32_t qn;
    uint32_t cnt[kMaxShards];              // pass 1: k-mers of the block per owner
    uint32_t fill[kMaxShards];             // pass 2: slots handed out so far
    unsigned long long base[kMaxShards];   // first slot of the block's segment
    uint32_t queue[BLK * kItems];          // read << 20 | chunk << 16 | mask of <= 4 windows of the chunk
    // staged pass 2: one round's k-mers (<= 4 per thread) sorted by owner before they leave the SM
    uint32_t tcnt[2][kMaxShards];          // k-mers of the round per owner (double-buffered: no barrier to clear it)
    uint32_t sfill[2][kMaxShards];         // slots handed out before the round
    uint64_t stage_h[BLK * 4];
    uint8_t  stage_r[BLK * 4];
};

// the (up to four) mixed k-mers of an item: windows given by the set bits of the item's mask, lowest first
template <int BLK>
__device__ __forceinline__ void item_hashes(const RouteSmem<BLK>& sm, const uint16_t* __restrict__ containers, uint32_t item,
                                            int k, uint64_t kmsk, uint64_t (&h)[4], bool (&has)[4])
{
    const uint32_t lr = item >> 20, ch = (item >> 16) & 15u;
    uint32_t m4 = item & 0xFFFFu;
    const uint32_t cb = sm.beg[lr];
    uint64_t bhi = 0, blo = 0;
    const int s0 = 1 + 2 * (int)ch;
    const int nsl = 1 + ((sm.len[lr] + 7) >> 3);
#pragma unroll
    for (int j = 0; j < 4; j++) bhi = (bhi << 16) | (s0 + j < nsl ? (uint64_t)__ldg(containers + cb + s0 + j) : 0ull);
#pragma unroll
    for (int j = 4; j < 6; j++) blo = (blo << 16) | (s0 + j < nsl ? (uint64_t)__ldg(containers + cb + s0 + j) : 0ull);
    blo <<= 32;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        has[j] = m4 != 0;
        h[j] = 0;
        if (has[j]) {
            const int o = __ffs(m4) - 1;
            m4 &= m4 - 1;
            const uint64_t w64 = o ? (bhi << (2 * o)) | (blo >> (64 - 2 * o)) : bhi;
            const uint64_t x = w64 >> (64 - 2 * k);
            const uint64_t rc = revcomp_brev(x, k);
            h[j] = mix2k(x < rc ? x : rc, k, kmsk);
        }
    }
}

}  // namespace

// STAGED: pass 2 sorts every round's k-mers by owner in shared memory and stores them as contiguous runs, so that one
// warp instruction writes 256 bytes to one owner.  Unstaged, lanes bound for one owner take consecutive slots straight
// from the registers: 32 / G lanes x 8 bytes per warp instruction and owner -- 128 bytes at G = 2, but 32 bytes at G = 8,
// and the fabric between the GPUs then carries 32-byte writes (measured at 8 GPUs: 280 GB/s per GPU).
template <int BLK, bool STAGED>
__global__ void __launch_bounds__(BLK, 3) k_route_block(const HashParams hp, const QueryArgs a, const BlockRouteArgs ra)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    RouteSmem<BLK>& sm = *reinterpret_cast<RouteSmem<BLK>*>(smem_raw);
    const int k = hp.k;
    const uint64_t kmsk = hp.mask;
    const int tid = threadIdx.x, lane = tid & 31;
    const uint32_t lt = (1u << lane) - 1u;
    const uint32_t G = ra.n_shards;
    const uint32_t nblocks = (a.n_reads + BLK - 1) / BLK;
    if (tid == 0) sm.qn = 0;
    __syncthreads();
    for (uint32_t blk = blockIdx.x; blk < nblocks; blk += gridDim.x) {
        const uint32_t r = blk * BLK + tid;
        // ---- the read: one part, or two (a pair joined by N, an N inside); anything else goes to the generic path ----
        uint32_t rbeg = 0; int plen0 = 0, plen1 = 0, pn0 = 0; bool mine = false;
        if (r < a.n_reads) {
            rbeg = __ldg(a.offsets + r);
            const int rslots = (int)(__ldg(a.offsets + r + 1) - rbeg);
            uint8_t st = 0;
            if (rslots > 0) {
                plen0 = (int)__ldg(a.containers + rbeg); pn0 = 1 + ((plen0 + 7) >> 3);
                mine = plen0 >= k && plen0 <= kMaxBases && pn0 <= rslots;
                if (mine && pn0 < rslots) {
                    plen1 = (int)__ldg(a.containers + rbeg + (uint32_t)pn0);
                    mine = plen1 >= k && plen1 <= kMaxBases && pn0 + 1 + ((plen1 + 7) >> 3) == rslots;
                }
                st = mine ? 1 : 2;
                if (!mine) ra.redo_list[atomicAdd(ra.redo_count, 1u)] = r;
            }
            ra.rstate[r] = st;
        }
        for (int part = 0; ; part++) {
            // ---- queue the part's windows: items of <= 4 windows of one 16-window chunk ----
            const int len = part ? plen1 : plen0;
            if (mine && len > 0) {
                sm.beg[tid] = rbeg + (part ? (uint32_t)pn0 : 0u); sm.len[tid] = (uint16_t)len;
                const int npos = len - k + 1;
                const uint32_t nitems = (uint32_t)(npos + 3) >> 2;          // chunks are multiples of four windows
                uint32_t at = atomicAdd(&sm.qn, nitems);
                for (int p = 0; p < npos; p += 4) {
                    const int left = npos - p;
                    const uint32_t m = (left >= 4 ? 0xFu : ((1u << left) - 1u)) << (p & 15);
                    sm.queue[at++] = ((uint32_t)tid << 20) | ((uint32_t)(p >> 4) << 16) | m;
                }
            }
            if (tid < kMaxShards) sm.cnt[tid] = 0;
            __syncthreads();
            const uint32_t qn = sm.qn;
            // ---- pass 1: how many k-mers does the block send to each owner ----
            {
                uint64_t c0 = 0, c1 = 0;                                   // 4 x 16-bit counters each: owners 0-3, 4-7
                for (uint32_t it = tid; it < qn; it += BLK) {
                    uint64_t h[4]; bool has[4];
                    item_hashes(sm, a.containers, sm.queue[it], k, kmsk, h, has);
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        if (has[j]) {
                            const uint32_t g = owner_of(h[j], hp.shift, G);
                            const uint64_t one = 1ull << (16 * (g & 3u));
                            if (g & 4u) c1 += one; else c0 += one;
                        }
                    }
                }
#pragma unroll
                for (int o = 16; o; o >>= 1) { c0 += __shfl_xor_sync(0xFFFFFFFFu, c0, o); c1 += __shfl_xor_sync(0xFFFFFFFFu, c1, o); }
                if (lane < (int)G) {
                    const uint32_t v = (uint32_t)(((lane & 4) ? c1 : c0) >> (16 * (lane & 3))) & 0xFFFFu;
                    if (v) atomicAdd(&sm.cnt[lane], v);
                }
            }
            __syncthreads();
            // ---- one reservation per owner for the whole block ----
            if (tid < (int)G) {
                const uint32_t c = sm.cnt[tid];
                const unsigned long long b = c ? atomicAdd(ra.cursor + tid, (unsigned long long)c) : 0ull;
                sm.base[tid] = b; sm.fill[tid] = 0;
                sm.tcnt[0][tid] = 0; sm.tcnt[1][tid] = 0; sm.sfill[0][tid] = 0;
                uint32_t* seg = ra.blkseg + (((size_t)blk * 2 + part) * kMaxShards + tid) * 2;
                seg[0] = (uint32_t)(b < ra.cap ? b : ra.cap); seg[1] = c;
                if (b + c > ra.cap) atomicExch(ra.error, 1u);            // the excess is dropped below
            }
            __syncthreads();
            // ---- pass 2: the k-mers again, each to its slot; every lane runs every round (warp-wide matching) ----
            const uint32_t rounds = (qn + BLK - 1) / BLK;
            for (uint32_t rd = 0; rd < rounds; rd++) {
                const uint32_t it = rd * BLK + tid;
                uint64_t h[4]; bool has[4] = {false, false, false, false};
                uint32_t lr = 0;
                if (it < qn) {
                    const uint32_t item = sm.queue[it];
                    lr = item >> 20;
                    item_hashes(sm, a.containers, item, k, kmsk, h, has);
                }
                if (!STAGED) {
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        const uint32_t g = has[j] ? owner_of(h[j], hp.shift, G) : 0x100u + (uint32_t)lane;   // idle lanes match nobody
                        const uint32_t m = __match_any_sync(0xFFFFFFFFu, g);
                        const int leader = __ffs(m) - 1;
                        uint32_t start = 0;
                        if (has[j] && lane == leader) start = atomicAdd(&sm.fill[g], (uint32_t)__popc(m));
                        start = __shfl_sync(0xFFFFFFFFu, start, leader);
                        if (has[j]) {
                            const unsigned long long idx = sm.base[g] + start + (uint32_t)__popc(m & lt);
                            if (idx < ra.cap) { ra.dst[g][idx] = h[j]; ra.rid[g][idx] = (uint8_t)lr; }
                        }
                    }
                } else {
                    // A: rank inside the round, per owner
                    const int cur = (int)(rd & 1u);
                    uint32_t og[4], loc[4];
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        const uint32_t g = has[j] ? owner_of(h[j], hp.shift, G) : 0x100u + (uint32_t)lane;
                        const uint32_t m = __match_any_sync(0xFFFFFFFFu, g);
                        const int leader = __ffs(m) - 1;
                        uint32_t start = 0;
                        if (has[j] && lane == leader) start = atomicAdd(&sm.tcnt[cur][g], (uint32_t)__popc(m));
                        start = __shfl_sync(0xFFFFFFFFu, start, leader);
                        og[j] = g; loc[j] = start + (uint32_t)__popc(m & lt);
                    }
                    __syncthreads();
                    // B: the round's k-mers into shared memory, owner by owner
                    uint32_t tc[kMaxShards];
#pragma unroll
                    for (int gg = 0; gg < kMaxShards; gg++) tc[gg] = gg < (int)G ? sm.tcnt[cur][gg] : 0u;
                    if (tid < (int)G) {                                        // for the next round (read after its first barrier)
                        sm.sfill[cur ^ 1][tid] = sm.sfill[cur][tid] + sm.tcnt[cur][tid];
                        sm.tcnt[cur ^ 1][tid] = 0;
                    }
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        if (has[j]) {
                            uint32_t off = 0;
#pragma unroll
                            for (int gg = 0; gg < kMaxShards - 1; gg++) off += gg < (int)og[j] ? tc[gg] : 0u;
                            sm.stage_h[off + loc[j]] = h[j]; sm.stage_r[off + loc[j]] = (uint8_t)lr;
                        }
                    }
                    __syncthreads();
                    // C: out, as contiguous runs
                    uint32_t total = 0;
#pragma unroll
                    for (int gg = 0; gg < kMaxShards; gg++) total += tc[gg];
                    for (uint32_t i = (uint32_t)tid; i < total; i += BLK) {
                        uint32_t g = 0, o = 0, nxt = 0;
#pragma unroll
                        for (int gg = 0; gg < kMaxShards - 1; gg++) {
                            nxt += tc[gg];
                            if (i >= nxt) { g = (uint32_t)gg + 1u; o = nxt; }
                        }
                        const unsigned long long idx = sm.base[g] + sm.sfill[cur][g] + (i - o);
                        if (idx < ra.cap) { ra.dst[g][idx] = sm.stage_h[i]; ra.rid[g][idx] = sm.stage_r[i]; }
                    }
                }
            }
            // next part?  (also the barrier that lets the queue be rebuilt)
            const int more = __syncthreads_or(part == 0 && mine && plen1 > 0);
            if (tid == 0) sm.qn = 0;
            if (!more) {
                // a block without second parts: its part-1 segments are empty
                if (part == 0 && tid < (int)G) {
                    uint32_t* seg = ra.blkseg + (((size_t)blk * 2 + 1) * kMaxShards + tid) * 2;
                    seg[0] = 0; seg[1] = 0;
                }
                __syncthreads();
                break;
            }
            __syncthreads();
        }
    }
    __threadfence_system();                                             // the stores may have gone to a peer GPU
}

int launch_route_block(const HashParams& hp, const QueryArgs& a, const BlockRouteArgs& ra, cudaStream_t s)
{
    if (a.n_reads == 0) return 0;
    if (ra.n_shards < 1 || ra.n_shards > (uint32_t)kMaxShards) return -60;
    constexpr int BLK = kRouteBlk;
    const size_t smem = sizeof(RouteSmem<BLK>);
    // grid: many more blocks than are resident (three per SM) so that the block scheduler evens out the tail (measured:
    // 11.2 ms per 10 M reads against 13.0 ms with 3 per SM and 14.0 ms with 2; small persistent grids meant to let another
    // batch's probe run beside the route were tried and cost more than the overlap returned -- DESIGN.md section 9)
    uint32_t ctas = (a.n_reads + BLK - 1) / BLK;
    static const uint32_t per_sm = getenv("CCB_ROUTE_CTAS_PER_SM") ? (uint32_t)atoi(getenv("CCB_ROUTE_CTAS_PER_SM")) : 64u;
    if (ctas > 148u * per_sm) ctas = 148u * per_sm;
    // staged stores pay two barriers per round and win where the runs would otherwise be short (see k_route_block)
    const char* env = getenv("CCB_ROUTE_STAGED");
    const bool staged = env ? atoi(env) != 0 : ra.n_shards >= 3;
    auto kern = staged ? k_route_block<BLK, true> : k_route_block<BLK, false>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    kern<<<ctas, BLK, smem, s>>>(hp, a, ra);
    return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------
// vote: the labels of a block's segments -> four (label,count) pairs per read -> result
// ---------------------------------------------------------------------------
template <int BLK>
__global__ void __launch_bounds__(BLK) k_vote_block(const QueryArgs a, const BlockVoteArgs va)
{
    __shared__ uint32_t acc_lab[BLK][kAcc];
    __shared__ uint32_t acc_cnt[BLK][kAcc];
    __shared__ uint8_t  state[BLK];
    const int tid = threadIdx.x;
    const uint32_t nblocks = (a.n_reads + BLK - 1) / BLK;
    for (uint32_t blk = blockIdx.x; blk < nblocks; blk += gridDim.x) {
        const uint32_t r = blk * BLK + tid;
#pragma unroll
        for (int s = 0; s < kAcc; s++) { acc_lab[tid][s] = kLabelNone; acc_cnt[tid][s] = 0; }
        state[tid] = r < a.n_reads ? va.rstate[r] : (uint8_t)3;
        __syncthreads();
        for (int part = 0; part < 2; part++) {
            for (uint32_t g = 0; g < va.n_shards; g++) {
                const uint32_t* seg = va.blkseg + (((size_t)blk * 2 + part) * kMaxShards + g) * 2;
                const uint64_t b = seg[0];
                uint64_t c = seg[1];
                if (b + c > va.cap) c = va.cap > b ? va.cap - b : 0;        // dropped at routing time (error flag is up)
                const uint16_t* __restrict__ lab = va.src[g] + b;
                const uint8_t* __restrict__ rid = va.rid[g] + b;
                for (uint32_t i = tid; i < (uint32_t)c; i += BLK) {
                    const uint32_t l = lab[i];
                    if (l >= kLabelTomb) continue;
                    const uint32_t lr = rid[i];
                    bool done = false;
#pragma unroll
                    for (int s = 0; s < kAcc; s++) {
                        if (!done) {
                            const uint32_t old = atomicCAS(&acc_lab[lr][s], kLabelNone, l);
                            if (old == kLabelNone || old == l) { atomicAdd(&acc_cnt[lr][s], 1u); done = true; }
                        }
                    }
                    if (!done) state[lr] = 4;                               // a fifth target: dense re-vote
                }
            }
        }
        __syncthreads();
        if (r < a.n_reads) {
            const uint8_t st = state[tid];
            uint16_t* o5 = a.out5 + (size_t)r * 5;
            if (st == 0) {
                o5[0] = o5[1] = o5[2] = o5[3] = o5[4] = 0;
                if (a.rows) a.rows[(size_t)r * a.row_stride] = 0;
            } else if (st == 4) {
                a.ovf_list[atomicAdd(a.ovf_count, 1u)] = r;
            } else if (st == 1) {
                uint32_t sum = 0, k1 = 0, k2 = 0; int n = 0;
#pragma unroll
                for (int i = 0; i < kAcc; i++) {
                    const uint32_t lab = acc_lab[tid][i];
                    if (lab == kLabelNone) continue;
                    n++;
                    const uint32_t c16 = acc_cnt[tid][i] & 0xFFFFu;
                    if (!c16) continue;
                    sum += c16;
                    const uint32_t key = (c16 << 16) | (0xFFFFu - lab);
                    if (key > k1) { k2 = k1; k1 = key; } else if (key > k2) k2 = key;
                }
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
                    for (int i = 0; i < kAcc; i++) {
                        const uint32_t lab = acc_lab[tid][i];
                        if (lab == kLabelNone) continue;
                        int rank = 0;
#pragma unroll
                        for (int j = 0; j < kAcc; j++) { const uint32_t o = acc_lab[tid][j]; if (o != kLabelNone && o < lab) rank++; }
                        if (2 * rank + 2 < a.row_stride) { row[1 + 2 * rank] = (uint16_t)lab; row[2 + 2 * rank] = (uint16_t)acc_cnt[tid][i]; }
                    }
                }
            }                                                               // st == 2: the generic path writes this read
        }
        __syncthreads();
    }
}

int launch_vote_block(const QueryArgs& a, const BlockVoteArgs& va, cudaStream_t s)
{
    if (a.n_reads == 0) return 0;
    constexpr int BLK = kRouteBlk;
    uint32_t ctas = (a.n_reads + BLK - 1) / BLK;
    static const uint32_t per_sm = getenv("CCB_VOTE_CTAS_PER_SM") ? (uint32_t)atoi(getenv("CCB_VOTE_CTAS_PER_SM")) : 64u;
    if (ctas > 148u * per_sm) ctas = 148u * per_sm;
    k_vote_block<BLK><<<ctas, BLK, 0, s>>>(a, va);
    return (int)cudaGetLastError();
}

// exact re-vote of a read that met more than four targets: its labels are the entries of its block's segments that
// carry its read byte; dense counters in global scratch (one CTA per listed read)
template <int BLK>
__global__ void __launch_bounds__(256) k_vote_block_dense(const QueryArgs a, const BlockVoteArgs va, const uint32_t* read_list,
                                                          const uint32_t* d_n_list, uint32_t max_list, uint32_t n_targets, uint32_t* scratch)
{
    __shared__ uint32_t s_key[2][256];
    __shared__ uint32_t s_sum[256];
    const uint32_t n_list = min(*d_n_list, max_list);
    uint32_t* cnt = scratch + (size_t)blockIdx.x * n_targets;
    for (uint32_t li = blockIdx.x; li < n_list; li += gridDim.x) {
        const uint32_t r = read_list[li];
        const uint32_t blk = r / BLK, me = r % BLK;
        for (uint32_t i = threadIdx.x; i < n_targets; i += blockDim.x) cnt[i] = 0;
        __syncthreads();
        for (int part = 0; part < 2; part++) {
            for (uint32_t g = 0; g < va.n_shards; g++) {
                const uint32_t* seg = va.blkseg + (((size_t)blk * 2 + part) * kMaxShards + g) * 2;
                const uint64_t b = seg[0];
                uint64_t c = seg[1];
                if (b + c > va.cap) c = va.cap > b ? va.cap - b : 0;
                for (uint32_t i = threadIdx.x; i < (uint32_t)c; i += blockDim.x) {
                    if (va.rid[g][b + i] != me) continue;
                    const uint32_t l = va.src[g][b + i];
                    if (l < kLabelTomb && l < n_targets) atomicAdd(cnt + l, 1u);
                }
            }
        }
        __syncthreads();
        dense_result(a, r, cnt, n_targets, s_key, s_sum);
        __syncthreads();
    }
}

int launch_vote_block_dense(const QueryArgs& a, const BlockVoteArgs& va, const uint32_t* read_list, const uint32_t* d_n_list,
                            uint32_t max_list, uint32_t n_targets, uint32_t* scratch, uint32_t n_scratch_blocks, cudaStream_t s)
{
    if (max_list == 0 || n_scratch_blocks == 0) return 0;
    const uint32_t g = max_list < n_scratch_blocks ? max_list : n_scratch_blocks;
    k_vote_block_dense<kRouteBlk><<<g, 256, 0, s>>>(a, va, read_list, d_n_list, max_list, n_targets, scratch);
    return (int)cudaGetLastError();
}

}  // namespace ccb
