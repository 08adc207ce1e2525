// query_lane.cu -- thread-per-read classification of short reads, block-cooperative.
//
// k_query (query.cu) gives a whole warp to one read; most of its instructions coordinate the 32 lanes
// (tile staging, validity masks, shuffle-merged votes).  For the reads that dominate a metagenomic
// sample -- one or two parts (a pair joined by N) of <= 192 bases, a handful of targets hit -- none of that
// is needed.  k_query_block gives every thread of a block one read and runs three phases per block of reads
// (phases A and B once more for the second parts, when any read of the block has one):
//   A  per read: (with the anchor index, query.cuh AnchorView) roll over the read until two sampled k-mers
//      are met, look both up in the sampled table at once, compare the read with the target text base by
//      base; every window that matches is answered from the text's own bits.  The windows that remain are
//      cut into work items of at most four windows of one 16-window chunk and queued in shared memory.
//   B  the items are dealt out over the block: six containers give the 46 bases a chunk's windows can
//      touch, each window's k-mer is cut out of them directly (no rolling state, no divergence between
//      lanes), four probes are in flight per thread, votes go to the read's four (label,count) pairs in
//      shared memory.
//   C  per read: best / second best / sum, sparse row; queue and vote pairs are left clear for the next block.
// Anything else -- three or more parts, longer parts, a fifth distinct target -- is handed back on a list and
// classified by k_query.  Results are the same u16[5] / sparse rows either way.
// Measurements, the rejected variants (L2 prefetch of the next item, TMA staging of the containers = STAGE)
// and the block shapes are in DESIGN.md section 11.
#include "query.cuh"
#include <stdlib.h>
#include <string.h>

namespace ccb {

constexpr int kLaneMaxBases = 192;
constexpr int kLaneAcc = 4;
constexpr int kItemsPerRead = 48;          // <= 191 windows -> 12 chunks of 16 -> <= 4 items each

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
};

template <int BLK>
struct BlockSmem {
    uint32_t beg[BLK];                           // first container of the read
    uint32_t acc_lab[BLK][kLaneAcc];
    uint32_t acc_cnt[BLK][kLaneAcc];
    uint16_t len[BLK];
    uint8_t  state[BLK];                         // 0 nothing to do, 1 handled here, 2 overflow
    uint32_t qn;
    uint32_t queue[BLK * kItemsPerRead];         // read << 20 | chunk << 16 | mask of <= 4 windows of the chunk
};

template <int BLK>
__device__ __forceinline__ void vote_smem(BlockSmem<BLK>& sm, uint32_t lr, uint32_t label, uint32_t count)
{
#pragma unroll
    for (int s = 0; s < kLaneAcc; s++) {
        const uint32_t old = atomicCAS(&sm.acc_lab[lr][s], kLabelNone, label);
        if (old == kLabelNone || old == label) { atomicAdd(&sm.acc_cnt[lr][s], count); return; }
    }
    sm.state[lr] = 2;
}

// Staging of a block's containers (STAGE): the packed reads of a block are one contiguous run of the batch, so one
// elected thread asks the TMA engine for the whole run (cp.async.bulk, completion on an mbarrier) one block ahead;
// phases A and B then read containers from shared memory instead of issuing 2-byte global loads.
template <int BLK>
struct StageSmem {
    static constexpr int kCap = BLK * 44;        // u16 per buffer: two parts of 150 bases per read on average (else: global loads)
    alignas(16) uint16_t cbuf[2][kCap];
    unsigned long long bar[2];
    int      sbase[2];                           // container index of cbuf[b][0]
    uint32_t staged[2];
};
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// entry of sector e whose fingerprint is fp (~0 when there is none)
__device__ __forceinline__ uint64_t match_anchor(const uint64_t (&e)[4], uint32_t fp)
{
    uint64_t ae = ~0ull;
#pragma unroll
    for (int sl = 3; sl >= 0; sl--) if ((uint32_t)(e[sl] >> 49) == fp && e[sl] != ~0ull) ae = e[sl];
    return ae;
}

}  // namespace

// Phase A for one part of a read (with the anchor index): mark the sampled k-mers among the part's first 64 windows, look
// the first two up in the sampled table at once (then, if neither has an entry, the next two), compare the part with the
// target text base by base.  W = the windows answered without
// a table probe; vote_lab / vote_hits = the label those windows vote for and how many of them are stored k-mers.
// ldc(i) = container i of the batch.
template <bool STATS, typename LDC>
__device__ __forceinline__ void anchor_part(const TableView& t, const AnchorView& av, LDC&& ldc, uint32_t beg, int len, int nslots,
                                            int k, uint64_t kmsk, int s2, Bits192& W, uint32_t& vote_lab, uint32_t& vote_hits,
                                            unsigned long long& st_lookups, unsigned long long& st_sectors)
{
    // Pass 1, in lockstep: which of the part's first 64 windows hold a sampled k-mer?  One bit per window, no branch that
    // depends on the read -- every lane of the warp runs the same 64 iterations.  (Rolling "until two sampled k-mers were
    // met" made each warp wait for its unluckiest lane, twice: 80 iterations per warp for 21 per read, 8 of 32 lanes
    // active, two thirds of the kernel's instructions: profiles/r02_final_ncu_details.csv.)
    uint64_t smask = 0;
    {
        uint64_t x, rc;
        {
            uint64_t v = 0;
#pragma unroll
            for (int j = 0; j < 4; j++) v = (v << 16) | (1 + j < nslots ? (uint64_t)ldc(beg + 1 + j) : 0ull);
            x = k > 1 ? v >> (64 - s2) : 0ull;
            rc = revcomp_brev(x, k);
        }
        uint32_t cw = 0;
        int i = k - 1;
        if (i & 7) cw = ldc(beg + 1 + (i >> 3));
        const int stop = min(len, k - 1 + 64);
        for (; i < stop; i++) {
            if ((i & 7) == 0) cw = ldc(beg + 1 + (i >> 3));
            const uint64_t code = (cw >> (14 - 2 * (i & 7))) & 3u;
            x = ((x << 2) | code) & kmsk;
            rc = (rc >> 2) | ((3ull - code) << s2);
            smask |= (uint64_t)(anchor_sampled(x < rc ? x : rc) ? 1u : 0u) << (i - (k - 1));
        }
    }
    // canonical k-mer of window p, cut out of five containers
    auto window = [&](int p, uint64_t& cn, uint32_t& fwd) {
        const int c = 1 + (p >> 3), o = p & 7;
        uint64_t hi = 0;
#pragma unroll
        for (int j = 0; j < 4; j++) hi = (hi << 16) | (c + j < nslots ? (uint64_t)ldc(beg + c + j) : 0ull);
        const uint64_t lo = c + 4 < nslots ? (uint64_t)ldc(beg + c + 4) : 0ull;
        const uint64_t w = o ? (hi << (2 * o)) | (lo >> (16 - 2 * o)) : hi;
        const uint64_t x = w >> (64 - 2 * k);
        const uint64_t rc = revcomp_brev(x, k);
        cn = x < rc ? x : rc; fwd = x <= rc ? 1u : 0u;
    };
    uint64_t ae = ~0ull; int apos = 0; uint32_t afwd = 0;
    // up to two rounds; a round = the next two sampled k-mers of the part, looked up together
    for (int round = 0; round < 2 && ae == ~0ull && smask; round++) {
        uint64_t c0 = 0, c1 = 0; int p0 = 0, p1 = 0; uint32_t f0 = 0, f1 = 0; int nc = 1;
        p0 = __ffsll((long long)smask) - 1; smask &= smask - 1;
        window(p0, c0, f0);
        if (smask) { p1 = __ffsll((long long)smask) - 1; smask &= smask - 1; window(p1, c1, f1); nc = 2; }
        const uint64_t h0 = mix2k(c0, k, kmsk), h1 = mix2k(c1, k, kmsk);
        const uint64_t b0 = mulhi64(h0 << t.hp.shift, av.nb), b1 = mulhi64(h1 << t.hp.shift, av.nb);
        const uint32_t fp0 = (uint32_t)h0 & 0x7FFFu, fp1 = (uint32_t)h1 & 0x7FFFu;
        uint64_t e0[4], e1[4];
        load_bucket64(av.buckets + b0, e0);                      // a full sector overflows into b ^ 1: the same 64 bytes
        if (nc > 1) load_bucket64(av.buckets + b1, e1);
        if (STATS) st_sectors += nc;
        uint64_t m = match_anchor(e0, fp0);
        if (m == ~0ull && e0[3] != ~0ull) {                       // full sector: its neighbour may hold it
            load_bucket64(av.buckets + (b0 ^ 1ull), e0);
            if (STATS) st_sectors++;
            m = match_anchor(e0, fp0);
        }
        if (m != ~0ull) { ae = m; apos = p0; afwd = f0; }
        else if (nc > 1) {
            m = match_anchor(e1, fp1);
            if (m == ~0ull && e1[3] != ~0ull) {
                load_bucket64(av.buckets + (b1 ^ 1ull), e1);
                if (STATS) st_sectors++;
                m = match_anchor(e1, fp1);
            }
            if (m != ~0ull) { ae = m; apos = p1; afwd = f1; }
        }
    }
    if (ae != ~0ull) {
        const uint32_t alab = (uint32_t)(ae >> 32) & 0xFFFFu;
        const uint64_t tpos = ae & 0xFFFFFFFFull;
        const bool same = afwd == ((uint32_t)(ae >> 48) & 1u);
        // text position of read base 0 (F half, or the reverse-complement half)
        const int64_t g0 = same ? (int64_t)tpos - apos : (int64_t)(2 * av.nF - tpos - (uint64_t)k) - apos;
        if (g0 >= 0 && (uint64_t)g0 + kLaneMaxBases + 64 <= 2 * av.nF) {
            const int sft = (int)(g0 & 31);
            const int ngr = (sft + len + 31) >> 5;                                    // text groups the read spans (<= 7)
            const uint32_t* tg = av.txt + (size_t)kTextWords * ((uint64_t)g0 >> 5);
            uint32_t pw0, pw1, pv, pd, po;                                            // group u
            pw0 = __ldg(tg); pw1 = __ldg(tg + 1); pv = __ldg(tg + 2); pd = __ldg(tg + 3); po = __ldg(tg + 4);
            // the span covers two 128-byte lines (seldom three): ask for the last group now, so that the second line is
            // on its way while the first is compared instead of one more full latency in the middle of the loop
            const uint32_t last_w0 = ngr > 1 ? __ldg(tg + kTextWords * (ngr - 1)) : 0u;
            Bits192 good, D, O;
            good.w[0] = good.w[1] = good.w[2] = 0; D = good; O = good;
            bool span_ok = true;
#pragma unroll
            for (int u = 0; u < 6; u++) {
                uint32_t nw0 = 0, nw1 = 0, nv = 0, nd = 0, no = 0;                    // group u + 1
                if (u + 1 < ngr) {
                    const uint32_t* g = tg + kTextWords * (u + 1);
                    nw0 = (u + 1 == ngr - 1) ? last_w0 : __ldg(g);
                    nw1 = __ldg(g + 1); nv = __ldg(g + 2); nd = __ldg(g + 3); no = __ldg(g + 4);
                }
                const int here = len - 32 * u;                                        // read bases from 32u on
                if (here > 0) {
                    // text bases at read positions 32u .. 32u+31
                    const uint64_t tcur = ((uint64_t)pw0 << 32) | pw1, tnxt = ((uint64_t)nw0 << 32) | nw1;
                    const uint64_t tt = sft ? (tcur << (2 * sft)) | (tnxt >> (64 - 2 * sft)) : tcur;
                    uint64_t rd = 0;
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        const int sl = 1 + 4 * u + j;
                        rd = (rd << 16) | (sl < nslots ? (uint64_t)ldc(beg + sl) : 0ull);
                    }
                    const uint64_t df = tt ^ rd;
                    const uint32_t mis = mismatch16((uint32_t)(df >> 32)) | (mismatch16((uint32_t)df) << 16);
                    const uint32_t gv = __funnelshift_r(pv, nv, sft), gd = __funnelshift_r(pd, nd, sft),
                                   go = __funnelshift_r(po, no, sft);
                    const uint32_t need = here >= 32 ? 0xFFFFFFFFu : (1u << here) - 1u;
                    if ((gv & need) != need) span_ok = false;     // a separator inside the span: another record may follow
                    const uint32_t g32 = ~mis & gv & need;
                    if (u & 1) { good.w[u >> 1] |= (uint64_t)g32 << 32; D.w[u >> 1] |= (uint64_t)gd << 32; O.w[u >> 1] |= (uint64_t)go << 32; }
                    else       { good.w[u >> 1] = g32; D.w[u >> 1] = gd; O.w[u >> 1] = go; }
                }
                pw0 = nw0; pw1 = nw1; pv = nv; pd = nd; po = no;
            }
            if (STATS) st_sectors += (uint32_t)(ngr * kTextWords * 4 + 31) / 32u;
            if (span_ok) {
                Bits192 run = good; int have = 1;
                while (2 * have <= k) { run.and_with(run.shr(have)); have *= 2; }
                if (have < k) run.and_with(run.shr(k - have));
                run.w[0] &= ~O.w[0]; run.w[1] &= ~O.w[1]; run.w[2] &= ~O.w[2];       // stored under another label: probe
                W = run;
                const uint32_t hits = __popcll(W.w[0] & D.w[0]) + __popcll(W.w[1] & D.w[1]) + __popcll(W.w[2] & D.w[2]);
                if (hits) { vote_lab = alab; vote_hits = hits; }
                if (STATS) st_lookups += __popcll(W.w[0]) + __popcll(W.w[1]) + __popcll(W.w[2]);
            }
        }
    }
}

// what k_anchor_scan leaves for k_query_block<PRE>: per read and part, the windows already answered and their vote
struct alignas(64) AnchorPre {
    uint64_t w[2][3];
    uint32_t vote[2];          // label << 16 | stored windows among the answered ones (0: none)
    uint32_t pad[2];
};
static_assert(sizeof(AnchorPre) == kAnchorPreBytes, "AnchorPre layout");

// The anchor pass as a kernel of its own: a thread per read, no shared memory, no barriers -- a chain of dependent loads
// (offsets -> containers -> sampled table -> text) that only needs many reads in flight.  The probe kernel that follows
// (k_query_block<PRE>) then runs at the table-only kernel's occupancy instead of carrying this pass's 125 registers.
// K: the k-mer length as a compile-time constant (31 full mode, 27 light mode: every shift of the rolling loop becomes an
// immediate), 0 = taken from the table at run time.
template <bool STATS, int K>
__global__ void __launch_bounds__(128) k_anchor_scan(const TableView t, const QueryArgs a, const AnchorView av, AnchorPre* __restrict__ pre)
{
    const int k = K ? K : t.hp.k;
    const uint64_t kmsk = K ? (K < 32 ? (1ull << (2 * (K & 31))) - 1ull : ~0ull) : t.hp.mask;
    const int s2 = 2 * (k - 1);
    unsigned long long st_lookups = 0, st_sectors = 0;
    auto ldc = [&](uint32_t idx) -> uint32_t { return (uint32_t)__ldg(a.containers + idx); };
    // the offsets, the header container and the far end of the NEXT read of this thread are asked for one read ahead:
    // one dependent trip to memory less in every read's chain
    const uint32_t stride = gridDim.x * blockDim.x;
    uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t nx_beg = 0, nx_end = 0, nx_hdr = 0, keep = 0;
    if (r < a.n_reads) { nx_beg = __ldg(a.offsets + r); nx_end = __ldg(a.offsets + r + 1); if (nx_end > nx_beg) nx_hdr = ldc(nx_beg); }
    for (; r < a.n_reads; r += stride) {
        const uint32_t rbeg = nx_beg;
        const int rslots = (int)(nx_end - nx_beg);
        const int plen0 = (int)nx_hdr;
        if (r + stride < a.n_reads) {
            nx_beg = __ldg(a.offsets + r + stride); nx_end = __ldg(a.offsets + r + stride + 1);
            if (nx_end > nx_beg) { nx_hdr = ldc(nx_beg); keep ^= ldc(nx_end - 1); }
        }
        if (rslots == 0) continue;
        const int pn0 = 1 + ((plen0 + 7) >> 3);
        int plen1 = 0;
        bool mine = plen0 >= k && plen0 <= kLaneMaxBases && pn0 <= rslots;
        if (mine && pn0 < rslots) {
            plen1 = (int)ldc(rbeg + (uint32_t)pn0);
            mine = plen1 >= k && plen1 <= kLaneMaxBases && pn0 + 1 + ((plen1 + 7) >> 3) == rslots;
        }
        if (!mine) continue;                          // k_query_block hands these reads to k_query and never looks at pre[r]
        AnchorPre o;
#pragma unroll
        for (int part = 0; part < 2; part++) {
            Bits192 W; W.w[0] = W.w[1] = W.w[2] = 0;
            uint32_t vlab = kLabelNone, vhits = 0;
            const int len = part ? plen1 : plen0;
            if (len - k + 1 >= 16)
                anchor_part<STATS>(t, av, ldc, rbeg + (part ? (uint32_t)pn0 : 0u), len, 1 + ((len + 7) >> 3), k, kmsk, s2, W, vlab, vhits,
                                   st_lookups, st_sectors);
            o.w[part][0] = W.w[0]; o.w[part][1] = W.w[1]; o.w[part][2] = W.w[2];
            o.vote[part] = vhits ? (vlab << 16) | (vhits & 0xFFFFu) : 0u;
        }
        o.pad[0] = o.pad[1] = 0;
        pre[r] = o;
    }
    if (keep == 0xABCDu) pre[0].pad[1] = keep;             // (keeps the look-ahead loads alive; pad is never read)
    if (STATS) {
#pragma unroll
        for (int o = 16; o; o >>= 1) {
            st_lookups += __shfl_xor_sync(0xFFFFFFFFu, st_lookups, o);
            st_sectors += __shfl_xor_sync(0xFFFFFFFFu, st_sectors, o);
        }
        if ((threadIdx.x & 31) == 0) { atomicAdd(a.stats + 0, st_lookups); atomicAdd(a.stats + 1, st_sectors); }
    }
}

template <int BLK, int MINB, bool ANCHOR, bool STATS, bool STAGE, bool PRE = false>
__global__ void __launch_bounds__(BLK, MINB) k_query_block(const TableView t, const QueryArgs a, const AnchorView av,
                                                          uint32_t* redo_list, uint32_t* redo_count)
{
    const AnchorPre* __restrict__ pre = reinterpret_cast<const AnchorPre*>(a.anchor_pre);
    extern __shared__ __align__(16) unsigned char smem_raw[];
    BlockSmem<BLK>& sm = *reinterpret_cast<BlockSmem<BLK>*>(smem_raw);
    StageSmem<BLK>& st = *reinterpret_cast<StageSmem<BLK>*>(smem_raw + ((sizeof(BlockSmem<BLK>) + 15) & ~(size_t)15));
    const int k = t.hp.k;
    const uint64_t kmsk = t.hp.mask;
    const int s2 = 2 * (k - 1);
    const int tid = threadIdx.x;
    unsigned long long st_lookups = 0, st_sectors = 0, st_table = 0;
    const uint32_t nblocks = (a.n_reads + BLK - 1) / BLK;
    // the queue is empty and every read's vote pairs are clear at the top of an iteration: phase C leaves them so
    // (three barriers per iteration: A|B, B|C, C|next A)
    if (tid == 0) sm.qn = 0;
#pragma unroll
    for (int s = 0; s < kLaneAcc; s++) { sm.acc_lab[tid][s] = kLabelNone; sm.acc_cnt[tid][s] = 0; }
    // thread 0: ask for the containers of block nb into buffer b (16-byte granules around the run)
    auto stage_issue = [&](int b, uint32_t nb) {
        const uint32_t r0 = nb * BLK, r1 = min(r0 + (uint32_t)BLK, a.n_reads);
        const uint32_t o0 = __ldg(a.offsets + r0), o1 = __ldg(a.offsets + r1);
        const uintptr_t base = (uintptr_t)a.containers;
        const uintptr_t p0 = (base + 2ull * o0) & ~(uintptr_t)15, p1 = (base + 2ull * o1 + 15) & ~(uintptr_t)15;
        const uint32_t nbytes = (uint32_t)(p1 - p0);
        if (o1 > o0 && p0 >= base && nbytes <= (uint32_t)StageSmem<BLK>::kCap * 2u) {
            st.sbase[b] = (int)((p0 - base) >> 1);
            st.staged[b] = 1;
            const uint32_t bar = smem_addr(&st.bar[b]);
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(nbytes) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         :: "r"(smem_addr(st.cbuf[b])), "l"(p0), "r"(nbytes), "r"(bar) : "memory");
        } else {
            st.staged[b] = 0;
        }
    };
    if (STAGE && tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_addr(&st.bar[0])));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_addr(&st.bar[1])));
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        if (blockIdx.x < nblocks) stage_issue(0, blockIdx.x);
    }
    __syncthreads();
    uint32_t stage_it = 0, stage_par = 0;
    for (uint32_t blk = blockIdx.x; blk < nblocks; blk += gridDim.x) {
        const uint32_t r = blk * BLK + tid;
        // containers of this block: from the staged copy when there is one
        const uint16_t* sb = nullptr; int sbase = 0;
        if (STAGE) {
            const int cur = (int)(stage_it & 1u);
            if (tid == 0 && blk + gridDim.x < nblocks) stage_issue(cur ^ 1, blk + gridDim.x);
            if (st.staged[cur]) {
                const uint32_t bar = smem_addr(&st.bar[cur]);
                uint32_t ok = 0;
                while (!ok)
                    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                                 : "=r"(ok) : "r"(bar), "r"((stage_par >> cur) & 1u) : "memory");
                stage_par ^= 1u << cur;
                sb = st.cbuf[cur]; sbase = st.sbase[cur];
            }
            stage_it++;
        }
        auto ldc = [&](uint32_t idx) -> uint32_t {
            return (STAGE && sb) ? (uint32_t)sb[(int)idx - sbase] : (uint32_t)__ldg(a.containers + idx);
        };
        // ---------------- the read: one part, or two (a pair joined by N, an N inside) ----------------
        uint32_t rbeg = 0; int plen0 = 0, plen1 = 0, pn0 = 0; bool mine = false;
        if (r < a.n_reads) {
            rbeg = __ldg(a.offsets + r);
            const int rslots = (int)(__ldg(a.offsets + r + 1) - rbeg);
            if (rslots == 0) {
                uint16_t* o5 = a.out5 + (size_t)r * 5;
                o5[0] = o5[1] = o5[2] = o5[3] = o5[4] = 0;
                if (a.rows) a.rows[(size_t)r * a.row_stride] = 0;
            } else {
                plen0 = (int)ldc(rbeg); pn0 = 1 + ((plen0 + 7) >> 3);
                mine = plen0 >= k && plen0 <= kLaneMaxBases && pn0 <= rslots;
                if (mine && pn0 < rslots) {
                    plen1 = (int)ldc(rbeg + (uint32_t)pn0);
                    mine = plen1 >= k && plen1 <= kLaneMaxBases && pn0 + 1 + ((plen1 + 7) >> 3) == rslots;
                }
                if (!mine) redo_list[atomicAdd(redo_count, 1u)] = r;      // more parts, longer, shorter than k
            }
        }
        sm.state[tid] = mine ? 1 : 0;
        for (int part = 0; ; part++) {
        // ---------------- phase A ----------------
        {
            const int len = part ? plen1 : plen0;
            if (mine && len > 0) {
                const uint32_t beg = rbeg + (part ? (uint32_t)pn0 : 0u);
                const int nslots = 1 + ((len + 7) >> 3);
                sm.beg[tid] = beg; sm.len[tid] = (uint16_t)len;
                const int npos = len - k + 1;
                Bits192 W; W.w[0] = W.w[1] = W.w[2] = 0;            // windows answered without a table probe
                if (ANCHOR && !PRE && npos >= 16) {      // fewer windows: the index's own line fetches (3-4) outweigh what it can save
                    uint32_t vlab = kLabelNone, vhits = 0;
                    anchor_part<STATS>(t, av, ldc, beg, len, nslots, k, kmsk, s2, W, vlab, vhits, st_lookups, st_sectors);
                    if (vhits) vote_smem(sm, (uint32_t)tid, vlab, vhits);
                }
                if (PRE) {                               // the anchor pass ran as a kernel of its own (k_anchor_scan)
                    const AnchorPre& ap = pre[r];
                    W.w[0] = ap.w[part][0]; W.w[1] = ap.w[part][1]; W.w[2] = ap.w[part][2];
                    const uint32_t v = ap.vote[part];
                    if (v & 0xFFFFu) vote_smem(sm, (uint32_t)tid, v >> 16, v & 0xFFFFu);
                }
                // ---- the windows still to probe: items of <= 4 windows of one 16-window chunk ----
                uint32_t nitems = 0;
#pragma unroll
                for (int ch = 0; ch < 12; ch++) {
                    const int p = 16 * ch;
                    if (p < npos) {
                        const uint64_t wv = ch < 4 ? W.w[0] : (ch < 8 ? W.w[1] : W.w[2]);
                        const int left = npos - p;
                        const uint32_t in = left >= 16 ? 0xFFFFu : ((1u << left) - 1u);
                        nitems += (__popc(~(uint32_t)(wv >> (16 * (ch & 3))) & in) + 3) >> 2;
                    }
                }
                if (nitems) {
                    uint32_t at = atomicAdd(&sm.qn, nitems);
#pragma unroll
                    for (int ch = 0; ch < 12; ch++) {
                        const int p = 16 * ch;
                        if (p < npos) {
                            const uint64_t wv = ch < 4 ? W.w[0] : (ch < 8 ? W.w[1] : W.w[2]);
                            const int left = npos - p;
                            const uint32_t in = left >= 16 ? 0xFFFFu : ((1u << left) - 1u);
                            uint32_t needm = ~(uint32_t)(wv >> (16 * (ch & 3))) & in;
                            while (needm) {
                                uint32_t rest = needm & (needm - 1);
                                rest &= rest - 1; rest &= rest - 1; rest &= rest - 1;                 // lowest four set bits removed
                                sm.queue[at++] = ((uint32_t)tid << 20) | ((uint32_t)ch << 16) | (needm ^ rest);
                                needm = rest;
                            }
                        }
                    }
                }
            }
        }
        __syncthreads();
        // ---------------- phase B ----------------
        const uint32_t qn = sm.qn;
        for (uint32_t it = tid; it < qn; it += BLK) {
            const uint32_t item = sm.queue[it];
            const uint32_t lr = item >> 20, ch = (item >> 16) & 15u;
            uint32_t m4 = item & 0xFFFFu;
            const uint32_t cb = sm.beg[lr];
            // the 48 bases from the chunk's first window on
            uint64_t bhi = 0, blo = 0;
            {
                const int s0 = 1 + 2 * (int)ch;
                const int nsl = 1 + ((sm.len[lr] + 7) >> 3);
#pragma unroll
                for (int j = 0; j < 4; j++) bhi = (bhi << 16) | (s0 + j < nsl ? (uint64_t)ldc(cb + s0 + j) : 0ull);
#pragma unroll
                for (int j = 4; j < 6; j++) blo = (blo << 16) | (s0 + j < nsl ? (uint64_t)ldc(cb + s0 + j) : 0ull);
                blo <<= 32;
            }
            uint64_t e[4][4], home[4], q[4]; bool need[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                need[j] = m4 != 0;
                if (need[j]) {
                    const int o = __ffs(m4) - 1;
                    m4 &= m4 - 1;
                    const uint64_t w64 = o ? (bhi << (2 * o)) | (blo >> (64 - 2 * o)) : bhi;
                    const uint64_t x = w64 >> (64 - 2 * k);
                    const uint64_t rc = revcomp_brev(x, k);
                    const uint64_t h = mix2k(x < rc ? x : rc, k, kmsk);
                    home[j] = home_of(h, t.hp); q[j] = key48_of(h) << 16;
                    load_bucket(t.buckets + home[j], e[j]);
                }
            }
            uint32_t run_lab = kLabelNone, run_cnt = 0;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                if (need[j]) {
                    uint32_t probes = 1;
                    const uint32_t l = resolve(t, home[j], q[j], e[j], STATS ? &probes : nullptr);
                    if (STATS) { st_lookups++; st_table++; st_sectors += probes; }
                    if (l < kLabelTomb) {
                        if (l == run_lab) run_cnt++;
                        else { if (run_cnt) vote_smem(sm, lr, run_lab, run_cnt); run_lab = l; run_cnt = 1; }
                    }
                }
            }
            if (run_cnt) vote_smem(sm, lr, run_lab, run_cnt);
        }
        // B|C barrier, and at the same time: does any read of the block have a second part?
        if (!__syncthreads_or(part == 0 && mine && plen1 > 0)) break;
        if (tid == 0) sm.qn = 0;
        __syncthreads();
        }   // part
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
#pragma unroll
        for (int s = 0; s < kLaneAcc; s++) { sm.acc_lab[tid][s] = kLabelNone; sm.acc_cnt[tid][s] = 0; }
        if (tid == 0) sm.qn = 0;
        __syncthreads();
    }
    if (STATS) {
#pragma unroll
        for (int o = 16; o; o >>= 1) {
            st_lookups += __shfl_xor_sync(0xFFFFFFFFu, st_lookups, o);
            st_sectors += __shfl_xor_sync(0xFFFFFFFFu, st_sectors, o);
            st_table += __shfl_xor_sync(0xFFFFFFFFu, st_table, o);
        }
        if ((threadIdx.x & 31) == 0) {
            atomicAdd(a.stats + 0, st_lookups); atomicAdd(a.stats + 1, st_sectors);
            if (a.stats_words > 2) atomicAdd(a.stats + 2, st_table);
        }
    }
}

namespace {

template <int BLK, int MINB, bool STAGE = false, bool PRE = false>
int launch_block(const TableView& t, const QueryArgs& a, const AnchorView* av, uint32_t* redo_list, uint32_t* redo_count,
                 cudaStream_t s)
{
    // the staged copy moves 16-byte granules: it needs the batch's containers to start on one
    if (STAGE && ((uintptr_t)a.containers & 15u)) return launch_block<BLK, MINB, false, PRE>(t, a, av, redo_list, redo_count, s);
    const size_t smem = ((sizeof(BlockSmem<BLK>) + 15) & ~(size_t)15) + (STAGE ? sizeof(StageSmem<BLK>) : 0);
    uint32_t ctas = (a.n_reads + BLK - 1) / BLK;
    const uint32_t cap = 148u * 64u * (256u / BLK);
    if (ctas > cap) ctas = cap;
    const AnchorView none{};
#define CCB_LAUNCH_BLOCK(AN, ST, AV)                                                                                   \
    do {   /* the attribute is per device: set it on every launch (microseconds) rather than remember who has it */   \
        cudaError_t e = cudaFuncSetAttribute(k_query_block<BLK, MINB, AN, ST, STAGE, PRE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
        if (e != cudaSuccess) return (int)e;                                                                           \
        k_query_block<BLK, MINB, AN, ST, STAGE, PRE><<<ctas, BLK, smem, s>>>(t, a, AV, redo_list, redo_count);                      \
    } while (0)
    if (PRE)     { if (a.stats) CCB_LAUNCH_BLOCK(false, true, none); else CCB_LAUNCH_BLOCK(false, false, none); }
    else if (av) { if (a.stats) CCB_LAUNCH_BLOCK(true, true, *av); else CCB_LAUNCH_BLOCK(true, false, *av); }
    else         { if (a.stats) CCB_LAUNCH_BLOCK(false, true, none); else CCB_LAUNCH_BLOCK(false, false, none); }
#undef CCB_LAUNCH_BLOCK
    return (int)cudaGetLastError();
}

}  // namespace

int launch_query_lane(const TableView& t, const QueryArgs& a, const AnchorView* av, uint32_t* redo_list,
                      uint32_t* redo_count, cudaStream_t s)
{
    if (a.n_reads == 0) return 0;
    // block shape (measured): with the anchor index 128 threads x 4 blocks per SM (120 registers;
    // more blocks spill and are slower, 256 x 2 is 1-2 % slower at the barriers), table only 256 x 3 (79 registers);
    // CCB_BLOCK_SHAPE overrides in measurement builds
#ifdef CCB_EXPERIMENTS     // the other shapes (and TMA staging of the containers) are measurement builds: make VARIANT=exp EXTRA=-DCCB_EXPERIMENTS
    static const char* shape = getenv("CCB_BLOCK_SHAPE");
    if (shape && !strcmp(shape, "128x4")) return launch_block<128, 4>(t, a, av, redo_list, redo_count, s);
    if (shape && !strcmp(shape, "256x3")) return launch_block<256, 3>(t, a, av, redo_list, redo_count, s);
    if (shape && !strcmp(shape, "128x6")) return launch_block<128, 6>(t, a, av, redo_list, redo_count, s);
    if (shape && !strcmp(shape, "256x2")) return launch_block<256, 2>(t, a, av, redo_list, redo_count, s);
    if (shape && !strcmp(shape, "128x4s")) return launch_block<128, 4, true>(t, a, av, redo_list, redo_count, s);
    if (shape && !strcmp(shape, "64x8s")) return launch_block<64, 8, true>(t, a, av, redo_list, redo_count, s);
    if (shape && !strcmp(shape, "256x2s")) return launch_block<256, 2, true>(t, a, av, redo_list, redo_count, s);
    if (shape && !strcmp(shape, "64x8")) return launch_block<64, 8>(t, a, av, redo_list, redo_count, s);
    if (shape && !strcmp(shape, "128x5")) return launch_block<128, 5>(t, a, av, redo_list, redo_count, s);
#endif
    // With the anchor index and scratch for its results (QueryArgs::anchor_pre): the anchor pass as a kernel of its own,
    // then the probe kernel at the table-only shape.  CCB_FUSED_ANCHOR=1 (or no scratch) keeps both in one kernel.
    const char* fe = getenv("CCB_FUSED_ANCHOR");
    const bool fused = fe && atoi(fe);
    if (av && a.anchor_pre && !fused) {
        uint32_t g = (a.n_reads + 127) / 128;
        if (g > 148u * 64u) g = 148u * 64u;
        AnchorPre* pre = reinterpret_cast<AnchorPre*>(a.anchor_pre);
#define CCB_SCAN(K) do { if (a.stats) k_anchor_scan<true, K><<<g, 128, 0, s>>>(t, a, *av, pre); else k_anchor_scan<false, K><<<g, 128, 0, s>>>(t, a, *av, pre); } while (0)
        if (t.hp.k == 31) CCB_SCAN(31); else if (t.hp.k == 27) CCB_SCAN(27); else CCB_SCAN(0);
#undef CCB_SCAN
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return (int)e;
        return launch_block<256, 3, false, true>(t, a, nullptr, redo_list, redo_count, s);
    }
    return av ? launch_block<128, 4>(t, a, av, redo_list, redo_count, s) : launch_block<256, 3>(t, a, av, redo_list, redo_count, s);
}

}  // namespace ccb
