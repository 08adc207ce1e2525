// query.cu -- fused k-mer extraction -> canonicalise -> minimizer -> line fetch -> probe -> vote.
//
// Replaces the reference's queryKernel + mergeKernel + resultKernel
// (src/CuClarkDB.cu:1000-1183, 1261-1355, 1361-1411) by one kernel built around
// what a B200 charges for random access: ~45 G 128-byte
// line requests/s, however little of the line is used.
//
//  * the table is addressed by a minimizer (table.cuh), so the consecutive
//    k-mers of a read mostly share a line;
//  * a lane owns 16 CONSECUTIVE k-mer start positions of one read: forward and
//    reverse-complement words roll, the 27 m-mer hashes of its window come out
//    of the same registers, the sliding minimum is a van Herk prefix/suffix
//    scan in registers (no divergence);
//  * short reads are packed several to a warp (a 150-bp read takes 8 lanes),
//    long reads take the whole warp for several tiles;
//  * each run of positions with the same line issues ONE 128-byte
//    cp.async.bulk (TMA path) into the warp's shared-memory staging area,
//    completion on a per-warp mbarrier; adjacent lanes that continue the same
//    run share the line;
//  * probes then compare the 12 keys of the staged line in shared memory; a k-mer whose home
//    line says "look further" (overflow) is queued and rides along with the
//    NEXT tile's fetch, asking for the next line of its probe sequence, so a
//    tile pays one memory round trip however long the chains are;
//  * votes: a lane run-length-merges its labels, then adds (label,count) to an
//    8-slot accumulator of its read in shared memory; best / second best are
//    taken on the key (count desc, target asc) == the reference's ascending
//    scan (src/CuClarkDB.cu:1380-1399).  Reads with more than 8 distinct
//    targets are re-done exactly by k_query_dense.
#include "query.cuh"

namespace ccb {

constexpr int kWarps      = 1;             // warps per CTA (shared memory, not the CTA shape, bounds residency)
constexpr int kCtasPerSm  = 8;
constexpr int kSeg        = 16;            // k-mer start positions per lane
constexpr int kLinesMax   = 128;           // lines staged per warp tile
constexpr int kLineStride = 144;           // bytes between staged lines (128 + 16: conflict-free sector reads)
constexpr int kAccSlots   = 8;
constexpr int kQueue      = 160;           // k-mers that wait for the next line of their probe sequence
constexpr int kItemRows   = kQueue / 32;
constexpr uint32_t kAccEmpty = 0xFFFFFFFFu;
constexpr uint32_t kNoLine   = 0xFFFFFFFFu;

// a k-mer that was not settled by its home line: it asks for line `hop` of its probe sequence in the NEXT
// tile's fetch, so a tile never waits for more than one memory round trip
struct alignas(16) Item { uint64_t c; uint32_t home; uint16_t hop; uint8_t rid; uint8_t s0; };

struct alignas(128) WarpSmem {
    unsigned char lines[kLinesMax * kLineStride];
    Item     items[2][kQueue];
    uint32_t nq[2];
    uint32_t acc_lab[32][kAccSlots];
    uint32_t acc_cnt[32][kAccSlots];
    uint32_t acc_ovf[32];
    unsigned long long bar;
};

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// add (label, count) to the accumulator of read `rid` of this warp's chunk
static __device__ __noinline__ void acc_add(WarpSmem& ws, int rid, uint32_t lab, uint32_t cnt)
{
    for (int s = 0; s < kAccSlots; s++) {
        uint32_t cur = *(volatile uint32_t*)&ws.acc_lab[rid][s];
        if (cur == kAccEmpty) {
            cur = atomicCAS(&ws.acc_lab[rid][s], kAccEmpty, lab);
            if (cur == kAccEmpty) cur = lab;
        }
        if (cur == lab) { atomicAdd(&ws.acc_cnt[rid][s], cnt); return; }
    }
    ws.acc_ovf[rid] = 1;
}

__device__ __forceinline__ uint64_t min64(uint64_t a, uint64_t b) { return a < b ? a : b; }

// Canonical k-mers and home lines of the 16 positions of a lane.  w0..w2 = 48 bases starting at
// the lane's first position, 16 per word, first base in the top bits.
template <int W>
__device__ __forceinline__ void lane_kmers(uint32_t w0, uint32_t w1, uint32_t w2, int k, uint64_t kmsk, uint32_t n_lines,
                                           uint64_t (&cn)[kSeg], uint32_t (&line)[kSeg])
{
    static_assert(W == 1 || (W == kWin && kSeg > W && kSeg + W - 1 <= 3 * W), "window structure");
    const int s2 = 2 * (k - 1);
    uint64_t x = ((((uint64_t)w0) << 32) | w1) >> (64 - s2);          // first k-1 bases
    uint64_t rc = revcomp(x, k);
    // bases k-1 .. k+14: the 32 bits at bit offset s2 of the window
    const uint32_t r0 = s2 < 32 ? __funnelshift_l(w1, w0, s2) : __funnelshift_l(w2, w1, s2 - 32);
    constexpr int MS = 2 * (W - 1);
    const uint64_t mmsk = kmask(k - (W - 1));
    uint64_t ph[kSeg + W - 1];
#pragma unroll
    for (int i = 0; i < kSeg; i++) {
        const uint64_t code = (r0 >> (30 - 2 * i)) & 3u;
        x = ((x << 2) | code) & kmsk;
        rc = (rc >> 2) | ((3ull - code) << s2);
        cn[i] = min64(x, rc);
        if (W == 1) {
            ph[i] = phi64(cn[i]);
        } else {
            // the k-mer at i holds the m-mers at i (its first m bases) and i+W-1 (its last m bases)
            if (i < W - 1) ph[i] = phi64(min64(x >> MS, rc & mmsk));
            ph[i + W - 1] = phi64(min64(x & mmsk, rc >> MS));
        }
    }
    uint64_t mn[kSeg];
    if (W == 1) {
#pragma unroll
        for (int i = 0; i < kSeg; i++) mn[i] = ph[i];
    } else {
        // sliding minimum over W m-mers: suffix/prefix minima inside blocks of W (van Herk / Gil-Werman)
        constexpr int R = kSeg + W - 1 - 2 * W;
        uint64_t suf0[W], pre1[W], suf1[W], pre2[R > 0 ? R : 1];
        suf0[W - 1] = ph[W - 1];
#pragma unroll
        for (int j = W - 2; j >= 0; j--) suf0[j] = min64(ph[j], suf0[j + 1]);
        pre1[0] = ph[W];
#pragma unroll
        for (int j = 1; j < W; j++) pre1[j] = min64(pre1[j - 1], ph[W + j]);
        suf1[W - 1] = ph[2 * W - 1];
#pragma unroll
        for (int j = W - 2; j >= 0; j--) suf1[j] = min64(ph[W + j], suf1[j + 1]);
        pre2[0] = ph[2 * W];
#pragma unroll
        for (int j = 1; j < R; j++) pre2[j] = min64(pre2[j - 1], ph[2 * W + j]);
        mn[0] = suf0[0];
#pragma unroll
        for (int i = 1; i < W; i++) mn[i] = min64(suf0[i], pre1[i - 1]);
        mn[W] = suf1[0];
#pragma unroll
        for (int i = W + 1; i < kSeg; i++) mn[i] = min64(suf1[i - W], pre2[i - W - 1]);
    }
#pragma unroll
    for (int i = 0; i < kSeg; i++) line[i] = line_of(mn[i], n_lines);
}

// valid k-mer start positions among [P0, P0+16) of a read (bit i <=> position P0+i), from the
// part headers (src/CuClarkDB.cu:1042-1047): a k-mer must lie inside one part.
__device__ __forceinline__ uint32_t valid_mask(const uint16_t* c, int nsl, int P0, int k)
{
    uint32_t vm = 0;
    int ps = 0;
    while (ps < nsl) {
        const int pl = (int)c[ps];
        const int lo = 8 * ps, hi = lo + pl - k;             // valid starts lo..hi (position 0 = first base of slot 1)
        if (lo >= P0 + kSeg) break;
        const int a = max(lo, P0) - P0, b = min(hi, P0 + kSeg - 1) - P0;
        if (a <= b) vm |= ((2u << b) - 1u) & ~((1u << a) - 1u);
        ps += 1 + ((pl + 7) >> 3);
    }
    return vm;
}

// Search a staged line (shared memory, byte address `base`) for c: all 12 keys, straight-line code.
// Returns 0 = resolved (lab = raw label or kLabelNone), 1 = the key may have left for the next
// line of its probe sequence (its bit in the line's filter of leavers is set).
__device__ __forceinline__ int probe_staged(uint32_t base, uint64_t c, uint32_t& lab)
{
    const uint32_t lb = leave_bit(c);
    uint32_t meta = 0;
    lab = kLabelNone;
#pragma unroll
    for (int q = 0; q < 4; q++) {
        uint64_t e[4];
        asm volatile("ld.shared.v2.u64 {%0,%1}, [%2];" : "=l"(e[0]), "=l"(e[1]) : "r"(base + q * 32));
        asm volatile("ld.shared.v2.u64 {%0,%1}, [%2];" : "=l"(e[2]), "=l"(e[3]) : "r"(base + q * 32 + 16));
        if (e[0] == c) lab = (uint32_t)e[3] & 0xFFFFu;
        if (e[1] == c) lab = (uint32_t)(e[3] >> 16) & 0xFFFFu;
        if (e[2] == c) lab = (uint32_t)(e[3] >> 32) & 0xFFFFu;
        if ((lb >> 4) == (uint32_t)q) meta = (uint32_t)(e[3] >> 48);
    }
    return (lab == kLabelNone && ((meta >> (lb & 15u)) & 1u)) ? 1 : 0;
}

template <int W, bool STATS>
__global__ void __launch_bounds__(kWarps * 32, kCtasPerSm)
k_query(const TableView t, const QueryArgs a)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31;
    const int wid  = threadIdx.x >> 5;
    WarpSmem& ws = reinterpret_cast<WarpSmem*>(smem_raw)[wid];
    const int k = t.hp.k;
    const uint64_t kmsk = t.hp.kmsk;
    const uint32_t n_lines = t.hp.n_lines;
    const uint32_t FULL = 0xFFFFFFFFu;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const uint32_t bar_a = smem_addr(&ws.bar);
    const uint32_t lines_a = smem_addr(ws.lines);
    if (lane == 0) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(bar_a));
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    uint32_t parity = 0;
    unsigned long long st_lookups = 0, st_sectors = 0, st_lines = 0;

    const uint32_t n_chunks = (a.n_reads + 31u) / 32u;
    for (uint32_t chunk = blockIdx.x * kWarps + wid; chunk < n_chunks; chunk += gridDim.x * kWarps) {
        // ---- the chunk's reads, one per lane -------------------------------------------------
        const uint32_t r_mine = chunk * 32u + lane;
        const int cnt = (int)min(32u, a.n_reads - chunk * 32u);
        uint32_t beg_mine = 0; int nsl_mine = 0, nseg_mine = 0;
        if (lane < cnt) {
            beg_mine = a.offsets[r_mine];
            nsl_mine = (int)(a.offsets[r_mine + 1] - beg_mine);
            const int tp = 8 * (nsl_mine - 1) - (k - 1);               // k-mer start positions of the container stream
            nseg_mine = tp > 0 ? (tp + kSeg - 1) / kSeg : 0;
        }
#pragma unroll
        for (int s = 0; s < kAccSlots; s++) { ws.acc_lab[lane][s] = kAccEmpty; ws.acc_cnt[lane][s] = 0; }
        ws.acc_ovf[lane] = 0;
        if (lane < 2) ws.nq[lane] = 0;
        __syncwarp();

        int done = 0, long_tile = 0, qsel = 0;
        for (;;) {
            const int n_it = (int)*(volatile uint32_t*)&ws.nq[qsel];       // items carried over from the previous tile
            if (done >= cnt && n_it == 0) break;
            const Item* qin = ws.items[qsel];
            Item* qout = ws.items[qsel ^ 1];

            // ---- tile: which (read, segment) does each lane work on ---------------------------
            int src = 0, seg = 0, n_fin = 0;        // holder lane of my read, my segment, reads whose positions end here
            bool active = false;
            if (done < cnt) {
                const int nseg_first = __shfl_sync(FULL, nseg_mine, done);
                if (nseg_first > 32) {               // long read: the whole warp, several tiles
                    src = done; seg = long_tile * 32 + lane; active = seg < nseg_first;
                    long_tile++;
                    n_fin = long_tile * 32 >= nseg_first ? 1 : 0;
                    if (n_fin) long_tile = 0;
                } else {                             // pack as many consecutive reads as fit in 32 lanes
                    const bool cand = lane >= done && lane < cnt;
                    int incl = cand ? nseg_mine : 0;
#pragma unroll
                    for (int d = 1; d < 32; d <<= 1) { int v = __shfl_up_sync(FULL, incl, d); if (lane >= d) incl += v; }
                    const bool fits = cand && incl <= 32;
                    n_fin = __popc(__ballot_sync(FULL, fits));
                    const int used = __shfl_sync(FULL, incl, done + n_fin - 1);
                    const bool owner = fits && nseg_mine > 0;
                    const uint32_t startmask = __reduce_or_sync(FULL, owner ? (1u << (incl - nseg_mine)) : 0u);
                    const uint32_t nz = __ballot_sync(FULL, owner);
                    active = lane < used;
                    const int idx = __popc(startmask & (lt_mask | (1u << lane)));      // reads starting at or before my lane
                    src = active ? __fns(nz, 0, idx) : done;
                    if (src < 0 || src > 31) src = done;
                    const int a0 = __shfl_sync(FULL, incl - nseg_mine, src);
                    seg = lane - a0;
                }
            }
            const uint32_t rbeg = __shfl_sync(FULL, beg_mine, src);
            const int rnsl = __shfl_sync(FULL, nsl_mine, src);
            const uint16_t* c = a.containers + rbeg;

            // ---- phase A: window, k-mers, minimizers, lines, runs -----------------------------
            uint64_t cn[kSeg]; uint32_t line[kSeg];
            uint32_t vm = 0;
            if (active) {
                const int P0 = kSeg * seg;
                vm = valid_mask(c, rnsl, P0, k);
                const int s0 = 1 + 2 * seg;
                uint32_t h[6];
#pragma unroll
                for (int j = 0; j < 6; j++) h[j] = s0 + j < rnsl ? (uint32_t)__ldg(c + s0 + j) : 0u;
                lane_kmers<W>((h[0] << 16) | h[1], (h[2] << 16) | h[3], (h[4] << 16) | h[5], k, kmsk, n_lines, cn, line);
            }
            // runs: a position starts a run iff its line differs from the previous valid position's line,
            // in lane-major order over the whole tile (so adjacent lanes share a line)
            uint32_t starts = 0, last = kNoLine; int first_i = -1;
#pragma unroll
            for (int i = 0; i < kSeg; i++) {
                if ((vm >> i) & 1u) {
                    if (first_i < 0) first_i = i; else if (line[i] != last) starts |= 1u << i;
                    last = line[i];
                }
            }
            uint32_t carry = last;                   // last line of the nearest lane <= me that has a valid position
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { uint32_t v = __shfl_up_sync(FULL, carry, d); if (lane >= d && carry == kNoLine) carry = v; }
            uint32_t inh = __shfl_up_sync(FULL, carry, 1);
            if (lane == 0) inh = kNoLine;
            if (first_i >= 0) {
                uint32_t fl = kNoLine;
#pragma unroll
                for (int i = 0; i < kSeg; i++) if (i == first_i) fl = line[i];
                if (fl != inh) starts |= 1u << first_i;
            }
            const int my_lines = __popc(starts);
            int off = my_lines;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { int v = __shfl_up_sync(FULL, off, d); if (lane >= d) off += v; }
            const int pos_lines = __shfl_sync(FULL, off, 31);
            off -= my_lines;                          // exclusive

            // ---- carried items: the next line of their probe sequence --------------------------
            // item j is handled by lane j & 31 in row j >> 5; items of one run are adjacent in the queue
            uint32_t it_line[kItemRows]; int it_slot[kItemRows]; uint32_t it_start = 0;
            int item_lines = 0;
#pragma unroll
            for (int r = 0; r < kItemRows; r++) {
                const int j = r * 32 + lane;
                bool st = false;
                it_line[r] = kNoLine;
                if (j < n_it) {
                    it_line[r] = hop_line(qin[j].home, qin[j].hop, n_lines);
                    st = j == 0 || hop_line(qin[j - 1].home, qin[j - 1].hop, n_lines) != it_line[r];
                }
                const uint32_t b = __ballot_sync(FULL, st);
                it_slot[r] = pos_lines + item_lines + __popc(b & (lt_mask | (1u << lane))) - 1;
                if (st) it_start |= 1u << r;
                item_lines += __popc(b);
            }
            const int total_lines = pos_lines + item_lines;
            const int n_stage = min(total_lines, kLinesMax);

            // ---- phase B: one 128-byte bulk copy per run into the staging area ---------------
            if (n_stage > 0) {
                if (lane == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar_a), "r"(n_stage * 128) : "memory");
                __syncwarp();
#pragma unroll
                for (int i = 0; i < kSeg; i++) {
                    if ((starts >> i) & 1u) {
                        const int slot = off + __popc(starts & ((1u << i) - 1u));
                        if (slot < kLinesMax)
                            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], 128, [%2];"
                                         :: "r"(lines_a + slot * kLineStride), "l"(t.sectors + 4ull * line[i]), "r"(bar_a) : "memory");
                    }
                }
#pragma unroll
                for (int r = 0; r < kItemRows; r++) {
                    if (((it_start >> r) & 1u) && it_slot[r] < kLinesMax)
                        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], 128, [%2];"
                                     :: "r"(lines_a + it_slot[r] * kLineStride), "l"(t.sectors + 4ull * it_line[r]), "r"(bar_a) : "memory");
                }
                uint32_t ok = 0;
                while (!ok)
                    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                                 : "=r"(ok) : "r"(bar_a), "r"(parity) : "memory");
                parity ^= 1u;
            }
            if (STATS) st_lines += my_lines + __popc(it_start);

            // ---- phase C: probes from shared memory, votes ------------------------------------
            uint32_t cur_lab = kLabelNone, cur_cnt = 0, pend = 0;
#pragma unroll
            for (int i = 0; i < kSeg; i++) {
                if ((vm >> i) & 1u) {
                    const int slot = off + __popc(starts & ((2u << i) - 1u)) - 1;
                    uint32_t lab = kLabelNone, nsec = 4;
                    if (slot < kLinesMax) {
                        if (probe_staged(lines_a + slot * kLineStride, cn[i], lab)) pend |= 1u << i;
                    } else {
                        nsec = 0;
                        lab = probe_chain(t, line[i], cn[i], 0, STATS ? &nsec : nullptr);
                    }
                    if (STATS) { st_lookups++; st_sectors += nsec; }
                    if (lab < kLabelTomb) {
                        if (lab == cur_lab) cur_cnt++;
                        else { if (cur_cnt) acc_add(ws, src, cur_lab, cur_cnt); cur_lab = lab; cur_cnt = 1; }
                    }
                }
            }
            if (cur_cnt) acc_add(ws, src, cur_lab, cur_cnt);

            // carried items against the line they asked for
            uint32_t requeue = 0;
#pragma unroll
            for (int r = 0; r < kItemRows; r++) {
                const int j = r * 32 + lane;
                if (j < n_it) {
                    const Item it = qin[j];
                    uint32_t lab = kLabelNone, nsec = 4;
                    if (it_slot[r] < kLinesMax) {
                        if (probe_staged(lines_a + it_slot[r] * kLineStride, it.c, lab)) {
                            if (it.hop + 1 < kMaxHops) requeue |= 1u << r;
                        }
                    } else {
                        nsec = 0;
                        lab = probe_chain(t, it.home, it.c, it.hop, STATS ? &nsec : nullptr);
                    }
                    if (STATS) st_sectors += nsec;
                    if (lab < kLabelTomb) acc_add(ws, it.rid, lab, 1);
                }
            }

            // ---- queue for the next tile: my unresolved positions (lane-major, so the k-mers of one
            //      run stay adjacent), then the items that have to hop again ----------------------
            const int my_pend = __popc(pend);
            int poff = my_pend;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { int v = __shfl_up_sync(FULL, poff, d); if (lane >= d) poff += v; }
            const int pos_items = __shfl_sync(FULL, poff, 31);
            poff -= my_pend;
#pragma unroll
            for (int i = 0; i < kSeg; i++) {
                if ((pend >> i) & 1u) {
                    const int idx = poff + __popc(pend & ((1u << i) - 1u));
                    if (idx < kQueue) {
                        Item it; it.c = cn[i]; it.home = line[i]; it.hop = 1; it.rid = (uint8_t)src; it.s0 = 0;
                        qout[idx] = it;
                    } else {                         // queue full: finish this one in global memory now
                        uint32_t nsec = 0;
                        const uint32_t lab = probe_chain(t, line[i], cn[i], 1, STATS ? &nsec : nullptr);
                        if (STATS) st_sectors += nsec;
                        if (lab < kLabelTomb) acc_add(ws, src, lab, 1);
                    }
                }
            }
            int n_out = pos_items;
#pragma unroll
            for (int r = 0; r < kItemRows; r++) {
                const bool rq = (requeue >> r) & 1u;
                const uint32_t b = __ballot_sync(FULL, rq);
                if (rq) {
                    Item it = qin[r * 32 + lane];
                    it.hop++;
                    const int idx = n_out + __popc(b & lt_mask);
                    if (idx < kQueue) qout[idx] = it;
                    else {
                        uint32_t nsec = 0;
                        const uint32_t lab = probe_chain(t, it.home, it.c, it.hop, STATS ? &nsec : nullptr);
                        if (STATS) st_sectors += nsec;
                        if (lab < kLabelTomb) acc_add(ws, it.rid, lab, 1);
                    }
                }
                n_out += __popc(b);
            }
            __syncwarp();
            if (lane == 0) { ws.nq[qsel] = 0; ws.nq[qsel ^ 1] = (uint32_t)min(n_out, kQueue); }
            qsel ^= 1;
            done += n_fin;
            __syncwarp();
        }

        // ---- results of the chunk's reads (src/CuClarkDB.cu:1370-1405) --------------------------
        if (lane < cnt) {
            uint32_t sum = 0, k1 = 0, k2 = 0; int nnz = 0;
            for (int s = 0; s < kAccSlots; s++) {
                const uint32_t lab = ws.acc_lab[lane][s];
                if (lab == kAccEmpty) break;
                const uint32_t c16 = ws.acc_cnt[lane][s] & 0xFFFFu;      // reference counters are u16
                if (!c16) continue;
                nnz++; sum += c16;
                const uint32_t key = (c16 << 16) | (0xFFFFu - lab);
                if (key > k1) { k2 = k1; k1 = key; } else if (key > k2) k2 = key;
            }
            uint16_t* o5 = a.out5 + (size_t)r_mine * 5;
            o5[0] = (uint16_t)sum;
            o5[1] = (uint16_t)(k1 ? (0xFFFFu - (k1 & 0xFFFFu)) + 1u : 0u);
            o5[2] = (uint16_t)(k1 >> 16);
            o5[3] = (uint16_t)(k2 ? (0xFFFFu - (k2 & 0xFFFFu)) + 1u : 0u);
            o5[4] = (uint16_t)(k2 >> 16);
            if (a.rows) {
                uint16_t* row = a.rows + (size_t)r_mine * a.row_stride;
                const int fit = (a.row_stride - 1) / 2;
                row[0] = (uint16_t)min(nnz, fit);
                for (int s = 0; s < kAccSlots; s++) {
                    const uint32_t lab = ws.acc_lab[lane][s];
                    if (lab == kAccEmpty) break;
                    const uint32_t c16 = ws.acc_cnt[lane][s] & 0xFFFFu;
                    if (!c16) continue;
                    int rank = 0;
                    for (int u = 0; u < kAccSlots; u++) {
                        const uint32_t l2 = ws.acc_lab[lane][u];
                        if (l2 != kAccEmpty && (ws.acc_cnt[lane][u] & 0xFFFFu) && l2 < lab) rank++;
                    }
                    if (rank < fit) { row[1 + 2 * rank] = (uint16_t)lab; row[2 + 2 * rank] = (uint16_t)c16; }
                }
            }
            if (ws.acc_ovf[lane] && a.ovf_list) {
                const uint32_t slot = atomicAdd(a.ovf_count, 1u);
                a.ovf_list[slot] = r_mine;
            }
        }
        __syncwarp();
    }
    if (STATS) {
#pragma unroll
        for (int o = 16; o; o >>= 1) {
            st_lookups += __shfl_xor_sync(FULL, st_lookups, o);
            st_sectors += __shfl_xor_sync(FULL, st_sectors, o);
            st_lines += __shfl_xor_sync(FULL, st_lines, o);
        }
        if (lane == 0) { atomicAdd(a.stats + 0, st_lookups); atomicAdd(a.stats + 1, st_sectors); atomicAdd(a.stats + 2, st_lines); }
    }
}

template <int W, bool STATS>
static int launch_one(const TableView& t, const QueryArgs& a, cudaStream_t s)
{
    static bool configured = false;          // per instantiation; the attribute is per device function and idempotent
    const int smem = kWarps * (int)sizeof(WarpSmem);
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(k_query<W, STATS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return (int)e;
        configured = true;
    }
    const uint32_t n_chunks = (a.n_reads + 31u) / 32u;
    uint32_t ctas = (n_chunks + kWarps - 1) / kWarps;
    const uint32_t cap = 148u * kCtasPerSm;         // resident CTAs, warps stride over chunks of 32 reads
    if (ctas > cap) ctas = cap;
    k_query<W, STATS><<<ctas, kWarps * 32, smem, s>>>(t, a);
    return (int)cudaGetLastError();
}

int launch_query(const TableView& t, const QueryArgs& a, cudaStream_t s)
{
    if (a.n_reads == 0) return 0;
    const bool win = t.hp.m != t.hp.k;              // k >= 23: window of kWin m-mers
    if (win) return a.stats ? launch_one<kWin, true>(t, a, s) : launch_one<kWin, false>(t, a, s);
    return a.stats ? launch_one<1, true>(t, a, s) : launch_one<1, false>(t, a, s);
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
                uint64_t x = kmer_at(data, j, k, t.hp.kmsk);
                uint32_t lab = lookup(t, canonical(x, k));
                if (lab != kLabelNone && lab < n_targets) atomicAdd(cnt + lab, 1u);
            }
        }
        __syncthreads();
        // top-2 under (count desc, target asc) + sum
        uint32_t k1 = 0, k2 = 0, sum = 0;
        for (uint32_t i = threadIdx.x; i < n_targets; i += blockDim.x) {
            uint32_t c16 = cnt[i] & 0xFFFFu;
            if (!c16) continue;
            sum += c16;
            uint32_t key = (c16 << 16) | (0xFFFFu - i);
            if (key > k1) { k2 = k1; k1 = key; } else if (key > k2) k2 = key;
        }
        s_key[0][threadIdx.x] = k1; s_key[1][threadIdx.x] = k2; s_sum[threadIdx.x] = sum;
        __syncthreads();
        for (int o = 128; o; o >>= 1) {
            if ((int)threadIdx.x < o) {
                uint32_t a1 = s_key[0][threadIdx.x], a2 = s_key[1][threadIdx.x];
                uint32_t b1 = s_key[0][threadIdx.x + o], b2 = s_key[1][threadIdx.x + o];
                uint32_t m1 = max(a1, b1);
                uint32_t m2 = max(min(a1, b1), max(a2, b2));
                s_key[0][threadIdx.x] = m1; s_key[1][threadIdx.x] = m2;
                s_sum[threadIdx.x] += s_sum[threadIdx.x + o];
            }
            __syncthreads();
        }
        if (threadIdx.x == 0) {
            k1 = s_key[0][0]; k2 = s_key[1][0];
            uint16_t* o5 = a.out5 + (size_t)r * 5;
            o5[0] = (uint16_t)s_sum[0];
            o5[1] = (uint16_t)(k1 ? (0xFFFFu - (k1 & 0xFFFFu)) + 1u : 0u);
            o5[2] = (uint16_t)(k1 >> 16);
            o5[3] = (uint16_t)(k2 ? (0xFFFFu - (k2 & 0xFFFFu)) + 1u : 0u);
            o5[4] = (uint16_t)(k2 >> 16);
            if (a.rows) {
                uint16_t* row = a.rows + (size_t)r * a.row_stride;
                int n = 0, fit = (a.row_stride - 1) / 2;
                for (uint32_t i = 0; i < n_targets && n < fit; i++)
                    if (cnt[i] & 0xFFFFu) { row[1 + 2 * n] = (uint16_t)i; row[2 + 2 * n] = (uint16_t)cnt[i]; n++; }
                row[0] = (uint16_t)n;
            }
        }
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

}  // namespace ccb
