// query.cuh -- launch interface of the classification kernels.
#pragma once
#include "table.cuh"

namespace ccb {

struct QueryArgs {
    // batch in the reference's packed layout (src/CuCLARK_hh.hh:1615-1715):
    // offsets[n_reads+1] in units of u16, each read = parts [len][ceil(len/8) containers]
    const uint32_t* offsets;
    const uint16_t* containers;
    uint32_t        n_reads;
    // results: u16[5] per read {sumN, idxBest, best, idxSecond, second}
    // (src/CuClarkDB.cu:1401-1405); optional sparse rows {n,(t,c)*} ascending t
    uint16_t*       out5;
    uint16_t*       rows;
    int             row_stride;      // u16 per row
    // reads that hit more than 32 distinct targets are re-done by the dense kernel
    uint32_t*       ovf_list;
    uint32_t*       ovf_count;
    // optional accounting [0]=k-mers probed [1]=32-byte sectors read
    unsigned long long* stats;
    int             stats_words;     // 2, or 4: [2] = k-mers probed in the k-mer table (not answered by the anchor index)
    // optional: classify only the reads listed (read_list[0 .. min(*n_list, n_reads)) ), e.g. the ones the
    // thread-per-read kernel handed back
    const uint32_t* read_list;
    const uint32_t* n_list;
    // optional scratch, n_reads x kAnchorPreBytes: with it the anchor pass runs as a kernel of its own (query_lane.cu)
    void*           anchor_pre;
};
constexpr int kAnchorPreBytes = 64;

// Sharded table: each read's mixed k-mers h = mix2k(canonical) are sent
// to the shard that owns them.  Per read r and owner g the k-mers occupy
//   dst[g][pos[g * n_reads + r] - seg_base[g] ...)      (pos = exclusive scan of counts, [owner][read])
// and the labels come back at the same indices of src[g].
constexpr int kMaxShards = 8;
struct RouteArgs {
    uint32_t        n_shards;
    const unsigned long long* dst_off;      // optional (device): owner g's segment starts dst_off[g] entries into dst[g]
                                            // (the block path's k-mers come first, shard_block.cu)
    uint32_t*       counts;                 // count pass: u32[n_shards * n_reads]
    const uint32_t* pos;                    // u32[n_shards * n_reads + 1]
    uint64_t        seg_base[kMaxShards];   // pos of owner g's first k-mer (= pos[g * n_reads])
    uint64_t*       dst[kMaxShards];        // route pass: owner g's segment (local send buffer or peer inbox)
    uint32_t        derive_base;            // 1: take seg_base[g] from pos[g * n_reads] on the device (no host round trip)
    uint64_t        cap;                    // elements a segment may hold; stores beyond it are dropped (0 = unlimited)
};
int launch_route(const TableView& t, const QueryArgs& a, const RouteArgs& ra, bool write, cudaStream_t s, uint32_t max_ctas = 0);
// owner side: labels[i] = label of mixed k-mer hashes[i] (0xFFFF: absent or not discriminative)
int launch_probe_list(const TableView& t, const uint64_t* hashes, uint64_t n, uint16_t* labels,
                      unsigned long long* stats, cudaStream_t s);
// requester side: vote from the returned label lists; overflowing reads go to ovf_list as in launch_query
struct VoteArgs {
    uint32_t        n_shards;
    const unsigned long long* src_off;      // optional (device): as RouteArgs::dst_off
    uint64_t        cap;                    // entries a region holds (0 = unlimited): labels beyond it were never written
                                            // (the route pass dropped the k-mers and raised the error word) and are not read
    const uint32_t* pos;
    uint64_t        seg_base[kMaxShards];
    const uint16_t* src[kMaxShards];
    uint32_t        derive_base;
};
// Peer exchange (no NCCL on the data path): after the route pass, tell every owner how many k-mers it
// got from this rank (count_dst[g] points into owner g's memory); over-capacity sets *error.
// `first` (optional, device): entries the block path already put into each region
int launch_post_counts(const uint32_t* pos, uint32_t n_reads, uint32_t n_shards, uint64_t cap,
                       unsigned long long* const* count_dst, uint32_t* error, cudaStream_t s,
                       const unsigned long long* first = nullptr);
// owner side over an inbox of n_shards regions of `cap` k-mers each, region s filled by rank s with
// counts_in[s] k-mers; the label of inbox[s][i] is stored to lab_dst[s][i] (rank s's memory)
struct InboxArgs {
    uint32_t        n_shards;
    uint64_t        cap;
    const uint64_t* inbox;
    const unsigned long long* counts_in;
    uint16_t*       lab_dst[kMaxShards];
};
int launch_probe_inbox(const TableView& t, const InboxArgs& ia, unsigned long long* stats, cudaStream_t s);
int launch_vote_lists(const QueryArgs& a, const VoteArgs& va, cudaStream_t s);
int launch_vote_lists_dense(const QueryArgs& a, const VoteArgs& va, const uint32_t* read_list, const uint32_t* d_n_list,
                            uint32_t max_list, uint32_t n_targets, uint32_t* scratch, uint32_t n_scratch_blocks, cudaStream_t s);

// One-pass routing and block-wise voting for short reads (shard_block.cu): a block of kRouteBlk reads owns one
// contiguous segment per owner, reserved with one atomicAdd on the requester's cursor.
constexpr int kRouteBlk = 256;
struct BlockRouteArgs {
    uint32_t  n_shards;
    uint64_t* dst[kMaxShards];        // owner g's region for this requester (local send segment or a peer's inbox)
    uint8_t*  rid[kMaxShards];        // local, same indices as dst[g]: which read of its block the k-mer belongs to
    uint64_t  cap;                    // entries a region holds (< 2^32); the excess is dropped and *error set
    unsigned long long* cursor;       // [kMaxShards] entries reserved per owner so far
    uint32_t* blkseg;                 // [blocks][2 parts][kMaxShards][2]: first entry, entries of (block, part, owner)
    uint8_t*  rstate;                 // per read: 0 no k-mers, 1 routed by the block kernel, 2 handed to the generic path
    uint32_t* redo_list; uint32_t* redo_count;
    uint32_t* error;
};
int launch_route_block(const HashParams& hp, const QueryArgs& a, const BlockRouteArgs& ra, cudaStream_t s);
struct BlockVoteArgs {
    uint32_t  n_shards;
    const uint16_t* src[kMaxShards];  // labels, same indices as BlockRouteArgs::dst
    const uint8_t*  rid[kMaxShards];
    uint64_t  cap;
    const uint32_t* blkseg;
    const uint8_t*  rstate;
};
// reads with rstate 1 get their result; a read that met a fifth target goes to a.ovf_list / a.ovf_count
int launch_vote_block(const QueryArgs& a, const BlockVoteArgs& va, cudaStream_t s);
int launch_vote_block_dense(const QueryArgs& a, const BlockVoteArgs& va, const uint32_t* read_list, const uint32_t* d_n_list,
                            uint32_t max_list, uint32_t n_targets, uint32_t* scratch, uint32_t n_scratch_blocks, cudaStream_t s);

// Anchor index (DESIGN.md "Anchor and verify"): the target sequences themselves, both strands, plus a small
// table of sampled k-mers -> text position.  A read whose sampled k-mer is found there is compared base by
// base with the text; a window that matches needs no table probe -- the text knows whether the k-mer starting
// there is stored in the table under the record's own label ("same"), under another label ("other": probed
// as usual) or not at all.  Windows that do not match are probed as usual, so the result is exact whatever
// the sampled table holds.
constexpr int kTextWords = 5;   // per 32 text positions
struct AnchorView {
    // text T = F . revcomp(F), group g = positions [32g, 32g+32), 5 words:
    //   [0] bases 0..15, [1] bases 16..31   (2 bits per base, first base in the top bits)
    //   [2] valid  bit i: position 32g+i is a base (0: separator / N / padding)
    //   [3] same   bit i: the k-mer starting there is in the k-mer table under the label of its record
    //   [4] other  bit i: ... is in the k-mer table under a different label
    const uint32_t*   txt;
    uint64_t          nF;       // bases in F (multiple of 32); T has 2 * nF
    // sampled k-mers of F: 32-byte sectors of 4 x [fp15 | fwd1 | label16 | pos32]; sector b overflows into b ^ 1
    // (same 128-byte line), a k-mer that finds both full is not indexed
    const ulonglong4* buckets;
    uint64_t          nb;       // sectors (even)
};
#if defined(__CUDACC__)
// which canonical k-mers are sampled (1 in 8): cheap enough to evaluate for every window of a read
__device__ __forceinline__ bool anchor_sampled(uint64_t c)
{
    return ((((uint32_t)c ^ (uint32_t)(c >> 32)) * 0x9E3779B1u) >> 29) == 0u;
}
#endif

// Thread-per-read kernel for short single-part reads (<= 192 bases, <= 4 distinct targets hit): phase A, every
// thread takes one read: looks for an anchor, compares the read with the text; phase B, the windows still to
// probe are dealt out over the block four at a time; phase C, one result per thread.  Reads it does not take
// (several parts, long, many targets) are appended to redo_list / *redo_count and must be classified by
// launch_query with read_list = redo_list afterwards.  av may be null (no anchor index).
int launch_query_lane(const TableView& t, const QueryArgs& a, const AnchorView* av, uint32_t* redo_list,
                      uint32_t* redo_count, cudaStream_t s);

// fast path: one warp per read, votes kept in registers
// max_ctas: grid cap (0 = the default, 128 CTAs per SM); a pass over a short read_list wants a small grid --
// 19 k CTAs that find nothing to do still cost 56 us
int launch_query(const TableView& t, const QueryArgs& a, cudaStream_t s, uint32_t max_ctas = 0);
// exact path for any number of distinct targets: one CTA per listed read, dense
// per-CTA counters in global scratch (n_scratch_blocks * n_targets u32)
int launch_query_dense(const TableView& t, const QueryArgs& a, const uint32_t* read_list,
                       const uint32_t* d_n_list, uint32_t max_list, uint32_t n_targets,
                       uint32_t* scratch, uint32_t n_scratch_blocks, cudaStream_t s);

}  // namespace ccb
