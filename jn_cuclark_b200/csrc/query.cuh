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
    // optional: classify only the reads listed (read_list[0 .. min(*n_list, n_reads)) ), e.g. the ones the
    // thread-per-read kernel handed back
    const uint32_t* read_list;
    const uint32_t* n_list;
};

// Sharded table: each read's mixed k-mers h = mix2k(canonical) are sent
// to the shard that owns them.  Per read r and owner g the k-mers occupy
//   dst[g][pos[g * n_reads + r] - seg_base[g] ...)      (pos = exclusive scan of counts, [owner][read])
// and the labels come back at the same indices of src[g].
constexpr int kMaxShards = 8;
struct RouteArgs {
    uint32_t        n_shards;
    uint32_t*       counts;                 // count pass: u32[n_shards * n_reads]
    const uint32_t* pos;                    // u32[n_shards * n_reads + 1]
    uint64_t        seg_base[kMaxShards];   // pos of owner g's first k-mer (= pos[g * n_reads])
    uint64_t*       dst[kMaxShards];        // route pass: owner g's segment (local send buffer or peer inbox)
    uint32_t        derive_base;            // 1: take seg_base[g] from pos[g * n_reads] on the device (no host round trip)
    uint64_t        cap;                    // elements a segment may hold; stores beyond it are dropped (0 = unlimited)
};
int launch_route(const TableView& t, const QueryArgs& a, const RouteArgs& ra, bool write, cudaStream_t s);
// owner side: labels[i] = label of mixed k-mer hashes[i] (0xFFFF: absent or not discriminative)
int launch_probe_list(const TableView& t, const uint64_t* hashes, uint64_t n, uint16_t* labels,
                      unsigned long long* stats, cudaStream_t s);
// requester side: vote from the returned label lists; overflowing reads go to ovf_list as in launch_query
struct VoteArgs {
    uint32_t        n_shards;
    const uint32_t* pos;
    uint64_t        seg_base[kMaxShards];
    const uint16_t* src[kMaxShards];
    uint32_t        derive_base;
};
// Peer exchange (no NCCL on the data path): after the route pass, tell every owner how many k-mers it
// got from this rank (count_dst[g] points into owner g's memory); over-capacity sets *error.
int launch_post_counts(const uint32_t* pos, uint32_t n_reads, uint32_t n_shards, uint64_t cap,
                       unsigned long long* const* count_dst, uint32_t* error, cudaStream_t s);
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

// Anchor index (DESIGN.md "Anchor and verify"): the target sequences themselves, both strands, 2 bits per
// base, plus per-position bits, plus a small table of sampled k-mers -> text position.  A read whose
// sampled k-mer is found there is compared base by base with the text; every window that matches needs no
// table probe -- its answer is the per-position bit "this text k-mer is in the table" and the label of the
// record.  Windows that do not match are probed as usual, so the result is exact whatever the anchors say.
struct AnchorView {
    const uint32_t*   seq;      // text T = F . revcomp(F): 16 bases per word, first base in the top bits
    const uint32_t*   valid;    // bit p: base p of T is a base (0: separator / N / padding)
    const uint32_t*   disc;     // bit p: the k-mer starting at p is stored in the k-mer table
    uint64_t          nF;       // bases in F (multiple of 32); T has 2 * nF
    const ulonglong4* buckets;  // sampled k-mers: 4 x [fp15 | fwd1 | label16 | pos32], no overflow chains
    uint64_t          nb;
};
// same as launch_query, but reads are first tried against the anchor index
int launch_query_anchor(const TableView& t, const QueryArgs& a, const AnchorView& av, cudaStream_t s);

// Thread-per-read kernel for short single-part reads (<= 192 bases, <= 4 distinct targets hit): every lane
// walks its own read, so nothing is coordinated across lanes.  Reads it does not take (several parts, long,
// many targets) are appended to redo_list / *redo_count and must be classified by launch_query with
// read_list = redo_list afterwards.  av may be null (no anchor index).
int launch_query_lane(const TableView& t, const QueryArgs& a, const AnchorView* av, uint32_t* redo_list,
                      uint32_t* redo_count, cudaStream_t s);

// fast path: one warp per read, votes kept in registers
int launch_query(const TableView& t, const QueryArgs& a, cudaStream_t s);
// exact path for any number of distinct targets: one CTA per listed read, dense
// per-CTA counters in global scratch (n_scratch_blocks * n_targets u32)
int launch_query_dense(const TableView& t, const QueryArgs& a, const uint32_t* read_list,
                       const uint32_t* d_n_list, uint32_t max_list, uint32_t n_targets,
                       uint32_t* scratch, uint32_t n_scratch_blocks, cudaStream_t s);

}  // namespace ccb
