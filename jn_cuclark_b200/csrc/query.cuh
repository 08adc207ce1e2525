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
};

// fast path: one warp per read, votes kept in registers
int launch_query(const TableView& t, const QueryArgs& a, cudaStream_t s);
// exact path for any number of distinct targets: one CTA per listed read, dense
// per-CTA counters in global scratch (n_scratch_blocks * n_targets u32)
int launch_query_dense(const TableView& t, const QueryArgs& a, const uint32_t* read_list,
                       const uint32_t* d_n_list, uint32_t max_list, uint32_t n_targets,
                       uint32_t* scratch, uint32_t n_scratch_blocks, cudaStream_t s);

}  // namespace ccb
