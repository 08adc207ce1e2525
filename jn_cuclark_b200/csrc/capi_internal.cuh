// capi_internal.cuh -- state behind the opaque ccb_handle (shared by capi.cu and synth.cu).
#pragma once
#include "../../include/cuclark_b200.h"
#include "query.cuh"
#include "table.cuh"
#include "hostnuma.hpp"
#include <mutex>
#include <string>
#include <vector>

namespace ccb {

struct Slot {
    cudaStream_t stream = nullptr;
    cudaEvent_t  done = nullptr;
    // pinned host
    uint32_t* h_offsets = nullptr; uint16_t* h_cont = nullptr;
    uint16_t* h_final5 = nullptr;  uint16_t* h_rows = nullptr;
    // device
    uint32_t* d_offsets = nullptr; uint16_t* d_cont = nullptr;
    uint16_t* d_final5 = nullptr;  uint16_t* d_rows = nullptr;
    uint32_t* d_ovf_list = nullptr; uint32_t* d_ovf_count = nullptr;
    uint32_t* d_dense_scratch = nullptr;
    void*     d_anchor_pre = nullptr;          // max_reads x kAnchorPreBytes: results of the anchor pass (query_lane.cu)
    unsigned long long* d_stats = nullptr;     // lookups, sectors
    uint32_t* h_ovf_count = nullptr;           // pinned
    unsigned long long* h_stats = nullptr;     // pinned
    bool      in_flight = false;
    bool      shared_results = false;          // h_final5 / h_rows point into the handle's contiguous result arrays
    uint32_t  n_reads = 0;
    int       flags = 0;
};

constexpr uint32_t kDenseBlocks = 64;

// the calling thread on the CPUs next to `device` while this lives: pinned memory allocated meanwhile is local to the GPU
struct DeviceNodeBind {
    char id[32];
    ccbh::NodeBind bind;
    static const char* bus_id(int device, char (&buf)[32])
    {
        buf[0] = 0;
        if (cudaDeviceGetPCIBusId(buf, (int)sizeof buf, device) != cudaSuccess) { cudaGetLastError(); buf[0] = 0; }
        return buf;
    }
    explicit DeviceNodeBind(int device) : bind(bus_id(device, id)) {}
};

// anchor index owner (anchor.cu)
struct Anchor {
    uint32_t* d_txt = nullptr;
    ulonglong4* d_buckets = nullptr;
    uint64_t nF = 0, nb = 0, entries = 0, dropped = 0, bytes = 0;
    bool ready = false;
    AnchorView view() const { return AnchorView{d_txt, nF, d_buckets, nb}; }
};
void anchor_free(Anchor& a);
struct AnchorBuild;                  // anchor.cu: target texts being collected (ccb_anchor_begin .. ccb_anchor_end)
void anchor_build_free(ccb_handle* h);
struct BuildState;                   // dbbuild.cu
// shard.cu: exclusive prefix sum of u32 in place (three small kernels); tmp holds scan_tmp_words(n) u32
size_t scan_tmp_words(size_t n);
int    exclusive_scan_u32(uint32_t* d, size_t n, uint32_t* tmp, cudaStream_t s);
// group.cu: several devices, one sharded table, behind one leader handle
struct Group;
void group_destroy(ccb_handle* h);
int  group_db_load_files(ccb_handle* h, const char* prefix, uint64_t htsize, int key_bytes, unsigned sampling, uint64_t* n_loaded);
int  group_db_get_info(const ccb_handle* h, ccb_db_info* info);
int  group_batch_alloc(ccb_handle* h, uint32_t max_reads, uint64_t max_containers, int extended);
int  group_batch_buffers(ccb_handle* h, int slot, uint32_t** offsets, uint16_t** containers, uint16_t** final5, uint16_t** rows);
int  group_batch_submit(ccb_handle* h, int slot, uint32_t n_reads, uint64_t n_containers, int flags);
int  group_batch_wait(ccb_handle* h, int slot);
int  group_batch_poll(ccb_handle* h, int slot);
int  group_batch_free(ccb_handle* h);
int  group_sync(ccb_handle* h);

}  // namespace ccb

struct ccb_handle {
    ccb_config cfg{};
    ccb::Table table;
    bool       db_ready = false, db_building = false;
    cudaStream_t db_stream = nullptr;
    std::vector<ccb::Slot> slots;
    uint32_t   max_reads = 0; uint64_t max_cont = 0; bool batch_extended = false;
    // ccb_batch_alloc_ref: ONE pinned final / full result array for all reads of a file, as CuClarkDB::malloc hands out
    uint16_t*  ref_final = nullptr; uint16_t* ref_full = nullptr;
    int        row_stride = 0;
    ccb_stats  stats{};
    uint64_t   tombstones = 0; uint32_t max_disp = 0;
    std::string err;
    std::mutex  mu;                  // guards err and stats
    int        l2_fetch = 0;          // cudaLimitMaxL2FetchGranularity in effect
    // scratch for ccb_query_device
    uint32_t* q_ovf_list = nullptr; uint32_t* q_ovf_count = nullptr; uint32_t* q_dense = nullptr;
    uint32_t  q_ovf_cap = 0;
    void*     q_anchor_pre = nullptr; uint32_t q_pre_cap = 0;
    // sharded table (ccb_shard_*): this handle holds shard `shard` of `n_shards`
    uint32_t   n_shards = 1, shard = 0;
    // scratch of ccb_shard_count / ccb_shard_vote, one set per stream they were called on (batches of a
    // sharded pipeline overlap on two streams); never shared with ccb_query_device's q_* buffers
    struct ShardScratch { void* stream = nullptr; bool used = false; void* scan_tmp = nullptr; size_t scan_bytes = 0;
                          uint32_t* ovf_list = nullptr; uint32_t ovf_cap = 0; uint32_t* ovf_count = nullptr; uint32_t* dense = nullptr; };
    ShardScratch shard_scratch[4];
    ccb::Group* group = nullptr;     // leader of a device group (ccb_group_create): the batch calls go to group.cu
    ccb::Anchor anchor;              // optional: sequences + sampled k-mer positions in front of the table
    ccb::AnchorBuild* anchor_build = nullptr;
    // database construction in progress (ccb_build_*)
    ccb::BuildState* build = nullptr;
    std::mutex  build_mu;
};


// records the message on the handle (or thread-locally when h == NULL), returns code (never 0)
int ccb_fail(ccb_handle* h, int code, const char* fmt, ...);
