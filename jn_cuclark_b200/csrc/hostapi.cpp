// hostapi.cpp -- C ABI (include/ccb_host.h) over hostpipe.cpp
#include "../../include/ccb_host.h"
#include "hostpipe.hpp"
#include "hostnuma.hpp"
#include <stdlib.h>
#include <string.h>

struct ccb_reads { ccbh::ReadSet rs; };

extern "C" ccb_reads* ccb_host_index(const uint8_t* map, size_t nb, int threads)
{
    ccb_reads* r = new ccb_reads();
    if (ccbh::index_reads(map, nb, threads, r->rs) != 0) { delete r; return nullptr; }
    return r;
}
extern "C" int ccb_host_index_fastq_chunked(const uint8_t* map, size_t nb, size_t chunk, ccb_reads** out)
{
    if (!out) return -1;
    *out = nullptr;
    ccb_reads* r = new ccb_reads();
    const int rc = ccbh::index_fastq_chunked(map, nb, chunk, r->rs);
    if (rc) { delete r; return rc; }
    *out = r;
    return 0;
}
extern "C" void ccb_host_reads_free(ccb_reads* r) { delete r; }
extern "C" size_t ccb_host_reads_count(const ccb_reads* r) { return r ? r->rs.n() : 0; }
extern "C" int ccb_host_reads_is_fasta(const ccb_reads* r) { return r && r->rs.fasta; }
extern "C" const uint64_t* ccb_host_reads_field(const ccb_reads* r, int field)
{
    if (!r) return nullptr;
    switch (field) {
    case 0: return r->rs.name_s.data();
    case 1: return r->rs.name_e.data();
    case 2: return r->rs.seq_s.data();
    case 3: return r->rs.seq_e.data();
    case 4: return r->rs.len.data();
    default: return nullptr;
    }
}
extern "C" size_t ccb_host_pack_bound(const ccb_reads* r, size_t first, size_t count, int k)
{
    return ccbh::pack_bound(r->rs, first, count, k);
}
extern "C" int64_t ccb_host_pack(const ccb_reads* r, size_t first, size_t count, int k,
                                 uint32_t* offsets, uint16_t* containers, size_t cap)
{
    if (!r || first + count > r->rs.n()) return -3;
    return ccbh::pack_reads(r->rs, first, count, k, offsets, containers, cap);
}
extern "C" int ccb_host_join_pairs(const char* f1, const char* f2, char** out, size_t* len, char* err, size_t errcap)
{
    ccbh::TextBuf text; std::string e;
    int rc = ccbh::join_pairs(f1, f2, text, e);
    if (rc) { if (err && errcap) { strncpy(err, e.c_str(), errcap - 1); err[errcap - 1] = 0; } return rc; }
    *out = (char*)malloc(text.size() ? text.size() : 1);
    memcpy(*out, text.data(), text.size());
    *len = text.size();
    return 0;
}
extern "C" int ccb_host_csv(const ccb_reads* r, size_t first, size_t count, int k, int paired, int extended, int with_header,
                            const char* const* names, int n_names, const uint16_t* final5, const uint16_t* rows, int row_stride,
                            char** out, size_t* len)
{
    if (!r || first + count > r->rs.n() || n_names < 1) return -1;
    std::vector<std::string> nm(names, names + n_names);
    std::string s;
    if (with_header) ccbh::csv_header(s, extended != 0, nm);
    ccbh::csv_rows(s, r->rs, first, count, k, paired != 0, extended != 0, nm, final5, rows, row_stride);
    *out = (char*)malloc(s.size() ? s.size() : 1);
    memcpy(*out, s.data(), s.size());
    *len = s.size();
    return 0;
}
extern "C" void ccb_host_free(void* p) { free(p); }

extern "C" int ccb_host_numa_probe(const char* pci_bus_id, int* cpus_while_bound, int* cpus_after)
{
    int node;
    cpu_set_t m;
    {
        ccbh::NodeBind b(pci_bus_id);
        node = b.node();
        if (cpus_while_bound) *cpus_while_bound = sched_getaffinity(0, sizeof m, &m) == 0 ? CPU_COUNT(&m) : -1;
    }
    if (cpus_after) *cpus_after = sched_getaffinity(0, sizeof m, &m) == 0 ? CPU_COUNT(&m) : -1;
    return node;
}

extern "C" int ccb_host_numa_node_cpus(int node)
{
    cpu_set_t m;
    return ccbh::node_cpus(node, &m) ? CPU_COUNT(&m) : -1;
}
