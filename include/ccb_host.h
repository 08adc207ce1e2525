/*
 * ccb_host.h -- C ABI of the host side of the classification path (libcuclark_b200.so):
 * read indexing, 2-bit packing, paired-end join and CSV formatting as used by
 * bin/cuCLARK[-l].  Exposed so that other hosts (and the CPU-only tests) can drive
 * the same code.  Each entry names the reference code it replaces.
 */
#ifndef CCB_HOST_H
#define CCB_HOST_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct ccb_reads ccb_reads;

/* Read indexing, src/CuCLARK_hh.hh:1340-1534.  map must stay valid while the index is used.
 * NULL when the text starts with neither '>' nor '@'. */
ccb_reads* ccb_host_index(const uint8_t* map, size_t nb, int threads);
void       ccb_host_reads_free(ccb_reads* r);
size_t     ccb_host_reads_count(const ccb_reads* r);
int        ccb_host_reads_is_fasta(const ccb_reads* r);
/* field: 0 name_s, 1 name_e, 2 seq_s, 3 seq_e, 4 len; array of ccb_host_reads_count u64 */
const uint64_t* ccb_host_reads_field(const ccb_reads* r, int field);

/* FASTQ only: the index the streaming pipeline of bin/cuCLARK arrives at when it cuts the text into chunks of `chunk` bytes and
 * lets every chunk find its first record by itself (no counterpart in the reference, which indexes sequentially,
 * src/CuCLARK_hh.hh:1496-1524; the result must equal ccb_host_index).  *out = the records when the pieces fit together
 * (return 0); return 1 = not streamable (the pipeline hands such a file to its two-phase path), -1 = not FASTQ. */
int ccb_host_index_fastq_chunked(const uint8_t* map, size_t nb, size_t chunk, ccb_reads** out);

/* 2-bit packing, src/CuCLARK_hh.hh:1615-1715.  Returns #containers, -1 cap too small, -2 > 2^32. */
size_t  ccb_host_pack_bound(const ccb_reads* r, size_t first, size_t count, int k);
int64_t ccb_host_pack(const ccb_reads* r, size_t first, size_t count, int k,
                      uint32_t* offsets, uint16_t* containers, size_t cap);

/* Paired-end join, src/file.cc:205-268, in memory.  *out is malloc'd (ccb_host_free). 0 = ok. */
int  ccb_host_join_pairs(const char* file1, const char* file2, char** out, size_t* len, char* err, size_t errcap);

/* CSV text, src/CuCLARK_hh.hh:1945-2122: header (with_header) + one row per read.
 * names[0] = "NA", names[1..n_names-1] = targets.  *out is malloc'd (ccb_host_free). */
int  ccb_host_csv(const ccb_reads* r, size_t first, size_t count, int k, int paired, int extended, int with_header,
                  const char* const* names, int n_names, const uint16_t* final5, const uint16_t* rows, int row_stride,
                  char** out, size_t* len);
void ccb_host_free(void* p);

/* Pinned batch buffers next to their GPU (csrc/hostnuma.hpp; the reference allocates wherever the thread runs,
 * src/CuClarkDB.cu:340-365).  Probe for tests and diagnostics: binds the calling thread to the memory node of the PCI device
 * `pci_bus_id` ("0000:1b:00.0") the way ccb_batch_alloc does around its cudaMallocHost calls, reports the node (-1 = nothing
 * to do: one node, unknown device, CCB_NUMA=0), the number of CPUs the thread may run on while bound and after the
 * binding was undone. */
int  ccb_host_numa_probe(const char* pci_bus_id, int* cpus_while_bound, int* cpus_after);
/* CPUs of memory node `node` (/sys/devices/system/node/node<N>/cpulist), -1 = no such node */
int  ccb_host_numa_node_cpus(int node);

#ifdef __cplusplus
}
#endif
#endif
