// hostpipe.hpp -- host side of the classification path: read indexing, 2-bit
// packing, in-memory paired-end join, CSV formatting and the batch pipeline
// that feeds the C ABI (include/cuclark_b200.h).
//
// Replaces the host engine pieces of the reference that sit on the hot path
// (src/CuCLARK_hh.hh:1335-1783 index/pack/enqueue, :1945-2122 print,
// src/file.cc:205-268 pair join).  Rules that decide results are the
// reference's; the organisation is not: indexing is parallel over byte ranges,
// batches are packed by a pool of workers into library-owned pinned slots, the
// CSV text of a batch is formatted by the worker that packed it and appended in
// input order, paired files are joined in memory instead of through a
// temporary FASTA file.
#pragma once
#include <stddef.h>
#include <stdint.h>
#include <stdio.h>
#include <memory>
#include <string>
#include <vector>

struct ccb_handle;

namespace ccbh {

// std::vector whose resize() leaves new elements uninitialised: the index of 10 M reads is 400 MB, filling it
// with zeros on one thread before the parallel copy costs as much as the copy
template <typename T>
struct NoInitAlloc : std::allocator<T> {
    template <typename U> struct rebind { using other = NoInitAlloc<U>; };
    NoInitAlloc() = default;
    template <typename U> NoInitAlloc(const NoInitAlloc<U>&) {}
    template <typename U> void construct(U* p) noexcept { ::new ((void*)p) U; }
    template <typename U, typename... A> void construct(U* p, A&&... a) { ::new ((void*)p) U(std::forward<A>(a)...); }
};
using OffsetVec = std::vector<uint64_t, NoInitAlloc<uint64_t>>;
using TextBuf = std::vector<char, NoInitAlloc<char>>;      // the joined text of 10 M pairs is 3.3 GB: not to be zero-filled first

// One entry per read, byte offsets into the mapped text (A1, src/CuCLARK_hh.hh:1364-1394, 1496-1524)
struct ReadSet {
    const uint8_t* map = nullptr;
    size_t nb = 0;
    bool fasta = true;
    OffsetVec name_s, name_e;   // name bytes [name_s, name_e)
    OffsetVec seq_s, seq_e;     // sequence bytes [seq_s, seq_e) (may contain newlines)
    OffsetVec len;              // length used for gamma
    size_t n() const { return len.size(); }
    void clear() { name_s.clear(); name_e.clear(); seq_s.clear(); seq_e.clear(); len.clear(); }
};

// 0 = ok, -1 = neither '>' nor '@' first ("Failed to recognize the format of the file.")
int index_reads(const uint8_t* map, size_t nb, int threads, ReadSet& out);

// FASTQ only: the index the streaming pipeline arrives at when the text is cut into chunks of `chunk` bytes, each chunk
// finding its first record by itself (hostpipe.cpp fastq_chunk_walk).  0 = the pieces fit together, out = the records
// (equal to index_reads' by construction); 1 = not streamable (the pipeline gives such a file to the two-phase path).
int index_fastq_chunked(const uint8_t* map, size_t nb, size_t chunk, ReadSet& out);

// upper bound on u16 containers reads [first, first+count) can occupy
size_t pack_bound(const ReadSet& r, size_t first, size_t count, int k);
// A2 (src/CuCLARK_hh.hh:1615-1715): returns the number of containers written, -1 if cap is too small.
// Parts longer than 65528 bases are split with k-1 bases of overlap (the reference's u16 part
// length wraps there); everything else follows the reference's rules.
int64_t pack_reads(const ReadSet& r, size_t first, size_t count, int k,
                   uint32_t* offsets, uint16_t* containers, size_t cap);

// A10 (src/file.cc:205-268): ">id\n<mate1>N<mate2>\n" per pair, in memory.  0 = ok, otherwise err is set.
int join_pairs(const char* file1, const char* file2, TextBuf& out, std::string& err, int threads = 0);
// the same on two open streams, line by line as the reference reads them; both are closed
int join_pairs_streams(FILE* a, FILE* b, TextBuf& out, std::string& err);
// the same on two texts in memory (mapped files, or texts inflated from .gz).  Files that are what they should be -- the
// same number of lines, a multiple of four, every fourth line of both starting with '@' -- are joined by `threads` workers
// on ranges of records (0 = one per hardware thread); anything else goes line by line through join_pairs_streams, so
// every result and every error message is the sequential join's.
int join_pairs_texts(const uint8_t* a, size_t na, const uint8_t* b, size_t nb, TextBuf& out, std::string& err, int threads = 0);

// A7 (src/CuCLARK_hh.hh:1952-1967, 2096-2118)
void csv_header(std::string& out, bool extended, const std::vector<std::string>& target_names);
void csv_rows(std::string& out, const ReadSet& r, size_t first, size_t count, int k, bool paired, bool extended,
              const std::vector<std::string>& target_names, const uint16_t* final5, const uint16_t* rows, int row_stride);

struct ClassifyOptions {
    int  k = 31;
    int  threads = 1;
    bool paired = false;
    bool extended = false;
    bool verbose = false;
    int  row_stride = 32;
    const char* dump_final5 = nullptr;     // optional: raw u16[5] per read (second best is not in the CSV)
};

struct ClassifyStats {
    uint64_t reads = 0, batches = 0;
    double   seconds_index = 0, seconds_total = 0;
    // where the wall time of one file went: slot allocation, and summed over the workers:
    // packing, waiting for the device, formatting rows, waiting for the ordered writer
    double   seconds_alloc = 0, cpu_pack = 0, cpu_wait = 0, cpu_format = 0, cpu_sink = 0, cpu_index = 0;
};

// Allocate every handle's pinned slots for the largest batch classify_text forms (0.1-0.25 s of pinning): a caller
// that knows a large input is coming does this while the database loads; classify_text then finds them in place.
int reserve_slots(const std::vector<ccb_handle*>& devs, int k, bool extended);

// Classify the reads of one mapped text with the given device handles (one per GPU, replicated
// table; batches are dealt round-robin) and write <result_csv>.  0 = ok; err set otherwise.
int classify_text(const std::vector<ccb_handle*>& devs, const uint8_t* map, size_t nb, const char* result_csv,
                  const std::vector<std::string>& target_names, const ClassifyOptions& opt,
                  ClassifyStats& stats, std::string& err);

}  // namespace ccbh
