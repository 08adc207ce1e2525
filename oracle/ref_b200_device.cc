/*
 * ref_b200_device.cc -- the reference's UNMODIFIED host engine on top of libcuclark_b200.so
 * (test infrastructure: the executed proof of the in-process seam).
 *
 * The reference reaches the GPU only through `class CuClarkDB<HKMERr>` (src/CuClarkDB.cuh:99-151).
 * This file is the replacement a maintainer would drop in for src/CuClarkDB.cu:
 * every member function is a call of the C ABI (include/cuclark_b200.h).  The class declaration stays the
 * reference's own header; main.cc, CuCLARK_hh.hh, file.cc ... are compiled from /root/reference/src where
 * they lie, unmodified (the same recipe as ref_cpu_device.cc), so
 *
 *   _ref/cuCLARK_b200, _ref/cuCLARK-l_b200  =  reference host engine (flags, targets, read indexing, packing,
 *                                               pair merge, CSV printing)  +  this shim  +  libcuclark_b200.so
 *
 * tests/test_ref_shim_gpu.py runs them on a B200 and compares their CSVs byte for byte with the goldens the
 * reference's own device code produced (tests/golden/ref_csv, ref_gpu).
 *
 * The class has no spare member for the handle, so the per-object state lives in a side table keyed by `this`.
 */
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "main.cc"                  /* reference, via -I; brings CuClarkDB.cuh, parameters.hh (HTSIZE, MAXHITS) */
#include "cuclark_b200.h"

namespace {

struct ShimState {
    ccb_handle* h = nullptr;
    std::vector<size_t> n_reads, n_cont;
    bool swapped = false;
};

std::mutex g_mu;
std::map<const void*, ShimState*>& table()
{
    static std::map<const void*, ShimState*> t;
    return t;
}
ShimState* state_of(const void* self)
{
    std::lock_guard<std::mutex> g(g_mu);
    return table()[self];
}

/* CUERR (src/CuClarkDB.cu:45-53): print and exit(1) -- the library itself never exits */
void fail(ccb_handle* h)
{
    std::cerr << ccb_last_error(h) << std::endl;
    exit(1);
}

}  // namespace

template <typename HKMERr>
CuClarkDB<HKMERr>::CuClarkDB(const size_t _numDevices, const uint8_t _k, const size_t _numBatches, const size_t _numTargets, bool _verbose)
    : m_k(_k), m_numTargets(_numTargets), m_numBatches(_numBatches), m_verbose(_verbose), d_resultsFinal(nullptr)
{   /* src/CuClarkDB.cu:95-241 */
    m_numDevices = (int)_numDevices;
    ShimState* s = new ShimState();
    ccb_config c;
    memset(&c, 0, sizeof c);
    c.device = 0; c.k = _k; c.n_targets = (int)_numTargets; c.max_hits = MAXHITS; c.n_slots = (int)_numBatches; c.verbose = _verbose ? 1 : 0;
    if (ccb_create(&c, &s->h)) fail(nullptr);
    s->n_reads.resize(_numBatches, 0); s->n_cont.resize(_numBatches, 0);
    std::lock_guard<std::mutex> g(g_mu);
    table()[this] = s;
}

template <typename HKMERr>
CuClarkDB<HKMERr>::~CuClarkDB()
{   /* :248-279 */
    ShimState* s = state_of(this);
    if (s) { ccb_destroy(s->h); delete s; }
    std::lock_guard<std::mutex> g(g_mu);
    table().erase(this);
}

template <typename HKMERr>
void CuClarkDB<HKMERr>::freeBatchMemory()
{   /* :285-316 */
    ShimState* s = state_of(this);
    if (s && s->h) ccb_batch_free(s->h);
}

template <typename HKMERr>
bool CuClarkDB<HKMERr>::read(const char* _filename, size_t& _fileSize, size_t& _dbParts, const ITYPE& _modCollision)
{   /* :464-770 -- the whole table goes to the device here, so there are no parts to swap later */
    ShimState* s = state_of(this);
    uint64_t n = 0;
    if (ccb_db_load_files(s->h, _filename, HTSIZE, (int)sizeof(HKMERr), (unsigned)_modCollision, &n)) {
        std::cerr << ccb_last_error(s->h) << std::endl;
        return false;
    }
    _fileSize = n; _dbParts = 1;
    m_dbParts = 1;
    return true;
}

template <typename HKMERr>
size_t CuClarkDB<HKMERr>::malloc(size_t _numReads, size_t _maxReads, size_t _maxReadsInContainers,
                                 std::vector<ITYPE>& _indexBatches, RESULTS*& _fullResults, size_t _resultRowSize,
                                 RESULTS*& _finalResults, size_t _finalResultsRowSize, bool _isExtended,
                                 std::vector<uint32_t*>& _readsPointer, std::vector<CONTAINER*>& _readsInCon)
{   /* :322-421 -- the library owns the pinned buffers; ONE final / full result array indexed by global read number */
    ShimState* s = state_of(this);
    if (_resultRowSize != (size_t)(2 * MAXHITS + 2) || _finalResultsRowSize != 5) {
        std::cerr << "ref_b200_device: unexpected row sizes " << _resultRowSize << " / " << _finalResultsRowSize << std::endl;
        exit(1);
    }
    std::vector<uint64_t> index(_indexBatches.begin(), _indexBatches.end());
    index.resize(m_numBatches, (uint64_t)_numReads);
    uint16_t* full = nullptr; uint16_t* fin = nullptr;
    if (ccb_batch_alloc_ref(s->h, _numReads, (uint32_t)_maxReads, _maxReadsInContainers, index.data(), (int)m_numBatches,
                            _isExtended ? 1 : 0, &full, &fin)) fail(s->h);
    _fullResults = (RESULTS*)full; _finalResults = (RESULTS*)fin;
    _readsPointer.resize(m_numBatches); _readsInCon.resize(m_numBatches);
    for (size_t b = 0; b < m_numBatches; b++) {
        uint32_t* off = nullptr; uint16_t* cont = nullptr;
        if (ccb_batch_buffers(s->h, (int)b, &off, &cont, nullptr, nullptr)) fail(s->h);
        _readsPointer[b] = off; _readsInCon[b] = (CONTAINER*)cont;
    }
    return 0;
}

template <typename HKMERr> bool CuClarkDB<HKMERr>::sync()
{   /* :427-436 */
    ShimState* s = state_of(this);
    if (ccb_sync(s->h)) fail(s->h);
    return true;
}

template <typename HKMERr> bool CuClarkDB<HKMERr>::waitForBatch(size_t i)
{   /* :442-446 */
    ShimState* s = state_of(this);
    if (ccb_batch_wait(s->h, (int)i)) fail(s->h);
    return true;
}

template <typename HKMERr> bool CuClarkDB<HKMERr>::checkBatch(size_t i)
{   /* :450-456 */
    ShimState* s = state_of(this);
    const int r = ccb_batch_poll(s->h, (int)i);
    if (r < 0) fail(s->h);
    return r == 0;
}

/* :776-815 -- called once before the first query (CuCLARK_hh.hh:516) and then until it returns false (:1765);
 * with the whole table resident there is never a second part */
template <typename HKMERr> bool CuClarkDB<HKMERr>::swapDbParts() { return false; }

template <typename HKMERr>
bool CuClarkDB<HKMERr>::readyBatch(const size_t _batchId, const size_t _numReads, const size_t _containerCount)
{   /* :821-828 -- called concurrently from the OpenMP workers on distinct ids */
    ShimState* s = state_of(this);
    s->n_reads[_batchId] = _numReads; s->n_cont[_batchId] = _containerCount;
    return true;
}

template <typename HKMERr>
bool CuClarkDB<HKMERr>::queryBatch(const size_t _batchId, const bool _isExtended, const bool)
{   /* :836-987 -- returns true when the final result of the batch has been scheduled */
    ShimState* s = state_of(this);
    if (ccb_batch_submit(s->h, (int)_batchId, (uint32_t)s->n_reads[_batchId], s->n_cont[_batchId], _isExtended ? CCB_EXTENDED : 0)) fail(s->h);
    return true;
}

template class CuClarkDB<T16>;
template class CuClarkDB<T32>;
template class CuClarkDB<T64>;
