/*
 * ref_cpu_device.cc -- CPU stand-in for the reference's device layer
 * (test infrastructure, NOT product code).
 *
 * The reference's GPU binaries need a B200 (and, there, a number of targets that is a
 * multiple of 32: tests/golden/ref_gpu/README.md); their CSVs are the ref_gpu goldens.
 * For everything that can be pinned WITHOUT a GPU this file defines the member functions of the reference's `class CuClarkDB<HKMERr>`
 * -- declared in the reference's own, unmodified src/CuClarkDB.cuh -- on the CPU,
 * so that the reference's UNMODIFIED host engine (src/main.cc, src/CuCLARK_hh.hh,
 * src/file.cc, ...: flag parsing, targets, DB build/load decision, read indexing,
 * 2-bit packing, paired-end merge, CSV printing) links and runs anywhere:
 *
 *   _ref/cuCLARK_cpu, _ref/cuCLARK-l_cpu  =  reference host engine
 *                                          + reference EHashtable::Read / queryElement (lookup)
 *                                          + the oracle's restatement of the three kernels
 *                                            (k-mer assembly, counting, result; cuclark_oracle.c)
 *
 * Rows A1, A2, A4, A5, A7, A8, A9, A10 of SURVEY.md section 8 are then the
 * reference's own code; only A3 and A6 (device kernels) are restated.
 * tests/golden/make_golden.py csv runs these binaries to produce
 * tests/golden/ref_csv/.
 */
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

/* The reference's implementation headers carry no include guards and define non-inline functions,
 * so its main.cc (and through it CuCLARK_hh.hh, HashTableStorage_hh.hh, CuClarkDB.cuh) is compiled
 * as part of this translation unit -- from where it lies, unmodified. */
#include "main.cc"                  /* reference, via -I */
#include "cuclark_oracle.h"

namespace {

template <typename HKMERr>
struct State {
    EHashtable<HKMERr, rElement>* ht = nullptr;
    std::vector<ITYPE> index;                 /* first read of each batch */
    RESULTS* full = nullptr; size_t row = 0;
    RESULTS* fin = nullptr;
    std::vector<uint32_t*> off;
    std::vector<CONTAINER*> cont;
    std::vector<size_t> n_reads, n_cont;
    bool extended = false;
};

template <typename HKMERr>
std::map<const void*, State<HKMERr>*>& states()
{
    static std::map<const void*, State<HKMERr>*> s;
    return s;
}

template <typename HKMERr>
int find_cb(void* ctx, uint64_t x, uint16_t* label)
{
    EHashtable<HKMERr, rElement>* ht = (EHashtable<HKMERr, rElement>*)ctx;
    ILBL l = 0;
    if (ht->queryElement(x, l)) { *label = l; return 1; }
    return 0;
}

}  // namespace

template <typename HKMERr>
CuClarkDB<HKMERr>::CuClarkDB(const size_t, const uint8_t _k, const size_t _numBatches, const size_t _numTargets, bool _verbose)
    : m_k(_k), m_numTargets(_numTargets), m_numBatches(_numBatches), m_verbose(_verbose), d_resultsFinal(nullptr)
{
    m_numDevices = 0;
    State<HKMERr>* s = new State<HKMERr>();
    s->off.resize(_numBatches, nullptr); s->cont.resize(_numBatches, nullptr);
    s->n_reads.resize(_numBatches, 0); s->n_cont.resize(_numBatches, 0);
    states<HKMERr>()[this] = s;
}

template <typename HKMERr>
CuClarkDB<HKMERr>::~CuClarkDB()
{
    freeBatchMemory();
    State<HKMERr>* s = states<HKMERr>()[this];
    if (s) { delete s->ht; delete s; }
    states<HKMERr>().erase(this);
}

template <typename HKMERr>
void CuClarkDB<HKMERr>::freeBatchMemory()
{
    State<HKMERr>* s = states<HKMERr>()[this];
    if (!s) return;
    for (size_t b = 0; b < s->off.size(); b++) { free(s->off[b]); free(s->cont[b]); s->off[b] = nullptr; s->cont[b] = nullptr; }
    free(s->full); free(s->fin); s->full = s->fin = nullptr;
}

template <typename HKMERr>
bool CuClarkDB<HKMERr>::read(const char* _filename, size_t& _fileSize, size_t& _dbParts, const ITYPE& _modCollision)
{
    State<HKMERr>* s = states<HKMERr>()[this];
    std::vector<std::string> labels(m_numTargets, "x"), labels_c;
    s->ht = new EHashtable<HKMERr, rElement>(m_k, labels, labels_c);
    _dbParts = 1;
    return s->ht->Read(_filename, _fileSize, 1, _modCollision, false);
}

template <typename HKMERr>
size_t CuClarkDB<HKMERr>::malloc(size_t _numReads, size_t _maxReads, size_t _maxReadsInContainers,
                                 std::vector<ITYPE>& _indexBatches, RESULTS*& _fullResults, size_t _resultRowSize,
                                 RESULTS*& _finalResults, size_t _finalResultsRowSize, bool _isExtended,
                                 std::vector<uint32_t*>& _readsPointer, std::vector<CONTAINER*>& _readsInCon)
{
    State<HKMERr>* s = states<HKMERr>()[this];
    freeBatchMemory();
    s->index = _indexBatches; s->row = _resultRowSize; s->extended = _isExtended;
    s->full = (RESULTS*)calloc(_numReads * _resultRowSize + 1, sizeof(RESULTS));
    s->fin = (RESULTS*)calloc(_numReads * _finalResultsRowSize + 1, sizeof(RESULTS));
    _fullResults = s->full; _finalResults = s->fin;
    _readsPointer.resize(m_numBatches); _readsInCon.resize(m_numBatches);
    for (size_t b = 0; b < m_numBatches; b++) {
        s->off[b] = (uint32_t*)calloc(_maxReads + 1, sizeof(uint32_t));
        s->cont[b] = (CONTAINER*)calloc(_maxReadsInContainers + 1, sizeof(CONTAINER));
        _readsPointer[b] = s->off[b]; _readsInCon[b] = s->cont[b];
    }
    return 0;
}

template <typename HKMERr> bool CuClarkDB<HKMERr>::sync() { return true; }
template <typename HKMERr> bool CuClarkDB<HKMERr>::waitForBatch(size_t) { return true; }
template <typename HKMERr> bool CuClarkDB<HKMERr>::checkBatch(size_t) { return true; }
template <typename HKMERr> bool CuClarkDB<HKMERr>::swapDbParts() { return false; }

template <typename HKMERr>
bool CuClarkDB<HKMERr>::readyBatch(const size_t _batchId, const size_t _numReads, const size_t _containerCount)
{
    State<HKMERr>* s = states<HKMERr>()[this];
    s->n_reads[_batchId] = _numReads; s->n_cont[_batchId] = _containerCount;
    return true;
}

template <typename HKMERr>
bool CuClarkDB<HKMERr>::queryBatch(const size_t _batchId, const bool _isExtended, const bool)
{
    State<HKMERr>* s = states<HKMERr>()[this];
    const size_t first = s->index[_batchId];
    ccbo_query_packed_cb(find_cb<HKMERr>, s->ht, m_k, s->off[_batchId], s->cont[_batchId], s->n_reads[_batchId],
                         s->fin + 5 * first, _isExtended ? s->full + s->row * first : nullptr, (int)s->row, nullptr, 1);
    return true;
}

template class CuClarkDB<T16>;
template class CuClarkDB<T32>;
template class CuClarkDB<T64>;
