/*
 * cuclark_oracle.h -- CPU ORACLE (test infrastructure, NOT product code).
 *
 * A plain-C restatement of the cuCLARK / cuCLARK-l classification hot path of
 * ardaicoz/jn_cuclark.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may call into this library; the product (jn_cuclark_b200/) never does.
 *
 * Every function cites the reference file:line it restates (paths relative to
 * the reference checkout, e.g. src/CuClarkDB.cu:1190).
 *
 * Parity pin: see oracle/README.md.  The database half (A8/A9) is pinned
 * byte-for-byte against databases written by the unmodified reference binary
 * (oracle/_ref/cuCLARK, cuCLARK-l); the lookup (A4/A5) against the reference's
 * own hTable::find via oracle/_ref/ref_harness; parse/pack/pair-merge/print
 * (A1, A2, A7, A10) byte-for-byte against CSVs written by the reference's
 * unmodified host engine (tests/golden/ref_csv/, oracle/ref_cpu_device.cc).
 * The device kernels' k-mer assembly, counting and result (A3, A6) are pinned
 * byte-for-byte against CSVs the reference's own CUDA kernels wrote on a B200
 * (tests/golden/ref_gpu/w32_*.csv; see the README there).
 */
#ifndef CUCLARK_ORACLE_H
#define CUCLARK_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* src/parameters.hh:37-48, src/parameters_light_hh:37-49 */
#define CCBO_HTSIZE_FULL   1610612741ULL
#define CCBO_HTSIZE_LIGHT  57777779ULL
#define CCBO_MAXHITS_FULL  15
#define CCBO_MAXHITS_LIGHT 23
#define CCBO_OBJECTNAMEMAX 40
#define CCBO_NBN           1

/* ---- A3/A4: k-mer encoding ------------------------------------------- */
/* src/CuCLARK_hh.hh:286-300 (m_table / m_rTable).  Returns the x-encoding
 * code A=3 C=2 G=1 T/U=0 (either case), -10 for '\n', -2 for '>', -1 else. */
int      ccbo_rcode(unsigned char c);
/* src/CuClarkDB.cu:1195-1203 == src/hashTable_hh.hh:362-369 */
uint64_t ccbo_revcomp(uint64_t x, int k);
/* src/CuClarkDB.cu:1206 */
uint64_t ccbo_canonical(uint64_t x, int k);
/* key width picked by src/main.cc:251-286 */
int      ccbo_key_bytes(int k, uint64_t htsize);

/* ---- A8/A9: database -------------------------------------------------- */
typedef struct {
    uint64_t  htsize;      /* number of buckets                              */
    int       key_bytes;   /* 2, 4 or 8                                      */
    int       k;
    uint64_t  n;           /* number of stored k-mers                        */
    uint8_t  *sz;          /* [htsize] bucket sizes (.sz)                    */
    uint64_t *keys;        /* [n] quotient c / htsize, bucket order (.ky)    */
    uint16_t *labels;      /* [n] 0-based target index (.lb)                 */
    uint32_t *ptr;         /* [htsize+1] exclusive prefix of sz, built lazily  */
} ccbo_db;

typedef struct {
    int       n_files;
    char    **files;       /* [n_files] FASTA paths                          */
    uint16_t *file_label;  /* [n_files] index into labels                    */
    int       n_labels;
    char    **labels;      /* distinct labels, first-appearance order        */
} ccbo_targets;

/* src/CuCLARK_hh.hh:1790-1878 (getTargetsData): parse "<file> <label>" lines */
int  ccbo_targets_read(const char *path, ccbo_targets *t);
void ccbo_targets_free(ccbo_targets *t);

/* src/CuCLARK_hh.hh:581-592 (getdbName) */
void ccbo_db_name(char *out, size_t cap, const char *folder, int k, int n_labels,
                  uint64_t htsize, unsigned min_t, int light, unsigned gap);

/* Collect every (canonical k-mer, label) occurrence of one FASTA text.
 * full : src/CuCLARK_hh.hh:1129-1198 (every overlapping k-mer)
 * light: src/CuCLARK_hh.hh:711-767   (non-overlapping, every gap-th; *iter is
 *        the per-file counter of line 709 and must start at 0 per file)
 * Appends to *out (realloc'd), returns 0 on success. */
typedef struct { uint64_t c; uint16_t label; } ccbo_occ;
typedef struct { ccbo_occ *v; size_t n, cap; } ccbo_occ_vec;
int ccbo_extract_fasta(const unsigned char *txt, size_t len, int k, int light,
                       unsigned gap, uint16_t label, ccbo_occ_vec *out);

/* src/HashTableStorage_hh.hh:422-461 (addElement) + :230-280 (RemoveCommon) +
 * src/hashTable_hh.hh:399-410 (updateElement): keep c iff every occurrence
 * carries the same label and the saturating occurrence count is > min_t.
 * Then lay out like src/hashTable_hh.hh:474-546 (write).  Consumes occ. */
int  ccbo_db_from_occ(ccbo_occ_vec *occ, int k, uint64_t htsize, unsigned min_t,
                      ccbo_db *db);
int  ccbo_db_build(const ccbo_targets *t, int k, uint64_t htsize, int light,
                   unsigned gap, unsigned min_t, ccbo_db *db);
/* src/hashTable_hh.hh:474-546 */
int  ccbo_db_write(const ccbo_db *db, const char *prefix);
/* src/CuClarkDB.cu:464-741 / src/hashTable_hh.hh:549-829; sampling = -s */
int  ccbo_db_read(const char *prefix, uint64_t htsize, int key_bytes, int k,
                  unsigned sampling, ccbo_db *db);
void ccbo_db_free(ccbo_db *db);
void ccbo_db_index(ccbo_db *db);   /* build ptr (idempotent; call before threaded finds) */
/* src/hashTable_hh.hh:359-396 == src/CuClarkDB.cu:1190-1254; x = x-encoded k-mer */
int  ccbo_db_find(const ccbo_db *db, uint64_t x, uint16_t *label);
/* canonical value of stored entry i in bucket r: key*htsize + r (A9) */

/* ---- A1: read indexing ------------------------------------------------ */
typedef struct {
    size_t    n;
    uint64_t *name_s, *name_e;   /* [n] name byte range in the file          */
    uint64_t *seq_s,  *seq_e;    /* [n] sequence byte range (A1)             */
    uint64_t *len;               /* [n] length used for gamma                */
    int       is_fasta;
} ccbo_reads;
/* src/CuCLARK_hh.hh:1340-1534, restated for ONE batch covering the file
 * (-b 1); see oracle/README.md for why multi-batch splitting is not part of
 * the result contract.  Returns 0, or -1 if the format is not recognised. */
int  ccbo_index_reads(const unsigned char *map, size_t nb, ccbo_reads *r);
void ccbo_reads_free(ccbo_reads *r);

/* ---- A2: 2-bit packing ------------------------------------------------ */
/* src/CuCLARK_hh.hh:1615-1715.  offsets[n+1] (u32, units of u16),
 * containers: caller-provided capacity cap; returns #containers or -1. */
int64_t ccbo_pack(const unsigned char *map, const ccbo_reads *r, size_t first,
                  size_t count, int k, uint32_t *offsets, uint16_t *containers,
                  size_t cap);

/* ---- A3-A6: query + vote over a packed batch --------------------------- */
/* src/CuClarkDB.cu:1042-1117 (k-mers), :1190-1254 (probe), :1361-1411 (vote).
 * out5[n*5] = {sumN, idxBest, best, idxSecond, second}.
 * rows (optional, may be NULL): row_stride u16 per read = {n,(t,c)*n} with the
 * targets ascending (src/CuClarkDB.cu:1122-1182); rows longer than the stride
 * are truncated (reference behaviour there is undefined).
 * n_lookups (optional) receives the number of k-mers probed.
 * threads<=1 : scalar loop; else OpenMP over reads. */
int ccbo_query_packed(const ccbo_db *db, int k, const uint32_t *offsets,
                      const uint16_t *containers, size_t n_reads, int n_targets,
                      uint16_t *out5, uint16_t *rows, int row_stride,
                      uint64_t *n_lookups, int threads);

/* Same walk/vote with a caller-supplied probe (x = x-encoded k-mer, returns 1
 * on hit).  oracle/ref_harness.cc plugs the reference's own hTable::find in. */
typedef int (*ccbo_find_fn)(void *ctx, uint64_t x, uint16_t *label);
int ccbo_query_packed_cb(ccbo_find_fn find, void *ctx, int k, const uint32_t *offsets,
                         const uint16_t *containers, size_t n_reads,
                         uint16_t *out5, uint16_t *rows, int row_stride,
                         uint64_t *n_lookups, int threads);

/* ---- A7: CSV ---------------------------------------------------------- */
/* src/CuCLARK_hh.hh:1945-2122.  names[0]="NA", names[1..n_targets]. rows may
 * be NULL unless extended. Appends header + one line per read to path. */
int ccbo_write_csv(const char *path, const unsigned char *map, const ccbo_reads *r,
                   int k, int paired, int extended, int n_targets, char **names,
                   const uint16_t *out5, const uint16_t *rows, int row_stride);

/* ---- A10: paired-end join -------------------------------------------- */
/* src/file.cc:205-268 */
int ccbo_merge_paired(const char *f1, const char *f2, const char *out);

/* ---- synthetic benchmark workload (include/ccb_synth.h) ---------------- */
int ccbo_synth_db(uint64_t db_seed, uint32_t n_genomes, uint64_t genome_len, uint32_t n_labels,
                  int k, uint64_t htsize, ccbo_db *db);
int ccbo_synth_db_ex(uint64_t db_seed, uint32_t n_genomes, uint64_t genome_len, uint32_t n_labels,
                     int k, uint64_t htsize, int light, unsigned gap, ccbo_db *db);
struct ccb_workload;
int ccbo_synth2_reads_fasta(const char *path, const struct ccb_workload *w, uint64_t reads_seed, uint64_t db_seed, uint64_t first,
                            uint32_t n, uint32_t n_genomes, uint64_t genome_len);
int ccbo_synth_reads_fasta(const char *path, uint64_t reads_seed, uint64_t db_seed, uint64_t first,
                           uint32_t n, uint32_t read_len, uint32_t n_genomes, uint64_t genome_len);

#ifdef __cplusplus
}
#endif
#endif
