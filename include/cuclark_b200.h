/*
 * cuclark_b200.h -- C ABI of the B200-native cuCLARK / cuCLARK-l device layer.
 *
 * This is the drop-in seam for the reference's `class CuClarkDB<HKMERr>`
 * (reference src/CuClarkDB.cuh:99-151), the only thing its host engine
 * (src/CuCLARK_hh.hh) calls to reach the GPU.  Each entry point names the
 * reference method it replaces.  Differences by design:
 *   - opaque handle, plain pointers and sizes, int return codes (0 = ok,
 *     otherwise see ccb_last_error); the library never calls exit()
 *     (reference: CUERR -> exit(1), src/CuClarkDB.cu:45-53);
 *   - one handle drives ONE device; multi-GPU = one handle (thread or
 *     process) per GPU over a replicated table, reads partitioned by the caller;
 *   - the key width (reference template HKMERr) is a run-time argument.
 *
 * Batch data layout is the reference's (src/CuCLARK_hh.hh:1615-1715):
 *   offsets[n_reads+1]  u32, in units of u16 containers
 *   containers[]        per read: parts  [len:u16][ceil(len/8) x u16], 8 bases
 *                       per u16, first base in the top two bits, A=3 C=2 G=1 T=0
 * Result layout is the reference's (src/CuClarkDB.cu:1401-1405):
 *   final5[n_reads*5]   u16 {sumN, idxBest, best, idxSecond, second}; idx = target+1, 0 = none
 *   rows[n_reads*(2*max_hits+2)] u16 {n,(target,count)*n} targets ascending
 *                       (src/CuClarkDB.cu:1122-1182), only when extended
 */
#ifndef CUCLARK_B200_H
#define CUCLARK_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ccb_handle ccb_handle;

typedef struct ccb_config {
    int    device;        /* CUDA device ordinal                                           */
    int    k;             /* k-mer length, 2..32 (src/main.cc:106-110)                      */
    int    n_targets;     /* number of targets = labels 0..n_targets-1 (<= 65534)           */
    int    max_hits;      /* MAXHITS of the variant: 15 full / 23 light (parameters.hh:44)  */
    int    n_slots;       /* batch slots (reference: -b batches); >= 1, 2+ overlaps copies  */
    double load_factor;   /* table fill, 0 = default 0.5                                    */
    int    verbose;
} ccb_config;

typedef struct ccb_db_info {
    uint64_t entries;       /* occupied slots (incl. tombstones)        */
    uint64_t buckets;       /* home buckets                             */
    uint64_t bytes;         /* device bytes of the table                */
    uint64_t tombstones;    /* k-mers seen under more than one label    */
    uint32_t max_displacement;
} ccb_db_info;

typedef struct ccb_stats {
    uint64_t reads;         /* reads submitted                          */
    uint64_t lookups;       /* k-mers probed   (only with CCB_STATS)    */
    uint64_t sectors;       /* 32-byte table sectors read (CCB_STATS)   */
    uint64_t dense_reads;   /* reads re-done by the dense exact path    */
    uint64_t table_probes;  /* k-mers that went to the k-mer table (the rest were answered by the anchor index); CCB_STATS */
} ccb_stats;

enum { CCB_EXTENDED = 1, CCB_STATS = 2 /* lookups = k-mers answered, sectors = 32-byte sectors requested from the table, the
                                          anchor table and the target text together */,
       CCB_FORCE_DENSE = 4, CCB_NO_ANCHOR = 8,
       CCB_STATS_EX = 64 /* with CCB_STATS in ccb_query_device: d_stats holds 4 x u64 {lookups, sectors, table probes, 0} */,
       CCB_WARP_KERNEL = 32 /* warp-per-read kernel for every read (no block kernel, hence no anchor index) */ };

/* CuClarkDB::CuClarkDB (src/CuClarkDB.cu:95-241) / ~CuClarkDB (:248-279) */
int  ccb_create(const ccb_config* cfg, ccb_handle** out);
void ccb_destroy(ccb_handle* h);
/* message of the last failing call on this handle (or of ccb_create when h == NULL) */
const char* ccb_last_error(const ccb_handle* h);

/* CuClarkDB::read + swapDbParts (src/CuClarkDB.cu:464-770, 776-815): load
 * <prefix>.sz/.ky/.lb (reference on-disk format, src/hashTable_hh.hh:474-546)
 * into the device table.  sampling = -s factor (1 = all buckets). */
int  ccb_db_load_files(ccb_handle* h, const char* prefix, uint64_t htsize, int key_bytes,
                       unsigned sampling, uint64_t* n_loaded);
/* Same table from (canonical k-mer, label) pairs; a k-mer added under two
 * different labels becomes non-discriminative (HashTableStorage_hh.hh:244). */
int  ccb_db_begin(ccb_handle* h, uint64_t expected_entries);
int  ccb_db_add_pairs(ccb_handle* h, const uint64_t* canon, const uint16_t* labels, uint64_t n);
int  ccb_db_add_pairs_device(ccb_handle* h, const void* d_canon, const void* d_labels, uint64_t n);
int  ccb_db_end(ccb_handle* h);
int  ccb_db_get_info(const ccb_handle* h, ccb_db_info* info);

/* Database construction from target FASTA files -- replaces the reference's
 * single-threaded makeSpecificTargetSets (src/CuCLARK_hh.hh:691-1329: scan loops
 * :1129-1198 full, :711-767 light), EHashtable::addElement / RemoveCommon
 * (src/HashTableStorage_hh.hh:422-461, 230-280) and hTable::write
 * (src/hashTable_hh.hh:474-546).  Full mode: every overlapping k-mer of every
 * ACGTU run; light mode: non-overlapping k-mers, every gap-th of them per file.
 * A canonical k-mer is stored iff all its occurrences carry one label and its
 * occurrence count (saturating at 254) exceeds min_count.  ccb_build_end writes
 * <prefix>.sz/.ky/.lb byte-for-byte as the reference does (prefix may be NULL)
 * and, when load_table != 0, leaves the k-mer table ready for queries.
 * k comes from ccb_config.  max_bases bounds the bases all add_fasta calls
 * together may carry (the sum of the file sizes will do).  ccb_build_add_fasta
 * may be called from several host threads. */
typedef struct ccb_build_info {
    uint64_t bases;          /* ACGTU bases scanned */
    uint64_t occurrences;    /* k-mers emitted before de-duplication */
    uint64_t distinct;       /* distinct canonical k-mers */
    uint64_t stored;         /* discriminative k-mers written / loaded */
    uint32_t max_bucket;     /* largest reference bucket (must be <= 255) */
    uint32_t reserved;
} ccb_build_info;
int  ccb_build_begin(ccb_handle* h, uint64_t htsize, int key_bytes, int light, unsigned gap,
                     unsigned min_count, uint64_t max_bases);
int  ccb_build_add_fasta(ccb_handle* h, const unsigned char* text, uint64_t len, uint16_t label);
int  ccb_build_end(ccb_handle* h, const char* prefix, int load_table, ccb_build_info* info);

/* Sharded table -- a database larger than one device's memory.  The reference cuts the bucket
 * range into parts and swaps them through one GPU (CuClarkDB::read / swapDbParts,
 * src/CuClarkDB.cu:552-559, 776-815, queryKernel's range filter :1211-1214, mergeKernel :1261-1355).
 * Here shard g of n_shards (<= 8) lives on GPU g and a batch takes one exchange each way:
 *   ccb_shard_config  before the database is loaded/built/synthesised: keep only this shard's k-mers
 *   ccb_shard_count   d_pos = u32[n_shards*n_reads+1]: exclusive scan, laid out [owner][read], of the
 *                     number of k-mers each read sends to each owner ([n_shards*n_reads] = total)
 *   ccb_shard_route   read r's mixed k-mers (u64) for owner g go to dst[g][d_pos[g*n_reads+r]-seg_base[g] ...);
 *                     dst[g] is a device pointer this process may store to (a local send segment or a
 *                     peer GPU's inbox); seg_base and dst are HOST arrays of n_shards entries
 *   ccb_shard_probe   owner side: d_labels[i] (u16) = label of mixed k-mer d_hashes[i], 0xFFFF = none;
 *                     d_stats (optional) = 2 x u64 {lookups, sectors}, accumulated
 *   ccb_shard_vote    requester side: labels of (owner g, read r) are read at
 *                     src[g][d_pos[g*n_reads+r]-seg_base[g] ...); writes u16[5] per read (+ rows)
 * Results are identical to ccb_query_device on the unsharded table. */
int  ccb_shard_config(ccb_handle* h, int n_shards, int shard);
int  ccb_shard_count(ccb_handle* h, const void* d_offsets, const void* d_containers, uint32_t n_reads,
                     int n_shards, void* d_pos, void* stream);
int  ccb_shard_route(ccb_handle* h, const void* d_offsets, const void* d_containers, uint32_t n_reads,
                     int n_shards, const void* d_pos, const uint64_t* seg_base, void* const* dst, void* stream);
int  ccb_shard_probe(ccb_handle* h, const void* d_hashes, uint64_t n, void* d_labels, void* d_stats, void* stream);
int  ccb_shard_vote(ccb_handle* h, uint32_t n_reads, int n_shards, const void* d_pos, const uint64_t* seg_base,
                    const void* const* src, void* d_final5, void* d_rows, void* stream);
/* Peer exchange -- NVLink stores instead of a collective and no host round trip (route side: ccb_shard_route2 below):
 *   ccb_shard_probe_inbox  inbox = n_shards regions of cap k-mers, region s holds d_counts_in[s] (u64) k-mers
 *                          from rank s; the label of inbox[s][i] is stored at lab_dst[s][i] (rank s's memory)
 *   ccb_device_alloc / ccb_ipc_export / ccb_ipc_open / ccb_ipc_close
 *                          buffers another process may map (cudaIpc*; 64-byte handles)
 * A barrier across the ranks (any stream-ordered one, e.g. an NCCL all-reduce of one word) must separate
 * the route from probe_inbox and probe_inbox from the vote. */
int  ccb_shard_probe_inbox(ccb_handle* h, int n_shards, const void* d_inbox, uint64_t cap, const void* d_counts_in,
                           void* const* lab_dst, void* d_stats, void* stream);
/* One-pass pipeline (the product path of the sharded mode; the calls above remain for callers that want every
 * (read, owner) pair in its own segment).  A block of 256 reads shares one contiguous segment per owner, reserved
 * with one atomic on a cursor in the requester's memory, so the reads are walked once and no scan is needed; reads
 * the block kernel does not take (three or more parts, parts over 192 bases) follow behind through count/route.
 *   ccb_shardws_create   device scratch of ONE batch in flight: max_reads reads, n_shards regions of cap k-mers
 *   ccb_shard_route2     dst[g] = this requester's region of owner g (cap x u64: a local send segment or a peer's
 *                        inbox), count_dst[g] = where owner g reads how many k-mers it got (u64, owner's memory
 *                        or local); n_containers = u16 units of the batch (-1 unknown).  More than cap k-mers for
 *                        one owner: the excess is dropped and the workspace's error word set (ccb_shardws_error)
 *   ccb_shard_probe_inbox / ccb_shard_probe   owner side, unchanged: labels come back at the same indices
 *   ccb_shard_vote2      src[g] = labels of region g (cap x u16); writes u16[5] per read (+ rows)
 * The same stream-ordered barriers as above separate route2 from the probe and the probe from vote2. */
typedef struct ccb_shardws ccb_shardws;
int  ccb_shardws_create(ccb_handle* h, uint32_t max_reads, int n_shards, uint64_t cap, ccb_shardws** out);
void ccb_shardws_destroy(ccb_handle* h, ccb_shardws* ws);
int  ccb_shardws_error(ccb_handle* h, ccb_shardws* ws, void* stream, int* out);
int  ccb_shard_route2(ccb_handle* h, ccb_shardws* ws, const void* d_offsets, const void* d_containers, uint32_t n_reads,
                      int64_t n_containers, void* const* dst, void* const* count_dst, void* stream);
int  ccb_shard_vote2(ccb_handle* h, ccb_shardws* ws, uint32_t n_reads, const void* const* src, void* d_final5, void* d_rows,
                     void* stream);
int  ccb_device_alloc(ccb_handle* h, uint64_t bytes, void** ptr);
int  ccb_ipc_export(ccb_handle* h, void* ptr, unsigned char out[64]);
int  ccb_ipc_open(ccb_handle* h, const unsigned char in[64], void** ptr);
int  ccb_ipc_close(ccb_handle* h, void* ptr);

/* Device group: ONE process, n_devices GPUs, the table sharded over them (shard g on devices[g]) -- what
 * bin/cuCLARK --sharded (or a database that does not fit one GPU) uses in place of the reference's part swapping.
 * ccb_group_create returns a LEADER handle; on it ccb_db_load_files loads every shard, and ccb_batch_alloc /
 * ccb_batch_buffers / ccb_batch_submit / ccb_batch_wait / ccb_batch_poll / ccb_batch_free / ccb_sync / ccb_destroy work as
 * on a one-device handle: slot s is homed on devices[s mod n]; its k-mers travel to the owners' inboxes and the labels
 * back through peer mappings (NVLink), the stages chained by events, nothing synchronised on the host in between.
 * cfg->device is ignored.  The devices must be able to map each other's memory (a device may be listed more than
 * once: several shards then share it -- tests on a one-GPU box).  Results are identical to a
 * one-device handle holding the whole table.  A batch whose k-mers for one owner exceed the exchange region fails
 * in ccb_batch_wait with the reference's "increase the number of batches" message (nothing is read or written out of
 * bounds).  Not available on a leader: ccb_query_device, ccb_batch_alloc_ref, the anchor index, ccb_build_*. */
int  ccb_group_create(const int* devices, int n_devices, const ccb_config* cfg, ccb_handle** out);
int  ccb_group_size(const ccb_handle* h);                 /* 0 for an ordinary handle */
ccb_handle* ccb_group_member(ccb_handle* h, int i);       /* the handle that holds shard i (owned by the leader) */

/* Anchor index (no counterpart in the reference; results are unchanged by it).  The target sequences
 * themselves are kept on the device (both strands, 2 bits per base) with one bit per position saying
 * whether the k-mer starting there is in the table, and a small table of sampled k-mers -> position.
 * A read whose sampled k-mer is found is compared base by base with the sequence; every k-mer window
 * that matches is answered from the position bits without a table probe, every other window is probed
 * as usual -- so the answer is exact whatever the index contains.  Built after the table is loaded;
 * dropped when the table is replaced.  CCB_NO_ANCHOR in the query flags bypasses it.
 *   ccb_anchor_begin / ccb_anchor_add_fasta / ccb_anchor_end
 *                      from the target FASTA texts the database was built from (the files of the targets
 *                      definition, src/CuCLARK_hh.hh:1129-1198 gives the scan rules); label as in
 *                      ccb_build_add_fasta; add_fasta may be called from several host threads; max_bases bounds
 *                      the bytes of all texts together and must stay below 2^32 (positions are 32-bit)
 *   ccb_anchor_synth   for the synthetic workload of ccb_synth_db (same seed / genomes / labels) */
typedef struct ccb_anchor_info {
    int      ready;
    uint64_t bases;       /* bases per strand, incl. separators and padding */
    uint64_t entries;     /* sampled k-mers stored   */
    uint64_t dropped;     /* sampled k-mers that found their bucket full */
    uint64_t bytes;       /* device bytes */
} ccb_anchor_info;
int  ccb_anchor_begin(ccb_handle* h, uint64_t max_bases);
int  ccb_anchor_add_fasta(ccb_handle* h, const unsigned char* text, uint64_t len, uint16_t label);
int  ccb_anchor_end(ccb_handle* h, ccb_anchor_info* info);
int  ccb_anchor_synth(ccb_handle* h, uint64_t db_seed, uint32_t n_genomes, uint64_t genome_len, uint32_t n_labels);
int  ccb_anchor_drop(ccb_handle* h);
int  ccb_anchor_get_info(const ccb_handle* h, ccb_anchor_info* info);

/* CuClarkDB::malloc (src/CuClarkDB.cu:322-421): pinned host + device buffers for
 * every slot, owned by the library; the caller fills offsets/containers in
 * place and reads final5/rows in place after ccb_batch_wait.  Slots that are
 * already large enough (and idle) are kept: the call is then free. */
int  ccb_batch_alloc(ccb_handle* h, uint32_t max_reads, uint64_t max_containers, int extended);
int  ccb_batch_buffers(ccb_handle* h, int slot, uint32_t** offsets, uint16_t** containers,
                       uint16_t** final5, uint16_t** rows);
/* CuClarkDB::malloc with its own result layout (src/CuClarkDB.cu:322-421): ONE pinned final-result array of
 * num_reads x 5 u16 (and, when extended, ONE full-result array of num_reads x (2*max_hits+2) u16) for the whole file;
 * slot b (= batch b, n_batches == ccb_config.n_slots) delivers its results at row index_batches[b]
 * (h_resultsFinal[i] = _finalResults + rowSize * _indexBatches[i], :412-416), so the reference's print loop
 * (src/CuCLARK_hh.hh:2096-2118) indexes by global read number unchanged.  full_results may be NULL.  The arrays
 * belong to the library (ccb_batch_free); ccb_batch_buffers still hands out the slots' input buffers. */
int  ccb_batch_alloc_ref(ccb_handle* h, uint64_t num_reads, uint32_t max_reads, uint64_t max_containers,
                         const uint64_t* index_batches, int n_batches, int extended,
                         uint16_t** full_results, uint16_t** final_results);
/* CuClarkDB::readyBatch + queryBatch (src/CuClarkDB.cu:821-828, 836-987):
 * async H2D, query, D2H on the slot's stream; returns immediately. */
int  ccb_batch_submit(ccb_handle* h, int slot, uint32_t n_reads, uint64_t n_containers, int flags);
/* CuClarkDB::waitForBatch / checkBatch (src/CuClarkDB.cu:442-456) */
int  ccb_batch_wait(ccb_handle* h, int slot);
int  ccb_batch_poll(ccb_handle* h, int slot);          /* 0 = finished, 1 = still running */
/* CuClarkDB::freeBatchMemory (src/CuClarkDB.cu:285-316) */
int  ccb_batch_free(ccb_handle* h);
/* CuClarkDB::sync (src/CuClarkDB.cu:427-436) */
int  ccb_sync(ccb_handle* h);

/* Device-resident variant of queryBatch for callers that already hold the batch
 * in HBM (multi-GPU pipelines, kernel-only timing): all pointers are device
 * pointers, stream is a cudaStream_t (NULL = default stream).  d_rows may be NULL.
 * d_stats (optional) = 2 x u64 device counters {lookups, sectors}, accumulated (4 x u64 with CCB_STATS_EX:
 * + {k-mers probed in the k-mer table, 0}). */
int  ccb_query_device(ccb_handle* h, const void* d_offsets, const void* d_containers,
                      uint32_t n_reads, void* d_final5, void* d_rows, void* d_stats,
                      int flags, void* stream);

/* ccb_query_device for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  No counterpart in the reference.
 * Same arguments, flags and results as ccb_query_device, but the scratch (overflow / redo lists, the anchor pass's
 * per-read records, the dense path's target counters) is a fixed device array inside the library -- 6.6 MB that every GPU
 * loading the library holds until exit -- so the call makes no cudaMalloc / cudaFree, at the price of limits:
 *   n_reads <= CCB_NOALLOC_MAX_READS per call and a handle created with n_targets <= CCB_NOALLOC_MAX_TARGETS;
 *   beyond either the call fails (-70 / -71, message in ccb_last_error) and nothing is launched.
 * The array's address is looked up on the first call on each device (a call with n_reads = 0 does only that: make
 * it once before capturing; the CUDA driver also loads each kernel's code on its first launch, so run one batch of
 * every kind -- flags, read lengths -- before a region that must not touch the allocator).  ALL calls on one GPU share the array: two calls whose kernels overlap on the GPU (different
 * streams, or a captured graph replayed beside another call) race on it -- order them.  The array keeps the previous
 * call's contents; what has to start at zero is cleared on the caller's stream by the call itself. */
#define CCB_NOALLOC_MAX_READS   65536u
#define CCB_NOALLOC_MAX_TARGETS 8192u
int  ccb_query_device_noalloc(ccb_handle* h, const void* d_offsets, const void* d_containers,
                              uint32_t n_reads, void* d_final5, void* d_rows, void* d_stats,
                              int flags, void* stream);

int  ccb_get_stats(const ccb_handle* h, ccb_stats* out);

/* ---- synthetic workload + roofline probe (bench.py and GPU tests; include/ccb_synth.h) ----
 * ccb_synth_db: build the table from n_genomes counter-generated genomes (label = genome % n_labels).
 * ccb_synth_reads_packed: reads [first_read, first_read+n_reads) in the packed batch layout, in
 *   device memory allocated by the library (release with ccb_device_free).
 * ccb_synth_reads_text: the same reads as text, read_len+1 bytes each ('\n' terminated), to host.
 * ccb_bench_gather: n_probes independent random aligned loads of 32/64/128 B over the table. */
int  ccb_synth_db(ccb_handle* h, uint64_t db_seed, uint32_t n_genomes, uint64_t genome_len,
                  uint32_t n_labels, uint64_t* n_kmers);
int  ccb_synth_reads_packed(ccb_handle* h, uint64_t reads_seed, uint64_t db_seed, uint64_t first_read,
                            uint32_t n_reads, uint32_t read_len, uint32_t n_genomes, uint64_t genome_len,
                            void** d_offsets, void** d_containers, uint64_t* n_containers);
int  ccb_synth_reads_text(ccb_handle* h, uint64_t reads_seed, uint64_t db_seed, uint64_t first_read,
                          uint32_t n_reads, uint32_t read_len, uint32_t n_genomes, uint64_t genome_len,
                          char* host_out);
/* the other configurations (include/ccb_synth.h): a light database (every gap-th non-overlapping k-mer per genome) and
 * reads of `struct ccb_workload` (kind 1 pairs joined by N, kind 2 long reads with indels), packed on the device */
struct ccb_workload;
int  ccb_synth_db_light(ccb_handle* h, uint64_t db_seed, uint32_t n_genomes, uint64_t genome_len,
                        uint32_t n_labels, uint32_t gap, uint64_t* n_kmers);
int  ccb_synth2_reads_packed(ccb_handle* h, const struct ccb_workload* w, uint64_t reads_seed, uint64_t db_seed,
                             uint64_t first_read, uint32_t n_reads, uint32_t n_genomes, uint64_t genome_len,
                             void** d_offsets, void** d_containers, uint64_t* n_containers);
int  ccb_device_free(ccb_handle* h, void* d_ptr);
int  ccb_bench_gather(ccb_handle* h, uint64_t n_probes, int bytes, void* stream);
/* mode 0 uniform / 1 one 2 MiB page per warp load / 2 lane pairs share a 64-B block / 3 lane quads
 * share a 128-B line; footprint = bytes of the table the probes are drawn from (0 = whole table) */
int  ccb_bench_gather_ex(ccb_handle* h, uint64_t n_probes, int bytes, int mode, uint64_t footprint, void* stream);

/* number of CUDA devices visible (0 when there is none or no driver); slots of a handle */
int  ccb_device_count(void);
/* free and total bytes of a device's memory (cudaMemGetInfo); 0 = ok */
int  ccb_device_memory(int device, uint64_t* free_bytes, uint64_t* total_bytes);
int  ccb_batch_slots(const ccb_handle* h);

/* cudaLimitMaxL2FetchGranularity in effect on the handle's device (left at the driver's default;
 * env CCB_L2_FETCH=32|64|128 sets it for measurements -- it changes nothing on B200) */
int  ccb_l2_fetch_granularity(const ccb_handle* h);

/* library / build identification, e.g. "cuclark_b200 0.1 sm_100a" */
const char* ccb_version(void);

#ifdef __cplusplus
}
#endif
#endif
