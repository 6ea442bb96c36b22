/*
 * bucketmap_b200.h -- C-ABI of the B200-native BucketMap hot path (libbucketmap_b200.so).
 *
 * The reference (GZHoffie/bucket-map) has no FFI: its hot path sits behind three C++ class interfaces that only
 * main.cpp consumes.  Every entry point below names the reference member it replaces (paths relative to
 * /root/reference/bucket_map/).  The host classes in bucket-map_b200/host/ (bm::bucket_hash_indexer,
 * bm::q_gram_mapper, bm::bucket_locator) keep the reference signatures and call only these functions.
 *
 * Conventions
 *   - every function returns 0 on success or a negative bm_status; bm_last_error(ctx) gives the text.
 *   - the caller owns every host buffer (pinned memory recommended); the library owns all device memory in bm_ctx.
 *   - one bm_ctx per GPU, used by one host thread at a time; no global state; no exceptions cross the ABI.
 *   - there is NO CPU fallback: without a usable CUDA device bm_ctx_create fails with BM_E_CUDA.
 *
 * Shared data conventions
 *   - reference genome: all FASTA records concatenated, 2 bits per base (A0 C1 G2 T3), base i in byte i>>2 at
 *     bits 2*(i&3).  Buckets are views: bucket b = bases [bucket_start[b], bucket_start[b]+bucket_len[b]).
 *   - q-gram bit matrix ("rows"): the .qgram file layout of indexer/bucket_indexer.h:64-100 -- row-major, one row
 *     per sampled q-gram, W=(NB+7)>>3 bytes per row, bucket j at byte j>>3 bit j&7, no header, no padding.
 *   - reads: concatenated ASCII bases and ASCII phred+33 qualities; read i = [read_off[i], read_off[i+1]).
 *   - candidate lists: cand_off[2*n_seg+1]; list 2*s = forward strand of segment s, 2*s+1 = reverse complement.
 */
#ifndef BUCKETMAP_B200_H
#define BUCKETMAP_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct bm_ctx bm_ctx;

typedef enum bm_status {
    BM_OK = 0,
    BM_E_ARG = -1,      /* bad argument / call order (e.g. map before index load) */
    BM_E_CUDA = -2,     /* CUDA runtime error, text in bm_last_error */
    BM_E_NOMEM = -3,    /* device or host allocation failed */
    BM_E_LIMIT = -4     /* parameter outside what the kernels support (k>16, samples>64, ...) */
} bm_status;

/* Run-time parameters: main.cpp:12-39 (cmd_arguments) + the integer thresholds main.cpp:207 and
 * locator/bucket_locator.h:419-420 derive from them in float. */
typedef struct bm_params {
    uint32_t bucket_len;          /* BM_BUCKET_LEN (CMakeLists.txt:7-58) */
    uint32_t read_len;            /* -r */
    uint32_t q;                   /* -k index seed length */
    uint32_t k;                   /* -l query seed length */
    uint32_t mapper_samples;      /* -s */
    uint32_t mapper_fault;        /* ceil(s*e), main.cpp:207 */
    float    distinguishability;  /* -d */
    uint32_t avg_base_quality;    /* -b */
    uint32_t max_candidates;      /* 30, mapper/q_gram_mapper.h:285 */
    uint32_t num_segment_samples; /* 5,  mapper/q_gram_mapper.h:286 */
    uint32_t locator_samples;     /* -p */
    uint32_t allowed_mismatch;    /* ceil(e*p), locator/bucket_locator.h:419 */
    uint32_t allowed_indel;       /* ceil(n*r), locator/bucket_locator.h:420 */
    float    indel_rate;          /* -n */
    uint32_t quality_threshold;   /* -u */
} bm_params;

/* locate_t of locator/bucket_locator.h:143-148 */
typedef struct bm_loc {
    uint32_t bucket;
    int32_t  offset;       /* read start inside the bucket */
    uint32_t seg_offset;   /* segment start inside the read (0 after the best-location merge) */
    uint32_t votes;
    uint32_t is_original;  /* 1 = forward strand, 0 = reverse complement */
} bm_loc;

/* result of one bucketmap_align verification, locator/bucket_locator.h:560-589 */
typedef struct bm_aln {
    int32_t  score;        /* -(edit distance) */
    uint32_t begin;        /* sequence1_begin_position inside the (possibly reverse-complemented) text window */
    uint32_t map_qual;     /* 60u + score with 32-bit wrap (:570) */
    uint32_t kept;         /* map_qual >= quality_threshold (:571) */
    uint32_t cigar_off;    /* first op in the cigar array */
    uint32_t cigar_len;    /* number of run-length ops; op = (len<<4)|code, code M=0 I=1 D=2 */
} bm_aln;

/* kernel ids for bm_kernel_stats */
enum { BM_K_ROWPOP = 0, BM_K_INDEX = 1, BM_K_SAMPLE = 2, BM_K_MAP = 3, BM_K_LOCSAMPLE = 4, BM_K_LOCATE = 5,
       BM_K_MERGE = 6, BM_K_ALIGN = 7, BM_K_AUX = 8, BM_K_KTABLE = 9, BM_K_GIDX = 10, BM_K_LOCSCAN = 11, BM_K_MAPOVER = 12, BM_K_PACK = 13, BM_K_GATHER = 14, BM_K_COUNT = 15 };

/* ---- context -------------------------------------------------------------------------------------------------- */
int  bm_ctx_create(int device, bm_ctx** out);
void bm_ctx_destroy(bm_ctx* ctx);
const char* bm_last_error(const bm_ctx* ctx);
int  bm_version(void);

/* main.cpp:12-39,207 + bucket_locator.h:419-420: defaults and float-derived thresholds */
void bm_params_derive(bm_params* p, uint32_t bucket_len, uint32_t read_len, uint32_t q, uint32_t k, uint32_t s,
                      float d, uint32_t b, float e, float n, uint32_t pl, uint32_t u);
void bm_params_default(bm_params* p, uint32_t bucket_len);

/* ---- reference genome (replaces the two FASTA passes: utils.h:60-102 called from bucket_indexer.h:198 and
 *      bucket_locator.h:151-160) ---------------------------------------------------------------------------------- */
int bm_ref_upload(bm_ctx* ctx, const uint8_t* packed2bit, uint64_t n_bases, const uint64_t* bucket_start,
                  const uint32_t* bucket_len, uint32_t n_emitted, uint32_t NB);

/* ---- indexer (replaces bucket_indexer::_insert_into_bucket/_store_q_gram_index, bucket_indexer.h:49-100,171-216) - */
/* K1: builds the bit matrix on the device from the uploaded reference; leaves it resident (ready for bm_map_*);
 * if out_rows != NULL also writes the .qgram byte image (n_rows * W bytes) to host memory. */
int bm_index_build(bm_ctx* ctx, const int32_t* kmer_to_index, uint32_t q, uint32_t n_rows, uint8_t* out_rows);
/* replaces q_gram_mapper::load (q_gram_mapper.h:318-372): .qgram bytes + .kmers_index table -> device */
int bm_index_load(bm_ctx* ctx, const int32_t* kmer_to_index, uint32_t q, const uint8_t* rows, uint32_t n_rows,
                  uint32_t NB);
/* copies the resident matrix back in .qgram layout */
int bm_index_download(bm_ctx* ctx, uint8_t* out_rows);
/* distinguishability_filter::read (q_gram_mapper.h:171-187): zeros[] per row; *valid = the logged counter */
int bm_index_zeros(bm_ctx* ctx, float distinguishability, uint32_t* zeros_out /* n_rows or NULL */, uint32_t* valid);
/* coarse k-mer table (no reference counterpart: a B200 memory-layout choice).  For every possible query k-mer x the
 * bitset q_gram_mapper::query builds per sample (q_gram_mapper.h:399-407: the AND of the rows of x's indexed q-grams)
 * is precomputed on the device and OR-reduced over groups of G buckets: 4^k rows of NB/G bits (4^12 x 256 B = 4.3 GB
 * for G=16 on the 1.7 Gbp reference).  The mapper counts missed samples per GROUP from one coarse row per sample (a
 * bucket with < F misses lies in a group with < F coarse misses) and recounts the few passing groups exactly from the
 * .qgram matrix.  The table is derived from the resident matrix at the first mapper call for a given k; index files are
 * unchanged and results are identical.  mode 1 (default): build it when it fits in max_bytes (0 = half of the free
 * device memory) and pays off, G chosen by a cost model from the measured row density; mode 8/16/32: that G, always; mode 0: never (the mapper gathers the k-q+1 q-gram rows of
 * every sample over the full row width). */
int bm_index_kmer_table(bm_ctx* ctx, int mode, uint64_t max_bytes);
/* With the coarse table the mapper recounts exactly every group of buckets that passes the coarse vote.  A read made of
 * repeat k-mers passes almost everywhere; once more than max_groups groups of a list have passed, the list is handed to
 * the full-width q-gram gather (k_map_vote) instead.  -1 (default): max(16, groups per row / 8).  Results are identical
 * for any value. */
int bm_map_handover(bm_ctx* ctx, int max_groups);
/* k the resident table was built for (0: none) and its size */
int bm_index_kmer_table_info(bm_ctx* ctx, uint32_t* k, uint64_t* bytes);
/* What mode 1 decided at the last mapper call: *buckets_per_bit = 8 / 16 / 32 of the resident coarse table, or 0 when no
 * table is in use (switched off, does not fit, or the cost model found the full-width gather cheaper for this index: e.g. a
 * FracMinHash-sparse index, where a third of all k-mers has no indexed q-gram and every coarse bit would be set);
 * *density = measured share of set bits in the k-mer rows that fed the decision (0 if none was made). */
int bm_index_kmer_table_choice(bm_ctx* ctx, uint32_t* buckets_per_bit, double* density);
/* full-resolution k-mer table (no reference counterpart): the same precomputed AND, one bit per BUCKET, 4^k rows of
 * roundup(W,64) bytes (54.8 GB on the 1.7 Gbp reference).  With it the exact recount of a passing group reads ONE 32-bit
 * word per sample instead of the k-q+1 scattered .qgram words (each of which costs a 128-B DRAM line).  Built together
 * with the coarse table, only when it fits beside it with room for batches and the locator index; mode 0: never,
 * 1 (default): when it fits.  Results are identical either way. */
int bm_index_fine_table(bm_ctx* ctx, int mode);
int bm_index_fine_table_info(bm_ctx* ctx, uint32_t* k, uint64_t* bytes);
/* mapper::reset (q_gram_mapper.h:374-378) */
int bm_index_free(bm_ctx* ctx);
/* multi-GPU replication (no reference counterpart: the reference is one process).  bm_index_alloc sizes an empty
 * resident matrix for the given table; bm_index_device / bm_ref_device expose the device buffers (padded rows:
 * n_rows * Wp bytes, Wp = roundup(W,64); packed genome) so the caller can ncclBroadcast into / out of them. */
int bm_index_alloc(bm_ctx* ctx, const int32_t* kmer_to_index, uint32_t q, uint32_t n_rows, uint32_t NB);
int bm_index_device(bm_ctx* ctx, void** dptr, uint64_t* bytes);
int bm_ref_device(bm_ctx* ctx, void** dptr, uint64_t* bytes);

/* ---- reads ---------------------------------------------------------------------------------------------------- */
/* H2D of one read batch + segment table (q_gram_mapper.h:509-523 / bucket_locator.h:302-313).  The batch is turned into
 * its device form on arrival (k_pack_reads): 2-bit base planes and one bit per k-mer window "phred sum >= b*k"
 * (mapper/quality_filter.h:531-534 as used at q_gram_mapper.h:437-442 and bucket_locator.h:323-325), for p->k and
 * p->avg_base_quality; a later call with other values re-derives the plane from the resident qualities. */
int bm_reads_upload(bm_ctx* ctx, const bm_params* p, const uint8_t* seq, const uint8_t* qual,
                    const uint64_t* read_off, uint32_t n_reads);
/* The packed form, made on the HOST by the FASTQ parser (the reference's parser does the same conversion: seqan3 reads
 * FASTQ into dna4 / phred ranks, q_gram_mapper.h:493).  Layout: read i owns chunks [pl_off[i], pl_off[i+1]) with
 * pl_off[i+1] - pl_off[i] = ceil(len_i / 32); chunk c of a read = one uint64: bits 0..31 the low bits and bits 32..63
 * the high bits of the ranks (A0 C1 G2 T3, IUPAC as bm_reads_upload) of bases 32c..32c+31, bit t = base 32c+t, bases past
 * the end 0.  qk (same indexing, one uint32 per chunk): bit t = the k bases starting at 32c+t have phred sum >= b*k.
 *   bm_pack_layout   pl_off[0..n_reads] from read_off.
 *   bm_pack_reads    packs reads [r0, r1) (callable from several host threads on disjoint ranges).  qual may be NULL
 *                    (planes only); qk may be NULL (only *n_low is wanted).  *n_low = reads of the range with at least
 *                    one k-mer window below b*k: when it is 0 for the whole batch the quality plane is all-pass and
 *                    need not travel (pass qk = NULL to the calls below). */
int bm_pack_layout(const uint64_t* read_off, uint32_t n_reads, uint32_t* pl_off);
int bm_pack_reads(const uint8_t* seq, const uint8_t* qual, const uint64_t* read_off, const uint32_t* pl_off,
                  uint32_t r0, uint32_t r1, uint32_t k, uint32_t b, uint64_t* planes, uint32_t* qk, uint32_t* n_low);
/* the same for reads that are not concatenated (views into a FASTQ buffer): read i has read_off[i+1] - read_off[i] bases at
 * seq[i] and qual_len[i] (NULL: as many as bases; missing ones count as '!') quality characters at qual[i] */
int bm_pack_reads_v(const uint8_t* const* seq, const uint8_t* const* qual, const uint32_t* qual_len,
                    const uint64_t* read_off, const uint32_t* pl_off, uint32_t r0, uint32_t r1, uint32_t k, uint32_t b,
                    uint64_t* planes, uint32_t* qk, uint32_t* n_low);
/* Builds everything the first batch would otherwise build on arrival (row zeros for p->distinguishability, the k-mer
 * tables and the locator index for p->k), so that a pipeline can do it while its parser fills the first batch. */
int bm_prepare(bm_ctx* ctx, const bm_params* p);
/* bm_reads_upload for a packed batch.  qk == NULL: every k-mer window passes the quality test for p->avg_base_quality. */
int bm_reads_upload_packed(bm_ctx* ctx, const bm_params* p, const uint64_t* planes, const uint32_t* qk,
                           const uint32_t* pl_off, const uint64_t* read_off, uint32_t n_reads);
int bm_reads_segments(bm_ctx* ctx, uint32_t* n_seg, uint32_t* seg_read, uint32_t* seg_start, uint32_t* seg_len);

/* ---- mapper (replaces q_gram_mapper::query_sequence/query + fault_tolerate_filter, q_gram_mapper.h:39-102,380-480) */
/* K2a+K2 on the resident batch; results stay on the device. */
int bm_map_device(bm_ctx* ctx, const bm_params* p);
/* cand_off: 2*n_seg+1 entries; cand: capacity cand_cap entries (2*max_candidates*n_seg always suffices).
 * stats: 8 entries. [0]=segments with >=1 candidate, [1]=sum forward, [2]=sum reverse, [3]=rows the reference's
 * cascade ANDs for this batch (indexed q-grams of both strands: R of SURVEY 8d), [4]=bytes K2 actually gathered
 * (less than [3]*Wp: tiles whose buckets have all reached F misses are not gathered further; with the k-mer tables: the
 * bytes the fused kernel's loads ask for), [5]=lists the fused kernel handed to the full-width gather (repeats),
 * [6]=coarse groups that passed the coarse vote and were recounted exactly, [7]=candidates whose vote map outgrew the
 * index locator's arrays and went to the bucket-scan kernel (valid after bm_locate_device). */
int bm_map_download(bm_ctx* ctx, uint32_t* cand_off, uint32_t* cand, uint64_t cand_cap, uint64_t* stats);
/* host-buffer form: upload + device + download (what q_gram_mapper::map does per FASTQ batch) */
int bm_map_reads(bm_ctx* ctx, const bm_params* p, const uint8_t* seq, const uint8_t* qual, const uint64_t* read_off,
                 uint32_t n_reads, uint32_t* cand_off, uint32_t* cand, uint64_t cand_cap, uint64_t* stats);

/* ---- locator (replaces bucket_locator::_prepare_read_query/_create_kmer_index/_find_offset/_locate/
 *      _filter_best_locations, bucket_locator.h:162-405,613-705) ------------------------------------------------- */
/* genome k-mer position index (no reference counterpart: a B200 memory-layout choice).  The reference rebuilds an
 * unordered_multimap of a bucket's k-mers for every non-empty bucket (bucket_locator.h:162-177); here the start positions
 * of ALL k-mers of the reference are kept sorted by (hash, position) (4 B per base + 4^k offsets; 6.9 GB for the 1.7 Gbp
 * reference), built on the device at the first locator call for a given k, and a bucket's occurrences of a k-mer are a
 * contiguous run found by binary search.  mode 1 (default): build it when it fits (k <= 14, < 2^32 bases); mode 0:
 * never (the locator scans the text of each candidate bucket).  Results are identical. */
int bm_locator_index(bm_ctx* ctx, int mode);
/* K3a+K3 (+K3b when filter_best != 0) on the resident batch and the candidates of the last bm_map_device. */
int bm_locate_device(bm_ctx* ctx, const bm_params* p, int filter_best);
/* loc_off: n_reads+1; locs: capacity loc_cap (cand_off[2*n_seg] always suffices) */
int bm_locate_download(bm_ctx* ctx, uint32_t* loc_off, bm_loc* locs, uint64_t loc_cap);
/* number of locations the last bm_locate_device produced */
int bm_locate_count(bm_ctx* ctx, uint64_t* n_locs);
/* device-resident results (loc_off: n_reads+1 uint32; locs: n_locs bm_loc) for a device-side gather over NCCL;
 * valid until the next bm_locate_device / bm_reads_upload on this context */
int bm_locate_device_buffers(bm_ctx* ctx, void** loc_off, void** locs, uint64_t* n_locs);
/* host-buffer form of the whole alignment-free path: reads in, merged best locations out */
int bm_map_locate(bm_ctx* ctx, const bm_params* p, const uint8_t* seq, const uint8_t* qual,
                  const uint64_t* read_off, uint32_t n_reads, int filter_best, uint32_t* loc_off, bm_loc* locs,
                  uint64_t loc_cap, uint64_t* stats);
/* bm_map_locate for a packed batch (92 instead of 608 bytes per 300-bp read over PCIe); same records */
int bm_map_locate_packed(bm_ctx* ctx, const bm_params* p, const uint64_t* planes, const uint32_t* qk,
                         const uint32_t* pl_off, const uint64_t* read_off, uint32_t n_reads, int filter_best,
                         uint32_t* loc_off, bm_loc* locs, uint64_t loc_cap, uint64_t* stats);
/* Opt-in: bm_map_locate cuts batches of at least min_reads reads (default and 0: never) in two halves that run on two
 * pipeline lanes at once, so that the locator / merge / result copy of one half run under the host->device copy of the
 * other.  Same records in the same order; after such a call no batch is resident in ctx.  Not faster on a PCIe-bound
 * box (see bm_pipeline.cu), kept for hosts with a faster link. */
int bm_pipeline_lanes(bm_ctx* ctx, uint32_t min_reads);
/* A lane: a second context on ctx's GPU that views ctx's reference, index and derived tables (bm_prepare(ctx, p) is run
 * first) and owns only streams and batch buffers.  Every call of the path returns with its results, so one context keeps
 * one batch in flight; a pipeline that drives ctx and a lane from two host threads (what locator/bucket_locator.h:505-604
 * would be with two batches in flight) hides one batch's host<->device copies under the other's kernels.  The lane is
 * valid until the parent's index, reference or table settings change; destroy it (bm_ctx_destroy) before the parent.
 * Index-changing calls on a lane are undefined. */
int bm_ctx_create_lane(bm_ctx* ctx, const bm_params* p, bm_ctx** lane);

/* ---- several GPUs (no reference counterpart: the reference is one single-threaded process; main.cpp:181-225 is the call
 *      site a multi-GPU bucketmap shards).  Reads are sharded over the ranks, the index is replicated.  One
 *      bm_ctx per GPU: one process per GPU (torchrun, MPI) or several contexts in one process (bucketmap --devices). -- */
#define BM_COMM_ID_BYTES 128      /* ncclUniqueId */
#define BM_PEER_HANDLE_BYTES 128
/* NCCL (dlopen'ed libnccl.so.2): bm_comm_unique_id on one rank, its bytes to every rank by any means, bm_comm_init on all;
 * or bm_comm_attach with a communicator (ncclComm_t) the caller already has. */
int bm_comm_unique_id(uint8_t* id);
int bm_comm_init(bm_ctx* ctx, const uint8_t* id, int nranks, int rank);
int bm_comm_attach(bm_ctx* ctx, void* nccl_comm, int nranks, int rank);
int bm_comm_destroy(bm_ctx* ctx);
/* ncclBroadcast of the resident index (geometry, kmer_to_index, bit matrix) / of the reference (2-bit text + bucket table)
 * from `root` into every rank's context; derived tables are rebuilt locally (cheaper than moving them).  Collective. */
int bm_bcast_index(bm_ctx* ctx, int root);
int bm_bcast_ref(bm_ctx* ctx, int root);
/* bm_index_build (indexer/bucket_indexer.h:171-216, the loop over the buckets) sharded over the ranks: every rank, holding the
 * reference (bm_ref_upload / bm_bcast_ref), indexes its contiguous share of the buckets; ncclAllReduce sums the matrices (a bit
 * belongs to one bucket, so the sum is the OR).  Every rank ends up with the whole index.  Collective, same table everywhere. */
int bm_index_build_sharded(bm_ctx* ctx, const int32_t* kmer_to_index, uint32_t q, uint32_t n_rows);
/* Gather of the location records to `root`, in rank order = read order (contiguous shards).  Not an NCCL call: each rank's
 * records are written into the root's memory over NVLink (peer mapping) behind its locator pass, flagged, and packed by
 * the root -- no host synchronisation on any rank, two batches in flight.
 *   bm_gather_open      on every rank: max_reads / max_locs = largest batch / record count of any rank; writes this rank's
 *                       handle (BM_PEER_HANDLE_BYTES).  The caller passes the handles of all ranks, in rank order, to
 *   bm_gather_connect   on every rank.
 *   bm_gather_results   on every rank after bm_locate_device, the same number of times.  Root with host pointers: blocks
 *                       until the packed records (loc_off: sum of n_reads + 1 entries) are in host memory; with NULLs (and
 *                       on the other ranks) nothing waits.
 *   bm_gather_sync      host waits for this rank's gather stream;  bm_gather_join: the library stream waits for it (so
 *                       bm_timer_stop covers it);  bm_gather_device_buffers: the root's packed result on the device. */
int bm_gather_open(bm_ctx* ctx, int rank, int nranks, int root, uint32_t max_reads, uint64_t max_locs, uint8_t* handle);
int bm_gather_connect(bm_ctx* ctx, const uint8_t* handles);
int bm_gather_results(bm_ctx* ctx, uint32_t* loc_off, bm_loc* locs, uint64_t loc_cap, uint64_t* n_reads_total,
                      uint64_t* n_locs_total);
int bm_gather_sync(bm_ctx* ctx);
/* root: on != 0 -- a bm_gather_results call with host pointers returns once the totals are known; the arrays are complete
 * after bm_gather_sync or the next bm_gather_results (the copy runs under the next batch).  Default 0. */
int bm_gather_host_async(bm_ctx* ctx, int on);
int bm_gather_join(bm_ctx* ctx);
int bm_gather_device_buffers(bm_ctx* ctx, void** loc_off, void** locs, uint64_t* n_reads, uint64_t* n_locs);
int bm_gather_download(bm_ctx* ctx, uint32_t* loc_off, bm_loc* locs, uint64_t loc_cap, uint64_t* n_reads, uint64_t* n_locs);
int bm_gather_close(bm_ctx* ctx);

/* ---- bucketmap_align verification (replaces seqan3::align_pairwise + cigar_from_alignment at
 *      bucket_locator.h:520-528,560-589) -------------------------------------------------------------------------- */
/* K4 on the (unmerged) locations of the last bm_locate_device(filter_best=0). */
int bm_align_device(bm_ctx* ctx, const bm_params* p);
/* aln: one per location; cigar: capacity cigar_cap ops; *cigar_used = ops needed */
int bm_align_download(bm_ctx* ctx, bm_aln* aln, uint32_t* cigar, uint64_t cigar_cap, uint64_t* cigar_used);

/* K4 for callers that cannot allocate.  No counterpart in the reference: added (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The call allocates nothing -- no
 * cudaMalloc / cudaFree, no allocating library routine: the traceback records of the verifier live in a static device array
 * of BM_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until the process exits.
 *   Inputs, all DEVICE pointers in the layouts of the calls above: the 2-bit text + bucket tables (bm_ref_upload), the reads'
 * base planes + chunk offsets (bm_pack_layout / bm_pack_reads) and base offsets, loc_off[n_reads + 1] + locs (what
 * bm_locate_download delivers); max_read_len = the longest read (<= 1024).  n_aln = loc_off[n_reads] alignments are made.
 *   Outputs, DEVICE pointers: aln[n_aln] (same fields as bm_align_device; cigar_off = i * BM_NOALLOC_CIGAR_OPS), cigar[n_aln *
 * BM_NOALLOC_CIGAR_OPS] (the ops of alignment i, first op first, in its slot; an alignment with more than
 * BM_NOALLOC_CIGAR_OPS runs reports its cigar_len but no ops: use bm_align_device for those), cigar_cnt[n_aln].
 *   Limit: ceil(n_aln / 128) * 128 * max_cols * 8 * words bytes of records must fit the array, max_cols = max_read_len + 3 +
 * (uint32_t)(p->indel_rate * max_read_len), words = 4 / 6 / 8 / 10 / 12 / 18 / 36 for reads of up to 64 / 128 / 192 / 256 /
 * 320 / 512 / 1024 bases: 1,024 alignments of 300-base reads, 3,328 of 150-base reads.  More is refused with BM_E_LIMIT and
 * a message that names the bound; split the batch.
 *   The kernels are enqueued on ctx's stream and the call returns without waiting (bm_sync).  All calls share the array: calls
 * that overlap on the GPU race; nothing in it must be cleared between calls. */
#define BM_NOALLOC_SCRATCH_BYTES (32u << 20)
#define BM_NOALLOC_CIGAR_OPS 16
int bm_align_noalloc(bm_ctx* ctx, const bm_params* p, const uint8_t* d_packed, const uint64_t* d_bstart,
                     const uint32_t* d_blen, const uint64_t* d_planes, const uint32_t* d_pl_off,
                     const uint64_t* d_read_off, uint32_t n_reads, uint32_t max_read_len,
                     const uint32_t* d_loc_off, const bm_loc* d_locs, uint32_t n_aln,
                     bm_aln* d_aln, uint32_t* d_cigar, uint32_t* d_cigar_cnt);

/* ---- measurement ----------------------------------------------------------------------------------------------- */
/* CUDA events on the library's stream around whatever is enqueued between start and stop. */
int bm_timer_start(bm_ctx* ctx);
int bm_timer_stop(bm_ctx* ctx, float* ms);
/* accumulated per-kernel device time (CUDA events around each launch) and launch counts since the last reset;
 * timing each launch costs two event records, so it is off unless enabled. */
int bm_kernel_timing(bm_ctx* ctx, int enable);
int bm_kernel_stats(bm_ctx* ctx, int kernel_id, double* total_ms, uint64_t* launches);
int bm_kernel_stats_reset(bm_ctx* ctx);
/* writes `bytes` bytes of device scratch (L2 flush between timed iterations) */
int bm_flush_l2(bm_ctx* ctx, uint64_t bytes);
int bm_sync(bm_ctx* ctx);
/* pinned host memory for the caller's batches */
void* bm_host_alloc(uint64_t bytes);
void  bm_host_free(void* p);

#ifdef __cplusplus
}
#endif
#endif
