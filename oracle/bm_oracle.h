/*
 * bm_oracle.h -- CPU restatement of BucketMap's candidate-bucket mapping path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under bucket-map_b200/ may include, link or call this.
 * Allowed users: tests/, __graft_entry__.smoke(), and bench.py's cpu_baseline / --impl reference legs.
 *
 * Every function cites the reference file:line it restates (paths relative to
 * /root/reference/bucket_map/).  Parity pinning status: see oracle/README.md.
 *
 * Data conventions shared with the product (but implemented independently):
 *   - genome: all FASTA records concatenated, 2 bits per base (A0 C1 G2 T3), base i in byte i>>2 at bits 2*(i&3).
 *   - bucket table: b_start[] = global base coordinate of each EMITTED bucket, b_len[] = its length in bases.
 *   - reads: concatenated ASCII bases / ASCII phred+33 qualities, read i = [read_off[i], read_off[i+1]).
 *   - segments: (seg_read, seg_start, seg_len); one per read unless len > 2*read_len (then num_segment_samples).
 *   - candidates: cand_off[2*n_seg+1]; list 2*s = forward strand of segment s, list 2*s+1 = reverse complement.
 */
#ifndef BM_ORACLE_H
#define BM_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_params {
    uint32_t bucket_len;          /* BM_BUCKET_LEN */
    uint32_t read_len;            /* -r */
    uint32_t q;                   /* -k index seed length */
    uint32_t k;                   /* -l query seed length */
    uint32_t mapper_samples;      /* -s */
    uint32_t mapper_fault;        /* ceil(s*e) in float, main.cpp:207 */
    float    distinguishability;  /* -d */
    uint32_t avg_base_quality;    /* -b */
    uint32_t max_candidates;      /* 30, q_gram_mapper.h:285 */
    uint32_t num_segment_samples; /* 5,  q_gram_mapper.h:286 */
    uint32_t locator_samples;     /* -p */
    uint32_t allowed_mismatch;    /* ceil(e*p) in float, bucket_locator.h:419 */
    uint32_t allowed_indel;       /* ceil(n*r) in float, bucket_locator.h:420 */
    float    indel_rate;          /* -n */
    uint32_t quality_threshold;   /* -u */
} orc_params;

typedef struct orc_loc {          /* locate_t, bucket_locator.h:143-148 */
    uint32_t bucket;
    int32_t  offset;
    uint32_t seg_offset;
    uint32_t votes;
    uint32_t is_original;
} orc_loc;

typedef struct orc_aln {          /* one per location in the BM_ALIGN path, bucket_locator.h:560-589 */
    int32_t  score;               /* -(edit distance) */
    uint32_t begin;               /* sequence1_begin_position inside the (possibly reverse-complemented) text window */
    uint32_t map_qual;            /* 60u + score, 32-bit wrap */
    uint32_t kept;                /* map_qual >= quality_threshold */
    uint32_t cigar_off;           /* into the cigar op array */
    uint32_t cigar_len;           /* number of run-length ops; op = (len<<4)|code, code M=0 I=1 D=2 */
} orc_aln;

/* defaults of main.cpp:12-39 with the derived integer thresholds computed in float like the reference */
void orc_default_params(orc_params* p, uint32_t bucket_len);
void orc_derive_params(orc_params* p, uint32_t bucket_len, uint32_t read_len, uint32_t q, uint32_t k, uint32_t s,
                       float d, uint32_t b, float e, float n, uint32_t pl, uint32_t u);

/* tools/hash_function_generator.h:95-117 + indexer/bucket_indexer.h:147-159. returns num_valid_kmers */
uint32_t orc_fracminhash(uint32_t q, uint64_t x, uint64_t y, float frac, int32_t* kmer_to_index);

/* CMakeLists.txt:13-50 (awk): NB = sum ceil(len/L) */
uint32_t orc_num_buckets(const uint64_t* rec_len, uint32_t n_rec, uint32_t L);
/* utils.h:60-102: emitted buckets. arrays sized >= NB. returns emitted count */
uint32_t orc_bucketize(const uint64_t* rec_len, uint32_t n_rec, uint32_t L, uint32_t r,
                       uint64_t* b_start, uint32_t* b_len, uint32_t* b_rec);

/* indexer/bucket_indexer.h:49-100: rows = n_rows * W bytes, .qgram layout (zeroed by callee) */
void orc_index_build(const uint8_t* packed, const uint64_t* b_start, const uint32_t* b_len, uint32_t n_emit,
                     uint32_t NB, uint32_t q, const int32_t* kmer_to_index, uint32_t n_rows, uint8_t* rows,
                     uint32_t threads);

/* mapper/q_gram_mapper.h:171-187 zeros[]; returns number of rows with NB-ones > threshold */
uint32_t orc_zeros(const uint8_t* rows, uint32_t n_rows, uint32_t NB, float d, uint32_t* zeros);

/* mapper/q_gram_mapper.h:509-523 / locator/bucket_locator.h:302-313. arrays sized n_reads*max(1,num_segment_samples) */
uint32_t orc_segments(const orc_params* p, const uint64_t* read_off, uint32_t n_reads,
                      uint32_t* seg_read, uint32_t* seg_start, uint32_t* seg_len);

/* utils.h:160-178 (ub==0 defined as n zeros; reference behaviour there is stale/UB) */
void orc_sample(uint32_t n, uint32_t ub, uint32_t* out);

/* mapper/q_gram_mapper.h:380-480.  cand sized 2*max_candidates*n_seg. stats[0]=segments with >=1 candidate,
 * stats[1]=sum fwd, stats[2]=sum rev, stats[3]=total rows ANDed (R of SURVEY 8d). */
void orc_map_segments(const orc_params* p, uint32_t NB, const int32_t* kmer_to_index, const uint8_t* rows,
                      uint32_t n_rows, const uint8_t* seq, const uint8_t* qual, const uint64_t* read_off,
                      const uint32_t* seg_read, const uint32_t* seg_start, const uint32_t* seg_len, uint32_t n_seg,
                      uint32_t* cand_off, uint32_t* cand, uint64_t* stats, uint32_t threads);

/* the same with the loaded index kept between calls (what q_gram_mapper::load leaves in memory), so a timed
 * mapping pass does not include the index load */
typedef struct orc_index orc_index;
orc_index* orc_index_open(const orc_params* p, uint32_t NB, const int32_t* kmer_to_index, const uint8_t* rows,
                          uint32_t n_rows);
void orc_index_close(orc_index* ix);
void orc_map_segments_h(const orc_index* ix, const orc_params* p, const uint8_t* seq, const uint8_t* qual,
                        const uint64_t* read_off, const uint32_t* seg_read, const uint32_t* seg_start,
                        const uint32_t* seg_len, uint32_t n_seg, uint32_t* cand_off, uint32_t* cand, uint64_t* stats,
                        uint32_t threads);

/* locator/bucket_locator.h:292-347: P (index,hash) pairs per segment */
void orc_prepare_query(const orc_params* p, const uint8_t* seq, const uint8_t* qual, const uint64_t* read_off,
                       const uint32_t* seg_read, const uint32_t* seg_start, const uint32_t* seg_len, uint32_t n_seg,
                       uint16_t* samp_idx, uint32_t* samp_hash);

/* locator/bucket_locator.h:162-290, 613-705.  locs sized cand_off[2*n_seg]; loc_off sized n_reads+1 */
void orc_locate(const orc_params* p, const uint8_t* packed, const uint64_t* b_start, const uint32_t* b_len,
                uint32_t n_emit, const uint8_t* seq, const uint8_t* qual, const uint64_t* read_off, uint32_t n_reads,
                const uint32_t* seg_read, const uint32_t* seg_start, const uint32_t* seg_len, uint32_t n_seg,
                const uint32_t* cand_off, const uint32_t* cand, uint32_t* loc_off, orc_loc* locs, uint32_t threads);

/* locator/bucket_locator.h:350-405. out sized like locs */
void orc_filter_best(const orc_params* p, const uint64_t* read_off, uint32_t n_reads, const uint32_t* loc_off,
                     const orc_loc* locs, uint32_t* out_off, orc_loc* out);

/* locator/bucket_locator.h:520-528,548-589 (seqan3 align_pairwise restated, see README for tie-breaks).
 * cigar sized cigar_cap ops; returns ops used (or needed). */
uint64_t orc_align(const orc_params* p, const uint8_t* packed, const uint64_t* b_start, const uint32_t* b_len,
                   const uint8_t* seq, const uint64_t* read_off, uint32_t n_reads, const uint32_t* loc_off,
                   const orc_loc* locs, orc_aln* aln, uint32_t* cigar, uint64_t cigar_cap, uint32_t threads);

/* whole CLI flows on files (main.cpp:135-233).  x,y = FracMinHash hash parameters (the reference draws them
 * from rand()).  align != 0 selects the bucketmap_align flavour.  return 0 on success. */
int orc_cli_index(const char* fasta, uint32_t bucket_len, uint32_t read_len, uint32_t q, float frac, uint64_t x,
                  uint64_t y, const char* index_dir, const char* indicator, uint32_t threads);
int orc_cli_map(const char* fasta, const orc_params* p, const char* index_dir, const char* indicator,
                const char* fastq, const char* sam, int align, uint32_t threads);

#ifdef __cplusplus
}
#endif
#endif
