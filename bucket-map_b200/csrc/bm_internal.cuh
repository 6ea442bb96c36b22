// bm_internal.cuh -- context layout and small device helpers shared by the kernels of libbucketmap_b200.so.
// Product code: nothing here (or in any file of csrc/) includes, links or calls oracle/.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <vector>

#include "../../include/bucketmap_b200.h"

// ---------------------------------------------------------------------------------------------------------------
// growable device buffer
struct DevBuf {
    void*  p = nullptr;
    size_t cap = 0;
    bool   owned = true;                 // false: a view of another context's buffer (pipeline lanes share the index)
    void view_of(const DevBuf& o) { if (owned && p) cudaFree(p); p = o.p; cap = o.cap; owned = false; }
    cudaError_t ensure(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (!owned) { p = nullptr; cap = 0; owned = true; }
        if (p) { cudaFree(p); p = nullptr; cap = 0; }
        size_t want = bytes + (bytes >> 3) + 256;   // slack so slowly growing batches do not realloc every call
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    // no slack: for the one buffer that is sized in tens of GB (the k-mer row table)
    cudaError_t ensure_exact(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        release();
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e == cudaSuccess) cap = bytes;
        return e;
    }
    void release() { if (p && owned) cudaFree(p); p = nullptr; cap = 0; owned = true; }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

// growable pinned host buffer (staging tables the copy engine reads asynchronously)
struct HostBuf {
    void*  p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        void* np_ = nullptr;
        size_t want = bytes + (bytes >> 3) + 256;
        cudaError_t e = cudaMallocHost(&np_, want);
        if (e != cudaSuccess) return e;
        if (p) cudaFreeHost(p);
        p = np_; cap = want;
        return cudaSuccess;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

struct KStat { double ms = 0; uint64_t launches = 0; };
struct PendingEv { int kid; cudaEvent_t a, b; };

struct bm_gather;

struct bm_ctx {
    int device = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;
    std::string err;

    // reference genome (2-bit) + bucket table
    DevBuf d_packed, d_bstart, d_blen;
    uint64_t n_bases = 0;
    uint32_t n_emit = 0, NB = 0, max_blen = 0;

    // q-gram bit matrix, rows padded to Wp bytes (multiple of 64 B: whole DRAM bursts, 16-B aligned vector loads)
    DevBuf d_rows, d_k2i, d_zeros, d_distok, d_distok_p, d_stage;   // d_distok_p: distok in bit-plane order (K2f)
    uint32_t q = 0, n_rows = 0, W = 0, Wp = 0;
    bool have_index = false;
    float zeros_for_d = -1.f;     // distinguishability the distok bitmap was built for (<0: none)
    uint32_t dist_threshold = 0, dist_valid = 0;

    // coarse k-mer table: for every possible query k-mer x, the bitset q_gram_mapper::query builds for a sample with hash
    // x (AND of the rows of x's indexed q-grams), OR-reduced over groups of 32 buckets: coarse bit g of row x = "some
    // bucket of 32g..32g+31 contains all q-grams of x".  4^k rows of Wcs 32-bit words (Wcs*4 = 128 B on the 1.7 Gbp
    // reference: ONE cache line per sample, 2.1 GB in all), derived on the device from d_rows.  A bucket with < F missed
    // samples lies in a group with < F coarse misses, so the mapper counts coarse misses first and computes exact miss
    // counts only for the (rare) groups that pass.  krows_k = the k it was built for (0: not resident).
    DevBuf d_krows;
    uint32_t krows_k = 0, Wc = 0, Wcs = 0;   // coarse words in use per row / row stride in words (64-B multiple)
    uint32_t ktable_logg = 4, krows_logg = 0;   // log2 of the buckets per coarse bit: wanted / of the resident table
    int ktable_mode = 1;          // 0 never, 1 when it fits (bm_index_kmer_table)
    // mode 1 chooses the buckets per coarse bit (8 / 16 / 32) -- or no table at all -- from a cost model fed with the measured
    // density of the k-mer rows (bm_index.cu: kmer_table_choose); modes 8 / 16 / 32 force a group size
    bool ktable_auto = true;
    float auto_d = -1.f;                                      // ... and the -d it was made for
    uint32_t auto_k = 0, auto_S = 0, auto_F = 0, auto_logg = 4;   // the decision and what it was made for (auto_logg 0xFF: no table)
    double auto_density = 0.0;
    uint64_t ktable_max_bytes = 0;   // 0: half of the free device memory at build time
    bool use_ktable = false;      // decided by bm_internal_kmer_table_prepare for the current mapper call
    // locator samples written by the fused mapper kernel for the resident batch (0: none; else the (k, P, b) they
    // were drawn with) -- bm_locate_device then skips K3a
    uint32_t lsamp_k = 0, lsamp_P = 0, lsamp_b = 0;

    // genome k-mer position index for the locator (bm_locidx.cu): start positions of all k-mers of the reference sorted
    // by (hash, position) + 4^k+1 run offsets; gidx_k = the k it was built for (0: not resident)
    DevBuf d_gidx_pos, d_gidx_off;
    uint32_t gidx_k = 0, gidx_lb = 0, gidx_bshift = 0;   // hash offsets are kept per genome block: 2^lb blocks of 2^bshift bases
    int gidx_mode = 1;            // 0 never, 1 when it fits (bm_locator_index)
    bool use_gidx = false;

    // full-resolution k-mer table (optional, bm_index_fine_table): the same AND as the coarse table but one bit per bucket,
    // 4^k rows of Wp bytes (54.8 GB on the 1.7 Gbp reference).  The fused mapper recounts a passing group from ONE word
    // per sample of this table instead of k-q+1 scattered .qgram words.  Built together with the coarse table when it
    // fits beside everything else; frows_k = the k it was built for (0: not resident).
    DevBuf d_frows;
    uint32_t frows_k = 0, fine_tried_k = 0;   // fine_tried_k: the k a build was last attempted for (not retried per call)
    int ftable_mode = 1;          // 0 never, 1 when it fits
    bool use_ftable = false;

    // read batch.  Canonical device form (what every kernel reads): base bit planes, 32 bases per chunk -- chunk c of
    // read i is d_pl[pl_off[i] + c] = {low bits, high bits} of the ranks of bases 32c..32c+31 (A0 C1 G2 T3, bit t = base
    // 32c+t, bases past the end 0) -- and, when some read of the batch has a base below -b, the k-mer quality plane
    // d_qk (bit t of chunk c: the k bases from 32c+t on have phred sum >= b*k, quality_filter.h:531-534).  ASCII batches
    // (d_seq, d_qual) are turned into this form by k_pack_reads right after their copy.
    DevBuf d_seq, d_qual, d_read_off, d_seg_read, d_seg_start, d_seg_len, d_read_seg0;
    DevBuf d_pl, d_qk, d_pl_off;
    HostBuf h_read_seg0, h_seg_read, h_seg_start, h_seg_len, h_off_rel, h_pl_off;   // pinned; filled by bm_internal_build_segments
    bool have_qk = false;         // false: every k-mer window of the batch passes the quality test (d_qk unused)
    bool ascii_resident = false;  // d_seq / d_qual hold the batch (it came as ASCII): the planes can be re-derived
    bool want_ascii = true;       // the batch being set up arrives as ASCII (sizes d_seq / d_qual)
    uint32_t pack_k = 0, pack_b = 0;   // the (k, b) d_qk was computed for
    uint64_t n_pl_chunks = 0;
    uint32_t n_reads = 0, n_seg = 0, n_seg_total = 0, max_seg_len = 0, max_read_len = 0, batch_max_len = 0;
    bool h_seg_valid = true;      // the pinned host copies of the segment tables describe the resident batch
    uint64_t n_read_bytes = 0;
    bool have_reads = false;

    // mapper results
    DevBuf d_map_over;            // lists the fused mapper kernel hands to K2 (too many groups pass: repeats) + count
    uint32_t map_over_cap = 0;
    int map_max_pass = -1;        // hand-over threshold in passing groups (-1: max(16, n_groups/8); bm_map_handover)
    DevBuf d_rowidx, d_segok, d_cand_fixed, d_cand_cnt, d_cand_off, d_cand, d_stats, d_scan_tmp;
    uint32_t map_max_cand = 0;
    uint64_t n_cand = 0;          // total candidates of the last bm_map_device (valid after bm_map_finish)
    bool have_map = false;

    // locator results
    DevBuf d_lsamp_idx, d_lsamp_hash, d_cres_pos, d_cres_votes, d_cseg, d_loc_cnt, d_loc_off, d_locs, d_locs2,
           d_loc_off2, d_vote_scratch, d_locs3, d_loc_pool, d_loc_flags, d_loc_over;
    uint64_t n_locs = 0, n_locs_best = 0;
    bool have_locs = false, locs_filtered = false, n_locs_known = false;

    // alignment results
    DevBuf d_aln, d_cigar_fixed, d_cigar_cnt, d_cigar_off, d_cigar, d_aln_scratch;
    uint64_t n_aln = 0, n_cigar = 0;
    bool have_aln = false;

    // L2 flush scratch
    DevBuf d_flush;

    // bm_map_locate pipeline: copy stream + one event per chunk; second lane (a child context viewing this one's index)
    cudaStream_t copy_stream = nullptr;
    std::vector<cudaEvent_t> chunk_ev;
    bm_ctx* lane = nullptr;
    cudaEvent_t lane_ev = nullptr;
    bool lanes_off = false;       // true in the child lane itself
    uint32_t lane_min_reads = 0xFFFFFFFFu;   // batches from this size on run on two lanes (bm_pipeline_lanes; default: never)
    HostBuf h_lane_off;

    // multi-GPU (bm_comm.cu): NCCL communicator for the index broadcast; peer-memory gather of the location records
    void* comm = nullptr;
    bool comm_owned = false;
    int comm_rank = 0, comm_nranks = 0;
    bm_gather* gather = nullptr;

    // measurement
    cudaEvent_t ev_a = nullptr, ev_b = nullptr;
    bool ktiming = false;
    KStat kstat[BM_K_COUNT];
    std::vector<PendingEv> pending;
    std::vector<cudaEvent_t> ev_pool;
};

#define BM_CK(ctx, call)                                                                        \
    do {                                                                                        \
        cudaError_t _e = (call);                                                                \
        if (_e != cudaSuccess) {                                                                \
            char _b[512];                                                                       \
            snprintf(_b, sizeof _b, "%s:%d: %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(_e)); \
            (ctx)->err = _b;                                                                    \
            return (_e == cudaErrorMemoryAllocation) ? BM_E_NOMEM : BM_E_CUDA;                  \
        }                                                                                       \
    } while (0)

#define BM_FAIL(ctx, code, msg) do { (ctx)->err = (msg); return (code); } while (0)

// per-kernel timing scope: records events around a launch when ctx->ktiming is on
struct KScope {
    bm_ctx* c; int kid; cudaEvent_t a = nullptr, b = nullptr;
    static cudaEvent_t get(bm_ctx* c) {
        if (!c->ev_pool.empty()) { cudaEvent_t e = c->ev_pool.back(); c->ev_pool.pop_back(); return e; }
        cudaEvent_t e; cudaEventCreate(&e); return e;
    }
    KScope(bm_ctx* c_, int kid_, int own_kernels = 1) : c(c_), kid(kid_) {
        c->kstat[kid].launches += own_kernels;   // hand-written kernels launched inside the scope (library kernels not counted)
        if (c->ktiming) { a = get(c); b = get(c); cudaEventRecord(a, c->stream); }
    }
    ~KScope() {
        if (a) { cudaEventRecord(b, c->stream); c->pending.push_back({kid, a, b}); }
    }
};

int bm_internal_gather_fence(bm_ctx* c);                       // bm_comm.cu: a gather copy still reads the result buffers
int bm_internal_index_prepare(bm_ctx* ctx, float d);          // K0: zeros + distok bitmap for threshold d
int bm_internal_kmer_table_prepare(bm_ctx* ctx, const bm_params* p);  // K0b: k-mer tables for p->k (if they fit and pay off)
int bm_internal_gidx_prepare(bm_ctx* ctx, uint32_t k);        // genome k-mer position index for seed length k (if it fits)
int bm_internal_locate_idx(bm_ctx* c, const bm_params* p, uint32_t n, uint32_t* over_list, uint32_t* d_n_over, uint32_t* n_over);
int bm_internal_exclusive_scan(bm_ctx* ctx, const uint32_t* d_in, uint32_t* d_out, uint32_t n);
int bm_internal_h2d(bm_ctx* c, void* dst, const void* src, size_t bytes);   // large copy from possibly pageable memory
// reads: size the batch buffers (host tables + device), then append the segment table of reads [r0, r1)
int bm_internal_index_build_range(bm_ctx* c, const int32_t* k2i, uint32_t q, uint32_t n_rows, uint32_t b0, uint32_t b1);
int bm_internal_reads_begin(bm_ctx* c, const bm_params* p, const uint64_t* read_off, uint32_t n_reads, const uint32_t* pl_off,
                            bool lazy = false);
// k_pack_reads over reads [r0, r1) of the resident ASCII batch (d_seq, d_qual, d_read_off, d_pl_off) on c->stream
int bm_internal_pack_range(bm_ctx* c, const bm_params* p, uint32_t r0, uint32_t r1);
// makes sure the quality plane of the resident batch matches (k, b) of this call (re-packs an ASCII batch if not)
int bm_internal_pack_check(bm_ctx* c, const bm_params* p);
int bm_internal_build_segments(bm_ctx* c, const bm_params* p, const uint64_t* read_off, uint32_t r0, uint32_t r1);
int bm_internal_build_unit_range(bm_ctx* c, const bm_params* p, const uint64_t* read_off, const uint32_t* pl_off, uint32_t r0, uint32_t r1,
                                 bool scan = false);
// mapper in three steps so that K2a/K2 can run on a segment range as soon as its bytes are resident
int bm_internal_map_begin(bm_ctx* c, const bm_params* p, uint32_t n_seg_cap, uint32_t max_seg_len);
int bm_internal_map_range(bm_ctx* c, const bm_params* p, uint32_t seg0, uint32_t seg1, uint32_t max_seg_len);
int bm_internal_map_finish(bm_ctx* c, const bm_params* p);
int bm_internal_map_fused(bm_ctx* c, const bm_params* p, uint32_t seg0, uint32_t seg1, uint32_t max_seg_len);   // K2f
int bm_internal_map_locate_serial(bm_ctx* c, const bm_params* p, const uint8_t* seq, const uint8_t* qual,
                                  const uint64_t* read_off, uint32_t n_reads, int filter_best, uint32_t* loc_off,
                                  bm_loc* locs, uint64_t loc_cap, uint64_t* stats);

// ---------------------------------------------------------------------------------------------------------------
// device helpers
#ifdef __CUDACC__

// seqan3 dna4 char->rank as used for FASTQ input (utils.h:192-204).  IUPAC codes map to their first base
// (unpinned; synthetic inputs are ACGT only).
__device__ __forceinline__ uint32_t bm_rank_dna4(uint32_t c) {
    c &= 0xDFu;   // upper-case
    uint32_t r = 0;
    r = (c == 'C' || c == 'Y' || c == 'S' || c == 'B') ? 1u : r;
    r = (c == 'G' || c == 'K') ? 2u : r;
    r = (c == 'T' || c == 'U') ? 3u : r;
    return r;
}

// utils.h:291-302: reverse the k base-4 digits and complement each
__device__ __forceinline__ uint32_t bm_revcomp_hash(uint32_t h, uint32_t k) {
    uint32_t x = ~h;
    x = __brev(x);                                              // reverses bits; digits now reversed with swapped bit pairs
    x = ((x >> 1) & 0x55555555u) | ((x & 0x55555555u) << 1);    // restore bit order inside each digit
    return (k >= 16) ? x : (x >> (32 - 2 * k));
}

// base i of the 2-bit packed genome
__device__ __forceinline__ uint32_t bm_gbase(const uint8_t* __restrict__ p, uint64_t i) {
    return (p[i >> 2] >> ((i & 3) << 1)) & 3u;
}

// the resident read batch as the kernels see it
struct ReadPlanes {
    const uint2* pl;          // {lo, hi} per chunk of 32 bases
    const uint32_t* qk;       // k-mer quality plane per chunk, or nullptr: every window passes
    const uint32_t* pl_off;   // first chunk of read i; pl_off[i+1] - pl_off[i] = ceil(len_i / 32)
};

// chunk c (segment positions 32c .. 32c+31) of the segment that starts at base s of the read with chunks [c0, c1)
__device__ __forceinline__ void seg_chunk(const ReadPlanes& rp, uint32_t c0, uint32_t c1, uint32_t s, uint32_t c,
                                          uint32_t& lo, uint32_t& hi, uint32_t& qk) {
    const uint32_t i = c0 + (s >> 5) + c, sh = s & 31u;
    uint2 a = make_uint2(0u, 0u), b = make_uint2(0u, 0u);
    uint32_t qa = 0u, qb = 0u;
    if (i < c1) { a = __ldg(rp.pl + i); qa = rp.qk ? __ldg(rp.qk + i) : 0xFFFFFFFFu; }
    if (sh && i + 1 < c1) { b = __ldg(rp.pl + i + 1); qb = rp.qk ? __ldg(rp.qk + i + 1) : 0xFFFFFFFFu; }
    lo = __funnelshift_r(a.x, b.x, sh); hi = __funnelshift_r(a.y, b.y, sh); qk = __funnelshift_r(qa, qb, sh);
}

// shared by K2a and K3a: stage one segment in shared memory and evaluate its k-mers
struct SegStage {
    uint8_t* base;    // 2-bit ranks, one per byte
    uint32_t* qkm;    // k-mer quality bits, one word per chunk of 32 k-mer positions
    uint32_t* mask;   // one ballot word per chunk of 32 k-mer positions
};

__device__ __forceinline__ uint32_t kmer_hash_at(const uint8_t* sb, uint32_t j, uint32_t k) {
    uint32_t h = 0;   // seqan3 views::kmer_hash(ungapped{k}): first base most significant
    for (uint32_t t = 0; t < k; t++) h = (h << 2) | sb[j + t];
    return h;
}
// utils.h:160-178 with upper_bound = G-1 (see bm_api.cu sample_det for the ub==0 convention)
__device__ __forceinline__ uint32_t sample_index(uint32_t i, uint32_t n, uint32_t G) {
    if (G <= 1) return 0;
    if (i + 1 == n) return G - 1;
    double delta = __ddiv_rn((double)G, (double)(n - 1));
    return (uint32_t)floor(__dmul_rn((double)(int)i, delta));
}

// position of the x-th (0-based) set bit of m (x < popc(m)): binary search on popcounts (__fns is a slow library loop)
__device__ __forceinline__ uint32_t select_bit(uint32_t m, uint32_t x) {
    uint32_t pos = 0, c;
    c = __popc(m & 0xFFFFu); if (x >= c) { x -= c; pos = 16; m >>= 16; }
    c = __popc(m & 0xFFu);   if (x >= c) { x -= c; pos += 8; m >>= 8; }
    c = __popc(m & 0xFu);    if (x >= c) { x -= c; pos += 4; m >>= 4; }
    c = __popc(m & 0x3u);    if (x >= c) { x -= c; pos += 2; m >>= 2; }
    if (x >= (m & 1u)) pos += 1;
    return pos;
}

// position of the x-th (0-based) set bit over the chunk masks
__device__ __forceinline__ uint32_t select_good(const uint32_t* mask, uint32_t nchunk, uint32_t x) {
    uint32_t c = 0;
    for (; c + 1 < nchunk; c++) {
        uint32_t n = __popc(mask[c]);
        if (x < n) break;
        x -= n;
    }
    return c * 32 + select_bit(mask[c], x);
}

// segment (read chunks [c0, c1), start s, n bases) -> ranks, one per byte, and the k-mer quality words
__device__ __forceinline__ void stage_segment(const SegStage& st, const ReadPlanes& rp, uint32_t c0, uint32_t c1,
                                              uint32_t s, uint32_t n, uint32_t lane) {
    const uint32_t nch = (n + 31) / 32;
    for (uint32_t c = 0; c < nch; c++) {
        uint32_t lo, hi, qk;
        seg_chunk(rp, c0, c1, s, c, lo, hi, qk);          // the same words in every lane (broadcast loads)
        const uint32_t i = c * 32 + lane;
        if (i < n) st.base[i] = (uint8_t)((((hi >> lane) & 1u) << 1) | ((lo >> lane) & 1u));
        if (lane == 0) st.qkm[c] = qk;
    }
    __syncwarp();
}
__device__ __forceinline__ bool kmer_qual_ok(const SegStage& st, uint32_t j) { return (st.qkm[j >> 5] >> (j & 31u)) & 1u; }

__device__ __forceinline__ uint4 bm_ldg_stream(const uint4* p) {
    uint4 v;
    asm("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}
#endif
