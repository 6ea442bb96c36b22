// bm_align.cu -- bucketmap_align's verification (K4): semi-global unit-cost edit distance of the whole read against the
// text window of every located candidate, with begin position and CIGAR.
//
// Replaces seqan3::align_pairwise(edit_scheme, free end gaps in sequence1=text) + cigar_from_alignment at
// locator/bucket_locator.h:520-528,548-589.
//
// One thread per alignment.  Forward pass: Myers' bit-vector recurrence in its block form (64 query rows per word, NW
// words in registers, horizontal carry between words), one text column per step; after each column the thread stores
// the vertical delta vectors (Pv, Mv) and the score at every word boundary to a column-major scratch record, so the
// warp's stores are contiguous.  The best end column is the LAST column attaining the minimum of the bottom row.
// Traceback: any cell value is boundary score + popcount of the masked deltas of ONE word, so the thread walks back
// from (m, end) reading two records per step and prefers diagonal, then up (query base consumed: I), then left (text
// base consumed: D) -- the conventions of the oracle (oracle/bm_oracle.cpp align_semi_global; seqan3's own tie-breaks
// are not available in this environment).
//
// Two kernels.  k_align_band (the common case) keeps, per text column, only the two 64-row words of the delta vectors that
// the optimal path can reach: with e edits the path stays on diagonals j - i in [-e, (n-m)+e], so for every e up to
// E = (63 - (n-m))/2 (28 for a 300-base read and its 307-column window) the rows [j-(n-m)-E, j+E+1] of column j lie in words
// w0(j), w0(j)+1 -- 40 bytes per column instead of 96, written plane by plane so that a warp's stores are contiguous, with the
// pattern masks in shared memory instead of 40 registers per thread.  Alignments with more than E edits (false candidates)
// go to k_align, which records whole columns.
#include <algorithm>
#include <mutex>

#include "bm_internal.cuh"

namespace {

constexpr int kAlnThreads = 128;
constexpr uint32_t kCigTmp = 16;       // ops kept per alignment in the first pass

struct AlignArgs {
    const uint8_t* packed; const uint64_t* bstart; const uint32_t* blen;
    ReadPlanes rp; const uint64_t* read_off; uint32_t n_reads;
    const uint32_t* loc_off; const bm_loc* locs;
    float indel_rate; uint32_t quality_threshold;
    unsigned long long* scratch; uint32_t chunk_n, max_cols;
    bm_aln* aln; uint32_t* cig_tmp; uint32_t* cig_cnt;
    const uint32_t* cig_off; uint32_t* cigar; const uint32_t* over_list;
    const uint32_t* id_list;          // MODE 0 over a list (alignments the band kernel passed on) or nullptr: first + t
    uint32_t* band_over; uint32_t* n_band_over;
};

template <int NW> struct Rec { static constexpr int kWords = (2 * NW + (NW + 3) / 4 + 1) & ~1; };

__device__ __forceinline__ uint32_t read_of_loc(const uint32_t* loc_off, uint32_t n_reads, uint32_t id) {
    uint32_t lo = 0, hi = n_reads;            // last r with loc_off[r] <= id
    while (hi - lo > 1) { uint32_t mid = (lo + hi) >> 1; if (loc_off[mid] <= id) lo = mid; else hi = mid; }
    return lo;
}

template <int NW>
struct Tracer {
    const unsigned long long* scratch; size_t stride; uint32_t slot;
    __device__ __forceinline__ unsigned long long word(uint32_t col, uint32_t w) const {
        return scratch[((size_t)col * stride + slot) * Rec<NW>::kWords + w];
    }
    // D[i][col]
    __device__ __forceinline__ int cell(uint32_t col, uint32_t i) const {
        if (i == 0) return 0;
        const uint32_t w = (i - 1) >> 6, nb = i - 64 * w;
        const unsigned long long mask = nb == 64 ? ~0ull : ((1ull << nb) - 1ull);
        const unsigned long long pv = word(col, w), mv = word(col, NW + w);
        const int bs = (int)((word(col, 2 * NW + (w >> 2)) >> (16 * (w & 3))) & 0xFFFFull);
        return bs + __popcll(pv & mask) - __popcll(mv & mask);
    }
    // D[i][col] - D[i-1][col], i >= 1
    __device__ __forceinline__ int vdelta(uint32_t col, uint32_t i) const {
        const uint32_t w = (i - 1) >> 6, b = (i - 1) & 63;
        return (int)((word(col, w) >> b) & 1ull) - (int)((word(col, NW + w) >> b) & 1ull);
    }
};

// MODE 0: first pass (score, begin, op count, first kCigTmp ops reversed); MODE 1: overflow list, ops written in place
template <int NW, int MODE>
__global__ void __launch_bounds__(kAlnThreads)
k_align(AlignArgs a, uint32_t first, uint32_t count) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= count) return;
    const uint32_t id = MODE ? a.over_list[first + t] : (a.id_list ? a.id_list[first + t] : first + t);
    const bm_loc l = a.locs[id];
    const uint32_t r = read_of_loc(a.loc_off, a.n_reads, id);
    const uint2* qpl = a.rp.pl + a.rp.pl_off[r];            // bit planes of the whole read (32 bases per chunk)
    const uint32_t qch = a.rp.pl_off[r + 1] - a.rp.pl_off[r];
    const uint32_t m = (uint32_t)(a.read_off[r + 1] - a.read_off[r]);
    const uint32_t blen = a.blen[l.bucket];
    bm_aln out;
    out.score = 0; out.begin = 0; out.map_qual = 0; out.kept = 0; out.cigar_off = 0; out.cigar_len = 0;
    // a negative or out-of-bucket offset makes the reference slice its bucket out of range (undefined there):
    // such a location is reported as not kept with an empty alignment
    if (l.offset < 0 || (uint32_t)l.offset >= blen || m == 0 || m > 64u * NW) {
        if (!MODE) { a.aln[id] = out; a.cig_cnt[id] = 0; }
        return;
    }
    // :550  width = min(|query| + 1 + (size_t)(rate * |read|), |bucket| - offset)
    const uint32_t extra = (uint32_t)(unsigned long long)__fmul_rn(a.indel_rate, (float)m);
    const uint32_t n = min(m + 1u + extra, blen - (uint32_t)l.offset);
    const uint64_t tbase = a.bstart[l.bucket] + (uint32_t)l.offset;
    const bool rev = l.is_original == 0;

    unsigned long long P0[NW], P1[NW], P2[NW], P3[NW];
#pragma unroll
    for (int w = 0; w < NW; w++) {
        const uint32_t lim = (m > 64u * w) ? min(64u, m - 64u * w) : 0u;
        const unsigned long long mask = lim == 64u ? ~0ull : ((1ull << lim) - 1ull);
        uint2 c0 = make_uint2(0u, 0u), c1 = make_uint2(0u, 0u);
        if (2u * w < qch) c0 = __ldg(qpl + 2 * w);
        if (2u * w + 1u < qch) c1 = __ldg(qpl + 2 * w + 1);
        const unsigned long long lo = (unsigned long long)c0.x | ((unsigned long long)c1.x << 32);
        const unsigned long long hi = (unsigned long long)c0.y | ((unsigned long long)c1.y << 32);
        P0[w] = ~lo & ~hi & mask; P1[w] = lo & ~hi & mask; P2[w] = ~lo & hi & mask; P3[w] = lo & hi & mask;
    }
    unsigned long long Pv[NW], Mv[NW];
    int bs[NW];
#pragma unroll
    for (int w = 0; w < NW; w++) { Pv[w] = ~0ull; Mv[w] = 0ull; bs[w] = 64 * w; }
    const uint32_t lastw = (m - 1) >> 6, lastb = (m - 1) & 63;
    const size_t stride = a.chunk_n;
    unsigned long long* rec0 = a.scratch + (size_t)t * Rec<NW>::kWords;
    auto store = [&](uint32_t col) {
        unsigned long long rec[Rec<NW>::kWords];
#pragma unroll
        for (int w = 0; w < Rec<NW>::kWords; w++) rec[w] = 0ull;
#pragma unroll
        for (int w = 0; w < NW; w++) {
            rec[w] = Pv[w]; rec[NW + w] = Mv[w];
            rec[2 * NW + (w >> 2)] |= (unsigned long long)(bs[w] & 0xFFFF) << (16 * (w & 3));
        }
        ulonglong2* dst = reinterpret_cast<ulonglong2*>(rec0 + (size_t)col * stride * Rec<NW>::kWords);
#pragma unroll
        for (int w = 0; w < Rec<NW>::kWords / 2; w++) dst[w] = make_ulonglong2(rec[2 * w], rec[2 * w + 1]);
    };
    store(0);
    int score = (int)m, best = (int)m;
    uint32_t bj = 0;
    for (uint32_t j = 1; j <= n; j++) {
        const uint32_t tc = rev ? 3u - bm_gbase(a.packed, tbase + (n - j)) : bm_gbase(a.packed, tbase + (j - 1));   // :563-567
        int hin = 0;
#pragma unroll
        for (int w = 0; w < NW; w++) {
            unsigned long long Eq = tc == 0 ? P0[w] : (tc == 1 ? P1[w] : (tc == 2 ? P2[w] : P3[w]));
            const unsigned long long hneg = hin < 0 ? 1ull : 0ull, hpos = hin > 0 ? 1ull : 0ull;
            const unsigned long long Xv = Eq | Mv[w];
            Eq |= hneg;
            const unsigned long long Xh = (((Eq & Pv[w]) + Pv[w]) ^ Pv[w]) | Eq;
            unsigned long long Ph = Mv[w] | ~(Xh | Pv[w]);
            unsigned long long Mh = Pv[w] & Xh;
            if (w == (int)lastw) score += (int)((Ph >> lastb) & 1ull) - (int)((Mh >> lastb) & 1ull);
            const int hout = (int)(Ph >> 63) - (int)(Mh >> 63);
            Ph = (Ph << 1) | hpos;
            Mh = (Mh << 1) | hneg;
            Pv[w] = Mh | ~(Xv | Ph);
            Mv[w] = Ph & Xv;
            if (w + 1 < NW) bs[w + 1] += hout;
            hin = hout;
        }
        store(j);
        if (score <= best) { best = score; bj = j; }        // last column attaining the minimum
    }
    // traceback from (m, bj)
    Tracer<NW> tr{a.scratch, stride, t};
    uint32_t i = m, j = bj, nops = 0, cur_op = 3, cur_len = 0;
    int cur = best;
    uint32_t* tmp = a.cig_tmp + (size_t)id * kCigTmp;
    uint32_t wpos = 0;
    if (MODE) wpos = a.cig_off[id] + a.aln[id].cigar_len;    // ops are produced last-to-first: fill backwards
    auto flush = [&]() {
        if (cur_len) {
            const uint32_t op = (cur_len << 4) | cur_op;
            if (MODE) a.cigar[--wpos] = op;
            else if (nops < kCigTmp) tmp[nops] = op;
            nops++;
        }
    };
    while (i > 0) {
        uint32_t op;
        int nxt;
        const int dg = (j > 0) ? tr.cell(j - 1, i - 1) : 0;
        bool mism = true;
        if (j > 0) {
            const uint32_t tc = rev ? 3u - bm_gbase(a.packed, tbase + (n - j)) : bm_gbase(a.packed, tbase + (j - 1));
            const uint2 qc = __ldg(qpl + ((i - 1) >> 5));
            mism = ((((qc.y >> ((i - 1) & 31u)) & 1u) << 1) | ((qc.x >> ((i - 1) & 31u)) & 1u)) != tc;
        }
        if (j > 0 && cur == dg + (mism ? 1 : 0)) { op = 0; nxt = dg; }
        else {
            const int up = cur - tr.vdelta(j, i);
            if (cur == up + 1) { op = 1; nxt = up; }
            else { op = 2; nxt = tr.cell(j - 1, i); }
        }
        if (op != cur_op) { flush(); cur_op = op; cur_len = 0; }
        cur_len++;
        cur = nxt;
        if (op == 0) { i--; j--; } else if (op == 1) { i--; } else { j--; }
    }
    flush();
    if (!MODE) {
        out.score = -best;
        out.begin = j;
        out.map_qual = 60u + (uint32_t)out.score;                                    // :570, 32-bit wrap
        out.kept = ((unsigned long long)out.map_qual >= (unsigned long long)a.quality_threshold) ? 1u : 0u;
        out.cigar_len = nops;
        a.aln[id] = out;
        a.cig_cnt[id] = nops;
    }
}

// ---- band-limited first pass ------------------------------------------------------------------------------------------------
// 32 bases of the 2-bit genome from base `pos` on (base t at bits 2t): two aligned 8-byte loads and a funnel shift.  The
// text of an alignment is read 32 columns at a time, one block ahead, instead of one dependent byte load per column.
__device__ __forceinline__ unsigned long long load_bases32(const uint8_t* __restrict__ packed, uint64_t pos) {
    const uint64_t byte = pos >> 2;
    const uint32_t s = (uint32_t)(byte & 7u) * 8u + (uint32_t)(pos & 3u) * 2u;
    const unsigned long long* p = reinterpret_cast<const unsigned long long*>(packed + (byte & ~(uint64_t)7));
    const unsigned long long w0 = __ldg(p), w1 = __ldg(p + 1);
    return s ? ((w0 >> s) | (w1 << (64u - s))) : w0;
}

constexpr int kBandPlanes = 5;         // Pv[w0], Pv[w0+1], Mv[w0], Mv[w0+1], boundary score of word w0

// first word kept for column j: the word of row max(1, j - nm - E), at most NW-2 (two words are kept)
template <int NW>
__device__ __forceinline__ uint32_t band_w0(int j, int nm, int E) {
    if (NW == 1) return 0u;
    int r = j - nm - E;
    if (r < 1) r = 1;
    uint32_t w = (uint32_t)(r - 1) >> 6;
    return w > (uint32_t)(NW - 2) ? (uint32_t)(NW - 2) : w;
}

template <int NW>
struct BandTracer {
    const unsigned long long* rec; int nm, E;                     // rec: this lane's word of column 0, plane 0
    __device__ __forceinline__ unsigned long long plane(uint32_t col, uint32_t pl) const {
        return rec[((size_t)col * kBandPlanes + pl) * 32];
    }
    // the next columns the walk will need, into L1 (a warp's five planes of one column are 1,280 contiguous bytes)
    __device__ __forceinline__ void prefetch(uint32_t col) const {
#pragma unroll
        for (int pl = 0; pl < kBandPlanes; pl++)
            asm volatile("prefetch.global.L1 [%0];" ::"l"(rec + ((size_t)col * kBandPlanes + pl) * 32));
    }
    // D[i][col]
    __device__ __forceinline__ int cell(uint32_t col, uint32_t i) const {
        if (i == 0) return 0;
        const uint32_t w0 = band_w0<NW>((int)col, nm, E);
        const uint32_t w = (i - 1) >> 6, nb = i - 64 * w;
        const unsigned long long mask = nb == 64 ? ~0ull : ((1ull << nb) - 1ull);
        int bs = (int)(long long)plane(col, 4);
        if (w != w0) {                                            // second kept word: its boundary score follows from the first
            bs += __popcll(plane(col, 0)) - __popcll(plane(col, 2));
            return bs + __popcll(plane(col, 1) & mask) - __popcll(plane(col, 3) & mask);
        }
        return bs + __popcll(plane(col, 0) & mask) - __popcll(plane(col, 2) & mask);
    }
    // D[i][col] - D[i-1][col], i >= 1
    __device__ __forceinline__ int vdelta(uint32_t col, uint32_t i) const {
        const uint32_t w0 = band_w0<NW>((int)col, nm, E);
        const uint32_t w = (i - 1) >> 6, b = (i - 1) & 63;
        const uint32_t sel = (w != w0) ? 1u : 0u;
        return (int)((plane(col, sel) >> b) & 1ull) - (int)((plane(col, 2 + sel) >> b) & 1ull);
    }
};

template <int NW>
__global__ void __launch_bounds__(kAlnThreads)
k_align_band(AlignArgs a, uint32_t first, uint32_t count) {
    extern __shared__ unsigned long long s_eq[];                  // [4][NW][kAlnThreads] pattern masks
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= count) return;
    const uint32_t id = first + t;
    const bm_loc l = a.locs[id];
    const uint32_t r = read_of_loc(a.loc_off, a.n_reads, id);
    const uint2* qpl = a.rp.pl + a.rp.pl_off[r];
    const uint32_t qch = a.rp.pl_off[r + 1] - a.rp.pl_off[r];
    const uint32_t m = (uint32_t)(a.read_off[r + 1] - a.read_off[r]);
    const uint32_t blen = a.blen[l.bucket];
    bm_aln out;
    out.score = 0; out.begin = 0; out.map_qual = 0; out.kept = 0; out.cigar_off = 0; out.cigar_len = 0;
    if (l.offset < 0 || (uint32_t)l.offset >= blen || m == 0 || m > 64u * NW) {
        a.aln[id] = out; a.cig_cnt[id] = 0;
        return;
    }
    const uint32_t extra = (uint32_t)(unsigned long long)__fmul_rn(a.indel_rate, (float)m);
    const uint32_t n = min(m + 1u + extra, blen - (uint32_t)l.offset);          // :550
    const uint64_t tbase = a.bstart[l.bucket] + (uint32_t)l.offset;
    const bool rev = l.is_original == 0;
    const int nm = (int)n - (int)m;
    const int E = (63 - (nm > 0 ? nm : 0)) / 2;

    unsigned long long* eq = s_eq + threadIdx.x;
#define EQ(c, w) eq[((c) * NW + (w)) * kAlnThreads]
#pragma unroll
    for (int w = 0; w < NW; w++) {
        const uint32_t lim = (m > 64u * w) ? min(64u, m - 64u * w) : 0u;
        const unsigned long long mask = lim == 64u ? ~0ull : ((1ull << lim) - 1ull);
        uint2 c0 = make_uint2(0u, 0u), c1 = make_uint2(0u, 0u);
        if (2u * w < qch) c0 = __ldg(qpl + 2 * w);
        if (2u * w + 1u < qch) c1 = __ldg(qpl + 2 * w + 1);
        const unsigned long long lo = (unsigned long long)c0.x | ((unsigned long long)c1.x << 32);
        const unsigned long long hi = (unsigned long long)c0.y | ((unsigned long long)c1.y << 32);
        EQ(0, w) = ~lo & ~hi & mask; EQ(1, w) = lo & ~hi & mask; EQ(2, w) = ~lo & hi & mask; EQ(3, w) = lo & hi & mask;
    }
    unsigned long long Pv[NW], Mv[NW];
    int bs[NW];
#pragma unroll
    for (int w = 0; w < NW; w++) { Pv[w] = ~0ull; Mv[w] = 0ull; bs[w] = 64 * w; }
    const uint32_t lastw = (m - 1) >> 6, lastb = (m - 1) & 63;
    // Scratch layout [warp][column][plane][lane]: a warp writes 1,280 contiguous bytes per column and its columns follow
    // each other, so the forward pass streams ~400 KB per warp front to back and the traceback walks the same region back
    // (open DRAM pages, not one page per store as a [column][plane][alignment] layout would touch).
    constexpr size_t stride = 32;
    unsigned long long* rec0 = a.scratch + (size_t)(t >> 5) * a.max_cols * kBandPlanes * 32 + (t & 31u);
    constexpr int W1 = NW > 1 ? 1 : 0;
    // column 0
    rec0[0 * stride] = Pv[0]; rec0[1 * stride] = Pv[W1]; rec0[2 * stride] = Mv[0]; rec0[3 * stride] = Mv[W1];
    rec0[4 * stride] = (unsigned long long)(long long)bs[0];
    int score = (int)m, best = (int)m;
    uint32_t bj = 0;
    uint32_t j = 1;
    // text of column j: base idx = j-1 of the window, or, reverse strand, the complement of base n-j (:563-567); the block
    // of 32 bases in use and the one the walk enters next are kept in registers
    const uint32_t idx0 = rev ? n - 1 : 0u;
    uint32_t tblk = idx0 >> 5;
    unsigned long long tw = load_bases32(a.packed, tbase + 32ull * tblk);
    unsigned long long tw_next = (rev ? tblk > 0 : (uint64_t)32 * (tblk + 1) < n)
                                     ? load_bases32(a.packed, tbase + 32ull * (rev ? tblk - 1 : tblk + 1)) : 0ull;
    // the columns are walked in NW-1 phases: inside phase ww the kept words are ww, ww+1 (compile-time register indices)
#pragma unroll
    for (int ww = 0; ww < (NW > 1 ? NW - 1 : 1); ww++) {
        uint32_t jend = n;
        if (ww < NW - 2) { const int e2 = 64 * (ww + 1) + nm + E; jend = e2 < 0 ? 0u : ((uint32_t)e2 < n ? (uint32_t)e2 : n); }
        for (; j <= jend; j++) {
            const uint32_t idx = rev ? n - j : j - 1;
            if ((idx >> 5) != tblk) {                            // entered the next block: it is there already; fetch the one after
                tblk = idx >> 5;
                tw = tw_next;
                if (rev ? tblk > 0 : (uint64_t)32 * (tblk + 1) < n)
                    tw_next = load_bases32(a.packed, tbase + 32ull * (rev ? tblk - 1 : tblk + 1));
            }
            const uint32_t tb = (uint32_t)(tw >> ((idx & 31u) * 2u)) & 3u;
            const uint32_t tc = rev ? 3u - tb : tb;
            int hin = 0;
#pragma unroll
            for (int w = 0; w < NW; w++) {
                unsigned long long Eq = EQ(tc, w);
                const unsigned long long hneg = hin < 0 ? 1ull : 0ull, hpos = hin > 0 ? 1ull : 0ull;
                const unsigned long long Xv = Eq | Mv[w];
                Eq |= hneg;
                const unsigned long long Xh = (((Eq & Pv[w]) + Pv[w]) ^ Pv[w]) | Eq;
                unsigned long long Ph = Mv[w] | ~(Xh | Pv[w]);
                unsigned long long Mh = Pv[w] & Xh;
                if (w == (int)lastw) score += (int)((Ph >> lastb) & 1ull) - (int)((Mh >> lastb) & 1ull);
                const int hout = (int)(Ph >> 63) - (int)(Mh >> 63);
                Ph = (Ph << 1) | hpos;
                Mh = (Mh << 1) | hneg;
                Pv[w] = Mh | ~(Xv | Ph);
                Mv[w] = Ph & Xv;
                if (w + 1 < NW) bs[w + 1] += hout;
                hin = hout;
            }
            unsigned long long* rc = rec0 + (size_t)j * kBandPlanes * stride;
            rc[0 * stride] = Pv[ww]; rc[1 * stride] = Pv[ww + W1]; rc[2 * stride] = Mv[ww]; rc[3 * stride] = Mv[ww + W1];
            rc[4 * stride] = (unsigned long long)(long long)bs[ww];
            if (score <= best) { best = score; bj = j; }        // last column attaining the minimum
        }
    }
#undef EQ
    out.score = -best;
    out.map_qual = 60u + (uint32_t)out.score;                                    // :570, 32-bit wrap
    out.kept = ((unsigned long long)out.map_qual >= (unsigned long long)a.quality_threshold) ? 1u : 0u;
    if (best > E) {                                              // the path may leave the kept words: full-column kernel
        a.aln[id] = out; a.cig_cnt[id] = 0;
        a.band_over[atomicAdd(a.n_band_over, 1u)] = id;
        return;
    }
    // traceback from (m, bj): diagonal, then up (I), then left (D)
    BandTracer<NW> tr{rec0, nm, E};
    uint32_t i = m, nops = 0, cur_op = 3, cur_len = 0;
    j = bj;
    int cur = best;
    uint32_t tb_blk = 0xFFFFFFFFu, q_blk = 0xFFFFFFFFu;          // cached block of the text / chunk of the read
    unsigned long long tb_w = 0ull;
    uint2 q_w = make_uint2(0u, 0u);
    for (uint32_t d = 1; d <= 3 && d <= j; d++) tr.prefetch(j - d);
    uint32_t* tmp = a.cig_tmp + (size_t)id * kCigTmp;
    auto flush = [&]() {
        if (cur_len) {
            if (nops < kCigTmp) tmp[nops] = (cur_len << 4) | cur_op;
            nops++;
        }
    };
    while (i > 0) {
        uint32_t op;
        int nxt;
        const int dg = (j > 0) ? tr.cell(j - 1, i - 1) : 0;
        bool mism = true;
        if (j > 0) {
            const uint32_t idx = rev ? n - j : j - 1;
            if ((idx >> 5) != tb_blk) { tb_blk = idx >> 5; tb_w = load_bases32(a.packed, tbase + 32ull * tb_blk); }
            const uint32_t tb = (uint32_t)(tb_w >> ((idx & 31u) * 2u)) & 3u;
            const uint32_t tc = rev ? 3u - tb : tb;
            if (((i - 1) >> 5) != q_blk) { q_blk = (i - 1) >> 5; q_w = __ldg(qpl + q_blk); }
            mism = ((((q_w.y >> ((i - 1) & 31u)) & 1u) << 1) | ((q_w.x >> ((i - 1) & 31u)) & 1u)) != tc;
        }
        if (j > 0 && cur == dg + (mism ? 1 : 0)) { op = 0; nxt = dg; }
        else {
            const int up = cur - tr.vdelta(j, i);
            if (cur == up + 1) { op = 1; nxt = up; }
            else { op = 2; nxt = tr.cell(j - 1, i); }
        }
        if (op != cur_op) { flush(); cur_op = op; cur_len = 0; }
        cur_len++;
        cur = nxt;
        if (op == 0) { i--; j--; } else if (op == 1) { i--; } else { j--; }
        if (op != 1 && j >= 4) tr.prefetch(j - 4);                // the column the walk reaches three steps from now
    }
    flush();
    out.begin = j;
    out.cigar_len = nops;
    a.aln[id] = out;
    a.cig_cnt[id] = nops;
}

// final CIGAR arrays: reverse the first-pass ops into place; alignments with more ops than kCigTmp go to the overflow list
__global__ void k_cigar_write(bm_aln* __restrict__ aln, const uint32_t* __restrict__ cig_off,
                              const uint32_t* __restrict__ cig_tmp, uint32_t n, uint32_t* __restrict__ cigar,
                              uint32_t* __restrict__ over_list, uint32_t* __restrict__ n_over) {
    uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= n) return;
    const uint32_t len = aln[id].cigar_len, off = cig_off[id];
    aln[id].cigar_off = off;
    if (len > kCigTmp) { over_list[atomicAdd(n_over, 1u)] = id; return; }
    for (uint32_t t = 0; t < len; t++) cigar[off + t] = cig_tmp[(size_t)id * kCigTmp + (len - 1 - t)];
}

template <int NW>
cudaError_t run_align(bm_ctx* c, AlignArgs& a, uint32_t n_aln, uint32_t max_cols) {
    size_t free_b = 0, total_b = 0;
    cudaError_t e = cudaMemGetInfo(&free_b, &total_b);
    if (e != cudaSuccess) return e;
    size_t budget = free_b / 2 + c->d_aln_scratch.cap;
    if (budget > ((size_t)16 << 30)) budget = (size_t)16 << 30;
    // ---- band kernel on everything, in chunks sized by the scratch budget ----
    e = c->d_loc_over.ensure((size_t)(n_aln + 2) * 4 + 64);
    if (e != cudaSuccess) return e;
    a.band_over = c->d_loc_over.as<uint32_t>();
    a.n_band_over = a.band_over + n_aln;
    e = cudaMemsetAsync(a.n_band_over, 0, 4, c->stream);
    if (e != cudaSuccess) return e;
    {
        const size_t per = (size_t)kBandPlanes * 8 * max_cols;
        uint32_t chunk = (uint32_t)std::min<size_t>(n_aln, std::max<size_t>(budget / per, kAlnThreads));
        chunk = (chunk + kAlnThreads - 1) / kAlnThreads * kAlnThreads;
        e = c->d_aln_scratch.ensure((size_t)chunk * per);
        if (e != cudaSuccess) return e;
        a.scratch = c->d_aln_scratch.as<unsigned long long>();
        a.chunk_n = chunk; a.max_cols = max_cols;
        const size_t smem = (size_t)4 * NW * kAlnThreads * 8;
        if (smem > 48 * 1024) {
            e = cudaFuncSetAttribute(k_align_band<NW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e != cudaSuccess) return e;
        }
        for (uint32_t first = 0; first < n_aln; first += chunk) {
            const uint32_t cnt = std::min(chunk, n_aln - first);
            KScope ks(c, BM_K_ALIGN);
            k_align_band<NW><<<(cnt + kAlnThreads - 1) / kAlnThreads, kAlnThreads, smem, c->stream>>>(a, first, cnt);
        }
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    // ---- what left the band (more than E edits): whole-column records ----
    uint32_t n_over = 0;
    e = cudaMemcpyAsync(&n_over, a.n_band_over, 4, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) return e;
    const size_t per = (size_t)Rec<NW>::kWords * 8 * max_cols;
    uint32_t chunk = (uint32_t)std::min<size_t>(n_over ? n_over : 1, std::max<size_t>(budget / per, kAlnThreads));
    chunk = (chunk + kAlnThreads - 1) / kAlnThreads * kAlnThreads;
    e = c->d_aln_scratch.ensure((size_t)chunk * per);            // also what the CIGAR overflow pass (MODE 1) uses
    if (e != cudaSuccess) return e;
    a.scratch = c->d_aln_scratch.as<unsigned long long>();
    a.chunk_n = chunk;
    a.id_list = a.band_over;
    for (uint32_t first = 0; first < n_over; first += chunk) {
        const uint32_t cnt = std::min(chunk, n_over - first);
        KScope ks(c, BM_K_ALIGN);
        k_align<NW, 0><<<(cnt + kAlnThreads - 1) / kAlnThreads, kAlnThreads, 0, c->stream>>>(a, first, cnt);
    }
    a.id_list = nullptr;
    return cudaGetLastError();
}

template <int NW>
cudaError_t run_align_over(bm_ctx* c, AlignArgs& a, uint32_t n_over) {
    for (uint32_t first = 0; first < n_over; first += a.chunk_n) {
        const uint32_t cnt = std::min(a.chunk_n, n_over - first);
        KScope ks(c, BM_K_ALIGN);
        k_align<NW, 1><<<(cnt + kAlnThreads - 1) / kAlnThreads, kAlnThreads, 0, c->stream>>>(a, first, cnt);
    }
    return cudaGetLastError();
}

}  // namespace

extern "C" {

int bm_align_device(bm_ctx* c, const bm_params* p) {
    if (!c) return BM_E_ARG;
    if (!p) BM_FAIL(c, BM_E_ARG, "bm_align_device: null parameters");
    if (!c->have_locs || c->locs_filtered)
        BM_FAIL(c, BM_E_ARG, "bm_align_device: bm_locate_device(filter_best=0) first (bucketmap_align verifies every location)");
    cudaSetDevice(c->device);
    c->have_aln = false;
    uint64_t n64 = 0;
    int r = bm_locate_count(c, &n64);
    if (r) return r;
    const uint32_t n_aln = (uint32_t)n64;
    c->n_aln = n_aln; c->n_cigar = 0;
    BM_CK(c, c->d_aln.ensure((size_t)(n_aln + 1) * sizeof(bm_aln)));
    BM_CK(c, c->d_cigar_cnt.ensure((size_t)(n_aln + 1) * 4));
    BM_CK(c, c->d_cigar_off.ensure((size_t)(n_aln + 1) * 4));
    if (n_aln == 0) { c->have_aln = true; return BM_OK; }
    const uint32_t max_len = c->max_read_len;
    if (max_len > 1024) BM_FAIL(c, BM_E_LIMIT, "bm_align_device: reads longer than 1024 bases are not supported by the verifier");
    const uint32_t max_cols = max_len + 2 + (uint32_t)(p->indel_rate * (float)max_len) + 1;
    BM_CK(c, c->d_cigar_fixed.ensure((size_t)n_aln * kCigTmp * 4));
    BM_CK(c, cudaMemsetAsync((uint32_t*)c->d_cigar_cnt.p + n_aln, 0, 4, c->stream));
    AlignArgs a;
    a.packed = c->d_packed.as<uint8_t>(); a.bstart = c->d_bstart.as<uint64_t>(); a.blen = c->d_blen.as<uint32_t>();
    a.rp = ReadPlanes{c->d_pl.as<uint2>(), nullptr, c->d_pl_off.as<uint32_t>()};
    a.read_off = c->d_read_off.as<uint64_t>(); a.n_reads = c->n_reads;
    a.loc_off = c->d_loc_off.as<uint32_t>(); a.locs = c->d_locs.as<bm_loc>();
    a.indel_rate = p->indel_rate; a.quality_threshold = p->quality_threshold;
    a.scratch = nullptr; a.chunk_n = 0;
    a.aln = c->d_aln.as<bm_aln>(); a.cig_tmp = c->d_cigar_fixed.as<uint32_t>(); a.cig_cnt = c->d_cigar_cnt.as<uint32_t>();
    a.cig_off = c->d_cigar_off.as<uint32_t>(); a.cigar = nullptr; a.over_list = nullptr;
    a.id_list = nullptr; a.band_over = nullptr; a.n_band_over = nullptr;
    const uint32_t nw = (max_len + 63) / 64;
    cudaError_t e;
    if (nw <= 1) e = run_align<1>(c, a, n_aln, max_cols);
    else if (nw <= 2) e = run_align<2>(c, a, n_aln, max_cols);
    else if (nw <= 3) e = run_align<3>(c, a, n_aln, max_cols);
    else if (nw <= 4) e = run_align<4>(c, a, n_aln, max_cols);
    else if (nw <= 5) e = run_align<5>(c, a, n_aln, max_cols);
    else if (nw <= 8) e = run_align<8>(c, a, n_aln, max_cols);
    else e = run_align<16>(c, a, n_aln, max_cols);
    BM_CK(c, e);
    r = bm_internal_exclusive_scan(c, c->d_cigar_cnt.as<uint32_t>(), c->d_cigar_off.as<uint32_t>(), n_aln + 1);
    if (r) return r;
    uint32_t total = 0;
    BM_CK(c, cudaMemcpyAsync(&total, c->d_cigar_off.as<uint32_t>() + n_aln, 4, cudaMemcpyDeviceToHost, c->stream));
    BM_CK(c, cudaStreamSynchronize(c->stream));
    c->n_cigar = total;
    BM_CK(c, c->d_cigar.ensure((size_t)total * 4 + 16));
    BM_CK(c, c->d_vote_scratch.ensure((size_t)(n_aln + 2) * 4));
    uint32_t* over_list = c->d_vote_scratch.as<uint32_t>();
    uint32_t* n_over = over_list + n_aln;
    BM_CK(c, cudaMemsetAsync(n_over, 0, 4, c->stream));
    a.cigar = c->d_cigar.as<uint32_t>(); a.over_list = over_list;
    {
        KScope ks(c, BM_K_AUX);
        k_cigar_write<<<(n_aln + 255) / 256, 256, 0, c->stream>>>(c->d_aln.as<bm_aln>(), c->d_cigar_off.as<uint32_t>(),
                                                                   c->d_cigar_fixed.as<uint32_t>(), n_aln,
                                                                   c->d_cigar.as<uint32_t>(), over_list, n_over);
    }
    BM_CK(c, cudaGetLastError());
    uint32_t h_over = 0;
    BM_CK(c, cudaMemcpyAsync(&h_over, n_over, 4, cudaMemcpyDeviceToHost, c->stream));
    BM_CK(c, cudaStreamSynchronize(c->stream));
    if (h_over) {
        if (nw <= 1) e = run_align_over<1>(c, a, h_over);
        else if (nw <= 2) e = run_align_over<2>(c, a, h_over);
        else if (nw <= 3) e = run_align_over<3>(c, a, h_over);
        else if (nw <= 4) e = run_align_over<4>(c, a, h_over);
        else if (nw <= 5) e = run_align_over<5>(c, a, h_over);
        else if (nw <= 8) e = run_align_over<8>(c, a, h_over);
        else e = run_align_over<16>(c, a, h_over);
        BM_CK(c, e);
    }
    c->have_aln = true;
    return BM_OK;
}

int bm_align_download(bm_ctx* c, bm_aln* aln, uint32_t* cigar, uint64_t cigar_cap, uint64_t* cigar_used) {
    if (!c) return BM_E_ARG;
    if (!c->have_aln) BM_FAIL(c, BM_E_ARG, "bm_align_download: bm_align_device first");
    cudaSetDevice(c->device);
    if (cigar_used) *cigar_used = c->n_cigar;
    if (aln && c->n_aln) BM_CK(c, cudaMemcpyAsync(aln, c->d_aln.p, (size_t)c->n_aln * sizeof(bm_aln), cudaMemcpyDeviceToHost, c->stream));
    if (cigar && c->n_cigar) {
        if (cigar_cap < c->n_cigar) BM_FAIL(c, BM_E_ARG, "bm_align_download: cigar buffer too small");
        BM_CK(c, cudaMemcpyAsync(cigar, c->d_cigar.p, (size_t)c->n_cigar * 4, cudaMemcpyDeviceToHost, c->stream));
    }
    BM_CK(c, cudaStreamSynchronize(c->stream));
    return BM_OK;
}

}  // extern "C"

// ---- bm_align_noalloc: K4 without a single allocation -------------------------------------------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops, harnesses that forbid it).  It is k_align<NW, 0> -- the whole-column verifier
// above, unchanged -- with its traceback records in a file-scope device array instead of ctx's grown-on-demand buffer; inputs
// and outputs are the caller's device pointers.  Every GPU that loads the library holds the array until the process exits.
//
// All calls share the array: two calls that overlap on the GPU (different contexts, lanes or streams) race.  Nothing in it has
// to start at zero -- a thread's traceback reads only records (columns 0..n of its own slot) that the same thread stored
// earlier in the same launch -- so the previous call's contents are simply overwritten.
namespace {

constexpr size_t kNoAllocWords = (size_t)BM_NOALLOC_SCRATCH_BYTES / 8;
static __device__ __align__(256) unsigned long long g_noalloc_scratch[kNoAllocWords];

// The array's name is a host-side stand-in in host code, not a device address: the address the kernels get is asked for once
// per device, with that device current.
constexpr int kNoAllocMaxDevices = 64;
std::mutex g_noalloc_mu;
unsigned long long* g_noalloc_ptr[kNoAllocMaxDevices] = {nullptr};

cudaError_t noalloc_scratch(int device, unsigned long long** out) {
    if (device < 0 || device >= kNoAllocMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        g_noalloc_ptr[device] = static_cast<unsigned long long*>(p);
    }
    *out = g_noalloc_ptr[device];
    return cudaSuccess;
}

// ops of alignment id, first op first, at cigar[id * BM_NOALLOC_CIGAR_OPS ..]; k_align leaves them last op first
__global__ void k_noalloc_finish(bm_aln* __restrict__ aln, uint32_t* __restrict__ cigar, uint32_t n) {
    const uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= n) return;
    const uint32_t len = aln[id].cigar_len;
    aln[id].cigar_off = id * kCigTmp;
    if (len > kCigTmp) return;                 // more runs than a slot holds: the count is reported, the ops are not
    uint32_t* s = cigar + (size_t)id * kCigTmp;
    for (uint32_t a = 0, b = len; a + 1 < b; a++) { --b; const uint32_t t = s[a]; s[a] = s[b]; s[b] = t; }
}

template <int NW>
cudaError_t noalloc_launch(cudaStream_t st, AlignArgs& a, uint32_t n_aln, uint32_t max_cols, size_t* need, uint32_t* most) {
    const uint32_t padded = (n_aln + kAlnThreads - 1) / kAlnThreads * kAlnThreads;
    const size_t per = (size_t)Rec<NW>::kWords * 8 * max_cols;
    *need = (size_t)padded * per;
    *most = (uint32_t)((size_t)BM_NOALLOC_SCRATCH_BYTES / per / kAlnThreads * kAlnThreads);
    if (*need > (size_t)BM_NOALLOC_SCRATCH_BYTES) return cudaErrorInvalidValue;
    a.chunk_n = padded; a.max_cols = max_cols;
    k_align<NW, 0><<<padded / kAlnThreads, kAlnThreads, 0, st>>>(a, 0u, n_aln);
    return cudaGetLastError();
}

}  // namespace

extern "C" int bm_align_noalloc(bm_ctx* c, const bm_params* p, const uint8_t* d_packed, const uint64_t* d_bstart,
                                const uint32_t* d_blen, const uint64_t* d_planes, const uint32_t* d_pl_off,
                                const uint64_t* d_read_off, uint32_t n_reads, uint32_t max_read_len,
                                const uint32_t* d_loc_off, const bm_loc* d_locs, uint32_t n_aln,
                                bm_aln* d_aln, uint32_t* d_cigar, uint32_t* d_cigar_cnt) {
    if (!c) return BM_E_ARG;
    if (!p || !d_packed || !d_bstart || !d_blen || !d_planes || !d_pl_off || !d_read_off || !d_loc_off || !d_locs ||
        !d_aln || !d_cigar || !d_cigar_cnt)
        BM_FAIL(c, BM_E_ARG, "bm_align_noalloc: null argument");
    if (n_aln == 0) return BM_OK;
    if (n_reads == 0 || max_read_len == 0) BM_FAIL(c, BM_E_ARG, "bm_align_noalloc: locations without reads");
    if (max_read_len > 1024) BM_FAIL(c, BM_E_LIMIT, "bm_align_noalloc: reads longer than 1024 bases are not supported by the verifier");
    BM_CK(c, cudaSetDevice(c->device));
    AlignArgs a;
    BM_CK(c, noalloc_scratch(c->device, &a.scratch));
    a.packed = d_packed; a.bstart = d_bstart; a.blen = d_blen;
    a.rp = ReadPlanes{reinterpret_cast<const uint2*>(d_planes), nullptr, d_pl_off};
    a.read_off = d_read_off; a.n_reads = n_reads;
    a.loc_off = d_loc_off; a.locs = d_locs;
    a.indel_rate = p->indel_rate; a.quality_threshold = p->quality_threshold;
    a.aln = d_aln; a.cig_tmp = d_cigar; a.cig_cnt = d_cigar_cnt;
    a.cig_off = nullptr; a.cigar = nullptr; a.over_list = nullptr; a.id_list = nullptr;
    a.band_over = nullptr; a.n_band_over = nullptr;
    const uint32_t max_cols = max_read_len + 2 + (uint32_t)(p->indel_rate * (float)max_read_len) + 1;
    const uint32_t nw = (max_read_len + 63) / 64;
    size_t need = 0;
    uint32_t most = 0;
    cudaError_t e;
    if (nw <= 1) e = noalloc_launch<1>(c->stream, a, n_aln, max_cols, &need, &most);
    else if (nw <= 2) e = noalloc_launch<2>(c->stream, a, n_aln, max_cols, &need, &most);
    else if (nw <= 3) e = noalloc_launch<3>(c->stream, a, n_aln, max_cols, &need, &most);
    else if (nw <= 4) e = noalloc_launch<4>(c->stream, a, n_aln, max_cols, &need, &most);
    else if (nw <= 5) e = noalloc_launch<5>(c->stream, a, n_aln, max_cols, &need, &most);
    else if (nw <= 8) e = noalloc_launch<8>(c->stream, a, n_aln, max_cols, &need, &most);
    else e = noalloc_launch<16>(c->stream, a, n_aln, max_cols, &need, &most);
    if (need > (size_t)BM_NOALLOC_SCRATCH_BYTES) {
        char msg[256];
        snprintf(msg, sizeof msg, "bm_align_noalloc: %u alignments of reads up to %u bases need %zu bytes of traceback records, "
                 "the static array holds %u: at most %u alignments per call at this read length", n_aln, max_read_len, need,
                 (unsigned)BM_NOALLOC_SCRATCH_BYTES, most);
        BM_FAIL(c, BM_E_LIMIT, msg);
    }
    BM_CK(c, e);
    k_noalloc_finish<<<(n_aln + 255) / 256, 256, 0, c->stream>>>(d_aln, d_cigar, n_aln);
    BM_CK(c, cudaGetLastError());
    return BM_OK;
}
