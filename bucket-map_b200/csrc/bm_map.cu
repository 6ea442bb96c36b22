// bm_map.cu -- the mapper: k-mer filtering + sampling (K2a) and the gather / miss-count vote (K2).
//
// Replaces q_gram_mapper::query_sequence / query and fault_tolerate_filter (mapper/q_gram_mapper.h:39-102,380-480).
//
// Closed form of the reference's cascade (q_gram_mapper.h:75-102): after all S samples, filters[F-1-j] holds the
// buckets with <= j missed samples, and best_results() returns the set bits of the highest non-empty level.  That is
// "all buckets whose miss count equals the minimum miss count m*, if m* <= F-1, else nothing".  K2 therefore keeps
// one saturating bit-sliced miss counter per bucket (NP bit planes, 128 buckets per lane) instead of F full bitsets.
//
// K2 work split: one warp per (segment, strand).  A warp walks the row in 512-B column tiles (32 lanes x 16 B); per
// tile every lane issues the S*(k-q+1) independent 128-bit loads of its column, B samples at a time, ANDs the rows of
// one sample, and adds the complement to its counters.  Everything after the loads is register-only; cross-lane work
// is one ballot per tile (is any bucket still under F misses?) plus, for the rare tiles that hold a contender, a
// warp min / prefix sum.  No shared-memory traffic on the critical path, no block-level barrier.
#include <cub/device/device_scan.cuh>

#include "bm_internal.cuh"

namespace {

constexpr int kWarpsPerCta = 4;

// ---- K2a: good k-mers -> S samples -> row ids of both strands ------------------------------------------------------------
// q_gram_mapper.h:431-469.  rowidx[((2*seg+strand)*S + s)*NQ + i] = kmer_to_index[(h_s >> 2i) & Q]  (or -1)
__global__ void __launch_bounds__(kWarpsPerCta * 32)
k_map_sample(const uint8_t* __restrict__ seq, const uint8_t* __restrict__ qual, const uint64_t* __restrict__ read_off,
             const uint32_t* __restrict__ seg_read, const uint32_t* __restrict__ seg_start,
             const uint32_t* __restrict__ seg_len, uint32_t seg0, uint32_t n_seg, uint32_t max_seg_len,
             const int32_t* __restrict__ k2i, const uint32_t* __restrict__ distok, uint32_t k, uint32_t q, uint32_t S,
             uint32_t minq, uint32_t min_good, uint32_t ktable, int32_t* __restrict__ rowidx, uint8_t* __restrict__ segok,
             unsigned long long* __restrict__ stats) {
    extern __shared__ __align__(16) uint8_t smem[];
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t seg = seg0 + blockIdx.x * kWarpsPerCta + warp;   // segments [seg0, n_seg) of the batch
    if (seg >= n_seg) return;
    const uint32_t nchunk_max = (max_seg_len + 31) / 32 + 1;
    const uint32_t per_warp = ((2 * max_seg_len + 3) & ~3u) + nchunk_max * 4;
    SegStage st;
    st.base = smem + (size_t)warp * per_warp;
    st.qual = st.base + max_seg_len;
    st.mask = reinterpret_cast<uint32_t*>(st.base + ((2 * max_seg_len + 3) & ~3u));

    const uint32_t n = seg_len[seg];
    const uint64_t off = read_off[seg_read[seg]] + seg_start[seg];
    stage_segment(st, seq, qual, off, n, lane);

    const uint32_t NQ = k - q + 1, Q = (1u << (2 * q)) - 1u;
    const uint32_t nk = (n >= k) ? n - k + 1 : 0;
    const uint32_t nchunk = (nk + 31) / 32;
    uint32_t G = 0;
    for (uint32_t c = 0; c < nchunk; c++) {
        uint32_t j = c * 32 + lane;
        bool good = false;
        if (j < nk) {
            uint32_t h = kmer_hash_at(st.base, j, k);
            bool dist = false;                                    // q_gram_mapper.h:189-196
            for (uint32_t i = 0; i < NQ; i++) {
                uint32_t g = (h >> (2 * i)) & Q;
                dist |= (__ldg(distok + (g >> 5)) >> (g & 31)) & 1u;
            }
            good = dist && kmer_qual_at(st.qual, j, k) >= minq;   // :437-442
        }
        uint32_t m = __ballot_sync(0xFFFFFFFFu, good);
        if (lane == 0) st.mask[c] = m;
        G += __popc(m);
    }
    __syncwarp();
    const bool ok = G >= min_good && G > 0;                        // :445-447
    if (lane == 0) segok[seg] = ok ? 1 : 0;
    if (!ok) return;
    uint32_t gathered = 0;
    for (uint32_t s = lane; s < S; s += 32) {
        uint32_t pos = select_good(st.mask, nchunk, sample_index(s, S, G));   // :457-460
        uint32_t h = kmer_hash_at(st.base, pos, k);
        uint32_t hr = bm_revcomp_hash(h, k);                                  // :465-468
        const uint32_t NO = ktable ? 1u : NQ;                                // row ids written per sample
        int32_t* o0 = rowidx + ((size_t)(2 * seg) * S + s) * NO;
        int32_t* o1 = rowidx + ((size_t)(2 * seg + 1) * S + s) * NO;
        if (ktable) { o0[0] = (int32_t)h; o1[0] = (int32_t)hr; }             // k-mer row table: the hash is the row id
        for (uint32_t i = 0; i < NQ; i++) {
            int32_t a = __ldg(k2i + ((h >> (2 * i)) & Q)), b = __ldg(k2i + ((hr >> (2 * i)) & Q));
            if (!ktable) { o0[i] = a; o1[i] = b; }
            gathered += (a >= 0) + (b >= 0);
        }
    }
    for (int o = 16; o; o >>= 1) gathered += __shfl_xor_sync(0xFFFFFFFFu, gathered, o);
    if (lane == 0 && gathered) atomicAdd(stats + 3, (unsigned long long)gathered);
}

// ---- K2: gather + vote --------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 and4(uint4 a, uint4 b) { return make_uint4(a.x & b.x, a.y & b.y, a.z & b.z, a.w & b.w); }
__device__ __forceinline__ uint4 or4(uint4 a, uint4 b) { return make_uint4(a.x | b.x, a.y | b.y, a.z | b.z, a.w | b.w); }
__device__ __forceinline__ uint4 xor4(uint4 a, uint4 b) { return make_uint4(a.x ^ b.x, a.y ^ b.y, a.z ^ b.z, a.w ^ b.w); }
__device__ __forceinline__ uint4 andn4(uint4 a, uint4 b) { return make_uint4(a.x & ~b.x, a.y & ~b.y, a.z & ~b.z, a.w & ~b.w); }
__device__ __forceinline__ uint4 not4(uint4 a) { return make_uint4(~a.x, ~a.y, ~a.z, ~a.w); }
__device__ __forceinline__ bool any4(uint4 a) { return (a.x | a.y | a.z | a.w) != 0u; }
__device__ __forceinline__ uint32_t popc4(uint4 a) { return __popc(a.x) + __popc(a.y) + __popc(a.z) + __popc(a.w); }

// bits of column `col` (128 buckets) that are real buckets (< NB)
__device__ __forceinline__ uint4 valid_bits(uint32_t col, uint32_t NB) {
    uint32_t lo = col * 128u;
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        uint32_t b = lo + 32u * i;
        w[i] = (b >= NB) ? 0u : (NB - b >= 32u ? 0xFFFFFFFFu : ((1u << (NB - b)) - 1u));
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
}

// one group of BB samples: BB*NQ independent 128-bit loads per lane, then the bit-sliced miss counters are advanced
template <int NQ, int NP, int BB>
__device__ __forceinline__ uint32_t vote_group(const uint4* __restrict__ colp, uint32_t Wp16, const int32_t* idx, bool load_ok,
                                               uint4 (&p)[NP]) {
    const uint4 ONES = make_uint4(~0u, ~0u, ~0u, ~0u);
    uint4 v[BB][NQ];
    uint32_t nloads = 0;
#pragma unroll
    for (int b = 0; b < BB; b++)
#pragma unroll
        for (int i = 0; i < NQ; i++) {
            const int32_t r = idx[b * NQ + i];
            v[b][i] = ONES;                               // un-sampled q-gram: skipped in the AND (:404-405)
            if (r >= 0 && load_ok) { v[b][i] = bm_ldg_stream(colp + (size_t)r * Wp16); nloads++; }
        }
#pragma unroll
    for (int b = 0; b < BB; b++) {
        uint4 bf = v[b][0];
#pragma unroll
        for (int i = 1; i < NQ; i++) bf = and4(bf, v[b][i]);
        uint4 carry = not4(bf);                           // 1 = this sample misses the bucket
#pragma unroll
        for (int i = 0; i < NP; i++) { uint4 t = and4(p[i], carry); p[i] = xor4(p[i], carry); carry = t; }
#pragma unroll
        for (int i = 0; i < NP; i++) p[i] = or4(p[i], carry);   // saturate at 2^NP-1 (>= F by dispatch)
    }
    return nloads;
}

// buckets of this lane's column whose miss count is still < F
template <int NP>
__device__ __forceinline__ uint4 below_f(const uint4 (&p)[NP], uint32_t F) {
    uint4 lt = make_uint4(0, 0, 0, 0), eq = make_uint4(~0u, ~0u, ~0u, ~0u);
#pragma unroll
    for (int i = NP - 1; i >= 0; i--) {
        if ((F >> i) & 1u) { lt = or4(lt, andn4(eq, p[i])); eq = and4(eq, p[i]); }
        else eq = andn4(eq, p[i]);
    }
    return lt;
}

// Dead-tile early exit.  A bucket that has missed F samples can never be reported (best_results() only looks at the F
// cascade levels, q_gram_mapper.h:90-102) and miss counts only grow, so once every bucket of a 4096-bucket tile has
// reached F misses the remaining samples' rows need not be gathered for that tile; inside a live tile, lanes whose 128
// buckets are all dead stop loading too.  On a random genome a 12-mer hits a foreign bucket with probability ~0.2 %, so
// almost every tile dies after F+1 or F+2 of the S samples: the kernel moves about half of the S*(k-q+1)*W bytes per
// list that the reference's cascade touches, with identical results.  Only the buckets attaining the MINIMUM miss count
// are reported, so once a finished tile holds a bucket with `best` misses the bar for later tiles drops to best+1.
template <int NQ, int NP>
__global__ void __launch_bounds__(kWarpsPerCta * 32)
k_map_vote(const uint4* __restrict__ rows, uint32_t Wp16, uint32_t NB, const int32_t* __restrict__ rowidx,
           const uint8_t* __restrict__ segok, uint32_t list0, uint32_t n_lists, uint32_t S, uint32_t F, uint32_t max_cand,
           uint32_t* __restrict__ cand_fixed, uint32_t* __restrict__ cand_cnt, unsigned long long* __restrict__ stats) {
    constexpr int B = (NQ >= 8) ? 1 : (8 / NQ);           // samples in flight per lane while no tile can be dead yet
    // One row per sample (k-mer row table): a tile is practically never dead after exactly T samples (that needs all
    // 4096 buckets to miss every one of them), so the (T+1)-th row is requested together with the first T.
    constexpr uint32_t EXTRA = (NQ == 1) ? 1u : 0u;
    extern __shared__ int32_t s_idx[];                    // kWarpsPerCta * S*NQ row ids
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t list = list0 + blockIdx.x * kWarpsPerCta + warp;   // = 2*segment + strand; lists [list0, n_lists)
    if (list >= n_lists) return;
    if (!segok[list >> 1]) {                              // q_gram_mapper.h:445-447: too few good k-mers
        if (lane == 0) cand_cnt[list] = 0;
        return;
    }
    int32_t* my_idx = s_idx + (size_t)warp * S * NQ;
    for (uint32_t t = lane; t < S * NQ; t += 32) my_idx[t] = rowidx[(size_t)list * S * NQ + t];
    __syncwarp();

    const uint4 ZERO = make_uint4(0, 0, 0, 0);
    uint32_t best = 0xFFFFFFFFu, count = 0, nloads = 0;
    uint32_t* out = cand_fixed + (size_t)list * max_cand;
    const uint32_t ntile = (Wp16 + 31) / 32;
    for (uint32_t tile = 0; tile < ntile; tile++) {
        const uint32_t col = tile * 32 + lane;
        const bool active = col < Wp16;
        const uint4* colp = rows + (active ? col : 0);
        // a bucket is out of the race once it has missed T samples: T = F, or best+1 once an earlier tile holds a
        // bucket that finished with `best` misses (only buckets attaining the minimum are reported)
        const uint32_t T = (best < F) ? best + 1u : F;
        const uint32_t blind = min(S, T + EXTRA);         // fewer than T samples seen: nothing can be out yet
        uint4 p[NP];
#pragma unroll
        for (int i = 0; i < NP; i++) p[i] = ZERO;
        uint32_t s = 0;
        if constexpr (B >= 8) for (; s + 8 <= blind; s += 8) nloads += vote_group<NQ, NP, 8>(colp, Wp16, my_idx + s * NQ, active, p);
        if constexpr (B >= 4) for (; s + 4 <= blind; s += 4) nloads += vote_group<NQ, NP, 4>(colp, Wp16, my_idx + s * NQ, active, p);
        if constexpr (B >= 2) for (; s + 2 <= blind; s += 2) nloads += vote_group<NQ, NP, 2>(colp, Wp16, my_idx + s * NQ, active, p);
        for (; s < blind; s++) nloads += vote_group<NQ, NP, 1>(colp, Wp16, my_idx + s * NQ, active, p);
        bool tile_alive = true;
        for (; s < S; s++) {
            const bool lane_alive = active && any4(below_f<NP>(p, T));
            if (!__any_sync(0xFFFFFFFFu, lane_alive)) { tile_alive = false; break; }
            nloads += vote_group<NQ, NP, 1>(colp, Wp16, my_idx + s * NQ, lane_alive, p);
        }
        if (!tile_alive) continue;
        // buckets with miss count < T
        const uint4 lt = below_f<NP>(p, T);
        uint4 alive = active ? and4(lt, valid_bits(col, NB)) : ZERO;
        if (!__any_sync(0xFFFFFFFFu, any4(alive))) continue;
        // minimum miss count inside this lane's 128 buckets, and the buckets attaining it
        uint32_t lvl = 0xFFFFu;
        uint4 cand = alive;
        if (any4(alive)) {
            lvl = 0;
#pragma unroll
            for (int i = NP - 1; i >= 0; i--) {
                uint4 t = andn4(cand, p[i]);
                if (any4(t)) cand = t; else lvl |= 1u << i;
            }
        }
        const uint32_t tmin = __reduce_min_sync(0xFFFFFFFFu, lvl);
        if (tmin < best) { best = tmin; count = 0; }
        if (tmin == best) {
            uint32_t n = (lvl == best) ? popc4(cand) : 0u;
            uint32_t incl = n;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o); if (lane >= o) incl += t; }
            uint32_t pos = count + incl - n;
            if (n) {
                uint32_t w[4] = {cand.x, cand.y, cand.z, cand.w};
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    uint32_t x = w[i];
                    while (x) {
                        uint32_t bit = __ffs(x) - 1; x &= x - 1;
                        if (pos < max_cand) out[pos] = col * 128u + 32u * i + bit;   // ascending, like _Find_next
                        pos++;
                    }
                }
            }
            count += __shfl_sync(0xFFFFFFFFu, incl, 31);
        }
    }
    // :471-476: a list longer than the cap is emptied, not truncated
    if (lane == 0) cand_cnt[list] = (best < F && count <= max_cand) ? count : 0u;
    nloads = __reduce_add_sync(0xFFFFFFFFu, nloads);
    if (lane == 0) atomicAdd(stats + 4, (unsigned long long)nloads * 16ull);   // bytes actually gathered
}

// ---- K2c: coarse vote over the k-mer table + exact count of the groups that pass ---------------------------------------------
// One warp per (segment, strand).  Lane w holds coarse word w of the row (32 groups = 1,024 buckets); the S sample rows
// are S independent, fully coalesced 4-B loads per lane (one 128-B line per sample on the 1.7 Gbp reference), all in
// flight together.  The same bit-sliced saturating counter as K2 counts COARSE misses per group; a bucket with < F
// misses lies in a group with < F coarse misses (a coarse hit is the OR of its buckets' hits), so only passing groups
// can hold a reportable bucket.  For those the exact per-bucket miss counts come from the .qgram matrix itself: lane s
// ANDs the k-q+1 32-bit words of sample s (q_gram_mapper.h:399-407), 32 ballots transpose them into one count per lane.
// Reported set, order and the > max_candidates rule are those of best_results() / query_sequence (:90-102, :471-476).
// up to BB samples at once: every load is issued before the first counter update waits for one
template <int NP, int BB>
__device__ __forceinline__ void coarse_group(const uint32_t* __restrict__ colp, uint32_t Wcs, const int32_t* h, uint32_t n,
                                             bool act, uint32_t (&p)[NP]) {
    uint32_t v[BB];
#pragma unroll
    for (int b = 0; b < BB; b++) {
        v[b] = 0xFFFFFFFFu;                                // b >= n: no sample, nothing missed
        if (b < n) v[b] = act ? __ldg(colp + (size_t)(uint32_t)h[b] * Wcs) : 0u;
    }
#pragma unroll
    for (int b = 0; b < BB; b++) {
        uint32_t carry = ~v[b];                            // 1 = no bucket of the group holds this sample
#pragma unroll
        for (int i = 0; i < NP; i++) { uint32_t t = p[i] & carry; p[i] ^= carry; carry = t; }
#pragma unroll
        for (int i = 0; i < NP; i++) p[i] |= carry;        // saturate at 2^NP-1 (>= F by dispatch)
    }
}

template <int NP>
__global__ void __launch_bounds__(kWarpsPerCta * 32)
k_map_coarse(const uint32_t* __restrict__ crows, uint32_t Wcs, uint32_t Wc, uint32_t gshift, const uint32_t* __restrict__ rows32,
             uint32_t Wp32, uint32_t NB, const int32_t* __restrict__ k2i, uint32_t NQ, uint32_t Q,
             const int32_t* __restrict__ khash, const uint8_t* __restrict__ segok, uint32_t list0, uint32_t n_lists,
             uint32_t S, uint32_t F, uint32_t max_cand, uint32_t* __restrict__ cand_fixed, uint32_t* __restrict__ cand_cnt,
             unsigned long long* __restrict__ stats) {
    extern __shared__ int32_t s_idx[];                    // kWarpsPerCta * S sample hashes
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t list = list0 + blockIdx.x * kWarpsPerCta + warp;   // = 2*segment + strand
    if (list >= n_lists) return;
    if (!segok[list >> 1]) {                              // q_gram_mapper.h:445-447: too few good k-mers
        if (lane == 0) cand_cnt[list] = 0;
        return;
    }
    int32_t* my_h = s_idx + (size_t)warp * S;
    for (uint32_t t = lane; t < S; t += 32) my_h[t] = khash[(size_t)list * S + t];
    __syncwarp();

    uint32_t best = 0xFFFFFFFFu, count = 0, nfine = 0;
    uint32_t* out = cand_fixed + (size_t)list * max_cand;
    for (uint32_t w0 = 0; w0 < Wc; w0 += 32) {
        const uint32_t w = w0 + lane;
        const bool act = w < Wc;
        const uint32_t* colp = crows + (act ? w : 0);
        uint32_t p[NP];
#pragma unroll
        for (int i = 0; i < NP; i++) p[i] = 0u;
        for (uint32_t s = 0; s < S; s += 16) coarse_group<NP, 16>(colp, Wcs, my_h + s, min(16u, S - s), act, p);
        // groups with fewer than F coarse misses
        uint32_t lt = 0u, eq = 0xFFFFFFFFu;
#pragma unroll
        for (int i = NP - 1; i >= 0; i--) {
            if ((F >> i) & 1u) { lt |= eq & ~p[i]; eq &= p[i]; }
            else eq &= ~p[i];
        }
        const uint32_t pass = act ? lt : 0u;
        uint32_t any = __ballot_sync(0xFFFFFFFFu, pass != 0u);
        while (any) {
            const uint32_t l = __ffs(any) - 1; any &= any - 1;
            uint32_t pw = __shfl_sync(0xFFFFFFFFu, pass, l);
            uint32_t last_gi = 0xFFFFFFFFu;
            while (pw) {
                const uint32_t g = __ffs(pw) - 1; pw &= pw - 1;
                // coarse bit -> the 32-bit word of a full row that holds its buckets (32*gi .. 32*gi+31); with 8 or 16
                // buckets per coarse bit several bits share a word, which is evaluated once
                const uint32_t gi = ((w0 + l) * 32 + g) >> gshift;
                if (gi >= Wp32 || gi == last_gi) continue;
                last_gi = gi;
                const uint32_t b0 = gi * 32u;
                const uint32_t vmask = (b0 >= NB) ? 0u : (NB - b0 >= 32u ? 0xFFFFFFFFu : ((1u << (NB - b0)) - 1u));
                uint32_t hits = 0;                         // lane b: samples that hit bucket b0 + b
                for (uint32_t sb = 0; sb < S; sb += 32) {
                    const uint32_t si = sb + lane;
                    uint32_t word = 0u;
                    if (si < S) {
                        const uint32_t h = (uint32_t)my_h[si];
                        word = vmask;
                        for (uint32_t i = 0; i < NQ; i++) {
                            const int32_t r = __ldg(k2i + ((h >> (2 * i)) & Q));
                            if (r >= 0) { word &= __ldg(rows32 + (size_t)r * Wp32 + gi); nfine++; }
                        }
                    }
#pragma unroll
                    for (uint32_t b = 0; b < 32; b++) {
                        const uint32_t cnt = __popc(__ballot_sync(0xFFFFFFFFu, (word >> b) & 1u));
                        if (lane == b) hits += cnt;
                    }
                }
                const uint32_t miss = S - hits;
                const uint32_t lvl = (b0 + lane < NB && miss < F) ? miss : 0xFFFFu;
                const uint32_t tmin = __reduce_min_sync(0xFFFFFFFFu, lvl);
                if (tmin >= F) continue;
                if (tmin < best) { best = tmin; count = 0; }
                if (tmin == best) {
                    const uint32_t m = __ballot_sync(0xFFFFFFFFu, lvl == best);
                    const uint32_t pos = count + __popc(m & ((1u << lane) - 1u));
                    if (lvl == best && pos < max_cand) out[pos] = b0 + lane;   // ascending, like _Find_next
                    count += __popc(m);
                }
            }
        }
    }
    // :471-476: a list longer than the cap is emptied, not truncated
    if (lane == 0) cand_cnt[list] = (best < F && count <= max_cand) ? count : 0u;
    nfine = __reduce_add_sync(0xFFFFFFFFu, nfine);
    if (lane == 0) atomicAdd(stats + 4, (unsigned long long)S * Wc * 4ull + (unsigned long long)nfine * 4ull);   // bytes gathered
}

// ---- compaction of the fixed-slot lists into cand_off / cand ---------------------------------------------------------------
__global__ void k_cand_compact(const uint32_t* __restrict__ cand_fixed, const uint32_t* __restrict__ cand_cnt,
                               const uint32_t* __restrict__ cand_off, uint32_t n_lists, uint32_t max_cand,
                               uint32_t* __restrict__ cand, uint32_t* __restrict__ cand_list,
                               unsigned long long* __restrict__ stats) {
    uint32_t seg = blockIdx.x * blockDim.x + threadIdx.x;
    if (2 * seg >= n_lists) return;
    uint32_t tot = 0;
    for (uint32_t d = 0; d < 2; d++) {
        uint32_t l = 2 * seg + d, n = cand_cnt[l], o = cand_off[l];
        for (uint32_t i = 0; i < n; i++) { cand[o + i] = cand_fixed[(size_t)l * max_cand + i]; cand_list[o + i] = l; }
        if (n) atomicAdd(stats + 1 + d, (unsigned long long)n);
        tot += n;
    }
    if (tot) atomicAdd(stats + 0, 1ull);
}

template <int NQ>
cudaError_t launch_vote_np(bm_ctx* c, const bm_params* p, uint32_t NP, uint32_t list0, uint32_t n_lists, uint32_t blocks, size_t smem) {
    const uint4* rows = c->d_rows.as<uint4>();
#define BM_VOTE(NPv)                                                                                               \
    k_map_vote<NQ, NPv><<<blocks, kWarpsPerCta * 32, smem, c->stream>>>(                                           \
        rows, c->Wp / 16, c->NB, c->d_rowidx.as<int32_t>(), c->d_segok.as<uint8_t>(), list0, n_lists, \
        p->mapper_samples, p->mapper_fault, p->max_candidates, c->d_cand_fixed.as<uint32_t>(),                     \
        c->d_cand_cnt.as<uint32_t>(), c->d_stats.as<unsigned long long>())
    if (NP == 3) BM_VOTE(3); else if (NP == 4) BM_VOTE(4); else BM_VOTE(6);
#undef BM_VOTE
    return cudaGetLastError();
}

}  // namespace

extern "C" {

int bm_map_device(bm_ctx* c, const bm_params* p) {
    if (!c) return BM_E_ARG;
    if (!p) BM_FAIL(c, BM_E_ARG, "bm_map_device: null parameters");
    if (!c->have_reads) BM_FAIL(c, BM_E_ARG, "bm_map_device: no read batch resident");
    int r = bm_internal_map_begin(c, p, c->n_seg, c->max_seg_len);
    if (r) return r;
    r = bm_internal_map_range(c, p, 0, c->n_seg, c->max_seg_len);
    if (r) return r;
    return bm_internal_map_finish(c, p);
}

int bm_map_download(bm_ctx* c, uint32_t* cand_off, uint32_t* cand, uint64_t cand_cap, uint64_t* stats) {
    if (!c) return BM_E_ARG;
    if (!c->have_map) BM_FAIL(c, BM_E_ARG, "bm_map_download: bm_map_device first");
    cudaSetDevice(c->device);
    const uint32_t n_lists = 2 * c->n_seg;
    if (cand_off) BM_CK(c, cudaMemcpyAsync(cand_off, c->d_cand_off.p, (size_t)(n_lists + 1) * 4, cudaMemcpyDeviceToHost, c->stream));
    if (cand && c->n_cand) {
        if (cand_cap < c->n_cand) BM_FAIL(c, BM_E_ARG, "bm_map_download: candidate buffer too small");
        BM_CK(c, cudaMemcpyAsync(cand, c->d_cand.p, (size_t)c->n_cand * 4, cudaMemcpyDeviceToHost, c->stream));
    }
    if (stats) BM_CK(c, cudaMemcpyAsync(stats, c->d_stats.p, 8 * 8, cudaMemcpyDeviceToHost, c->stream));
    BM_CK(c, cudaStreamSynchronize(c->stream));
    return BM_OK;
}

int bm_map_reads(bm_ctx* c, const bm_params* p, const uint8_t* seq, const uint8_t* qual, const uint64_t* read_off,
                 uint32_t n_reads, uint32_t* cand_off, uint32_t* cand, uint64_t cand_cap, uint64_t* stats) {
    int r = bm_reads_upload(c, p, seq, qual, read_off, n_reads);
    if (r) return r;
    r = bm_map_device(c, p);
    if (r) return r;
    return bm_map_download(c, cand_off, cand, cand_cap, stats);
}

}  // extern "C"

// ---- the mapper in three steps ---------------------------------------------------------------------------------------------
int bm_internal_map_begin(bm_ctx* c, const bm_params* p, uint32_t n_seg_cap, uint32_t max_seg_len) {
    if (!c->have_index) BM_FAIL(c, BM_E_ARG, "bm_map_device: no index resident (the q-gram index is empty)");
    if (p->q != c->q) BM_FAIL(c, BM_E_ARG, "bm_map_device: index seed length differs from the resident index");
    if (p->k < p->q || p->k > 16) BM_FAIL(c, BM_E_LIMIT, "query seed length must satisfy q <= k <= 16");
    const uint32_t NQ = p->k - p->q + 1, S = p->mapper_samples, F = p->mapper_fault;
    if (NQ > 8) BM_FAIL(c, BM_E_LIMIT, "k - q must be <= 7");
    if (S == 0 || S > 256) BM_FAIL(c, BM_E_LIMIT, "mapper samples must be in [1,256]");
    if (F == 0 || F > 63) BM_FAIL(c, BM_E_LIMIT, "mapper fault-tolerance levels must be in [1,63]");
    if (p->max_candidates == 0 || p->max_candidates > 1024) BM_FAIL(c, BM_E_LIMIT, "max candidates must be in [1,1024]");
    if (max_seg_len > 8192) BM_FAIL(c, BM_E_LIMIT, "segments longer than 8192 bases are not supported (-r)");
    cudaSetDevice(c->device);
    c->have_map = false; c->have_locs = false; c->have_aln = false;
    int r = bm_internal_index_prepare(c, p->distinguishability);
    if (r) return r;
    r = bm_internal_kmer_table_prepare(c, p->k);          // decides c->use_ktable for this call
    if (r) return r;
    const size_t n_lists = 2 * (size_t)n_seg_cap;
    BM_CK(c, c->d_rowidx.ensure(n_lists * S * NQ * 4 + 16));
    BM_CK(c, c->d_segok.ensure((size_t)n_seg_cap + 16));
    BM_CK(c, c->d_cand_fixed.ensure(n_lists * p->max_candidates * 4 + 16));
    BM_CK(c, c->d_cand_cnt.ensure((n_lists + 1) * 4));
    BM_CK(c, c->d_cand_off.ensure((n_lists + 1) * 4));
    BM_CK(c, c->d_stats.ensure(128 * 8));
    BM_CK(c, cudaMemsetAsync(c->d_stats.p, 0, 8 * 8, c->stream));
    c->map_max_cand = p->max_candidates;
    c->n_cand = 0;
    c->lsamp_k = c->lsamp_P = c->lsamp_b = 0;
    if (c->use_ktable && p->locator_samples && p->locator_samples <= 1024) {
        // K2f also draws the locator's samples (same k-mers, quality filter only: bucket_locator.h:315-341)
        BM_CK(c, c->d_lsamp_idx.ensure((size_t)n_seg_cap * p->locator_samples * 2 + 16));
        BM_CK(c, c->d_lsamp_hash.ensure((size_t)n_seg_cap * p->locator_samples * 4 + 16));
        c->lsamp_k = p->k; c->lsamp_P = p->locator_samples; c->lsamp_b = p->avg_base_quality;
    }
    const uint32_t msl = max_seg_len ? max_seg_len : 1;
    size_t smem_a = (size_t)kWarpsPerCta * ((((size_t)2 * msl + 3) & ~(size_t)3) + ((msl + 31) / 32 + 1) * 4);
    if (smem_a > 48 * 1024)
        BM_CK(c, cudaFuncSetAttribute(k_map_sample, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_a));
    return BM_OK;
}

// K2a + K2 over segments [seg0, seg1); everything they read must already be resident (or ordered before on c->stream)
int bm_internal_map_range(bm_ctx* c, const bm_params* p, uint32_t seg0, uint32_t seg1, uint32_t max_seg_len) {
    if (seg1 <= seg0) return BM_OK;
    if (c->use_ktable && p->k <= 16 && p->mapper_samples <= 256) return bm_internal_map_fused(c, p, seg0, seg1, max_seg_len);
    const uint32_t NQ = p->k - p->q + 1, S = p->mapper_samples, F = p->mapper_fault;
    // smallest G with !(G < 0.2*S), the comparison of q_gram_mapper.h:445 in double
    uint32_t min_good = 0;
    while ((double)min_good < 0.2 * (double)S) min_good++;
    const uint32_t msl = max_seg_len ? max_seg_len : 1;
    size_t smem_a = (size_t)kWarpsPerCta * ((((size_t)2 * msl + 3) & ~(size_t)3) + ((msl + 31) / 32 + 1) * 4);
    const uint32_t blocks_a = (seg1 - seg0 + kWarpsPerCta - 1) / kWarpsPerCta;
    {
        KScope ks(c, BM_K_SAMPLE);
        k_map_sample<<<blocks_a, kWarpsPerCta * 32, smem_a, c->stream>>>(
            c->d_seq.as<uint8_t>(), c->d_qual.as<uint8_t>(), c->d_read_off.as<uint64_t>(),
            c->d_seg_read.as<uint32_t>(), c->d_seg_start.as<uint32_t>(), c->d_seg_len.as<uint32_t>(), seg0, seg1, msl,
            c->d_k2i.as<int32_t>(), c->d_distok.as<uint32_t>(), p->k, p->q, S, p->avg_base_quality * p->k, min_good,
            c->use_ktable ? 1u : 0u, c->d_rowidx.as<int32_t>(), c->d_segok.as<uint8_t>(), c->d_stats.as<unsigned long long>());
    }
    BM_CK(c, cudaGetLastError());
    const uint32_t NP = (F <= 7) ? 3 : (F <= 15 ? 4 : 6);
    const uint32_t NQv = c->use_ktable ? 1u : NQ;         // rows gathered per sample
    size_t smem_v = (size_t)kWarpsPerCta * S * NQv * 4;
    const uint32_t list0 = 2 * seg0, list1 = 2 * seg1;
    const uint32_t blocks_v = (list1 - list0 + kWarpsPerCta - 1) / kWarpsPerCta;
    cudaError_t e = cudaSuccess;
    if (c->use_ktable) {
        KScope ks(c, BM_K_MAP);
        const uint32_t Q = (p->q >= 16) ? 0xFFFFFFFFu : ((1u << (2 * p->q)) - 1u);
#define BM_COARSE(NPv)                                                                                              \
        k_map_coarse<NPv><<<blocks_v, kWarpsPerCta * 32, smem_v, c->stream>>>(                                      \
            c->d_krows.as<uint32_t>(), c->Wcs, c->Wc, 5u - c->krows_logg, c->d_rows.as<uint32_t>(), c->Wp / 4, c->NB, c->d_k2i.as<int32_t>(), \
            NQ, Q, c->d_rowidx.as<int32_t>(), c->d_segok.as<uint8_t>(), list0, list1, S, F, p->max_candidates,      \
            c->d_cand_fixed.as<uint32_t>(), c->d_cand_cnt.as<uint32_t>(), c->d_stats.as<unsigned long long>())
        if (NP == 3) BM_COARSE(3); else if (NP == 4) BM_COARSE(4); else BM_COARSE(6);
#undef BM_COARSE
        e = cudaGetLastError();
    } else {
        KScope ks(c, BM_K_MAP);
        switch (NQv) {
            case 1: e = launch_vote_np<1>(c, p, NP, list0, list1, blocks_v, smem_v); break;
            case 2: e = launch_vote_np<2>(c, p, NP, list0, list1, blocks_v, smem_v); break;
            case 3: e = launch_vote_np<3>(c, p, NP, list0, list1, blocks_v, smem_v); break;
            case 4: e = launch_vote_np<4>(c, p, NP, list0, list1, blocks_v, smem_v); break;
            case 5: e = launch_vote_np<5>(c, p, NP, list0, list1, blocks_v, smem_v); break;
            case 6: e = launch_vote_np<6>(c, p, NP, list0, list1, blocks_v, smem_v); break;
            case 7: e = launch_vote_np<7>(c, p, NP, list0, list1, blocks_v, smem_v); break;
            default: e = launch_vote_np<8>(c, p, NP, list0, list1, blocks_v, smem_v); break;
        }
    }
    BM_CK(c, e);
    return BM_OK;
}

// fixed-slot lists of all c->n_seg segments -> cand_off / cand
int bm_internal_map_finish(bm_ctx* c, const bm_params* p) {
    const uint32_t n_seg = c->n_seg, n_lists = 2 * n_seg;
    if (n_seg) {
        int r = bm_internal_exclusive_scan(c, c->d_cand_cnt.as<uint32_t>(), c->d_cand_off.as<uint32_t>(), n_lists + 1);
        if (r) return r;
        // the compacted arrays are at most max_candidates per list, but sizing them exactly needs the total
        uint32_t total = 0;
        BM_CK(c, cudaMemcpyAsync(&total, c->d_cand_off.as<uint32_t>() + n_lists, 4, cudaMemcpyDeviceToHost, c->stream));
        BM_CK(c, cudaStreamSynchronize(c->stream));
        c->n_cand = total;
        BM_CK(c, c->d_cand.ensure((size_t)total * 4 + 16));
        BM_CK(c, c->d_cseg.ensure((size_t)total * 4 + 16));
        {
            KScope ks(c, BM_K_AUX);
            k_cand_compact<<<(n_seg + 255) / 256, 256, 0, c->stream>>>(
                c->d_cand_fixed.as<uint32_t>(), c->d_cand_cnt.as<uint32_t>(), c->d_cand_off.as<uint32_t>(), n_lists,
                p->max_candidates, c->d_cand.as<uint32_t>(), c->d_cseg.as<uint32_t>(),
                c->d_stats.as<unsigned long long>());
        }
        BM_CK(c, cudaGetLastError());
    } else {
        BM_CK(c, cudaMemsetAsync(c->d_cand_off.p, 0, 4, c->stream));
    }
    c->have_map = true;
    return BM_OK;
}
