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
k_map_sample(const ReadPlanes rp, const uint32_t* __restrict__ seg_read, const uint32_t* __restrict__ seg_start,
             const uint32_t* __restrict__ seg_len, uint32_t seg0, uint32_t n_seg, uint32_t max_seg_len,
             const int32_t* __restrict__ k2i, const uint32_t* __restrict__ distok, uint32_t k, uint32_t q, uint32_t S,
             uint32_t min_good, uint32_t ktable, int32_t* __restrict__ rowidx, uint8_t* __restrict__ segok,
             unsigned long long* __restrict__ stats) {
    extern __shared__ __align__(16) uint8_t smem[];
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t seg = seg0 + blockIdx.x * kWarpsPerCta + warp;   // segments [seg0, n_seg) of the batch
    if (seg >= n_seg) return;
    const uint32_t nchunk_max = (max_seg_len + 31) / 32 + 1;
    const uint32_t per_warp = ((max_seg_len + 3) & ~3u) + 2 * nchunk_max * 4;
    SegStage st;
    st.base = smem + (size_t)warp * per_warp;
    st.qkm = reinterpret_cast<uint32_t*>(st.base + ((max_seg_len + 3) & ~3u));
    st.mask = st.qkm + nchunk_max;

    const uint32_t n = seg_len[seg], rd = seg_read[seg];
    stage_segment(st, rp, rp.pl_off[rd], rp.pl_off[rd + 1], seg_start[seg], n, lane);

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
            good = dist && kmer_qual_ok(st, j);                   // :437-442
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
__device__ __forceinline__ void vote_list(const uint4* __restrict__ rows, uint32_t Wp16, uint32_t NB,
                                          const int32_t* __restrict__ rowidx, const uint8_t* __restrict__ segok,
                                          uint32_t list, uint32_t S, uint32_t F, uint32_t max_cand,
                                          uint32_t* __restrict__ cand_fixed, uint32_t* __restrict__ cand_cnt,
                                          unsigned long long* __restrict__ stats, int32_t* s_idx) {
    constexpr int B = (NQ >= 8) ? 1 : (8 / NQ);           // samples in flight per lane while no tile can be dead yet
    // One row per sample (k-mer row table): a tile is practically never dead after exactly T samples (that needs all
    // 4096 buckets to miss every one of them), so the (T+1)-th row is requested together with the first T.
    constexpr uint32_t EXTRA = (NQ == 1) ? 1u : 0u;
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
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

// one warp per list of [list0, n_lists)
template <int NQ, int NP>
__global__ void __launch_bounds__(kWarpsPerCta * 32)
k_map_vote(const uint4* __restrict__ rows, uint32_t Wp16, uint32_t NB, const int32_t* __restrict__ rowidx,
           const uint8_t* __restrict__ segok, uint32_t list0, uint32_t n_lists, uint32_t S, uint32_t F, uint32_t max_cand,
           uint32_t* __restrict__ cand_fixed, uint32_t* __restrict__ cand_cnt, unsigned long long* __restrict__ stats) {
    extern __shared__ int32_t s_idx[];                    // kWarpsPerCta * S*NQ row ids
    const uint32_t list = list0 + blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5);   // = 2*segment + strand
    if (list >= n_lists) return;
    vote_list<NQ, NP>(rows, Wp16, NB, rowidx, segok, list, S, F, max_cand, cand_fixed, cand_cnt, stats, s_idx);
}

// the lists the fused kernel handed over (too many groups passed its coarse vote: repeats); fixed grid, the count is
// read on the device so that no host round trip is needed
template <int NQ, int NP>
__global__ void __launch_bounds__(kWarpsPerCta * 32)
k_map_vote_lists(const uint4* __restrict__ rows, uint32_t Wp16, uint32_t NB, const int32_t* __restrict__ rowidx,
                 const uint8_t* __restrict__ segok, const uint32_t* __restrict__ list_ids, const uint32_t* __restrict__ n_ids,
                 uint32_t S, uint32_t F, uint32_t max_cand, uint32_t* __restrict__ cand_fixed,
                 uint32_t* __restrict__ cand_cnt, unsigned long long* __restrict__ stats) {
    extern __shared__ int32_t s_idx[];
    const uint32_t n = *n_ids;
    for (uint32_t i = blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5); i < n; i += gridDim.x * kWarpsPerCta) {
        vote_list<NQ, NP>(rows, Wp16, NB, rowidx, segok, list_ids[i], S, F, max_cand, cand_fixed, cand_cnt, stats, s_idx);
        __syncwarp();
    }
}

// K2f spreads its three counters over 60 slots (stats[8 + 3s ..]); fold them into stats[3], stats[4], stats[6];
// stats[5] = lists K2f handed to the full-width gather
__global__ void k_stats_fold(unsigned long long* __restrict__ stats, const uint32_t* __restrict__ n_over) {
    unsigned long long a = 0, b = 0, c = 0;
    if (threadIdx.x < 60) {
        a = stats[8 + 3 * threadIdx.x]; b = stats[9 + 3 * threadIdx.x]; c = stats[10 + 3 * threadIdx.x];
        stats[8 + 3 * threadIdx.x] = 0; stats[9 + 3 * threadIdx.x] = 0; stats[10 + 3 * threadIdx.x] = 0;
    }
    for (int o = 16; o; o >>= 1) {
        a += __shfl_xor_sync(0xFFFFFFFFu, a, o); b += __shfl_xor_sync(0xFFFFFFFFu, b, o); c += __shfl_xor_sync(0xFFFFFFFFu, c, o);
    }
    if ((threadIdx.x & 31) == 0) { if (a) atomicAdd(stats + 3, a); if (b) atomicAdd(stats + 4, b); if (c) atomicAdd(stats + 6, c); }
    if (threadIdx.x == 0) stats[5] = *n_over;
}

// ---- compaction of the fixed-slot lists into cand_off / cand ---------------------------------------------------------------
__global__ void k_cand_compact(const uint32_t* __restrict__ cand_fixed, const uint32_t* __restrict__ cand_cnt,
                               const uint32_t* __restrict__ cand_off, uint32_t n_lists, uint32_t max_cand,
                               uint32_t* __restrict__ cand, uint32_t* __restrict__ cand_list,
                               unsigned long long* __restrict__ stats, uint32_t count_reads) {
    uint32_t seg = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t nd[2] = {0, 0};
    if (2 * seg < n_lists) {
        for (uint32_t d = 0; d < 2; d++) {
            uint32_t l = 2 * seg + d, n = cand_cnt[l], o = cand_off[l];
            for (uint32_t i = 0; i < n; i++) { cand[o + i] = cand_fixed[(size_t)l * max_cand + i]; cand_list[o + i] = l; }
            nd[d] = n;
        }
    }
    // the three counters of the batch: one atomic per warp, not per segment (3 M adds on one address cost ~0.5 ms)
    // stats[0] counts READS with a candidate (q_gram_mapper.h:527-545); a segment is a read unless the batch holds reads
    // longer than 2r (then k_count_mapped_reads counts them)
    const uint32_t s0 = count_reads ? __reduce_add_sync(0xFFFFFFFFu, (nd[0] + nd[1]) ? 1u : 0u) : 0u;
    const uint32_t s1 = __reduce_add_sync(0xFFFFFFFFu, nd[0]), s2 = __reduce_add_sync(0xFFFFFFFFu, nd[1]);
    if ((threadIdx.x & 31) == 0) {
        if (s0) atomicAdd(stats + 0, (unsigned long long)s0);
        if (s1) atomicAdd(stats + 1, (unsigned long long)s1);
        if (s2) atomicAdd(stats + 2, (unsigned long long)s2);
    }
}

// reads (not segments) with at least one candidate, for batches with long reads (5 segments per read)
__global__ void k_count_mapped_reads(const uint32_t* __restrict__ read_seg0, const uint32_t* __restrict__ cand_off,
                                     uint32_t n_reads, unsigned long long* __restrict__ stats) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    const bool has = r < n_reads && cand_off[2 * read_seg0[r + 1]] > cand_off[2 * read_seg0[r]];
    const uint32_t s0 = __reduce_add_sync(0xFFFFFFFFu, has ? 1u : 0u);
    if ((threadIdx.x & 31) == 0 && s0) atomicAdd(stats + 0, (unsigned long long)s0);
}

template <int NQ>
cudaError_t launch_vote_lists_np(bm_ctx* c, const bm_params* p, uint32_t NP, const uint32_t* ids, const uint32_t* n_ids,
                                 uint32_t blocks, size_t smem) {
#define BM_VOTEL(NPv)                                                                                              \
    k_map_vote_lists<NQ, NPv><<<blocks, kWarpsPerCta * 32, smem, c->stream>>>(                                     \
        c->d_rows.as<uint4>(), c->Wp / 16, c->NB, c->d_rowidx.as<int32_t>(), c->d_segok.as<uint8_t>(), ids, n_ids,  \
        p->mapper_samples, p->mapper_fault, p->max_candidates, c->d_cand_fixed.as<uint32_t>(),                     \
        c->d_cand_cnt.as<uint32_t>(), c->d_stats.as<unsigned long long>())
    if (NP == 3) BM_VOTEL(3); else if (NP == 4) BM_VOTEL(4); else BM_VOTEL(6);
#undef BM_VOTEL
    return cudaGetLastError();
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
    cudaSetDevice(c->device);
    int r = bm_internal_pack_check(c, p);                 // the k-mer quality plane is per (-l, -b)
    if (r) return r;
    r = bm_internal_map_begin(c, p, c->n_seg, c->max_seg_len);
    if (r) return r;
    r = bm_internal_map_range(c, p, 0, c->n_seg, c->max_seg_len);
    if (r) return r;
    return bm_internal_map_finish(c, p);
}

int bm_prepare(bm_ctx* c, const bm_params* p) {
    if (!c) return BM_E_ARG;
    if (!p) BM_FAIL(c, BM_E_ARG, "bm_prepare: null parameters");
    if (!c->have_index) BM_FAIL(c, BM_E_ARG, "bm_prepare: no index resident");
    cudaSetDevice(c->device);
    int r = bm_internal_index_prepare(c, p->distinguishability);
    if (r) return r;
    if (p->k >= p->q && p->k <= 16) { r = bm_internal_kmer_table_prepare(c, p); if (r) return r; }   // builds the locator index first
    if (c->d_packed.p) { r = bm_internal_gidx_prepare(c, p->k); if (r) return r; }
    BM_CK(c, cudaStreamSynchronize(c->stream));
    return BM_OK;
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
    r = bm_internal_kmer_table_prepare(c, p);             // decides c->use_ktable for this call
    if (r) return r;
    const size_t n_lists = 2 * (size_t)n_seg_cap;
    BM_CK(c, c->d_rowidx.ensure(n_lists * S * NQ * 4 + 16));
    BM_CK(c, c->d_segok.ensure((size_t)n_seg_cap + 16));
    BM_CK(c, c->d_cand_fixed.ensure(n_lists * p->max_candidates * 4 + 16));
    BM_CK(c, c->d_cand_cnt.ensure((n_lists + 1) * 4));
    BM_CK(c, c->d_cand_off.ensure((n_lists + 1) * 4));
    BM_CK(c, c->d_stats.ensure(208 * 8));
    BM_CK(c, cudaMemsetAsync(c->d_stats.p, 0, 192 * 8, c->stream));   // [0..7] the counters, [8..187] K2f's slots
    c->map_max_cand = p->max_candidates;
    c->n_cand = 0;
    if (c->use_ktable) {                                  // lists K2f hands over to K2 (+ their count)
        BM_CK(c, c->d_map_over.ensure((n_lists + 2) * 4));
        BM_CK(c, cudaMemsetAsync(c->d_map_over.as<uint32_t>() + n_lists, 0, 4, c->stream));
        c->map_over_cap = (uint32_t)n_lists;
    }
    c->lsamp_k = c->lsamp_P = c->lsamp_b = 0;
    if (c->use_ktable && p->locator_samples && p->locator_samples <= 1024) {
        // K2f also draws the locator's samples (same k-mers, quality filter only: bucket_locator.h:315-341)
        BM_CK(c, c->d_lsamp_idx.ensure((size_t)n_seg_cap * p->locator_samples * 2 + 16));
        BM_CK(c, c->d_lsamp_hash.ensure((size_t)n_seg_cap * p->locator_samples * 4 + 16));
        c->lsamp_k = p->k; c->lsamp_P = p->locator_samples; c->lsamp_b = p->avg_base_quality;
    }
    const uint32_t msl = max_seg_len ? max_seg_len : 1;
    size_t smem_a = (size_t)kWarpsPerCta * ((((size_t)msl + 3) & ~(size_t)3) + 2 * ((msl + 31) / 32 + 1) * 4);
    if (smem_a > 48 * 1024)
        BM_CK(c, cudaFuncSetAttribute(k_map_sample, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_a));
    return BM_OK;
}

// K2a + K2 over segments [seg0, seg1); everything they read must already be resident (or ordered before on c->stream)
int bm_internal_map_range(bm_ctx* c, const bm_params* p, uint32_t seg0, uint32_t seg1, uint32_t max_seg_len) {
    if (seg1 <= seg0) return BM_OK;
    if (c->use_ktable) return bm_internal_map_fused(c, p, seg0, seg1, max_seg_len);   // K2f
    const uint32_t NQ = p->k - p->q + 1, S = p->mapper_samples, F = p->mapper_fault;
    // smallest G with !(G < 0.2*S), the comparison of q_gram_mapper.h:445 in double
    uint32_t min_good = 0;
    while ((double)min_good < 0.2 * (double)S) min_good++;
    const uint32_t msl = max_seg_len ? max_seg_len : 1;
    size_t smem_a = (size_t)kWarpsPerCta * ((((size_t)msl + 3) & ~(size_t)3) + 2 * ((msl + 31) / 32 + 1) * 4);
    const uint32_t blocks_a = (seg1 - seg0 + kWarpsPerCta - 1) / kWarpsPerCta;
    {
        KScope ks(c, BM_K_SAMPLE);
        const ReadPlanes rp{c->d_pl.as<uint2>(), c->have_qk ? c->d_qk.as<uint32_t>() : nullptr, c->d_pl_off.as<uint32_t>()};
        k_map_sample<<<blocks_a, kWarpsPerCta * 32, smem_a, c->stream>>>(
            rp, c->d_seg_read.as<uint32_t>(), c->d_seg_start.as<uint32_t>(), c->d_seg_len.as<uint32_t>(), seg0, seg1, msl,
            c->d_k2i.as<int32_t>(), c->d_distok.as<uint32_t>(), p->k, p->q, S, min_good,
            0u, c->d_rowidx.as<int32_t>(), c->d_segok.as<uint8_t>(), c->d_stats.as<unsigned long long>());
    }
    BM_CK(c, cudaGetLastError());
    const uint32_t NP = (F <= 7) ? 3 : (F <= 15 ? 4 : 6);
    const uint32_t NQv = NQ;                              // rows gathered per sample
    size_t smem_v = (size_t)kWarpsPerCta * S * NQv * 4;
    const uint32_t list0 = 2 * seg0, list1 = 2 * seg1;
    const uint32_t blocks_v = (list1 - list0 + kWarpsPerCta - 1) / kWarpsPerCta;
    cudaError_t e = cudaSuccess;
    {
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
    if (c->use_ktable && n_seg) {
        const uint32_t NQ = p->k - p->q + 1, S = p->mapper_samples, F = p->mapper_fault;
        const uint32_t NP = (F <= 7) ? 3 : (F <= 15 ? 4 : 6);
        const size_t smem_v = (size_t)kWarpsPerCta * S * NQ * 4;
        const uint32_t* ids = c->d_map_over.as<uint32_t>();
        const uint32_t* n_ids = ids + c->map_over_cap;
        const uint32_t blocks = (uint32_t)c->sm_count * 8;
        cudaError_t e = cudaSuccess;
        {
            KScope ks(c, BM_K_MAPOVER);
            switch (NQ) {
                case 1: e = launch_vote_lists_np<1>(c, p, NP, ids, n_ids, blocks, smem_v); break;
                case 2: e = launch_vote_lists_np<2>(c, p, NP, ids, n_ids, blocks, smem_v); break;
                case 3: e = launch_vote_lists_np<3>(c, p, NP, ids, n_ids, blocks, smem_v); break;
                case 4: e = launch_vote_lists_np<4>(c, p, NP, ids, n_ids, blocks, smem_v); break;
                case 5: e = launch_vote_lists_np<5>(c, p, NP, ids, n_ids, blocks, smem_v); break;
                case 6: e = launch_vote_lists_np<6>(c, p, NP, ids, n_ids, blocks, smem_v); break;
                case 7: e = launch_vote_lists_np<7>(c, p, NP, ids, n_ids, blocks, smem_v); break;
                default: e = launch_vote_lists_np<8>(c, p, NP, ids, n_ids, blocks, smem_v); break;
            }
        }
        BM_CK(c, e);
        k_stats_fold<<<1, 64, 0, c->stream>>>(c->d_stats.as<unsigned long long>(), n_ids);
    }
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
                c->d_stats.as<unsigned long long>(), n_seg == c->n_reads ? 1u : 0u);
            if (n_seg != c->n_reads && c->n_reads)
                k_count_mapped_reads<<<(c->n_reads + 255) / 256, 256, 0, c->stream>>>(
                    c->d_read_seg0.as<uint32_t>(), c->d_cand_off.as<uint32_t>(), c->n_reads,
                    c->d_stats.as<unsigned long long>());
        }
        BM_CK(c, cudaGetLastError());
    } else {
        BM_CK(c, cudaMemsetAsync(c->d_cand_off.p, 0, 4, c->stream));
    }
    c->have_map = true;
    return BM_OK;
}
