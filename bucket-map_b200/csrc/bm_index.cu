// bm_index.cu -- q-gram bit matrix: build on device (K1), load/download in .qgram layout, row popcounts (K0).
//
// Device layout: n_rows rows of Wp bytes, Wp = roundup((NB+7)>>3, 64); bucket j of a row at byte j>>3 bit j&7 exactly
// like the .qgram file (indexer/bucket_indexer.h:64-73), padding bytes zero.  64-B padded rows make every row start
// on a DRAM burst boundary and keep the mapper's 16-B vector loads aligned.
#include <algorithm>
#include <cmath>

#include "bm_internal.cuh"

namespace {

// ---- K1: index build ---------------------------------------------------------------------------------------------
// One CTA per emitted bucket (utils.h:82-96); each thread walks a contiguous slice of the bucket with a rolling
// base-4 hash (seqan3 views::kmer_hash(ungapped{q}), first base most significant) and sets bit (row, bucket) for every
// sampled q-gram (bucket_indexer.h:57-60).  Bits of 32 neighbouring buckets share a word, so the set is a RED.OR.
// Consecutive CTAs touch the same 32-B sector column of the matrix, which keeps the read-modify-write traffic in L2.
__global__ void __launch_bounds__(256)
k_index_build(const uint8_t* __restrict__ packed, const uint64_t* __restrict__ bstart, const uint32_t* __restrict__ blen,
              const int32_t* __restrict__ k2i, uint32_t q, uint32_t* __restrict__ rows32, uint32_t Wp32, uint32_t b0) {
    const uint32_t b = b0 + blockIdx.x;       // b0: first bucket of this launch (a rank's share in the sharded build)
    const uint64_t start = bstart[b];
    const uint32_t len = blen[b];
    if (len < q) return;
    const uint32_t npos = len - q + 1;
    const uint32_t per = (npos + blockDim.x - 1) / blockDim.x;
    uint32_t lo = threadIdx.x * per;
    if (lo >= npos) return;
    uint32_t hi = min(lo + per, npos);
    const uint32_t mask = (q >= 16) ? 0xFFFFFFFFu : ((1u << (2 * q)) - 1u);
    const uint32_t word = b >> 5, bit = 1u << (b & 31);
    uint32_t h = 0;
    uint64_t g = start + lo;
    for (uint32_t i = 0; i + 1 < q; i++) h = (h << 2) | bm_gbase(packed, g + i);
    g += q - 1;
    for (uint32_t pos = lo; pos < hi; pos++, g++) {
        h = ((h << 2) | bm_gbase(packed, g)) & mask;
        int idx = __ldg(k2i + h);
        if (idx >= 0) atomicOr(rows32 + (size_t)idx * Wp32 + word, bit);
    }
}

// ---- layout conversion between the unpadded .qgram image and the padded device rows ----------------------------------
__global__ void k_pad_rows(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, uint32_t W, uint32_t Wp,
                           uint64_t total /* n_rows*Wp */) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (; i < total; i += stride) {
        uint64_t r = i / Wp; uint32_t c = (uint32_t)(i - r * Wp);
        dst[i] = (c < W) ? src[r * W + c] : (uint8_t)0;
    }
}
__global__ void k_unpad_rows(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, uint32_t W, uint32_t Wp,
                             uint64_t total /* n_rows*W */) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (; i < total; i += stride) {
        uint64_t r = i / W; uint32_t c = (uint32_t)(i - r * W);
        dst[i] = src[r * Wp + c];
    }
}

// ---- K0: zeros[row] = NB - popcount(row)  (distinguishability_filter::read, q_gram_mapper.h:171-184) ------------------
__global__ void k_row_zeros(const uint4* __restrict__ rows, uint32_t n_rows, uint32_t Wp16, uint32_t NB, uint32_t thr,
                            uint32_t* __restrict__ zeros, uint32_t* __restrict__ valid) {
    uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= n_rows) return;
    const uint4* r = rows + (size_t)warp * Wp16;
    uint32_t ones = 0;
    for (uint32_t c = lane; c < Wp16; c += 32) {
        uint4 v = __ldg(r + c);
        ones += __popc(v.x) + __popc(v.y) + __popc(v.z) + __popc(v.w);
    }
    for (int o = 16; o; o >>= 1) ones += __shfl_xor_sync(0xFFFFFFFFu, ones, o);
    if (lane == 0) {
        zeros[warp] = NB - ones;                  // both branches of :178-183 give NB - ones
        if (NB - ones > thr) atomicAdd(valid, 1u);   // the logged counter (:175-177)
    }
}

// distok bit g = (kmer_to_index[g] >= 0 && zeros[idx] >= threshold): one q-gram's share of is_highly_distinguishable
// (q_gram_mapper.h:189-196), folded into a 4^q-bit table the sampler kernel keeps hot in L1.
__global__ void k_distok(const int32_t* __restrict__ k2i, const uint32_t* __restrict__ zeros, uint32_t n_codes,
                         uint32_t thr, uint32_t* __restrict__ distok) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    bool ok = false;
    if (g < n_codes) { int idx = k2i[g]; ok = idx >= 0 && zeros[idx] >= thr; }
    uint32_t m = __ballot_sync(0xFFFFFFFFu, ok);
    if ((threadIdx.x & 31) == 0 && g < n_codes) distok[g >> 5] = m;
}

// ---- K0b: coarse k-mer table -----------------------------------------------------------------------------------------------
// For k-mer code x: v = AND over i in [0, k-q] of rows[kmer_to_index[(x >> 2i) & Q]] (un-indexed q-grams skipped), i.e.
// exactly the bf_res that q_gram_mapper::query builds for sample hash x (q_gram_mapper.h:399-407) before filter->read();
// coarse bit g of row x = (32-bit word g of v != 0).  A CTA takes 64 consecutive k-mer codes -- they share all but the
// lowest three base-4 digits, so the rows of the higher q-grams are re-read from L1 -- one thread owns one 16-B column
// (4 coarse bits) and 8 neighbouring lanes assemble one coarse word.  Reads ~1.3x (4^k x Wp) bytes, writes 4^k x 128 B.
__global__ void __launch_bounds__(256)
k_kmer_coarse(const uint4* __restrict__ rows, uint32_t Wp16, uint32_t NB, const int32_t* __restrict__ k2i, uint32_t NQ,
              uint32_t Q, uint64_t n_kmers, uint32_t* __restrict__ crows, uint32_t Wcs, uint32_t logg,
              uint4* __restrict__ frows /* full-resolution rows (4^k x Wp16) or nullptr */) {
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t nbits = 128u >> logg;          // coarse bits per 16-B column: 4 (G=32), 8 (G=16), 16 (G=8)
    const uint32_t lpw = 32u / nbits;             // lanes that share one coarse word: 8, 4, 2
    for (uint64_t x0 = (uint64_t)blockIdx.x * 64; x0 < n_kmers; x0 += (uint64_t)gridDim.x * 64) {
        const uint32_t nx = (uint32_t)min((uint64_t)64, n_kmers - x0);
        for (uint32_t c0 = warp * 32; c0 < Wp16; c0 += 256) {
            const uint32_t c = c0 + lane;
            const bool act = c < Wp16;
            // a k-mer none of whose q-grams is indexed matches every REAL bucket; padding bits stay zero
            uint32_t w[4];
#pragma unroll
            for (int i = 0; i < 4; i++) {
                uint32_t b = c * 128u + 32u * i;
                w[i] = (!act || b >= NB) ? 0u : (NB - b >= 32u ? 0xFFFFFFFFu : ((1u << (NB - b)) - 1u));
            }
            const uint4 all = make_uint4(w[0], w[1], w[2], w[3]);
            const uint32_t widx = c / lpw;
#pragma unroll 4
            for (uint32_t j = 0; j < nx; j++) {
                const uint32_t x = (uint32_t)(x0 + j);
                uint4 v = all;
                for (uint32_t i = 0; i < NQ; i++) {
                    const int32_t r = __ldg(k2i + ((x >> (2 * i)) & Q));
                    if (r >= 0 && act) {
                        const uint4 t = __ldg(rows + (size_t)r * Wp16 + c);
                        v.x &= t.x; v.y &= t.y; v.z &= t.z; v.w &= t.w;
                    }
                }
                if (frows && act) frows[(x0 + j) * Wp16 + c] = v;
                const uint32_t vw[4] = {v.x, v.y, v.z, v.w};
                uint32_t nib = 0;
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    const uint32_t t = vw[i];
                    if (logg == 5) nib |= (t ? 1u : 0u) << i;
                    else if (logg == 4) nib |= ((t & 0xFFFFu ? 1u : 0u) | (t >> 16 ? 2u : 0u)) << (2 * i);
                    else nib |= ((t & 0xFFu ? 1u : 0u) | (t & 0xFF00u ? 2u : 0u) | (t & 0xFF0000u ? 4u : 0u) | (t >> 24 ? 8u : 0u)) << (4 * i);
                }
                uint32_t cw = nib << (nbits * (lane & (lpw - 1)));
                cw |= __shfl_xor_sync(0xFFFFFFFFu, cw, 1);
                if (lpw > 2) cw |= __shfl_xor_sync(0xFFFFFFFFu, cw, 2);
                if (lpw > 4) cw |= __shfl_xor_sync(0xFFFFFFFFu, cw, 4);
                if ((lane & (lpw - 1)) == 0 && widx < Wcs) crows[(x0 + j) * Wcs + widx] = cw;
            }
        }
    }
}

// Density of the k-mer rows: set bits of the AND row of `n_samples` pseudo-random k-mer codes (one CTA each).  Feeds the
// choice of the coarse group size.
__global__ void __launch_bounds__(128)
k_kmer_density(const uint4* __restrict__ rows, uint32_t Wp16, uint32_t NB, const int32_t* __restrict__ k2i, uint32_t NQ,
               uint32_t Q, uint32_t kmask, unsigned long long* __restrict__ total) {
    const uint32_t x = (blockIdx.x * 0x9E3779B1u) & kmask;
    int32_t r[8];
    for (uint32_t i = 0; i < NQ; i++) r[i] = __ldg(k2i + ((x >> (2 * i)) & Q));
    uint32_t ones = 0;
    for (uint32_t c = threadIdx.x; c < Wp16; c += blockDim.x) {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const uint32_t b = c * 128u + 32u * i;
            w[i] = (b >= NB) ? 0u : (NB - b >= 32u ? 0xFFFFFFFFu : ((1u << (NB - b)) - 1u));
        }
        for (uint32_t i = 0; i < NQ; i++)
            if (r[i] >= 0) {
                const uint4 t = __ldg(rows + (size_t)r[i] * Wp16 + c);
                w[0] &= t.x; w[1] &= t.y; w[2] &= t.z; w[3] &= t.w;
            }
        ones += __popc(w[0]) + __popc(w[1]) + __popc(w[2]) + __popc(w[3]);
    }
    ones = __reduce_add_sync(0xFFFFFFFFu, ones);
    if ((threadIdx.x & 31) == 0 && ones) atomicAdd(total, (unsigned long long)ones);
}

// the same bits indexed by the two q-bit planes of a q-gram as the fused mapper kernel reads them out of its ballot
// words: index = (hi << q) | lo, bit t of lo / hi = low / high bit of the rank of base t (first base = bit 0)
__global__ void k_distok_planes(const uint32_t* __restrict__ distok, uint32_t q, uint32_t n_codes, uint32_t* __restrict__ out) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n_codes || !((distok[g >> 5] >> (g & 31)) & 1u)) return;
    uint32_t lo = 0, hi = 0;
    for (uint32_t t = 0; t < q; t++) {
        uint32_t base = (g >> (2 * (q - 1 - t))) & 3u;
        lo |= (base & 1u) << t; hi |= (base >> 1) << t;
    }
    uint32_t idx = (hi << q) | lo;
    atomicOr(out + (idx >> 5), 1u << (idx & 31));
}

}  // namespace

// sizes the resident matrix for a kmer_to_index table (used by build, load, alloc and the index broadcast)
int bm_internal_index_set_table(bm_ctx* c, const int32_t* k2i, uint32_t q, uint32_t n_rows, uint32_t NB) {
    if (q == 0 || q > 14) BM_FAIL(c, BM_E_LIMIT, "index seed length q must be in [1,14]");
    if (NB == 0) BM_FAIL(c, BM_E_ARG, "NB must be > 0");
    uint64_t n_codes = 1ull << (2 * q);
    uint32_t cnt = 0;
    for (uint64_t i = 0; i < n_codes; i++) {
        if (k2i[i] >= 0) { if ((uint32_t)k2i[i] >= n_rows) BM_FAIL(c, BM_E_ARG, "kmer_to_index entry >= n_rows"); cnt++; }
    }
    (void)cnt;
    c->have_index = false; c->zeros_for_d = -1.f; c->krows_k = 0; c->frows_k = 0; c->fine_tried_k = 0; c->auto_k = 0;
    c->q = q; c->n_rows = n_rows; c->NB = NB;
    c->W = (NB + 7) >> 3; c->Wp = (c->W + 63u) & ~63u;
    BM_CK(c, c->d_k2i.ensure(n_codes * 4));
    BM_CK(c, cudaMemcpyAsync(c->d_k2i.p, k2i, n_codes * 4, cudaMemcpyHostToDevice, c->stream));
    BM_CK(c, c->d_rows.ensure((size_t)n_rows * c->Wp + 64));
    BM_CK(c, c->d_zeros.ensure((size_t)(n_rows + 1) * 4));
    BM_CK(c, c->d_distok.ensure((size_t)((n_codes + 31) / 32) * 4 + 64));
    BM_CK(c, c->d_distok_p.ensure((size_t)((n_codes + 31) / 32) * 4 + 64));
    return BM_OK;
}

namespace {
inline int set_table(bm_ctx* c, const int32_t* k2i, uint32_t q, uint32_t n_rows, uint32_t NB) {
    return bm_internal_index_set_table(c, k2i, q, n_rows, NB);
}
}  // namespace

int bm_internal_index_prepare(bm_ctx* c, float d) {
    if (!c->have_index) BM_FAIL(c, BM_E_ARG, "no index resident (bm_index_build / bm_index_load first)");
    if (c->zeros_for_d == d) return BM_OK;
    const uint32_t thr = (unsigned int)(d * c->NB);          // q_gram_mapper.h:163 (float * unsigned, truncated)
    BM_CK(c, c->d_stats.ensure(208 * 8));
    uint32_t* d_valid = c->d_stats.as<uint32_t>() + 400;     // scratch word past the mapper's stats and counter slots
    BM_CK(c, cudaMemsetAsync(d_valid, 0, 4, c->stream));
    if (c->n_rows) {
        KScope ks(c, BM_K_ROWPOP);
        uint32_t blocks = (c->n_rows + 7) / 8;
        k_row_zeros<<<blocks, 256, 0, c->stream>>>(c->d_rows.as<uint4>(), c->n_rows, c->Wp / 16, c->NB, thr,
                                                   c->d_zeros.as<uint32_t>(), d_valid);
    }
    uint32_t n_codes = 1u << (2 * c->q);
    {
        KScope ks(c, BM_K_AUX);
        k_distok<<<(n_codes + 255) / 256, 256, 0, c->stream>>>(c->d_k2i.as<int32_t>(), c->d_zeros.as<uint32_t>(), n_codes,
                                                               thr, c->d_distok.as<uint32_t>());
        BM_CK(c, cudaMemsetAsync(c->d_distok_p.p, 0, (size_t)((n_codes + 31) / 32) * 4, c->stream));
        k_distok_planes<<<(n_codes + 255) / 256, 256, 0, c->stream>>>(c->d_distok.as<uint32_t>(), c->q, n_codes,
                                                                      c->d_distok_p.as<uint32_t>());
    }
    BM_CK(c, cudaGetLastError());
    BM_CK(c, cudaMemcpyAsync(&c->dist_valid, d_valid, 4, cudaMemcpyDeviceToHost, c->stream));
    BM_CK(c, cudaStreamSynchronize(c->stream));
    c->dist_threshold = thr; c->zeros_for_d = d;
    return BM_OK;
}

// Which coarse table, if any, pays off for this index and these parameters.  Bytes per read at the DRAM pins (128-B lines):
//   coarse vote   2 S rows of ceil(NB / G / 8) bytes, whole lines
//   recount       every group that passes the coarse vote costs S table words (S * (k-q+1) .qgram words without the fine
//                 table), one line each; a group passes by chance when >= S-F+1 of the S samples hit one of its G buckets:
//                 binomial tail with p = 1 - (1 - d)^G, d = measured density of the k-mer rows
//   full gather   (no table, or a list with too many passing groups) the k-q+1 q-gram rows of every sample, about half of
//                 them with the dead-tile exit when every q-gram is indexed
// On the 1.7 Gbp / L=65536 / -f 1 reference d = 1.6 % and G = 16 wins (7.7 + 3 kB against 17 kB for G = 8, 86 kB for G = 32,
// 166 kB for the gather); with L = 262144 / -k 10 d = 3.8 % and G = 8 wins; with -f 0.25 a third of all k-mers has no indexed
// q-gram (d = 43 %): every coarse bit is set, every list would be handed over, and the gather alone is cheaper.
static int kmer_table_choose(bm_ctx* c, const bm_params* p, uint32_t* logg_out) {
    const uint32_t k = p->k, S = p->mapper_samples, F = p->mapper_fault, NQ = k - c->q + 1;
    if (c->auto_k == k && c->auto_S == S && c->auto_F == F && c->auto_d == p->distinguishability) { *logg_out = c->auto_logg; return BM_OK; }
    BM_CK(c, c->d_stats.ensure(208 * 8));
    unsigned long long* d_tot = c->d_stats.as<unsigned long long>() + 204;
    BM_CK(c, cudaMemsetAsync(d_tot, 0, 8, c->stream));
    const uint32_t n_samples = 4096;
    {
        KScope ks(c, BM_K_AUX);
        k_kmer_density<<<n_samples, 128, 0, c->stream>>>(c->d_rows.as<uint4>(), c->Wp / 16, c->NB, c->d_k2i.as<int32_t>(), NQ,
                                                         (1u << (2 * c->q)) - 1u, (k >= 16) ? 0xFFFFFFFFu : ((1u << (2 * k)) - 1u),
                                                         d_tot);
    }
    BM_CK(c, cudaGetLastError());
    unsigned long long tot = 0;
    BM_CK(c, cudaMemcpyAsync(&tot, d_tot, 8, cudaMemcpyDeviceToHost, c->stream));
    BM_CK(c, cudaStreamSynchronize(c->stream));
    const double d = (double)tot / ((double)n_samples * (double)c->NB);
    const double f_idx = (double)c->n_rows / (double)(1ull << (2 * c->q));
    const double gather = 2.0 * S * NQ * f_idx * c->Wp * (f_idx > 0.9 ? 0.45 : 0.8);
    const uint32_t need = (S >= F) ? S - F + 1 : 0;
    auto tail = [&](double pr) {                      // P(Bin(S, pr) >= need)
        double t = 0.0;
        for (uint32_t h = need; h <= S; h++) {
            double cmb = 1.0;
            for (uint32_t i = 0; i < h; i++) cmb = cmb * (double)(S - i) / (double)(i + 1);
            t += cmb * std::pow(pr, (double)h) * std::pow(1.0 - pr, (double)(S - h));
        }
        return t > 1.0 ? 1.0 : t;
    };
    const bool fine = c->ftable_mode != 0;
    double best = gather;
    uint32_t best_logg = 0xFF;
    // A bit matrix that lives in L2 (small references: C1's is 1 MB) costs no DRAM bytes either way, and the byte model has
    // nothing to say; there the fused kernel wins by being one pass instead of three (measured on C1 at -f 0.25: 0.33 ms
    // against 0.61 ms for sampler + gather + locator sampler per 100 k reads).
    const bool l2_resident = (uint64_t)c->n_rows * c->Wp <= ((uint64_t)32 << 20);
    // No q-gram passes -d (C4 at L = 262144, -k 9: every row is set in 63 % of the buckets): no k-mer survives the filter and
    // every segment leaves the mapper at that test, which the fused kernel reaches in one pass (0.8 against 3.0 ms per 1 M reads).
    const bool nothing_passes = c->zeros_for_d == p->distinguishability && c->dist_valid == 0;
    if ((l2_resident || nothing_passes) && need > 0) { best = 0.0; best_logg = 4; }
    for (uint32_t logg = 3; logg <= 5 && need > 0 && !l2_resident && !nothing_passes; logg++) {
        const double G = (double)(1u << logg);
        const double groups = std::ceil((double)c->NB / G);
        const double lines = std::ceil(std::ceil(groups / 8.0) / 128.0);
        const double coarse = 2.0 * S * lines * 128.0;
        const double pc = 1.0 - std::pow(1.0 - d, G);
        const double passes = 2.0 * groups * tail(pc) + 1.0;
        const double max_pass = std::max(16.0, groups / 8.0);
        const double recount = passes * S * (fine ? 1.0 : NQ * f_idx) * 128.0;
        const double cost = coarse + (passes > max_pass ? gather : std::min(recount, gather));
        if (cost < best) { best = cost; best_logg = logg; }
    }
    c->auto_k = k; c->auto_S = S; c->auto_F = F; c->auto_d = p->distinguishability; c->auto_logg = best_logg; c->auto_density = d;
    *logg_out = best_logg;
    return BM_OK;
}

// Builds (or re-uses) the k-mer tables for query seed length p->k; sets c->use_ktable / c->use_ftable.  Not built when
// switched off, when 4^k rows do not fit the memory budget, or (mode 1) when the cost model above says that the full-width
// gather is cheaper for this index: the mapper then gathers the k-q+1 q-gram rows per sample.  Both ways are CUDA paths
// with identical results.
int bm_internal_kmer_table_prepare(bm_ctx* c, const bm_params* p) {
    const uint32_t k = p->k;
    c->use_ktable = false; c->use_ftable = false;
    if (!c->have_index) BM_FAIL(c, BM_E_ARG, "no index resident (bm_index_build / bm_index_load first)");
    if (c->ktable_mode == 0 || k < c->q || k < 3 || k > 15 || k - c->q + 1 > 8 || c->n_rows == 0) {
        if (c->ktable_mode == 0 && c->d_krows.p) { c->d_krows.release(); c->krows_k = 0; c->d_frows.release(); c->frows_k = 0; }
        return BM_OK;
    }
    uint32_t want_logg = c->ktable_logg;
    if (c->ktable_auto) {
        int rc = kmer_table_choose(c, p, &want_logg);
        if (rc) return rc;
        if (want_logg == 0xFF) {                  // the gather is cheaper: no tables
            if (c->d_krows.p) { c->d_krows.release(); c->krows_k = 0; c->d_frows.release(); c->frows_k = 0; }
            return BM_OK;
        }
    }
    if (c->ftable_mode == 0 && c->d_frows.p) { c->d_frows.release(); c->frows_k = 0; }
    const bool fine_ok = c->ftable_mode == 0 || (c->frows_k == k && c->d_frows.p) || c->fine_tried_k == k;
    if (c->krows_k == k && c->krows_logg == want_logg && c->d_krows.p && fine_ok) {
        c->use_ktable = true; c->use_ftable = c->ftable_mode != 0 && c->frows_k == k && c->d_frows.p;
        return BM_OK;
    }
    const uint64_t n_kmers = 1ull << (2 * k);
    const uint32_t Wp32 = c->Wp / 4;
    const uint32_t logg = want_logg;                               // buckets per coarse bit = 1 << logg
    c->Wc = ((Wp32 << (5 - logg)) + 31) / 32;
    c->Wcs = (c->Wc + 15u) & ~15u;
    const uint64_t bytes = (n_kmers + 1) * c->Wcs * 4;             // + one all-zero row (the fused mapper's padding samples)
    c->d_krows.release(); c->krows_k = 0; c->d_frows.release(); c->frows_k = 0;
    // The locator's genome k-mer index needs ~16 B per base of scratch while it is built: build it BEFORE the tables take
    // their memory (it is wanted at the first locator call anyway), so that the fine table can use what is left.
    if (c->gidx_mode && c->d_packed.p) { int r = bm_internal_gidx_prepare(c, k); if (r) return r; }
    uint64_t budget = c->ktable_max_bytes;
    size_t fr = 0, tot = 0;
    BM_CK(c, cudaMemGetInfo(&fr, &tot));
    if (!budget) budget = fr / 2;                 // leave the other half to read batches, candidates and scratch
    if (bytes > budget) return BM_OK;
    cudaError_t e = c->d_krows.ensure_exact(bytes);
    if (e != cudaSuccess) { cudaGetLastError(); return BM_OK; }   // does not fit after all: q-gram gather
    // full-resolution table: only when it fits beside the coarse one with room for batches, results and scratch
    // (24 GB, or a quarter of the card) -- and for the locator index if that is still to come
    uint4* frows = nullptr;
    c->fine_tried_k = k;
    if (c->ftable_mode) {
        const uint64_t fbytes = n_kmers * (uint64_t)c->Wp;
        BM_CK(c, cudaMemGetInfo(&fr, &tot));
        uint64_t reserve = (uint64_t)24 << 30;
        if (reserve < tot / 8) reserve = tot / 8;
        if (c->gidx_mode && !c->d_gidx_pos.p && c->n_bases) reserve += c->n_bases * 20 + (n_kmers << 8);
        if (fbytes + reserve <= fr && c->d_frows.ensure_exact(fbytes) == cudaSuccess) frows = c->d_frows.as<uint4>();
        else cudaGetLastError();
    }
    BM_CK(c, cudaMemsetAsync(c->d_krows.p, 0, bytes, c->stream));   // stride padding words
    {
        KScope ks(c, BM_K_KTABLE);
        const uint64_t want = (n_kmers + 63) / 64;
        const uint32_t blocks = (uint32_t)(want < (uint64_t)c->sm_count * 64 ? want : (uint64_t)c->sm_count * 64);
        k_kmer_coarse<<<blocks, 256, 0, c->stream>>>(c->d_rows.as<uint4>(), c->Wp / 16, c->NB, c->d_k2i.as<int32_t>(),
                                                     k - c->q + 1, (c->q >= 16) ? 0xFFFFFFFFu : ((1u << (2 * c->q)) - 1u),
                                                     n_kmers, c->d_krows.as<uint32_t>(), c->Wcs, logg, frows);
    }
    BM_CK(c, cudaGetLastError());
    c->krows_k = k; c->krows_logg = logg; c->use_ktable = true;
    if (frows) { c->frows_k = k; c->use_ftable = true; }
    return BM_OK;
}

extern "C" {

int bm_index_kmer_table(bm_ctx* c, int mode, uint64_t max_bytes) {
    if (!c) return BM_E_ARG;
    if (mode != 0 && mode != 1 && mode != 8 && mode != 16 && mode != 32)
        BM_FAIL(c, BM_E_ARG, "bm_index_kmer_table: mode must be 0 (off), 1 (when it fits) or a group size 8/16/32");
    c->ktable_logg = (mode == 8) ? 3 : (mode == 32) ? 5 : 4;
    c->ktable_auto = mode == 1; c->auto_k = 0;
    c->ktable_mode = mode ? 1 : 0; c->ktable_max_bytes = max_bytes;
    if (mode == 0 || (max_bytes && c->d_krows.cap > max_bytes)) {
        cudaSetDevice(c->device);
        c->d_krows.release(); c->krows_k = 0; c->d_frows.release(); c->frows_k = 0;
    }
    c->fine_tried_k = 0;
    return BM_OK;
}

int bm_index_fine_table(bm_ctx* c, int mode) {
    if (!c) return BM_E_ARG;
    if (mode != 0 && mode != 1) BM_FAIL(c, BM_E_ARG, "bm_index_fine_table: mode must be 0 (off) or 1 (when it fits)");
    c->ftable_mode = mode; c->fine_tried_k = 0; c->auto_k = 0;
    if (!mode) { cudaSetDevice(c->device); c->d_frows.release(); c->frows_k = 0; }
    return BM_OK;
}

int bm_index_fine_table_info(bm_ctx* c, uint32_t* k, uint64_t* bytes) {
    if (!c) return BM_E_ARG;
    const bool on = c->d_frows.p && c->frows_k;
    if (k) *k = on ? c->frows_k : 0;
    if (bytes) *bytes = on ? ((1ull << (2 * c->frows_k)) * (uint64_t)c->Wp) : 0;
    return BM_OK;
}

int bm_index_kmer_table_choice(bm_ctx* c, uint32_t* buckets_per_bit, double* density) {
    if (!c) return BM_E_ARG;
    const bool on = c->d_krows.p && c->krows_k;
    if (buckets_per_bit) *buckets_per_bit = on ? (1u << c->krows_logg) : 0u;
    if (density) *density = c->auto_k ? c->auto_density : 0.0;
    return BM_OK;
}

int bm_index_kmer_table_info(bm_ctx* c, uint32_t* k, uint64_t* bytes) {
    if (!c) return BM_E_ARG;
    if (k) *k = (c->d_krows.p ? c->krows_k : 0);
    if (bytes) *bytes = (c->d_krows.p && c->krows_k) ? (((1ull << (2 * c->krows_k)) + 1) * c->Wcs * 4) : 0;
    return BM_OK;
}

int bm_index_build(bm_ctx* c, const int32_t* k2i, uint32_t q, uint32_t n_rows, uint8_t* out_rows) {
    if (!c) return BM_E_ARG;
    if (!k2i) BM_FAIL(c, BM_E_ARG, "bm_index_build: kmer_to_index is null");
    if (c->NB == 0 || !c->d_packed.p) BM_FAIL(c, BM_E_ARG, "bm_index_build: bm_ref_upload first");
    cudaSetDevice(c->device);
    int r = set_table(c, k2i, q, n_rows, c->NB);
    if (r) return r;
    BM_CK(c, cudaMemsetAsync(c->d_rows.p, 0, (size_t)n_rows * c->Wp + 64, c->stream));
    if (c->n_emit && n_rows) {
        KScope ks(c, BM_K_INDEX);
        k_index_build<<<c->n_emit, 256, 0, c->stream>>>(c->d_packed.as<uint8_t>(), c->d_bstart.as<uint64_t>(),
                                                        c->d_blen.as<uint32_t>(), c->d_k2i.as<int32_t>(), q,
                                                        c->d_rows.as<uint32_t>(), c->Wp / 4, 0u);
    }
    BM_CK(c, cudaGetLastError());
    BM_CK(c, cudaStreamSynchronize(c->stream));
    c->have_index = true;
    if (out_rows) return bm_index_download(c, out_rows);
    return BM_OK;
}

}  // extern "C"

// K1 over the buckets [b0, b1) only, into a zeroed matrix: one rank's share of bm_index_build_sharded (bm_comm.cu), which
// sums the ranks' matrices afterwards.  Enqueued on c->stream, not waited for; have_index stays false until the sum is in.
int bm_internal_index_build_range(bm_ctx* c, const int32_t* k2i, uint32_t q, uint32_t n_rows, uint32_t b0, uint32_t b1) {
    if (!k2i) BM_FAIL(c, BM_E_ARG, "bm_index_build_sharded: kmer_to_index is null");
    if (c->NB == 0 || !c->d_packed.p) BM_FAIL(c, BM_E_ARG, "bm_index_build_sharded: bm_ref_upload / bm_bcast_ref first");
    if (b0 > b1 || b1 > c->n_emit) BM_FAIL(c, BM_E_ARG, "bm_index_build_sharded: bad bucket range");
    cudaSetDevice(c->device);
    int r = set_table(c, k2i, q, n_rows, c->NB);
    if (r) return r;
    c->have_index = false;
    BM_CK(c, cudaMemsetAsync(c->d_rows.p, 0, (size_t)n_rows * c->Wp + 64, c->stream));
    if (b1 > b0 && n_rows) {
        KScope ks(c, BM_K_INDEX);
        k_index_build<<<b1 - b0, 256, 0, c->stream>>>(c->d_packed.as<uint8_t>(), c->d_bstart.as<uint64_t>(),
                                                      c->d_blen.as<uint32_t>(), c->d_k2i.as<int32_t>(), q,
                                                      c->d_rows.as<uint32_t>(), c->Wp / 4, b0);
    }
    BM_CK(c, cudaGetLastError());
    return BM_OK;
}

extern "C" {

int bm_index_load(bm_ctx* c, const int32_t* k2i, uint32_t q, const uint8_t* rows, uint32_t n_rows, uint32_t NB) {
    if (!c) return BM_E_ARG;
    if (!k2i || (n_rows && !rows)) BM_FAIL(c, BM_E_ARG, "bm_index_load: null argument");
    cudaSetDevice(c->device);
    if (c->d_packed.p && c->NB && c->NB != NB) BM_FAIL(c, BM_E_ARG, "bm_index_load: NB differs from the uploaded reference");
    int r = set_table(c, k2i, q, n_rows, NB);
    if (r) return r;
    uint64_t raw = (uint64_t)n_rows * c->W, padded = (uint64_t)n_rows * c->Wp;
    if (n_rows) {
        BM_CK(c, c->d_stage.ensure(raw));
        { int rh = bm_internal_h2d(c, c->d_stage.p, rows, raw); if (rh) return rh; }
        KScope ks(c, BM_K_AUX);
        k_pad_rows<<<c->sm_count * 8, 256, 0, c->stream>>>(c->d_stage.as<uint8_t>(), c->d_rows.as<uint8_t>(), c->W, c->Wp, padded);
    }
    BM_CK(c, cudaMemsetAsync((uint8_t*)c->d_rows.p + padded, 0, 64, c->stream));
    BM_CK(c, cudaGetLastError());
    BM_CK(c, cudaStreamSynchronize(c->stream));
    c->d_stage.release();
    c->have_index = true;
    return BM_OK;
}

int bm_index_download(bm_ctx* c, uint8_t* out_rows) {
    if (!c) return BM_E_ARG;
    if (!c->have_index) BM_FAIL(c, BM_E_ARG, "bm_index_download: no index resident");
    if (!out_rows) BM_FAIL(c, BM_E_ARG, "bm_index_download: null output");
    cudaSetDevice(c->device);
    uint64_t raw = (uint64_t)c->n_rows * c->W;
    if (!raw) return BM_OK;
    BM_CK(c, c->d_stage.ensure(raw));
    {
        KScope ks(c, BM_K_AUX);
        k_unpad_rows<<<c->sm_count * 8, 256, 0, c->stream>>>(c->d_rows.as<uint8_t>(), c->d_stage.as<uint8_t>(), c->W, c->Wp, raw);
    }
    BM_CK(c, cudaGetLastError());
    BM_CK(c, cudaMemcpyAsync(out_rows, c->d_stage.p, raw, cudaMemcpyDeviceToHost, c->stream));
    BM_CK(c, cudaStreamSynchronize(c->stream));
    c->d_stage.release();
    return BM_OK;
}

int bm_index_zeros(bm_ctx* c, float d, uint32_t* zeros_out, uint32_t* valid) {
    if (!c) return BM_E_ARG;
    cudaSetDevice(c->device);
    int r = bm_internal_index_prepare(c, d);
    if (r) return r;
    if (zeros_out && c->n_rows) {
        BM_CK(c, cudaMemcpyAsync(zeros_out, c->d_zeros.p, (size_t)c->n_rows * 4, cudaMemcpyDeviceToHost, c->stream));
        BM_CK(c, cudaStreamSynchronize(c->stream));
    }
    if (valid) *valid = c->dist_valid;
    return BM_OK;
}

int bm_index_alloc(bm_ctx* c, const int32_t* k2i, uint32_t q, uint32_t n_rows, uint32_t NB) {
    if (!c) return BM_E_ARG;
    if (!k2i) BM_FAIL(c, BM_E_ARG, "bm_index_alloc: kmer_to_index is null");
    cudaSetDevice(c->device);
    int r = set_table(c, k2i, q, n_rows, NB);
    if (r) return r;
    BM_CK(c, cudaMemsetAsync(c->d_rows.p, 0, (size_t)n_rows * c->Wp + 64, c->stream));
    BM_CK(c, cudaStreamSynchronize(c->stream));
    c->have_index = true;
    return BM_OK;
}

int bm_index_device(bm_ctx* c, void** dptr, uint64_t* bytes) {
    if (!c) return BM_E_ARG;
    if (!c->have_index) BM_FAIL(c, BM_E_ARG, "bm_index_device: no index resident");
    c->zeros_for_d = -1.f; c->krows_k = 0; c->frows_k = 0; c->fine_tried_k = 0; c->auto_k = 0;   // the caller may overwrite the rows: recompute on next use
    if (dptr) *dptr = c->d_rows.p;
    if (bytes) *bytes = (uint64_t)c->n_rows * c->Wp;
    return BM_OK;
}

int bm_ref_device(bm_ctx* c, void** dptr, uint64_t* bytes) {
    if (!c) return BM_E_ARG;
    if (!c->d_packed.p) BM_FAIL(c, BM_E_ARG, "bm_ref_device: bm_ref_upload first");
    c->d_gidx_pos.release(); c->d_gidx_off.release(); c->gidx_k = 0;   // the caller may overwrite the text
    if (dptr) *dptr = c->d_packed.p;
    if (bytes) *bytes = (c->n_bases + 3) / 4;
    return BM_OK;
}

int bm_index_free(bm_ctx* c) {
    if (!c) return BM_E_ARG;
    cudaSetDevice(c->device);
    c->d_krows.release(); c->d_frows.release(); c->frows_k = 0; c->fine_tried_k = 0;
    c->d_rows.release(); c->d_zeros.release(); c->d_distok.release(); c->d_distok_p.release(); c->d_k2i.release(); c->d_stage.release();
    c->have_index = false; c->zeros_for_d = -1.f; c->krows_k = 0; c->have_map = false;
    return BM_OK;
}

}  // extern "C"
