// bm_mapfused.cu -- K2f: the mapper in ONE kernel when the coarse k-mer table is resident.
//
// One warp per segment does everything q_gram_mapper::query_sequence does for it (q_gram_mapper.h:414-480) and, because
// the locator samples the same k-mers of the same segment with the quality filter alone (bucket_locator.h:315-341), the
// locator's P samples as well.  The segment arrives as bit planes (bm_reads.cu): 12 bytes per 32 bases.
//
//   1. lane c takes chunk c of the segment (planes + k-mer quality word) into shared memory; per chunk, the lane that
//      owns position j funnel-shifts its q-gram window out of the planes and looks up ONE bit of the (plane-ordered)
//      distinguishability table, a ballot makes it a word; is_highly_distinguishable (:189-196) is then an OR of k-q+1
//      shifted copies of those words and the k-mer tests of :437-442 two ANDs, 32 k-mers per lane operation.
//   2. the Sampler picks S good k-mers (utils.h:160-178); their hashes (both strands) go to shared memory.
//   3. per strand: S coarse rows, 8 B per lane, all loads in flight together; the samples that hit each group of buckets
//      are counted with a carry-save adder tree on bit planes (7 inputs -> 3 bits in 4 full adders).
//   4. groups with < F coarse misses (rare) are recounted exactly -- from ONE word per sample of the full-resolution
//      k-mer table when it is resident, else from the k-q+1 .qgram words of every sample; best_results()' answer
//      (:90-102) is written as the candidate list.
//
// No intermediate arrays: the packed read comes in, candidate lists and locator samples go out.
#include "bm_internal.cuh"

namespace {

constexpr int kFW = 4;   // warps per CTA

struct FusedArgs {
    ReadPlanes rp;
    const uint32_t* seg_read; const uint32_t* seg_start; const uint32_t* seg_len;
    uint32_t seg0, n_seg, NC;                 // NC: plane/mask words per warp in shared memory
    const int32_t* k2i; const uint32_t* distok_p;
    uint32_t k, q, NQ, S, min_good;
    const uint2* crows; uint32_t Wcs2, Wc, gshift;
    const uint32_t* rows32; const uint32_t* frows32; uint32_t Wp32, NB, F, max_cand;   // frows32: fine k-mer table or nullptr
    uint32_t* cand_fixed; uint32_t* cand_cnt; uint8_t* segok; unsigned long long* stats;
    uint32_t P; uint16_t* sidx; uint32_t* shash;
    uint32_t max_pass; int32_t* rowidx; uint32_t* over_list; uint32_t* n_over;   // hand-over to K2 (repeats)
};

__device__ __forceinline__ void fa(uint32_t a, uint32_t b, uint32_t c, uint32_t& s, uint32_t& cy) {
    s = a ^ b ^ c;
    cy = (a & b) | (c & (a ^ b));
}
// p += x0 + 2*x1 + 4*x2 (bit-sliced, no overflow by dispatch: NP planes hold S)
template <int NP>
__device__ __forceinline__ void add_w(uint32_t (&p)[NP], uint32_t x0, uint32_t x1, uint32_t x2) {
    uint32_t k0 = p[0] & x0; p[0] ^= x0;
    uint32_t k1, k2;
    fa(p[1], x1, k0, p[1], k1);
    fa(p[2], x2, k1, p[2], k2);
#pragma unroll
    for (int i = 3; i < NP; i++) { uint32_t t = p[i] & k2; p[i] ^= k2; k2 = t; }
}
template <int NP, int O>
__device__ __forceinline__ void add7(uint32_t (&p)[NP], const uint32_t (&m)[16]) {
    uint32_t s1, c1, s2, c2, s3, c3, s4, c4;
    fa(m[O + 0], m[O + 1], m[O + 2], s1, c1);
    fa(m[O + 3], m[O + 4], m[O + 5], s2, c2);
    fa(s1, s2, m[O + 6], s3, c3);
    fa(c1, c2, c3, s4, c4);
    add_w<NP>(p, s3, s4, c4);
}
template <int NP>
__device__ __forceinline__ void add16(uint32_t (&p)[NP], const uint32_t (&m)[16]) {
    add7<NP, 0>(p, m);
    add7<NP, 7>(p, m);
    add_w<NP>(p, m[14] ^ m[15], m[14] & m[15], 0u);
}
template <int NP>
__device__ __forceinline__ uint32_t below(const uint32_t (&p)[NP], uint32_t F) {
    uint32_t lt = 0u, eq = 0xFFFFFFFFu;
#pragma unroll
    for (int i = NP - 1; i >= 0; i--) {
        if ((F >> i) & 1u) { lt |= eq & ~p[i]; eq &= p[i]; }
        else eq &= ~p[i];
    }
    return (F >> NP) ? 0xFFFFFFFFu : lt;
}

// zero bit between neighbouring bits of a 16-bit value
__device__ __forceinline__ uint32_t spread16(uint32_t x) {
    x = (x | (x << 8)) & 0x00FF00FFu;
    x = (x | (x << 4)) & 0x0F0F0F0Fu;
    x = (x | (x << 2)) & 0x33333333u;
    x = (x | (x << 1)) & 0x55555555u;
    return x;
}
// seqan3 views::kmer_hash(ungapped{k}) of the k-mer at `pos`, from the two bit planes (bit t of a plane word = base t)
__device__ __forceinline__ uint32_t hash_from_planes(const uint32_t* lo, const uint32_t* hi, uint32_t pos, uint32_t k) {
    const uint32_t c = pos >> 5, sh = pos & 31;
    const uint32_t lw = __funnelshift_r(lo[c], lo[c + 1], sh), hw = __funnelshift_r(hi[c], hi[c + 1], sh);
    return (spread16(__brev(hw) >> (32 - k)) << 1) | spread16(__brev(lw) >> (32 - k));   // first base most significant
}

// 32x32 bit-matrix transpose across the warp: lane i holds row i (bit j = column j) and ends up with column i
__device__ __forceinline__ uint32_t transpose32(uint32_t x, uint32_t lane) {
    uint32_t m = 0x0000FFFFu;
#pragma unroll
    for (int j = 16; j; j >>= 1) {
        const uint32_t y = __shfl_xor_sync(0xFFFFFFFFu, x, j);
        x = (lane & j) ? ((x & ~m) | ((y >> j) & m)) : ((x & m) | ((y & m) << j));
        m ^= m << (j >> 1);
    }
    return x;
}

template <int NP, int MINB>
__global__ void __launch_bounds__(kFW * 32, MINB)
k_map_fused(const FusedArgs a) {
    extern __shared__ uint32_t sm[];
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t seg = a.seg0 + blockIdx.x * kFW + warp;
    if (seg >= a.n_seg) return;
    const uint32_t S = a.S, NQ = a.NQ, k = a.k, q = a.q;
    uint32_t* lo = sm + (size_t)warp * (5 * a.NC + 2 * S + 2 * S * NQ);
    uint32_t* hi = lo + a.NC;
    uint32_t* Dq = hi + a.NC;                 // q-gram distinguishability bits
    uint32_t* mA = Dq + a.NC;                 // mapper: quality AND distinguishability
    uint32_t* mB = mA + a.NC;                 // locator: quality only
    uint32_t* hsh = mB + a.NC;                // [2][S] sample hashes, forward then reverse complement
    int32_t* rid = reinterpret_cast<int32_t*>(hsh + 2 * S);   // [2][S][NQ] q-gram row ids (only without the fine table)

    const uint32_t n = a.seg_len[seg], rd = a.seg_read[seg], sstart = a.seg_start[seg];
    const uint32_t c0 = __ldg(a.rp.pl_off + rd), c1 = __ldg(a.rp.pl_off + rd + 1);
    const uint32_t nk = (n >= k) ? n - k + 1 : 0;
    const uint32_t nchunk = (nk + 31) / 32, nbch = (n + 31) / 32;
    const uint32_t nqg = (n >= q) ? n - q + 1 : 0, nqch = (nqg + 31) / 32;      // q-gram start positions
    const uint32_t qmask = (1u << q) - 1u;

    // ---- 1a. the segment's chunks -> shared memory (chunk nbch is read by the funnel shifts of the last positions) --------
    for (uint32_t c = lane; c <= nbch; c += 32) {
        uint32_t l, h, qq;
        seg_chunk(a.rp, c0, c1, sstart, c, l, h, qq);
        lo[c] = l; hi[c] = h; mB[c] = qq;
    }
    __syncwarp();
    // ---- 1b. one distinguishability bit per q-gram start (q_gram_mapper.h:189-196, one table bit per q-gram) ---------------
    for (uint32_t c = 0; c < nqch; c++) {
        const uint32_t lw = __funnelshift_r(lo[c], lo[c + 1], lane) & qmask, hw = __funnelshift_r(hi[c], hi[c + 1], lane) & qmask;
        const uint32_t idx = (hw << q) | lw;
        const uint32_t Dc = __ballot_sync(0xFFFFFFFFu, (__ldg(a.distok_p + (idx >> 5)) >> (idx & 31)) & 1u);
        if (lane == 0) Dq[c] = Dc;
    }
    if (lane == 0) Dq[nqch] = 0u;
    __syncwarp();
    // ---- 1c. good k-mers, 32 per lane operation: a k-mer is distinguishable if any of its k-q+1 q-grams is (:189-196) and
    //          passes the quality test if its window's bit of the quality plane is set (:437-442)
    uint32_t GA = 0, GB = 0;
    for (uint32_t c = lane; c < nchunk; c += 32) {
        const uint32_t d0 = Dq[c], d1 = Dq[c + 1];
        uint32_t m = d0;
        for (uint32_t i = 1; i < NQ; i++) m |= __funnelshift_r(d0, d1, i);
        const uint32_t rem = nk - c * 32;
        const uint32_t b = mB[c] & (rem >= 32 ? 0xFFFFFFFFu : ((1u << rem) - 1u));
        m &= b;
        mA[c] = m; mB[c] = b;
        GA += __popc(m); GB += __popc(b);
    }
    GA = __reduce_add_sync(0xFFFFFFFFu, GA); GB = __reduce_add_sync(0xFFFFFFFFu, GB);
    __syncwarp();

    // ---- 2. samples: the locator's P (quality filter only, bucket_locator.h:315-341) and the mapper's S
    //         (q_gram_mapper.h:445-469).  With S, P <= 16 the two samplers run side by side in the two half-warps.
    if (a.P && GB == 0) {                              // :328-330: no good k-mer -> all k-mers
        for (uint32_t c = lane; c < nchunk; c += 32) {
            const uint32_t rem = nk - c * 32;
            mB[c] = rem >= 32 ? 0xFFFFFFFFu : ((1u << rem) - 1u);
        }
        GB = nk;
        __syncwarp();
    }
    const bool ok = GA >= a.min_good && GA > 0;
    const bool fine = a.frows32 != nullptr;
    const uint32_t Q = (q >= 16) ? 0xFFFFFFFFu : ((1u << (2 * q)) - 1u);
    uint32_t gathered = 0;
    if (lane == 0) a.segok[seg] = ok ? 1 : 0;
    if (a.P && a.P <= 16 && S <= 16) {
        const bool isB = lane >= 16;
        const uint32_t s = lane & 15;
        const bool want = isB ? (s < a.P) : (ok && s < S);
        uint32_t pos = 0, h = 0xFFFFFFFFu;             // segment shorter than k: no usable locator sample
        if (want && nk) {
            pos = select_good(isB ? mB : mA, nchunk, sample_index(s, isB ? a.P : S, isB ? GB : GA));
            h = hash_from_planes(lo, hi, pos, k);
        }
        if (isB && want) {
            a.sidx[(size_t)seg * a.P + s] = (uint16_t)pos;
            a.shash[(size_t)seg * a.P + s] = h;
        }
        if (!ok) {
            if (lane < 2) a.cand_cnt[2 * seg + lane] = 0;
            return;
        }
        if (!isB && want) {
            const uint32_t hr = bm_revcomp_hash(h, k);
            hsh[s] = h; hsh[S + s] = hr;
            for (uint32_t i = 0; i < NQ; i++) {
                const int32_t r0 = __ldg(a.k2i + ((h >> (2 * i)) & Q)), r1 = __ldg(a.k2i + ((hr >> (2 * i)) & Q));
                if (!fine) { rid[s * NQ + i] = r0; rid[(S + s) * NQ + i] = r1; }
                gathered += (r0 >= 0) + (r1 >= 0);
            }
        }
    } else {
        for (uint32_t s = lane; s < a.P; s += 32) {
            uint32_t pos = 0, h = 0xFFFFFFFFu;
            if (nk) {
                pos = select_good(mB, nchunk, sample_index(s, a.P, GB));
                h = hash_from_planes(lo, hi, pos, k);
            }
            a.sidx[(size_t)seg * a.P + s] = (uint16_t)pos;
            a.shash[(size_t)seg * a.P + s] = h;
        }
        if (!ok) {
            if (lane < 2) a.cand_cnt[2 * seg + lane] = 0;
            return;
        }
        for (uint32_t s = lane; s < S; s += 32) {
            const uint32_t pos = select_good(mA, nchunk, sample_index(s, S, GA));
            const uint32_t h = hash_from_planes(lo, hi, pos, k);
            const uint32_t hr = bm_revcomp_hash(h, k);
            hsh[s] = h; hsh[S + s] = hr;
            for (uint32_t i = 0; i < NQ; i++) {
                const int32_t r0 = __ldg(a.k2i + ((h >> (2 * i)) & Q)), r1 = __ldg(a.k2i + ((hr >> (2 * i)) & Q));
                if (!fine) { rid[s * NQ + i] = r0; rid[(S + s) * NQ + i] = r1; }
                gathered += (r0 >= 0) + (r1 >= 0);
            }
        }
    }
    __syncwarp();

    // ---- 3. + 4. per strand: coarse vote, exact recount of passing groups ----------------------------------------------------
    uint32_t nfine = 0;
    for (uint32_t strand = 0; strand < 2; strand++) {
        const uint32_t* hs = hsh + strand * S;
        const int32_t* rs = rid + (size_t)strand * S * NQ;
        const uint32_t list = 2 * seg + strand;
        uint32_t best = 0xFFFFFFFFu, count = 0, npass = 0;
        bool handover = false;
        uint32_t* out = a.cand_fixed + (size_t)list * a.max_cand;
        for (uint32_t w0 = 0; 2 * w0 < a.Wc; w0 += 32) {
            const uint32_t w = w0 + lane;              // this lane: coarse words 2w, 2w+1
            const bool act = 2 * w < a.Wc;
            const uint2* colp = a.crows + (act ? w : 0);
            uint32_t px[NP], py[NP];
#pragma unroll
            for (int i = 0; i < NP; i++) { px[i] = 0u; py[i] = 0u; }
            for (uint32_t s0 = 0; s0 < S; s0 += 16) {
                uint32_t mx[16], my[16];               // 1 = some bucket of the group holds the sample
#pragma unroll
                for (int b = 0; b < 16; b++) {
                    mx[b] = 0u; my[b] = 0u;            // no sample / no such word: no hit
                    if (s0 + b < S && act) {
                        const uint2 v = __ldg(colp + (size_t)hs[s0 + b] * a.Wcs2);
                        mx[b] = v.x; my[b] = v.y;
                    }
                }
                add16<NP>(px, mx);
                add16<NP>(py, my);
            }
            // px, py count HITS per group: fewer than F misses <=> at least S-F+1 hits
            const uint32_t need = (S >= a.F) ? S - a.F + 1u : 0u;
            const uint32_t passx = act ? ~below<NP>(px, need) : 0u, passy = act ? ~below<NP>(py, need) : 0u;
            // On a repeat-rich reference a read made of repeat k-mers passes the coarse vote almost everywhere; recounting
            // group after group would then cost more than gathering the full-width rows once: such lists go to K2.
            npass += __reduce_add_sync(0xFFFFFFFFu, __popc(passx) + __popc(passy));
            if (npass > a.max_pass) { handover = true; break; }
            uint32_t any = __ballot_sync(0xFFFFFFFFu, (passx | passy) != 0u);
            while (any) {
                const uint32_t l = __ffs(any) - 1; any &= any - 1;
                const uint32_t pwx = __shfl_sync(0xFFFFFFFFu, passx, l), pwy = __shfl_sync(0xFFFFFFFFu, passy, l);
                uint32_t last_gi = 0xFFFFFFFFu;
                for (uint32_t half = 0; half < 2; half++) {
                    uint32_t pw = half ? pwy : pwx;
                    const uint32_t cw = 2 * (w0 + l) + half;                 // coarse word index
                    while (pw) {
                        const uint32_t g = __ffs(pw) - 1; pw &= pw - 1;
                        // coarse bit -> the 32-bit word of a full row that holds its buckets (32*gi .. 32*gi+31);
                        // with 8 or 16 buckets per coarse bit several bits share a word, which is evaluated once
                        const uint32_t gi = (cw * 32 + g) >> a.gshift;
                        if (gi >= a.Wp32 || gi == last_gi) continue;
                        last_gi = gi;
                        const uint32_t b0 = gi * 32u;
                        const uint32_t vmask = (b0 >= a.NB) ? 0u : (a.NB - b0 >= 32u ? 0xFFFFFFFFu : ((1u << (a.NB - b0)) - 1u));
                        uint32_t hits = 0;             // lane b: samples that hit bucket b0 + b
                        for (uint32_t sb = 0; sb < S; sb += 32) {
                            const uint32_t si = sb + lane;
                            uint32_t word = 0u;
                            if (si < S) {
                                word = vmask;          // q_gram_mapper.h:399-407
                                if (fine) { word &= __ldg(a.frows32 + (size_t)hs[si] * a.Wp32 + gi); nfine++; }
                                else for (uint32_t i = 0; i < NQ; i++) {
                                    const int32_t r = rs[si * NQ + i];
                                    if (r >= 0) { word &= __ldg(a.rows32 + (size_t)r * a.Wp32 + gi); nfine++; }
                                }
                            }
                            hits += __popc(transpose32(word, lane));   // lane b: which of these 32 samples hit bucket b0+b
                        }
                        const uint32_t miss = S - hits;
                        const uint32_t lvl = (b0 + lane < a.NB && miss < a.F) ? miss : 0xFFFFu;
                        const uint32_t tmin = __reduce_min_sync(0xFFFFFFFFu, lvl);
                        if (tmin >= a.F) continue;
                        if (tmin < best) { best = tmin; count = 0; }
                        if (tmin == best) {
                            const uint32_t m = __ballot_sync(0xFFFFFFFFu, lvl == best);
                            const uint32_t pos = count + __popc(m & ((1u << lane) - 1u));
                            if (lvl == best && pos < a.max_cand) out[pos] = b0 + lane;   // ascending, like _Find_next
                            count += __popc(m);
                        }
                    }
                }
            }
        }
        if (handover) {
            for (uint32_t t = lane; t < S * NQ; t += 32) {
                int32_t r = 0;
                if (fine) { const uint32_t si = t / NQ, i = t - si * NQ; r = __ldg(a.k2i + ((hs[si] >> (2 * i)) & Q)); }
                else r = rs[t];
                a.rowidx[(size_t)list * S * NQ + t] = r;
            }
            if (lane == 0) { a.cand_cnt[list] = 0u; a.over_list[atomicAdd(a.n_over, 1u)] = list; }
            continue;
        }
        // :471-476: a list longer than the cap is emptied, not truncated
        if (lane == 0) a.cand_cnt[list] = (best < a.F && count <= a.max_cand) ? count : 0u;
    }
    gathered = __reduce_add_sync(0xFFFFFFFFu, gathered);
    nfine = __reduce_add_sync(0xFFFFFFFFu, nfine);
    if (lane == 0) {
        // 60 counter slots picked by CTA (a million adds on one address would queue up in one L2 slice); folded into
        // stats[3], stats[4] by bm_internal_map_finish
        unsigned long long* slot = a.stats + 8 + 2 * (blockIdx.x % 60u);
        if (gathered) atomicAdd(slot, (unsigned long long)gathered);
        atomicAdd(slot + 1, 2ull * S * a.Wc * 4ull + (unsigned long long)nfine * 4ull + (a.rp.qk ? 12ull : 8ull) * nbch);   // bytes loaded
    }
}

}  // namespace

extern "C" int bm_map_handover(bm_ctx* c, int max_groups) {
    if (!c) return BM_E_ARG;
    c->map_max_pass = max_groups < 0 ? -1 : max_groups;
    return BM_OK;
}

// K2f over segments [seg0, seg1); also leaves the locator's samples for the batch (c->lsamp_*)
int bm_internal_map_fused(bm_ctx* c, const bm_params* p, uint32_t seg0, uint32_t seg1, uint32_t max_seg_len) {
    if (seg1 <= seg0) return BM_OK;
    const uint32_t S = p->mapper_samples, NQ = p->k - p->q + 1, P = p->locator_samples;
    uint32_t min_good = 0;                            // smallest G with !(G < 0.2*S), q_gram_mapper.h:445 in double
    while ((double)min_good < 0.2 * (double)S) min_good++;
    const uint32_t msl = max_seg_len ? max_seg_len : 1;
    FusedArgs a;
    a.rp = ReadPlanes{c->d_pl.as<uint2>(), c->have_qk ? c->d_qk.as<uint32_t>() : nullptr, c->d_pl_off.as<uint32_t>()};
    a.seg_read = c->d_seg_read.as<uint32_t>(); a.seg_start = c->d_seg_start.as<uint32_t>(); a.seg_len = c->d_seg_len.as<uint32_t>();
    a.seg0 = seg0; a.n_seg = seg1; a.NC = (msl + 31) / 32 + 3;
    a.k2i = c->d_k2i.as<int32_t>(); a.distok_p = c->d_distok_p.as<uint32_t>();
    a.k = p->k; a.q = p->q; a.NQ = NQ; a.S = S; a.min_good = min_good;
    a.crows = c->d_krows.as<uint2>(); a.Wcs2 = c->Wcs / 2; a.Wc = c->Wc; a.gshift = 5u - c->krows_logg;
    a.rows32 = c->d_rows.as<uint32_t>(); a.frows32 = c->use_ftable ? c->d_frows.as<uint32_t>() : nullptr; a.Wp32 = c->Wp / 4; a.NB = c->NB; a.F = p->mapper_fault; a.max_cand = p->max_candidates;
    a.cand_fixed = c->d_cand_fixed.as<uint32_t>(); a.cand_cnt = c->d_cand_cnt.as<uint32_t>(); a.segok = c->d_segok.as<uint8_t>();
    a.stats = c->d_stats.as<unsigned long long>();
    a.P = (P && P <= 1024 && c->d_lsamp_idx.p && c->d_lsamp_hash.p) ? P : 0;
    a.sidx = c->d_lsamp_idx.as<uint16_t>(); a.shash = c->d_lsamp_hash.as<uint32_t>();
    const uint32_t n_groups = c->Wc * 32;             // coarse bits per row
    a.max_pass = c->map_max_pass >= 0 ? (uint32_t)c->map_max_pass : (n_groups / 8 > 16 ? n_groups / 8 : 16);
    a.rowidx = c->d_rowidx.as<int32_t>(); a.over_list = c->d_map_over.as<uint32_t>(); a.n_over = a.over_list + c->map_over_cap;
    const size_t smem = (size_t)kFW * (5 * a.NC + 2 * S + 2 * S * NQ) * 4;
    const uint32_t blocks = (seg1 - seg0 + kFW - 1) / kFW;
    const int NP = (S <= 15) ? 4 : (S <= 63 ? 6 : 9);
    KScope ks(c, BM_K_MAP);
#define BM_FUSED(NPv, MB)                                                                                          \
    do {                                                                                                           \
        if (smem > 48 * 1024)                                                                                      \
            BM_CK(c, cudaFuncSetAttribute(k_map_fused<NPv, MB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        k_map_fused<NPv, MB><<<blocks, kFW * 32, smem, c->stream>>>(a);                                            \
    } while (0)
    // min CTAs per SM: 9 x 4 warps at 56 registers; forcing 12 (40 registers) spills and was measured no faster
    if (NP == 4) BM_FUSED(4, 9); else if (NP == 6) BM_FUSED(6, 8); else BM_FUSED(9, 6);
#undef BM_FUSED
    BM_CK(c, cudaGetLastError());
    return BM_OK;
}
