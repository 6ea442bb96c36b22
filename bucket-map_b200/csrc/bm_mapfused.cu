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
#include <cstdlib>

#include "bm_internal.cuh"

namespace {

constexpr int kFW = 4;   // warps per CTA

struct FusedArgs {
    ReadPlanes rp;
    const uint32_t* seg_read; const uint32_t* seg_start; const uint32_t* seg_len;
    uint32_t seg0, n_seg, NC, S4, SW;         // NC: plane/mask words per warp in shared memory; S4: hash row stride; SW: words per warp
    const int32_t* k2i; const uint32_t* distok_p;
    uint32_t k, q, NQ, S, min_good;
    const uint2* crows; uint32_t Wcs2, Wc, gshift, zero_row;   // zero_row: index of the all-zero row behind the coarse table
    const uint32_t* rows32; const uint32_t* frows32; uint32_t Wp32, NB, F, max_cand;   // frows32: fine k-mer table or nullptr
    uint32_t* cand_fixed; uint32_t* cand_cnt; uint8_t* segok; unsigned long long* stats;
    uint32_t P; uint16_t* sidx; uint32_t* shash;
    uint32_t ldmode, pf_lines, pf_inv;
    uint32_t fm[9];                           // bit planes of the hits a group needs: S - F + 1
    uint32_t all_distok;                      // every q-gram is indexed and passes -d: no distinguishability look-ups needed
    uint32_t all_indexed;                     // every q-gram has a row (-f 1): no kmer_to_index look-ups needed for the counter
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
// hits of up to 15 samples per group as 4 bit planes: a carry-save tree of 11 full adders (m[15] is ignored)
__device__ __forceinline__ void count15(const uint32_t (&m)[16], uint32_t (&p)[4]) {
    uint32_t sa, ca, sb, cb, sc, cc, sd, cd, se, ce, sf, cf, cg, sh, ch, si, ci, cj;
    fa(m[0], m[1], m[2], sa, ca);
    fa(m[3], m[4], m[5], sb, cb);
    fa(m[6], m[7], m[8], sc, cc);
    fa(m[9], m[10], m[11], sd, cd);
    fa(m[12], m[13], m[14], se, ce);
    fa(sa, sb, sc, sf, cf);
    fa(sd, se, sf, p[0], cg);
    fa(ca, cb, cc, sh, ch);
    fa(cd, ce, cf, si, ci);
    fa(sh, si, cg, p[1], cj);
    fa(ch, ci, cj, p[2], p[3]);
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

// groups with at least `need` hits, need given as one all-ones / all-zeros mask per bit plane (fm[i] = bit i of need):
// the bit-sliced form of hits >= need, two logic operations per plane
template <int NP>
__device__ __forceinline__ uint32_t at_least(const uint32_t (&p)[NP], const uint32_t* fm) {
    uint32_t lt = 0u, eq = 0xFFFFFFFFu;
#pragma unroll
    for (int i = NP - 1; i >= 0; i--) {
        lt |= eq & ~p[i] & fm[i];
        eq &= ~(p[i] ^ fm[i]);
    }
    return ~lt;
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
__device__ __forceinline__ uint32_t hash_from_planes(const uint2* pl, uint32_t pos, uint32_t k) {
    const uint32_t c = pos >> 5, sh = pos & 31;
    const uint2 p0 = pl[c], p1 = pl[c + 1];
    const uint32_t lw = __funnelshift_r(p0.x, p1.x, sh), hw = __funnelshift_r(p0.y, p1.y, sh);
    return (spread16(__brev(hw) >> (32 - k)) << 1) | spread16(__brev(lw) >> (32 - k));   // first base most significant
}

// one scattered 32-bit word of the full-resolution table (tuning: the L2 prefetch-size hint decides how much of the 128-B
// line a miss brings in from DRAM)
__device__ __forceinline__ uint32_t ld_word(const uint32_t* p, uint32_t mode) {
    uint32_t v;
    if (mode == 1) asm volatile("ld.global.nc.L2::64B.u32 %0, [%1];" : "=r"(v) : "l"(p));
    else if (mode == 2) asm volatile("ld.global.L2::64B.u32 %0, [%1];" : "=r"(v) : "l"(p));
    else if (mode == 3) asm volatile("ld.global.nc.L1::no_allocate.L2::64B.u32 %0, [%1];" : "=r"(v) : "l"(p));
    else v = __ldg(p);
    return v;
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
    extern __shared__ uint4 sm4[];
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t seg = a.seg0 + blockIdx.x * kFW + warp;
    if (seg >= a.n_seg) return;
    const uint32_t S = a.S, NQ = a.NQ, k = a.k, q = a.q;
    // per-warp shared memory (a.SW words, a multiple of 4): [2][S4] sample hashes (16-B aligned rows) | planes | D | mA | mB | rid
    uint32_t* hsh = reinterpret_cast<uint32_t*>(sm4) + (size_t)warp * a.SW;
    uint2* pls = reinterpret_cast<uint2*>(hsh + 2 * a.S4);   // {lo, hi} bit planes of the segment, 32 bases per entry
    uint32_t* Dq = reinterpret_cast<uint32_t*>(pls + a.NC);   // q-gram distinguishability bits
    uint32_t* mA = Dq + a.NC;                 // mapper: quality AND distinguishability
    uint32_t* mB = mA + a.NC;                 // locator: quality only
    int32_t* rid = reinterpret_cast<int32_t*>(mB + a.NC);     // [2][S][NQ] q-gram row ids (only without the fine table)

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
        pls[c] = make_uint2(l, h); mB[c] = qq;
    }
    __syncwarp();
    // ---- 1b. one distinguishability bit per q-gram start (q_gram_mapper.h:189-196, one table bit per q-gram) ---------------
    if (a.all_distok) {
        // every one of the 4^q table bits is set (the usual case at -f 1): the bits of positions past the last q-gram start are
        // never looked at (k-mer p < nk reads q-gram bits p .. p+k-q <= n-q), so whole words of ones do
        for (uint32_t c = lane; c <= nqch; c += 32) Dq[c] = 0xFFFFFFFFu;
    } else {
        uint2 p0 = pls[0];
        for (uint32_t c = 0; c < nqch; c += 4) {
            uint32_t bit[4];
#pragma unroll
            for (int t = 0; t < 4; t++) {
                bit[t] = 0u;
                if (c + t < nqch) {
                    const uint2 p1 = pls[c + t + 1];
                    const uint32_t lw = __funnelshift_r(p0.x, p1.x, lane) & qmask, hw = __funnelshift_r(p0.y, p1.y, lane) & qmask;
                    const uint32_t idx = hw * (qmask + 1u) + lw;
                    bit[t] = (__ldg(a.distok_p + (idx >> 5)) >> (idx & 31)) & 1u;
                    p0 = p1;
                }
            }
#pragma unroll
            for (int t = 0; t < 4; t++) {
                const uint32_t Dc = __ballot_sync(0xFFFFFFFFu, bit[t]);
                if (lane == (uint32_t)t && c + t < nqch) Dq[c + t] = Dc;
            }
        }
        if (lane == 0) Dq[nqch] = 0u;
    }
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
    if (lane >= S && lane < a.S4) { hsh[lane] = a.zero_row; hsh[a.S4 + lane] = a.zero_row; }   // S <= 15: see the coarse vote
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
            h = hash_from_planes(pls, pos, k);
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
            hsh[s] = h; hsh[a.S4 + s] = hr;
            if (fine && a.all_indexed) gathered += 2 * NQ;
            else for (uint32_t i = 0; i < NQ; i++) {
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
                h = hash_from_planes(pls, pos, k);
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
            const uint32_t h = hash_from_planes(pls, pos, k);
            const uint32_t hr = bm_revcomp_hash(h, k);
            hsh[s] = h; hsh[a.S4 + s] = hr;
            if (fine && a.all_indexed) gathered += 2 * NQ;
            else for (uint32_t i = 0; i < NQ; i++) {
                const int32_t r0 = __ldg(a.k2i + ((h >> (2 * i)) & Q)), r1 = __ldg(a.k2i + ((hr >> (2 * i)) & Q));
                if (!fine) { rid[s * NQ + i] = r0; rid[(S + s) * NQ + i] = r1; }
                gathered += (r0 >= 0) + (r1 >= 0);
            }
        }
    }
    __syncwarp();

    // ---- 3. + 4. per strand: coarse vote, exact recount of passing groups ----------------------------------------------------
    // The reverse strand's coarse rows are requested into L2 now, one (sample, 128-B line) per lane, so that they arrive
    // while the forward strand is voted on and recounted: one DRAM round trip less on the warp's critical path.
    {
        const uint32_t lines = a.pf_lines;             // 128-B lines per coarse row; t / lines == (t * pf_inv) >> 16
        for (uint32_t t = lane; t < S * lines; t += 32) {
            const uint32_t si = (t * a.pf_inv) >> 16, ln = t - si * lines;
            const uint2* pp = a.crows + (size_t)hsh[a.S4 + si] * a.Wcs2 + ln * 16u;
            asm volatile("prefetch.global.L2 [%0];" ::"l"(pp));
        }
    }
    uint32_t nfine = 0, npass_all = 0;
    for (uint32_t strand = 0; strand < 2; strand++) {
        const uint32_t* hs = hsh + strand * a.S4;
        const int32_t* rs = rid + (size_t)strand * S * NQ;
        const uint32_t list = 2 * seg + strand;
        uint32_t best = 0xFFFFFFFFu, count = 0, npass = 0;
        bool handover = false;
        uint32_t* out = a.cand_fixed + (size_t)list * a.max_cand;
        for (uint32_t w0 = 0; 2 * w0 < a.Wc; w0 += 32) {
            const uint32_t w = w0 + lane;              // this lane: coarse words 2w, 2w+1
            const bool act = 2 * w < a.Wc;
            const uint2* colp = a.crows + (act ? w : 0);   // idle lanes read column 0; their result is dropped below
            uint32_t px[NP], py[NP];
            if constexpr (NP == 4) {
                // S <= 15: all rows in flight at once, hashes fetched four at a time, one exact 15-input counter
                uint32_t mx[16], my[16];               // 1 = some bucket of the group holds the sample
                const uint4* h4 = reinterpret_cast<const uint4*>(hs);
#pragma unroll
                for (int b4 = 0; b4 < 4; b4++) {
                    const uint4 hv = h4[b4];           // entries past S name the all-zero row behind the table
                    const uint32_t hh[4] = {hv.x, hv.y, hv.z, hv.w};
#pragma unroll
                    for (int t = 0; t < 4; t++) {
                        const uint2 v = __ldg(colp + (size_t)hh[t] * a.Wcs2);
                        mx[4 * b4 + t] = v.x; my[4 * b4 + t] = v.y;
                    }
                }
                count15(mx, px);
                count15(my, py);
            } else {
#pragma unroll
                for (int i = 0; i < NP; i++) { px[i] = 0u; py[i] = 0u; }
                for (uint32_t s0 = 0; s0 < S; s0 += 16) {
                    uint32_t mx[16], my[16];
#pragma unroll
                    for (int b = 0; b < 16; b++) {
                        mx[b] = 0u; my[b] = 0u;        // no sample: no hit
                        if (s0 + b < S) {
                            const uint2 v = __ldg(colp + (size_t)hs[s0 + b] * a.Wcs2);
                            mx[b] = v.x; my[b] = v.y;
                        }
                    }
                    add16<NP>(px, mx);
                    add16<NP>(py, my);
                }
            }
            // px, py count HITS per group: fewer than F misses <=> at least S-F+1 hits
            const uint32_t passx = act ? at_least<NP>(px, a.fm) : 0u, passy = act ? at_least<NP>(py, a.fm) : 0u;
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
                                if (fine) { word &= ld_word(a.frows32 + (size_t)hs[si] * a.Wp32 + gi, a.ldmode); nfine++; }
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
        npass_all += npass;
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
        // stats[3], stats[4], stats[6] by bm_internal_map_finish
        unsigned long long* slot = a.stats + 8 + 3 * (blockIdx.x % 60u);
        if (gathered) atomicAdd(slot, (unsigned long long)gathered);
        if (npass_all) atomicAdd(slot + 2, (unsigned long long)npass_all);      // coarse groups that needed an exact recount
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
    a.zero_row = 1u << (2 * p->k);
    a.rows32 = c->d_rows.as<uint32_t>(); a.frows32 = c->use_ftable ? c->d_frows.as<uint32_t>() : nullptr; a.Wp32 = c->Wp / 4; a.NB = c->NB; a.F = p->mapper_fault; a.max_cand = p->max_candidates;
    a.cand_fixed = c->d_cand_fixed.as<uint32_t>(); a.cand_cnt = c->d_cand_cnt.as<uint32_t>(); a.segok = c->d_segok.as<uint8_t>();
    a.stats = c->d_stats.as<unsigned long long>();
    a.P = (P && P <= 1024 && c->d_lsamp_idx.p && c->d_lsamp_hash.p) ? P : 0;
    a.sidx = c->d_lsamp_idx.as<uint16_t>(); a.shash = c->d_lsamp_hash.as<uint32_t>();
    const uint32_t n_groups = c->Wc * 32;             // coarse bits per row
    a.max_pass = c->map_max_pass >= 0 ? (uint32_t)c->map_max_pass : (n_groups / 8 > 16 ? n_groups / 8 : 16);
    a.rowidx = c->d_rowidx.as<int32_t>(); a.over_list = c->d_map_over.as<uint32_t>(); a.n_over = a.over_list + c->map_over_cap;
    a.S4 = (S + 3u) & ~3u;
    if (a.S4 < 16 && S <= 15) a.S4 = 16;                  // the S <= 15 path reads four uint4 of hashes per strand
    a.SW = (2 * a.S4 + 5 * a.NC + 2 * S * NQ + 3u) & ~3u;
    const size_t smem = (size_t)kFW * a.SW * 4;
    const uint32_t blocks = (seg1 - seg0 + kFW - 1) / kFW;
    const int NP = (S <= 15) ? 4 : (S <= 63 ? 6 : 9);
    KScope ks(c, BM_K_MAP);
#define BM_FUSED(NPv, MB)                                                                                          \
    do {                                                                                                           \
        if (smem > 48 * 1024)                                                                                      \
            BM_CK(c, cudaFuncSetAttribute(k_map_fused<NPv, MB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        k_map_fused<NPv, MB><<<blocks, kFW * 32, smem, c->stream>>>(a);                                            \
    } while (0)
    // min CTAs per SM (occupancy against registers per thread); $BM_K2F_MINB overrides the default for tuning runs
    {
        const uint32_t need = (S >= p->mapper_fault) ? S - p->mapper_fault + 1u : 0u;   // fewer than F misses <=> >= S-F+1 hits
        for (int i = 0; i < 9; i++) a.fm[i] = ((need >> i) & 1u) ? 0xFFFFFFFFu : 0u;
    }
    a.all_indexed = (c->n_rows == (1u << (2 * c->q))) ? 1u : 0u;
    // dist_valid counts the rows with MORE zeros than the threshold (the logged counter); the table bit asks for >=
    a.all_distok = (a.all_indexed && c->dist_valid == c->n_rows && !getenv("BM_K2F_NO_ALLDIST")) ? 1u : 0u;
    a.pf_lines = (a.Wc * 4u + 127u) / 128u;
    a.pf_inv = 65535u / a.pf_lines + 1u;              // exact for t < 65536 / lines; S * lines <= 256 * 64 stays far below
    if ((uint64_t)S * a.pf_lines * a.pf_lines >= 65536) { a.pf_lines = 0; a.pf_inv = 0; }   // (t * inv) >> 16 would be off: no prefetch
    const char* lde = getenv("BM_K2F_LD");
    a.ldmode = lde ? (uint32_t)atoi(lde) : 3u;
    const char* mbe = getenv("BM_K2F_MINB");
    const int minb = mbe ? atoi(mbe) : 0;
    if (NP == 4) {
        if (minb == 8) BM_FUSED(4, 8); else if (minb == 10) BM_FUSED(4, 10); else if (minb == 9) BM_FUSED(4, 9);
        else if (minb == 14) BM_FUSED(4, 14);
        else BM_FUSED(4, 12);                          // 48 warps per SM at 40 registers: measured fastest (2.36 vs 2.55 ms at 9)
    } else if (NP == 6) BM_FUSED(6, 8); else BM_FUSED(9, 6);
#undef BM_FUSED
    BM_CK(c, cudaGetLastError());
    return BM_OK;
}
