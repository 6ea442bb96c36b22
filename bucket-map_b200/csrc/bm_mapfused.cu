// bm_mapfused.cu -- K2f: the mapper in ONE kernel when the coarse k-mer table is resident.
//
// One warp per segment does everything q_gram_mapper::query_sequence does for it (q_gram_mapper.h:414-480) and, because
// the locator samples the same k-mers of the same segment with the quality filter alone (bucket_locator.h:315-341), the
// locator's P samples as well:
//
//   1. streaming pass over the segment, 32 positions per step, everything in registers: base ranks become two ballot
//      bit planes, phred ranks a warp prefix sum; with a two-step lag the lane that owns position j reads its q-gram
//      window out of the planes by funnel shift, looks up ONE bit of the (plane-ordered) distinguishability table, and
//      the k-mer tests of :437-442 are an OR over k-q+1 neighbouring bits plus a prefix-sum difference.
//   2. the Sampler picks S good k-mers (utils.h:160-178); their hashes (both strands) and q-gram row ids go to shared
//      memory.
//   3. per strand: S coarse rows, 8 B per lane, all loads in flight together; the samples that hit each group of buckets
//      are counted with a carry-save adder tree on bit planes (7 inputs -> 3 bits in 4 full adders).
//   4. groups with < F coarse misses (rare) are recounted exactly from the .qgram matrix; best_results()' answer
//      (:90-102) is written as the candidate list.
//
// No intermediate arrays: the read bytes come in, candidate lists and locator samples go out.
#include "bm_internal.cuh"

namespace {

constexpr int kFW = 4;   // warps per CTA

struct FusedArgs {
    const uint8_t* seq; const uint8_t* qual; const uint64_t* read_off;
    const uint32_t* seg_read; const uint32_t* seg_start; const uint32_t* seg_len;
    uint32_t seg0, n_seg, NC;                 // NC: plane/mask words per warp in shared memory
    const int32_t* k2i; const uint32_t* distok_p;
    uint32_t k, q, NQ, S, minq, bq, min_good;
    const uint2* crows; uint32_t Wcs2, Wc, gshift;
    const uint32_t* rows32; uint32_t Wp32, NB, F, max_cand;
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

// bm_rank_dna4 (seqan3 dna4 char_to_rank incl. IUPAC first-base convention) as two 32-bit tables of 2-bit entries
// indexed by the low 5 bits of a letter: C,Y,S,B -> 1; G,K -> 2; T,U -> 3; everything else 0
__device__ __forceinline__ uint32_t rank_dna4_lut(uint32_t c) {
    // letters @ABCDEFGHIJKLMNO (0..15): B=2:1 C=3:1 G=7:2 K=11:2      PQRSTUVWXYZ (16..26): S=19:1 T=20:3 U=21:3 Y=25:1
    constexpr uint32_t LO = (1u << 4) | (1u << 6) | (2u << 14) | (2u << 22);
    constexpr uint32_t HI = (1u << 6) | (3u << 8) | (3u << 10) | (1u << 18);
    const uint32_t t = (c & 0x10u) ? HI : LO;
    const uint32_t r = (t >> ((c & 15u) * 2u)) & 3u;
    return ((c & 0xC0u) == 0x40u) ? r : 0u;           // only 0x40..0x7F are letters (either case)
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
    uint32_t* lo = sm + (size_t)warp * (4 * a.NC + 2 * S + 2 * S * NQ);
    uint32_t* hi = lo + a.NC;
    uint32_t* mA = hi + a.NC;                 // mapper: quality AND distinguishability
    uint32_t* mB = mA + a.NC;                 // locator: quality only
    uint32_t* hsh = mB + a.NC;                // [2][S] sample hashes, forward then reverse complement
    int32_t* rid = reinterpret_cast<int32_t*>(hsh + 2 * S);   // [2][S][NQ] q-gram row ids

    const uint32_t n = a.seg_len[seg];
    const uint64_t off = a.read_off[a.seg_read[seg]] + a.seg_start[seg];
    const uint32_t nk = (n >= k) ? n - k + 1 : 0;
    const uint32_t nchunk = (nk + 31) / 32, nbch = (n + 31) / 32;
    const uint32_t qmask = (1u << q) - 1u, nqmask = (1u << NQ) - 1u;

    // ---- 1. streaming pass ---------------------------------------------------------------------------------------------
    uint32_t lo1 = 0, hi1 = 0;                        // planes of chunk c-1
    uint32_t Dp = 0;                                  // q-gram distinguishability bits of chunk c-2
    uint32_t q2 = 0, q1 = 0;                          // phred rank of this lane's position in chunks c-2, c-1
    uint32_t mn2 = 0xFFFFFFFFu, mn1 = 0xFFFFFFFFu;    // lowest phred rank inside chunks c-2, c-1
    uint32_t GA = 0, GB = 0;
    uint32_t klo = 0, khi = 0, kA = 0, kB = 0;        // lane c keeps the plane / mask words of chunk c (c < 32)
    uint32_t nb = 'A', nq = 33;                       // bytes of the NEXT chunk, requested one step ahead
    if (lane < n) { nb = a.seq[off + lane]; nq = a.qual[off + lane]; }
    for (uint32_t c = 0; c <= nbch + 1; c++) {
        const uint32_t i = c * 32 + lane;
        const uint32_t cb = nb, cq = nq;
        nb = 'A'; nq = 33;
        if (i + 32 < n) { nb = a.seq[off + i + 32]; nq = a.qual[off + i + 32]; }
        const uint32_t r = rank_dna4_lut(cb), qv = cq - 33u;   // positions >= n: rank 0, phred 0
        const uint32_t lo0 = __ballot_sync(0xFFFFFFFFu, r & 1u), hi0 = __ballot_sync(0xFFFFFFFFu, r & 2u);
        const uint32_t mn0 = __reduce_min_sync(0xFFFFFFFFu, (i < n) ? qv : 0xFFFFFFFFu);
        if (c < 32) { if (lane == c) { klo = lo0; khi = hi0; } }
        else if (lane == 0) { lo[c] = lo0; hi[c] = hi0; }
        uint32_t Dc = 0;                               // ... of chunk c-1
        if (c >= 1) {                                  // q-grams starting in chunk c-1 (windows reach into chunk c)
            const uint32_t lw = __funnelshift_r(lo1, lo0, lane) & qmask, hw = __funnelshift_r(hi1, hi0, lane) & qmask;
            const uint32_t idx = (hw << q) | lw;
            Dc = __ballot_sync(0xFFFFFFFFu, (__ldg(a.distok_p + (idx >> 5)) >> (idx & 31)) & 1u);
        }
        if (c >= 2 && c - 2 < nchunk) {                // k-mers starting in chunk c-2
            const uint32_t j = (c - 2) * 32 + lane;
            const bool dist = (__funnelshift_r(Dp, Dc, lane) & nqmask) != 0u;     // q_gram_mapper.h:189-196
            bool qok = j < nk;                         // views::kmer_quality window sum >= b*k (:437-442)
            if (min(mn2, mn1) < a.bq) {                // else every base of both chunks has phred >= b: all windows pass
                uint32_t Pi2 = q2, Pi1 = q1;           // inclusive prefix sums of the two chunks
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t t2 = __shfl_up_sync(0xFFFFFFFFu, Pi2, o), t1 = __shfl_up_sync(0xFFFFFFFFu, Pi1, o);
                    if (lane >= o) { Pi2 += t2; Pi1 += t1; }
                }
                const uint32_t tot2 = __shfl_sync(0xFFFFFFFFu, Pi2, 31);
                const uint32_t e = lane + k - 1;
                const uint32_t pa = __shfl_sync(0xFFFFFFFFu, Pi2, e & 31), pb = __shfl_sync(0xFFFFFFFFu, Pi1, e & 31);
                const uint32_t qs = ((e < 32) ? pa : tot2 + pb) - (Pi2 - q2);
                qok = qok && qs >= a.minq;
            }
            const uint32_t b = __ballot_sync(0xFFFFFFFFu, qok), m = __ballot_sync(0xFFFFFFFFu, qok && dist);   // :437-442
            if (c - 2 < 32) { if (lane == c - 2) { kA = m; kB = b; } }
            else if (lane == 0) { mA[c - 2] = m; mB[c - 2] = b; }
            GA += __popc(m); GB += __popc(b);
        }
        lo1 = lo0; hi1 = hi0; Dp = Dc;
        q2 = q1; q1 = qv; mn2 = mn1; mn1 = mn0;
    }
    if (lane <= nbch + 1) { lo[lane] = klo; hi[lane] = khi; }      // chunks 0..31: one store per lane
    if (lane < nchunk) { mA[lane] = kA; mB[lane] = kB; }
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
                rid[s * NQ + i] = r0; rid[(S + s) * NQ + i] = r1;
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
                rid[s * NQ + i] = r0; rid[(S + s) * NQ + i] = r1;
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
                                for (uint32_t i = 0; i < NQ; i++) {
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
            for (uint32_t t = lane; t < S * NQ; t += 32) a.rowidx[(size_t)list * S * NQ + t] = rs[t];
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
        atomicAdd(slot + 1, 2ull * S * a.Wc * 4ull + (unsigned long long)nfine * 4ull + 2ull * n);   // bytes loaded
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
    a.seq = c->d_seq.as<uint8_t>(); a.qual = c->d_qual.as<uint8_t>(); a.read_off = c->d_read_off.as<uint64_t>();
    a.seg_read = c->d_seg_read.as<uint32_t>(); a.seg_start = c->d_seg_start.as<uint32_t>(); a.seg_len = c->d_seg_len.as<uint32_t>();
    a.seg0 = seg0; a.n_seg = seg1; a.NC = (msl + 31) / 32 + 3;
    a.k2i = c->d_k2i.as<int32_t>(); a.distok_p = c->d_distok_p.as<uint32_t>();
    a.k = p->k; a.q = p->q; a.NQ = NQ; a.S = S; a.minq = p->avg_base_quality * p->k; a.bq = p->avg_base_quality; a.min_good = min_good;
    a.crows = c->d_krows.as<uint2>(); a.Wcs2 = c->Wcs / 2; a.Wc = c->Wc; a.gshift = 5u - c->krows_logg;
    a.rows32 = c->d_rows.as<uint32_t>(); a.Wp32 = c->Wp / 4; a.NB = c->NB; a.F = p->mapper_fault; a.max_cand = p->max_candidates;
    a.cand_fixed = c->d_cand_fixed.as<uint32_t>(); a.cand_cnt = c->d_cand_cnt.as<uint32_t>(); a.segok = c->d_segok.as<uint8_t>();
    a.stats = c->d_stats.as<unsigned long long>();
    a.P = (P && P <= 1024 && c->d_lsamp_idx.p && c->d_lsamp_hash.p) ? P : 0;
    a.sidx = c->d_lsamp_idx.as<uint16_t>(); a.shash = c->d_lsamp_hash.as<uint32_t>();
    const uint32_t n_groups = c->Wc * 32;             // coarse bits per row
    a.max_pass = c->map_max_pass >= 0 ? (uint32_t)c->map_max_pass : (n_groups / 8 > 16 ? n_groups / 8 : 16);
    a.rowidx = c->d_rowidx.as<int32_t>(); a.over_list = c->d_map_over.as<uint32_t>(); a.n_over = a.over_list + c->map_over_cap;
    const size_t smem = (size_t)kFW * (4 * a.NC + 2 * S + 2 * S * NQ) * 4;
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
