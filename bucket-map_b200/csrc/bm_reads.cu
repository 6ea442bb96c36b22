// bm_reads.cu -- the read batch in its canonical device form (bit planes + k-mer quality plane) and how it gets there.
//
// The mapper's first act on a read is to turn every base into 2 bits and every k-mer window's phred sum into 1 bit
// (q_gram_mapper.h:431-442: views::kmer_hash, views::kmer_quality >= b*k).  That conversion is done ONCE per batch:
//   * ASCII batches (bm_reads_upload, bm_map_locate: bases + phred+33 qualities as the FASTQ holds them) are converted
//     on the device by k_pack_reads right behind their host->device copy -- one coalesced streaming pass, HBM-bound
//     (reads 2 B per base, writes 0.25-0.375 B per base);
//   * packed batches (bm_reads_upload_packed, bm_map_locate_packed) arrive in that form: the FASTQ parser threads of the
//     C++ host pack while they parse (bm_pack_reads below runs on the HOST and touches bytes the parser has in cache
//     anyway), which cuts the PCIe bytes of a 300-bp read from 608 to 92.
// Every kernel of the path (K2f, K2a, K3a, K4) reads only the packed form.
//
// bm_pack_reads is host code by nature (it is the parser's output stage, next to the FASTQ text); it holds no part of
// the mapping computation.
#include <cstring>

#include "bm_internal.cuh"

#if defined(__SSE2__) && !defined(__CUDA_ARCH__)
#include <emmintrin.h>
#define BM_PACK_SSE2 1
#else
#define BM_PACK_SSE2 0
#endif

namespace {

// seqan3 dna4 char_to_rank incl. the IUPAC first-base convention (see bm_rank_dna4): two 32-bit tables of 2-bit entries
// indexed by the low 5 bits of a letter: C,Y,S,B -> 1; G,K -> 2; T,U -> 3; everything else 0
__host__ __device__ __forceinline__ uint32_t rank_dna4_lut(uint32_t c) {
    const uint32_t LO = (1u << 4) | (1u << 6) | (2u << 14) | (2u << 22);
    const uint32_t HI = (1u << 6) | (3u << 8) | (3u << 10) | (1u << 18);
    const uint32_t t = (c & 0x10u) ? HI : LO;
    const uint32_t r = (t >> ((c & 15u) * 2u)) & 3u;
    return ((c & 0xC0u) == 0x40u) ? r : 0u;           // only 0x40..0x7F are letters (either case)
}

// One warp per read: 32 bases per step -> two ballot words; the quality word of chunk c needs the phred ranks of chunks
// c and c+1 (a window spans at most two chunks, k <= 32), so it is written with a lag of one step.  When no base of the
// two chunks is below b every window passes and no sum is formed.
__global__ void __launch_bounds__(128)
k_pack_reads(const uint8_t* __restrict__ seq, const uint8_t* __restrict__ qual, const uint64_t* __restrict__ read_off,
             const uint32_t* __restrict__ pl_off, uint32_t r0, uint32_t r1, uint32_t k, uint32_t bq,
             uint2* __restrict__ pl, uint32_t* __restrict__ qk) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t r = r0 + ((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    if (r >= r1) return;
    const uint64_t off = read_off[r];
    const uint32_t n = (uint32_t)(read_off[r + 1] - off);
    const uint32_t nch = (n + 31) / 32, c0 = pl_off[r];
    const uint32_t minq = bq * k;
    uint32_t q1 = 0, mn1 = 0xFFFFFFFFu;                  // phred rank of this lane's position / minimum of chunk c-1
    for (uint32_t c = 0; c <= nch; c++) {
        const uint32_t i = c * 32 + lane;
        uint32_t b = 'A', qv = 0;
        if (i < n) { b = seq[off + i]; qv = (uint32_t)qual[off + i] - 33u; }
        const uint32_t mn0 = __reduce_min_sync(0xFFFFFFFFu, (i < n) ? qv : 0xFFFFFFFFu);
        if (c < nch) {
            const uint32_t rk = rank_dna4_lut(b);
            const uint32_t lo = __ballot_sync(0xFFFFFFFFu, rk & 1u), hi = __ballot_sync(0xFFFFFFFFu, rk & 2u);
            if (lane == 0) pl[c0 + c] = make_uint2(lo, hi);
        }
        if (c >= 1) {                                    // quality word of chunk c-1
            const uint32_t j = (c - 1) * 32 + lane;      // window [j, j+k)
            uint32_t w = 0xFFFFFFFFu;
            if (min(mn1, mn0) < bq) {
                // inclusive prefix sums of chunks c-1 and c
                uint32_t P1 = q1, P0 = qv;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t t1 = __shfl_up_sync(0xFFFFFFFFu, P1, o), t0 = __shfl_up_sync(0xFFFFFFFFu, P0, o);
                    if (lane >= o) { P1 += t1; P0 += t0; }
                }
                const uint32_t tot1 = __shfl_sync(0xFFFFFFFFu, P1, 31);
                const uint32_t e = lane + k - 1;           // last position of the window, relative to chunk c-1
                const uint32_t pa = __shfl_sync(0xFFFFFFFFu, P1, e & 31), pb = __shfl_sync(0xFFFFFFFFu, P0, e & 31);
                const uint32_t qs = ((e < 32) ? pa : tot1 + pb) - (P1 - q1);
                w = __ballot_sync(0xFFFFFFFFu, qs >= minq);
            }
            // windows that run past the end of the read are not k-mers: their bits are never read (0 for determinism)
            const uint32_t nk = (n >= k) ? n - k + 1 : 0;
            const uint32_t first = (c - 1) * 32;
            const uint32_t valid = nk <= first ? 0u : (nk - first >= 32u ? 0xFFFFFFFFu : ((1u << (nk - first)) - 1u));
            if (lane == 0 && qk) qk[c0 + c - 1] = w & valid;
            (void)j;
        }
        q1 = qv; mn1 = mn0;
    }
}

}  // namespace

int bm_internal_pack_range(bm_ctx* c, const bm_params* p, uint32_t r0, uint32_t r1) {
    if (r1 <= r0) return BM_OK;
    if (p->k == 0 || p->k > 32) BM_FAIL(c, BM_E_LIMIT, "query seed length must be in [1,32]");
    const uint32_t warps = r1 - r0;
    KScope ks(c, BM_K_PACK);
    k_pack_reads<<<(warps + 3) / 4, 128, 0, c->stream>>>(c->d_seq.as<uint8_t>(), c->d_qual.as<uint8_t>(),
                                                         c->d_read_off.as<uint64_t>(), c->d_pl_off.as<uint32_t>(), r0, r1,
                                                         p->k, p->avg_base_quality, c->d_pl.as<uint2>(), c->d_qk.as<uint32_t>());
    BM_CK(c, cudaGetLastError());
    return BM_OK;
}

// The quality plane depends on (k, b).  An ASCII batch is re-packed when a call comes with other values; a batch that
// arrived packed cannot be (its qualities never reached the device).
int bm_internal_pack_check(bm_ctx* c, const bm_params* p) {
    if (!c->have_qk && !c->ascii_resident) {
        // packed batch declared "every window passes" for the b it was packed with: holds for any b' <= b
        if (p->avg_base_quality > c->pack_b)
            BM_FAIL(c, BM_E_ARG, "the resident batch was packed for a lower -b: pack it again for this -b");
        return BM_OK;
    }
    if (c->pack_k == p->k && c->pack_b == p->avg_base_quality) return BM_OK;
    if (!c->ascii_resident)
        BM_FAIL(c, BM_E_ARG, "the resident batch was packed for another (-l, -b): pack it again for these parameters");
    int r = bm_internal_pack_range(c, p, 0, c->n_reads);
    if (r) return r;
    c->pack_k = p->k; c->pack_b = p->avg_base_quality; c->have_qk = true;
    c->lsamp_k = c->lsamp_P = c->lsamp_b = 0;            // locator samples drawn with the old plane are stale
    return BM_OK;
}

extern "C" {

// ---- host-side packer (the FASTQ parser's output stage) ------------------------------------------------------------------
int bm_pack_layout(const uint64_t* read_off, uint32_t n_reads, uint32_t* pl_off) {
    if (!read_off || !pl_off) return BM_E_ARG;
    uint64_t o = 0;
    for (uint32_t i = 0; i < n_reads; i++) {
        pl_off[i] = (uint32_t)o;
        if (read_off[i + 1] < read_off[i]) return BM_E_ARG;
        o += (read_off[i + 1] - read_off[i] + 31) / 32;
        if (o > 0xFFFFFFFFull) return BM_E_LIMIT;
    }
    pl_off[n_reads] = (uint32_t)o;
    return BM_OK;
}

// one read: n bases at s, nq <= n quality characters at q (the rest count as '!', phred 0); returns 1 if a window fails
static inline uint32_t pack_one(const uint8_t* lut, const uint8_t* s, const uint8_t* q, uint32_t n, uint32_t nq, uint32_t k,
                                uint32_t b, uint64_t* out, uint32_t* qo) {
    const uint32_t nch = (n + 31) / 32;
    for (uint32_t c = 0; c < nch; c++) {
        const uint32_t lim = n - 32 * c < 32 ? n - 32 * c : 32;
        uint32_t lo = 0, hi = 0;
        const uint8_t* t = s + 32 * c;
        uint32_t i = 0;
#if BM_PACK_SSE2
        // 16 bases per step while they are plain A/C/G/T (either case): with c1, c2 = bits 1, 2 of the character the rank is
        // (c2, c1 ^ c2), and a byte-mask move collects one bit of 16 characters.  Anything else takes the table below.
        for (; i + 16 <= lim; i += 16) {
            const __m128i x = _mm_loadu_si128(reinterpret_cast<const __m128i*>(t + i));
            const __m128i u = _mm_and_si128(x, _mm_set1_epi8((char)0xDF));
            const __m128i ok = _mm_or_si128(_mm_or_si128(_mm_cmpeq_epi8(u, _mm_set1_epi8('A')), _mm_cmpeq_epi8(u, _mm_set1_epi8('C'))),
                                            _mm_or_si128(_mm_cmpeq_epi8(u, _mm_set1_epi8('G')), _mm_cmpeq_epi8(u, _mm_set1_epi8('T'))));
            if (_mm_movemask_epi8(ok) != 0xFFFF) break;
            const uint32_t m1 = (uint32_t)_mm_movemask_epi8(_mm_slli_epi16(x, 6));      // bit 1 of every character
            const uint32_t m2 = (uint32_t)_mm_movemask_epi8(_mm_slli_epi16(x, 5));      // bit 2
            lo |= (m1 ^ m2) << i; hi |= m2 << i;
        }
#endif
        for (; i < lim; i++) { const uint32_t v = lut[t[i]]; lo |= (v & 1u) << i; hi |= (v >> 1) << i; }
        out[c] = (uint64_t)lo | ((uint64_t)hi << 32);
    }
    if (!q) return 0;
    uint32_t mn = nq < n ? 0u : 0xFFFFFFFFu;
    {
        uint32_t i = 0;
#if BM_PACK_SSE2
        // smallest quality character, 16 at a time; a character below '!' counts as a huge phred in the loop below (32-bit
        // wrap), so a block that holds one is left to it
        __m128i m = _mm_set1_epi8((char)0xFF);
        uint32_t i16 = 0;
        for (; i16 + 16 <= nq; i16 += 16) m = _mm_min_epu8(m, _mm_loadu_si128(reinterpret_cast<const __m128i*>(q + i16)));
        if (i16) {
            m = _mm_min_epu8(m, _mm_srli_si128(m, 8)); m = _mm_min_epu8(m, _mm_srli_si128(m, 4));
            m = _mm_min_epu8(m, _mm_srli_si128(m, 2)); m = _mm_min_epu8(m, _mm_srli_si128(m, 1));
            const uint32_t lowest = (uint32_t)_mm_cvtsi128_si32(m) & 0xFFu;
            if (lowest >= 33u) { const uint32_t v = lowest - 33u; mn = v < mn ? v : mn; i = i16; }
        }
#endif
        for (; i < nq; i++) { const uint32_t v = (uint32_t)q[i] - 33u; mn = v < mn ? v : mn; }
    }
    const uint32_t nk = (n >= k) ? n - k + 1 : 0;
    if (mn >= b || nk == 0) {                                     // every window passes
        if (qo) for (uint32_t c = 0; c < nch; c++) {
            const uint32_t first = 32 * c;
            qo[c] = nk <= first ? 0u : (nk - first >= 32u ? 0xFFFFFFFFu : ((1u << (nk - first)) - 1u));
        }
        return 0;
    }
    const uint32_t minq = b * k;
    bool any_fail = false;
    if (qo) memset(qo, 0, (size_t)nch * 4);
    auto ph = [&](uint32_t i) -> uint32_t { return i < nq ? (uint32_t)q[i] - 33u : 0u; };
    uint32_t sum = 0;
    for (uint32_t i = 0; i + 1 < k; i++) sum += ph(i);
    for (uint32_t j = 0; j < nk; j++) {                           // quality_filter.h:531-534: window sum of phred ranks
        sum += ph(j + k - 1);
        if (sum >= minq) { if (qo) qo[j >> 5] |= 1u << (j & 31); } else any_fail = true;
        sum -= ph(j);
    }
    return any_fail ? 1u : 0u;
}

static const uint8_t* pack_lut() {
    static uint8_t lut[256];
    static bool ready = false;
    if (!ready) { for (int c = 0; c < 256; c++) lut[c] = (uint8_t)rank_dna4_lut((uint32_t)c); ready = true; }
    return lut;
}

int bm_pack_reads(const uint8_t* seq, const uint8_t* qual, const uint64_t* read_off, const uint32_t* pl_off, uint32_t r0,
                  uint32_t r1, uint32_t k, uint32_t b, uint64_t* planes, uint32_t* qk, uint32_t* n_low) {
    if (!seq || !read_off || !pl_off || !planes || k == 0 || k > 32) return BM_E_ARG;
    const uint8_t* lut = pack_lut();
    uint32_t low = 0;
    for (uint32_t r = r0; r < r1; r++) {
        const uint32_t n = (uint32_t)(read_off[r + 1] - read_off[r]);
        low += pack_one(lut, seq + read_off[r], qual ? qual + read_off[r] : nullptr, n, n, k, b, planes + pl_off[r],
                        qk ? qk + pl_off[r] : nullptr);
    }
    if (n_low) *n_low = low;
    return BM_OK;
}

// the same for reads that are not concatenated: read i has read_off[i+1] - read_off[i] bases at seq[i] and qual_len[i]
// (NULL: as many as bases) quality characters at qual[i] -- views into a FASTQ buffer, nothing is copied first
int bm_pack_reads_v(const uint8_t* const* seq, const uint8_t* const* qual, const uint32_t* qual_len, const uint64_t* read_off,
                    const uint32_t* pl_off, uint32_t r0, uint32_t r1, uint32_t k, uint32_t b, uint64_t* planes, uint32_t* qk,
                    uint32_t* n_low) {
    if (!seq || !read_off || !pl_off || !planes || k == 0 || k > 32) return BM_E_ARG;
    const uint8_t* lut = pack_lut();
    uint32_t low = 0;
    for (uint32_t r = r0; r < r1; r++) {
        const uint32_t n = (uint32_t)(read_off[r + 1] - read_off[r]);
        uint32_t nq = qual_len ? qual_len[r] : n;
        if (nq > n) nq = n;
        low += pack_one(lut, seq[r], qual ? qual[r] : nullptr, n, nq, k, b, planes + pl_off[r], qk ? qk + pl_off[r] : nullptr);
    }
    if (n_low) *n_low = low;
    return BM_OK;
}

}  // extern "C"
