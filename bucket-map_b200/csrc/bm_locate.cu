// bm_locate.cu -- the locator: read k-mer sampling (K3a), in-bucket occurrence search + ordered offset vote (K3),
// per-read ordering (gather) and the best-location merge (K3b).
//
// Replaces bucket_locator::_prepare_read_query / _create_kmer_index / _find_offset / _locate / _filter_best_locations
// (locator/bucket_locator.h:162-405,613-705).
//
// The reference rebuilds an unordered_multimap of all ~65.8 k k-mers of a bucket for every non-empty bucket (its
// largest single cost) and probes it with the P sampled k-mers of each read.  Here the roles are swapped: the
// candidates are grouped by bucket (radix sort), one CTA takes one bucket and up to CH of its candidates, puts their
// <= CH*P query k-mers in a small shared-memory hash set, streams the bucket's 2-bit text through it ONCE with a
// rolling hash, and collects the (query slot, offset) hits.  A bitonic sort orders the hits by (slot, descending
// offset) -- the order libstdc++'s unordered_multimap::equal_range yields equal keys (newest insertion first, SURVEY
// section 0) -- and one thread per candidate then replays the reference's order-dependent vote loop literally on a
// small sorted array.  Vote maps that outgrow the array are replayed on a dense per-position table in global memory.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include "bm_internal.cuh"

namespace {

constexpr int kLocThreads = 256;
constexpr int kTsBits = 11;
constexpr uint32_t kTS = 1u << kTsBits;        // hash-set slots
constexpr uint32_t kMaxQueries = 1024;         // CH * P
constexpr uint32_t kMaxCH = 128;               // candidates per CTA pass
constexpr uint32_t kHC = 4096;                 // hits kept in shared memory
constexpr uint32_t kVC = 8;                    // vote-map entries per candidate in shared memory
constexpr uint32_t kEmpty = 0xFFFFFFFFu;
constexpr uint32_t kNoSlot = 0xFFFFu;
constexpr uint32_t kSlotBits = kTsBits + 1;    // slot ids 0..kTS (kTS = the all-T k-mer when k == 16)
constexpr uint32_t kPoolSlots = 32;

constexpr int kSampWarps = 4;

// ---- K3a: P sampled (index, hash) pairs per segment (bucket_locator.h:315-341) -------------------------------------------
__global__ void __launch_bounds__(kSampWarps * 32)
k_loc_sample(const ReadPlanes rp, const uint32_t* __restrict__ seg_read, const uint32_t* __restrict__ seg_start,
             const uint32_t* __restrict__ seg_len, uint32_t n_seg, uint32_t max_seg_len, uint32_t k, uint32_t P,
             uint16_t* __restrict__ sidx, uint32_t* __restrict__ shash) {
    extern __shared__ __align__(16) uint8_t smem[];
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t seg = blockIdx.x * kSampWarps + warp;
    if (seg >= n_seg) return;
    const uint32_t nchunk_max = (max_seg_len + 31) / 32 + 1;
    const uint32_t per_warp = ((max_seg_len + 3) & ~3u) + 2 * nchunk_max * 4;
    SegStage st;
    st.base = smem + (size_t)warp * per_warp;
    st.qkm = reinterpret_cast<uint32_t*>(st.base + ((max_seg_len + 3) & ~3u));
    st.mask = st.qkm + nchunk_max;
    const uint32_t n = seg_len[seg], rd = seg_read[seg];
    stage_segment(st, rp, rp.pl_off[rd], rp.pl_off[rd + 1], seg_start[seg], n, lane);
    const uint32_t nk = (n >= k) ? n - k + 1 : 0;
    const uint32_t nchunk = (nk + 31) / 32;
    uint32_t G = 0;
    for (uint32_t c = 0; c < nchunk; c++) {
        uint32_t j = c * 32 + lane;
        bool good = j < nk && kmer_qual_ok(st, j);                      // :323-325 (quality only)
        uint32_t m = __ballot_sync(0xFFFFFFFFu, good);
        if (lane == 0) st.mask[c] = m;
        G += __popc(m);
    }
    __syncwarp();
    if (G == 0) {                                                       // :328-330: no good k-mer -> all k-mers
        for (uint32_t c = lane; c < nchunk; c += 32) {
            uint32_t rem = nk - c * 32;
            st.mask[c] = rem >= 32 ? 0xFFFFFFFFu : ((1u << rem) - 1u);
        }
        G = nk;
        __syncwarp();
    }
    for (uint32_t s = lane; s < P; s += 32) {
        uint32_t pos = 0, h = kEmpty;                                   // segment shorter than k: no usable sample
        if (nk) {
            pos = select_good(st.mask, nchunk, sample_index(s, P, G));  // :331-333
            h = kmer_hash_at(st.base, pos, k);
        }
        sidx[(size_t)seg * P + s] = (uint16_t)pos;                      // uint16_t like good_indices (:327)
        shash[(size_t)seg * P + s] = h;
    }
}

__global__ void k_gather_u32(const uint32_t* __restrict__ src, const uint32_t* __restrict__ idx, uint32_t n,
                             uint32_t* __restrict__ dst) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[idx[i]];
}

__global__ void k_iota(uint32_t* __restrict__ v, uint32_t n) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) v[i] = i;
}

// ---- work list: (bucket-aligned) chunks of <= CH candidates over the bucket-sorted candidate array -------------------------
__global__ void k_loc_chunks(const uint32_t* __restrict__ sbucket, uint32_t n, uint32_t CH, uint2* __restrict__ chunks,
                             uint32_t* __restrict__ n_chunks) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t b = sbucket[i];
    if (i > 0 && sbucket[i - 1] == b) return;              // not the head of its bucket run
    uint32_t lo = i, hi = n;                                // first index with sbucket > b
    while (lo < hi) { uint32_t mid = (lo + hi) >> 1; if (sbucket[mid] <= b) lo = mid + 1; else hi = mid; }
    const uint32_t end = lo, cnt = (end - i + CH - 1) / CH;
    uint32_t base = atomicAdd(n_chunks, cnt);
    for (uint32_t c = 0; c < cnt; c++) chunks[base + c] = make_uint2(i + c * CH, min(i + (c + 1) * CH, end));
}

// ---- K3: one CTA per chunk --------------------------------------------------------------------------------------------------
struct LocArgs {
    const uint8_t* packed; const uint64_t* bstart; const uint32_t* blen; uint32_t n_emit;
    const uint32_t* sbucket; const uint32_t* scand; const uint2* chunks; const uint32_t* n_chunks;
    const uint32_t* cseg; const uint32_t* seg_len; const uint16_t* sidx; const uint32_t* shash;
    uint32_t k, P; int A; uint32_t min_votes; uint32_t posbits;
    uint32_t* pool_flags; uint32_t* pool; uint32_t hits_cap, dense_cap;
    int32_t* res_pos; uint32_t* res_votes;
};

__device__ __forceinline__ void bitonic_sort(uint32_t* a, uint32_t n_pad) {
    for (uint32_t kk = 2; kk <= n_pad; kk <<= 1)
        for (uint32_t j = kk >> 1; j > 0; j >>= 1) {
            for (uint32_t i = threadIdx.x; i < n_pad; i += kLocThreads) {
                uint32_t x = i ^ j;
                if (x > i) {
                    uint32_t u = a[i], v = a[x];
                    bool up = (i & kk) == 0;
                    if ((u > v) == up) { a[i] = v; a[x] = u; }
                }
            }
            __syncthreads();
        }
}

__device__ __forceinline__ uint32_t lower_bound_u32(const uint32_t* a, uint32_t n, uint32_t key) {
    uint32_t lo = 0, hi = n;
    while (lo < hi) { uint32_t mid = (lo + hi) >> 1; if (a[mid] < key) lo = mid + 1; else hi = mid; }
    return lo;
}

// rolling k-mer hash over the bucket text; every position whose k-mer is in the query set appends (slot, offset)
__device__ __forceinline__ void scan_bucket(const LocArgs& a, const uint32_t* s_keys, uint32_t* s_nh, uint32_t* dst,
                                            uint32_t cap, uint64_t start, uint32_t npos, uint32_t kmask,
                                            uint32_t posmask, bool allT) {
    const uint32_t k = a.k;
    const uint32_t per = (npos + kLocThreads - 1) / kLocThreads;
    const uint32_t lo = threadIdx.x * per, hi = min(lo + per, npos);
    if (lo >= hi) return;
    uint64_t g = start + lo;
    uint32_t byte = a.packed[g >> 2];
    uint32_t h = 0;
    for (uint32_t i = 0; i + 1 < k; i++) {
        h = (h << 2) | ((byte >> ((g & 3) << 1)) & 3u);
        g++;
        if ((g & 3) == 0) byte = a.packed[g >> 2];
    }
    for (uint32_t pos = lo; pos < hi; pos++) {
        h = ((h << 2) | ((byte >> ((g & 3) << 1)) & 3u)) & kmask;
        g++;
        if ((g & 3) == 0) byte = a.packed[g >> 2];
        uint32_t slot = kNoSlot;
        if (h == kEmpty) { if (allT) slot = kTS; }
        else {
            uint32_t s = (h * 0x9E3779B1u) >> (32 - kTsBits);
            while (true) {
                uint32_t key = s_keys[s];
                if (key == h) { slot = s; break; }
                if (key == kEmpty) break;
                s = (s + 1) & (kTS - 1);
            }
        }
        if (slot != kNoSlot) {
            uint32_t j = atomicAdd(s_nh, 1u);
            if (j < cap) dst[j] = (slot << a.posbits) | (posmask - pos);
        }
    }
}

__global__ void __launch_bounds__(kLocThreads)
k_locate(LocArgs a) {
    __shared__ uint32_t s_keys[kTS];
    __shared__ uint32_t s_hits[kHC];
    __shared__ uint16_t s_qslot[kMaxQueries];
    __shared__ int32_t s_vpos[kMaxCH * kVC];
    __shared__ uint32_t s_vcnt[kMaxCH * kVC];
    __shared__ uint32_t s_over[kMaxCH];
    __shared__ uint32_t s_nh, s_allT, s_nover, s_pool;
    __shared__ unsigned long long s_best;

    const uint32_t tid = threadIdx.x;
    const uint32_t k = a.k, P = a.P;
    const uint32_t kmask = (k >= 16) ? 0xFFFFFFFFu : ((1u << (2 * k)) - 1u);
    const uint32_t posmask = (1u << a.posbits) - 1u;
    const uint32_t n_chunks = *a.n_chunks;

    for (uint32_t w = blockIdx.x; w < n_chunks; w += gridDim.x) {
        const uint2 ch = a.chunks[w];
        const uint32_t bucket = a.sbucket[ch.x];
        const uint32_t nc = ch.y - ch.x;
        if (bucket >= a.n_emit) {                  // phantom bucket (SURVEY 8a note 4): the reference would index out of range
            for (uint32_t c = tid; c < nc; c += kLocThreads) {
                uint32_t cand = a.scand[ch.x + c];
                a.res_pos[cand] = -1; a.res_votes[cand] = 0;
            }
            continue;
        }
        for (uint32_t i = tid; i < kTS; i += kLocThreads) s_keys[i] = kEmpty;
        if (tid == 0) { s_nh = 0; s_allT = 0; s_nover = 0; s_pool = kEmpty; }
        __syncthreads();

        // 1. query k-mers -> hash set (bucket_locator.h:239-243 gives the strand-adjusted k-mer)
        for (uint32_t qi = tid; qi < nc * P; qi += kLocThreads) {
            const uint32_t c = qi / P, si = qi - c * P;
            const uint32_t list = a.cseg[a.scand[ch.x + c]];
            uint32_t h = a.shash[(size_t)(list >> 1) * P + si];
            uint32_t slot = kNoSlot;
            if (h != kEmpty || k == 16) {
                if (list & 1u) h = bm_revcomp_hash(h, k);
                if (h == kEmpty) { s_allT = 1; slot = kTS; }
                else {
                    uint32_t s = (h * 0x9E3779B1u) >> (32 - kTsBits);
                    while (true) {
                        uint32_t old = atomicCAS(&s_keys[s], kEmpty, h);
                        if (old == kEmpty || old == h) break;
                        s = (s + 1) & (kTS - 1);
                    }
                    slot = s;
                }
            }
            s_qslot[qi] = (uint16_t)slot;
        }
        __syncthreads();

        // 2. stream the bucket text through the set (replaces _create_kmer_index, :162-177); hits that do not fit in
        //    shared memory are rare (repeats): the pass is then repeated into a global scratch region from the pool
        const uint64_t start = a.bstart[bucket];
        const uint32_t len = a.blen[bucket];
        const uint32_t npos = (len >= k) ? len - k + 1 : 0;
        scan_bucket(a, s_keys, &s_nh, s_hits, kHC, start, npos, kmask, posmask, s_allT != 0);
        __syncthreads();

        // 3. order the hits: slot ascending, offset descending
        const uint32_t nh = s_nh;
        uint32_t* H = s_hits;
        if (nh > kHC) {
            __syncthreads();
            if (tid == 0) {
                uint32_t t = blockIdx.x % kPoolSlots;
                while (atomicCAS(&a.pool_flags[t], 0u, 1u) != 0u) t = (t + 1) % kPoolSlots;
                __threadfence();
                s_pool = t; s_nh = 0;
            }
            __syncthreads();
            H = a.pool + (size_t)s_pool * (a.hits_cap + a.dense_cap);
            scan_bucket(a, s_keys, &s_nh, H, a.hits_cap, start, npos, kmask, posmask, s_allT != 0);
            __syncthreads();
        }
        uint32_t n_pad = 1;
        while (n_pad < nh) n_pad <<= 1;
        for (uint32_t i = nh + tid; i < n_pad; i += kLocThreads) H[i] = kEmpty;
        __syncthreads();
        if (nh > 1) bitonic_sort(H, n_pad);

        // 4. one thread per candidate replays _find_offset's vote loop (:233-287)
        if (tid < nc) {
            const uint32_t cand = a.scand[ch.x + tid];
            const uint32_t list = a.cseg[cand];
            const uint32_t seg = list >> 1;
            const bool reverse = (list & 1u) != 0;
            const uint32_t slen = a.seg_len[seg];
            int32_t* vp = s_vpos + tid * kVC;
            uint32_t* vc = s_vcnt + tid * kVC;
            uint32_t nv = 0;
            bool over = false;
            for (uint32_t i = 0; i < P && !over; i++) {
                const uint32_t si = reverse ? P - 1 - i : i;
                const uint32_t slot = s_qslot[tid * P + si];
                if (slot == kNoSlot) continue;
                uint32_t cur = a.sidx[(size_t)seg * P + si];
                if (reverse) cur = slen - k - cur;
                uint32_t x = lower_bound_u32(H, nh, slot << a.posbits);
                const bool first_mode = nv == 0;
                for (; x < nh && (H[x] >> a.posbits) == slot; x++) {
                    const int32_t position = (int32_t)((posmask - (H[x] & posmask)) - cur);
                    uint32_t e = 0;
                    bool voted = false;
                    if (first_mode) {
                        while (e < nv && vp[e] < position) e++;
                        if (e < nv && vp[e] == position) { vc[e]++; voted = true; }
                    } else {
                        while (e < nv && vp[e] < position - a.A) e++;
                        uint32_t f = e;
                        for (; f < nv && vp[f] <= position + a.A; f++) { vc[f]++; voted = true; }
                        while (e < nv && vp[e] < position) e++;   // insertion point if nothing voted
                    }
                    if (!voted) {
                        if (nv == kVC) { over = true; break; }
                        for (uint32_t t = nv; t > e; t--) { vp[t] = vp[t - 1]; vc[t] = vc[t - 1]; }
                        vp[e] = position; vc[e] = 1; nv++;
                    }
                }
            }
            if (over) s_over[atomicAdd(&s_nover, 1u)] = tid;
            else {
                int32_t bp = -1; uint32_t bv = 0;
                for (uint32_t e = 0; e < nv; e++) if (vc[e] > bv) { bv = vc[e]; bp = vp[e]; }   // ties: smallest key
                if (nv && bv >= a.min_votes && bp >= 0) { a.res_pos[cand] = bp; a.res_votes[cand] = bv; }
                else { a.res_pos[cand] = -1; a.res_votes[cand] = 0; }
            }
        }
        __syncthreads();

        // 5. vote maps that did not fit: dense per-position table in the pool region
        const uint32_t nover = s_nover;
        if (nover) {
            if (tid == 0 && s_pool == kEmpty) {
                uint32_t t = blockIdx.x % kPoolSlots;
                while (atomicCAS(&a.pool_flags[t], 0u, 1u) != 0u) t = (t + 1) % kPoolSlots;
                __threadfence();
                s_pool = t;
            }
            __syncthreads();
            uint32_t* dense = a.pool + (size_t)s_pool * (a.hits_cap + a.dense_cap) + a.hits_cap;
            for (uint32_t oi = 0; oi < nover; oi++) {
                const uint32_t ci = s_over[oi];
                const uint32_t cand = a.scand[ch.x + ci];
                const uint32_t list = a.cseg[cand];
                const uint32_t seg = list >> 1;
                const bool reverse = (list & 1u) != 0;
                const uint32_t slen = a.seg_len[seg];
                const int32_t shift = (int32_t)slen;                      // position >= -(slen - k)
                const uint32_t range = len + slen + 1;
                for (uint32_t i = tid; i < range; i += kLocThreads) dense[i] = 0;
                if (tid == 0) s_best = 0ull;
                __syncthreads();
                if (tid == 0) {
                    uint32_t nkeys = 0;
                    for (uint32_t i = 0; i < P; i++) {
                        const uint32_t si = reverse ? P - 1 - i : i;
                        const uint32_t slot = s_qslot[ci * P + si];
                        if (slot == kNoSlot) continue;
                        uint32_t cur = a.sidx[(size_t)seg * P + si];
                        if (reverse) cur = slen - k - cur;
                        uint32_t x = lower_bound_u32(H, nh, slot << a.posbits);
                        const bool first_mode = nkeys == 0;
                        for (; x < nh && (H[x] >> a.posbits) == slot; x++) {
                            const int32_t position = (int32_t)((posmask - (H[x] & posmask)) - cur);
                            const int32_t ip = position + shift;
                            if (first_mode) { if (dense[ip]++ == 0) nkeys++; }
                            else {
                                bool voted = false;
                                for (int32_t p = ip - a.A; p <= ip + a.A; p++)
                                    if (p >= 0 && p < (int32_t)range && dense[p]) { dense[p]++; voted = true; }
                                if (!voted) { dense[ip] = 1; nkeys++; }
                            }
                        }
                    }
                }
                __syncthreads();
                // argmax votes, ties -> smallest position: maximise (votes << 32) | (~index)
                unsigned long long best = 0ull;
                for (uint32_t i = tid; i < range; i += kLocThreads) {
                    uint32_t v = dense[i];
                    if (v) { unsigned long long key = ((unsigned long long)v << 32) | (0xFFFFFFFFu - i); if (key > best) best = key; }
                }
                if (best) atomicMax(&s_best, best);
                __syncthreads();
                if (tid == 0) {
                    unsigned long long b = s_best;
                    uint32_t bv = (uint32_t)(b >> 32);
                    int32_t bp = (int32_t)(0xFFFFFFFFu - (uint32_t)b) - shift;
                    if (b && bv >= a.min_votes && bp >= 0) { a.res_pos[cand] = bp; a.res_votes[cand] = bv; }
                    else { a.res_pos[cand] = -1; a.res_votes[cand] = 0; }
                }
                __syncthreads();
            }
        }
        if (tid == 0 && s_pool < kEmpty - 1) { __threadfence(); atomicExch(&a.pool_flags[s_pool], 0u); }
        __syncthreads();
    }
}

// ---- gather: per-read location lists in the reference's order (_locate, :651-695) ------------------------------------------
// ascending bucket; inside a bucket forward segments in segment order, then reverse-strand segments in reversed order.
struct GatherArgs {
    const uint64_t* read_off; const uint32_t* read_seg0; const uint32_t* seg_start; const uint32_t* seg_len;
    const uint32_t* cand_off; const uint32_t* cand; const uint32_t* cseg; const int32_t* res_pos;
    const uint32_t* res_votes; uint32_t n_reads;
};

__device__ __forceinline__ unsigned long long gather_key(uint32_t bucket, uint32_t list, uint32_t seg0) {
    uint32_t seg = (list >> 1) - seg0;
    uint32_t sk = (list & 1u) ? (0x80000000u | (0x7FFFFFFFu - seg)) : seg;
    return ((unsigned long long)bucket << 32) | sk;
}

template <bool WRITE>
__global__ void k_loc_gather(GatherArgs g, uint32_t* __restrict__ cnt, const uint32_t* __restrict__ loc_off,
                             bm_loc* __restrict__ locs) {
    uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= g.n_reads) return;
    const uint32_t s0 = g.read_seg0[r], s1 = g.read_seg0[r + 1];
    const uint32_t c0 = g.cand_off[2 * s0], c1 = g.cand_off[2 * s1];
    if (!WRITE) {
        uint32_t n = 0;
        for (uint32_t c = c0; c < c1; c++) n += g.res_pos[c] > 0;        // :674,686: offset 0 is dropped
        cnt[r] = n;
        return;
    }
    uint32_t o = loc_off[r];
    const uint32_t n = loc_off[r + 1] - o;
    const uint32_t rl = (uint32_t)(g.read_off[r + 1] - g.read_off[r]);
    unsigned long long last = 0ull; bool have_last = false;
    for (uint32_t e = 0; e < n; e++) {
        unsigned long long bk = ~0ull; uint32_t bc = 0;
        for (uint32_t c = c0; c < c1; c++) {
            if (g.res_pos[c] <= 0) continue;
            unsigned long long key = gather_key(g.cand[c], g.cseg[c], s0);
            if (have_last && key <= last) continue;
            if (key < bk) { bk = key; bc = c; }
        }
        last = bk; have_last = true;
        const uint32_t list = g.cseg[bc], seg = list >> 1;
        bm_loc l;
        l.bucket = g.cand[bc];
        l.seg_offset = g.seg_start[seg];
        l.votes = g.res_votes[bc];
        if (list & 1u) { l.offset = g.res_pos[bc] - (int32_t)(rl - g.seg_start[seg] - g.seg_len[seg]); l.is_original = 0; }   // :687-691
        else { l.offset = g.res_pos[bc] - (int32_t)g.seg_start[seg]; l.is_original = 1; }                                    // :675-679
        locs[o + e] = l;
    }
}

// ---- K3b: _filter_best_locations (:350-405), one thread per read ----------------------------------------------------------------
__device__ __forceinline__ bool loc_key_less(const bm_loc& x, uint32_t b, int32_t off, uint32_t orig) {
    if (x.bucket != b) return x.bucket < b;
    if (x.offset != off) return x.offset < off;
    return x.is_original < orig;          // std::tuple<unsigned,int,bool>: false (reverse) sorts first
}

__global__ void k_filter_best(const uint64_t* __restrict__ read_off, uint32_t n_reads, float indel_rate,
                              const uint32_t* __restrict__ loc_off, const bm_loc* __restrict__ locs,
                              bm_loc* __restrict__ work, uint32_t* __restrict__ cnt) {
    uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_reads) return;
    const uint32_t o = loc_off[r], n = loc_off[r + 1] - o;
    const uint32_t rl = (uint32_t)(read_off[r + 1] - read_off[r]);
    const float win = __fmul_rn((float)rl, indel_rate);                  // :365-366, float, no contraction
    bm_loc* m = work + o;                                                 // the map, kept sorted by key
    uint32_t nm = 0;
    for (uint32_t i = 0; i < n; i++) {
        const bm_loc l = locs[o + i];
        bool found = false;
        if (nm) {
            const int32_t lower = (int32_t)__fsub_rn((float)l.offset, win);
            const int32_t upper = (int32_t)__fadd_rn((float)l.offset, win);
            for (uint32_t e = 0; e < nm; e++)
                if (m[e].bucket == l.bucket && m[e].offset <= upper && m[e].offset >= lower && m[e].is_original == l.is_original) {
                    m[e].votes += l.votes; found = true;
                }
        }
        if (!found) {
            uint32_t e = 0;
            while (e < nm && loc_key_less(m[e], l.bucket, l.offset, l.is_original)) e++;
            for (uint32_t t = nm; t > e; t--) m[t] = m[t - 1];
            m[e] = l; m[e].seg_offset = 0; nm++;
        }
    }
    uint32_t mx = 0;
    for (uint32_t e = 0; e < nm; e++) if (m[e].votes > mx) mx = m[e].votes;
    uint32_t k = 0;
    for (uint32_t e = 0; e < nm; e++) if (m[e].votes == mx) { if (k != e) m[k] = m[e]; k++; }
    cnt[r] = k;
}

__global__ void k_loc_compact(const uint32_t* __restrict__ src_off, const uint32_t* __restrict__ dst_off,
                              const bm_loc* __restrict__ src, bm_loc* __restrict__ dst, uint32_t n_reads) {
    uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_reads) return;
    uint32_t s = src_off[r], d = dst_off[r], n = dst_off[r + 1] - d;
    for (uint32_t i = 0; i < n; i++) dst[d + i] = src[s + i];
}

}  // namespace

extern "C" {

int bm_locate_device(bm_ctx* c, const bm_params* p, int filter_best) {
    if (!c) return BM_E_ARG;
    if (!p) BM_FAIL(c, BM_E_ARG, "bm_locate_device: null parameters");
    if (!c->have_map) BM_FAIL(c, BM_E_ARG, "bm_locate_device: bm_map_device first");
    if (!c->d_packed.p) BM_FAIL(c, BM_E_ARG, "bm_locate_device: bm_ref_upload first");
    const uint32_t P = p->locator_samples, k = p->k;
    if (P == 0 || P > 64) BM_FAIL(c, BM_E_LIMIT, "locator samples must be in [1,64]");
    if (k == 0 || k > 16) BM_FAIL(c, BM_E_LIMIT, "query seed length must be <= 16");
    uint32_t posbits = 1;
    while ((1ull << posbits) < (uint64_t)c->max_blen + 1) posbits++;
    if (posbits + kSlotBits > 32) BM_FAIL(c, BM_E_LIMIT, "bucket length too large for the locator (max 2^20 bases)");
    cudaSetDevice(c->device);
    c->have_locs = false; c->have_aln = false;
    { int rp = bm_internal_pack_check(c, p); if (rp) return rp; }
    { int rg = bm_internal_gather_fence(c); if (rg) return rg; }
    const uint32_t n_seg = c->n_seg, n_reads = c->n_reads;
    const uint64_t n_cand = c->n_cand;

    BM_CK(c, c->d_lsamp_idx.ensure((size_t)n_seg * P * 2 + 16));
    BM_CK(c, c->d_lsamp_hash.ensure((size_t)n_seg * P * 4 + 16));
    BM_CK(c, c->d_loc_cnt.ensure((size_t)(n_reads + 1) * 4));
    BM_CK(c, c->d_loc_off.ensure((size_t)(n_reads + 1) * 4));
    BM_CK(c, c->d_locs.ensure((size_t)(n_cand + 1) * sizeof(bm_loc)));
    BM_CK(c, cudaMemsetAsync(c->d_loc_cnt.p, 0, (size_t)(n_reads + 1) * 4, c->stream));
    // the fused mapper kernel has already drawn these samples when it ran with the same (k, P, b) on this batch
    const bool have_samples = c->lsamp_k == k && c->lsamp_P == P && c->lsamp_b == p->avg_base_quality;
    if (n_seg && !have_samples) {
        const uint32_t msl = c->max_seg_len ? c->max_seg_len : 1;
        size_t smem = (size_t)kSampWarps * ((((size_t)msl + 3) & ~(size_t)3) + 2 * ((msl + 31) / 32 + 1) * 4);
        if (smem > 48 * 1024)
            BM_CK(c, cudaFuncSetAttribute(k_loc_sample, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        KScope ks(c, BM_K_LOCSAMPLE);
        const ReadPlanes rp{c->d_pl.as<uint2>(), c->have_qk ? c->d_qk.as<uint32_t>() : nullptr, c->d_pl_off.as<uint32_t>()};
        k_loc_sample<<<(n_seg + kSampWarps - 1) / kSampWarps, kSampWarps * 32, smem, c->stream>>>(
            rp, c->d_seg_read.as<uint32_t>(), c->d_seg_start.as<uint32_t>(), c->d_seg_len.as<uint32_t>(), n_seg, msl, k, P,
            c->d_lsamp_idx.as<uint16_t>(), c->d_lsamp_hash.as<uint32_t>());
    }
    BM_CK(c, cudaGetLastError());
    if (n_cand) {
        const uint32_t n_all = (uint32_t)n_cand;
        BM_CK(c, c->d_cres_pos.ensure((size_t)n_all * 4));
        BM_CK(c, c->d_cres_votes.ensure((size_t)n_all * 4));
        // K3i: occurrence search on the genome k-mer position index; what it cannot finish in its small vote arrays
        // (repeats) is left in over_list for the bucket-scan kernel below
        const uint32_t* keys_in = c->d_cand.as<uint32_t>();   // bucket of every candidate to scan
        const uint32_t* vals_in = nullptr;                    // their candidate ids (nullptr: 0..n-1)
        uint32_t n = n_all;
        int ri = bm_internal_gidx_prepare(c, k);
        if (ri) return ri;
        if (c->use_gidx) {
            BM_CK(c, c->d_loc_over.ensure((size_t)n_all * 8 + 64));
            uint32_t* over_list = c->d_loc_over.as<uint32_t>();
            uint32_t* over_bucket = over_list + n_all;
            uint32_t* d_n_over = over_bucket + n_all;
            uint32_t n_over = 0;
            ri = bm_internal_locate_idx(c, p, n_all, over_list, d_n_over, &n_over);
            if (ri) return ri;
            n = n_over;
            if (n) {
                KScope ks(c, BM_K_AUX);
                k_gather_u32<<<(n + 255) / 256, 256, 0, c->stream>>>(c->d_cand.as<uint32_t>(), over_list, n, over_bucket);
                keys_in = over_bucket; vals_in = over_list;
            }
        }
        if (n) {
        // scratch: [chunks (uint2 x n) | iota | sorted bucket | sorted cand | n_chunks]
        BM_CK(c, c->d_vote_scratch.ensure((size_t)n * 4 * 5 + 64));
        uint2* chunks = c->d_vote_scratch.as<uint2>();
        uint32_t* iota = reinterpret_cast<uint32_t*>(chunks + n);
        uint32_t* sbucket = iota + n;
        uint32_t* scand = sbucket + n;
        uint32_t* n_chunks = scand + n;
        BM_CK(c, cudaMemsetAsync(n_chunks, 0, 4, c->stream));
        int end_bit = 1;
        while (end_bit < 32 && (1ull << end_bit) <= (uint64_t)c->NB) end_bit++;
        size_t tmp = 0;
        BM_CK(c, cub::DeviceRadixSort::SortPairs(nullptr, tmp, keys_in, sbucket, iota, scand, (int)n, 0,
                                                 end_bit, c->stream));
        BM_CK(c, c->d_scan_tmp.ensure(tmp));
        const uint32_t CH = (kMaxQueries / P) < kMaxCH ? (kMaxQueries / P) : kMaxCH;
        const uint32_t hits_cap_min = c->max_blen + 1;
        uint32_t hits_cap = kHC * 2;
        while (hits_cap < hits_cap_min) hits_cap <<= 1;
        const uint32_t dense_cap = (c->max_blen + c->max_seg_len + 64 + 3) & ~3u;
        BM_CK(c, c->d_loc_pool.ensure((size_t)kPoolSlots * ((size_t)hits_cap + dense_cap) * 4));
        BM_CK(c, c->d_loc_flags.ensure(kPoolSlots * 4));
        uint32_t* pool_flags = c->d_loc_flags.as<uint32_t>();
        BM_CK(c, cudaMemsetAsync(pool_flags, 0, kPoolSlots * 4, c->stream));
        {
            KScope ks(c, BM_K_AUX, 2);
            if (!vals_in) { k_iota<<<(n + 255) / 256, 256, 0, c->stream>>>(iota, n); vals_in = iota; }
            BM_CK(c, cub::DeviceRadixSort::SortPairs(c->d_scan_tmp.p, tmp, keys_in, sbucket, vals_in, scand,
                                                     (int)n, 0, end_bit, c->stream));
            k_loc_chunks<<<(n + 255) / 256, 256, 0, c->stream>>>(sbucket, n, CH, chunks, n_chunks);
        }
        BM_CK(c, cudaGetLastError());
        LocArgs a;
        a.packed = c->d_packed.as<uint8_t>(); a.bstart = c->d_bstart.as<uint64_t>(); a.blen = c->d_blen.as<uint32_t>();
        a.n_emit = c->n_emit; a.sbucket = sbucket; a.scand = scand; a.chunks = chunks; a.n_chunks = n_chunks;
        a.cseg = c->d_cseg.as<uint32_t>(); a.seg_len = c->d_seg_len.as<uint32_t>();
        a.sidx = c->d_lsamp_idx.as<uint16_t>(); a.shash = c->d_lsamp_hash.as<uint32_t>();
        a.k = k; a.P = P; a.A = (int)p->allowed_indel;
        a.min_votes = (uint32_t)((int)P - (int)p->allowed_mismatch);   // :284, int arithmetic compared as unsigned
        a.posbits = posbits; a.pool_flags = pool_flags; a.pool = c->d_loc_pool.as<uint32_t>();
        a.hits_cap = hits_cap; a.dense_cap = dense_cap;
        a.res_pos = c->d_cres_pos.as<int32_t>(); a.res_votes = c->d_cres_votes.as<uint32_t>();
        {
            KScope ks(c, BM_K_LOCSCAN);
            uint32_t grid = (uint32_t)c->sm_count * 4;
            if (grid > n) grid = n;
            k_locate<<<grid, kLocThreads, 0, c->stream>>>(a);
        }
        BM_CK(c, cudaGetLastError());
        }
    }
    int r = 0;
    if (n_reads) {
        GatherArgs g;
        g.read_off = c->d_read_off.as<uint64_t>(); g.read_seg0 = c->d_read_seg0.as<uint32_t>();
        g.seg_start = c->d_seg_start.as<uint32_t>(); g.seg_len = c->d_seg_len.as<uint32_t>();
        g.cand_off = c->d_cand_off.as<uint32_t>(); g.cand = c->d_cand.as<uint32_t>(); g.cseg = c->d_cseg.as<uint32_t>();
        g.res_pos = c->d_cres_pos.as<int32_t>(); g.res_votes = c->d_cres_votes.as<uint32_t>(); g.n_reads = n_reads;
        const uint32_t blocks = (n_reads + 127) / 128;
        {
            KScope ks(c, BM_K_MERGE);
            k_loc_gather<false><<<blocks, 128, 0, c->stream>>>(g, c->d_loc_cnt.as<uint32_t>(), nullptr, nullptr);
        }
        r = bm_internal_exclusive_scan(c, c->d_loc_cnt.as<uint32_t>(), c->d_loc_off.as<uint32_t>(), n_reads + 1);
        if (r) return r;
        {
            KScope ks(c, BM_K_MERGE);
            k_loc_gather<true><<<blocks, 128, 0, c->stream>>>(g, nullptr, c->d_loc_off.as<uint32_t>(), c->d_locs.as<bm_loc>());
        }
        BM_CK(c, cudaGetLastError());
        if (filter_best) {
            BM_CK(c, c->d_locs2.ensure((size_t)(n_cand + 1) * sizeof(bm_loc)));
            BM_CK(c, c->d_locs3.ensure((size_t)(n_cand + 1) * sizeof(bm_loc)));
            BM_CK(c, c->d_loc_off2.ensure((size_t)(n_reads + 1) * 4));
            BM_CK(c, cudaMemsetAsync((uint32_t*)c->d_loc_cnt.p + n_reads, 0, 4, c->stream));
            {
                KScope ks(c, BM_K_MERGE);
                k_filter_best<<<blocks, 128, 0, c->stream>>>(c->d_read_off.as<uint64_t>(), n_reads, p->indel_rate,
                                                             c->d_loc_off.as<uint32_t>(), c->d_locs.as<bm_loc>(),
                                                             c->d_locs2.as<bm_loc>(), c->d_loc_cnt.as<uint32_t>());
            }
            r = bm_internal_exclusive_scan(c, c->d_loc_cnt.as<uint32_t>(), c->d_loc_off2.as<uint32_t>(), n_reads + 1);
            if (r) return r;
            {
                KScope ks(c, BM_K_MERGE);
                k_loc_compact<<<blocks, 128, 0, c->stream>>>(c->d_loc_off.as<uint32_t>(), c->d_loc_off2.as<uint32_t>(),
                                                             c->d_locs2.as<bm_loc>(), c->d_locs3.as<bm_loc>(), n_reads);
            }
            BM_CK(c, cudaGetLastError());
        }
    } else {
        BM_CK(c, cudaMemsetAsync(c->d_loc_off.p, 0, 4, c->stream));
        if (filter_best) {
            BM_CK(c, c->d_loc_off2.ensure(4));
            BM_CK(c, cudaMemsetAsync(c->d_loc_off2.p, 0, 4, c->stream));
        }
    }
    c->have_locs = true; c->locs_filtered = filter_best != 0; c->n_locs_known = false;
    return BM_OK;
}

int bm_locate_count(bm_ctx* c, uint64_t* n_locs) {
    if (!c) return BM_E_ARG;
    if (!c->have_locs) BM_FAIL(c, BM_E_ARG, "bm_locate_count: bm_locate_device first");
    cudaSetDevice(c->device);
    if (!c->n_locs_known) {
        uint32_t tot[2] = {0, 0};
        BM_CK(c, cudaMemcpyAsync(&tot[0], c->d_loc_off.as<uint32_t>() + c->n_reads, 4, cudaMemcpyDeviceToHost, c->stream));
        if (c->locs_filtered)
            BM_CK(c, cudaMemcpyAsync(&tot[1], c->d_loc_off2.as<uint32_t>() + c->n_reads, 4, cudaMemcpyDeviceToHost, c->stream));
        BM_CK(c, cudaStreamSynchronize(c->stream));
        c->n_locs = tot[0]; c->n_locs_best = tot[1]; c->n_locs_known = true;
    }
    if (n_locs) *n_locs = c->locs_filtered ? c->n_locs_best : c->n_locs;
    return BM_OK;
}

int bm_locate_device_buffers(bm_ctx* c, void** loc_off, void** locs, uint64_t* n_locs) {
    if (!c) return BM_E_ARG;
    uint64_t n = 0;
    int r = bm_locate_count(c, &n);
    if (r) return r;
    if (loc_off) *loc_off = c->locs_filtered ? c->d_loc_off2.p : c->d_loc_off.p;
    if (locs) *locs = c->locs_filtered ? c->d_locs3.p : c->d_locs.p;
    if (n_locs) *n_locs = n;
    return BM_OK;
}

int bm_locate_download(bm_ctx* c, uint32_t* loc_off, bm_loc* locs, uint64_t loc_cap) {
    if (!c) return BM_E_ARG;
    uint64_t n = 0;
    int r = bm_locate_count(c, &n);
    if (r) return r;
    const void* off = c->locs_filtered ? c->d_loc_off2.p : c->d_loc_off.p;
    const void* src = c->locs_filtered ? c->d_locs3.p : c->d_locs.p;
    if (loc_off) BM_CK(c, cudaMemcpyAsync(loc_off, off, (size_t)(c->n_reads + 1) * 4, cudaMemcpyDeviceToHost, c->stream));
    if (locs && n) {
        if (loc_cap < n) BM_FAIL(c, BM_E_ARG, "bm_locate_download: location buffer too small");
        BM_CK(c, cudaMemcpyAsync(locs, src, (size_t)n * sizeof(bm_loc), cudaMemcpyDeviceToHost, c->stream));
    }
    BM_CK(c, cudaStreamSynchronize(c->stream));
    return BM_OK;
}

}  // extern "C"

int bm_internal_map_locate_serial(bm_ctx* c, const bm_params* p, const uint8_t* seq, const uint8_t* qual,
                                  const uint64_t* read_off, uint32_t n_reads, int filter_best, uint32_t* loc_off,
                                  bm_loc* locs, uint64_t loc_cap, uint64_t* stats) {
    int r = bm_reads_upload(c, p, seq, qual, read_off, n_reads);
    if (r) return r;
    r = bm_map_device(c, p);
    if (r) return r;
    r = bm_locate_device(c, p, filter_best);
    if (r) return r;
    if (stats) {
        BM_CK(c, cudaMemcpyAsync(stats, c->d_stats.p, 8 * 8, cudaMemcpyDeviceToHost, c->stream));
    }
    return bm_locate_download(c, loc_off, locs, loc_cap);
}
