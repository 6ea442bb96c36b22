// bm_locidx.cu -- K3i: the locator's occurrence search on a genome-wide k-mer position index.
//
// The reference builds, for every non-empty bucket, an unordered_multimap of all k-mers of the bucket and probes it with
// the P sampled k-mers of each read (bucket_locator.h:162-177, 233-287).  What _find_offset needs from that map is
// "all offsets of k-mer h inside bucket b, newest insertion (= largest offset) first".  With 180 GB of HBM that answer
// can be kept for the WHOLE reference at once: `pos` holds the start positions of all k-mers of the 2-bit genome sorted
// by (k-mer hash, position) and `off[h]` the first entry of hash h (4^k + 1 entries).  The occurrences of h inside a
// bucket are then the contiguous run of pos[off[h] .. off[h+1]) that falls in [bucket_start, bucket_start + len - k],
// found by two binary searches and walked backwards.  One thread per candidate replays the reference's order-dependent
// vote loop on a small sorted array; candidates whose vote map outgrows it (repeats) are handed to the bucket-scan
// kernel (k_locate), which has the dense fallback.  No per-bucket work is left, so the cost no longer depends on how
// many buckets a batch touches.
//
// Built once per (reference, k) on the device: hash every position, radix-sort (hash, position) pairs, mark run starts.
// 4 B per base for `pos` (6.8 GB on the 1.7 Gbp reference) + 4^k * 4 B for `off`; ~27 GB of scratch during the build.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <thrust/iterator/reverse_iterator.h>

#include <cstdlib>

#include "bm_internal.cuh"

namespace {

constexpr uint32_t kEmptyHash = 0xFFFFFFFFu;
constexpr uint32_t kVCi = 8;                    // vote-map entries per candidate kept in shared memory (P <= 64 by bm_locate_device)

// seqan3 views::kmer_hash(ungapped{k}) of the k-mer starting at base i of the 2-bit genome (k <= 16)
__device__ __forceinline__ uint32_t genome_kmer(const uint8_t* __restrict__ packed, uint64_t i, uint32_t k) {
    const uint64_t b = i >> 2;
    uint64_t w = 0;
#pragma unroll
    for (int t = 0; t < 5; t++) w |= (uint64_t)packed[b + t] << (8 * t);     // 20 bases from base 4*b on (padded tail)
    uint32_t x = (uint32_t)(w >> (2 * (i & 3)));                             // base t of the k-mer at bits 2t
    x = __brev(x);                                                           // base t at the top, bit pairs swapped
    x = ((x >> 1) & 0x55555555u) | ((x & 0x55555555u) << 1);
    return x >> (32 - 2 * k);                                                // first base most significant
}

__global__ void k_gidx_keys(const uint8_t* __restrict__ packed, uint64_t n_pos, uint32_t k, uint32_t* __restrict__ keys,
                            uint32_t* __restrict__ vals) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (; i < n_pos; i += stride) { keys[i] = genome_kmer(packed, i, k); vals[i] = (uint32_t)i; }
}

// Offsets over the composite key (hash << lb) | (position >> bshift): the genome is cut into <= 2^lb blocks, so that the
// run of one hash inside the one or two blocks a bucket touches is a handful of entries instead of ~n_bases / 4^k.
// off[c] = first sorted entry with composite key >= c; runs of absent keys are filled by the thread at the next run start.
__device__ __forceinline__ uint32_t comp_key(uint32_t h, uint32_t pos, uint32_t lb, uint32_t bshift) {
    return (h << lb) | (pos >> bshift);
}
// run starts: off[c] = index of the first entry with composite key c (the table is pre-filled with 0xFFFFFFFF and
// off[n_comp] = n_pos); a reverse running minimum then turns it into "first entry with key >= c" for absent keys too
__global__ void k_gidx_mark(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ pos, uint64_t n_pos,
                            uint32_t n_comp, uint32_t lb, uint32_t bshift, uint32_t* __restrict__ off) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    if (i == 0) off[n_comp] = (uint32_t)n_pos;
    for (; i < n_pos; i += stride) {
        const uint32_t h = comp_key(keys[i], pos[i], lb, bshift);
        if (i == 0 || h != comp_key(keys[i - 1], pos[i - 1], lb, bshift)) off[h] = (uint32_t)i;
    }
}
struct MinOp { __device__ __forceinline__ uint32_t operator()(uint32_t a, uint32_t b) const { return a < b ? a : b; } };

struct IdxArgs {
    const uint32_t* pos; const uint32_t* off; uint32_t lb, bshift;
    const uint64_t* bstart; const uint32_t* blen; uint32_t n_emit;
    const uint32_t* cand; const uint32_t* cseg; const uint32_t* seg_len; const uint16_t* sidx; const uint32_t* shash;
    uint32_t n_cand, k, P; int A; uint32_t min_votes;
    int32_t* res_pos; uint32_t* res_votes; uint32_t* over_list; uint32_t* n_over;
};

// first index in [lo, hi) with pos[idx] >= key
__device__ __forceinline__ uint32_t lb_pos(const uint32_t* __restrict__ pos, uint32_t lo, uint32_t hi, uint64_t key) {
    while (lo < hi) { const uint32_t mid = lo + ((hi - lo) >> 1); if ((uint64_t)__ldg(pos + mid) < key) lo = mid + 1; else hi = mid; }
    return lo;
}

// 16 lanes per candidate run the P occurrence searches side by side (two offsets, then the handful of entries of the
// block(s) fetched together -- a binary search for the bucket start and a forward walk to its end only for longer runs:
// every step is a dependent load, so the searches of one candidate must not be serialised); the first threads of the
// CTA, one per candidate, then replay the vote loop over the runs in the reference's order.
constexpr int kSub = 16;
constexpr uint32_t kMaxP = 64;

template <int kIdxThreads>
__global__ void __launch_bounds__(kIdxThreads)
k_locate_idx(const IdxArgs a) {
    constexpr int kCandPerCta = kIdxThreads / kSub;
    // [sample][candidate] / [entry][candidate]: the vote loop runs in neighbouring lanes, one candidate each
    __shared__ uint32_t s_x0[kMaxP][kCandPerCta], s_x1[kMaxP][kCandPerCta], s_cur[kMaxP][kCandPerCta], s_p0[kMaxP][kCandPerCta];
    __shared__ int32_t s_vp[kVCi][kCandPerCta];
    __shared__ uint32_t s_vc[kVCi][kCandPerCta];
    const uint32_t tid = threadIdx.x, sub = tid % kSub;
    uint32_t grp = tid / kSub;
    uint32_t cand = blockIdx.x * kCandPerCta + grp;
    bool valid = cand < a.n_cand;
    const uint32_t k = a.k, P = a.P;
    uint32_t bucket = 0, list = 0, seg = 0, slen = 0, npos = 0;
    uint64_t start = 0;
    bool reverse = false, phantom = false;
    if (valid) {
        bucket = a.cand[cand];
        phantom = bucket >= a.n_emit;                                   // SURVEY 8a note 4
        list = a.cseg[cand]; seg = list >> 1; reverse = (list & 1u) != 0;
        slen = a.seg_len[seg];
        if (!phantom) {
            start = a.bstart[bucket];
            const uint32_t len = a.blen[bucket];
            npos = (len >= k) ? len - k + 1 : 0;
        }
    }
    for (uint32_t i = sub; i < P; i += kSub) {                          // bucket_locator.h:233-243
        uint32_t x0 = 0, x1 = 0, cur = 0, p0 = 0;                       // p0: the run's last entry, the first one voted on
        if (valid && !phantom && npos) {
            const uint32_t si = reverse ? P - 1 - i : i;
            uint32_t h = a.shash[(size_t)seg * P + si];
            if (h != kEmptyHash) {                                      // else: segment shorter than k
                if (reverse) h = bm_revcomp_hash(h, k);
                cur = a.sidx[(size_t)seg * P + si];
                if (reverse) cur = slen - k - cur;
                // the run of h inside the genome block(s) the bucket's k-mer starts fall into
                const uint64_t end = start + npos;
                const uint32_t c0 = comp_key(h, (uint32_t)start, a.lb, a.bshift), c1 = comp_key(h, (uint32_t)(end - 1), a.lb, a.bshift);
                const uint32_t o0 = __ldg(a.off + c0), o1 = __ldg(a.off + c1 + 1);
                if (o1 - o0 > 4u) {
                    x0 = lb_pos(a.pos, o0, o1, start);
                    x1 = x0;
                    while (x1 < o1 && (uint64_t)__ldg(a.pos + x1) < end) x1++;
                    if (x1 > x0) p0 = __ldg(a.pos + x1 - 1);
                } else if (o0 != o1) {
                    // the usual case, a handful of entries in the block(s): fetch them side by side instead of a
                    // binary search and a walk of dependent loads; the entries inside [start, end) are contiguous
                    uint32_t v[4];
#pragma unroll
                    for (uint32_t t = 0; t < 4; t++) v[t] = (o0 + t < o1) ? __ldg(a.pos + o0 + t) : 0xFFFFFFFFu;
                    uint32_t below = 0, inside = 0;
#pragma unroll
                    for (uint32_t t = 0; t < 4; t++) {
                        const bool live = o0 + t < o1;
                        below += (live && (uint64_t)v[t] < start) ? 1u : 0u;
                        const bool in = live && (uint64_t)v[t] >= start && (uint64_t)v[t] < end;
                        inside += in ? 1u : 0u;
                        if (in) p0 = v[t];                              // ascending: the last one inside stays
                    }
                    x0 = o0 + below; x1 = x0 + inside;
                }
            }
        }
        s_x0[i][grp] = x0; s_x1[i][grp] = x1; s_cur[i][grp] = cur; s_p0[i][grp] = p0;
    }
    __syncthreads();
    // the vote loops of the CTA's candidates run in the first kCandPerCta threads (one warp or less), not in one lane of
    // every 16: the loop is serial per candidate and costs an issue slot per warp instruction however few lanes are on
    if (tid >= kCandPerCta) return;
    grp = tid;
    cand = blockIdx.x * kCandPerCta + grp;
    if (cand >= a.n_cand) return;
    bucket = a.cand[cand];
    if (bucket >= a.n_emit) { a.res_pos[cand] = -1; a.res_votes[cand] = 0; return; }
    start = a.bstart[bucket];
#define VP(e) s_vp[e][grp]
#define VC(e) s_vc[e][grp]
    uint32_t nv = 0;
    bool over = false;
    for (uint32_t i = 0; i < P && !over; i++) {                         // :244-287
        const uint32_t x0 = s_x0[i][grp], x1 = s_x1[i][grp], cur = s_cur[i][grp];
        const bool first_mode = nv == 0;
        for (uint32_t x = x1; x > x0 && !over; x--) {                   // equal_range order: largest offset first
            const uint32_t gp = (x == x1) ? s_p0[i][grp] : __ldg(a.pos + x - 1);
            const int32_t position = (int32_t)((uint32_t)((uint64_t)gp - start) - cur);
            uint32_t e = 0;
            bool voted = false;
            if (first_mode) {
                while (e < nv && VP(e) < position) e++;
                if (e < nv && VP(e) == position) { VC(e)++; voted = true; }
            } else {
                while (e < nv && VP(e) < position - a.A) e++;
                for (uint32_t f = e; f < nv && VP(f) <= position + a.A; f++) { VC(f)++; voted = true; }
                while (e < nv && VP(e) < position) e++;                 // insertion point if nothing voted
            }
            if (!voted) {
                if (nv == kVCi) { over = true; break; }
                for (uint32_t t = nv; t > e; t--) { VP(t) = VP(t - 1); VC(t) = VC(t - 1); }
                VP(e) = position; VC(e) = 1; nv++;
            }
        }
    }
    if (over) { a.over_list[atomicAdd(a.n_over, 1u)] = cand; return; }   // repeats: bucket-scan kernel with the dense table
    int32_t bp = -1; uint32_t bv = 0;
    for (uint32_t e = 0; e < nv; e++) if (VC(e) > bv) { bv = VC(e); bp = VP(e); }   // ties: smallest key
    if (nv && bv >= a.min_votes && bp >= 0) { a.res_pos[cand] = bp; a.res_votes[cand] = bv; }
    else { a.res_pos[cand] = -1; a.res_votes[cand] = 0; }
#undef VP
#undef VC
}

}  // namespace

// Builds (or re-uses) the genome k-mer position index for seed length k; sets c->use_gidx.  Not built when switched off,
// for k > 14 (the offset table would need 4^k entries), for references of 2^32 bases or more, or when the scratch does
// not fit: the locator then scans each candidate bucket's text (k_locate).  Both ways give identical results.
int bm_internal_gidx_prepare(bm_ctx* c, uint32_t k) {
    c->use_gidx = false;
    if (c->gidx_mode == 0 || k == 0 || k > 14 || !c->d_packed.p || c->n_bases < k || c->n_bases >= 0xFFFFFFF0ull) return BM_OK;
    if (c->gidx_k == k && c->d_gidx_pos.p && c->d_gidx_off.p) { c->use_gidx = true; return BM_OK; }
    c->d_gidx_pos.release(); c->d_gidx_off.release(); c->gidx_k = 0;
    const uint64_t n_pos = c->n_bases - k + 1;
    const uint32_t lb = (30 - 2 * k) < 6 ? (30 - 2 * k) : 6;              // log2 of the genome blocks per hash (table <= 4 GB)
    uint32_t bshift = 0;
    while (((c->n_bases - 1) >> bshift) >= (1ull << lb)) bshift++;
    const uint32_t n_codes = 1u << (2 * k + lb);                          // composite keys
    size_t tmp = 0;
    {
        cub::DoubleBuffer<uint32_t> dk(nullptr, nullptr), dv(nullptr, nullptr);
        BM_CK(c, cub::DeviceRadixSort::SortPairs(nullptr, tmp, dk, dv, (uint64_t)n_pos, 0, (int)(2 * k), c->stream));
        auto rit = thrust::make_reverse_iterator((uint32_t*)nullptr);
        size_t tmp2 = 0;
        BM_CK(c, cub::DeviceScan::InclusiveScan(nullptr, tmp2, rit, rit, MinOp(), (int)(n_codes + 1), c->stream));
        if (tmp2 > tmp) tmp = tmp2;                   // one scratch buffer serves the sort, then the scan
    }
    size_t fr = 0, tot = 0;
    BM_CK(c, cudaMemGetInfo(&fr, &tot));
    const uint64_t need = n_pos * 16 + tmp + ((uint64_t)n_codes + 1) * 4 + (64ull << 20);
    if (need > fr / 2 + fr / 4) return BM_OK;                       // leave a quarter of the free memory alone
    DevBuf k0, k1, v0, v1, t;
    cudaError_t e = k0.ensure_exact(n_pos * 4 + 16);
    if (e == cudaSuccess) e = k1.ensure_exact(n_pos * 4 + 16);
    if (e == cudaSuccess) e = v0.ensure_exact(n_pos * 4 + 16);
    if (e == cudaSuccess) e = v1.ensure_exact(n_pos * 4 + 16);
    if (e == cudaSuccess) e = t.ensure_exact(tmp + 16);
    if (e == cudaSuccess) e = c->d_gidx_off.ensure_exact(((size_t)n_codes + 2) * 4);
    if (e != cudaSuccess) {
        cudaGetLastError();
        k0.release(); k1.release(); v0.release(); v1.release(); t.release(); c->d_gidx_off.release();
        return BM_OK;
    }
    int rc = BM_OK;
    {
        KScope ks(c, BM_K_GIDX, 2);   // k_gidx_keys + k_gidx_mark (the sort and the scan are library kernels)
        const uint32_t blocks = (uint32_t)c->sm_count * 16;
        k_gidx_keys<<<blocks, 256, 0, c->stream>>>(c->d_packed.as<uint8_t>(), n_pos, k, k0.as<uint32_t>(), v0.as<uint32_t>());
        cub::DoubleBuffer<uint32_t> dk(k0.as<uint32_t>(), k1.as<uint32_t>()), dv(v0.as<uint32_t>(), v1.as<uint32_t>());
        cudaError_t se = cub::DeviceRadixSort::SortPairs(t.p, tmp, dk, dv, (uint64_t)n_pos, 0, (int)(2 * k), c->stream);
        if (se == cudaSuccess) {
            uint32_t* off = c->d_gidx_off.as<uint32_t>();
            se = cudaMemsetAsync(off, 0xFF, ((size_t)n_codes + 1) * 4, c->stream);
            if (se == cudaSuccess) {
                k_gidx_mark<<<blocks, 256, 0, c->stream>>>(dk.Current(), dv.Current(), n_pos, n_codes, lb, bshift, off);
                se = cudaGetLastError();
            }
            if (se == cudaSuccess) {                  // reverse inclusive min-scan, in place; the sort's scratch is free again
                auto rit = thrust::make_reverse_iterator(off + (size_t)n_codes + 1);
                size_t tmp2 = tmp;
                se = cub::DeviceScan::InclusiveScan(t.p, tmp2, rit, rit, MinOp(), (int)(n_codes + 1), c->stream);
            }
        }
        if (se == cudaSuccess) se = cudaStreamSynchronize(c->stream);
        if (se != cudaSuccess) {
            c->err = std::string("genome k-mer index build: ") + cudaGetErrorString(se);
            rc = BM_E_CUDA;
        } else {
            DevBuf& keep = (dv.Current() == v0.as<uint32_t>()) ? v0 : v1;   // the sorted positions stay resident
            c->d_gidx_pos.p = keep.p; c->d_gidx_pos.cap = keep.cap; c->d_gidx_pos.owned = true;
            keep.p = nullptr; keep.cap = 0;
        }
    }
    k0.release(); k1.release(); v0.release(); v1.release(); t.release();
    if (rc) { c->d_gidx_off.release(); return rc; }
    c->gidx_k = k; c->gidx_lb = lb; c->gidx_bshift = bshift; c->use_gidx = true;
    return BM_OK;
}

// K3i over all n candidates of the last mapper call; *n_over = candidates left for the bucket-scan kernel (their ids are
// in over_list)
int bm_internal_locate_idx(bm_ctx* c, const bm_params* p, uint32_t n, uint32_t* over_list, uint32_t* d_n_over, uint32_t* n_over) {
    IdxArgs a;
    a.pos = c->d_gidx_pos.as<uint32_t>(); a.off = c->d_gidx_off.as<uint32_t>(); a.lb = c->gidx_lb; a.bshift = c->gidx_bshift;
    a.bstart = c->d_bstart.as<uint64_t>(); a.blen = c->d_blen.as<uint32_t>(); a.n_emit = c->n_emit;
    a.cand = c->d_cand.as<uint32_t>(); a.cseg = c->d_cseg.as<uint32_t>(); a.seg_len = c->d_seg_len.as<uint32_t>();
    a.sidx = c->d_lsamp_idx.as<uint16_t>(); a.shash = c->d_lsamp_hash.as<uint32_t>();
    a.n_cand = n; a.k = p->k; a.P = p->locator_samples; a.A = (int)p->allowed_indel;
    a.min_votes = (uint32_t)((int)p->locator_samples - (int)p->allowed_mismatch);   // :284, int arithmetic compared as unsigned
    a.res_pos = c->d_cres_pos.as<int32_t>(); a.res_votes = c->d_cres_votes.as<uint32_t>();
    a.over_list = over_list; a.n_over = d_n_over;
    BM_CK(c, cudaMemsetAsync(d_n_over, 0, 4, c->stream));
    {
        KScope ks(c, BM_K_LOCATE);
        // 16 candidates per CTA: 0.61 ms per 1 M candidates on C2 (8: 0.63, 32: 0.68)
        k_locate_idx<256><<<(n + 15) / 16, 256, 0, c->stream>>>(a);
    }
    BM_CK(c, cudaGetLastError());
    BM_CK(c, cudaMemcpyAsync(n_over, d_n_over, 4, cudaMemcpyDeviceToHost, c->stream));
    if (c->d_stats.p)   // stats[7]: candidates whose vote map outgrew K3i's arrays (finished by the bucket-scan kernel)
        BM_CK(c, cudaMemcpyAsync(c->d_stats.as<unsigned long long>() + 7, d_n_over, 4, cudaMemcpyDeviceToDevice, c->stream));
    BM_CK(c, cudaStreamSynchronize(c->stream));
    return BM_OK;
}
