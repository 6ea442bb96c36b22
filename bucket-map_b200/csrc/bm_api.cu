// bm_api.cu -- context, parameters, reference/read upload and measurement entry points of libbucketmap_b200.so.
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <new>
#include <set>
#include <thread>
#include <vector>

#include <cub/device/device_scan.cuh>

#include <sys/mman.h>

#include "bm_internal.cuh"

extern "C" {

int bm_version(void) { return 200; }

int bm_ctx_create(int device, bm_ctx** out) {
    if (!out) return BM_E_ARG;
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0 || device < 0 || device >= n) return BM_E_CUDA;   // no CPU fallback
    if (cudaSetDevice(device) != cudaSuccess) return BM_E_CUDA;
    bm_ctx* c = new (std::nothrow) bm_ctx();
    if (!c) return BM_E_NOMEM;
    c->device = device;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) c->sm_count = prop.multiProcessorCount;
    if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreate(&c->ev_a) != cudaSuccess || cudaEventCreate(&c->ev_b) != cudaSuccess) {
        delete c;
        return BM_E_CUDA;
    }
    *out = c;
    return BM_OK;
}

void bm_ctx_destroy(bm_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    bm_gather_close(c);
    bm_comm_destroy(c);
    DevBuf* bufs[] = {&c->d_packed, &c->d_bstart, &c->d_blen, &c->d_rows, &c->d_k2i, &c->d_zeros, &c->d_distok,
                      &c->d_stage, &c->d_seq, &c->d_qual, &c->d_read_off, &c->d_seg_read, &c->d_seg_start,
                      &c->d_seg_len, &c->d_rowidx, &c->d_segok, &c->d_cand_fixed, &c->d_cand_cnt, &c->d_cand_off,
                      &c->d_cand, &c->d_stats, &c->d_scan_tmp, &c->d_lsamp_idx, &c->d_lsamp_hash, &c->d_cres_pos,
                      &c->d_cres_votes, &c->d_cseg, &c->d_loc_cnt, &c->d_loc_off, &c->d_locs, &c->d_locs2,
                      &c->d_loc_off2, &c->d_vote_scratch, &c->d_locs3, &c->d_loc_pool, &c->d_loc_flags, &c->d_read_seg0, &c->d_aln, &c->d_cigar_fixed, &c->d_cigar_cnt,
                      &c->d_cigar_off, &c->d_cigar, &c->d_aln_scratch, &c->d_flush, &c->d_krows, &c->d_distok_p, &c->d_gidx_pos,
                      &c->d_gidx_off, &c->d_loc_over, &c->d_map_over, &c->d_frows, &c->d_pl, &c->d_qk, &c->d_pl_off};
    if (c->lane_ev) cudaEventDestroy(c->lane_ev);
    if (c->lane) { bm_ctx_destroy(c->lane); c->lane = nullptr; cudaSetDevice(c->device); }
    for (DevBuf* b : bufs) b->release();
    c->h_lane_off.release();
    for (HostBuf* h : {&c->h_read_seg0, &c->h_seg_read, &c->h_seg_start, &c->h_seg_len, &c->h_off_rel, &c->h_pl_off}) h->release();
    for (auto e : c->chunk_ev) cudaEventDestroy(e);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    for (auto& p : c->pending) { cudaEventDestroy(p.a); cudaEventDestroy(p.b); }
    for (auto e : c->ev_pool) cudaEventDestroy(e);
    if (c->ev_a) cudaEventDestroy(c->ev_a);
    if (c->ev_b) cudaEventDestroy(c->ev_b);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

const char* bm_last_error(const bm_ctx* c) { return c ? c->err.c_str() : "null context / no CUDA device"; }

// main.cpp:12-39 defaults; main.cpp:207 and bucket_locator.h:419-420 thresholds, computed in float like the reference
void bm_params_derive(bm_params* p, uint32_t bucket_len, uint32_t read_len, uint32_t q, uint32_t k, uint32_t s,
                      float d, uint32_t b, float e, float n, uint32_t pl, uint32_t u) {
    p->bucket_len = bucket_len; p->read_len = read_len; p->q = q; p->k = k;
    p->mapper_samples = s;
    p->mapper_fault = (uint32_t)std::ceil(s * e);
    p->distinguishability = d; p->avg_base_quality = b;
    p->max_candidates = 30; p->num_segment_samples = 5;
    p->locator_samples = pl;
    p->allowed_mismatch = (uint32_t)(int)std::ceil(e * pl);
    p->allowed_indel = (uint32_t)(int)std::ceil(n * read_len);
    p->indel_rate = n; p->quality_threshold = u;
}
void bm_params_default(bm_params* p, uint32_t bucket_len) {
    bm_params_derive(p, bucket_len, 300, 9, 12, 15, 0.5f, 25, 0.4f, 0.02f, 10, 40);
}

int bm_locator_index(bm_ctx* c, int mode) {
    if (!c) return BM_E_ARG;
    if (mode != 0 && mode != 1) BM_FAIL(c, BM_E_ARG, "bm_locator_index: mode must be 0 (off) or 1 (when it fits)");
    c->gidx_mode = mode;
    if (!mode) { cudaSetDevice(c->device); c->d_gidx_pos.release(); c->d_gidx_off.release(); c->gidx_k = 0; }
    return BM_OK;
}

int bm_ref_upload(bm_ctx* c, const uint8_t* packed, uint64_t n_bases, const uint64_t* bstart, const uint32_t* blen,
                  uint32_t n_emit, uint32_t NB) {
    if (!c) return BM_E_ARG;
    if (!packed || (n_emit && (!bstart || !blen)) || n_emit > NB) BM_FAIL(c, BM_E_ARG, "bm_ref_upload: bad arguments");
    cudaSetDevice(c->device);
    uint64_t nbytes = (n_bases + 3) / 4;
    c->d_gidx_pos.release(); c->d_gidx_off.release(); c->gidx_k = 0;          // the position index belongs to the old text
    BM_CK(c, c->d_packed.ensure(nbytes + 64));
    BM_CK(c, cudaMemsetAsync((uint8_t*)c->d_packed.p + nbytes, 0, 64, c->stream));   // kernels read whole words
    { int rh = bm_internal_h2d(c, c->d_packed.p, packed, nbytes); if (rh) return rh; }
    BM_CK(c, c->d_bstart.ensure((size_t)(n_emit + 1) * 8));
    BM_CK(c, c->d_blen.ensure((size_t)(n_emit + 1) * 4));
    uint32_t mx = 0;
    for (uint32_t i = 0; i < n_emit; i++) {
        if (bstart[i] + blen[i] > n_bases) BM_FAIL(c, BM_E_ARG, "bm_ref_upload: bucket outside the genome");
        if (blen[i] > mx) mx = blen[i];
    }
    if (n_emit) {
        BM_CK(c, cudaMemcpyAsync(c->d_bstart.p, bstart, (size_t)n_emit * 8, cudaMemcpyHostToDevice, c->stream));
        BM_CK(c, cudaMemcpyAsync(c->d_blen.p, blen, (size_t)n_emit * 4, cudaMemcpyHostToDevice, c->stream));
    }
    BM_CK(c, cudaStreamSynchronize(c->stream));
    c->n_bases = n_bases; c->n_emit = n_emit; c->NB = NB; c->max_blen = mx;
    return BM_OK;
}

// utils.h:160-178 (Sampler).  The reference's early return (last_upper_bound is never updated, so it fires only
// for upper_bound==0) leaves stale samples and then reads out of range; here upper_bound==0 gives n zeros.
static void sample_det(uint32_t n, uint32_t ub, std::vector<uint32_t>& out) {
    out.clear();
    if (n == 0) return;
    if (ub == 0) { out.assign(n, 0u); return; }
    double delta = (n == 1) ? 0.0 : static_cast<double>(ub + 1u) / (double)(n - 1u);
    for (uint32_t i = 0; i + 1 < n; i++) out.push_back((uint32_t)std::floor((double)(int)i * delta));
    out.push_back(ub);
}

}  // extern "C"

// Sizes the host tables and the device buffers of a batch of n_reads reads (n_seg is bounded by
// n_reads * num_segment_samples) and resets the per-batch state; the tables are filled by bm_internal_build_segments.
// lazy (packed batches only): the pass over the reads is left to bm_internal_build_unit_range(scan = true), chunk by chunk,
// under the assumption that no read is longer than 2r; the caller starts over without `lazy` when a chunk says otherwise.
int bm_internal_reads_begin(bm_ctx* c, const bm_params* p, const uint64_t* read_off, uint32_t n_reads, const uint32_t* pl_off,
                            bool lazy) {
    if (p->read_len == 0 || p->num_segment_samples == 0) BM_FAIL(c, BM_E_ARG, "bm_reads_upload: bad parameters");
    c->have_reads = false; c->have_map = false; c->have_locs = false; c->have_aln = false;
    c->lsamp_k = c->lsamp_P = c->lsamp_b = 0;
    if (read_off[n_reads] < read_off[0]) BM_FAIL(c, BM_E_ARG, "bm_reads_upload: read_off not monotone");
    const uint64_t nbytes = read_off[n_reads] - read_off[0];
    // long reads (> 2r) are cut into num_segment_samples segments; without one the table has one entry per read
    // exact segment count: reads longer than 2r are cut into num_segment_samples segments (one cheap pass)
    uint64_t n_long = 0, n_chunks = 0, max_len = 0;
    const uint64_t two_r = 2ull * p->read_len;
    if (lazy && pl_off) n_chunks = (uint64_t)pl_off[n_reads] - pl_off[0];
    else for (uint32_t i = 0; i < n_reads; i++) {          // one pass: long reads, chunks of the packed form, longest read
        const uint64_t len = read_off[i + 1] - read_off[i];
        n_long += len > two_r;
        n_chunks += (len + 31) / 32;
        max_len = len > max_len ? len : max_len;
        if (pl_off && (uint64_t)(pl_off[i + 1] - pl_off[i]) != (len + 31) / 32)
            BM_FAIL(c, BM_E_ARG, "packed batch: pl_off is not the layout of bm_pack_layout");
    }
    if (max_len > 0x7FFFFFFFull) BM_FAIL(c, BM_E_LIMIT, "bm_reads_upload: read longer than 2^31 (or read_off not monotone)");
    c->batch_max_len = (uint32_t)max_len;
    const uint64_t seg_cap = n_reads + n_long * (p->num_segment_samples - 1);
    if (seg_cap > 0x7FFFFFFFull) BM_FAIL(c, BM_E_LIMIT, "bm_reads_upload: too many segments in one batch");
    if (n_chunks > 0xFFFFFFF0ull) BM_FAIL(c, BM_E_LIMIT, "bm_reads_upload: more than 2^32 chunks of 32 bases in one batch");
    c->n_pl_chunks = n_chunks;
    BM_CK(c, c->h_pl_off.ensure((size_t)(n_reads + 1) * 4));
    BM_CK(c, c->d_pl_off.ensure((size_t)(n_reads + 1) * 4));
    BM_CK(c, c->d_pl.ensure((size_t)(n_chunks + 2) * 8));
    BM_CK(c, c->d_qk.ensure((size_t)(n_chunks + 2) * 4));
    BM_CK(c, c->h_off_rel.ensure((size_t)(n_reads + 1) * 8));
    BM_CK(c, c->h_read_seg0.ensure((size_t)(n_reads + 1) * 4));
    BM_CK(c, c->h_seg_read.ensure((size_t)(seg_cap + 1) * 4));
    BM_CK(c, c->h_seg_start.ensure((size_t)(seg_cap + 1) * 4));
    BM_CK(c, c->h_seg_len.ensure((size_t)(seg_cap + 1) * 4));
    BM_CK(c, c->d_read_seg0.ensure((size_t)(n_reads + 1) * 4));
    if (c->want_ascii) {
        BM_CK(c, c->d_seq.ensure(nbytes + 16));
        BM_CK(c, c->d_qual.ensure(nbytes + 16));
    }
    BM_CK(c, c->d_read_off.ensure((size_t)(n_reads + 1) * 8));
    BM_CK(c, c->d_seg_read.ensure((size_t)(seg_cap + 1) * 4));
    BM_CK(c, c->d_seg_start.ensure((size_t)(seg_cap + 1) * 4));
    BM_CK(c, c->d_seg_len.ensure((size_t)(seg_cap + 1) * 4));
    c->n_reads = n_reads; c->n_seg = 0; c->n_seg_total = (uint32_t)seg_cap; c->max_seg_len = 0; c->max_read_len = 0;
    c->n_read_bytes = nbytes;
    c->h_seg_valid = true;
    return BM_OK;
}

// Reads [r0, r1) of a batch WITHOUT long reads (one segment per read, segment index == read index): only the offset
// tables are made on the host -- copied from the caller's arrays where those can be used as they are -- and the segment
// tables are written on the device (k_unit_segments).  pl_off: the caller's chunk offsets (packed batches) or nullptr.
// scan: the checks bm_internal_reads_begin(lazy) left out, for these reads; returns 1 when one of them is longer than 2r.
int bm_internal_build_unit_range(bm_ctx* c, const bm_params* p, const uint64_t* read_off, const uint32_t* pl_off,
                                 uint32_t r0, uint32_t r1, bool scan) {
    if (scan) {
        const uint64_t two_r = 2ull * p->read_len;
        uint64_t max_len = c->batch_max_len, bad = 0;
        for (uint32_t i = r0; i < r1; i++) {
            const uint64_t len = read_off[i + 1] - read_off[i];          // a decreasing offset wraps to a huge length
            max_len = len > max_len ? len : max_len;
            bad |= (uint64_t)(pl_off[i + 1] - pl_off[i]) ^ ((len + 31) / 32);
        }
        if (max_len > two_r) {
            if (max_len > 0x7FFFFFFFull) BM_FAIL(c, BM_E_LIMIT, "bm_reads_upload: read longer than 2^31 (or read_off not monotone)");
            return 1;
        }
        if (bad) BM_FAIL(c, BM_E_ARG, "packed batch: pl_off is not the layout of bm_pack_layout");
        c->batch_max_len = (uint32_t)max_len;
    }
    uint64_t* off_rel = c->h_off_rel.as<uint64_t>();
    uint32_t* hpl = c->h_pl_off.as<uint32_t>();
    const uint64_t base = read_off[0];
    if (base == 0) memcpy(off_rel + r0, read_off + r0, (size_t)(r1 - r0 + 1) * 8);
    else for (uint32_t i = r0; i <= r1; i++) off_rel[i] = read_off[i] - base;
    if (pl_off) {
        const uint32_t pb = pl_off[0];
        if (pb == 0) memcpy(hpl + r0, pl_off + r0, (size_t)(r1 - r0 + 1) * 4);
        else for (uint32_t i = r0; i <= r1; i++) hpl[i] = pl_off[i] - pb;
    } else {
        uint32_t plo = r0 ? hpl[r0] : 0u;
        for (uint32_t i = r0; i < r1; i++) { hpl[i] = plo; plo += (uint32_t)((read_off[i + 1] - read_off[i] + 31) / 32); }
        hpl[r1] = plo;
    }
    c->n_seg = r1;
    c->max_read_len = c->batch_max_len;
    c->max_seg_len = c->batch_max_len < p->read_len ? c->batch_max_len : p->read_len;
    c->h_seg_valid = false;                      // the host copies of the segment tables were not made
    return BM_OK;
}

// segment table of reads [r0, r1): q_gram_mapper.h:509-523 / bucket_locator.h:302-313.  Reads must be appended in order.
int bm_internal_build_segments(bm_ctx* c, const bm_params* p, const uint64_t* read_off, uint32_t r0, uint32_t r1) {
    uint64_t* off_rel = c->h_off_rel.as<uint64_t>();
    uint32_t* seg0 = c->h_read_seg0.as<uint32_t>();
    uint32_t *sr = c->h_seg_read.as<uint32_t>(), *ss = c->h_seg_start.as<uint32_t>(), *sl = c->h_seg_len.as<uint32_t>();
    const uint64_t base = read_off[0];
    const uint32_t rl = p->read_len;
    uint32_t n_seg = c->n_seg, mx = c->max_seg_len, mxr = c->max_read_len;
    uint32_t* pl_off = c->h_pl_off.as<uint32_t>();
    uint32_t plo = r0 ? pl_off[r0] : 0u;
    std::vector<uint32_t> st;
    for (uint32_t i = r0; i < r1; i++) {
        if (read_off[i + 1] < read_off[i]) BM_FAIL(c, BM_E_ARG, "bm_reads_upload: read_off not monotone");
        off_rel[i] = read_off[i] - base;
        const uint64_t len64 = read_off[i + 1] - read_off[i];
        pl_off[i] = plo;
        plo += (uint32_t)((len64 + 31) / 32);
        if (len64 > 0x7FFFFFFFull) BM_FAIL(c, BM_E_LIMIT, "bm_reads_upload: read longer than 2^31");
        const uint32_t len = (uint32_t)len64;
        if (len > mxr) mxr = len;
        seg0[i] = n_seg;
        if (len > 2 * rl) {
            sample_det(p->num_segment_samples, len - rl - 1, st);
            for (uint32_t s : st) {
                const uint32_t e = s + rl < len ? s + rl : len;
                sr[n_seg] = i; ss[n_seg] = s; sl[n_seg] = e - s; n_seg++;
                if (e - s > mx) mx = e - s;
            }
        } else {
            const uint32_t e = rl < len ? rl : len;
            sr[n_seg] = i; ss[n_seg] = 0; sl[n_seg] = e; n_seg++;
            if (e > mx) mx = e;
        }
    }
    off_rel[r1] = read_off[r1] - base;
    pl_off[r1] = plo;
    seg0[r1] = n_seg;
    c->n_seg = n_seg; c->max_seg_len = mx; c->max_read_len = mxr;
    return BM_OK;
}

extern "C" {

static int upload_tables(bm_ctx* c, uint32_t n_reads) {
    const uint32_t n_seg = c->n_seg;
    BM_CK(c, cudaMemcpyAsync(c->d_read_seg0.p, c->h_read_seg0.p, (size_t)(n_reads + 1) * 4, cudaMemcpyHostToDevice, c->stream));
    BM_CK(c, cudaMemcpyAsync(c->d_read_off.p, c->h_off_rel.p, (size_t)(n_reads + 1) * 8, cudaMemcpyHostToDevice, c->stream));
    BM_CK(c, cudaMemcpyAsync(c->d_pl_off.p, c->h_pl_off.p, (size_t)(n_reads + 1) * 4, cudaMemcpyHostToDevice, c->stream));
    if (n_seg) {
        BM_CK(c, cudaMemcpyAsync(c->d_seg_read.p, c->h_seg_read.p, (size_t)n_seg * 4, cudaMemcpyHostToDevice, c->stream));
        BM_CK(c, cudaMemcpyAsync(c->d_seg_start.p, c->h_seg_start.p, (size_t)n_seg * 4, cudaMemcpyHostToDevice, c->stream));
        BM_CK(c, cudaMemcpyAsync(c->d_seg_len.p, c->h_seg_len.p, (size_t)n_seg * 4, cudaMemcpyHostToDevice, c->stream));
    }
    return BM_OK;
}

int bm_reads_upload(bm_ctx* c, const bm_params* p, const uint8_t* seq, const uint8_t* qual, const uint64_t* read_off,
                    uint32_t n_reads) {
    if (!c) return BM_E_ARG;
    if (!p || !read_off || (n_reads && (!seq || !qual))) BM_FAIL(c, BM_E_ARG, "bm_reads_upload: bad arguments");
    cudaSetDevice(c->device);
    // read i = seq[read_off[i], read_off[i+1]); read_off[0] may be non-zero (a slice of a larger batch)
    c->want_ascii = true;
    int r = bm_internal_reads_begin(c, p, read_off, n_reads, nullptr);
    if (r) return r;
    r = bm_internal_build_segments(c, p, read_off, 0, n_reads);
    if (r) return r;
    const uint64_t base = read_off[0], nbytes = c->n_read_bytes;
    if (nbytes) {
        BM_CK(c, cudaMemcpyAsync(c->d_seq.p, seq + base, nbytes, cudaMemcpyHostToDevice, c->stream));
        BM_CK(c, cudaMemcpyAsync(c->d_qual.p, qual + base, nbytes, cudaMemcpyHostToDevice, c->stream));
    }
    r = upload_tables(c, n_reads);
    if (r) return r;
    r = bm_internal_pack_range(c, p, 0, n_reads);              // ASCII -> bit planes + k-mer quality plane
    if (r) return r;
    BM_CK(c, cudaStreamSynchronize(c->stream));   // the caller's buffers may be reused after return
    c->ascii_resident = true; c->have_qk = true; c->pack_k = p->k; c->pack_b = p->avg_base_quality;
    c->have_reads = true;
    return BM_OK;
}

// the batch already in the device form (bm_pack_layout / bm_pack_reads made it on the host)
int bm_reads_upload_packed(bm_ctx* c, const bm_params* p, const uint64_t* planes, const uint32_t* qk,
                           const uint32_t* pl_off, const uint64_t* read_off, uint32_t n_reads) {
    if (!c) return BM_E_ARG;
    if (!p || !read_off || !pl_off || (n_reads && !planes)) BM_FAIL(c, BM_E_ARG, "bm_reads_upload_packed: bad arguments");
    cudaSetDevice(c->device);
    c->want_ascii = false;
    int r = bm_internal_reads_begin(c, p, read_off, n_reads, pl_off);
    if (r) return r;
    r = bm_internal_build_segments(c, p, read_off, 0, n_reads);
    if (r) return r;
    const uint64_t nch = c->n_pl_chunks;
    if (nch) {
        BM_CK(c, cudaMemcpyAsync(c->d_pl.p, planes + pl_off[0], nch * 8, cudaMemcpyHostToDevice, c->stream));
        if (qk) BM_CK(c, cudaMemcpyAsync(c->d_qk.p, qk + pl_off[0], nch * 4, cudaMemcpyHostToDevice, c->stream));
    }
    r = upload_tables(c, n_reads);
    if (r) return r;
    BM_CK(c, cudaStreamSynchronize(c->stream));
    c->ascii_resident = false; c->have_qk = qk != nullptr; c->pack_k = p->k; c->pack_b = p->avg_base_quality;
    c->have_reads = true;
    return BM_OK;
}

int bm_reads_segments(bm_ctx* c, uint32_t* n_seg, uint32_t* seg_read, uint32_t* seg_start, uint32_t* seg_len) {
    if (!c) return BM_E_ARG;
    if (!c->have_reads) BM_FAIL(c, BM_E_ARG, "bm_reads_segments: no read batch resident");
    if (n_seg) *n_seg = c->n_seg;
    size_t b = (size_t)c->n_seg * 4;
    if (!c->h_seg_valid && b && (seg_read || seg_start || seg_len)) {     // tables were written on the device: fetch them
        cudaSetDevice(c->device);
        BM_CK(c, cudaMemcpyAsync(c->h_seg_read.p, c->d_seg_read.p, b, cudaMemcpyDeviceToHost, c->stream));
        BM_CK(c, cudaMemcpyAsync(c->h_seg_start.p, c->d_seg_start.p, b, cudaMemcpyDeviceToHost, c->stream));
        BM_CK(c, cudaMemcpyAsync(c->h_seg_len.p, c->d_seg_len.p, b, cudaMemcpyDeviceToHost, c->stream));
        BM_CK(c, cudaStreamSynchronize(c->stream));
        c->h_seg_valid = true;
    }
    if (seg_read && b) memcpy(seg_read, c->h_seg_read.p, b);
    if (seg_start && b) memcpy(seg_start, c->h_seg_start.p, b);
    if (seg_len && b) memcpy(seg_len, c->h_seg_len.p, b);
    return BM_OK;
}

// ---- measurement ------------------------------------------------------------------------------------------------
int bm_timer_start(bm_ctx* c) {
    if (!c) return BM_E_ARG;
    cudaSetDevice(c->device);
    BM_CK(c, cudaEventRecord(c->ev_a, c->stream));
    return BM_OK;
}
int bm_timer_stop(bm_ctx* c, float* ms) {
    if (!c || !ms) return BM_E_ARG;
    cudaSetDevice(c->device);
    BM_CK(c, cudaEventRecord(c->ev_b, c->stream));
    BM_CK(c, cudaEventSynchronize(c->ev_b));
    BM_CK(c, cudaEventElapsedTime(ms, c->ev_a, c->ev_b));
    return BM_OK;
}
int bm_kernel_timing(bm_ctx* c, int enable) {
    if (!c) return BM_E_ARG;
    c->ktiming = enable != 0;
    return BM_OK;
}
static int drain_pending(bm_ctx* c) {
    if (c->pending.empty()) return BM_OK;
    BM_CK(c, cudaStreamSynchronize(c->stream));
    for (auto& p : c->pending) {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, p.a, p.b) == cudaSuccess) c->kstat[p.kid].ms += ms;
        c->ev_pool.push_back(p.a); c->ev_pool.push_back(p.b);
    }
    c->pending.clear();
    return BM_OK;
}
int bm_kernel_stats(bm_ctx* c, int kid, double* total_ms, uint64_t* launches) {
    if (!c || kid < 0 || kid >= BM_K_COUNT) return BM_E_ARG;
    cudaSetDevice(c->device);
    int r = drain_pending(c);
    if (r) return r;
    if (total_ms) *total_ms = c->kstat[kid].ms;
    if (launches) *launches = c->kstat[kid].launches;
    return BM_OK;
}
int bm_kernel_stats_reset(bm_ctx* c) {
    if (!c) return BM_E_ARG;
    cudaSetDevice(c->device);
    int r = drain_pending(c);
    if (r) return r;
    for (auto& k : c->kstat) k = KStat();
    return BM_OK;
}
int bm_flush_l2(bm_ctx* c, uint64_t bytes) {
    if (!c) return BM_E_ARG;
    cudaSetDevice(c->device);
    BM_CK(c, c->d_flush.ensure(bytes));
    BM_CK(c, cudaMemsetAsync(c->d_flush.p, 0x5a, bytes, c->stream));
    return BM_OK;
}
int bm_sync(bm_ctx* c) {
    if (!c) return BM_E_ARG;
    cudaSetDevice(c->device);
    BM_CK(c, cudaStreamSynchronize(c->stream));
    return BM_OK;
}
// Pinned memory for the caller's batches.  Large blocks are 2 MB-aligned anonymous memory with MADV_HUGEPAGE, touched by the
// caller's thread and then registered with the driver: on the measured host cudaMallocHost costs 0.4 ms per MB, all of it
// under the driver's lock (a parser thread getting its next batch buffer stalls the device thread's copies for that long),
// huge-page faults 0.12 ms per MB outside the lock and the registration 0.03 ms per MB.  On a machine without a usable CUDA
// device (the host-side unit tests) the memory stays unregistered; bm_host_free tells the kinds apart.
static std::mutex g_plain_m;
static std::map<void*, std::pair<size_t, bool>> g_plain;       // block -> (bytes, registered)

void* bm_host_alloc(uint64_t bytes) {
    void* p = nullptr;
    constexpr size_t kHuge = (size_t)2 << 20;
    if (bytes >= 2 * kHuge) {
        const size_t n = ((size_t)bytes + kHuge - 1) & ~(kHuge - 1);
        if (posix_memalign(&p, kHuge, n) == 0) {
            static const bool thp = !getenv("BM_HOST_THP") || atoi(getenv("BM_HOST_THP")) != 0;
            if (thp) madvise(p, n, MADV_HUGEPAGE);
            for (size_t i = 0; i < n; i += 4096) static_cast<volatile char*>(p)[i] = 0;
            const bool reg = cudaHostRegister(p, n, cudaHostRegisterPortable) == cudaSuccess;
            if (!reg) cudaGetLastError();
            std::lock_guard<std::mutex> g(g_plain_m);
            g_plain[p] = {n, reg};
            return p;
        }
        p = nullptr;
    }
    if (cudaMallocHost(&p, bytes ? bytes : 1) == cudaSuccess) return p;
    cudaGetLastError();
    if (posix_memalign(&p, 4096, bytes ? bytes : 1) != 0) return nullptr;
    std::lock_guard<std::mutex> g(g_plain_m);
    g_plain[p] = {(size_t)bytes, false};
    return p;
}
void bm_host_free(void* p) {
    if (!p) return;
    {
        std::unique_lock<std::mutex> g(g_plain_m);
        auto it = g_plain.find(p);
        if (it != g_plain.end()) {
            const bool reg = it->second.second;
            g_plain.erase(it);
            g.unlock();
            if (reg) cudaHostUnregister(p);
            free(p);
            return;
        }
    }
    cudaFreeHost(p);
}

}  // extern "C"

// Large host->device copy from memory that may be pageable (a memory-mapped index file, the FASTA packer's array).  One
// cudaMemcpyAsync: the driver stages it through its own pinned buffers.  (Tried: four host threads copying 16 MB pieces into
// their own pinned slots and handing each to the copy engine -- 0.24 s instead of 0.12 s for the 850 MB .qgram image of C2,
// the slots' cudaMallocHost and the page faults of the mapping cost more than the staging saves; not kept.)
int bm_internal_h2d(bm_ctx* c, void* dst, const void* src, size_t bytes) {
    BM_CK(c, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, c->stream));
    return BM_OK;
}

int bm_internal_exclusive_scan(bm_ctx* c, const uint32_t* d_in, uint32_t* d_out, uint32_t n) {
    size_t tmp = 0;
    BM_CK(c, cub::DeviceScan::ExclusiveSum(nullptr, tmp, d_in, d_out, (int)n, c->stream));
    BM_CK(c, c->d_scan_tmp.ensure(tmp));
    BM_CK(c, cub::DeviceScan::ExclusiveSum(c->d_scan_tmp.p, tmp, d_in, d_out, (int)n, c->stream));
    return BM_OK;
}
