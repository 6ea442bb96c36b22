// bm_pipeline.cu -- bm_map_locate: the host-buffer entry point of the alignment-free path (what bucket_locator::locate
// does per FASTQ batch), software-pipelined over two lanes.
//
// A batch of n reads is cut into chunks; two lanes (child contexts: own stream, own read/result buffers, the parent's
// index and reference viewed, not copied) take alternate chunks, each driven by its own host thread:
//     H2D(chunk) -> K2a -> K2 -> K3a -> K3 -> gather -> K3b -> D2H(records straight into the caller's arrays)
// so the copies and the host-side segment table of one chunk overlap the kernels of the other, and the latency-bound
// locator kernels of one lane fill the gaps of the HBM-bound vote kernel of the other.  Results are identical to the
// serial path for any chunking (reads are independent; tests/test_gpu_fullsize.py checks sharding invariance).
#include <condition_variable>
#include <mutex>
#include <thread>

#include "bm_internal.cuh"

namespace {

constexpr uint32_t kPipeMinReads = 32768;      // below this the serial path is used
constexpr uint32_t kChunkMin = 16384, kChunkMax = 131072;

// (re)point a lane at the parent's index + reference
int lane_sync(bm_ctx* parent, bm_ctx*& lane) {
    if (!lane) {
        int r = bm_ctx_create(parent->device, &lane);
        if (r) return r;
    }
    bm_ctx* l = lane;
    l->d_packed.view_of(parent->d_packed); l->d_bstart.view_of(parent->d_bstart); l->d_blen.view_of(parent->d_blen);
    l->n_bases = parent->n_bases; l->n_emit = parent->n_emit; l->NB = parent->NB; l->max_blen = parent->max_blen;
    l->d_rows.view_of(parent->d_rows); l->d_k2i.view_of(parent->d_k2i); l->d_zeros.view_of(parent->d_zeros);
    l->d_distok.view_of(parent->d_distok);
    l->q = parent->q; l->n_rows = parent->n_rows; l->W = parent->W; l->Wp = parent->Wp;
    l->have_index = parent->have_index; l->zeros_for_d = parent->zeros_for_d;
    l->dist_threshold = parent->dist_threshold; l->dist_valid = parent->dist_valid;
    l->ktiming = parent->ktiming;
    l->have_reads = l->have_map = l->have_locs = l->have_aln = false;
    return BM_OK;
}

struct Shared {
    std::mutex m;
    std::condition_variable cv;
    std::vector<int64_t> count;     // locations of chunk j, -1 until known
    int err = 0;
    std::string err_text;
};

}  // namespace

extern "C" int bm_map_locate(bm_ctx* c, const bm_params* p, const uint8_t* seq, const uint8_t* qual,
                             const uint64_t* read_off, uint32_t n_reads, int filter_best, uint32_t* loc_off, bm_loc* locs,
                             uint64_t loc_cap, uint64_t* stats) {
    if (!c) return BM_E_ARG;
    if (!p || !read_off) BM_FAIL(c, BM_E_ARG, "bm_map_locate: bad arguments");
    if (n_reads < kPipeMinReads || !loc_off)
        return bm_internal_map_locate_serial(c, p, seq, qual, read_off, n_reads, filter_best, loc_off, locs, loc_cap, stats);
    if (!c->have_index) BM_FAIL(c, BM_E_ARG, "bm_map_locate: no index resident (the q-gram index is empty)");
    if (!c->d_packed.p) BM_FAIL(c, BM_E_ARG, "bm_map_locate: bm_ref_upload first");
    cudaSetDevice(c->device);
    int r = bm_internal_index_prepare(c, p->distinguishability);      // K0 once, shared by both lanes
    if (r) return r;
    BM_CK(c, cudaStreamSynchronize(c->stream));
    for (auto& l : c->lane) {
        r = lane_sync(c, l);
        if (r) BM_FAIL(c, r, "bm_map_locate: cannot create a pipeline lane");
    }
    c->have_reads = c->have_map = c->have_locs = c->have_aln = false;

    uint32_t chunk = n_reads / 8;
    chunk = chunk < kChunkMin ? kChunkMin : (chunk > kChunkMax ? kChunkMax : chunk);
    const uint32_t n_chunks = (n_reads + chunk - 1) / chunk;
    Shared sh;
    sh.count.assign(n_chunks, -1);
    uint64_t lane_stats[2][8] = {{0}, {0}};

    auto fail = [&](int code, const char* text) {
        std::lock_guard<std::mutex> g(sh.m);
        if (!sh.err) { sh.err = code; sh.err_text = text ? text : ""; }
        sh.cv.notify_all();
    };
    auto worker = [&](int t) {
        bm_ctx* l = c->lane[t];
        cudaSetDevice(l->device);
        for (uint32_t j = (uint32_t)t; j < n_chunks; j += 2) {
            { std::lock_guard<std::mutex> g(sh.m); if (sh.err) return; }
            const uint32_t r0 = j * chunk, r1 = (r0 + chunk < n_reads) ? r0 + chunk : n_reads, nj = r1 - r0;
            int rc = bm_reads_upload(l, p, seq, qual, read_off + r0, nj);
            if (!rc) rc = bm_map_device(l, p);
            if (!rc) rc = bm_locate_device(l, p, filter_best);
            uint64_t nl = 0;
            if (!rc) rc = bm_locate_count(l, &nl);
            if (rc) { fail(rc, bm_last_error(l)); return; }
            // base of this chunk in the caller's arrays = locations of all earlier chunks
            uint64_t base = 0;
            {
                std::unique_lock<std::mutex> g(sh.m);
                sh.count[j] = (int64_t)nl;
                sh.cv.notify_all();
                sh.cv.wait(g, [&] {
                    if (sh.err) return true;
                    for (uint32_t i = 0; i < j; i++) if (sh.count[i] < 0) return false;
                    return true;
                });
                if (sh.err) return;
                for (uint32_t i = 0; i < j; i++) base += (uint64_t)sh.count[i];
            }
            if (locs && base + nl > loc_cap) { fail(BM_E_ARG, "bm_map_locate: location buffer too small"); return; }
            if (base + nl > 0xFFFFFFFFull) { fail(BM_E_LIMIT, "bm_map_locate: more than 2^32 locations in one batch"); return; }
            const void* d_off = l->locs_filtered ? l->d_loc_off2.p : l->d_loc_off.p;
            const void* d_loc = l->locs_filtered ? l->d_locs3.p : l->d_locs.p;
            cudaError_t e = cudaMemcpyAsync(loc_off + r0, d_off, (size_t)nj * 4, cudaMemcpyDeviceToHost, l->stream);
            if (e == cudaSuccess && locs && nl)
                e = cudaMemcpyAsync(locs + base, d_loc, (size_t)nl * sizeof(bm_loc), cudaMemcpyDeviceToHost, l->stream);
            uint64_t st[8] = {0};
            if (e == cudaSuccess) e = cudaMemcpyAsync(st, l->d_stats.p, 8 * 8, cudaMemcpyDeviceToHost, l->stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(l->stream);
            if (e != cudaSuccess) { fail(BM_E_CUDA, cudaGetErrorString(e)); return; }
            if (base) for (uint32_t i = 0; i < nj; i++) loc_off[r0 + i] += (uint32_t)base;
            for (int i = 0; i < 8; i++) lane_stats[t][i] += st[i];
        }
    };
    std::thread other(worker, 1);
    worker(0);
    other.join();
    if (sh.err) { c->err = sh.err_text; return sh.err; }
    uint64_t total = 0;
    for (auto v : sh.count) total += (uint64_t)v;
    loc_off[n_reads] = (uint32_t)total;
    if (stats) for (int i = 0; i < 8; i++) stats[i] = lane_stats[0][i] + lane_stats[1][i];
    // per-kernel accounting of the lanes is reported through the parent
    for (auto l : c->lane) {
        for (int k = 0; k < BM_K_COUNT; k++) {
            double ms = 0; uint64_t n = 0;
            bm_kernel_stats(l, k, &ms, &n);
            c->kstat[k].ms += ms; c->kstat[k].launches += n;
        }
        bm_kernel_stats_reset(l);
    }
    return BM_OK;
}
