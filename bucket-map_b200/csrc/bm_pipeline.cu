// bm_pipeline.cu -- bm_map_locate: the host-buffer entry point of the alignment-free path (what bucket_locator::locate
// does per FASTQ batch), with the host->device copy of the batch pipelined against the mapper kernels.
//
// The batch is cut into chunks of reads.  For each chunk the host appends its slice of the segment table (pinned
// staging), the copy stream moves the chunk's bases, qualities, offsets and table slice to the device and records an
// event; the compute stream waits for that event and runs K2a + K2 on the chunk's segment range.  The mapper (HBM-bound,
// ~3/4 of the device time) therefore starts after the first chunk has landed and hides the rest of the 600 B/read
// transfer.  The locator runs ONCE over the whole batch afterwards: its cost is per non-empty bucket (it streams each
// bucket's text once for all candidates in it), so it must not be split.  Results do not depend on the chunking.
//
// Opt-in (bm_pipeline_lanes): a batch is cut in two halves that run through this pipeline on two LANES at once: the
// context itself and a child context whose index buffers (bit matrix, coarse k-mer table, genome k-mer index, reference) are views of the
// parent's, each with its own streams and batch buffers, the child driven by a second host thread.  The host->device
// copies of both lanes share the link (the call is PCIe-bound), and the tail of a lane that cannot run under its own
// copy -- compaction, locator, merge, result copy -- runs under the other lane's.  The second lane writes its records
// behind the first lane's once that count is known.  Measured on C2 (1 M x 300 bp): 60.9 M reads/s against 63.8 M on one
// lane -- the call is bound by the 608 MB host->device copy (13 ms at the 47 GB/s the chunked copies reach, 55 GB/s
// link), and what the second lane hides of the 3.5 ms tail it loses again starting its copies late; hence off by default.
#include <atomic>
#include <chrono>
#include <cstdlib>
#include <thread>

#include "bm_internal.cuh"

namespace {
constexpr uint32_t kPipeMinReads = 32768;      // below this the serial path is used
constexpr uint32_t kChunkMin = 16384;

// what the second lane needs from the first: its record count (-1: not known yet, -2: the first lane failed)
struct LaneLink {
    std::atomic<long long> base{-1};
    // the link is shared: lane 1's copies are ordered behind lane 0's last one, so that lane 0 has the whole link first
    // and its tail then runs under lane 1's copies (0: not recorded yet, 1: h2d_done is recorded, -1: lane 0 failed)
    std::atomic<int> h2d_state{0};
    cudaEvent_t h2d_done = nullptr;
};

// the batch on the host: ASCII (seq, qual) or packed (planes, qk or nullptr, pl_off)
struct HostReads {
    const uint8_t* seq = nullptr; const uint8_t* qual = nullptr;
    const uint64_t* planes = nullptr; const uint32_t* qk = nullptr; const uint32_t* pl_off = nullptr;
    bool packed() const { return planes != nullptr; }
};

int pipeline_run(bm_ctx* c, const bm_params* p, const HostReads& src, const uint64_t* read_off,
                 uint32_t n_reads, int filter_best, uint32_t* loc_off, bm_loc* locs, uint64_t loc_cap, uint64_t* stats,
                 LaneLink* publish, LaneLink* wait_for, uint64_t* n_out, bool allow_lazy = true);

// the child lane sees the parent's index, tables and reference through views
void lane_share_index(bm_ctx* l, bm_ctx* c) {
    l->d_packed.view_of(c->d_packed); l->d_bstart.view_of(c->d_bstart); l->d_blen.view_of(c->d_blen);
    l->n_bases = c->n_bases; l->n_emit = c->n_emit; l->NB = c->NB; l->max_blen = c->max_blen;
    l->d_rows.view_of(c->d_rows); l->d_k2i.view_of(c->d_k2i); l->d_zeros.view_of(c->d_zeros);
    l->d_distok.view_of(c->d_distok); l->d_distok_p.view_of(c->d_distok_p);
    l->q = c->q; l->n_rows = c->n_rows; l->W = c->W; l->Wp = c->Wp; l->have_index = c->have_index;
    l->zeros_for_d = c->zeros_for_d; l->dist_threshold = c->dist_threshold; l->dist_valid = c->dist_valid;
    l->d_krows.view_of(c->d_krows); l->krows_k = c->krows_k; l->Wc = c->Wc; l->Wcs = c->Wcs;
    l->ktable_logg = c->krows_logg; l->krows_logg = c->krows_logg; l->ktable_auto = false;   // the lane takes the parent's table as it is
    l->d_frows.view_of(c->d_frows); l->frows_k = c->frows_k; l->fine_tried_k = c->fine_tried_k;
    l->ftable_mode = (c->d_frows.p && c->frows_k) ? 1 : 0;
    l->ktable_mode = c->d_krows.p ? 1 : 0;          // never build in the lane: use the parent's table or none
    l->ktable_max_bytes = c->ktable_max_bytes;
    l->d_gidx_pos.view_of(c->d_gidx_pos); l->d_gidx_off.view_of(c->d_gidx_off); l->gidx_k = c->gidx_k; l->gidx_lb = c->gidx_lb; l->gidx_bshift = c->gidx_bshift;
    l->gidx_mode = (c->d_gidx_pos.p && c->d_gidx_off.p) ? 1 : 0;
    l->ktiming = c->ktiming;
}
}  // namespace

namespace {
// segment tables of reads [r0, r1) when every read is one segment (q_gram_mapper.h:517-522 with |read| <= 2r)
__global__ void k_unit_segments(const uint64_t* __restrict__ read_off, uint32_t r0, uint32_t r1, uint32_t read_len,
                                uint32_t* __restrict__ seg_read, uint32_t* __restrict__ seg_start,
                                uint32_t* __restrict__ seg_len, uint32_t* __restrict__ read_seg0) {
    const uint32_t i = r0 + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= r1) return;
    const uint64_t len = read_off[i + 1] - read_off[i];
    seg_read[i] = i; seg_start[i] = 0u; seg_len[i] = len < read_len ? (uint32_t)len : read_len;
    read_seg0[i] = i;
    if (i + 1 == r1) read_seg0[r1] = r1;
}

int pipeline_run(bm_ctx* c, const bm_params* p, const HostReads& src, const uint64_t* read_off,
                 uint32_t n_reads, int filter_best, uint32_t* loc_off, bm_loc* locs, uint64_t loc_cap, uint64_t* stats,
                 LaneLink* publish, LaneLink* wait_for, uint64_t* n_out, bool allow_lazy) {
    cudaSetDevice(c->device);
    static const bool trace = getenv("BM_PIPE_TRACE") != nullptr;
    const auto t_begin = std::chrono::steady_clock::now();
    auto since = [&] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_begin).count(); };
    double t_pre = 0, t_first = 0, t_enq = 0, t_h2d = 0, t_fin = 0, t_loc = 0;
    if (!c->copy_stream) BM_CK(c, cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));

    // Chunk schedule: a small first chunk so that the mapper starts as early as possible, growing chunks while the link
    // and the mapper run side by side (every chunk costs a handful of small table copies and a launch), and a small last one
    // so that little mapper work is left when the last byte has landed: 1/16, 1/8, 3/16, 1/4, 1/4, 1/8 of the batch.
    std::vector<uint32_t> bounds(1, 0u);
    {
        static const uint32_t kSixteenths[] = {1, 2, 3, 4, 4, 2};
        uint32_t done16 = 0;
        for (uint32_t f : kSixteenths) {
            done16 += f;
            uint32_t b = (uint32_t)((uint64_t)n_reads * done16 / 16);
            if (done16 == 16) b = n_reads;
            if (b - bounds.back() >= kChunkMin || b == n_reads) { if (b > bounds.back()) bounds.push_back(b); }
        }
        if (bounds.back() != n_reads) bounds.push_back(n_reads);
    }
    const uint32_t n_chunks = (uint32_t)bounds.size() - 1;
    while (c->chunk_ev.size() < n_chunks) {
        cudaEvent_t e;
        BM_CK(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        c->chunk_ev.push_back(e);
    }
    // The link is the bottleneck: the bases and qualities of a chunk need nothing from the host-side tables, so their
    // copies are enqueued first (chunk 0's before anything else is prepared) and the segment tables of the chunk are
    // built while they are in flight.
    const uint64_t base = read_off[0];
    if (read_off[n_reads] < base) BM_FAIL(c, BM_E_ARG, "bm_reads_upload: read_off not monotone");
    const bool packed = src.packed();
    c->want_ascii = !packed;
    if (packed) {
        if (src.pl_off[n_reads] < src.pl_off[0]) BM_FAIL(c, BM_E_ARG, "bm_map_locate_packed: pl_off not monotone");
        const uint64_t nch = src.pl_off[n_reads] - src.pl_off[0];
        BM_CK(c, c->d_pl.ensure((nch + 2) * 8));
        BM_CK(c, c->d_qk.ensure((nch + 2) * 4));
    } else {
        const uint64_t nb = read_off[n_reads] - base;
        BM_CK(c, c->d_seq.ensure(nb + 16));
        BM_CK(c, c->d_qual.ensure(nb + 16));
        BM_CK(c, c->d_pl.ensure((nb / 32 + n_reads + 2) * 8));      // upper bound of sum ceil(len/32)
        BM_CK(c, c->d_qk.ensure((nb / 32 + n_reads + 2) * 4));
    }
    // the copy stream must not overwrite buffers an earlier call's kernels may still read
    BM_CK(c, cudaStreamSynchronize(c->stream));
    if (wait_for) {
        int st;
        while ((st = wait_for->h2d_state.load(std::memory_order_acquire)) == 0) std::this_thread::yield();
        if (st < 0) BM_FAIL(c, BM_E_ARG, "bm_map_locate: the first lane failed");
        BM_CK(c, cudaStreamWaitEvent(c->copy_stream, wait_for->h2d_done, 0));
    }
    auto copy_text = [&](uint32_t j) -> int {
        const uint32_t r0 = bounds[j], r1 = bounds[j + 1];
        if (read_off[r0] < base || read_off[r1] < read_off[r0] || read_off[r1] > read_off[n_reads])
            BM_FAIL(c, BM_E_ARG, "bm_reads_upload: read_off not monotone");
        const uint64_t b0 = read_off[r0] - base, nb = read_off[r1] - read_off[r0];
        if (packed) {
            if (src.pl_off[r0] < src.pl_off[0] || src.pl_off[r1] < src.pl_off[r0])
                BM_FAIL(c, BM_E_ARG, "bm_map_locate_packed: pl_off not monotone");
            const uint64_t c0 = src.pl_off[r0] - src.pl_off[0], nc = src.pl_off[r1] - src.pl_off[r0];
            if (nc) {
                BM_CK(c, cudaMemcpyAsync(c->d_pl.as<uint64_t>() + c0, src.planes + src.pl_off[r0], nc * 8, cudaMemcpyHostToDevice, c->copy_stream));
                if (src.qk)
                    BM_CK(c, cudaMemcpyAsync(c->d_qk.as<uint32_t>() + c0, src.qk + src.pl_off[r0], nc * 4, cudaMemcpyHostToDevice, c->copy_stream));
            }
        } else if (nb) {
            BM_CK(c, cudaMemcpyAsync(c->d_seq.as<uint8_t>() + b0, src.seq + base + b0, nb, cudaMemcpyHostToDevice, c->copy_stream));
            BM_CK(c, cudaMemcpyAsync(c->d_qual.as<uint8_t>() + b0, src.qual + base + b0, nb, cudaMemcpyHostToDevice, c->copy_stream));
        }
        return BM_OK;
    };
    int r = copy_text(0);
    if (r) return r;
    // Packed batches: the host's pass over the read offsets (long reads? layout valid? longest read) is made chunk by
    // chunk inside the loop instead of up front -- the mapper of chunk 0 starts ~0.8 ms earlier on a batch of 1 M reads.
    // It assumes what is almost always true, that no read is longer than 2r; the chunk that finds one starts the call over.
    const bool lazy = allow_lazy && packed && !publish && !wait_for;
    r = bm_internal_reads_begin(c, p, read_off, n_reads, packed ? src.pl_off : nullptr, lazy);
    if (r) { cudaStreamSynchronize(c->copy_stream); return r; }
    c->ascii_resident = !packed; c->have_qk = !packed || src.qk != nullptr;
    c->pack_k = p->k; c->pack_b = p->avg_base_quality;
    // segments are at most read_len long (q_gram_mapper.h:521), which bounds the sampler's shared memory up front
    const uint32_t msl = p->read_len;
    r = bm_internal_map_begin(c, p, c->n_seg_total, msl);   // K0 + buffers for the whole batch
    if (r) { cudaStreamSynchronize(c->copy_stream); return r; }
    // no read longer than 2r in the batch: one segment per read, segment index == read index, and the four segment
    // tables (16 B per read of link time) are written by a kernel from the read offsets instead of being copied
    const bool unit_segments = c->n_seg_total == n_reads;
    t_pre = since();
    for (uint32_t j = 0; j < n_chunks; j++) {
        const uint32_t r0 = bounds[j], r1 = bounds[j + 1];
        const uint32_t s0 = c->n_seg;
        if (j == 1) t_first = since();
        if (j) { r = copy_text(j); if (r) { cudaStreamSynchronize(c->copy_stream); cudaStreamSynchronize(c->stream); return r; } }
        r = unit_segments ? bm_internal_build_unit_range(c, p, read_off, packed ? src.pl_off : nullptr, r0, r1, lazy)
                          : bm_internal_build_segments(c, p, read_off, r0, r1);
        if (r == 1) {                            // a long read in a lazily scanned batch: once more, with the full pass
            cudaStreamSynchronize(c->copy_stream); cudaStreamSynchronize(c->stream);
            return pipeline_run(c, p, src, read_off, n_reads, filter_best, loc_off, locs, loc_cap, stats, publish, wait_for, n_out, false);
        }
        if (r) { cudaStreamSynchronize(c->copy_stream); cudaStreamSynchronize(c->stream); return r; }
        const uint32_t s1 = c->n_seg;
        cudaStream_t cs = c->copy_stream;
        // read r1's offset / first segment are rewritten by the next chunk with the same values
        BM_CK(c, cudaMemcpyAsync(c->d_read_off.as<uint64_t>() + r0, c->h_off_rel.as<uint64_t>() + r0, (size_t)(r1 - r0 + 1) * 8,
                                 cudaMemcpyHostToDevice, cs));
        BM_CK(c, cudaMemcpyAsync(c->d_pl_off.as<uint32_t>() + r0, c->h_pl_off.as<uint32_t>() + r0, (size_t)(r1 - r0 + 1) * 4,
                                 cudaMemcpyHostToDevice, cs));
        if (!unit_segments)
            BM_CK(c, cudaMemcpyAsync(c->d_read_seg0.as<uint32_t>() + r0, c->h_read_seg0.as<uint32_t>() + r0, (size_t)(r1 - r0 + 1) * 4,
                                     cudaMemcpyHostToDevice, cs));
        if (s1 > s0 && !unit_segments) {
            const size_t n4 = (size_t)(s1 - s0) * 4;
            BM_CK(c, cudaMemcpyAsync(c->d_seg_read.as<uint32_t>() + s0, c->h_seg_read.as<uint32_t>() + s0, n4, cudaMemcpyHostToDevice, cs));
            BM_CK(c, cudaMemcpyAsync(c->d_seg_start.as<uint32_t>() + s0, c->h_seg_start.as<uint32_t>() + s0, n4, cudaMemcpyHostToDevice, cs));
            BM_CK(c, cudaMemcpyAsync(c->d_seg_len.as<uint32_t>() + s0, c->h_seg_len.as<uint32_t>() + s0, n4, cudaMemcpyHostToDevice, cs));
        }
        BM_CK(c, cudaEventRecord(c->chunk_ev[j], cs));
        BM_CK(c, cudaStreamWaitEvent(c->stream, c->chunk_ev[j], 0));
        if (unit_segments && r1 > r0) {
            k_unit_segments<<<(r1 - r0 + 255) / 256, 256, 0, c->stream>>>(
                c->d_read_off.as<uint64_t>(), r0, r1, p->read_len, c->d_seg_read.as<uint32_t>(), c->d_seg_start.as<uint32_t>(),
                c->d_seg_len.as<uint32_t>(), c->d_read_seg0.as<uint32_t>());
            BM_CK(c, cudaGetLastError());
        }
        if (!packed) {                                                              // ASCII -> bit planes, on the device
            r = bm_internal_pack_range(c, p, r0, r1);
            if (r) { cudaStreamSynchronize(c->copy_stream); cudaStreamSynchronize(c->stream); return r; }
        }
        r = bm_internal_map_range(c, p, s0, s1, msl);                               // K2f (or K2a + K2) on this chunk
        if (r) { cudaStreamSynchronize(c->copy_stream); cudaStreamSynchronize(c->stream); return r; }
    }
    t_enq = since();
    if (trace) { cudaStreamSynchronize(c->copy_stream); t_h2d = since(); }
    if (publish) {
        BM_CK(c, cudaEventRecord(publish->h2d_done, c->copy_stream));
        publish->h2d_state.store(1, std::memory_order_release);
    }
    c->have_reads = true;
    r = bm_internal_map_finish(c, p);
    if (r) return r;
    if (trace) { cudaStreamSynchronize(c->stream); t_fin = since(); }
    r = bm_locate_device(c, p, filter_best);                                       // K3a, K3, gather, K3b: whole batch
    if (r) return r;
    if (trace) { cudaStreamSynchronize(c->stream); t_loc = since(); }
    if (stats) BM_CK(c, cudaMemcpyAsync(stats, c->d_stats.p, 8 * 8, cudaMemcpyDeviceToHost, c->stream));
    if (!publish && !wait_for) {
        r = bm_locate_download(c, loc_off, locs, loc_cap);
        if (trace) fprintf(stderr, "[pipe] reads %u: loop entered %.2f  first chunk enqueued %.2f  all enqueued %.2f  h2d done %.2f  map+finish %.2f  locate %.2f  download %.2f ms\n",
                           n_reads, t_pre, t_first, t_enq, t_h2d, t_fin, t_loc, since());
        return r;
    }
    // one of two lanes: lane 0 announces how many records it has, lane 1 appends its own behind them
    uint64_t n = 0;
    r = bm_locate_count(c, &n);
    if (r) return r;
    if (n_out) *n_out = n;
    if (publish) publish->base.store((long long)n, std::memory_order_release);
    long long lbase = 0;
    if (wait_for) {
        while ((lbase = wait_for->base.load(std::memory_order_acquire)) == -1) std::this_thread::yield();
        if (lbase < 0) BM_FAIL(c, BM_E_ARG, "bm_map_locate: the first lane failed");
    }
    if ((uint64_t)lbase + n > loc_cap) BM_FAIL(c, BM_E_ARG, "bm_map_locate: location buffer too small");
    const void* off = c->locs_filtered ? c->d_loc_off2.p : c->d_loc_off.p;
    const void* lsrc = c->locs_filtered ? c->d_locs3.p : c->d_locs.p;
    uint32_t* off_dst = loc_off;
    if (wait_for) {                                 // offsets are rebased on the host: stage them
        BM_CK(c, c->h_lane_off.ensure((size_t)(n_reads + 1) * 4));
        off_dst = c->h_lane_off.as<uint32_t>();
    }
    BM_CK(c, cudaMemcpyAsync(off_dst, off, (size_t)(n_reads + 1) * 4, cudaMemcpyDeviceToHost, c->stream));
    if (n) BM_CK(c, cudaMemcpyAsync(locs + lbase, lsrc, (size_t)n * sizeof(bm_loc), cudaMemcpyDeviceToHost, c->stream));
    BM_CK(c, cudaStreamSynchronize(c->stream));
    if (wait_for) for (uint32_t i = 1; i <= n_reads; i++) loc_off[i] = off_dst[i] + (uint32_t)lbase;   // [0] is lane 0's last
    if (trace) fprintf(stderr, "[pipe] lane %d reads %u: enqueued %.2f  h2d done %.2f  map+finish %.2f  locate %.2f  download %.2f ms\n",
                       wait_for ? 1 : 0, n_reads, t_enq, t_h2d, t_fin, t_loc, since());
    return BM_OK;
}
}  // namespace

extern "C" int bm_pipeline_lanes(bm_ctx* c, uint32_t min_reads) {
    if (!c) return BM_E_ARG;
    c->lane_min_reads = min_reads ? min_reads : 0xFFFFFFFFu;
    return BM_OK;
}

// A second context on the parent's GPU that VIEWS its reference, index and derived tables and owns only its streams and batch
// buffers: batches submitted from two host threads -- one per context -- overlap the host<->device copies of one with the
// kernels of the other (every call of the path waits for its own results, so one context alone cannot).
extern "C" int bm_ctx_create_lane(bm_ctx* c, const bm_params* p, bm_ctx** lane) {
    if (!c) return BM_E_ARG;
    if (!lane || !p) BM_FAIL(c, BM_E_ARG, "bm_ctx_create_lane: null argument");
    *lane = nullptr;
    if (c->lanes_off) BM_FAIL(c, BM_E_ARG, "bm_ctx_create_lane: the parent is a lane itself");
    int r = bm_prepare(c, p);                        // a lane never builds: everything it views exists from here on
    if (r) return r;
    bm_ctx* l = nullptr;
    r = bm_ctx_create(c->device, &l);
    if (r) BM_FAIL(c, r, "bm_ctx_create_lane: cannot create a context on the parent's device");
    l->lanes_off = true;
    lane_share_index(l, c);
    l->ktiming = false;
    *lane = l;
    return BM_OK;
}

extern "C" int bm_map_locate_packed(bm_ctx* c, const bm_params* p, const uint64_t* planes, const uint32_t* qk,
                                    const uint32_t* pl_off, const uint64_t* read_off, uint32_t n_reads, int filter_best,
                                    uint32_t* loc_off, bm_loc* locs, uint64_t loc_cap, uint64_t* stats) {
    if (!c) return BM_E_ARG;
    if (!p || !read_off || !pl_off || (n_reads && !planes)) BM_FAIL(c, BM_E_ARG, "bm_map_locate_packed: bad arguments");
    if (n_reads < kPipeMinReads) {
        int r = bm_reads_upload_packed(c, p, planes, qk, pl_off, read_off, n_reads);
        if (r) return r;
        r = bm_map_device(c, p);
        if (r) return r;
        r = bm_locate_device(c, p, filter_best);
        if (r) return r;
        if (stats) BM_CK(c, cudaMemcpyAsync(stats, c->d_stats.p, 8 * 8, cudaMemcpyDeviceToHost, c->stream));
        return bm_locate_download(c, loc_off, locs, loc_cap);
    }
    HostReads src;
    src.planes = planes; src.qk = qk; src.pl_off = pl_off;
    return pipeline_run(c, p, src, read_off, n_reads, filter_best, loc_off, locs, loc_cap, stats, nullptr, nullptr, nullptr);
}

extern "C" int bm_map_locate(bm_ctx* c, const bm_params* p, const uint8_t* seq, const uint8_t* qual,
                             const uint64_t* read_off, uint32_t n_reads, int filter_best, uint32_t* loc_off, bm_loc* locs,
                             uint64_t loc_cap, uint64_t* stats) {
    if (!c) return BM_E_ARG;
    if (!p || !read_off || (n_reads && (!seq || !qual))) BM_FAIL(c, BM_E_ARG, "bm_map_locate: bad arguments");
    if (n_reads < kPipeMinReads)
        return bm_internal_map_locate_serial(c, p, seq, qual, read_off, n_reads, filter_best, loc_off, locs, loc_cap, stats);
    HostReads src;
    src.seq = seq; src.qual = qual;
    if (n_reads < c->lane_min_reads || n_reads < 2 * kPipeMinReads || !loc_off || !locs || c->lanes_off)
        return pipeline_run(c, p, src, read_off, n_reads, filter_best, loc_off, locs, loc_cap, stats, nullptr, nullptr, nullptr);

    // ---- two lanes ----
    cudaSetDevice(c->device);
    if (!c->have_index) BM_FAIL(c, BM_E_ARG, "bm_map_device: no index resident (the q-gram index is empty)");
    // derived tables are built once, by the parent, before the lanes start (the child only ever views them)
    int r = bm_internal_index_prepare(c, p->distinguishability);
    if (r) return r;
    if (p->k >= p->q && p->k <= 16) { r = bm_internal_kmer_table_prepare(c, p); if (r) return r; }
    r = bm_internal_gidx_prepare(c, p->k);
    if (r) return r;
    BM_CK(c, cudaStreamSynchronize(c->stream));
    if (!c->lane) {
        r = bm_ctx_create(c->device, &c->lane);
        if (r) BM_FAIL(c, r, "bm_map_locate: cannot create the second pipeline lane");
        c->lane->lanes_off = true;
    }
    bm_ctx* l = c->lane;
    lane_share_index(l, c);
    const uint32_t h = n_reads / 2;
    LaneLink link;
    if (!c->lane_ev) BM_CK(c, cudaEventCreateWithFlags(&c->lane_ev, cudaEventDisableTiming));
    link.h2d_done = c->lane_ev;
    uint64_t st0[8] = {0}, st1[8] = {0}, n0 = 0, n1 = 0;
    int r1 = BM_OK;
    std::thread t([&] {
        r1 = pipeline_run(l, p, src, read_off + h, n_reads - h, filter_best, loc_off + h, locs, loc_cap, st1, nullptr,
                          &link, &n1);
    });
    int r0 = pipeline_run(c, p, src, read_off, h, filter_best, loc_off, locs, loc_cap, st0, &link, nullptr, &n0);
    if (r0) { link.h2d_state.store(-1, std::memory_order_release); link.base.store(-2, std::memory_order_release); }
    t.join();
    // the lane's kernel launches and timings count as this context's
    for (int k = 0; k < BM_K_COUNT; k++) {
        double ms = 0; uint64_t ln = 0;
        bm_kernel_stats(l, k, &ms, &ln);
        c->kstat[k].ms += ms; c->kstat[k].launches += ln;
    }
    bm_kernel_stats_reset(l);
    // no batch is resident as a whole after a two-lane call
    c->have_reads = false; c->have_map = false; c->have_locs = false; c->have_aln = false;
    if (r0) return r0;
    if (r1) { c->err = l->err; return r1; }
    if (stats) for (int i = 0; i < 8; i++) stats[i] = st0[i] + st1[i];
    return BM_OK;
}
