// bm_pipeline.cu -- bm_map_locate: the host-buffer entry point of the alignment-free path (what bucket_locator::locate
// does per FASTQ batch), with the host->device copy of the batch pipelined against the mapper kernels.
//
// The batch is cut into chunks of reads.  For each chunk the host appends its slice of the segment table (pinned
// staging), the copy stream moves the chunk's bases, qualities, offsets and table slice to the device and records an
// event; the compute stream waits for that event and runs K2a + K2 on the chunk's segment range.  The mapper (HBM-bound,
// ~3/4 of the device time) therefore starts after the first chunk has landed and hides the rest of the 600 B/read
// transfer.  The locator runs ONCE over the whole batch afterwards: its cost is per non-empty bucket (it streams each
// bucket's text once for all candidates in it), so it must not be split.  Results do not depend on the chunking.
#include "bm_internal.cuh"

namespace {
constexpr uint32_t kPipeMinReads = 32768;      // below this the serial path is used
constexpr uint32_t kChunkMin = 16384, kChunkMax = 131072;
}  // namespace

extern "C" int bm_map_locate(bm_ctx* c, const bm_params* p, const uint8_t* seq, const uint8_t* qual,
                             const uint64_t* read_off, uint32_t n_reads, int filter_best, uint32_t* loc_off, bm_loc* locs,
                             uint64_t loc_cap, uint64_t* stats) {
    if (!c) return BM_E_ARG;
    if (!p || !read_off || (n_reads && (!seq || !qual))) BM_FAIL(c, BM_E_ARG, "bm_map_locate: bad arguments");
    if (n_reads < kPipeMinReads)
        return bm_internal_map_locate_serial(c, p, seq, qual, read_off, n_reads, filter_best, loc_off, locs, loc_cap, stats);
    cudaSetDevice(c->device);
    if (!c->copy_stream) BM_CK(c, cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));

    uint32_t chunk = n_reads / 16;
    chunk = chunk < kChunkMin ? kChunkMin : (chunk > kChunkMax ? kChunkMax : chunk);
    const uint32_t n_chunks = (n_reads + chunk - 1) / chunk;
    while (c->chunk_ev.size() < n_chunks) {
        cudaEvent_t e;
        BM_CK(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        c->chunk_ev.push_back(e);
    }
    int r = bm_internal_reads_begin(c, p, read_off, n_reads);
    if (r) return r;
    // segments are at most read_len long (q_gram_mapper.h:521), which bounds the sampler's shared memory up front
    const uint32_t msl = p->read_len;
    r = bm_internal_map_begin(c, p, c->n_seg_total, msl);   // K0 + buffers for the whole batch
    if (r) return r;
    // the copy stream must not overwrite buffers an earlier call's kernels may still read
    BM_CK(c, cudaStreamSynchronize(c->stream));

    const uint64_t base = read_off[0];
    uint8_t *d_seq = c->d_seq.as<uint8_t>(), *d_qual = c->d_qual.as<uint8_t>();
    for (uint32_t j = 0; j < n_chunks; j++) {
        const uint32_t r0 = j * chunk, r1 = (r0 + chunk < n_reads) ? r0 + chunk : n_reads;
        const uint32_t s0 = c->n_seg;
        r = bm_internal_build_segments(c, p, read_off, r0, r1);
        if (r) { cudaStreamSynchronize(c->copy_stream); cudaStreamSynchronize(c->stream); return r; }
        const uint32_t s1 = c->n_seg;
        const uint64_t b0 = read_off[r0] - base, nb = read_off[r1] - read_off[r0];
        cudaStream_t cs = c->copy_stream;
        if (nb) {
            BM_CK(c, cudaMemcpyAsync(d_seq + b0, seq + base + b0, nb, cudaMemcpyHostToDevice, cs));
            BM_CK(c, cudaMemcpyAsync(d_qual + b0, qual + base + b0, nb, cudaMemcpyHostToDevice, cs));
        }
        // read r1's offset / first segment are rewritten by the next chunk with the same values
        BM_CK(c, cudaMemcpyAsync(c->d_read_off.as<uint64_t>() + r0, c->h_off_rel.as<uint64_t>() + r0, (size_t)(r1 - r0 + 1) * 8,
                                 cudaMemcpyHostToDevice, cs));
        BM_CK(c, cudaMemcpyAsync(c->d_read_seg0.as<uint32_t>() + r0, c->h_read_seg0.as<uint32_t>() + r0, (size_t)(r1 - r0 + 1) * 4,
                                 cudaMemcpyHostToDevice, cs));
        if (s1 > s0) {
            const size_t n4 = (size_t)(s1 - s0) * 4;
            BM_CK(c, cudaMemcpyAsync(c->d_seg_read.as<uint32_t>() + s0, c->h_seg_read.as<uint32_t>() + s0, n4, cudaMemcpyHostToDevice, cs));
            BM_CK(c, cudaMemcpyAsync(c->d_seg_start.as<uint32_t>() + s0, c->h_seg_start.as<uint32_t>() + s0, n4, cudaMemcpyHostToDevice, cs));
            BM_CK(c, cudaMemcpyAsync(c->d_seg_len.as<uint32_t>() + s0, c->h_seg_len.as<uint32_t>() + s0, n4, cudaMemcpyHostToDevice, cs));
        }
        BM_CK(c, cudaEventRecord(c->chunk_ev[j], cs));
        BM_CK(c, cudaStreamWaitEvent(c->stream, c->chunk_ev[j], 0));
        r = bm_internal_map_range(c, p, s0, s1, msl);                               // K2a + K2 on this chunk
        if (r) { cudaStreamSynchronize(c->copy_stream); cudaStreamSynchronize(c->stream); return r; }
    }
    c->have_reads = true;
    r = bm_internal_map_finish(c, p);
    if (r) return r;
    r = bm_locate_device(c, p, filter_best);                                       // K3a, K3, gather, K3b: whole batch
    if (r) return r;
    if (stats) BM_CK(c, cudaMemcpyAsync(stats, c->d_stats.p, 8 * 8, cudaMemcpyDeviceToHost, c->stream));
    return bm_locate_download(c, loc_off, locs, loc_cap);
}
