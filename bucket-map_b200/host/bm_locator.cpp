// bm_locator.cpp -- locator / bucket_locator of the C++ host: FASTQ batches -> device pipeline -> SAM.
//
// The reference's bucket_locator::locate (locator/bucket_locator.h:455-611) makes three passes over the FASTQ file
// (mapper, k-mer sampling, output) with a per-bucket hash-table rebuild in between.  Here every batch of reads goes
// through the device once: bm_map_locate (K2a, K2, K3a, K3, gather, K3b) for bucketmap, or map / locate / bm_align
// (K4) for bucketmap_align, and the SAM records are rendered from the returned location records in read order.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <condition_variable>
#include <deque>
#include <fstream>
#include <iostream>
#include <mutex>
#include <thread>

#include "bm_host.hpp"

namespace bm {

namespace fs = std::filesystem;

static double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// locator/locator.h:22-36
unsigned int locator::initialize(fs::path const& fasta_file_name, fs::path const& index_directory,
                                 std::string const& indicator) {
    auto res = _i->index(fasta_file_name, index_directory, indicator);
    _i->reset();
    return res;
}

bucket_locator::bucket_locator(indexer* ind, mapper* map, unsigned int bucket_len, unsigned int read_len, uint8_t seed_len,
                               float mismatch_rate, float indel_rate, unsigned int sample_size,
                               unsigned int quality_threshold, unsigned int num_segment_samples, bool align)
    : locator(ind), _m(map), bucket_length(bucket_len), read_length(read_len), k(seed_len), mismatch_rate_(mismatch_rate),
      indel_rate_(indel_rate), sample_size_(sample_size), avg_base_quality_(quality_threshold),
      num_segment_samples_(num_segment_samples), align_(align) {}

// locator/bucket_locator.h:440-453
unsigned int bucket_locator::initialize(fs::path const& fasta_file_name, fs::path const& index_directory,
                                        std::string const& indicator) {
    genome_file_name = fasta_file_name;
    auto res = locator::initialize(fasta_file_name, index_directory, indicator);
    if (auto* qm = dynamic_cast<q_gram_mapper*>(_m)) {
        if (qm->num_buckets() == 0) {
            genome const* g = load_genome_cached(fasta_file_name, bucket_length, read_length);
            if (g) qm->set_num_buckets(g->NB);
        }
    }
    _m->load(index_directory, indicator);
    return res;
}

namespace {

// blocking FIFO between two pipeline stages
template <class T>
struct batch_queue {
    std::mutex m; std::condition_variable cv; std::deque<T> q;
    void push(T v) { { std::lock_guard<std::mutex> g(m); q.push_back(v); } cv.notify_one(); }
    bool pop(T& v) { std::unique_lock<std::mutex> g(m); cv.wait(g, [&] { return !q.empty(); }); v = q.front(); q.pop_front(); return true; }
};

struct sam_writer {
    FILE* f = nullptr;
    std::string buf;
    explicit sam_writer(fs::path const& p) { f = fopen(p.c_str(), "wb"); buf.reserve(1 << 24); }
    ~sam_writer() { flush(); if (f) fclose(f); }
    void flush() { if (f && !buf.empty()) fwrite(buf.data(), 1, buf.size(), f); buf.clear(); }
    void maybe_flush() { if (buf.size() > (1 << 24) - (1 << 16)) flush(); }
    void num(uint64_t v) { char t[24]; int n = snprintf(t, sizeof t, "%llu", (unsigned long long)v); buf.append(t, (size_t)n); }
};

inline char dna4_char(uint8_t c) {         // FASTQ read as dna4 and printed back (seqan3 dna4 conversion)
    switch (c & 0xDF) {
        case 'C': case 'Y': case 'S': case 'B': return 'C';
        case 'G': case 'K': return 'G';
        case 'T': case 'U': return 'T';
        default: return 'A';
    }
}

}   // namespace

// locator/bucket_locator.h:455-611
void bucket_locator::locate(fs::path const& sequence_file, fs::path const& index_file, fs::path const& sam_file,
                            unsigned int quality_threshold) {
    auto* qm = dynamic_cast<q_gram_mapper*>(_m);
    if (!qm) {
        std::cerr << "[ERROR]\t\tbucket_locator needs a bm::q_gram_mapper (the device pipeline maps and locates in one pass).\n";
        return;
    }
    if (!qm->loaded()) { std::cerr << "[ERROR]\t\tThe q-gram index is empty. Cannot accept query.\n"; return; }
    bm_ctx* c = qm->ctx()->get();
    genome const* g = load_genome_cached(genome_file_name, bucket_length, read_length);
    if (!g) return;
    const uint32_t n_emit = (uint32_t)g->b_start.size();
    int rc = bm_ref_upload(c, g->packed.data(), g->n_bases, g->b_start.data(), g->b_len.data(), n_emit, g->NB);
    if (rc) { std::cerr << "[ERROR]\t\t" << bm_last_error(c) << "\n"; return; }

    bm_params p;
    qm->fill_params(p);
    p.k = k;
    p.locator_samples = sample_size_;
    p.allowed_mismatch = (uint32_t)(int)std::ceil(mismatch_rate_ * sample_size_);     // :419 (float * unsigned)
    p.allowed_indel = (uint32_t)(int)std::ceil(indel_rate_ * read_length);            // :420
    p.indel_rate = indel_rate_;
    p.avg_base_quality = avg_base_quality_;
    p.num_segment_samples = num_segment_samples_;
    p.quality_threshold = quality_threshold;

    // bucket names / offsets / @SQ lines (:473-503)
    std::vector<std::string> bucket_name, ref_ids;
    std::vector<unsigned int> bucket_offsets;
    std::vector<size_t> ref_lengths;
    {
        std::ifstream bucket_info(index_file);
        std::string name, last_bucket_name;
        unsigned int bucket_index = 0;
        for (uint32_t i = 0; i < n_emit; i++) {
            std::getline(bucket_info, name);
            name = name.substr(0, name.find(' '));
            if (name != last_bucket_name) {
                if (bucket_index != 0) { ref_ids.push_back(last_bucket_name); ref_lengths.push_back((size_t)bucket_index * bucket_length); }
                last_bucket_name = name;
                bucket_index = 0;
            }
            bucket_name.push_back(name);
            bucket_offsets.push_back(bucket_index * bucket_length);
            bucket_index++;
        }
        if (bucket_index != 0) { ref_ids.push_back(last_bucket_name); ref_lengths.push_back((size_t)bucket_index * bucket_length); }
    }
    fastq_reader fin(sequence_file);
    if (!fin.ok()) { std::cerr << "[ERROR]\t\tCannot read " << sequence_file << ".\n"; return; }
    sam_writer out(sam_file);
    if (!out.f) { std::cerr << "[ERROR]\t\tCannot write " << sam_file << ".\n"; return; }
    out.buf += "@HD\tVN:1.6\n";
    for (size_t i = 0; i < ref_ids.size(); i++) { out.buf += "@SQ\tSN:" + ref_ids[i] + "\tLN:"; out.num(ref_lengths[i]); out.buf += '\n'; }

    bm_kernel_timing(c, 1);
    bm_kernel_stats_reset(c);

    // Three stages on three threads, batches handed on through two queues: FASTQ parsing into pinned buffers, the device
    // pass (this thread: one context, one host thread at a time), SAM rendering.  Records leave in read order because
    // every stage is a single thread working through the batches in order.
    struct work {
        read_batch b;
        std::vector<uint32_t> loc_off, cigar;
        std::vector<bm_loc> locs;
        std::vector<bm_aln> aln;
    };
    constexpr int kDepth = 3;
    work pool[kDepth];
    batch_queue<work*> free_q, parsed_q, done_q;
    for (auto& w : pool) free_q.push(&w);
    const uint32_t batch_reads = batch_reads_;
    double parse_s = 0;
    const double t_pipe = now_s();
    std::thread reader([&] {
        work* w;
        while (free_q.pop(w) && w) {
            const double t = now_s();
            const bool more = fin.next(w->b, batch_reads, (uint64_t)1 << 31);
            parse_s += now_s() - t;
            if (!more) break;
            parsed_q.push(w);
        }
        parsed_q.push(nullptr);
    });

    std::string seq_text;
    unsigned int read_id = 0, mapped_locations = 0;
    double device_s = 0, output_s = 0;
    const bool align = align_;
    std::thread writer([&] {
        static const char* kDecode = nullptr;
        static char decode[256];
        if (!kDecode) { for (int v = 0; v < 256; v++) decode[v] = dna4_char((uint8_t)v); kDecode = decode; }
        work* w;
        while (done_q.pop(w) && w) {
            const double t1 = now_s();
            const read_batch& b = w->b;
            const uint32_t n = b.size();
            // SAM records in read order (:535-604)
            for (uint32_t i = 0; i < n; i++) {
                if (w->loc_off[i] == w->loc_off[i + 1]) continue;
                const uint8_t* s = &b.seq[b.off[i]];
                const size_t len = (size_t)(b.off[i + 1] - b.off[i]);
                seq_text.resize(len);
                for (size_t t = 0; t < len; t++) seq_text[t] = kDecode[s[t]];
                for (uint32_t j = w->loc_off[i]; j < w->loc_off[i + 1]; j++) {
                    const bm_loc& l = w->locs[j];
                    size_t map_qual, ref_offset;
                    if (!align) {
                        map_qual = std::min((unsigned int)60, 6 * l.votes);                  // :591
                        ref_offset = (unsigned int)(bucket_offsets[l.bucket] + l.offset);    // :592
                    } else {
                        if (!w->aln[j].kept) continue;                                       // :571-573
                        map_qual = (uint8_t)w->aln[j].map_qual;                              // mapq is a uint8_t field in SAM
                        ref_offset = (unsigned int)(w->aln[j].begin + bucket_offsets[l.bucket] + l.offset);   // :576
                    }
                    out.buf.append(b.id(i), b.id_len(i)); out.buf += '\t';
                    out.num(l.is_original ? 0 : 16); out.buf += '\t';
                    out.buf += bucket_name[l.bucket]; out.buf += '\t';
                    out.num(ref_offset + 1); out.buf += '\t';
                    out.num(map_qual); out.buf += '\t';
                    if (align && w->aln[j].cigar_len) {
                        for (uint32_t t = 0; t < w->aln[j].cigar_len; t++) {
                            const uint32_t op = w->cigar[w->aln[j].cigar_off + t];
                            out.num(op >> 4); out.buf += "MID"[op & 15u];
                        }
                    } else out.buf += '*';
                    out.buf += "\t*\t0\t0\t";
                    out.buf += seq_text; out.buf += '\t';
                    out.buf.append((const char*)&b.qual[b.off[i]], len);
                    out.buf += '\n';
                    mapped_locations++;
                }
                out.maybe_flush();
            }
            read_id += n;
            output_s += now_s() - t1;
            free_q.push(w);
        }
    });

    bool failed = false;
    {
        work* w;
        while (parsed_q.pop(w) && w) {
            read_batch& b = w->b;
            const uint32_t n = b.size();
            const double t0 = now_s();
            uint64_t stats[8] = {0};
            uint64_t n_locs = 0;
            w->loc_off.resize((size_t)n + 1);
            rc = BM_OK;
            bool done = false;
            if (!align_) {
                // one call: H2D chunks overlapped with the mapper.  The merged lists are almost always one record per
                // read; if this batch needs more room than 4 per read the staged calls below size the buffer exactly.
                const size_t cap = (size_t)4 * n + 1024;
                w->locs.resize(cap);
                int r1 = bm_map_locate(c, &p, b.seq.data(), b.qual.data(), b.off.data(), n, 1, w->loc_off.data(), w->locs.data(),
                                       cap, stats);
                if (r1 == BM_OK) { n_locs = w->loc_off[n]; done = true; }
            }
            if (!done) {
                rc = bm_reads_upload(c, &p, b.seq.data(), b.qual.data(), b.off.data(), n);
                if (!rc) rc = bm_map_device(c, &p);
                if (!rc) rc = bm_locate_device(c, &p, align_ ? 0 : 1);
                if (!rc) rc = bm_locate_count(c, &n_locs);
                if (!rc) { w->locs.resize(n_locs + 1); rc = bm_locate_download(c, w->loc_off.data(), w->locs.data(), n_locs + 1); }
                if (align_) {
                    if (!rc) rc = bm_align_device(c, &p);                                   // K4 (:560-589)
                    uint64_t used = 0;
                    w->aln.resize(n_locs + 1);
                    if (!rc) rc = bm_align_download(c, w->aln.data(), nullptr, 0, &used);
                    w->cigar.resize(used + 1);
                    if (!rc && used) rc = bm_align_download(c, w->aln.data(), w->cigar.data(), used, &used);
                }
                if (!rc) rc = bm_map_download(c, nullptr, nullptr, 0, stats);
            }
            if (rc) { std::cerr << "[ERROR]\t\t" << bm_last_error(c) << "\n"; failed = true; break; }
            qm->mapped_reads += (unsigned)stats[0]; qm->num_buckets_orig += (unsigned)stats[1]; qm->num_buckets_rev_comp += (unsigned)stats[2];
            qm->num_records += n;
            device_s += now_s() - t0;
            done_q.push(w);
        }
    }
    done_q.push(nullptr);
    writer.join();
    if (failed) { free_q.push(nullptr); work* w; while (parsed_q.pop(w) && w) free_q.push(w); }
    else free_q.push(nullptr);
    reader.join();
    if (failed) return;
    out.flush();

    // the reference's phase log (q_gram_mapper.h:548-555, bucket_locator.h:698-702,606-610), filled with device times
    double ms[BM_K_COUNT]; uint64_t launches[BM_K_COUNT];
    for (int i = 0; i < BM_K_COUNT; i++) { ms[i] = 0; launches[i] = 0; bm_kernel_stats(c, i, &ms[i], &launches[i]); }
    bm_kernel_timing(c, 0);
    if (qm->num_records) qm->print_map_benchmark((float)((ms[BM_K_SAMPLE] + ms[BM_K_MAP]) * 1e-3));
    std::cerr << "[BENCHMARK]\tTotal time used for building k-mer index for each bucket: 0 s.\n";
    const float time = (float)((ms[BM_K_LOCSAMPLE] + ms[BM_K_LOCATE] + ms[BM_K_MERGE]) * 1e-3);
    std::cerr << "[BENCHMARK]\tTotal time used for finding exact location of the sequences: " << time << " s ("
              << time * 1000 * 1000 / (read_id ? read_id : 1) << " μs/seq).\n";
    std::cerr << "[BENCHMARK]\tTotal mapped locations: " << mapped_locations << " (" << (float)mapped_locations / read_id
              << " per sequence).\n";
    const float es = (float)(ms[BM_K_ALIGN] * 1e-3 + output_s);
    std::cerr << "[BENCHMARK]\tTotal time used for alignment verification and output: " << es << " s ("
              << es / mapped_locations * 1000 * 1000 << " μs per pairwise alignment).\n";
    std::cerr << "[BENCHMARK]\tHost pipeline: FASTQ parsing " << (float)parse_s << " s, device passes " << (float)device_s
              << " s, SAM rendering " << (float)output_s << " s on three threads; wall " << (float)(now_s() - t_pipe) << " s.\n";
    std::cerr << "[BENCHMARK]\tDevice pipeline (H2D + kernels + D2H): " << (float)device_s << " s for " << read_id
              << " reads (" << (float)(read_id / (device_s > 0 ? device_s : 1)) << " reads/s).\n";
    _m->reset();                                                                       // :627
}

}   // namespace bm
