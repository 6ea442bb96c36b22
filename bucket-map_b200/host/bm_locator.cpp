// bm_locator.cpp -- locator / bucket_locator of the C++ host: FASTQ batches -> device pipeline -> SAM.
//
// The reference's bucket_locator::locate (locator/bucket_locator.h:455-611) makes three passes over the FASTQ file
// (mapper, k-mer sampling, output) with a per-bucket hash-table rebuild in between.  Here every batch of reads goes
// through the device once: bm_map_locate (K2a, K2, K3a, K3, gather, K3b) for bucketmap, or map / locate / bm_align
// (K4) for bucketmap_align, and the SAM records are rendered from the returned location records in read order.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <condition_variable>
#include <atomic>
#include <deque>
#include <fstream>
#include <iostream>
#include <map>
#include <memory>
#include <mutex>
#include <thread>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

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
    void push(T v) { { std::lock_guard<std::mutex> g(m); q.push_back(v); } cv.notify_all(); }
    bool pop(T& v) { std::unique_lock<std::mutex> g(m); cv.wait(g, [&] { return !q.empty(); }); v = q.front(); q.pop_front(); return true; }
};

inline char dna4_char(uint8_t c) {         // FASTQ read as dna4 and printed back (seqan3 dna4 conversion)
    switch (c & 0xDF) {
        case 'C': case 'Y': case 'S': case 'B': return 'C';
        case 'G': case 'K': return 'G';
        case 'T': case 'U': return 'T';
        default: return 'A';
    }
}

inline void put_num(std::string& out, uint64_t v) {
    char t[24];
    int n = 0;
    do { t[n++] = (char)('0' + v % 10); v /= 10; } while (v);
    while (n) out += t[--n];
}

// the two sinks of the SAM renderer: the size of the text, and the text
struct count_sink {
    uint64_t n = 0;
    void append(const char*, size_t len) { n += len; }
    void push(char) { n++; }
    void fill(size_t count, char) { n += count; }
    void num(uint64_t v) { do { n++; v /= 10; } while (v); }
    void bases(const uint8_t*, size_t len, const char*) { n += len; }
};
struct mem_sink {
    char* p;
    void append(const char* s, size_t len) { memcpy(p, s, len); p += len; }
    void push(char c) { *p++ = c; }
    void fill(size_t count, char c) { memset(p, c, count); p += count; }
    void num(uint64_t v) {
        char t[24];
        int n = 0;
        do { t[n++] = (char)('0' + v % 10); v /= 10; } while (v);
        while (n) *p++ = t[--n];
    }
    void bases(const uint8_t* s, size_t len, const char* decode) { for (size_t x = 0; x < len; x++) p[x] = decode[s[x]]; p += len; }
};

// one batch on its way through the pipeline
struct work {
    read_batch b;
    uint64_t seq_no = 0;
    pinned_bytes loc_off, locs;            // uint32[n+1], bm_loc[cap]: the device writes them directly
    size_t loc_cap = 0;
    std::vector<bm_aln> aln;
    std::vector<uint32_t> cigar;
    std::vector<std::string> text;         // SAM text of the batch, one piece per render thread
    bool failed = false;
};

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
    bm_ctx* c0 = qm->ctx()->get();
    genome const* g = load_genome_cached(genome_file_name, bucket_length, read_length);
    if (!g) return;
    const uint32_t n_emit = (uint32_t)g->b_start.size();
    int rc = bm_ref_upload(c0, g->packed.data(), g->n_bases, g->b_start.data(), g->b_len.data(), n_emit, g->NB);
    if (rc) { std::cerr << "[ERROR]\t\t" << bm_last_error(c0) << "\n"; return; }

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

    fastq_reader fin(sequence_file);
    if (!fin.ok()) { std::cerr << "[ERROR]\t\tCannot read " << sequence_file << ".\n"; return; }
    // The two big derived tables pay for themselves only on long runs: the locator's genome k-mer index (16 B per base of
    // scratch, ~50 ms on 1.7 Gbp, saves ~5 ms per million reads) and the full-resolution k-mer table (tens of GB, saves
    // ~1 ms per million reads).  A run of a few million reads is faster without building, allocating and freeing them;
    // $BM_LOCATOR_INDEX / $BM_FINE_TABLE decide explicitly.  Results are identical either way.
    int want_lidx = -1, want_fine = -1;            // -1: leave the context's setting alone
    {
        std::error_code ec;
        const uint64_t fq_bytes = (uint64_t)fs::file_size(sequence_file, ec);
        const uint64_t est_reads = ec ? 0 : fq_bytes / (2ull * read_length + 16);
        if (!std::getenv("BM_LOCATOR_INDEX")) want_lidx = est_reads >= 8000000ull ? 1 : 0;
        if (!std::getenv("BM_FINE_TABLE")) want_fine = est_reads >= 16000000ull ? 1 : 0;
        if (want_lidx >= 0) bm_locator_index(c0, want_lidx);
        if (want_fine >= 0) bm_index_fine_table(c0, want_fine);
    }
    bool out_mappable = true;                     // regular file opened read-write: the SAM text is copied into a mapping
    int out_fd = ::open(sam_file.c_str(), O_RDWR | O_CREAT | O_TRUNC, 0644);
    if (out_fd < 0) { out_mappable = false; out_fd = ::open(sam_file.c_str(), O_WRONLY | O_CREAT | O_TRUNC, 0644); }
    if (out_fd < 0) { std::cerr << "[ERROR]\t\tCannot write " << sam_file << ".\n"; return; }
    {
        struct stat st;
        if (fstat(out_fd, &st) != 0 || !S_ISREG(st.st_mode) || std::getenv("BM_SAM_PWRITE")) out_mappable = false;
    }
    const bool host_trace = std::getenv("BM_HOST_TRACE") != nullptr;

    // ---- the parser starts on the first batch now; the device side gets ready meanwhile ----------------------------------
    struct dev { bm_ctx* c; std::unique_ptr<context> own; };
    std::vector<dev> devs;
    devs.push_back({c0, nullptr});
    const size_t n_dev = devices_.size() > 1 ? devices_.size() : 1;
    const size_t depth = n_dev + 2;        // one batch being parsed, one per device, one being rendered
    std::vector<work> pool(depth);
    batch_queue<work*> free_q, parsed_q, done_q;
    for (auto& w : pool) free_q.push(&w);
    const uint32_t batch_reads = batch_reads_;
    double parse_s = 0;
    const double t_pipe = now_s();
    std::thread reader([&] {
        work* w;
        uint64_t seq_no = 0;
        while (free_q.pop(w) && w) {
            const double t = now_s();
            const bool more = fin.next(w->b, batch_reads, (uint64_t)1 << 31, p.k, p.avg_base_quality);
            parse_s += now_s() - t;
            if (!more) break;
            w->seq_no = seq_no++;
            parsed_q.push(w);
        }
        for (size_t i = 0; i < n_dev; i++) parsed_q.push(nullptr);
    });

    // several GPUs: one context per device; reference and index travel from the mapper's context over NCCL
    bool setup_failed = false;
    const double t_setup = now_s();
    if (n_dev > 1 && !setup_failed) {
        for (size_t i = 1; i < n_dev; i++) {
            std::unique_ptr<context> cx(new context(devices_[i]));
            bm_ctx* c = cx->get();
            if (want_lidx >= 0) bm_locator_index(c, want_lidx);
            if (want_fine >= 0) bm_index_fine_table(c, want_fine);
            devs.push_back({c, std::move(cx)});
        }
        uint8_t id[BM_COMM_ID_BYTES];
        if (bm_comm_unique_id(id) != BM_OK) { std::cerr << "[ERROR]\t\tNCCL (libnccl.so.2) is needed for --devices and cannot be loaded.\n"; setup_failed = true; }
        std::vector<int> rcs(n_dev, 0);
        if (!setup_failed) {
            parallel_parts((unsigned)n_dev, [&](unsigned i, unsigned) {
                bm_ctx* c = devs[i].c;
                int r = bm_comm_init(c, id, (int)n_dev, (int)i);
                if (!r) r = bm_bcast_ref(c, 0);
                if (!r) r = bm_bcast_index(c, 0);
                if (!r) r = bm_prepare(c, &p);
                rcs[i] = r;
            });
            for (size_t i = 0; i < n_dev; i++)
                if (rcs[i]) { std::cerr << "[ERROR]\t\t" << bm_last_error(devs[i].c) << "\n"; setup_failed = true; }
        }
    } else if (!setup_failed) {
        rc = bm_prepare(c0, &p);               // row zeros, k-mer tables, locator index: before the first batch arrives
        if (rc) { std::cerr << "[ERROR]\t\t" << bm_last_error(c0) << "\n"; setup_failed = true; }
    }
    const double setup_s = now_s() - t_setup;

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
    uint64_t file_pos = 0;
    {
        std::string hdr = "@HD\tVN:1.6\n";
        for (size_t i = 0; i < ref_ids.size(); i++) { hdr += "@SQ\tSN:" + ref_ids[i] + "\tLN:"; put_num(hdr, ref_lengths[i]); hdr += '\n'; }
        if (::pwrite(out_fd, hdr.data(), hdr.size(), 0) != (ssize_t)hdr.size()) setup_failed = true;
        file_pos = hdr.size();
    }
    for (auto& d : devs) { bm_kernel_timing(d.c, 1); bm_kernel_stats_reset(d.c); }

    // ---- stage 2: one thread per device -----------------------------------------------------------------------------------
    std::mutex stat_m;
    double device_s = 0;
    std::atomic<bool> failed{setup_failed};
    const bool align = align_;
    auto device_loop = [&](size_t di) {
        bm_ctx* c = devs[di].c;
        work* w;
        while (parsed_q.pop(w) && w) {
            if (failed.load()) { w->failed = true; done_q.push(w); continue; }
            read_batch& b = w->b;
            const uint32_t n = b.size();
            const double t0 = now_s();
            uint64_t stats[8] = {0};
            uint64_t n_locs = 0;
            const uint64_t* planes = b.planes.as<uint64_t>();
            const uint32_t* qkp = b.all_pass ? nullptr : b.qk.as<uint32_t>();
            const uint32_t* plo = b.pl_off.as<uint32_t>();
            const uint64_t* roff = b.read_off.as<uint64_t>();
            w->loc_off.reserve((size_t)(n + 1) * 4);
            int r = BM_OK;
            bool done = false;
            if (!align) {
                // one call: H2D chunks overlapped with the mapper.  The merged lists are almost always one record per
                // read; if this batch needs more room than 2 per read the staged calls below size the buffer exactly.
                w->loc_cap = (size_t)2 * n + 1024;
                w->locs.reserve(w->loc_cap * sizeof(bm_loc));
                int r1 = bm_map_locate_packed(c, &p, planes, qkp, plo, roff, n, 1, w->loc_off.as<uint32_t>(), w->locs.as<bm_loc>(),
                                              w->loc_cap, stats);
                if (r1 == BM_OK) { n_locs = w->loc_off.as<uint32_t>()[n]; done = true; }
            }
            if (!done) {
                r = bm_reads_upload_packed(c, &p, planes, qkp, plo, roff, n);
                if (!r) r = bm_map_device(c, &p);
                if (!r) r = bm_locate_device(c, &p, align ? 0 : 1);
                if (!r) r = bm_locate_count(c, &n_locs);
                if (!r) {
                    w->loc_cap = n_locs + 1;
                    w->locs.reserve(w->loc_cap * sizeof(bm_loc));
                    r = bm_locate_download(c, w->loc_off.as<uint32_t>(), w->locs.as<bm_loc>(), w->loc_cap);
                }
                if (align) {
                    if (!r) r = bm_align_device(c, &p);                                   // K4 (:560-589)
                    uint64_t used = 0;
                    w->aln.resize(n_locs + 1);
                    if (!r) r = bm_align_download(c, w->aln.data(), nullptr, 0, &used);
                    w->cigar.resize(used + 1);
                    if (!r && used) r = bm_align_download(c, w->aln.data(), w->cigar.data(), used, &used);
                }
                if (!r) r = bm_map_download(c, nullptr, nullptr, 0, stats);
            }
            if (r) {
                std::lock_guard<std::mutex> lk(stat_m);
                std::cerr << "[ERROR]\t\t" << bm_last_error(c) << "\n";
                failed.store(true); w->failed = true;
            } else {
                std::lock_guard<std::mutex> lk(stat_m);
                qm->mapped_reads += (unsigned)stats[0]; qm->num_buckets_orig += (unsigned)stats[1]; qm->num_buckets_rev_comp += (unsigned)stats[2];
                qm->num_records += n;
                device_s += now_s() - t0;
            }
            done_q.push(w);
        }
        done_q.push(nullptr);
    };
    std::vector<std::thread> dev_threads;
    for (size_t di = 0; di < n_dev; di++) dev_threads.emplace_back(device_loop, di);

    // ---- stage 3: SAM rendering on the host threads, records in read order (:535-604) ---------------------------------------
    unsigned int read_id = 0, mapped_locations = 0;
    double output_s = 0;
    {
        static char decode[256];
        for (int v = 0; v < 256; v++) decode[v] = dna4_char((uint8_t)v);
        std::map<uint64_t, work*> waiting;
        uint64_t next_seq = 0;
        size_t ended = 0;
        // The records of reads [r0, r1) into a sink: once into a counting sink (the exact size of every thread's piece, so that
        // the pieces have their places in the file before a byte is written), once into memory -- the mapping of the file's new
        // tail, or, for outputs that cannot be mapped, a buffer that pwrite sends.
        auto emit = [&](auto& out, work* w, uint32_t r0, uint32_t r1) -> unsigned {
            const read_batch& b = w->b;
            const uint32_t* loc_off = w->loc_off.as<uint32_t>();
            const bm_loc* locs = w->locs.as<bm_loc>();
            unsigned records = 0;
            for (uint32_t i = r0; i < r1; i++) {
                if (loc_off[i] == loc_off[i + 1]) continue;
                const size_t len = (size_t)(b.off[i + 1] - b.off[i]);
                for (uint32_t j = loc_off[i]; j < loc_off[i + 1]; j++) {
                    const bm_loc& l = locs[j];
                    size_t map_qual, ref_offset;
                    if (!align) {
                        map_qual = std::min((unsigned int)60, 6 * l.votes);                  // :591
                        ref_offset = (unsigned int)(bucket_offsets[l.bucket] + l.offset);    // :592
                    } else {
                        if (!w->aln[j].kept) continue;                                       // :571-573
                        map_qual = (uint8_t)w->aln[j].map_qual;                              // mapq is a uint8_t field in SAM
                        ref_offset = (unsigned int)(w->aln[j].begin + bucket_offsets[l.bucket] + l.offset);   // :576
                    }
                    out.append(b.id[i], b.id_len[i]); out.push('\t');
                    out.num(l.is_original ? 0 : 16); out.push('\t');
                    out.append(bucket_name[l.bucket].data(), bucket_name[l.bucket].size()); out.push('\t');
                    out.num(ref_offset + 1); out.push('\t');
                    out.num(map_qual); out.push('\t');
                    if (align && w->aln[j].cigar_len) {
                        for (uint32_t x = 0; x < w->aln[j].cigar_len; x++) {
                            const uint32_t op = w->cigar[w->aln[j].cigar_off + x];
                            out.num(op >> 4); out.push("MID"[op & 15u]);
                        }
                    } else out.push('*');
                    out.append("\t*\t0\t0\t", 7);
                    out.bases((const uint8_t*)b.seq[i], len, decode); out.push('\t');
                    out.append(b.qual[i], b.qual_len[i]);
                    if (b.qual_len[i] < len) out.fill(len - b.qual_len[i], '!');
                    out.push('\n');
                    records++;
                }
            }
            return records;
        };
        auto render = [&](work* w) {
            const double t1 = now_s();
            const uint32_t n = w->b.size();
            unsigned T = host_threads();
            if (n < 2048) T = 1;
            std::vector<uint64_t> bytes(T, 0), at(T + 1, file_pos);
            std::vector<unsigned> cnt(T, 0);
            auto range = [&](unsigned t, unsigned parts, uint32_t& r0, uint32_t& r1) {
                r0 = (uint32_t)((uint64_t)n * t / parts); r1 = (uint32_t)((uint64_t)n * (t + 1) / parts);
            };
            parallel_parts(T, [&](unsigned t, unsigned parts) {
                uint32_t r0, r1;
                range(t, parts, r0, r1);
                count_sink c;
                cnt[t] = emit(c, w, r0, r1);
                bytes[t] = c.n;
            });
            for (unsigned t = 0; t < T; t++) { at[t + 1] = at[t] + bytes[t]; mapped_locations += cnt[t]; }
            const double t_counted = now_s();
            bool ok = true, mapped = false;
            if (out_mappable && at[T] > file_pos) {
                const uint64_t page = (uint64_t)sysconf(_SC_PAGESIZE);
                const uint64_t a0 = file_pos & ~(page - 1);
                void* m = MAP_FAILED;
                if (::ftruncate(out_fd, (off_t)at[T]) == 0)
                    m = ::mmap(nullptr, (size_t)(at[T] - a0), PROT_READ | PROT_WRITE, MAP_SHARED, out_fd, (off_t)a0);
                if (m == MAP_FAILED) out_mappable = false;
                else {
                    parallel_parts(T, [&](unsigned t, unsigned parts) {
                        uint32_t r0, r1;
                        range(t, parts, r0, r1);
                        mem_sink dst{(char*)m + (at[t] - a0)};
                        emit(dst, w, r0, r1);
                    });
                    ::munmap(m, (size_t)(at[T] - a0));
                    mapped = true;
                }
            }
            if (!mapped) {
                if (w->text.size() < T) w->text.resize(T);
                parallel_parts(T, [&](unsigned t, unsigned parts) {
                    uint32_t r0, r1;
                    range(t, parts, r0, r1);
                    std::string& buf = w->text[t];
                    buf.resize((size_t)bytes[t]);
                    mem_sink dst{buf.data()};
                    emit(dst, w, r0, r1);
                    size_t done = 0;
                    while (done < buf.size()) {
                        ssize_t k = ::pwrite(out_fd, buf.data() + done, buf.size() - done, (off_t)(at[t] + done));
                        if (k <= 0) { ok = false; break; }
                        done += (size_t)k;
                    }
                });
            }
            if (host_trace) fprintf(stderr, "[host] SAM batch of %u reads: sizes %.2f  render + write (%s) %.2f ms\n", n,
                                    (t_counted - t1) * 1e3, mapped ? "mapped" : "pwrite", (now_s() - t_counted) * 1e3);
            if (!ok) { std::cerr << "[ERROR]\t\tWriting " << sam_file << " failed.\n"; failed.store(true); }
            file_pos = at[T];
            read_id += n;
            output_s += now_s() - t1;
        };
        work* w;
        while (ended < n_dev && done_q.pop(w)) {
            if (!w) { ended++; continue; }
            waiting[w->seq_no] = w;
            while (!waiting.empty() && waiting.begin()->first == next_seq) {
                work* x = waiting.begin()->second;
                waiting.erase(waiting.begin());
                if (!x->failed && !failed.load()) render(x);
                next_seq++;
                free_q.push(x);
            }
        }
    }
    for (auto& t : dev_threads) t.join();
    free_q.push(nullptr);
    reader.join();
    ::close(out_fd);
    if (failed.load()) return;

    // the reference's phase log (q_gram_mapper.h:548-555, bucket_locator.h:698-702,606-610), filled with device times
    double ms[BM_K_COUNT]; uint64_t launches[BM_K_COUNT];
    for (int i = 0; i < BM_K_COUNT; i++) {
        ms[i] = 0; launches[i] = 0;
        for (auto& d : devs) { double m1 = 0; uint64_t l1 = 0; bm_kernel_stats(d.c, i, &m1, &l1); ms[i] += m1; launches[i] += l1; }
    }
    for (auto& d : devs) bm_kernel_timing(d.c, 0);
    if (qm->num_records) qm->print_map_benchmark((float)((ms[BM_K_SAMPLE] + ms[BM_K_MAP] + ms[BM_K_MAPOVER]) * 1e-3));
    std::cerr << "[BENCHMARK]\tTotal time used for building k-mer index for each bucket: 0 s.\n";
    const float time = (float)((ms[BM_K_LOCSAMPLE] + ms[BM_K_LOCATE] + ms[BM_K_LOCSCAN] + ms[BM_K_MERGE]) * 1e-3);
    std::cerr << "[BENCHMARK]\tTotal time used for finding exact location of the sequences: " << time << " s ("
              << time * 1000 * 1000 / (read_id ? read_id : 1) << " μs/seq).\n";
    std::cerr << "[BENCHMARK]\tTotal mapped locations: " << mapped_locations << " (" << (float)mapped_locations / read_id
              << " per sequence).\n";
    const float es = (float)(ms[BM_K_ALIGN] * 1e-3 + output_s);
    std::cerr << "[BENCHMARK]\tTotal time used for alignment verification and output: " << es << " s ("
              << es / mapped_locations * 1000 * 1000 << " μs per pairwise alignment).\n";
    std::cerr << "[BENCHMARK]\tHost pipeline: FASTQ splitting + packing " << (float)parse_s << " s, device passes " << (float)device_s
              << " s on " << n_dev << " GPU(s), SAM rendering + writing " << (float)output_s << " s; " << host_threads()
              << " host threads; derived tables" << (n_dev > 1 ? " + NCCL replication" : "") << " " << (float)setup_s
              << " s (under the first batch's parse); wall " << (float)(now_s() - t_pipe) << " s.\n";
    std::cerr << "[BENCHMARK]\tDevice pipeline (H2D + kernels + D2H): " << (float)device_s << " s for " << read_id
              << " reads (" << (float)(read_id / (device_s > 0 ? device_s : 1)) << " reads/s).\n";
    _m->reset();                                                                       // :627
}

}   // namespace bm
