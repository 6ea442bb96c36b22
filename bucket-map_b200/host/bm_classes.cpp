// bm_classes.cpp -- bucket_hash_indexer and q_gram_mapper of the C++ host (the locator is in bm_locator.cpp).
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>

#include "bm_host.hpp"

namespace bm {

namespace fs = std::filesystem;

static double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// utils.h:104-144
static bool check_extension_in(fs::path const& dir, std::string const& ext) {
    if (!fs::create_directories(dir)) {
        for (const auto& entry : fs::directory_iterator(dir)) {
            if (entry.path().extension() == ext) {
                std::cerr << "[ERROR]\t\tThe file with extension " << ext << " already exists in directory: "
                          << entry.path() << ".\n";
                return false;
            }
        }
    }
    return true;
}
static bool check_filename_in(fs::path const& dir, std::string const& filename) {
    if (!fs::create_directories(dir)) {
        for (const auto& entry : fs::directory_iterator(dir)) {
            if (entry.path() == dir / filename) {
                std::cerr << "[ERROR]\t\tThe specified file already exists in directory: " << dir / filename << ".\n";
                return false;
            }
        }
    }
    return true;
}

// single-entry cache: the reference parses the FASTA twice per run (index build, locator); one parse is enough here
genome const* load_genome_cached(fs::path const& fasta, uint32_t bucket_len, uint32_t read_len) {
    static genome g;
    static std::string key;
    std::string k = fasta.string() + "|" + std::to_string(bucket_len) + "|" + std::to_string(read_len);
    if (k != key) {
        const auto t0 = std::chrono::steady_clock::now();
        if (!g.read_fasta(fasta)) {
            std::cerr << "[ERROR]\t\tCannot read the reference genome " << fasta << ".\n";
            key.clear();
            return nullptr;
        }
        g.bucketize(bucket_len, read_len);
        key = k;
        std::cerr << "[BENCHMARK]\tElapsed time for reading the reference genome (" << g.n_bases << " bases): "
                  << std::chrono::duration<float>(std::chrono::steady_clock::now() - t0).count() << " s.\n";
    }
    return &g;
}

// ---- bucket_hash_indexer (indexer/bucket_indexer.h:138-216) ----------------------------------------------------------------
bucket_hash_indexer::bucket_hash_indexer(unsigned int bucket_len, unsigned int read_len, uint8_t seed_len,
                                         std::function<std::size_t(std::size_t)> min_hash_function,
                                         unsigned int min_hash_threshold, context* ctx)
    : ctx_(ctx ? ctx : &context::default_context()), bucket_length(bucket_len), read_length(read_len), q(seed_len) {
    std::cerr << "[INFO]\t\tSet index seed length to be: " << (int)q << ".\n";
    int index = 0;
    const unsigned int n = 1u << (2 * q);
    kmer_to_index.reserve(n);
    for (unsigned int i = 0; i < n; i++) {                         // :147-157
        if (min_hash_function(i) <= min_hash_threshold) kmer_to_index.push_back(index++);
        else kmer_to_index.push_back(-1);
    }
    num_valid_kmers = (unsigned)index;
    std::cerr << "[INFO]\t\tNumber of remaining k-mers after FracMinHash is " << num_valid_kmers << " out of " << n
              << " (" << ((float)num_valid_kmers) / n * 100 << "%).\n";
}

unsigned int bucket_hash_indexer::index(fs::path const& fasta_file_name, fs::path const& index_directory,
                                        std::string const& indicator) {
    if (!check_extension_in(index_directory, EXTENSION) || !check_filename_in(index_directory, indicator + ".qgram"))
        return 0;
    const double t0 = now_s();
    genome const* g = load_genome_cached(fasta_file_name, bucket_length, read_length);
    if (!g) return 0;
    bm_ctx* c = ctx_->get();
    const uint32_t n_emit = (uint32_t)g->b_start.size();
    if (g->NB == 0) { std::cerr << "[ERROR]\t\tThe reference genome is empty.\n"; return 0; }
    int rc = bm_ref_upload(c, g->packed.data(), g->n_bases, g->b_start.data(), g->b_len.data(), n_emit, g->NB);
    if (rc) { std::cerr << "[ERROR]\t\t" << bm_last_error(c) << "\n"; return 0; }
    const uint32_t W = (g->NB + 7) >> 3;
    std::vector<uint8_t> rows((size_t)num_valid_kmers * W);
    if (num_valid_kmers == 0) {
        std::cerr << "[ERROR]\t\tThe q-gram index is empty. No file is created.\n";     // :81-84
    } else {
        rc = bm_index_build(c, kmer_to_index.data(), q, num_valid_kmers, rows.data());  // K1 (replaces :49-61)
        if (rc) { std::cerr << "[ERROR]\t\t" << bm_last_error(c) << "\n"; return 0; }
        if (check_filename_in(index_directory, indicator + ".qgram")) {                 // :76-100
            std::ofstream f(index_directory / (indicator + ".qgram"), std::ios::binary);
            f.write((const char*)rows.data(), (std::streamsize)rows.size());
            f.close();
            std::cerr << "[INFO]\t\tThe bucket q-gram index is stored in: " << index_directory / (indicator + ".qgram") << ".\n";
        }
    }
    bucket_id.clear();
    for (uint32_t b = 0; b < n_emit; b++) bucket_id.push_back(g->ids[g->b_rec[b]]);
    std::cerr << "[INFO]\t\tThe number of buckets: " << bucket_id.size() << ".\n";
    if (check_filename_in(index_directory, indicator + ".bucket_id")) {                 // :103-114
        std::ofstream f(index_directory / (indicator + ".bucket_id"));
        for (auto const& s : bucket_id) f << s << "\n";
        f.close();
        std::cerr << "[INFO]\t\tThe bucket ids are stored in: " << index_directory / (indicator + ".bucket_id") << ".\n";
    }
    if (check_filename_in(index_directory, indicator + ".kmers_index")) {               // :116-127
        std::string text;
        text.reserve(kmer_to_index.size() * 7);
        char tmp[16];
        for (int v : kmer_to_index) { int n = snprintf(tmp, sizeof tmp, "%d\n", v); text.append(tmp, (size_t)n); }
        std::ofstream f(index_directory / (indicator + ".kmers_index"), std::ios::binary);
        f.write(text.data(), (std::streamsize)text.size());
        f.close();
        std::cerr << "[INFO]\t\tThe kmer indexes are stored in: " << index_directory / (indicator + ".kmers_index") << ".\n";
    }
    std::cerr << "[BENCHMARK]\tElapsed time for creating and storing index files: " << (float)(now_s() - t0) << " s.\n";
    return (unsigned)bucket_id.size();
}

void bucket_hash_indexer::reset() {
    bucket_id.clear(); bucket_id.shrink_to_fit();
    kmer_to_index.clear(); kmer_to_index.shrink_to_fit();
}

// ---- q_gram_mapper (mapper/q_gram_mapper.h:281-557) -------------------------------------------------------------------------
q_gram_mapper::q_gram_mapper(unsigned int bucket_len, unsigned int read_len, uint8_t query_seed_length,
                             uint8_t index_seed_length, unsigned int samples, unsigned int fault, float distinguishability,
                             unsigned int quality_threshold, unsigned int num_candidate_buckets,
                             unsigned int num_segment_samples, context* ctx)
    : ctx_(ctx ? ctx : &context::default_context()), bucket_length(bucket_len), read_length(read_len),
      k(query_seed_length), q(index_seed_length), num_samples(samples), num_fault_tolerance(fault),
      allowed_max_candidate_buckets(num_candidate_buckets), num_segment_samples_(num_segment_samples),
      distinguishability_(distinguishability), avg_base_quality_(quality_threshold) {
    std::cerr << "[INFO]\t\tSet query seed length to be " << (int)k << ", and index seed length " << (int)q << ".\n";
#ifdef BM_BUCKET_NUM
    NB_ = BM_BUCKET_NUM;
#endif
}

void q_gram_mapper::fill_params(bm_params& p) const {
    memset(&p, 0, sizeof p);
    p.bucket_len = bucket_length; p.read_len = read_length; p.q = q; p.k = k;
    p.mapper_samples = num_samples; p.mapper_fault = num_fault_tolerance;
    p.distinguishability = distinguishability_; p.avg_base_quality = avg_base_quality_;
    p.max_candidates = allowed_max_candidate_buckets; p.num_segment_samples = num_segment_samples_;
}

void q_gram_mapper::set_num_buckets(uint32_t NB) { NB_ = NB; }

void q_gram_mapper::load(fs::path const& index_directory, const std::string& indicator) {
    if (loaded_) { std::cerr << "[ERROR]\t\tThe q-gram index is not empty. Terminating load.\n"; return; }   // :325-328
    if (NB_ == 0) {
        std::cerr << "[ERROR]\t\tThe number of buckets is unknown: call set_num_buckets() (bucket_locator::initialize "
                     "does) or build with -DBM_BUCKET_NUM.\n";
        return;
    }
    // .kmers_index (:331-342)
    unsigned int sampled_q_grams = 0;
    const size_t n_codes = (size_t)1 << (2 * q);
    kmer_to_index.clear();
    {
        std::ifstream f(index_directory / (indicator + ".kmers_index"), std::ios::binary);
        if (f) {
            std::string text((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
            const char* s = text.c_str();
            char* e = nullptr;
            kmer_to_index.reserve(n_codes);
            int last = 0;
            for (size_t i = 0; i < n_codes; i++) {
                long v = strtol(s, &e, 10);
                if (e == s) v = last; else last = (int)v;      // a short file repeats the last value, like operator>>
                s = e;
                if (v >= 0) sampled_q_grams++;
                kmer_to_index.push_back((int)v);
            }
            std::cerr << "[INFO]\t\tSuccessfully loaded " << index_directory / (indicator + ".kmers_index") << ".\n";
        } else {
            // Legacy index without a .kmers_index file (written before FracMinHash existed): the reference reads no table,
            // hence zero rows of the .qgram file (:331-357), and maps nothing -- a SAM with the header only.  Same here: no
            // q-gram has a row.
            kmer_to_index.assign(n_codes, -1);
        }
    }
    // .qgram (:344-366): the file is memory-mapped and handed to the device as it is
    const uint32_t W = (NB_ + 7) >> 3;
    file_map fm;
    if (fm.open(index_directory / (indicator + ".qgram")) && kmer_to_index.size() == n_codes) {
        const double t0 = now_s();
        const size_t need = (size_t)sampled_q_grams * W;
        std::vector<uint8_t> padded;
        const uint8_t* rows = (const uint8_t*)fm.data();
        if (fm.size() < need) {                                   // short file: missing rows read as zero, like the stream would
            padded.assign(need, 0);
            if (fm.size()) memcpy(padded.data(), fm.data(), fm.size());
            rows = padded.data();
        }
        bm_ctx* c = ctx_->get();
        int rc = bm_index_load(c, kmer_to_index.data(), q, rows, sampled_q_grams, NB_);
        uint32_t valid = 0;
        if (!rc) rc = bm_index_zeros(c, distinguishability_, nullptr, &valid);            // K0 (:171-187)
        if (rc) std::cerr << "[ERROR]\t\t" << bm_last_error(c) << "\n";
        else {
            loaded_ = true;
            const unsigned thr = (unsigned int)(distinguishability_ * NB_);
            std::cerr << "[BENCHMARK]\tNumber of Q-grams with distinguishability >= " << ((float)thr) / NB_ << ": " << valid
                      << " (" << ((float)valid) / sampled_q_grams * 100 << "%).\n";
            std::cerr << "[BENCHMARK]\tElapsed time for loading index files: " << (float)(now_s() - t0) << " s.\n";
            std::cerr << "[INFO]\t\tSuccessfully loaded " << index_directory / (indicator + ".qgram") << ".\n";
        }
    }
}

void q_gram_mapper::print_map_benchmark(float time) const {
    std::cerr << "[BENCHMARK]\tElapsed time for bucket mapping: " << time << " s (" << time * 1000 * 1000 / num_records
              << " μs/seq).\n";
    std::cerr << "[BENCHMARK]\tNumber of reads that have at least one candidate bucket: " << mapped_reads << "  ("
              << ((float)mapped_reads) / num_records * 100 << "%).\n";
    std::cerr << "[BENCHMARK]\tAverage number of buckets an original read is mapped to: "
              << ((float)num_buckets_orig) / mapped_reads << ".\n";
    std::cerr << "[BENCHMARK]\tAverage number of buckets a reverse complement of the read is mapped to: "
              << ((float)num_buckets_rev_comp) / mapped_reads << ".\n";
}

std::pair<segments_t, segments_t> q_gram_mapper::map(fs::path const& sequence_file) {
    segments_t res_orig(NB_), res_rev_comp(NB_);
    if (!loaded_) {
        std::cerr << "[ERROR]\t\tThe q-gram index is empty. Cannot accept query.\n";       // :389-393
        return std::make_pair(res_orig, res_rev_comp);
    }
    fastq_reader fin(sequence_file);
    if (!fin.ok()) { std::cerr << "[ERROR]\t\tCannot read " << sequence_file << ".\n"; return std::make_pair(res_orig, res_rev_comp); }
    bm_params p;
    fill_params(p);
    p.locator_samples = 1;
    bm_ctx* c = ctx_->get();
    read_batch b;
    std::vector<uint32_t> cand_off, cand, seg_read, seg_start, seg_len;
    const double t0 = now_s();
    while (fin.next(b, 1u << 20, (uint64_t)1 << 30, p.k, p.avg_base_quality)) {
        const uint32_t n = b.size();
        const size_t nseg_max = (size_t)n * (num_segment_samples_ ? num_segment_samples_ : 1);
        cand_off.resize(2 * nseg_max + 1);
        cand.resize(2 * nseg_max * allowed_max_candidate_buckets + 1);
        uint64_t stats[8] = {0};
        int rc = bm_reads_upload_packed(c, &p, b.planes.as<uint64_t>(), b.all_pass ? nullptr : b.qk.as<uint32_t>(),
                                        b.pl_off.as<uint32_t>(), b.read_off.as<uint64_t>(), n);
        if (!rc) rc = bm_map_device(c, &p);                                                 // K2f (or K2a + K2)
        if (!rc) rc = bm_map_download(c, cand_off.data(), cand.data(), cand.size(), stats);
        if (rc) { std::cerr << "[ERROR]\t\t" << bm_last_error(c) << "\n"; break; }
        uint32_t n_seg = 0;
        bm_reads_segments(c, &n_seg, nullptr, nullptr, nullptr);
        seg_read.resize(n_seg); seg_start.resize(n_seg); seg_len.resize(n_seg);
        bm_reads_segments(c, &n_seg, seg_read.data(), seg_start.data(), seg_len.data());
        std::vector<uint8_t> mapped(n, 0);
        for (uint32_t s = 0; s < n_seg; s++) {                                              // :527-540
            segment_info_t seg{num_records + seg_read[s], (int)seg_start[s]};
            for (uint32_t i = cand_off[2 * s]; i < cand_off[2 * s + 1]; i++) res_orig[cand[i]].push_back(seg);
            for (uint32_t i = cand_off[2 * s + 1]; i < cand_off[2 * s + 2]; i++) res_rev_comp[cand[i]].push_back(seg);
            const uint32_t no = cand_off[2 * s + 1] - cand_off[2 * s], nr = cand_off[2 * s + 2] - cand_off[2 * s + 1];
            if (no || nr) { mapped[seg_read[s]] = 1; num_buckets_orig += no; num_buckets_rev_comp += nr; }
        }
        for (uint32_t i = 0; i < n; i++) mapped_reads += mapped[i];
        num_records += n;
    }
    print_map_benchmark((float)(now_s() - t0));
    return std::make_pair(res_orig, res_rev_comp);
}

void q_gram_mapper::reset() {
    bm_index_free(ctx_->get());
    loaded_ = false;
}

}   // namespace bm
