// bm_host.hpp -- C++ host side of the B200-native BucketMap hot path.
//
// Mirrors the reference's three class interfaces (paths relative to /root/reference/bucket_map/):
//   indexer / bucket_hash_indexer   indexer/indexer.h:8-32, indexer/bucket_indexer.h:21-230, bucket_hash_indexer.h:8-46
//   mapper  / q_gram_mapper         mapper/mapper.h:4-34, mapper/q_gram_mapper.h:200-648
//   locator / bucket_locator        locator/locator.h:9-50, locator/bucket_locator.h:106-705
// with the same member names, argument meaning and error behaviour ([ERROR] line on stderr + early return, no
// exceptions), minus the NUM_BUCKETS template parameter: the bucket count is computed from the FASTA at run time with
// the formula the reference's CMake bakes in (CMakeLists.txt:13-50).  All computation goes through the C-ABI of
// include/bucketmap_b200.h (hand-written sm_100a kernels); there is no CPU implementation of the path here, and
// construction fails loudly when no CUDA device is usable.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <filesystem>
#include <functional>
#include <string>
#include <utility>
#include <vector>

#include "bucketmap_b200.h"

namespace bm {

// utils.h:309-311
typedef std::pair<unsigned int, int> segment_info_t;
typedef std::vector<std::vector<segment_info_t>> segments_t;

// ---- one GPU ---------------------------------------------------------------------------------------------------------
class context {
public:
    explicit context(int device);
    ~context();
    context(context const&) = delete;
    context& operator=(context const&) = delete;
    bm_ctx* get() const { return c_; }
    int device() const { return device_; }
    // process-wide default: device $BM_DEVICE (0 if unset); exits with an [ERROR] line if there is no CUDA device
    static context& default_context();

private:
    bm_ctx* c_ = nullptr;
    int device_ = 0;
};

// ---- reference genome: FASTA -> 2-bit packed bases + bucket table (utils.h:60-102) -------------------------------------------
struct genome {
    std::vector<std::string> ids;            // full FASTA header lines (after '>')
    std::vector<uint64_t> rec_len;
    std::vector<uint8_t> packed;             // base i in byte i>>2 at bits 2*(i&3); A0 C1 G2 T3 (other letters as dna5->dna4)
    uint64_t n_bases = 0;
    // bucket table
    uint32_t NB = 0;                         // sum ceil(len/L): includes skipped ("phantom") buckets
    std::vector<uint64_t> b_start;
    std::vector<uint32_t> b_len, b_rec;

    bool read_fasta(std::filesystem::path const& p);
    void bucketize(uint32_t bucket_len, uint32_t read_len);
};

// ---- reads: FASTQ -> batches of concatenated bases / qualities -------------------------------------------------------------------
// growable byte array in pinned host memory (bm_host_alloc): the copy engine reads it directly, so the host->device copy
// of a batch overlaps the mapper kernels inside bm_map_locate
class pinned_bytes {
public:
    pinned_bytes() = default;
    pinned_bytes(pinned_bytes const&) = delete;
    pinned_bytes& operator=(pinned_bytes const&) = delete;
    ~pinned_bytes() { if (p_) bm_host_free(p_); }
    uint8_t* data() { return p_; }
    const uint8_t* data() const { return p_; }
    size_t size() const { return n_; }
    void clear() { n_ = 0; }
    void truncate(size_t n) { if (n < n_) n_ = n; }
    uint8_t& operator[](size_t i) { return p_[i]; }
    const uint8_t& operator[](size_t i) const { return p_[i]; }
    void reserve(size_t cap) {
        if (cap <= cap_) return;
        size_t want = cap_ ? cap_ : (size_t)1 << 20;
        while (want < cap) want *= 2;
        uint8_t* q = (uint8_t*)bm_host_alloc(want);
        if (!q) { fprintf(stderr, "[ERROR]\t\tOut of pinned host memory (%zu bytes).\n", want); std::exit(2); }
        if (n_) memcpy(q, p_, n_);
        if (p_) bm_host_free(p_);
        p_ = q; cap_ = want;
    }
    void append(const void* src, size_t len) { reserve(n_ + len); memcpy(p_ + n_, src, len); n_ += len; }
    void append_fill(uint8_t v, size_t len) { reserve(n_ + len); memset(p_ + n_, v, len); n_ += len; }

private:
    uint8_t* p_ = nullptr;
    size_t n_ = 0, cap_ = 0;
};

struct read_batch {
    pinned_bytes seq, qual;                  // ASCII, concatenated
    std::vector<uint64_t> off{0};
    std::vector<char> id_pool;               // id lines (after '@'), concatenated
    std::vector<uint64_t> id_off{0};
    uint32_t size() const { return (uint32_t)off.size() - 1; }
    const char* id(uint32_t i) const { return id_pool.data() + id_off[i]; }
    size_t id_len(uint32_t i) const { return (size_t)(id_off[i + 1] - id_off[i]); }
    void clear() { seq.clear(); qual.clear(); off.assign(1, 0); id_pool.clear(); id_off.assign(1, 0); }
};

class fastq_reader {
public:
    explicit fastq_reader(std::filesystem::path const& p);
    ~fastq_reader();
    bool ok() const { return f_ != nullptr; }
    // appends up to max_reads reads (or until max_bases bases) to `b` (cleared first); false when the file is exhausted
    bool next(read_batch& b, uint32_t max_reads, uint64_t max_bases);

private:
    // next line as a view into the file buffer (no copy); false at end of file
    bool line(const char*& s, size_t& len);
    void* f_ = nullptr;
    std::vector<char> buf_;
    size_t pos_ = 0, end_ = 0;
    bool eof_ = false;
};

// FASTA parsed and bucketized once per (path, L, r); nullptr (after an [ERROR] line) if it cannot be read
genome const* load_genome_cached(std::filesystem::path const& fasta, uint32_t bucket_len, uint32_t read_len);

// tools/hash_function_generator.h:9-119; seed_x/seed_y >= 0 pin the two rand() draws (the reference seeds with time())
std::function<std::size_t(std::size_t)> generate_hash_function(std::size_t hash_table_size, long seed_x = -1, long seed_y = -1);

// ---- indexer -------------------------------------------------------------------------------------------------------------------
class indexer {
public:
    indexer() {}
    virtual ~indexer() = default;
    virtual unsigned int index(std::filesystem::path const& fasta_file_name, std::filesystem::path const& index_directory,
                               std::string const& indicator) = 0;
    virtual void reset() = 0;
};

class bucket_hash_indexer : public indexer {
public:
    bucket_hash_indexer(unsigned int bucket_len, unsigned int read_len, uint8_t seed_len,
                        std::function<std::size_t(std::size_t)> min_hash_function, unsigned int min_hash_threshold,
                        context* ctx = nullptr);
    unsigned int index(std::filesystem::path const& fasta_file_name, std::filesystem::path const& index_directory,
                       std::string const& indicator) override;
    void reset() override;

protected:
    context* ctx_;
    std::vector<std::string> bucket_id;
    unsigned int bucket_length, read_length;
    std::string EXTENSION = ".bhi";
    uint8_t q;
    std::vector<int> kmer_to_index;
    unsigned int num_valid_kmers = 0;
};

// ---- mapper --------------------------------------------------------------------------------------------------------------------
class mapper {
public:
    mapper() {}
    virtual ~mapper() = default;
    unsigned int num_records = 0;
    virtual void load(std::filesystem::path const& index_directory, const std::string& indicator) = 0;
    virtual std::pair<segments_t, segments_t> map(std::filesystem::path const& sequence_file) = 0;
    virtual void reset() = 0;
};

class q_gram_mapper : public mapper {
public:
    q_gram_mapper(unsigned int bucket_len, unsigned int read_len, uint8_t query_seed_length, uint8_t index_seed_length,
                  unsigned int samples, unsigned int fault, float distinguishability, unsigned int quality_threshold = 35,
                  unsigned int num_candidate_buckets = 30, unsigned int num_segment_samples = 5, context* ctx = nullptr);
    void load(std::filesystem::path const& index_directory, const std::string& indicator) override;
    std::pair<segments_t, segments_t> map(std::filesystem::path const& sequence_file) override;
    void reset() override;

    // what bucket_locator needs to run the fused device pipeline
    bool loaded() const { return loaded_; }
    uint32_t num_buckets() const { return NB_; }
    // NUM_BUCKETS of the reference's template (CMakeLists.txt:13-50); must be set before load()
    void set_num_buckets(uint32_t NB);
    unsigned int read_len() const { return read_length; }
    void fill_params(bm_params& p) const;
    context* ctx() const { return ctx_; }
    // the counters map() logs (q_gram_mapper.h:548-555); bucket_locator adds to them when it maps on the device
    unsigned int mapped_reads = 0, num_buckets_orig = 0, num_buckets_rev_comp = 0;
    void print_map_benchmark(float seconds) const;

private:
    context* ctx_;
    unsigned int bucket_length, read_length;
    uint8_t k, q;
    unsigned int num_samples, num_fault_tolerance, allowed_max_candidate_buckets, num_segment_samples_;
    float distinguishability_;
    unsigned int avg_base_quality_;
    std::vector<int> kmer_to_index;
    bool loaded_ = false;
    uint32_t NB_ = 0;
};

// ---- locator -------------------------------------------------------------------------------------------------------------------
class locator {
public:
    explicit locator(indexer* ind) : _i(ind) {}
    virtual ~locator() = default;
    unsigned int initialize(std::filesystem::path const& fasta_file_name, std::filesystem::path const& index_directory,
                            std::string const& indicator);
    virtual void locate(std::filesystem::path const& sequence_file, std::filesystem::path const& index_path,
                        std::filesystem::path const& sam_file, unsigned int quality_threshold = 30) {}

private:
    indexer* _i;
};

class bucket_locator : public locator {
public:
    bucket_locator(indexer* ind, mapper* map, unsigned int bucket_len, unsigned int read_len, uint8_t seed_len,
                   float mismatch_rate, float indel_rate, unsigned int sample_size, unsigned int quality_threshold,
                   unsigned int num_segment_samples = 5, bool align = false);
    unsigned int initialize(std::filesystem::path const& fasta_file_name, std::filesystem::path const& index_directory,
                            std::string const& indicator);
    void locate(std::filesystem::path const& sequence_file, std::filesystem::path const& index_file,
                std::filesystem::path const& sam_file, unsigned int quality_threshold = 30) override;
    // reads per device batch (default 1 Mi) -- host memory / latency knob, no effect on results
    void set_batch_reads(uint32_t n) { batch_reads_ = n ? n : 1; }

private:
    mapper* _m;
    unsigned int bucket_length, read_length;
    uint8_t k;
    float mismatch_rate_, indel_rate_;
    unsigned int sample_size_, avg_base_quality_, num_segment_samples_;
    bool align_;
    std::filesystem::path genome_file_name;
    uint32_t batch_reads_ = 1u << 18;   // reads per device batch: small enough for the parse / device / SAM stages to overlap
};

}   // namespace bm
