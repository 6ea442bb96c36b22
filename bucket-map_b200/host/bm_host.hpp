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
#include <memory>
#include <string>
#include <thread>
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
    // process-wide default: device $BM_DEVICE (0 if unset); exits with an [ERROR] line if there is no CUDA device.
    // Thread-safe: main() brings it up on a side thread while the FASTA is parsed.
    static context& default_context();

private:
    bm_ctx* c_ = nullptr;
    int device_ = 0;
};

// ---- reference genome: FASTA -> 2-bit packed bases + bucket table (utils.h:60-102) -------------------------------------------
// zero-initialised byte array from anonymous pages (first touched by whoever writes them: the FASTA packer's threads)
class page_buffer {
public:
    page_buffer() = default;
    page_buffer(page_buffer const&) = delete;
    page_buffer& operator=(page_buffer const&) = delete;
    ~page_buffer() { release(); }
    bool allocate(size_t n);
    void release();
    uint8_t* data() { return p_; }
    const uint8_t* data() const { return p_; }
    size_t size() const { return n_; }

private:
    uint8_t* p_ = nullptr;
    size_t n_ = 0;
};

// read-only memory map of a whole file
class file_map {
public:
    file_map() = default;
    file_map(file_map const&) = delete;
    file_map& operator=(file_map const&) = delete;
    ~file_map() { close(); }
    bool open(std::filesystem::path const& p);
    void close();
    const char* data() const { return p_; }
    size_t size() const { return n_; }
    bool ok() const { return ok_; }

private:
    const char* p_ = nullptr;
    size_t n_ = 0;
    bool ok_ = false;
};

// worker threads of the host pipeline (FASTA packing, FASTQ packing, SAM rendering): --threads / $BM_THREADS, default
// min(hardware threads, 16)
unsigned host_threads();
void set_host_threads(unsigned n);
// fn(part, parts) on `parts` threads (the calling thread takes part 0)
void parallel_parts(unsigned parts, std::function<void(unsigned, unsigned)> const& fn);

struct genome {
    std::vector<std::string> ids;            // full FASTA header lines (after '>')
    std::vector<uint64_t> rec_len;
    page_buffer packed;                      // base i in byte i>>2 at bits 2*(i&3); A0 C1 G2 T3 (other letters as dna5->dna4)
    uint64_t n_bases = 0;
    // bucket table
    uint32_t NB = 0;                         // sum ceil(len/L): includes skipped ("phantom") buckets
    std::vector<uint64_t> b_start;
    std::vector<uint32_t> b_len, b_rec;

    bool read_fasta(std::filesystem::path const& p);
    void bucketize(uint32_t bucket_len, uint32_t read_len);
};

// ---- reads: FASTQ -> batches in the device's packed form -------------------------------------------------------------------------
// growable array in pinned host memory (bm_host_alloc): the copy engine reads it directly, so the host->device copy
// of a batch overlaps the mapper kernels inside bm_map_locate_packed
class pinned_bytes {
public:
    pinned_bytes() = default;
    pinned_bytes(pinned_bytes const&) = delete;
    pinned_bytes& operator=(pinned_bytes const&) = delete;
    ~pinned_bytes() { if (p_) bm_host_free(p_); }
    uint8_t* data() { return p_; }
    const uint8_t* data() const { return p_; }
    template <class T> T* as() { return reinterpret_cast<T*>(p_); }
    size_t capacity() const { return cap_; }
    void reserve(size_t cap) {                   // contents are NOT kept: every batch is written from scratch
        if (cap <= cap_) return;
        size_t want = cap_ ? cap_ : (size_t)1 << 20;
        while (want < cap) want *= 2;
        uint8_t* q = (uint8_t*)bm_host_alloc(want);
        if (!q) { fprintf(stderr, "[ERROR]\t\tOut of pinned host memory (%zu bytes).\n", want); std::exit(2); }
        if (p_) bm_host_free(p_);
        p_ = q; cap_ = want;
    }

private:
    uint8_t* p_ = nullptr;
    size_t cap_ = 0;
};

// One batch of FASTQ records.  The text stays where it is (views into the memory-mapped file: SAM rendering copies
// SEQ / QUAL from there); what travels to the device is written by the parser threads in the packed form of
// include/bucketmap_b200.h (bm_pack_reads): 2-bit base planes + the k-mer quality plane when some window fails -b.
struct read_batch {
    std::vector<const char*> id, seq, qual;     // start of the id line (after '@'), of the bases, of the qualities
    std::vector<uint32_t> id_len, qual_len;
    std::vector<uint64_t> off{0};                // bases: read i = [off[i], off[i+1])
    pinned_bytes planes, qk, pl_off, read_off;   // uint64 / uint32 / uint32 / uint64 arrays for bm_map_locate_packed
    bool all_pass = true;                        // no k-mer window below -b: the quality plane need not travel
    uint32_t size() const { return (uint32_t)off.size() - 1; }
    void clear() { id.clear(); seq.clear(); qual.clear(); id_len.clear(); qual_len.clear(); off.assign(1, 0); all_pass = true; }
};

class fastq_reader {
public:
    explicit fastq_reader(std::filesystem::path const& p);
    bool ok() const { return map_.ok(); }
    // Next batch of up to max_reads reads (or max_bases bases): splits the records (line ends found by the host threads, one
    // window of the file at a time; the walk over the lines is serial and copies nothing), then packs them for query seed
    // length k and quality threshold b on the host threads.  false when the file is exhausted.
    bool next(read_batch& b, uint32_t max_reads, uint64_t max_bases, uint32_t k, uint32_t bq);

private:
    bool line(const char*& s, size_t& len);
    void index_window();
    file_map map_;
    size_t pos_ = 0;
    std::vector<uint64_t> nl_;      // positions of the line feeds in [.., win_hi_), ascending; nl_i_ = the next one to use
    size_t nl_i_ = 0, win_hi_ = 0;
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
    // several GPUs (--devices): the batches of one FASTQ are spread over the devices, the index is replicated from the
    // mapper's context by bm_bcast_ref / bm_bcast_index (NCCL), SAM records leave in read order
    void set_devices(std::vector<int> const& devices) { devices_ = devices; }

private:
    mapper* _m;
    unsigned int bucket_length, read_length;
    uint8_t k;
    float mismatch_rate_, indel_rate_;
    unsigned int sample_size_, avg_base_quality_, num_segment_samples_;
    bool align_;
    std::filesystem::path genome_file_name;
    uint32_t batch_reads_ = 1u << 18;   // reads per device batch: small enough for the parse / device / SAM stages to overlap
    std::vector<int> devices_;
};

}   // namespace bm
