// bm_io.cpp -- device context wrapper, FASTA / FASTQ readers, FracMinHash hash function of the C++ host.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>

#include "bm_host.hpp"

namespace bm {

// ---- context -------------------------------------------------------------------------------------------------------------
context::context(int device) : device_(device) {
    int rc = bm_ctx_create(device, &c_);
    if (rc != BM_OK) {
        fprintf(stderr, "[ERROR]\t\tNo usable CUDA device %d (bm_ctx_create returned %d). This build has no CPU path.\n",
                device, rc);
        std::exit(2);
    }
    // device-memory options: both on by default, results identical either way
    if (const char* e = std::getenv("BM_KMER_TABLE")) bm_index_kmer_table(c_, std::atoi(e), 0);
    if (const char* e = std::getenv("BM_LOCATOR_INDEX")) bm_locator_index(c_, std::atoi(e) ? 1 : 0);
}
context::~context() { bm_ctx_destroy(c_); }

context& context::default_context() {
    static context* c = nullptr;
    if (!c) {
        const char* e = std::getenv("BM_DEVICE");
        c = new context(e ? std::atoi(e) : 0);
    }
    return *c;
}

// ---- FASTA ---------------------------------------------------------------------------------------------------------------
// Letters other than ACGT follow the reference's reader: FASTA is read as dna5 (anything that is not A/C/G/T/U
// becomes N) and converted to dna4 (N -> A) at utils.h:91.
static inline uint8_t rank_fasta(unsigned char c) {
    switch (c & 0xDF) {
        case 'C': return 1;
        case 'G': return 2;
        case 'T': case 'U': return 3;
        default: return 0;
    }
}

bool genome::read_fasta(std::filesystem::path const& p) {
    FILE* f = fopen(p.c_str(), "rb");
    if (!f) return false;
    ids.clear(); rec_len.clear(); packed.clear(); n_bases = 0;
    std::error_code ec;
    auto fsz = std::filesystem::file_size(p, ec);
    if (!ec) packed.reserve((size_t)fsz / 4 + 64);
    std::vector<char> buf(1 << 22);
    bool in_header = false, line_start = true;
    std::string header;
    uint8_t cur = 0;
    size_t n;
    while ((n = fread(buf.data(), 1, buf.size(), f)) > 0) {
        for (size_t i = 0; i < n; i++) {
            const char c = buf[i];
            if (in_header) {
                if (c == '\n') {
                    if (!header.empty() && header.back() == '\r') header.pop_back();
                    ids.push_back(header); rec_len.push_back(0); header.clear();
                    in_header = false; line_start = true;
                } else header.push_back(c);
                continue;
            }
            if (c == '\n') { line_start = true; continue; }
            if (line_start && c == '>') { in_header = true; line_start = false; continue; }
            line_start = false;
            if (c == '\r' || c == ' ' || c == '\t' || ids.empty()) continue;
            cur |= (uint8_t)(rank_fasta((unsigned char)c) << ((n_bases & 3) << 1));
            n_bases++;
            rec_len.back()++;
            if ((n_bases & 3) == 0) { packed.push_back(cur); cur = 0; }
        }
    }
    if (in_header) { ids.push_back(header); rec_len.push_back(0); }
    if (n_bases & 3) packed.push_back(cur);
    packed.resize(packed.size() + 64, 0);
    fclose(f);
    return true;
}

void genome::bucketize(uint32_t bucket_len, uint32_t read_len) {
    b_start.clear(); b_len.clear(); b_rec.clear();
    uint64_t nb_total = 0, base = 0;
    for (size_t r = 0; r < rec_len.size(); r++) {
        const uint64_t len = rec_len[r];
        if (len) nb_total += (len + bucket_len - 1) / bucket_len;          // CMakeLists.txt:13-50 (awk)
        float total_length = (float)len;                                      // utils.h:74-75: float division
        int num_buckets = (int)std::ceil(total_length / (int)bucket_len);
        for (int i = 0; i < num_buckets; i++) {
            int64_t start = (int64_t)i * bucket_len, end = start + bucket_len + read_len;
            if (end > (int64_t)len) end = (int64_t)len;
            if (end - start <= (int64_t)read_len) continue;                   // utils.h:88-90
            b_start.push_back(base + (uint64_t)start);
            b_len.push_back((uint32_t)(end - start));
            b_rec.push_back((uint32_t)r);
        }
        base += len;
    }
    NB = (uint32_t)nb_total;
}

// ---- FASTQ ---------------------------------------------------------------------------------------------------------------
fastq_reader::fastq_reader(std::filesystem::path const& p) : buf_(1 << 24) { f_ = fopen(p.c_str(), "rb"); }
fastq_reader::~fastq_reader() { if (f_) fclose((FILE*)f_); }

bool fastq_reader::getline(std::string& s) {
    s.clear();
    while (true) {
        if (pos_ == end_) {
            if (eof_) return !s.empty();
            end_ = fread(buf_.data(), 1, buf_.size(), (FILE*)f_);
            pos_ = 0;
            if (end_ == 0) { eof_ = true; return !s.empty(); }
        }
        const char* b = buf_.data() + pos_;
        const char* nl = (const char*)memchr(b, '\n', end_ - pos_);
        if (nl) {
            s.append(b, nl - b);
            pos_ += (size_t)(nl - b) + 1;
            if (!s.empty() && s.back() == '\r') s.pop_back();
            return true;
        }
        s.append(b, end_ - pos_);
        pos_ = end_;
    }
}

bool fastq_reader::next(read_batch& b, uint32_t max_reads, uint64_t max_bases) {
    b.clear();
    if (!f_) return false;
    std::string id, s, plus, q;
    while (b.size() < max_reads && b.seq.size() < max_bases) {
        do { if (!getline(id)) return b.size() > 0; } while (id.empty());
        if (!getline(s) || !getline(plus) || !getline(q)) break;
        b.ids.emplace_back(id, 1, std::string::npos);                 // full id line after '@' (seqan3 default)
        b.seq.insert(b.seq.end(), s.begin(), s.end());
        q.resize(s.size(), '!');
        b.qual.insert(b.qual.end(), q.begin(), q.end());
        b.off.push_back(b.seq.size());
    }
    return b.size() > 0;
}

// ---- FracMinHash hash function (tools/hash_function_generator.h:25-117) ---------------------------------------------------------
static const std::size_t kPrimeSizes[] = {5ul, 11ul, 23ul, 47ul, 97ul, 199ul, 409ul, 823ul, 1741ul, 3469ul, 6949ul, 14033ul,
                                          28411ul, 57557ul, 116731ul, 236897ul, 480881ul, 976369ul, 1982627ul, 4026031ul,
                                          8175383ul, 16601593ul, 33712729ul, 68460391ul, 139022417ul, 282312799ul,
                                          573292817ul, 1164186217ul, 2364114217ul, 4294967291ul};

std::function<std::size_t(std::size_t)> generate_hash_function(std::size_t hash_table_size, long seed_x, long seed_y) {
    std::size_t p = 0;
    for (std::size_t c : kPrimeSizes) if (c > 10 * hash_table_size) { p = c; break; }
    if (seed_x < 0 || seed_y < 0) std::srand((unsigned)std::time(nullptr));
    std::size_t x = seed_x >= 0 ? (std::size_t)seed_x : (std::size_t)std::rand() % (p - 1) + 1;
    std::size_t y = seed_y >= 0 ? (std::size_t)seed_y : (std::size_t)std::rand() % p;
    return [x, y, p, hash_table_size](int input) { return (x * input + y) % p % hash_table_size; };
}

}   // namespace bm
