// bm_io.cpp -- device context wrapper, FASTA / FASTQ readers, FracMinHash hash function of the C++ host.
#include <algorithm>
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

// Line-oriented: memchr finds the end of the line, a sequence line is packed four bases per byte through a 256-entry
// table (whitespace inside a line falls back to the per-character path).  ~1 GB/s instead of a per-byte state machine.
bool genome::read_fasta(std::filesystem::path const& p) {
    FILE* f = fopen(p.c_str(), "rb");
    if (!f) return false;
    ids.clear(); rec_len.clear(); packed.clear(); n_bases = 0;
    std::error_code ec;
    auto fsz = std::filesystem::file_size(p, ec);
    packed.assign((ec ? (size_t)1 << 20 : (size_t)fsz / 4) + 128, 0);
    static uint8_t lut[256];
    static bool lut_ready = false;
    if (!lut_ready) {
        for (int c = 0; c < 256; c++) lut[c] = rank_fasta((unsigned char)c);
        lut[(unsigned char)'\r'] = lut[(unsigned char)' '] = lut[(unsigned char)'\t'] = 0xFF;   // skipped
        lut_ready = true;
    }
    std::vector<char> buf(1 << 24);
    bool in_header = false, line_start = true;
    std::string header;
    size_t n;
    auto put = [&](uint8_t r) {
        if ((n_bases >> 2) + 1 >= packed.size()) packed.resize(packed.size() * 2, 0);
        packed[n_bases >> 2] |= (uint8_t)(r << ((n_bases & 3) << 1));
        n_bases++;
    };
    while ((n = fread(buf.data(), 1, buf.size(), f)) > 0) {
        size_t i = 0;
        while (i < n) {
            const char* nl = (const char*)memchr(buf.data() + i, '\n', n - i);
            const size_t e = nl ? (size_t)(nl - buf.data()) : n;          // segment [i, e) lies inside one line
            if (in_header) header.append(buf.data() + i, e - i);
            else if (line_start && e > i && buf[i] == '>') { in_header = true; header.assign(buf.data() + i + 1, e - i - 1); }
            else if (!ids.empty() && e > i) {
                const unsigned char* s = (const unsigned char*)buf.data();
                const uint64_t before = n_bases;
                size_t t = i;
                while (t < e) {
                    if ((n_bases & 3) == 0 && t + 4 <= e) {
                        if ((n_bases >> 2) + 2 >= packed.size()) packed.resize(packed.size() * 2, 0);
                        const uint8_t v0 = lut[s[t]], v1 = lut[s[t + 1]], v2 = lut[s[t + 2]], v3 = lut[s[t + 3]];
                        if (!((v0 | v1 | v2 | v3) & 0x80)) {
                            packed[n_bases >> 2] = (uint8_t)(v0 | (v1 << 2) | (v2 << 4) | (v3 << 6));
                            n_bases += 4; t += 4;
                            continue;
                        }
                    }
                    const uint8_t v = lut[s[t++]];
                    if (v != 0xFF) put(v);
                }
                rec_len.back() += n_bases - before;
            }
            if (e > i || nl) line_start = false;
            if (nl) {
                if (in_header) {
                    if (!header.empty() && header.back() == '\r') header.pop_back();
                    ids.push_back(header); rec_len.push_back(0); header.clear();
                    in_header = false;
                }
                line_start = true;
                i = e + 1;
            } else i = e;
        }
    }
    if (in_header) { ids.push_back(header); rec_len.push_back(0); }
    packed.resize(((size_t)n_bases + 3) / 4 + 64, 0);
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

// Lines are handed out as views into the 16 MB file buffer; when a line is cut by the end of the buffer the unread tail
// is moved to the front and the buffer refilled (grown if a single line is longer than the buffer).
bool fastq_reader::line(const char*& s, size_t& len) {
    while (true) {
        const char* b = buf_.data() + pos_;
        const char* nl = (end_ > pos_) ? (const char*)memchr(b, '\n', end_ - pos_) : nullptr;
        if (nl) {
            s = b; len = (size_t)(nl - b);
            pos_ += len + 1;
            if (len && s[len - 1] == '\r') len--;
            return true;
        }
        if (eof_) {
            if (end_ == pos_) return false;
            s = b; len = end_ - pos_;                        // last line without a newline
            pos_ = end_;
            if (len && s[len - 1] == '\r') len--;
            return true;
        }
        const size_t rest = end_ - pos_;
        if (rest && pos_) memmove(buf_.data(), buf_.data() + pos_, rest);
        pos_ = 0; end_ = rest;
        if (end_ == buf_.size()) buf_.resize(buf_.size() * 2);
        const size_t got = fread(buf_.data() + end_, 1, buf_.size() - end_, (FILE*)f_);
        if (got == 0) eof_ = true;
        end_ += got;
    }
}

bool fastq_reader::next(read_batch& b, uint32_t max_reads, uint64_t max_bases) {
    b.clear();
    if (!f_) return false;
    const char *id, *s, *plus, *q;
    size_t nid, ns, np, nq;
    while (b.size() < max_reads && b.seq.size() < max_bases) {
        do { if (!line(id, nid)) return b.size() > 0; } while (nid == 0);
        // a view dies with the next refill: copy each line before asking for the next one
        const size_t id_mark = b.id_pool.size(), seq_mark = b.seq.size();
        b.id_pool.insert(b.id_pool.end(), id + 1, id + nid);     // full id line after '@' (seqan3 default)
        bool ok = line(s, ns);
        if (ok && b.size() == 0) {
            // pinned allocations are slow (and growing means copying): size the batch from its first read
            const uint64_t want = std::min<uint64_t>(max_bases + ns, (uint64_t)(ns + 8) * max_reads);
            b.seq.reserve((size_t)want); b.qual.reserve((size_t)want);
        }
        if (ok) { b.seq.append(s, ns); ok = line(plus, np) && line(q, nq); }
        if (!ok) { b.id_pool.resize(id_mark); b.seq.truncate(seq_mark); break; }   // truncated record: dropped
        if (nq >= ns) b.qual.append(q, ns);
        else { b.qual.append(q, nq); b.qual.append_fill('!', ns - nq); }
        b.id_off.push_back(b.id_pool.size());
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
