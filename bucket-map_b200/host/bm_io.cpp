// bm_io.cpp -- device context wrapper, FASTA / FASTQ readers, FracMinHash hash function of the C++ host.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <chrono>
#include <cstring>
#include <ctime>
#include <mutex>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

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
    if (const char* e = std::getenv("BM_FINE_TABLE")) bm_index_fine_table(c_, std::atoi(e) ? 1 : 0);
}
context::~context() { bm_ctx_destroy(c_); }

context& context::default_context() {
    static std::once_flag once;
    static context* c = nullptr;
    std::call_once(once, [] {
        const char* e = std::getenv("BM_DEVICE");
        c = new context(e ? std::atoi(e) : 0);
    });
    return *c;
}

// ---- small utilities ------------------------------------------------------------------------------------------------------
bool page_buffer::allocate(size_t n) {
    release();
    void* p = mmap(nullptr, n ? n : 1, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
    if (p == MAP_FAILED) return false;
    p_ = (uint8_t*)p; n_ = n ? n : 1;
    return true;
}
void page_buffer::release() { if (p_) munmap(p_, n_); p_ = nullptr; n_ = 0; }

bool file_map::open(std::filesystem::path const& p) {
    close();
    int fd = ::open(p.c_str(), O_RDONLY);
    if (fd < 0) return false;
    struct stat st;
    if (fstat(fd, &st) != 0) { ::close(fd); return false; }
    n_ = (size_t)st.st_size;
    ok_ = true;
    if (n_) {
        void* m = mmap(nullptr, n_, PROT_READ, MAP_PRIVATE, fd, 0);
        if (m == MAP_FAILED) { ::close(fd); n_ = 0; ok_ = false; return false; }
        madvise(m, n_, MADV_SEQUENTIAL);
        p_ = (const char*)m;
    }
    ::close(fd);
    return true;
}
void file_map::close() { if (p_) munmap((void*)p_, n_); p_ = nullptr; n_ = 0; ok_ = false; }

static unsigned g_host_threads = 0;
unsigned host_threads() {
    if (!g_host_threads) {
        unsigned n = 0;
        if (const char* e = std::getenv("BM_THREADS")) n = (unsigned)std::atoi(e);
        if (!n) { n = std::thread::hardware_concurrency(); if (n > 16) n = 16; }
        g_host_threads = n ? n : 1;
    }
    return g_host_threads;
}
void set_host_threads(unsigned n) { g_host_threads = n ? n : 1; }

void parallel_parts(unsigned parts, std::function<void(unsigned, unsigned)> const& fn) {
    if (parts <= 1) { fn(0, 1); return; }
    std::vector<std::thread> ts;
    ts.reserve(parts - 1);
    for (unsigned i = 1; i < parts; i++) ts.emplace_back([&fn, i, parts] { fn(i, parts); });
    fn(0, parts);
    for (auto& t : ts) t.join();
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

// The file is memory-mapped and cut into one byte range per host thread (at line starts).  Pass 1 counts the bases of every
// range and notes its header lines; a short serial step turns that into record lengths and the base offset of every
// range; pass 2 packs each range, four bases per byte through a 256-entry table, straight into its place of the 2-bit
// array (the two bytes a range shares with its neighbours are OR-ed in atomically).  Both passes are memory-bound.
bool genome::read_fasta(std::filesystem::path const& p) {
    file_map fm;
    if (!fm.open(p)) return false;
    ids.clear(); rec_len.clear(); packed.release(); n_bases = 0;
    static uint8_t lut[256];
    static bool lut_ready = false;
    if (!lut_ready) {
        for (int c = 0; c < 256; c++) lut[c] = rank_fasta((unsigned char)c);
        lut[(unsigned char)'\r'] = lut[(unsigned char)' '] = lut[(unsigned char)'\t'] = 0xFF;   // skipped
        lut_ready = true;
    }
    const char* d = fm.data();
    const size_t n = fm.size();
    unsigned T = host_threads();
    if (n < ((size_t)1 << 22)) T = 1;
    std::vector<size_t> cut(T + 1, n);
    cut[0] = 0;
    for (unsigned t = 1; t < T; t++) {
        size_t at = n / T * t;
        const char* nl = at < n ? (const char*)memchr(d + at, '\n', n - at) : nullptr;
        cut[t] = nl ? (size_t)(nl - d) + 1 : n;
        if (cut[t] < cut[t - 1]) cut[t] = cut[t - 1];
    }
    struct header { size_t bases_before; std::string text; };          // bases of the range that precede the header
    struct part { std::vector<header> hdr; uint64_t bases = 0; };
    std::vector<part> parts(T);
    auto count_line = [&](const char* s, size_t len) -> uint64_t {
        if (len && s[len - 1] == '\r') len--;
        if (!memchr(s, ' ', len) && !memchr(s, '\t', len) && !memchr(s, '\r', len)) return len;
        uint64_t c = 0;
        for (size_t i = 0; i < len; i++) c += lut[(unsigned char)s[i]] != 0xFF;
        return c;
    };
    parallel_parts(T, [&](unsigned t, unsigned) {
        part& P = parts[t];
        size_t i = cut[t];
        const size_t e = cut[t + 1];
        while (i < e) {
            const char* nl = (const char*)memchr(d + i, '\n', e - i);
            const size_t le = nl ? (size_t)(nl - d) : e;
            if (le > i && d[i] == '>') {
                size_t hl = le - i - 1;
                if (hl && d[le - 1] == '\r') hl--;
                P.hdr.push_back({(size_t)P.bases, std::string(d + i + 1, hl)});
            } else if (le > i) {
                P.bases += count_line(d + i, le - i);
            }
            i = le + 1;
        }
    });
    // serial: record lengths, base offset of every range.  Bases before the first header of the file belong to no record.
    std::vector<uint64_t> base0(T + 1, 0);
    bool open_rec = false;
    uint64_t total = 0;
    std::vector<uint64_t> drop(T, 0);          // bases at the start of range t that precede the file's first header
    for (unsigned t = 0; t < T; t++) {
        base0[t] = total;
        uint64_t prev = 0;
        for (auto const& h : parts[t].hdr) {
            const uint64_t seg = h.bases_before - prev;
            if (open_rec) { rec_len.back() += seg; total += seg; } else drop[t] += seg;
            prev = h.bases_before;
            ids.push_back(h.text); rec_len.push_back(0); open_rec = true;
        }
        const uint64_t seg = parts[t].bases - prev;
        if (open_rec) { rec_len.back() += seg; total += seg; } else drop[t] += seg;
    }
    base0[T] = total;
    n_bases = total;
    const size_t nbytes = ((size_t)n_bases + 3) / 4 + 64;
    if (!packed.allocate(nbytes)) return false;
    uint8_t* out = packed.data();
    parallel_parts(T, [&](unsigned t, unsigned) {
        uint64_t g = base0[t];
        uint64_t skip = drop[t];
        const uint64_t g_end = base0[t + 1];
        auto put = [&](uint8_t r) {
            if (skip) { skip--; return; }
            const uint8_t v = (uint8_t)(r << ((g & 3) << 1));
            const uint64_t byte = g >> 2;
            // the first and last byte of a range may be shared with the neighbouring range
            if (byte == (base0[t] >> 2) || byte == ((g_end - 1) >> 2)) __atomic_fetch_or(out + byte, v, __ATOMIC_RELAXED);
            else out[byte] |= v;
            g++;
        };
        size_t i = cut[t];
        const size_t e = cut[t + 1];
        const unsigned char* s = (const unsigned char*)d;
        while (i < e) {
            const char* nl = (const char*)memchr(d + i, '\n', e - i);
            const size_t le = nl ? (size_t)(nl - d) : e;
            if (le > i && d[i] != '>') {
                size_t x = i;
                while (x < le) {
                    if (!skip && (g & 3) == 0 && x + 4 <= le && g + 4 <= g_end && (g >> 2) != ((g_end - 1) >> 2) &&
                        (g >> 2) != (base0[t] >> 2)) {
                        const uint8_t v0 = lut[s[x]], v1 = lut[s[x + 1]], v2 = lut[s[x + 2]], v3 = lut[s[x + 3]];
                        if (!((v0 | v1 | v2 | v3) & 0x80)) {
                            out[g >> 2] = (uint8_t)(v0 | (v1 << 2) | (v2 << 4) | (v3 << 6));
                            g += 4; x += 4;
                            continue;
                        }
                    }
                    const uint8_t v = lut[s[x++]];
                    if (v != 0xFF) put(v);
                }
            }
            i = le + 1;
        }
    });
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
fastq_reader::fastq_reader(std::filesystem::path const& p) { map_.open(p); }

// next line as a view into the mapped file; false at end of file
// Line feeds of the next window of the file, found by the host threads (each scans its own byte range with memchr, which also
// spreads the first touch of the mapping's pages over the threads).  Called when every line feed known so far has been used,
// so everything between pos_ and win_hi_ is the head of one line.
static constexpr uint64_t kCrFlag = 1ull << 63;      // in an entry of the line-feed index: the line ends in "\r\n"

void fastq_reader::index_window() {
    constexpr size_t kWindow = (size_t)256 << 20;
    const size_t end = map_.size();
    const size_t lo = win_hi_, hi = end - lo > kWindow ? lo + kWindow : end;
    unsigned T = host_threads();
    if (hi - lo < ((size_t)1 << 20)) T = 1;
    std::vector<std::vector<uint64_t>> found(T);
    const char* d = map_.data();
    parallel_parts(T, [&](unsigned t, unsigned parts) {
        size_t i = lo + (size_t)((unsigned __int128)(hi - lo) * t / parts);
        const size_t e = lo + (size_t)((unsigned __int128)(hi - lo) * (t + 1) / parts);
        std::vector<uint64_t>& v = found[t];
        v.reserve((e - i) / 96 + 16);
        while (i < e) {
            const char* nl = (const char*)memchr(d + i, '\n', e - i);
            if (!nl) break;
            const uint64_t at = (uint64_t)(nl - d);
            v.push_back(at | ((at > 0 && nl[-1] == '\r') ? kCrFlag : 0ull));     // the walk over the lines reads no file byte
            i = (size_t)at + 1;
        }
    });
    size_t total = 0;
    for (auto const& v : found) total += v.size();
    nl_.resize(total);
    size_t at = 0;
    for (auto const& v : found) { if (!v.empty()) memcpy(nl_.data() + at, v.data(), v.size() * 8); at += v.size(); }
    nl_i_ = 0;
    win_hi_ = hi;
}

bool fastq_reader::line(const char*& s, size_t& len) {
    const size_t end = map_.size();
    if (pos_ >= end) return false;
    while (nl_i_ == nl_.size() && win_hi_ < end) index_window();
    const char* b = map_.data() + pos_;
    const bool has_nl = nl_i_ < nl_.size();
    const uint64_t ent = has_nl ? nl_[nl_i_++] : (uint64_t)end;
    const size_t e = (size_t)(ent & ~kCrFlag);
    s = b;
    len = e - pos_;
    pos_ = e + (has_nl ? 1 : 0);
    if (has_nl ? (ent & kCrFlag) != 0 && len : (len && s[len - 1] == '\r')) len--;
    return true;
}

bool fastq_reader::next(read_batch& b, uint32_t max_reads, uint64_t max_bases, uint32_t k, uint32_t bq) {
    b.clear();
    if (!map_.ok()) return false;
    static const bool trace = std::getenv("BM_HOST_TRACE") != nullptr;
    const auto t_begin = std::chrono::steady_clock::now();
    auto since = [&] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_begin).count(); };
    // 1. split the records (one thread: four memchr per record); the text is not copied
    const char *id, *s, *plus, *q;
    size_t nid, ns, np, nq;
    uint64_t bases = 0;
    while (b.size() < max_reads && bases < max_bases) {
        bool got = true;
        do { if (!line(id, nid)) { got = false; break; } } while (nid == 0);
        if (!got) break;
        if (!line(s, ns) || !line(plus, np) || !line(q, nq)) break;     // truncated record: dropped
        b.id.push_back(id + 1); b.id_len.push_back((uint32_t)(nid ? nid - 1 : 0));   // full id line after '@' (seqan3 default)
        b.seq.push_back(s); b.qual.push_back(q); b.qual_len.push_back((uint32_t)(nq < ns ? nq : ns));
        bases += ns;
        b.off.push_back(bases);
    }
    const uint32_t n = b.size();
    if (!n) return false;
    const double t_split = since();
    // 2. the device form, written by the host threads into pinned memory
    b.read_off.reserve((size_t)(n + 1) * 8);
    b.pl_off.reserve((size_t)(n + 1) * 4);
    memcpy(b.read_off.data(), b.off.data(), (size_t)(n + 1) * 8);
    uint32_t* plo = b.pl_off.as<uint32_t>();
    if (bm_pack_layout(b.off.data(), n, plo) != BM_OK) {
        fprintf(stderr, "[ERROR]\t\tA batch of %u reads is too large for the packed form; use a smaller --batch-reads.\n", n);
        std::exit(2);
    }
    const size_t nch = plo[n];
    b.planes.reserve((nch + 1) * 8);
    unsigned T = host_threads();
    if (n < 4096) T = 1;
    const double t_alloc = since();
    // The quality plane is made (and its pinned buffer taken) only for a batch that needs it: the first pass packs the bases
    // and reports whether any k-mer window fails -b; most batches of good reads stop there.
    std::vector<uint32_t> low(T * 16, 0);
    auto pack = [&](uint64_t* planes, uint32_t* qk) {
        parallel_parts(T, [&](unsigned t, unsigned parts) {
            const uint32_t r0 = (uint32_t)((uint64_t)n * t / parts), r1 = (uint32_t)((uint64_t)n * (t + 1) / parts);
            bm_pack_reads_v((const uint8_t* const*)b.seq.data(), (const uint8_t* const*)b.qual.data(), b.qual_len.data(),
                            b.off.data(), plo, r0, r1, k, bq, planes, qk, &low[t * 16]);
        });
    };
    pack(b.planes.as<uint64_t>(), nullptr);
    b.all_pass = true;
    for (unsigned t = 0; t < T; t++) if (low[t * 16]) b.all_pass = false;
    if (!b.all_pass) {
        b.qk.reserve((nch + 1) * 4);
        pack(b.planes.as<uint64_t>(), b.qk.as<uint32_t>());
    }
    if (trace) fprintf(stderr, "[host] fastq batch of %u reads: split %.2f  buffers %.2f  pack on %u threads %.2f ms\n", n, t_split,
                       t_alloc - t_split, T, since() - t_alloc);
    return true;
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
