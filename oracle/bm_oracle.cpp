/*
 * bm_oracle.cpp -- CPU restatement of BucketMap's hot path.  TEST INFRASTRUCTURE ONLY (see bm_oracle.h).
 *
 * Written from the reference's behaviour, not from its text: the data model is flat arrays + 2-bit genome,
 * there is no seqan3, and every libstdc++-dependent behaviour the results hinge on is kept by using the same
 * container (std::unordered_multimap for the per-bucket k-mer index, std::map for the vote / merge maps).
 * Citations "X.h:a-b" are relative to /root/reference/bucket_map/.
 */
#include "bm_oracle.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <filesystem>
#include <fstream>
#include <map>
#include <string>
#include <thread>
#include <tuple>
#include <unordered_map>
#include <vector>

namespace {

inline uint32_t gbase(const uint8_t* p, uint64_t i) { return (p[i >> 2] >> ((i & 3) << 1)) & 3u; }

/* seqan3 dna4 char->rank (FASTQ, _phred94_traits utils.h:194-204).  IUPAC codes map to their first base
 * (seqan3 behaviour restated from memory: unpinned; synthetic inputs are ACGT only). */
inline uint8_t rank_dna4(uint8_t c) {
    switch (c) {
        case 'A': case 'a': return 0;
        case 'C': case 'c': return 1;
        case 'G': case 'g': return 2;
        case 'T': case 't': case 'U': case 'u': return 3;
        case 'Y': case 'y': case 'S': case 's': case 'B': case 'b': return 1;
        case 'K': case 'k': return 2;
        default: return 0;
    }
}
/* FASTA goes through dna5 first (utils.h:70,91): everything that is not ACGTU becomes N, then N -> A. */
inline uint8_t rank_dna5_to_dna4(uint8_t c) {
    switch (c) {
        case 'C': case 'c': return 1;
        case 'G': case 'g': return 2;
        case 'T': case 't': case 'U': case 'u': return 3;
        default: return 0;
    }
}

template <class F>
void parallel_for(uint64_t n, uint32_t threads, F f) {
    if (threads <= 1 || n < 2) { f(0, n, 0u); return; }
    if (threads > n) threads = (uint32_t)n;
    std::vector<std::thread> th;
    for (uint32_t t = 0; t < threads; t++) {
        uint64_t lo = n * t / threads, hi = n * (t + 1) / threads;
        th.emplace_back([=] { f(lo, hi, t); });
    }
    for (auto& x : th) x.join();
}

/* tools/hash_function_generator.h:25-42 (the pb_ds prime table); only the 32-bit part is ever reached. */
const uint64_t kPrimes[] = {5ull, 11ull, 23ull, 47ull, 97ull, 199ull, 409ull, 823ull, 1741ull, 3469ull, 6949ull,
                            14033ull, 28411ull, 57557ull, 116731ull, 236897ull, 480881ull, 976369ull, 1982627ull,
                            4026031ull, 8175383ull, 16601593ull, 33712729ull, 68460391ull, 139022417ull,
                            282312799ull, 573292817ull, 1164186217ull, 2364114217ull, 4294967291ull};

/* utils.h:291-302 */
inline uint32_t revcomp_hash(uint32_t h, uint32_t k) {
    uint32_t r = 0;
    for (uint32_t i = 0; i < k; i++) {
        r |= (~h) & 3u;
        if (i != k - 1) { r <<= 2; h >>= 2; }
    }
    return r;
}

/* utils.h:160-178.  The reference never updates last_upper_bound (stays 0), so its early return fires only for
 * ub==0 and leaves the previous call's samples in place (out-of-range reads follow: UB).  Here ub==0 yields what
 * the formula itself gives: n zeros. */
void sample_det(uint32_t n, uint32_t ub, std::vector<uint32_t>& out) {
    out.clear();
    if (n == 0) return;
    if (ub == 0) { out.assign(n, 0u); return; }
    double delta = (n == 1) ? 0.0 : static_cast<double>(ub + 1u) / (double)(n - 1u);
    for (int i = 0; i + 1 < (int)n; i++) out.push_back((uint32_t)std::floor(i * delta));
    out.push_back(ub);
}

struct SegView {
    const uint8_t* seq;   /* ASCII */
    const uint8_t* qual;  /* ASCII phred+33 */
    uint32_t len;
};

/* seqan3 views::kmer_hash(ungapped{k}) (call sites q_gram_mapper.h:431, bucket_locator.h:318) and
 * views::kmer_quality (quality_filter.h:531-534,611-631): big-endian base-4 value, window sum of phred ranks. */
void kmers_and_quals(const SegView& s, uint32_t k, std::vector<uint32_t>& hash, std::vector<uint32_t>& qsum) {
    hash.clear(); qsum.clear();
    if (s.len < k) return;
    uint32_t nk = s.len - k + 1;
    hash.resize(nk); qsum.resize(nk);
    uint64_t mask = (k >= 16) ? 0xFFFFFFFFull : ((1ull << (2 * k)) - 1);
    uint64_t h = 0; uint32_t qs = 0;
    for (uint32_t i = 0; i < s.len; i++) {
        h = ((h << 2) | rank_dna4(s.seq[i])) & mask;
        qs += (uint32_t)(s.qual[i] - 33);
        if (i >= k) qs -= (uint32_t)(s.qual[i - k] - 33);
        if (i + 1 >= k) { hash[i + 1 - k] = (uint32_t)h; qsum[i + 1 - k] = qs; }
    }
}

struct Matrix {          /* q-gram bit matrix re-laid into 64-bit words for the bitset-style ops */
    uint32_t NB = 0, nw = 0, n_rows = 0;
    std::vector<uint64_t> w;   /* n_rows * nw */
    std::vector<uint64_t> ones; /* NB ones */
    void load(const uint8_t* rows, uint32_t n_rows_, uint32_t NB_) {
        NB = NB_; nw = (NB + 63) / 64; n_rows = n_rows_;
        uint32_t W = (NB + 7) >> 3;
        w.assign((size_t)n_rows * nw, 0);
        for (uint32_t r = 0; r < n_rows; r++)
            memcpy((uint8_t*)&w[(size_t)r * nw], rows + (size_t)r * W, W);   /* little-endian host */
        ones.assign(nw, ~0ull);
        if (NB & 63) ones[nw - 1] = (1ull << (NB & 63)) - 1;
    }
};

/* q_gram_mapper.h:39-102: the fault-tolerance cascade, literal. */
struct Cascade {
    uint32_t F, nw;
    std::vector<uint64_t> f;  /* F * nw */
    Cascade(uint32_t F_, uint32_t nw_) : F(F_), nw(nw_), f((size_t)F_ * nw_) {}
    void reset(const std::vector<uint64_t>& ones) {
        for (uint32_t i = 0; i < F; i++) memcpy(&f[(size_t)i * nw], ones.data(), nw * 8);
    }
    void read(const uint64_t* in) {
        for (uint32_t i = 0; i + 1 < F; i++) {
            uint64_t* a = &f[(size_t)i * nw]; const uint64_t* b = &f[(size_t)(i + 1) * nw];
            for (uint32_t j = 0; j < nw; j++) a[j] &= (b[j] | in[j]);
        }
        uint64_t* a = &f[(size_t)(F - 1) * nw];
        for (uint32_t j = 0; j < nw; j++) a[j] &= in[j];
    }
    void best(std::vector<uint32_t>& res) const {
        res.clear();
        for (int i = (int)F - 1; i >= 0; i--) {
            const uint64_t* a = &f[(size_t)i * nw];
            for (uint32_t j = 0; j < nw; j++) {
                uint64_t x = a[j];
                while (x) { res.push_back(j * 64 + (uint32_t)__builtin_ctzll(x)); x &= x - 1; }
            }
            if (!res.empty()) return;
        }
    }
};

struct MapperState {
    const orc_params* p;
    const Matrix* M;
    const int32_t* k2i;
    uint32_t n_k2i;
    const uint32_t* zeros;
    uint32_t threshold;
    Cascade casc;
    std::vector<uint64_t> bf;
    std::vector<uint32_t> hash, qsum, good, samp, so, sr;
    uint64_t rows_anded = 0;
    MapperState(const orc_params* p_, const Matrix* M_, const int32_t* k2i_, const uint32_t* zeros_, uint32_t thr)
        : p(p_), M(M_), k2i(k2i_), n_k2i(1u << (2 * p_->q)), zeros(zeros_), threshold(thr),
          casc(p_->mapper_fault, M_->nw), bf(M_->nw) {}

    int idx_of(uint32_t g) const { return g < n_k2i ? k2i[g] : -1; }

    /* q_gram_mapper.h:189-196 */
    bool distinguishable(uint32_t h) const {
        uint32_t Q = n_k2i - 1;
        for (uint32_t i = 0; i <= p->k - p->q; i++) {
            int idx = idx_of((h >> (2 * i)) & Q);
            if (idx >= 0 && zeros[idx] >= threshold) return true;
        }
        return false;
    }
    /* q_gram_mapper.h:380-412 */
    void query(const std::vector<uint32_t>& hs, std::vector<uint32_t>& res) {
        uint32_t Q = n_k2i - 1, nw = M->nw;
        casc.reset(M->ones);
        for (uint32_t h : hs) {
            memcpy(bf.data(), M->ones.data(), nw * 8);
            for (uint32_t i = 0; i <= p->k - p->q; i++) {
                int idx = idx_of((h >> (2 * i)) & Q);
                if (idx >= 0) {
                    const uint64_t* row = &M->w[(size_t)idx * nw];
                    for (uint32_t j = 0; j < nw; j++) bf[j] &= row[j];
                    rows_anded++;
                }
            }
            casc.read(bf.data());
        }
        casc.best(res);
    }
    /* q_gram_mapper.h:414-480 */
    void query_sequence(const SegView& s, std::vector<uint32_t>& co, std::vector<uint32_t>& cr) {
        co.clear(); cr.clear();
        kmers_and_quals(s, p->k, hash, qsum);
        uint32_t minq = p->avg_base_quality * p->k;  /* q_gram_mapper.h:303 */
        good.clear();
        for (uint32_t j = 0; j < hash.size(); j++)
            if (distinguishable(hash[j]) && qsum[j] >= minq) good.push_back(hash[j]);
        if ((double)good.size() < 0.2 * p->mapper_samples) return;
        sample_det(p->mapper_samples, (uint32_t)good.size() - 1u, samp);
        so.clear(); sr.clear();
        for (uint32_t x : samp) so.push_back(good[x]);
        for (uint32_t h : so) sr.push_back(revcomp_hash(h, p->k));
        query(so, co);
        query(sr, cr);
        if (co.size() > p->max_candidates) co.clear();
        if (cr.size() > p->max_candidates) cr.clear();
    }
};

uint32_t dist_threshold(float d, uint32_t NB) { return (unsigned int)(d * NB); }  /* q_gram_mapper.h:163 */

/* ------------------------------------------------------------------ locator ------------------------------- */

/* bucket_locator.h:209-290, literal containers. */
std::pair<int, unsigned> find_offset(const std::unordered_multimap<unsigned int, int>& index, const orc_params* p,
                                     const uint32_t* samp_hash, const uint16_t* samp_idx, uint32_t seg_len,
                                     bool reverse, std::map<int, unsigned int>& votes) {
    votes.clear();
    const uint32_t P = p->locator_samples, k = p->k;
    const int A = (int)p->allowed_indel;
    for (uint32_t i = 0; i < P; i++) {
        uint32_t si = reverse ? P - 1 - i : i;
        unsigned int kmer = samp_hash[si], cur = samp_idx[si];
        if (reverse) { kmer = revcomp_hash(kmer, k); cur = seg_len - k - cur; }
        auto range = index.equal_range(kmer);
        if (votes.empty()) {
            for (auto it = range.first; it != range.second; ++it) {
                auto position = it->second - cur;       /* unsigned wrap, then int key */
                votes[position]++;
            }
        } else {
            for (auto it = range.first; it != range.second; ++it) {
                bool voted = false;
                auto position = it->second - cur;
                auto lo = votes.lower_bound(position - A);
                auto hi = votes.upper_bound(position + A);
                for (auto jt = lo; jt != hi; ++jt) { jt->second++; voted = true; }
                if (!voted) votes[position]++;
            }
        }
    }
    if (!votes.empty()) {
        auto best = votes.begin();
        for (auto it = votes.begin(); it != votes.end(); ++it)
            if (it->second > best->second) best = it;   /* ties keep the smallest key (bucket_locator.h:281-283) */
        if (best->second >= (unsigned)((int)P - (int)p->allowed_mismatch) && best->first >= 0)
            return {best->first, best->second};
    }
    return {-1, 0u};
}

struct CandRef { uint32_t seg; };

/* ------------------------------------------------------------------ alignment ----------------------------- */

/* seqan3::align_pairwise(edit_scheme, method_global with free end gaps in sequence1 only) + cigar_from_alignment,
 * call site bucket_locator.h:520-528,569-577.  seqan3's source is not available here, so this is a restatement
 * of its published behaviour with these tie-breaks (UNPINNED, see oracle/README.md):
 *   - score = -min_j D[m][j]; the end column is the LAST j attaining the minimum (<= update while scanning);
 *   - traceback prefers diagonal, then up (consumes a query base: 'I'), then left (consumes a text base: 'D');
 *   - traceback stops on row 0 (free leading gap); column 0 with rows left is all 'I'. */
void align_semi_global(const uint8_t* text, uint32_t n, const uint8_t* query, uint32_t m, int32_t& score,
                       uint32_t& begin, std::vector<uint32_t>& ops /* run-length, forward order */) {
    std::vector<uint16_t> D((size_t)(m + 1) * (n + 1));
    auto at = [&](uint32_t i, uint32_t j) -> uint16_t& { return D[(size_t)i * (n + 1) + j]; };
    for (uint32_t j = 0; j <= n; j++) at(0, j) = 0;
    for (uint32_t i = 1; i <= m; i++) {
        at(i, 0) = (uint16_t)i;
        for (uint32_t j = 1; j <= n; j++) {
            uint32_t d = at(i - 1, j - 1) + (query[i - 1] != text[j - 1]);
            uint32_t u = at(i - 1, j) + 1u, l = at(i, j - 1) + 1u;
            at(i, j) = (uint16_t)std::min(d, std::min(u, l));
        }
    }
    uint32_t best = at(m, 0), bj = 0;
    for (uint32_t j = 1; j <= n; j++) if (at(m, j) <= best) { best = at(m, j); bj = j; }
    score = -(int32_t)best;
    std::vector<uint8_t> rev;
    uint32_t i = m, j = bj;
    while (i > 0) {
        if (j > 0 && at(i, j) == at(i - 1, j - 1) + (query[i - 1] != text[j - 1])) { rev.push_back(0); i--; j--; }
        else if (at(i, j) == at(i - 1, j) + 1u) { rev.push_back(1); i--; }
        else { rev.push_back(2); j--; }
    }
    begin = j;
    ops.clear();
    for (size_t t = rev.size(); t-- > 0;) {
        if (!ops.empty() && (ops.back() & 15u) == rev[t]) ops.back() += 16u;
        else ops.push_back(16u | rev[t]);
    }
}

/* ------------------------------------------------------------------ file I/O (CLI flows) ------------------ */

struct Fasta {
    std::vector<std::string> ids;
    std::vector<uint64_t> len;
    std::vector<uint8_t> packed;
    uint64_t total = 0;
};
bool read_fasta(const std::string& path, Fasta& g) {
    std::ifstream in(path);
    if (!in) return false;
    std::string line;
    auto push = [&](uint8_t r) {
        if ((g.total & 3) == 0) g.packed.push_back(0);
        g.packed.back() |= (uint8_t)(r << ((g.total & 3) << 1));
        g.total++;
    };
    while (std::getline(in, line)) {
        if (!line.empty() && line.back() == '\r') line.pop_back();
        if (line.empty()) continue;
        if (line[0] == '>') { g.ids.push_back(line.substr(1)); g.len.push_back(0); continue; }
        if (g.ids.empty()) continue;
        for (char c : line) { if (c == ' ' || c == '\t') continue; push(rank_dna5_to_dna4((uint8_t)c)); g.len.back()++; }
    }
    g.packed.resize(g.packed.size() + 16, 0);
    return true;
}
struct Fastq {
    std::vector<std::string> ids;
    std::vector<uint8_t> seq, qual;
    std::vector<uint64_t> off{0};
};
bool read_fastq(const std::string& path, Fastq& f) {
    std::ifstream in(path);
    if (!in) return false;
    std::string id, s, plus, q;
    while (std::getline(in, id)) {
        if (id.empty()) continue;
        if (!std::getline(in, s) || !std::getline(in, plus) || !std::getline(in, q)) break;
        f.ids.push_back(id.substr(1));
        f.seq.insert(f.seq.end(), s.begin(), s.end());
        q.resize(s.size(), '!');
        f.qual.insert(f.qual.end(), q.begin(), q.end());
        f.off.push_back(f.seq.size());
    }
    return true;
}

}  // namespace

/* ======================================================================== C ABI ========================== */
extern "C" {

void orc_derive_params(orc_params* p, uint32_t bucket_len, uint32_t read_len, uint32_t q, uint32_t k, uint32_t s,
                       float d, uint32_t b, float e, float n, uint32_t pl, uint32_t u) {
    p->bucket_len = bucket_len; p->read_len = read_len; p->q = q; p->k = k;
    p->mapper_samples = s;
    p->mapper_fault = (uint32_t)std::ceil(s * e);                 /* main.cpp:207 (unsigned*float) */
    p->distinguishability = d; p->avg_base_quality = b;
    p->max_candidates = 30; p->num_segment_samples = 5;
    p->locator_samples = pl;
    p->allowed_mismatch = (uint32_t)(int)std::ceil(e * pl);       /* bucket_locator.h:419 */
    p->allowed_indel = (uint32_t)(int)std::ceil(n * read_len);    /* bucket_locator.h:420 */
    p->indel_rate = n; p->quality_threshold = u;
}
void orc_default_params(orc_params* p, uint32_t bucket_len) {
    orc_derive_params(p, bucket_len, 300, 9, 12, 15, 0.5f, 25, 0.4f, 0.02f, 10, 40);  /* main.cpp:12-39 */
}

uint32_t orc_fracminhash(uint32_t q, uint64_t x, uint64_t y, float frac, int32_t* k2i) {
    const uint64_t table = 10000;                                   /* main.cpp:176 */
    uint64_t pr = 0;
    for (uint64_t c : kPrimes) if (c > 10 * table) { pr = c; break; }
    unsigned int thr = (unsigned int)(table * frac);                /* main.cpp:185 (size_t*float) */
    int32_t next = 0;
    uint64_t n = 1ull << (2 * q);
    for (uint64_t i = 0; i < n; i++) {
        uint64_t h = (x * (uint64_t)(int)i + y) % pr % table;       /* hash_function_generator.h:114 */
        k2i[i] = (h <= thr) ? next++ : -1;                          /* bucket_indexer.h:149-156 */
    }
    return (uint32_t)next;
}

uint32_t orc_num_buckets(const uint64_t* rec_len, uint32_t n_rec, uint32_t L) {
    uint64_t nb = 0;
    for (uint32_t i = 0; i < n_rec; i++) if (rec_len[i]) nb += (rec_len[i] + L - 1) / L;
    return (uint32_t)nb;
}

uint32_t orc_bucketize(const uint64_t* rec_len, uint32_t n_rec, uint32_t L, uint32_t r, uint64_t* b_start,
                       uint32_t* b_len, uint32_t* b_rec) {
    uint32_t n = 0; uint64_t base = 0;
    for (uint32_t rec = 0; rec < n_rec; rec++) {
        float total = (float)rec_len[rec];                          /* utils.h:74-75: float division */
        int nb = (int)std::ceil(total / (int)L);
        for (int i = 0; i < nb; i++) {
            int64_t start = (int64_t)i * L, end = start + L + r;
            if (end > (int64_t)rec_len[rec]) end = (int64_t)rec_len[rec];
            if (end - start <= (int64_t)r) continue;                /* utils.h:88-90 */
            b_start[n] = base + (uint64_t)start; b_len[n] = (uint32_t)(end - start); b_rec[n] = rec; n++;
        }
        base += rec_len[rec];
    }
    return n;
}

void orc_index_build(const uint8_t* packed, const uint64_t* b_start, const uint32_t* b_len, uint32_t n_emit,
                     uint32_t NB, uint32_t q, const int32_t* k2i, uint32_t n_rows, uint8_t* rows, uint32_t threads) {
    const uint32_t W = (NB + 7) >> 3;
    memset(rows, 0, (size_t)n_rows * W);
    const uint32_t mask = (1u << (2 * q)) - 1;
    /* threads own disjoint byte columns (8 buckets each) so no two threads touch the same byte */
    uint32_t groups = (n_emit + 7) / 8;
    parallel_for(groups, threads, [&](uint64_t glo, uint64_t ghi, uint32_t) {
        for (uint64_t b = glo * 8; b < std::min<uint64_t>(ghi * 8, n_emit); b++) {
            uint32_t h = 0;
            for (uint32_t i = 0; i < b_len[b]; i++) {
                h = ((h << 2) | gbase(packed, b_start[b] + i)) & mask;
                if (i + 1 >= q) {
                    int idx = k2i[h];                               /* bucket_indexer.h:57-60 */
                    if (idx >= 0) rows[(size_t)idx * W + (b >> 3)] |= (uint8_t)(1u << (b & 7));  /* :64-73 */
                }
            }
        }
    });
}

uint32_t orc_zeros(const uint8_t* rows, uint32_t n_rows, uint32_t NB, float d, uint32_t* zeros) {
    const uint32_t W = (NB + 7) >> 3;
    const uint32_t thr = dist_threshold(d, NB);
    uint32_t valid = 0;
    for (uint32_t r = 0; r < n_rows; r++) {
        uint32_t ones = 0;
        for (uint32_t j = 0; j < W; j++) ones += (uint32_t)__builtin_popcount(rows[(size_t)r * W + j]);
        if (NB - ones > thr) valid++;
        zeros[r] = NB - ones;                                       /* q_gram_mapper.h:177-183 (both branches) */
    }
    return valid;
}

uint32_t orc_segments(const orc_params* p, const uint64_t* read_off, uint32_t n_reads, uint32_t* seg_read,
                      uint32_t* seg_start, uint32_t* seg_len) {
    uint32_t n = 0;
    std::vector<uint32_t> st;
    for (uint32_t i = 0; i < n_reads; i++) {
        uint32_t len = (uint32_t)(read_off[i + 1] - read_off[i]);
        st.assign(1, 0u);
        if (len > 2 * p->read_len) sample_det(p->num_segment_samples, len - p->read_len - 1, st);
        for (uint32_t s : st) {
            seg_read[n] = i; seg_start[n] = s;
            seg_len[n] = std::min(s + p->read_len, len) - s;        /* q_gram_mapper.h:520-521 */
            n++;
        }
    }
    return n;
}

void orc_sample(uint32_t n, uint32_t ub, uint32_t* out) {
    std::vector<uint32_t> v; sample_det(n, ub, v);
    std::copy(v.begin(), v.end(), out);
}

struct orc_index {                 /* what q_gram_mapper::load leaves in memory (q_gram_mapper.h:318-372) */
    Matrix M;
    std::vector<uint32_t> zeros;
    std::vector<int32_t> k2i;
    uint32_t thr = 0;
};

orc_index* orc_index_open(const orc_params* p, uint32_t NB, const int32_t* k2i, const uint8_t* rows, uint32_t n_rows) {
    orc_index* ix = new orc_index();
    ix->M.load(rows, n_rows, NB);
    ix->zeros.resize(n_rows);
    orc_zeros(rows, n_rows, NB, p->distinguishability, ix->zeros.data());
    ix->thr = dist_threshold(p->distinguishability, NB);
    ix->k2i.assign(k2i, k2i + (1ull << (2 * p->q)));
    return ix;
}
void orc_index_close(orc_index* ix) { delete ix; }

void orc_map_segments_h(const orc_index* ix, const orc_params* p, const uint8_t* seq, const uint8_t* qual,
                        const uint64_t* read_off, const uint32_t* seg_read, const uint32_t* seg_start,
                        const uint32_t* seg_len, uint32_t n_seg, uint32_t* cand_off, uint32_t* cand, uint64_t* stats,
                        uint32_t threads) {
    std::vector<std::vector<uint32_t>> lists((size_t)2 * n_seg);
    if (threads == 0) threads = 1;
    std::vector<uint64_t> anded(threads, 0);
    parallel_for(n_seg, threads, [&](uint64_t lo, uint64_t hi, uint32_t t) {
        MapperState st(p, &ix->M, ix->k2i.data(), ix->zeros.data(), ix->thr);
        for (uint64_t s = lo; s < hi; s++) {
            uint64_t o = read_off[seg_read[s]] + seg_start[s];
            SegView v{seq + o, qual + o, seg_len[s]};
            st.query_sequence(v, lists[2 * s], lists[2 * s + 1]);
        }
        anded[t] = st.rows_anded;
    });
    uint64_t mapped = 0, so = 0, sr = 0, tot = 0; uint32_t pos = 0;
    int64_t last_mapped_read = -1;
    for (uint32_t s = 0; s < n_seg; s++) {
        for (int d = 0; d < 2; d++) {
            cand_off[2 * s + d] = pos;
            for (uint32_t b : lists[2 * s + d]) cand[pos++] = b;
        }
        if (!lists[2 * s].empty() || !lists[2 * s + 1].empty()) {
            /* q_gram_mapper.h:527-545: mapped_reads counts READS (the segments of a read are consecutive) */
            if (last_mapped_read != (int64_t)seg_read[s]) { mapped++; last_mapped_read = (int64_t)seg_read[s]; }
            so += lists[2 * s].size(); sr += lists[2 * s + 1].size();
        }
    }
    cand_off[2 * n_seg] = pos;
    for (uint64_t a : anded) tot += a;
    if (stats) { stats[0] = mapped; stats[1] = so; stats[2] = sr; stats[3] = tot; }
}

void orc_map_segments(const orc_params* p, uint32_t NB, const int32_t* k2i, const uint8_t* rows, uint32_t n_rows,
                      const uint8_t* seq, const uint8_t* qual, const uint64_t* read_off, const uint32_t* seg_read,
                      const uint32_t* seg_start, const uint32_t* seg_len, uint32_t n_seg, uint32_t* cand_off,
                      uint32_t* cand, uint64_t* stats, uint32_t threads) {
    orc_index* ix = orc_index_open(p, NB, k2i, rows, n_rows);
    orc_map_segments_h(ix, p, seq, qual, read_off, seg_read, seg_start, seg_len, n_seg, cand_off, cand, stats, threads);
    orc_index_close(ix);
}

void orc_prepare_query(const orc_params* p, const uint8_t* seq, const uint8_t* qual, const uint64_t* read_off,
                       const uint32_t* seg_read, const uint32_t* seg_start, const uint32_t* seg_len, uint32_t n_seg,
                       uint16_t* samp_idx, uint32_t* samp_hash) {
    const uint32_t P = p->locator_samples, minq = p->avg_base_quality * p->k;  /* bucket_locator.h:431 */
    std::vector<uint32_t> hash, qsum, samp; std::vector<uint16_t> good;
    for (uint32_t s = 0; s < n_seg; s++) {
        uint64_t o = read_off[seg_read[s]] + seg_start[s];
        SegView v{seq + o, qual + o, seg_len[s]};
        kmers_and_quals(v, p->k, hash, qsum);
        good.clear();
        for (uint32_t j = 0; j < hash.size(); j++) if (qsum[j] >= minq) good.push_back((uint16_t)j);
        if (good.empty()) for (uint32_t j = 0; j < hash.size(); j++) good.push_back((uint16_t)j);
        if (good.empty()) {  /* segment shorter than k: reference indexes out of range (UB); defined as no samples */
            for (uint32_t i = 0; i < P; i++) { samp_idx[(size_t)s * P + i] = 0; samp_hash[(size_t)s * P + i] = 0xFFFFFFFFu; }
            continue;
        }
        sample_det(P, (uint32_t)good.size() - 1u, samp);
        for (uint32_t i = 0; i < P; i++) {
            uint16_t j = good[samp[i]];
            samp_idx[(size_t)s * P + i] = j; samp_hash[(size_t)s * P + i] = hash[j];
        }
    }
}

void orc_locate(const orc_params* p, const uint8_t* packed, const uint64_t* b_start, const uint32_t* b_len,
                uint32_t n_emit, const uint8_t* seq, const uint8_t* qual, const uint64_t* read_off, uint32_t n_reads,
                const uint32_t* seg_read, const uint32_t* seg_start, const uint32_t* seg_len, uint32_t n_seg,
                const uint32_t* cand_off, const uint32_t* cand, uint32_t* loc_off, orc_loc* locs, uint32_t threads) {
    const uint32_t P = p->locator_samples, k = p->k;
    std::vector<uint16_t> sidx((size_t)n_seg * P); std::vector<uint32_t> shash((size_t)n_seg * P);
    orc_prepare_query(p, seq, qual, read_off, seg_read, seg_start, seg_len, n_seg, sidx.data(), shash.data());
    /* bucket-major inverted lists, segment order within bucket (q_gram_mapper.h:527-533) */
    std::vector<std::vector<uint32_t>> fo(n_emit), fr(n_emit);
    for (uint32_t s = 0; s < n_seg; s++) {
        for (uint32_t c = cand_off[2 * s]; c < cand_off[2 * s + 1]; c++) if (cand[c] < n_emit) fo[cand[c]].push_back(s);
        for (uint32_t c = cand_off[2 * s + 1]; c < cand_off[2 * s + 2]; c++) if (cand[c] < n_emit) fr[cand[c]].push_back(s);
    }
    if (threads == 0) threads = 1;
    std::vector<std::vector<std::pair<uint32_t, orc_loc>>> out(threads);
    parallel_for(n_emit, threads, [&](uint64_t lo, uint64_t hi, uint32_t t) {
        std::unordered_multimap<unsigned int, int> index;
        std::map<int, unsigned int> votes;
        const uint64_t mask = (k >= 16) ? 0xFFFFFFFFull : ((1ull << (2 * k)) - 1);
        for (uint64_t b = lo; b < hi; b++) {
            if (fo[b].empty() && fr[b].empty()) continue;           /* bucket_locator.h:657-659 */
            index.clear(); index.reserve(p->bucket_len);            /* :168-169 */
            uint64_t h = 0; int offset = 0;
            for (uint32_t i = 0; i < b_len[b]; i++) {
                h = ((h << 2) | gbase(packed, b_start[b] + i)) & mask;
                if (i + 1 >= k) { index.emplace((unsigned int)h, offset); offset++; }  /* :172-176 */
            }
            for (uint32_t s : fo[b]) {                              /* :671-680 */
                auto r = find_offset(index, p, &shash[(size_t)s * P], &sidx[(size_t)s * P], seg_len[s], false, votes);
                if (r.first > 0)
                    out[t].push_back({seg_read[s], orc_loc{(uint32_t)b, r.first - (int)seg_start[s], seg_start[s], r.second, 1u}});
            }
            for (size_t z = fr[b].size(); z-- > 0;) {               /* :683-693, reversed list order */
                uint32_t s = fr[b][z];
                auto r = find_offset(index, p, &shash[(size_t)s * P], &sidx[(size_t)s * P], seg_len[s], true, votes);
                if (r.first > 0) {
                    uint32_t rl = (uint32_t)(read_off[seg_read[s] + 1] - read_off[seg_read[s]]);
                    int so = (int)(rl - seg_start[s] - seg_len[s]);
                    out[t].push_back({seg_read[s], orc_loc{(uint32_t)b, r.first - so, seg_start[s], r.second, 0u}});
                }
            }
        }
    });
    std::vector<uint32_t> cnt(n_reads + 1, 0);
    for (auto& v : out) for (auto& e : v) cnt[e.first + 1]++;
    for (uint32_t i = 0; i < n_reads; i++) cnt[i + 1] += cnt[i];
    memcpy(loc_off, cnt.data(), (n_reads + 1) * sizeof(uint32_t));
    std::vector<uint32_t> cur(cnt.begin(), cnt.end() - 1);
    for (auto& v : out) for (auto& e : v) locs[cur[e.first]++] = e.second;   /* thread order == bucket order */
}

void orc_filter_best(const orc_params* p, const uint64_t* read_off, uint32_t n_reads, const uint32_t* loc_off,
                     const orc_loc* locs, uint32_t* out_off, orc_loc* out) {
    uint32_t pos = 0;
    for (uint32_t i = 0; i < n_reads; i++) {
        out_off[i] = pos;
        unsigned int read_length = (unsigned int)(read_off[i + 1] - read_off[i]);
        std::map<std::tuple<unsigned int, int, bool>, unsigned int> lv;
        for (uint32_t c = loc_off[i]; c < loc_off[i + 1]; c++) {
            const orc_loc& l = locs[c];
            bool orig = l.is_original != 0;
            if (lv.empty()) { lv[{l.bucket, l.offset, orig}] = l.votes; continue; }
            bool found = false;
            int lower = l.offset - read_length * p->indel_rate;     /* bucket_locator.h:365-366: float, truncated */
            int upper = l.offset + read_length * p->indel_rate;
            for (auto it = lv.begin(); it != lv.end(); ++it) {
                int po = std::get<1>(it->first);
                if (l.bucket == std::get<0>(it->first) && po <= upper && po >= lower && std::get<2>(it->first) == orig) {
                    it->second += l.votes; found = true;
                }
            }
            if (!found) lv[{l.bucket, l.offset, orig}] = l.votes;
        }
        unsigned int mx = 0; uint32_t first = pos;
        for (auto it = lv.begin(); it != lv.end(); ++it) {
            if (it->second > mx) { pos = first; mx = it->second; }
            if (it->second == mx)
                out[pos++] = orc_loc{std::get<0>(it->first), std::get<1>(it->first), 0u, it->second, std::get<2>(it->first) ? 1u : 0u};
        }
    }
    out_off[n_reads] = pos;
}

uint64_t orc_align(const orc_params* p, const uint8_t* packed, const uint64_t* b_start, const uint32_t* b_len,
                   const uint8_t* seq, const uint64_t* read_off, uint32_t n_reads, const uint32_t* loc_off,
                   const orc_loc* locs, orc_aln* aln, uint32_t* cigar, uint64_t cigar_cap, uint32_t threads) {
    if (threads == 0) threads = 1;
    std::vector<std::vector<std::vector<uint32_t>>> tops(threads);
    std::vector<uint64_t> tlo(threads + 1, 0);
    parallel_for(n_reads, threads, [&](uint64_t lo, uint64_t hi, uint32_t t) {
        std::vector<uint8_t> text, query; std::vector<uint32_t> ops;
        tlo[t] = loc_off[lo];
        for (uint64_t i = lo; i < hi; i++) {
            uint32_t m = (uint32_t)(read_off[i + 1] - read_off[i]);
            query.resize(m);
            for (uint32_t j = 0; j < m; j++) query[j] = rank_dna4(seq[read_off[i] + j]);
            for (uint32_t c = loc_off[i]; c < loc_off[i + 1]; c++) {
                const orc_loc& l = locs[c];
                size_t width = std::min((size_t)m + 1 + (std::size_t)(p->indel_rate * m),      /* :550 */
                                        (size_t)b_len[l.bucket] - (size_t)(int64_t)l.offset);
                text.resize(width);
                for (size_t j = 0; j < width; j++) text[j] = (uint8_t)gbase(packed, b_start[l.bucket] + l.offset + j);
                if (!l.is_original) {                               /* :563-567 */
                    std::reverse(text.begin(), text.end());
                    for (auto& x : text) x = 3 - x;
                }
                orc_aln a{};
                align_semi_global(text.data(), (uint32_t)width, query.data(), m, a.score, a.begin, ops);
                a.map_qual = 60u + (uint32_t)a.score;               /* :570, unsigned 32-bit wrap */
                a.kept = ((size_t)a.map_qual >= (size_t)p->quality_threshold) ? 1u : 0u;
                a.cigar_len = (uint32_t)ops.size();
                aln[c] = a;
                tops[t].push_back(ops);
            }
        }
    });
    uint64_t pos = 0;
    for (uint32_t t = 0; t < threads; t++) {
        uint64_t c = tlo[t];
        for (auto& ops : tops[t]) {
            aln[c].cigar_off = (uint32_t)pos;
            for (uint32_t o : ops) { if (pos < cigar_cap) cigar[pos] = o; pos++; }
            c++;
        }
    }
    return pos;
}

/* ------------------------------------------------------------------ CLI flows ----------------------------- */

int orc_cli_index(const char* fasta, uint32_t L, uint32_t r, uint32_t q, float frac, uint64_t x, uint64_t y,
                  const char* index_dir, const char* indicator, uint32_t threads) {
    namespace fs = std::filesystem;
    fs::path dir(index_dir); std::string ind(indicator);
    fs::create_directories(dir);
    if (fs::exists(dir / (ind + ".qgram"))) {                       /* bucket_indexer.h:182-184 */
        fprintf(stderr, "[ERROR]\t\tThe specified file already exists in directory: %s.\n", (dir / (ind + ".qgram")).c_str());
        return 0;
    }
    Fasta g;
    if (!read_fasta(fasta, g)) return -1;
    uint32_t NB = orc_num_buckets(g.len.data(), (uint32_t)g.len.size(), L);
    std::vector<uint64_t> bs(NB + 1); std::vector<uint32_t> bl(NB + 1), br(NB + 1);
    uint32_t n_emit = orc_bucketize(g.len.data(), (uint32_t)g.len.size(), L, r, bs.data(), bl.data(), br.data());
    std::vector<int32_t> k2i(1ull << (2 * q));
    uint32_t n_rows = orc_fracminhash(q, x, y, frac, k2i.data());
    uint32_t W = (NB + 7) >> 3;
    std::vector<uint8_t> rows((size_t)n_rows * W);
    orc_index_build(g.packed.data(), bs.data(), bl.data(), n_emit, NB, q, k2i.data(), n_rows, rows.data(), threads);
    { std::ofstream o(dir / (ind + ".qgram"), std::ios::binary); o.write((const char*)rows.data(), rows.size()); }
    { std::ofstream o(dir / (ind + ".bucket_id")); for (uint32_t b = 0; b < n_emit; b++) o << g.ids[br[b]] << "\n"; }
    { std::ofstream o(dir / (ind + ".kmers_index")); for (int32_t v : k2i) o << v << "\n"; }
    return (int)n_emit;
}

int orc_cli_map(const char* fasta, const orc_params* p, const char* index_dir, const char* indicator,
                const char* fastq, const char* sam, int align, uint32_t threads) {
    namespace fs = std::filesystem;
    fs::path dir(index_dir); std::string ind(indicator);
    Fasta g;
    if (!read_fasta(fasta, g)) return -1;
    const uint32_t L = p->bucket_len;
    uint32_t NB = orc_num_buckets(g.len.data(), (uint32_t)g.len.size(), L);
    std::vector<uint64_t> bs(NB + 1); std::vector<uint32_t> bl(NB + 1), br(NB + 1);
    uint32_t n_emit = orc_bucketize(g.len.data(), (uint32_t)g.len.size(), L, p->read_len, bs.data(), bl.data(), br.data());
    /* q_gram_mapper.h:318-372 */
    std::vector<int32_t> k2i(1ull << (2 * p->q), -1);
    uint32_t n_rows = 0;
    { std::ifstream in(dir / (ind + ".kmers_index")); if (!in) return -2;
      for (auto& v : k2i) { in >> v; if (v >= 0) n_rows++; } }
    uint32_t W = (NB + 7) >> 3;
    std::vector<uint8_t> rows((size_t)n_rows * W);
    { std::ifstream in(dir / (ind + ".qgram"), std::ios::binary); if (!in) return -3; in.read((char*)rows.data(), rows.size()); }
    Fastq f;
    if (!read_fastq(fastq, f)) return -4;
    uint32_t n_reads = (uint32_t)f.ids.size();
    std::vector<uint32_t> sr((size_t)n_reads * std::max(1u, p->num_segment_samples)), ss(sr.size()), sl(sr.size());
    uint32_t n_seg = orc_segments(p, f.off.data(), n_reads, sr.data(), ss.data(), sl.data());
    std::vector<uint32_t> coff(2 * (size_t)n_seg + 1), cand((size_t)2 * p->max_candidates * n_seg + 1);
    uint64_t stats[4];
    orc_map_segments(p, NB, k2i.data(), rows.data(), n_rows, f.seq.data(), f.qual.data(), f.off.data(), sr.data(),
                     ss.data(), sl.data(), n_seg, coff.data(), cand.data(), stats, threads);
    std::vector<uint32_t> loff(n_reads + 1); std::vector<orc_loc> locs(coff[2 * n_seg] + 1);
    orc_locate(p, g.packed.data(), bs.data(), bl.data(), n_emit, f.seq.data(), f.qual.data(), f.off.data(), n_reads,
               sr.data(), ss.data(), sl.data(), n_seg, coff.data(), cand.data(), loff.data(), locs.data(), threads);
    /* bucket_locator.h:473-503 */
    std::vector<std::string> bname; std::vector<unsigned int> boff;
    std::vector<std::string> ref_ids; std::vector<size_t> ref_len;
    { std::ifstream in(dir / (ind + ".bucket_id")); std::string name, last; unsigned int bi = 0;
      for (uint32_t i = 0; i < n_emit; i++) {
          std::getline(in, name); name = name.substr(0, name.find(' '));
          if (name != last) { if (bi != 0) { ref_ids.push_back(last); ref_len.push_back((size_t)bi * L); } last = name; bi = 0; }
          bname.push_back(name); boff.push_back(bi * L); bi++;
      }
      if (bi != 0) { ref_ids.push_back(last); ref_len.push_back((size_t)bi * L); } }
    std::ofstream out(sam);
    out << "@HD\tVN:1.6\n";
    for (size_t i = 0; i < ref_ids.size(); i++) out << "@SQ\tSN:" << ref_ids[i] << "\tLN:" << ref_len[i] << "\n";
    std::vector<uint32_t> foff(n_reads + 1); std::vector<orc_loc> best(locs.size());
    std::vector<orc_aln> aln; std::vector<uint32_t> cig;
    const uint32_t* use_off = loff.data(); const orc_loc* use = locs.data();
    if (!align) {
        orc_filter_best(p, f.off.data(), n_reads, loff.data(), locs.data(), foff.data(), best.data());
        use_off = foff.data(); use = best.data();
    } else {
        aln.resize(locs.size()); cig.resize((size_t)locs.size() * 64 + 1);
        uint64_t need = orc_align(p, g.packed.data(), bs.data(), bl.data(), f.seq.data(), f.off.data(), n_reads,
                                  loff.data(), locs.data(), aln.data(), cig.data(), cig.size(), threads);
        if (need > cig.size()) {
            cig.resize(need);
            orc_align(p, g.packed.data(), bs.data(), bl.data(), f.seq.data(), f.off.data(), n_reads, loff.data(),
                      locs.data(), aln.data(), cig.data(), cig.size(), threads);
        }
    }
    for (uint32_t i = 0; i < n_reads; i++) {
        std::string s((const char*)&f.seq[f.off[i]], f.off[i + 1] - f.off[i]);
        for (auto& c : s) c = "ACGT"[rank_dna4((uint8_t)c)];           /* dna4 round trip of the FASTQ reader */
        std::string qs((const char*)&f.qual[f.off[i]], f.off[i + 1] - f.off[i]);
        for (uint32_t c = use_off[i]; c < use_off[i + 1]; c++) {
            const orc_loc& l = use[c];
            unsigned flag = l.is_original ? 0 : 16;
            if (!align) {
                size_t mq = std::min((unsigned int)60, 6 * l.votes);   /* :591 */
                size_t pos0 = (unsigned int)(boff[l.bucket] + l.offset);  /* :592 */
                out << f.ids[i] << '\t' << flag << '\t' << bname[l.bucket] << '\t' << (pos0 + 1) << '\t' << mq
                    << "\t*\t*\t0\t0\t" << s << '\t' << qs << '\n';
            } else {
                const orc_aln& a = aln[c];
                if (!a.kept) continue;
                size_t pos0 = (unsigned int)(a.begin + boff[l.bucket] + l.offset);  /* :576 */
                std::string cg;
                for (uint32_t t = 0; t < a.cigar_len; t++) {
                    uint32_t o = cig[a.cigar_off + t];
                    cg += std::to_string(o >> 4); cg += "MID"[o & 15u];
                }
                out << f.ids[i] << '\t' << flag << '\t' << bname[l.bucket] << '\t' << (pos0 + 1) << '\t'
                    << (unsigned)(uint8_t)a.map_qual << '\t' << cg << "\t*\t0\t0\t" << s << '\t' << qs << '\n';
            }
        }
    }
    return 0;
}

}  // extern "C"
