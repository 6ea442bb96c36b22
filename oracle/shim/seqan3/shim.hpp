// oracle/shim/seqan3/shim.hpp -- TEST INFRASTRUCTURE ONLY.
//
// A minimal stand-in for the parts of SeqAn3 (un-vendored, unpinned `main` dependency of GZHoffie/bucket-map,
// bucket_map/CMakeLists.txt:69-80) that the reference's own headers use, so that /root/reference/bucket_map/main.cpp
// can be compiled UNMODIFIED here into oracle/_ref/bucketmap[_align] (see oracle/Makefile).  Every result-defining
// piece of the path (std::bitset cascade, unordered_multimap order, std::map vote merges, float thresholds, Sampler)
// is then the reference's own code; this file only supplies the library boundary it sits on:
//   alphabets (dna4/dna5/phred94), bitpacked_sequence, shape/ungapped, views::kmer_hash, FASTA/FASTQ input,
//   SAM output, cigar, semi-global edit-distance alignment, debug_stream.
// What is restated from memory of SeqAn3 and therefore NOT pinned by the reference: IUPAC->dna4 conversion, SAM
// text conventions, the edit-distance end-column / traceback tie-breaks (documented at align_pairwise below).
// Nothing under bucket-map_b200/ includes this.
#pragma once
#include <algorithm>
#include <array>
#include <bitset>
#include <chrono>
#include <cstdlib>
#include <ctime>
#include <random>
#include <span>
#include <sstream>
#include <cassert>
#include <cmath>
#include <compare>
#include <concepts>
#include <cstdint>
#include <filesystem>
#include <fstream>
#include <functional>
#include <iostream>
#include <iterator>
#include <map>
#include <ranges>
#include <stdexcept>
#include <string>
#include <tuple>
#include <type_traits>
#include <unordered_map>
#include <unordered_set>
#include <utility>
#include <vector>

#define SEQAN3_DOXYGEN_ONLY(x)

namespace seqan3 {

// ---- alphabets ---------------------------------------------------------------------------------------------------
struct dna5;
struct dna4 {
    uint8_t r = 0;
    constexpr dna4() = default;
    explicit constexpr dna4(dna5 const& o);
    static constexpr uint8_t alphabet_size = 4;
    constexpr uint8_t to_rank() const { return r; }
    constexpr char to_char() const { return "ACGT"[r]; }
    constexpr dna4& assign_rank(uint8_t v) { r = v; return *this; }
    // IUPAC codes collapse to the first base they can stand for (seqan3 dna4 conversion table; N -> A)
    constexpr dna4& assign_char(char c) {
        switch (c & 0xDF) {
            case 'C': case 'Y': case 'S': case 'B': r = 1; break;
            case 'G': case 'K': r = 2; break;
            case 'T': case 'U': r = 3; break;
            default: r = 0;
        }
        return *this;
    }
    constexpr dna4 complement() const { dna4 x; x.r = 3 - r; return x; }
    friend constexpr auto operator<=>(dna4 const&, dna4 const&) = default;
};
struct dna5 {
    uint8_t r = 0;   // A C G N T
    constexpr dna5() = default;
    static constexpr uint8_t alphabet_size = 5;
    constexpr uint8_t to_rank() const { return r; }
    constexpr char to_char() const { return "ACGNT"[r]; }
    constexpr dna5& assign_rank(uint8_t v) { r = v; return *this; }
    constexpr dna5& assign_char(char c) {
        switch (c & 0xDF) {
            case 'A': r = 0; break;
            case 'C': r = 1; break;
            case 'G': r = 2; break;
            case 'T': case 'U': r = 4; break;
            default: r = 3;
        }
        return *this;
    }
    friend constexpr auto operator<=>(dna5 const&, dna5 const&) = default;
};
constexpr dna4::dna4(dna5 const& o) { assign_char(o.to_char()); }

template <uint8_t MAXQ>
struct phred_base {
    uint8_t r = 0;
    static constexpr uint8_t alphabet_size = MAXQ;
    constexpr uint8_t to_rank() const { return r; }
    constexpr char to_char() const { return (char)(r + 33); }
    constexpr phred_base& assign_rank(uint8_t v) { r = v; return *this; }
    constexpr phred_base& assign_char(char c) {
        int v = (int)(unsigned char)c - 33;
        r = (uint8_t)(v < 0 ? 0 : (v >= MAXQ ? MAXQ - 1 : v));
        return *this;
    }
    friend constexpr auto operator<=>(phred_base const&, phred_base const&) = default;
};
using phred94 = phred_base<94>;
using phred42 = phred_base<42>;
using phred63 = phred_base<63>;

using dna4_vector = std::vector<dna4>;
using dna5_vector = std::vector<dna5>;

constexpr dna4 operator""_dna4(char c) { return dna4{}.assign_char(c); }
constexpr dna5 operator""_dna5(char c) { return dna5{}.assign_char(c); }
constexpr phred42 operator""_phred42(char c) { return phred42{}.assign_char(c); }
inline dna4_vector operator""_dna4(char const* s, std::size_t n) {
    dna4_vector v(n);
    for (std::size_t i = 0; i < n; i++) v[i].assign_char(s[i]);
    return v;
}

template <class T>
concept semialphabet = requires(std::remove_cvref_t<T> const& t) {
    { t.to_rank() };
    { std::remove_cvref_t<T>::alphabet_size };
};
template <class T>
inline constexpr auto alphabet_size = std::remove_cvref_t<T>::alphabet_size;
template <class T>
    requires semialphabet<T>
constexpr auto to_rank(T const& t) { return t.to_rank(); }
template <class T>
    requires requires(T const& t) { t.to_char(); }
constexpr char to_char(T const& t) { return t.to_char(); }

// ---- bitpacked_sequence<dna4>: 2 bits per base ---------------------------------------------------------------------
template <class alph_t>
class bitpacked_sequence {
    std::vector<uint64_t> w;
    std::size_t n = 0;

public:
    using value_type = alph_t;
    using size_type = std::size_t;
    class const_iterator {
        bitpacked_sequence const* s = nullptr;
        std::ptrdiff_t i = 0;

    public:
        using iterator_category = std::random_access_iterator_tag;
        using iterator_concept = std::random_access_iterator_tag;
        using value_type = alph_t;
        using difference_type = std::ptrdiff_t;
        using pointer = void;
        using reference = alph_t;
        const_iterator() = default;
        const_iterator(bitpacked_sequence const* s_, std::ptrdiff_t i_) : s(s_), i(i_) {}
        reference operator*() const { return (*s)[(std::size_t)i]; }
        reference operator[](difference_type d) const { return (*s)[(std::size_t)(i + d)]; }
        const_iterator& operator++() { ++i; return *this; }
        const_iterator operator++(int) { auto t = *this; ++i; return t; }
        const_iterator& operator--() { --i; return *this; }
        const_iterator operator--(int) { auto t = *this; --i; return t; }
        const_iterator& operator+=(difference_type d) { i += d; return *this; }
        const_iterator& operator-=(difference_type d) { i -= d; return *this; }
        friend const_iterator operator+(const_iterator a, difference_type d) { a.i += d; return a; }
        friend const_iterator operator+(difference_type d, const_iterator a) { a.i += d; return a; }
        friend const_iterator operator-(const_iterator a, difference_type d) { a.i -= d; return a; }
        friend difference_type operator-(const_iterator const& a, const_iterator const& b) { return a.i - b.i; }
        friend bool operator==(const_iterator const& a, const_iterator const& b) { return a.i == b.i; }
        friend auto operator<=>(const_iterator const& a, const_iterator const& b) { return a.i <=> b.i; }
    };
    using iterator = const_iterator;

    bitpacked_sequence() = default;
    template <std::ranges::input_range rng_t>
        requires (!std::same_as<std::remove_cvref_t<rng_t>, bitpacked_sequence>)
    explicit bitpacked_sequence(rng_t&& r) {
        for (auto&& c : r) push_back(alph_t{c});
    }
    void push_back(alph_t c) {
        if ((n & 31) == 0) w.push_back(0);
        w[n >> 5] |= (uint64_t)c.to_rank() << ((n & 31) << 1);
        n++;
    }
    alph_t operator[](std::size_t i) const {
        alph_t a;
        a.assign_rank((uint8_t)((w[i >> 5] >> ((i & 31) << 1)) & 3u));
        return a;
    }
    std::size_t size() const { return n; }
    bool empty() const { return n == 0; }
    const_iterator begin() const { return {this, 0}; }
    const_iterator end() const { return {this, (std::ptrdiff_t)n}; }
    void clear() { w.clear(); n = 0; }
};

// ---- shape ---------------------------------------------------------------------------------------------------------
struct ungapped { uint8_t value; };
struct bin_literal { uint64_t value; };

class shape {
    std::vector<uint8_t> bits;   // bits[i] = position i takes part (shape is read right to left; irrelevant for ungapped)

public:
    shape() = default;
    shape(ungapped u) : bits(u.value, 1) {}
    shape(bin_literal b) {
        uint64_t v = b.value;
        while (v) { bits.push_back((uint8_t)(v & 1)); v >>= 1; }
        std::reverse(bits.begin(), bits.end());
    }
    std::size_t size() const { return bits.size(); }
    bool all() const { return std::all_of(bits.begin(), bits.end(), [](uint8_t b) { return b != 0; }); }
    std::size_t count() const { return (std::size_t)std::count(bits.begin(), bits.end(), 1); }
    bool operator[](std::size_t i) const { return bits[i] != 0; }
    auto begin() const { return bits.begin(); }
    auto end() const { return bits.end(); }
    friend bool operator==(shape const& a, shape const& b) { return a.bits == b.bits; }
    friend auto operator<=>(shape const& a, shape const& b) { return a.bits <=> b.bits; }
};
inline shape operator""_shape(char const* s, std::size_t n) {
    uint64_t v = 0;
    for (std::size_t i = 0; i < n; i++) v = (v << 1) | (uint64_t)(s[i] == '1');
    return shape{bin_literal{v}};
}

// ---- views::kmer_hash ------------------------------------------------------------------------------------------------
// seqan3 views::kmer_hash for an ungapped shape: hash = sum_i rank(text[j+i]) * sigma^(k-1-i)  (first base most
// significant; the arithmetic is the one printed in the reference's own cachegrind log of seqan3's kmer_hash.hpp).
// Range length max(n+1,k)-k.  Iteration rolls; operator[] recomputes the full hash (as seqan3's random access does).
namespace detail {
template <class rng_t>
class kmer_hash_range {
    rng_t const* r = nullptr;
    std::size_t k = 0, n = 0;

    std::size_t full(std::size_t j) const {
        std::size_t h = 0;
        constexpr std::size_t sigma = alphabet_size<std::ranges::range_value_t<rng_t>>;
        auto it = std::ranges::begin(*r) + (std::ptrdiff_t)j;
        for (std::size_t i = 0; i < k; i++, ++it) h = h * sigma + (std::size_t)(*it).to_rank();
        return h;
    }

public:
    class iterator {
        kmer_hash_range const* p = nullptr;
        std::size_t j = 0, h = 0, top = 1;

    public:
        using iterator_category = std::input_iterator_tag;
        using iterator_concept = std::forward_iterator_tag;
        using value_type = std::size_t;
        using difference_type = std::ptrdiff_t;
        using pointer = void;
        using reference = std::size_t;
        iterator() = default;
        iterator(kmer_hash_range const* p_, std::size_t j_) : p(p_), j(j_) {
            constexpr std::size_t sigma = alphabet_size<std::ranges::range_value_t<rng_t>>;
            for (std::size_t i = 1; i < p->k; i++) top *= sigma;
            if (j < p->size()) h = p->full(j);
        }
        std::size_t operator*() const { return h; }
        iterator& operator++() {
            constexpr std::size_t sigma = alphabet_size<std::ranges::range_value_t<rng_t>>;
            ++j;
            if (j < p->size()) {
                auto b = std::ranges::begin(*p->r);
                h = (h - top * (std::size_t)(*(b + (std::ptrdiff_t)(j - 1))).to_rank()) * sigma +
                    (std::size_t)(*(b + (std::ptrdiff_t)(j + p->k - 1))).to_rank();
            }
            return *this;
        }
        iterator operator++(int) { auto t = *this; ++*this; return t; }
        friend bool operator==(iterator const& a, iterator const& b) { return a.j == b.j; }
    };
    kmer_hash_range() = default;
    kmer_hash_range(rng_t const& r_, std::size_t k_) : r(&r_), k(k_), n(std::ranges::size(r_)) {}
    std::size_t size() const { return std::max<std::size_t>(n + 1, k) - k; }
    bool empty() const { return size() == 0; }
    std::size_t operator[](std::size_t j) const { return full(j); }
    iterator begin() const { return iterator{this, 0}; }
    iterator end() const { return iterator{this, size()}; }
};
struct kmer_hash_closure {
    shape s;
    template <class rng_t>
    friend auto operator|(rng_t const& r, kmer_hash_closure const& c) {
        if (!c.s.all()) throw std::invalid_argument("shim: only ungapped shapes are supported");
        return kmer_hash_range<rng_t>{r, c.s.size()};
    }
};
struct kmer_hash_fn {
    kmer_hash_closure operator()(shape const& s) const { return {s}; }
};
}   // namespace detail

namespace views {
inline constexpr detail::kmer_hash_fn kmer_hash{};
inline constexpr auto complement = std::views::transform([](auto const& c) { return c.complement(); });
}   // namespace views

// ---- seqan3::detail pieces the reference's own copy of the k-mer view (mapper/quality_filter.h) is written against -----
namespace detail {
template <class rng_t>
concept const_iterable_range =
    std::ranges::input_range<std::remove_const_t<rng_t>> && std::ranges::input_range<rng_t const>;

template <bool const_range, class rng_t>
using maybe_const_range_t = std::conditional_t<const_range, rng_t const, rng_t>;
template <bool const_range, class rng_t>
using maybe_const_iterator_t = std::ranges::iterator_t<maybe_const_range_t<const_range, rng_t>>;
template <bool const_range, class rng_t>
using maybe_const_sentinel_t = std::ranges::sentinel_t<maybe_const_range_t<const_range, rng_t>>;

template <class it_t>
struct maybe_iterator_category {};
template <class it_t>
    requires requires { typename std::iterator_traits<it_t>::iterator_category; }
struct maybe_iterator_category<it_t> {
    using iterator_category = typename std::iterator_traits<it_t>::iterator_category;
};

template <class it_t>
using iterator_concept_tag_t = std::conditional_t<
    std::contiguous_iterator<it_t>, std::contiguous_iterator_tag,
    std::conditional_t<std::random_access_iterator<it_t>, std::random_access_iterator_tag,
                       std::conditional_t<std::bidirectional_iterator<it_t>, std::bidirectional_iterator_tag,
                                          std::conditional_t<std::forward_iterator<it_t>, std::forward_iterator_tag,
                                                             std::input_iterator_tag>>>>;

// range adaptor closure built from a functor and its bound arguments:  range | adaptor  ==  functor(range, args...)
template <class functor_t, class... args_t>
class adaptor_from_functor {
    functor_t f;
    std::tuple<args_t...> args;

public:
    constexpr adaptor_from_functor(functor_t f_, args_t... a) : f(std::move(f_)), args(std::move(a)...) {}
    template <std::ranges::range rng_t>
    constexpr auto operator()(rng_t&& r) const {
        return std::apply([&](auto const&... a) { return f(std::forward<rng_t>(r), a...); }, args);
    }
    template <std::ranges::range rng_t>
    friend constexpr auto operator|(rng_t&& r, adaptor_from_functor const& me) {
        return me(std::forward<rng_t>(r));
    }
};
}   // namespace detail

// ---- debug_stream (stderr; integers of char width print as numbers like seqan3's) ---------------------------------------
class debug_stream_type {
public:
    template <class T>
    debug_stream_type& operator<<(T const& v) {
        if constexpr (std::is_same_v<T, uint8_t> || std::is_same_v<T, int8_t>) std::cerr << (int)v;
        else if constexpr (std::is_same_v<T, std::filesystem::path>) std::cerr << '"' << v.string() << '"';
        else if constexpr (requires { std::cerr << v; }) std::cerr << v;
        else std::cerr << "<?>";
        return *this;
    }
    debug_stream_type& operator<<(char const* s) { std::cerr << s; return *this; }
};
inline debug_stream_type debug_stream{};

// ---- FASTA / FASTQ input -------------------------------------------------------------------------------------------------
struct sequence_file_input_default_traits_dna {
    using sequence_alphabet = dna5;
    using quality_alphabet = phred42;
    template <class alph>
    using sequence_container = std::vector<alph>;
    template <class alph>
    using quality_container = std::vector<alph>;
};

template <class traits_t = sequence_file_input_default_traits_dna>
class sequence_file_input {
public:
    using seq_t = typename traits_t::template sequence_container<typename traits_t::sequence_alphabet>;
    using qual_t = std::vector<typename traits_t::quality_alphabet>;
    class record_type {
        friend class sequence_file_input;
        std::string id_;
        seq_t seq_;
        qual_t qual_;

    public:
        std::string& id() { return id_; }
        std::string const& id() const { return id_; }
        seq_t& sequence() { return seq_; }
        seq_t const& sequence() const { return seq_; }
        qual_t& base_qualities() { return qual_; }
        qual_t const& base_qualities() const { return qual_; }
    };

private:
    std::ifstream in;
    bool fastq = false, at_end = false;
    record_type rec;
    std::string line, pending;
    bool have_pending = false;

    bool next_line(std::string& s) {
        if (have_pending) { s = std::move(pending); have_pending = false; return true; }
        if (!std::getline(in, s)) return false;
        if (!s.empty() && s.back() == '\r') s.pop_back();
        return true;
    }
    void read_next() {
        rec.id_.clear(); rec.seq_.clear(); rec.qual_.clear();
        std::string s;
        do { if (!next_line(s)) { at_end = true; return; } } while (s.empty());
        typename traits_t::sequence_alphabet a;
        typename traits_t::quality_alphabet q;
        if (fastq) {
            if (s[0] != '@') throw std::runtime_error("shim: FASTQ record does not start with '@'");
            rec.id_ = s.substr(1);
            std::string bases;
            while (next_line(s) && (s.empty() || s[0] != '+')) bases += s;
            for (char c : bases) rec.seq_.push_back(a.assign_char(c));
            std::size_t need = bases.size();
            while (rec.qual_.size() < need && next_line(s))
                for (char c : s) rec.qual_.push_back(q.assign_char(c));
        } else {
            if (s[0] != '>') throw std::runtime_error("shim: FASTA record does not start with '>'");
            rec.id_ = s.substr(1);
            while (next_line(s)) {
                if (!s.empty() && s[0] == '>') { pending = s; have_pending = true; break; }
                for (char c : s)
                    if (!std::isspace((unsigned char)c)) rec.seq_.push_back(a.assign_char(c));
            }
        }
    }

public:
    explicit sequence_file_input(std::filesystem::path const& p) : in(p) {
        if (!in) throw std::runtime_error("shim: cannot open " + p.string());
        std::string ext = p.extension().string();
        fastq = (ext == ".fq" || ext == ".fastq");
    }
    sequence_file_input(sequence_file_input const&) = delete;
    struct sentinel {};
    class iterator {
        sequence_file_input* f;

    public:
        explicit iterator(sequence_file_input* f_) : f(f_) {}
        record_type& operator*() const { return f->rec; }
        iterator& operator++() { f->read_next(); return *this; }
        bool operator==(sentinel) const { return f->at_end; }
        bool operator!=(sentinel) const { return !f->at_end; }
    };
    iterator begin() { read_next(); return iterator{this}; }
    sentinel end() { return {}; }
};
sequence_file_input(std::filesystem::path const&) -> sequence_file_input<sequence_file_input_default_traits_dna>;

// ---- cigar ---------------------------------------------------------------------------------------------------------------
class cigar {
public:
    class operation {
        char c = 'M';

    public:
        constexpr operation() = default;
        constexpr operation& assign_char(char ch) { c = ch; return *this; }
        constexpr char to_char() const { return c; }
        friend constexpr bool operator==(operation const&, operation const&) = default;
    };
    uint32_t count = 0;
    operation op{};
    constexpr cigar() = default;
    constexpr cigar(uint32_t n, operation o) : count(n), op(o) {}
    friend constexpr bool operator==(cigar const&, cigar const&) = default;
};
constexpr cigar::operation operator""_cigar_operation(char c) { return cigar::operation{}.assign_char(c); }
template <std::size_t I>
constexpr auto get(cigar const& c) {
    if constexpr (I == 0) return c.count;
    else return c.op;
}

// ---- SAM output ------------------------------------------------------------------------------------------------------------
enum class sam_flag : uint16_t { none = 0, paired = 1, proper_pair = 2, unmapped = 4, mate_unmapped = 8, on_reverse_strand = 16 };
constexpr sam_flag operator|(sam_flag a, sam_flag b) { return (sam_flag)((uint16_t)a | (uint16_t)b); }
constexpr sam_flag& operator|=(sam_flag& a, sam_flag b) { a = a | b; return a; }

enum class field { seq, id, ref_id, ref_offset, cigar, qual, flag, mapq, offset, mate, tags, header_ptr };
template <field... fs>
struct fields {};

// Text conventions (restated from the SAM specification / seqan3's format_sam; not pinned by the reference):
// "@HD\tVN:1.6", one "@SQ\tSN:<id>\tLN:<len>" per reference, then QNAME FLAG RNAME POS(1-based) MAPQ CIGAR(* if
// none) RNEXT(*) PNEXT(0) TLEN(0) SEQ QUAL.
template <class fields_t>
class sam_file_output;
template <field... fs>
class sam_file_output<fields<fs...>> {
    std::ofstream out;
    std::unordered_set<std::string> known;

    struct rec_t {
        std::string seq = "*", id = "*", ref = "*", cig = "*", qual = "*";
        std::size_t pos = 0; bool has_pos = false;
        unsigned flag = 0, mapq = 0;
    };
    template <class rng_t>
    static std::string chars(rng_t const& r) {
        std::string s;
        for (auto const& c : r) s.push_back(c.to_char());
        return s.empty() ? std::string("*") : s;
    }
    template <field f, class T>
    void set(rec_t& r, T const& v) {
        if constexpr (f == field::seq) r.seq = chars(v);
        else if constexpr (f == field::id) r.id = v;
        else if constexpr (f == field::ref_id) r.ref = v;
        else if constexpr (f == field::ref_offset) { r.pos = (std::size_t)v; r.has_pos = true; }
        else if constexpr (f == field::cigar) {
            std::string s;
            for (auto const& c : v) { s += std::to_string(c.count); s.push_back(c.op.to_char()); }
            r.cig = s.empty() ? std::string("*") : s;
        }
        else if constexpr (f == field::qual) r.qual = chars(v);
        else if constexpr (f == field::flag) r.flag = (unsigned)v;
        else if constexpr (f == field::mapq) r.mapq = (unsigned)(uint8_t)v;      // seqan3's mapq field is uint8_t
    }

public:
    template <class ids_t, class lens_t>
    sam_file_output(std::filesystem::path const& p, ids_t const& ref_ids, lens_t const& ref_lengths, fields<fs...>) : out(p) {
        if (!out) throw std::runtime_error("shim: cannot open " + p.string());
        out << "@HD\tVN:1.6\n";
        auto li = std::ranges::begin(ref_lengths);
        for (auto const& id : ref_ids) { out << "@SQ\tSN:" << id << "\tLN:" << *li << "\n"; ++li; known.insert(id); }
    }
    template <class... args_t>
    void emplace_back(args_t const&... args) {
        static_assert(sizeof...(args_t) == sizeof...(fs));
        rec_t r;
        (set<fs>(r, args), ...);
        if (r.ref != "*" && !known.count(r.ref)) throw std::runtime_error("shim: unknown reference id " + r.ref);
        out << r.id << '\t' << r.flag << '\t' << r.ref << '\t' << (r.has_pos ? r.pos + 1 : 0) << '\t' << r.mapq << '\t'
            << r.cig << "\t*\t0\t0\t" << r.seq << '\t' << r.qual << '\n';
    }
};
template <class ids_t, class lens_t, field... fs>
sam_file_output(std::filesystem::path const&, ids_t const&, lens_t const&, fields<fs...>) -> sam_file_output<fields<fs...>>;

// ---- pairwise alignment: only what locator/bucket_locator.h:520-528,569-577 configures -----------------------------------
struct cfg_base {};   // lets ADL find the operator| overloads below for align_cfg types
namespace align_cfg {
struct free_end_gaps_sequence1_leading { bool v; };
struct free_end_gaps_sequence2_leading { bool v; };
struct free_end_gaps_sequence1_trailing { bool v; };
struct free_end_gaps_sequence2_trailing { bool v; };
struct method_global : cfg_base {
    bool s1_lead = false, s2_lead = false, s1_trail = false, s2_trail = false;
    constexpr method_global() = default;
    constexpr method_global(free_end_gaps_sequence1_leading a, free_end_gaps_sequence2_leading b,
                            free_end_gaps_sequence1_trailing c, free_end_gaps_sequence2_trailing d)
        : s1_lead(a.v), s2_lead(b.v), s1_trail(c.v), s2_trail(d.v) {}
};
struct edit_scheme_t : cfg_base {};
inline constexpr edit_scheme_t edit_scheme{};
struct output_alignment {};
struct output_begin_position {};
struct output_end_position {};
struct output_score {};
}   // namespace align_cfg

struct configuration {
    align_cfg::method_global method{};
    bool edit = false;
    constexpr configuration() = default;
    constexpr configuration(align_cfg::method_global m) : method(m) {}
};
constexpr configuration operator|(configuration c, align_cfg::edit_scheme_t) { c.edit = true; return c; }
constexpr configuration operator|(configuration c, align_cfg::output_alignment) { return c; }
constexpr configuration operator|(configuration c, align_cfg::output_begin_position) { return c; }
constexpr configuration operator|(configuration c, align_cfg::output_end_position) { return c; }
constexpr configuration operator|(configuration c, align_cfg::output_score) { return c; }
constexpr configuration operator|(align_cfg::method_global m, align_cfg::edit_scheme_t) { configuration c{m}; c.edit = true; return c; }

// one aligned pair: op[i] in {'M' (column with two bases), 'I' (base of sequence2 against a gap in sequence1),
// 'D' (base of sequence1 against a gap in sequence2)}
struct alignment_columns { std::string ops; };

class alignment_result {
public:
    int score_ = 0;
    std::size_t begin1 = 0, begin2 = 0, end1 = 0, end2 = 0;
    alignment_columns aln;
    int score() const { return score_; }
    std::size_t sequence1_begin_position() const { return begin1; }
    std::size_t sequence2_begin_position() const { return begin2; }
    std::size_t sequence1_end_position() const { return end1; }
    std::size_t sequence2_end_position() const { return end2; }
    alignment_columns const& alignment() const { return aln; }
};

// Unit-cost edit distance, sequence2 (the read) aligned end to end, leading and trailing gaps in sequence1 (the text
// window) free -- the only configuration the reference uses.  Conventions where SeqAn3's source would decide
// (edit_distance_unbanded + its trace matrix; restated from memory, unpinned):
//   * score = -(minimum over all end columns j of D[m][j]); among equal minima the LAST column wins (SeqAn3 updates
//     its best column with `score <= best`);
//   * traceback from (m, j*) prefers diagonal, then up (consumes a read base: gap in the text = 'I'), then left
//     (consumes a text base: gap in the read = 'D'); it stops at row 0, whose column is sequence1_begin_position.
template <class seq1_t, class seq2_t>
std::vector<alignment_result> align_pairwise(std::tuple<seq1_t&, seq2_t&> const& seqs, configuration const& cfg) {
    auto const& t = std::get<0>(seqs);
    auto const& q = std::get<1>(seqs);
    if (!(cfg.edit && cfg.method.s1_lead && cfg.method.s1_trail && !cfg.method.s2_lead && !cfg.method.s2_trail))
        throw std::invalid_argument("shim: align_pairwise supports only the semi-global edit configuration of bucket_locator.h");
    const std::size_t n = std::ranges::size(t), m = std::ranges::size(q);
    std::vector<uint16_t> D((n + 1) * (m + 1));
    auto at = [&](std::size_t i, std::size_t j) -> uint16_t& { return D[i * (n + 1) + j]; };
    for (std::size_t j = 0; j <= n; j++) at(0, j) = 0;
    for (std::size_t i = 1; i <= m; i++) {
        at(i, 0) = (uint16_t)i;
        for (std::size_t j = 1; j <= n; j++) {
            unsigned d = at(i - 1, j - 1) + (q[i - 1] == t[j - 1] ? 0u : 1u);
            unsigned u = at(i - 1, j) + 1u, l = at(i, j - 1) + 1u;
            at(i, j) = (uint16_t)std::min(d, std::min(u, l));
        }
    }
    std::size_t bj = 0;
    unsigned best = at(m, 0);
    for (std::size_t j = 1; j <= n; j++)
        if (at(m, j) <= best) { best = at(m, j); bj = j; }
    alignment_result r;
    r.score_ = -(int)best;
    r.end1 = bj; r.end2 = m;
    std::size_t i = m, j = bj;
    std::string ops;
    while (i > 0) {
        if (j > 0 && at(i, j) == at(i - 1, j - 1) + (q[i - 1] == t[j - 1] ? 0u : 1u)) { ops.push_back('M'); i--; j--; }
        else if (at(i, j) == at(i - 1, j) + 1u) { ops.push_back('I'); i--; }
        else { ops.push_back('D'); j--; }
    }
    std::reverse(ops.begin(), ops.end());
    r.begin1 = j; r.begin2 = 0;
    r.aln.ops = std::move(ops);
    return {r};
}

inline std::vector<cigar> cigar_from_alignment(alignment_columns const& a) {
    std::vector<cigar> out;
    for (char c : a.ops) {
        if (!out.empty() && out.back().op.to_char() == c) out.back().count++;
        else out.emplace_back(1u, cigar::operation{}.assign_char(c));
    }
    return out;
}

// ---- never-instantiated leftovers (indexer/fm_indexer.h, bucket_fm_indexer.h: dead code that must still parse) ---------
class bi_fm_index {
public:
    template <class T>
    explicit bi_fm_index(T const&) {}
};

}   // namespace seqan3
