// bucketmap_analyze: accuracy of a SAM file against a ground truth -- the evaluation tool of SURVEY 8(f) rank 3.
//
// Restates the metrics of the reference's benchmark/sam_file_analyzer.cpp (class sam_analyzer, :6-358; the numbers of
// the README accuracy table come out of its benchmark(), :250-358) as a command-line tool: the reference hard-codes its
// input paths in main() (:371-383), here they are arguments.  Host-only evaluation tooling: it reads text files and
// never touches the GPU path or its results.
//
//   bucketmap_analyze --reads R.fastq [--dwgsim --genome G.fa] [--truth T] [--best GOOD.sam] [--tolerance 10]
//                     [--json] OUT.sam|DIR ...
//
//   --truth   `<origin> <position> <reverse> <cigar>` per read in file order (what tools/short_read_simulator.h:237-240
//             writes as .position_ground_truth; the reference tests for the extension ".ground_truth", :133, which that
//             name does not have -- both are accepted here), or a pbsim .maf (:147-176)
//   --best    another mapper's SAM as the answer (:86-123);  --dwgsim: the answer is in the read names (:201-226)
//
// Kept exactly: read names are compared after ' ' -> '_' and with everything from the first '/' cut (:33-57); SAM
// positions are 0-based (seqan3 reference_position) while the simulator's are 1-based -- the tolerance absorbs it as it
// does in the reference; "mapped" / "correct" count only reads the tested SAM gives at most one record for (:306-314).
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <filesystem>
#include <fstream>
#include <iostream>
#include <sstream>
#include <string>
#include <unordered_map>
#include <vector>

namespace {

struct MapPosition {  // sam_file_analyzer.cpp:9-14
    bool reverse_complement = false;
    unsigned sequence_id = 0;
    unsigned offset = 0;
    bool is_random = false;
};

std::string rename_id(std::string id) {  // :33-57, in the order the callers apply them
    std::replace(id.begin(), id.end(), ' ', '_');
    size_t cut = std::min(id.find('/'), id.find(' '));
    if (cut != std::string::npos) id.resize(cut);
    return id;
}

struct Metrics {
    std::string sam;
    size_t total_reads = 0, random_reads = 0, answer_unique = 0;
    size_t both_unique = 0, answer_only_unique = 0, test_only_unique = 0, none_unique = 0;
    size_t mapped_reads = 0, correct_reads = 0, mapped_random = 0, mapped_locations = 0, acceptable = 0;
    float mapped_pct() const { return (float)mapped_reads / (total_reads - random_reads) * 100; }
    float sensitivity_pct() const { return (float)correct_reads / answer_unique * 100; }
    float false_positive_pct() const { return (float)mapped_random / random_reads * 100; }
    float per_read() const { return (float)mapped_locations / mapped_reads; }
    float precision_pct() const { return (float)acceptable / mapped_locations * 100; }
};

class Analyzer {
    std::unordered_map<std::string, unsigned> read_id_to_index_, sequence_id_to_index_;
    std::vector<bool> answer_is_unique_, is_random_;
    std::vector<std::vector<MapPosition>> answer_;
    unsigned tolerance_;

    void refresh_unique() {  // :112-119, :178-185
        for (size_t i = 0; i < answer_is_unique_.size(); ++i) answer_is_unique_[i] = answer_[i].size() == 1;
    }

    // one SAM record -> (read index, flag, reference index, 0-based position); false = skip the line
    struct SamNames {
        std::unordered_map<std::string, unsigned> ref;
    };
    static bool parse_record(const std::string& line, std::string& qname, unsigned& flag, std::string& rname, long& pos) {
        size_t a = line.find('\t');
        if (a == std::string::npos) return false;
        size_t b = line.find('\t', a + 1);
        if (b == std::string::npos) return false;
        size_t c = line.find('\t', b + 1);
        if (c == std::string::npos) return false;
        size_t d = line.find('\t', c + 1);
        qname.assign(line, 0, a);
        flag = (unsigned)std::strtoul(line.c_str() + a + 1, nullptr, 10);
        rname.assign(line, b + 1, c - b - 1);
        pos = std::strtol(line.substr(c + 1, d == std::string::npos ? d : d - c - 1).c_str(), nullptr, 10) - 1;
        return true;
    }
    template <class F>
    void for_each_mapped(const std::string& path, F&& f) {
        std::ifstream is(path);
        if (!is) throw std::runtime_error("cannot open " + path);
        SamNames names;
        std::string line, qname, rname;
        while (std::getline(is, line)) {
            if (!line.empty() && line.back() == '\r') line.pop_back();
            if (line.empty()) continue;
            if (line[0] == '@') {
                if (line.compare(0, 3, "@SQ") == 0) {
                    size_t p = line.find("\tSN:");
                    if (p != std::string::npos) {
                        size_t e = line.find('\t', p + 4);
                        std::string sn = line.substr(p + 4, e == std::string::npos ? e : e - p - 4);
                        names.ref.emplace(sn, (unsigned)names.ref.size());
                    }
                }
                continue;
            }
            unsigned flag;
            long pos;
            if (!parse_record(line, qname, flag, rname, pos)) continue;
            auto it = read_id_to_index_.find(rename_id(qname));
            if (it == read_id_to_index_.end()) continue;  // the reference's catch (out_of_range)
            if (flag & 0x4u) continue;                     // unmapped
            auto rt = names.ref.find(rname);
            if (rt == names.ref.end() || pos < 0) continue;  // seqan3 would throw format_error / bad optional: skipped
            f(it->second, (flag & 0x10u) != 0, rt->second, (unsigned)pos);
        }
    }

public:
    explicit Analyzer(unsigned tolerance) : tolerance_(tolerance) {}

    void read_fasta_names(const std::string& path) {  // :68-77
        std::ifstream is(path);
        if (!is) throw std::runtime_error("cannot open " + path);
        std::string line;
        unsigned index = 0;
        while (std::getline(is, line))
            if (!line.empty() && line[0] == '>') {
                if (line.back() == '\r') line.pop_back();
                std::string id = line.substr(1);
                size_t cut = std::min(id.find('/'), id.find(' '));
                if (cut != std::string::npos) id.resize(cut);
                sequence_id_to_index_[id] = index++;
            }
    }

    void read_sequence_file(const std::string& path, bool is_dwgsim) {  // :197-245
        std::ifstream is(path);
        if (!is) throw std::runtime_error("cannot open " + path);
        std::string id, seq, plus, qual;
        unsigned index = 0;
        while (std::getline(is, id)) {
            if (id.empty()) continue;
            if (!std::getline(is, seq) || !std::getline(is, plus) || !std::getline(is, qual)) break;
            if (id.back() == '\r') id.pop_back();
            std::string renamed = rename_id(id.substr(1));
            read_id_to_index_.emplace(renamed, index);
            if (is_dwgsim) {
                // <contig name with one '_'>_<start>_<..>_<strand>_<..>_<random>...: split at '_' and ':' (:206-224)
                std::vector<std::string> parts;
                size_t pos, prev = 0;
                while ((pos = renamed.find_first_of("_:", prev)) != std::string::npos) {
                    if (pos > prev) parts.push_back(renamed.substr(prev, pos - prev));
                    prev = pos + 1;
                }
                if (prev < renamed.size()) parts.push_back(renamed.substr(prev));
                if (parts.size() < 7) throw std::runtime_error("not a dwgsim read name: " + renamed);
                MapPosition gt;
                gt.reverse_complement = std::stoi(parts[4]) != 0;
                gt.sequence_id = sequence_id_to_index_[parts[0] + "_" + parts[1]];
                gt.offset = (unsigned)std::stoi(parts[2]);
                gt.is_random = std::stoi(parts[6]) != 0;
                is_random_.push_back(gt.is_random);
                answer_.push_back({gt});
            }
            ++index;
        }
        for (unsigned i = 0; i < index; ++i) {
            answer_is_unique_.push_back(true);
            if (!is_dwgsim) {
                answer_.emplace_back();
                is_random_.push_back(false);
            }
        }
    }

    size_t read_best_alignment_file(const std::string& path) {  // :86-123
        size_t recorded = 0;
        for_each_mapped(path, [&](unsigned read, bool rev, unsigned ref, unsigned pos) {
            MapPosition p;
            p.reverse_complement = rev;
            p.sequence_id = ref;
            p.offset = pos;
            answer_[read].push_back(p);
            ++recorded;
        });
        refresh_unique();
        return recorded;
    }

    void read_ground_truth_file(const std::string& path) {  // :130-186
        std::ifstream is(path);
        if (!is) throw std::runtime_error("cannot open " + path);
        std::string ext = std::filesystem::path(path).extension().string();
        if (ext == ".maf") {
            std::string ig, read_name;
            unsigned offset;
            char rev;
            while (is >> ig >> ig >> ig >> offset >> ig >> ig >> ig >> ig >> ig >> read_name >> ig >> ig >> rev >> ig >> ig) {
                auto it = read_id_to_index_.find(read_name);
                if (it == read_id_to_index_.end()) continue;
                MapPosition p;
                p.reverse_complement = rev == '-';
                size_t start = read_name.find('S') + 1, end = read_name.find('_');
                p.sequence_id = (unsigned)std::stoi(read_name.substr(start, end - start)) - 1;
                p.offset = offset;
                answer_[it->second].push_back(p);
            }
        } else {  // the simulator's own format, one line per read in file order
            unsigned origin, position, rev;
            std::string cigar;
            size_t index = 0;
            while (is >> origin >> position >> rev >> cigar) {
                if (index >= answer_.size()) break;
                MapPosition p;
                p.reverse_complement = rev != 0;
                p.sequence_id = origin;
                p.offset = position;
                answer_[index++].push_back(p);
            }
        }
        refresh_unique();
    }

    Metrics benchmark(const std::string& sam_path) {  // :250-358
        Metrics m;
        m.sam = sam_path;
        size_t n = answer_.size();
        std::vector<bool> mapped(n, false), unique_map(n, true), correct(n, false), mapped_random(n, false);
        for_each_mapped(sam_path, [&](unsigned read, bool rev, unsigned ref, unsigned pos) {
            if (mapped[read]) unique_map[read] = false;
            mapped[read] = true;
            ++m.mapped_locations;
            if (is_random_[read]) {
                mapped_random[read] = true;
                return;
            }
            bool acceptable = false;
            for (const MapPosition& a : answer_[read])
                if (rev == a.reverse_complement && ref == a.sequence_id &&
                    (unsigned)std::abs((int)pos - (int)a.offset) <= tolerance_) {
                    correct[read] = true;
                    acceptable = true;
                }
            if (acceptable) ++m.acceptable;
        });
        m.total_reads = read_id_to_index_.size();
        m.random_reads = (size_t)std::count(is_random_.begin(), is_random_.end(), true);
        m.mapped_random = (size_t)std::count(mapped_random.begin(), mapped_random.end(), true);
        m.answer_unique = (size_t)std::count(answer_is_unique_.begin(), answer_is_unique_.end(), true);
        for (size_t i = 0; i < n; ++i) {
            if (unique_map[i]) {
                (answer_[i].size() == 1 ? m.both_unique : m.test_only_unique) += 1;
                if (mapped[i]) ++m.mapped_reads;
                if (correct[i]) ++m.correct_reads;
            } else {
                (answer_[i].size() == 1 ? m.answer_only_unique : m.none_unique) += 1;
            }
        }
        return m;
    }
};

void print_text(const Metrics& m) {  // the reference's lines, :331-355
    std::ostream& o = std::cout;
    o << "[BENCHMARK]\t============ Benchmarking sam file \"" << m.sam << "\" ============\n";
    o << "[BENCHMARK]\tTotal number of reads: " << m.total_reads << ".\n";
    o << "[BENCHMARK]\tTotal number of random reads: " << m.random_reads << ".\n";
    o << "[BENCHMARK]\tNumber of Uniquely mapped reads: " << m.answer_unique << ".\n";
    o << m.both_unique << " " << m.answer_only_unique << " " << m.test_only_unique << " " << m.none_unique << ".\n";
    o << "[BENCHMARK]\tTotal number of mapped reads: " << m.mapped_reads << " (" << m.mapped_pct() << "%).\n";
    o << "[BENCHMARK]\tTotal number of correctly mapped reads (sensitivity): " << m.correct_reads << " ("
      << m.sensitivity_pct() << "%).\n";
    o << "[BENCHMARK]\tTotal number of mapped random reads (false positives): " << m.mapped_random << " ("
      << m.false_positive_pct() << "%).\n";
    o << "[BENCHMARK]\tNumber of mapped locations returned: " << m.mapped_locations << " (" << m.per_read()
      << " per mapped read).\n";
    o << "[BENCHMARK]\tTotal number of acceptable mapped locations (precision): " << m.acceptable
      << " (precision: " << m.precision_pct() << "%).\n";
}

void print_json(const Metrics& m, bool last) {
    std::printf("  {\"sam\": \"%s\", \"reads\": %zu, \"random_reads\": %zu, \"answer_unique\": %zu, "
                "\"unique\": [%zu, %zu, %zu, %zu], \"mapped_reads\": %zu, \"correct_reads\": %zu, "
                "\"mapped_random\": %zu, \"mapped_locations\": %zu, \"acceptable_locations\": %zu}%s\n",
                m.sam.c_str(), m.total_reads, m.random_reads, m.answer_unique, m.both_unique, m.answer_only_unique,
                m.test_only_unique, m.none_unique, m.mapped_reads, m.correct_reads, m.mapped_random,
                m.mapped_locations, m.acceptable, last ? "" : ",");
}

}  // namespace

int main(int argc, char** argv) {
    std::string reads, genome, truth, best;
    std::vector<std::string> sams;
    bool dwgsim = false, json = false;
    unsigned tolerance = 10;  // main() of the reference: sam_analyzer analyzer(10)
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto need = [&](const char* what) -> std::string {
            if (i + 1 >= argc) {
                std::fprintf(stderr, "[ERROR]\t\t%s needs a value\n", what);
                std::exit(2);
            }
            return argv[++i];
        };
        if (a == "--reads") reads = need("--reads");
        else if (a == "--genome") genome = need("--genome");
        else if (a == "--truth") truth = need("--truth");
        else if (a == "--best") best = need("--best");
        else if (a == "--tolerance") tolerance = (unsigned)std::stoul(need("--tolerance"));
        else if (a == "--dwgsim") dwgsim = true;
        else if (a == "--json") json = true;
        else if (a == "-h" || a == "--help") {
            std::puts("bucketmap_analyze --reads R.fastq [--dwgsim --genome G.fa] [--truth T] [--best GOOD.sam] "
                      "[--tolerance 10] [--json] OUT.sam|DIR ...");
            return 0;
        } else sams.push_back(a);
    }
    if (reads.empty() || sams.empty() || (dwgsim && genome.empty())) {
        std::fprintf(stderr, "[ERROR]\t\t--reads and at least one SAM file are required (--dwgsim also needs --genome)\n");
        return 2;
    }
    try {
        Analyzer an(tolerance);
        if (!genome.empty()) an.read_fasta_names(genome);
        an.read_sequence_file(reads, dwgsim);
        if (!best.empty()) {
            size_t n = an.read_best_alignment_file(best);
            if (!json) std::cout << "[INFO]\t\tMapped sequences in the best alignment file: " << n << ".\n";
        }
        if (!truth.empty()) an.read_ground_truth_file(truth);
        std::vector<std::string> files;
        for (const std::string& s : sams) {
            if (std::filesystem::is_directory(s)) {  // benchmark_directory, :360-366
                std::vector<std::string> in_dir;
                for (const auto& e : std::filesystem::directory_iterator(s))
                    if (e.path().extension() == ".sam") in_dir.push_back(e.path().string());
                std::sort(in_dir.begin(), in_dir.end());
                files.insert(files.end(), in_dir.begin(), in_dir.end());
            } else files.push_back(s);
        }
        if (json) std::puts("[");
        for (size_t i = 0; i < files.size(); ++i) {
            Metrics m = an.benchmark(files[i]);
            if (json) print_json(m, i + 1 == files.size());
            else print_text(m);
        }
        if (json) std::puts("]");
    } catch (const std::exception& e) {
        std::fprintf(stderr, "[ERROR]\t\t%s\n", e.what());
        return 1;
    }
    return 0;
}
