// main.cpp -- `bucketmap` / `bucketmap_align` command line of the B200-native build.
//
// Same options, defaults, checks and exit codes as the reference's bucket_map/main.cpp:12-233 (sharg parser rewritten
// by hand).  Three things the reference bakes in at compile time (CMakeLists.txt:7-58) are run-time options here, with
// the compile-time macros honoured as defaults when defined:
//     -g/--genome <fasta>       (BM_GENOME_PATH)      -L/--bucket-len <n>   (BM_BUCKET_LEN, default 65536)
//     the bucket count          (BM_BUCKET_NUM)       computed from the FASTA
// and the FracMinHash hash parameters, which the reference draws from rand() seeded with time(), can be pinned with
// --seed-x / --seed-y so that index builds are reproducible.  --device selects the GPU, --devices several (0-7 or 0,2,5: the
// batches of one FASTQ are spread over them, the index is replicated over NCCL, SAM records leave in read order);
// --batch-reads the reads per device batch; --threads the host threads of the FASTA / FASTQ / SAM stages; --align switches
// the bucketmap_align behaviour on in either binary.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <thread>
#include <vector>

#include "bm_host.hpp"

namespace fs = std::filesystem;

struct cmd_arguments {                     // main.cpp:12-39
    bool only_indexer = false;
    fs::path fastq_path{};
    std::string index_indicator;
    fs::path output_sam_path{};
    uint8_t query_seed_length = 12;
    uint8_t index_seed_length = 9;
    unsigned int max_read_length = 300;
    unsigned int mapper_sample_size = 15;
    float mapper_distinguishability_threshold = 0.5;
    unsigned int average_base_quality = 25;
    float allowed_seed_miss_rate = 0.4;
    float locator_allowed_indel_rate = 0.02;
    float locator_sample_size = 10;
    unsigned int locator_quality_threshold = 40;
    float frac_min_hash = 0.25;
    // run-time replacements of the reference's compile-time configuration
    fs::path genome_path{};
    unsigned int bucket_len = 65536;
    long seed_x = -1, seed_y = -1;
    int device = -1;
    std::vector<int> devices;            // --devices 0-3 / 0,2,5: the batches of the FASTQ are spread over these GPUs
    unsigned int threads = 0;            // host threads of the FASTA / FASTQ / SAM stages (0: min(cores, 16))
    unsigned int batch_reads = 1u << 18;
#ifdef BM_ALIGN
    bool align = true;
#else
    bool align = false;
#endif
};

struct parser_error { std::string what; };

static bool has_ext(fs::path const& p, std::initializer_list<const char*> exts) {
    std::string e = p.extension().string();
    if (!e.empty()) e = e.substr(1);
    for (auto x : exts) if (e == x) return true;
    return false;
}

static void parse(int argc, char** argv, cmd_arguments& a) {
    bool have_indicator = false;
    auto need = [&](int& i, std::string const& opt, std::string const& inline_v, bool has_inline) -> std::string {
        if (has_inline) return inline_v;
        if (i + 1 >= argc) throw parser_error{"Missing value for option " + opt};
        return argv[++i];
    };
    auto to_uint = [](std::string const& s, std::string const& opt) -> unsigned long {
        char* e = nullptr;
        if (s.empty() || s[0] == '-') throw parser_error{"Value parse failed for " + opt + ": '" + s + "'"};
        unsigned long v = strtoul(s.c_str(), &e, 10);
        if (*e) throw parser_error{"Value parse failed for " + opt + ": '" + s + "'"};
        return v;
    };
    auto to_float = [](std::string const& s, std::string const& opt) -> float {
        char* e = nullptr;
        float v = strtof(s.c_str(), &e);
        if (s.empty() || *e) throw parser_error{"Value parse failed for " + opt + ": '" + s + "'"};
        return v;
    };
    for (int i = 1; i < argc; i++) {
        std::string arg = argv[i], name, val;
        bool has_inline = false;
        if (arg.rfind("--", 0) == 0) {
            name = arg.substr(2);
            auto eq = name.find('=');
            if (eq != std::string::npos) { val = name.substr(eq + 1); name = name.substr(0, eq); has_inline = true; }
        } else if (arg.size() >= 2 && arg[0] == '-') {
            name = arg.substr(1, 1);
            if (arg.size() > 2) { val = arg.substr(2); has_inline = true; }
        } else throw parser_error{"Unknown option " + arg + "."};
        auto is = [&](const char* s, const char* l) { return name == s || name == l; };
        if (is("x", "run-index")) a.only_indexer = true;
        else if (is("q", "query-file")) {
            a.fastq_path = need(i, arg, val, has_inline);
            if (!fs::exists(a.fastq_path)) throw parser_error{"Validation failed for option -q/--query-file: The file " + a.fastq_path.string() + " does not exist!"};
            if (!has_ext(a.fastq_path, {"fq", "fastq"})) throw parser_error{"Validation failed for option -q/--query-file: Expected one of the following valid extensions: [fq,fastq]!"};
        } else if (is("i", "index-indicator")) { a.index_indicator = need(i, arg, val, has_inline); have_indicator = true; }
        else if (is("o", "output-file")) {
            a.output_sam_path = need(i, arg, val, has_inline);
            if (fs::exists(a.output_sam_path)) throw parser_error{"Validation failed for option -o/--output-file: The file " + a.output_sam_path.string() + " already exists!"};
            if (!has_ext(a.output_sam_path, {"sam"})) throw parser_error{"Validation failed for option -o/--output-file: Expected one of the following valid extensions: [sam]!"};
        } else if (is("k", "index-seed")) a.index_seed_length = (uint8_t)to_uint(need(i, arg, val, has_inline), arg);
        else if (is("b", "average-base-quality")) a.average_base_quality = (unsigned)to_uint(need(i, arg, val, has_inline), arg);
        else if (is("l", "query-seed")) a.query_seed_length = (uint8_t)to_uint(need(i, arg, val, has_inline), arg);
        else if (is("r", "read-len")) a.max_read_length = (unsigned)to_uint(need(i, arg, val, has_inline), arg);
        else if (is("s", "mapper-samples")) a.mapper_sample_size = (unsigned)to_uint(need(i, arg, val, has_inline), arg);
        else if (is("d", "distinguishability")) a.mapper_distinguishability_threshold = to_float(need(i, arg, val, has_inline), arg);
        else if (is("e", "max-error-rate")) a.allowed_seed_miss_rate = to_float(need(i, arg, val, has_inline), arg);
        else if (is("n", "max-indel-rate")) a.locator_allowed_indel_rate = to_float(need(i, arg, val, has_inline), arg);
        else if (is("p", "locator-samples")) a.locator_sample_size = to_float(need(i, arg, val, has_inline), arg);
        else if (is("u", "quality")) a.locator_quality_threshold = (unsigned)to_uint(need(i, arg, val, has_inline), arg);
        else if (is("f", "kmer-frac")) a.frac_min_hash = to_float(need(i, arg, val, has_inline), arg);
        else if (is("g", "genome")) a.genome_path = need(i, arg, val, has_inline);
        else if (is("L", "bucket-len")) a.bucket_len = (unsigned)to_uint(need(i, arg, val, has_inline), arg);
        else if (name == "seed-x") a.seed_x = (long)to_uint(need(i, arg, val, has_inline), arg);
        else if (name == "seed-y") a.seed_y = (long)to_uint(need(i, arg, val, has_inline), arg);
        else if (name == "device") a.device = (int)to_uint(need(i, arg, val, has_inline), arg);
        else if (name == "devices") {
            std::string v = need(i, arg, val, has_inline);
            a.devices.clear();
            size_t pos = 0;
            while (pos <= v.size()) {
                size_t e = v.find(',', pos);
                if (e == std::string::npos) e = v.size();
                std::string tok = v.substr(pos, e - pos);
                size_t dash = tok.find('-');
                if (dash != std::string::npos) {
                    unsigned long lo = to_uint(tok.substr(0, dash), arg), hi = to_uint(tok.substr(dash + 1), arg);
                    if (hi < lo || hi - lo > 63) throw parser_error{"Value parse failed for " + arg + ": '" + v + "'"};
                    for (unsigned long d = lo; d <= hi; d++) a.devices.push_back((int)d);
                } else a.devices.push_back((int)to_uint(tok, arg));
                pos = e + 1;
            }
            if (a.devices.empty()) throw parser_error{"Value parse failed for " + arg + ": '" + v + "'"};
        }
        else if (name == "threads") a.threads = (unsigned)to_uint(need(i, arg, val, has_inline), arg);
        else if (name == "batch-reads") a.batch_reads = (unsigned)to_uint(need(i, arg, val, has_inline), arg);
        else if (name == "align") a.align = true;
        else if (name == "version-check") (void)need(i, arg, val, has_inline);     // accepted and ignored (benchmark_map.sh:31)
        else throw parser_error{"Unknown option " + arg + ". In case this is meant to be a non-option/argument/parameter, "
                                "please specify the start of non-options with '--'."};
    }
    if (!have_indicator) throw parser_error{"Option -i/--index-indicator is required but not set."};
}

static double wall_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

int main(int argc, char** argv) {
    const double t_start = wall_s();
    cmd_arguments args{};
#ifdef BM_GENOME_PATH
    args.genome_path = BM_GENOME_PATH;
#endif
#ifdef BM_BUCKET_LEN
    args.bucket_len = BM_BUCKET_LEN;
#endif
    if (const char* e = std::getenv("BM_GENOME")) if (args.genome_path.empty()) args.genome_path = e;
    try {
        parse(argc, argv, args);
    } catch (parser_error const& ext) {
        std::cerr << "[ERROR]\t\t" << ext.what << "\n";
        return -1;
    }
    if (args.align) std::cerr << "[INFO]\t\tAllowing Smith-Waterman for alignment verifications.\n";
    else std::cerr << "[INFO]\t\tNot using Smith-Waterman for alignment verifications.\n";
    if (args.genome_path.empty()) {
        std::cerr << "[ERROR]\t\tThe reference genome is not set. Use '-g <fasta>' (or the BM_GENOME environment variable).\n";
        return -1;
    }
    if (!args.devices.empty()) args.device = args.devices[0];
    if (args.device >= 0) setenv("BM_DEVICE", std::to_string(args.device).c_str(), 1);
    if (args.threads) bm::set_host_threads(args.threads);

    // the CUDA context comes up on a side thread while the host threads parse the FASTA
    std::thread ctx_up([] { bm::context::default_context(); });
    bm::genome const* g = bm::load_genome_cached(args.genome_path, args.bucket_len, args.max_read_length);
    const double t_fasta = wall_s();
    ctx_up.join();
    const double t_ctx = wall_s();
    if (!g) return -1;
    std::cerr << "[INFO]\t\tInitializing indexer and mapper with bucket length: " << args.bucket_len
              << ", and number of buckets: " << g->NB << ".\n";

    std::size_t HASH_TABLE_SIZE = 10000;                                                 // main.cpp:176-185
    auto min_hash_function = bm::generate_hash_function(HASH_TABLE_SIZE, args.seed_x, args.seed_y);
    bm::bucket_hash_indexer ind(args.bucket_len, args.max_read_length, args.index_seed_length, min_hash_function,
                                (unsigned int)(HASH_TABLE_SIZE * args.frac_min_hash));
    if (args.only_indexer) {
        ind.index(args.genome_path, fs::current_path(), args.index_indicator);
        return 0;
    }
    if (args.output_sam_path.empty()) {
        std::cerr << "[ERROR]\t\tThe output sam file is not set. Please set the output path using '-o' option.\n";
        return 1;
    }
    if (args.query_seed_length < args.index_seed_length) {
        std::cerr << "[ERROR]\t\tThe query seed length (currently set to " << (int)args.query_seed_length
                  << ") should be larger thanthe index seed length (currently set to " << (int)args.index_seed_length << ").\n";
        return 1;
    }
    bm::q_gram_mapper map(args.bucket_len, args.max_read_length, args.query_seed_length, args.index_seed_length,
                          args.mapper_sample_size,
                          (unsigned int)std::ceil(args.mapper_sample_size * args.allowed_seed_miss_rate),   // main.cpp:207
                          args.mapper_distinguishability_threshold, args.average_base_quality);
    map.set_num_buckets(g->NB);
    bm::bucket_locator loc(&ind, &map, args.bucket_len, args.max_read_length, args.query_seed_length,
                           args.allowed_seed_miss_rate, args.locator_allowed_indel_rate,
                           (unsigned int)args.locator_sample_size, args.average_base_quality, 5, args.align);
    loc.set_batch_reads(args.batch_reads);
    if (args.devices.size() > 1) loc.set_devices(args.devices);
    const double t_init0 = wall_s();
    loc.initialize(args.genome_path, fs::current_path(), args.index_indicator);
    const double t_init1 = wall_s();
    loc.locate(args.fastq_path, fs::current_path() / (args.index_indicator + ".bucket_id"), args.output_sam_path,
               args.locator_quality_threshold);
    const double t_end = wall_s();
    std::cerr << "[BENCHMARK]\tWall clock of this run: FASTA " << (float)(t_fasta - t_start) << " s (CUDA context up "
              << (float)(t_ctx - t_start) << " s after start, on a side thread), index files + upload " << (float)(t_init1 - t_init0)
              << " s, reads -> SAM " << (float)(t_end - t_init1) << " s, total " << (float)(t_end - t_start) << " s.\n";
    // everything is written and closed: leave without tearing the device allocations down one by one (the driver frees
    // them with the process)
    std::cerr.flush();
    fflush(nullptr);
    std::_Exit(0);
}
