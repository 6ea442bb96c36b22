// Test harness for the host I/O stage (no GPU needed): the multi-threaded FASTA packer and the FASTQ splitter / packer.
//   io_harness fasta <file> <L> <r> <threads>      -> n_bases, NB, records, buckets, FNV-1a of the packed bytes
//   io_harness fastq <file> <batch> <k> <b> <threads> -> per batch: reads, chunks, all_pass, FNV-1a of planes / qk / offsets / ids
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "bm_host.hpp"

static uint64_t fnv(const void* p, size_t n, uint64_t h = 1469598103934665603ull) {
    const uint8_t* b = (const uint8_t*)p;
    for (size_t i = 0; i < n; i++) { h ^= b[i]; h *= 1099511628211ull; }
    return h;
}

int main(int argc, char** argv) {
    if (argc < 3) return 2;
    if (!strcmp(argv[1], "fasta") && argc == 6) {
        bm::set_host_threads((unsigned)atoi(argv[5]));
        bm::genome g;
        if (!g.read_fasta(argv[2])) return 3;
        g.bucketize((uint32_t)atoi(argv[3]), (uint32_t)atoi(argv[4]));
        printf("n_bases %llu NB %u records %zu buckets %zu\n", (unsigned long long)g.n_bases, g.NB, g.ids.size(), g.b_start.size());
        for (size_t i = 0; i < g.ids.size(); i++) printf("rec %llu %s\n", (unsigned long long)g.rec_len[i], g.ids[i].c_str());
        printf("packed %016llx\n", (unsigned long long)fnv(g.packed.data(), (size_t)(g.n_bases + 3) / 4));
        for (size_t i = 0; i < g.b_start.size(); i++) printf("bucket %llu %u %u\n", (unsigned long long)g.b_start[i], g.b_len[i], g.b_rec[i]);
        return 0;
    }
    if (!strcmp(argv[1], "fastq") && argc == 7) {
        bm::set_host_threads((unsigned)atoi(argv[6]));
        bm::fastq_reader fin(argv[2]);
        if (!fin.ok()) return 3;
        bm::read_batch b;
        const uint32_t k = (uint32_t)atoi(argv[4]), bq = (uint32_t)atoi(argv[5]);
        while (fin.next(b, (uint32_t)atoi(argv[3]), (uint64_t)1 << 31, k, bq)) {
            const uint32_t n = b.size();
            const uint32_t* plo = b.pl_off.as<uint32_t>();
            uint64_t hid = 1469598103934665603ull, hq = hid;
            for (uint32_t i = 0; i < n; i++) { hid = fnv(b.id[i], b.id_len[i], hid); hq = fnv(b.qual[i], b.qual_len[i], hq); }
            printf("batch %u chunks %u all_pass %d planes %016llx qk %016llx off %016llx pl_off %016llx ids %016llx quals %016llx\n", n,
                   plo[n], b.all_pass ? 1 : 0, (unsigned long long)fnv(b.planes.data(), (size_t)plo[n] * 8),
                   (unsigned long long)(b.all_pass ? 0ull : fnv(b.qk.data(), (size_t)plo[n] * 4)),   // no quality plane is made for a batch that passes
                   (unsigned long long)fnv(b.read_off.data(), (size_t)(n + 1) * 8), (unsigned long long)fnv(plo, (size_t)(n + 1) * 4),
                   (unsigned long long)hid, (unsigned long long)hq);
        }
        return 0;
    }
    if (!strcmp(argv[1], "fastq_time") && argc == 7) {            // wall time of the splitter + packer alone
        bm::set_host_threads((unsigned)atoi(argv[6]));
        bm::fastq_reader fin(argv[2]);
        if (!fin.ok()) return 3;
        bm::read_batch b;
        uint64_t reads = 0;
        const auto t0 = std::chrono::steady_clock::now();
        while (fin.next(b, (uint32_t)atoi(argv[3]), (uint64_t)1 << 31, (uint32_t)atoi(argv[4]), (uint32_t)atoi(argv[5]))) reads += b.size();
        const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        printf("reads %llu seconds %.4f reads_per_s %.0f\n", (unsigned long long)reads, dt, reads / dt);
        return 0;
    }
    return 2;
}
