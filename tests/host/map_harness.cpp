// Test harness for the C++ boundary: bm::q_gram_mapper::load + map() (mapper/mapper.h:12-33), printed as text.
// usage: map_harness <fasta> <L> <r> <k> <q> <samples> <fault> <d> <b> <index_dir> <indicator> <fastq>
#include <cstdio>
#include <cstdlib>

#include "bm_host.hpp"

int main(int argc, char** argv) {
    if (argc != 13) return 2;
    const unsigned L = atoi(argv[2]), r = atoi(argv[3]), k = atoi(argv[4]), q = atoi(argv[5]), s = atoi(argv[6]), f = atoi(argv[7]);
    const float d = (float)atof(argv[8]);
    const unsigned b = atoi(argv[9]);
    bm::genome const* g = bm::load_genome_cached(argv[1], L, r);
    if (!g) return 3;
    bm::q_gram_mapper m(L, r, (uint8_t)k, (uint8_t)q, s, f, d, b);
    m.set_num_buckets(g->NB);
    m.load(argv[10], argv[11]);
    auto res = m.map(argv[12]);
    printf("num_records %u\n", m.num_records);
    printf("buckets %zu %zu\n", res.first.size(), res.second.size());
    for (int strand = 0; strand < 2; strand++) {
        auto const& lists = strand ? res.second : res.first;
        for (size_t bk = 0; bk < lists.size(); bk++) {
            if (lists[bk].empty()) continue;
            printf("%c %zu", strand ? 'R' : 'O', bk);
            for (auto const& sg : lists[bk]) printf(" %u:%d", sg.first, sg.second);
            printf("\n");
        }
    }
    // mapper::reset (q_gram_mapper.h:374-378), then a second map() must refuse: the index is gone
    m.reset();
    auto again = m.map(argv[12]);
    size_t left = 0;
    for (auto const& v : again.first) left += v.size();
    printf("after_reset %zu\n", left);
    return 0;
}
