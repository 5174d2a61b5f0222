// TEST INFRASTRUCTURE ONLY (never linked into the product).
// Writes the three index files kSpider::pairwise reads, using the REAL parallel-hashmap
// containers vendored in the reference (/root/reference/lib/parallel-hashmap, v1.3.8), the same
// way the reference indexers dump them (src/sourmash_indexing.cpp:283-310, src/index.cpp:336-363).
// Used to (1) produce fixtures whose bytes come from phmap itself, so the product's own
// phmap-format parser/writer can be validated against them, (2) feed oracle/_ref/pairwise_ref.
//
// Input: a "KSBRAW01" little-endian file
//   u64 magic = 0x313057415242534b ("KSBRAW01")
//   u64 n_totals ; n_totals x { u32 sampleID ; u32 total }
//   u64 n_counts ; n_counts x { u64 colorID  ; u64 count }
//   u64 n_colors ; n_colors x { u64 colorID  ; u32 s ; u32 member[s] }
// Usage: index_writer_ref <raw_file> <index_prefix>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>
#include "parallel_hashmap/phmap.h"
#include "parallel_hashmap/phmap_dump.h"

static bool rd(FILE* f, void* p, size_t n) { return fread(p, 1, n, f) == n; }

int main(int argc, char** argv) {
    if (argc < 3) { fprintf(stderr, "usage: %s raw prefix\n", argv[0]); return 2; }
    FILE* f = fopen(argv[1], "rb");
    if (!f) { perror("open raw"); return 1; }
    std::string prefix = argv[2];
    uint64_t magic = 0, n = 0;
    if (!rd(f, &magic, 8) || magic != 0x313057415242534bULL) { fprintf(stderr, "bad magic\n"); return 1; }

    {   // per-sample totals: flat_hash_map<u32,u32>::phmap_dump
        phmap::flat_hash_map<uint32_t, uint32_t> totals;
        if (!rd(f, &n, 8)) return 1;
        for (uint64_t i = 0; i < n; i++) { uint32_t kv[2]; if (!rd(f, kv, 8)) return 1; totals[kv[0]] = kv[1]; }
        phmap::BinaryOutputArchive ar((prefix + "_groupID_to_kmerCount.bin").c_str());
        totals.phmap_dump(ar);
    }
    {   // per-colour counts: flat_hash_map<u64,u64>::phmap_dump
        phmap::flat_hash_map<uint64_t, uint64_t> counts;
        if (!rd(f, &n, 8)) return 1;
        for (uint64_t i = 0; i < n; i++) { uint64_t kv[2]; if (!rd(f, kv, 16)) return 1; counts[kv[0]] = kv[1]; }
        phmap::BinaryOutputArchive ar((prefix + "_color_count.bin").c_str());
        counts.phmap_dump(ar);
    }
    {   // colour -> sources: size_t n, then n x { u64 colour ; flat_hash_set<u32> dump }
        if (!rd(f, &n, 8)) return 1;
        phmap::BinaryOutputArchive ar((prefix + "_color_to_sources.bin").c_str());
        ar.saveBinary((size_t)n);
        std::vector<uint32_t> m;
        for (uint64_t i = 0; i < n; i++) {
            uint64_t color; uint32_t s;
            if (!rd(f, &color, 8) || !rd(f, &s, 4)) return 1;
            m.resize(s);
            if (s && !rd(f, m.data(), 4ull * s)) return 1;
            phmap::flat_hash_set<uint32_t> set(m.begin(), m.end());
            ar.saveBinary(color);
            ar.saveBinary(set);
        }
    }
    fclose(f);
    return 0;
}
