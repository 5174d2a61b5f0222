// TEST INFRASTRUCTURE ONLY -- reader / writer of kProcessor `colored_kDataFrame` file sets built from the
// REFERENCE's own code: /root/reference/lib/kProcessor/src/colorTable.cpp is compiled unmodified where it lies
// (oracle/Makefile; intVectorsTable(folder) :218-242 and ::save :256-282 are the real ones) and the k-mer map is
// kProcessor's own type (include/kProcessor/kDataFrame.hpp:14,23-26: parallel_flat_hash_map<u64,u64,...,6>)
// dumped / loaded by the vendored phmap (kDataFramePHMAP::save / load, src/kDataFrames/kDataFramePHMAP.cpp:208-252).
// kDataFrame.hpp itself is not included (it pulls kmerDecoder / MQF, unbuildable here): its map alias and the four
// `.extra` lines are restated below, and colored_kDataFrame::save's namesMap loop (colored_kDataFrame.cpp:107-113).
//
//   ckf_ref write <in.raw> <prefix>   KSBCKF01 description -> <prefix>.phmap, .extra, colorTableConfig.json,
//                                     colors.intvectors, .namesMap  (what colored_kDataFrame::save(prefix) leaves)
//   ckf_ref dump <prefix> <out.txt>   loads them back with the real loaders and prints what the pairwise pass needs:
//                                     "C <colour> <#k-mers> <size> <members...>" (ascending colour id),
//                                     "T <sample> <total>" (total = sum of the counts of the colours holding the sample)
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <fstream>
#include <iostream>
#include <map>
#include <string>
#include <vector>

#include <memory>
#include <sstream>
#include "sdsl/vectors.hpp"
#include <parallel_hashmap/phmap.h>
#include <parallel_hashmap/phmap_dump.h>
// everything colorTable.hpp includes is in by now (include guards): only the reference's own classes see the macro.
// The dump enumerates intVectorsTable::colors, which has no public iterator; the class is otherwise used as is.
#define private public
#include "colorTable.hpp"
#undef private

using KMER_MAP = phmap::parallel_flat_hash_map<uint64_t, uint64_t, std::hash<uint64_t>, std::equal_to<uint64_t>,
                                               std::allocator<std::pair<const uint64_t, uint64_t>>, 6>;

namespace kProcessor { namespace utils {  // Utils/utils.cpp needs MQF headers; colorTable.cpp uses this one function
bool FileExists(std::string f) { std::ifstream s(f.c_str()); return s.good(); }
} }

template <class T> static T rd(std::ifstream& f) { T v; f.read(reinterpret_cast<char*>(&v), sizeof v); return v; }

static int do_write(const char* raw, const std::string& prefix) {
    std::ifstream f(raw, std::ios::binary);
    if (!f || rd<uint64_t>(f) != 0x3130464B43425343ULL /* "CSBCKF01" */) { std::fprintf(stderr, "bad raw file\n"); return 1; }
    const uint64_t ksize = rd<uint64_t>(f);
    KMER_MAP kmers;
    const uint64_t nk = rd<uint64_t>(f);
    for (uint64_t i = 0; i < nk; i++) { const uint64_t h = rd<uint64_t>(f), c = rd<uint64_t>(f); kmers[h] = c; }
    intVectorsTable table;
    const uint64_t nc = rd<uint64_t>(f);
    uint32_t max_sample = 0;
    for (uint64_t i = 0; i < nc; i++) {
        const uint64_t id = rd<uint64_t>(f);
        const uint32_t s = rd<uint32_t>(f);
        std::vector<uint32_t> m(s);
        for (auto& x : m) { x = rd<uint32_t>(f); max_sample = std::max(max_sample, x); }
        table.setColor(id, m);
    }
    table.numSamples = max_sample;
    table.numColors = (uint32_t)nc;
    const uint64_t nn = rd<uint64_t>(f);
    std::vector<std::pair<uint32_t, std::string>> names;
    for (uint64_t i = 0; i < nn; i++) {
        const uint32_t id = rd<uint32_t>(f);
        const uint32_t len = rd<uint32_t>(f);
        std::string s(len, ' ');
        f.read(&s[0], len);
        names.emplace_back(id, s);
    }
    {   // kDataFramePHMAP::save
        std::ofstream extra(prefix + ".extra");
        extra << ksize << std::endl << 1 << std::endl << 1 << std::endl << "kSize:" << ksize << std::endl;
        phmap::BinaryOutputArchive ar((prefix + ".phmap").c_str());
        kmers.phmap_dump(ar);
    }
    table.save(prefix);
    std::ofstream nm(prefix + ".namesMap");
    nm << names.size() << std::endl;
    for (auto& it : names) nm << it.first << " " << it.second << std::endl;
    return 0;
}

static int do_dump(const std::string& prefix, const char* out) {
    KMER_MAP kmers;
    {
        phmap::BinaryInputArchive ar((prefix + ".phmap").c_str());
        if (!kmers.phmap_load(ar)) return 1;
    }
    colorTable* ct = colorTable::load(prefix);
    intVectorsTable* table = dynamic_cast<intVectorsTable*>(ct);
    if (!table) { std::fprintf(stderr, "not an intVectorsTable\n"); return 1; }
    std::map<uint64_t, uint64_t> count;
    for (auto& it : table->colors) count[it.first] = 0;
    uint64_t orphan = 0;
    for (auto& kv : kmers) {
        auto it = count.find(kv.second);
        if (it == count.end()) orphan++; else it->second++;
    }
    std::map<uint32_t, uint64_t> totals;
    FILE* f = std::fopen(out, "w");
    if (!f) return 1;
    std::fprintf(f, "K %zu %llu\n", (size_t)kmers.size(), (unsigned long long)orphan);
    for (auto& cc : count) {
        std::vector<uint32_t> m;
        table->getSamples(cc.first, m);
        std::fprintf(f, "C %llu %llu %zu", (unsigned long long)cc.first, (unsigned long long)cc.second, m.size());
        for (uint32_t x : m) { std::fprintf(f, " %u", x); totals[x] += cc.second; }
        std::fprintf(f, "\n");
    }
    for (auto& t : totals) std::fprintf(f, "T %u %llu\n", t.first, (unsigned long long)t.second);
    std::fclose(f);
    return 0;
}

int main(int argc, char** argv) {
    if (argc == 4 && std::string(argv[1]) == "write") return do_write(argv[2], argv[3]);
    if (argc == 4 && std::string(argv[1]) == "dump") return do_dump(argv[2], argv[3]);
    std::fprintf(stderr, "usage: ckf_ref write <in.raw> <prefix> | ckf_ref dump <prefix> <out.txt>\n");
    return 2;
}
