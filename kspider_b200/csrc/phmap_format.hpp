// Byte-level reader/writer of parallel-hashmap v1.3.8 raw_hash_set dumps — the on-disk format of
// the three files kSpider::pairwise reads (src/pairwise.cpp:95-121,166-170).  Written from the
// format description in SURVEY.md Appendix A (phmap_dump.h:52-83; ctrl bytes phmap.h:284-296;
// layout phmap.h:878-881); phmap itself is NOT linked into the product.
#pragma once
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

namespace ksb {

constexpr uint64_t kGroupWidth = 16;  // SSE2 build of phmap (phmap.h:395,522-526)

struct DumpView {
    uint64_t size = 0, capacity = 0;
    const int8_t* ctrl = nullptr;
    const uint8_t* slots = nullptr;
    uint64_t bytes = 0;  // total bytes of the dump in the file
};

// Opens one dump at p with `avail` bytes left.  Returns false (with set_error) if malformed.
bool open_dump(const uint8_t* p, uint64_t avail, uint64_t slot_bytes, DumpView* v);

// Read-only mapping of a whole file.
struct MappedFile {
    const uint8_t* data = nullptr;
    uint64_t size = 0;
    int fd = -1;
    bool open(const std::string& path);
    void close();
    ~MappedFile() { close(); }
};

// Writer: a flat_hash_{set,map} dump of `n` unique keys with the capacity a range-constructed
// phmap container would have, keys placed by phmap's own hash (phmap_mix<8> of the integer,
// phmap_utils.h:66-79), H1/H2 split and quadratic group probing (phmap.h:144-167,339-354,2143-2157),
// so that lookups by the reference (operator[] on groupID_to_kmerCount, src/pairwise.cpp:257-258)
// find them.  key_bytes is 4 or 8; slot = key followed by value_bytes of value.
void append_dump(std::vector<uint8_t>& out, const void* keys, const void* vals, uint64_t n,
                 uint32_t key_bytes, uint32_t value_bytes);

}  // namespace ksb
