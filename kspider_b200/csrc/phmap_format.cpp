#include "phmap_format.hpp"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <functional>

#include "ksb_common.hpp"

namespace ksb {

bool open_dump(const uint8_t* p, uint64_t avail, uint64_t slot_bytes, DumpView* v) {
    if (avail < 16) { set_error("phmap dump: truncated header (%llu bytes left)", (unsigned long long)avail); return false; }
    std::memcpy(&v->size, p, 8);
    std::memcpy(&v->capacity, p + 8, 8);
    v->ctrl = nullptr;
    v->slots = nullptr;
    if (v->size == 0) { v->bytes = 16; return true; }
    if (v->capacity == 0 || ((v->capacity + 1) & v->capacity) != 0 || v->size > v->capacity) {
        set_error("phmap dump: invalid size/capacity %llu/%llu", (unsigned long long)v->size, (unsigned long long)v->capacity);
        return false;
    }
    const uint64_t ctrl_bytes = v->capacity + kGroupWidth + 1;
    if (slot_bytes && v->capacity > (UINT64_MAX - 16 - ctrl_bytes) / slot_bytes) { set_error("phmap dump: capacity overflow"); return false; }
    const uint64_t need = 16 + ctrl_bytes + slot_bytes * v->capacity;
    if (avail < need) {
        set_error("phmap dump: needs %llu bytes, %llu left (file truncated or not an SSE2-build dump)",
                  (unsigned long long)need, (unsigned long long)avail);
        return false;
    }
    v->ctrl = reinterpret_cast<const int8_t*>(p + 16);
    v->slots = p + 16 + ctrl_bytes;
    v->bytes = need;
    return true;
}

bool MappedFile::open(const std::string& path) {
    close();
    fd = ::open(path.c_str(), O_RDONLY);
    if (fd < 0) { set_error("cannot open %s", path.c_str()); return false; }
    struct stat st;
    if (fstat(fd, &st) != 0) { set_error("cannot stat %s", path.c_str()); close(); return false; }
    size = (uint64_t)st.st_size;
    if (size == 0) { data = nullptr; return true; }
    void* m = mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
    if (m == MAP_FAILED) { set_error("cannot mmap %s", path.c_str()); close(); return false; }
    madvise(m, size, MADV_SEQUENTIAL);
    data = static_cast<const uint8_t*>(m);
    return true;
}

void MappedFile::close() {
    if (data) munmap(const_cast<uint8_t*>(data), size);
    if (fd >= 0) ::close(fd);
    data = nullptr; size = 0; fd = -1;
}

static inline uint64_t phmap_mix8(uint64_t a) {
    const unsigned __int128 m = (unsigned __int128)a * 0xde5fb9d2630458e9ULL;
    return (uint64_t)(m >> 64) + (uint64_t)m;
}

static inline uint64_t normalize_capacity(uint64_t n) { return n ? (~0ULL >> __builtin_clzll(n)) : 1; }
static inline uint64_t growth_to_lowerbound_capacity(uint64_t g) { return g + (uint64_t)(((int64_t)g - 1) / 7); }

void append_dump(std::vector<uint8_t>& out, const void* keys, const void* vals, uint64_t n,
                 uint32_t key_bytes, uint32_t value_bytes) {
    const uint64_t slot_bytes = key_bytes + value_bytes;
    uint64_t cap = 0;
    if (n) cap = normalize_capacity(growth_to_lowerbound_capacity(n));
    const size_t base = out.size();
    out.resize(base + 16);
    std::memcpy(&out[base], &n, 8);
    std::memcpy(&out[base + 8], &cap, 8);
    if (n == 0) return;
    const uint64_t ctrl_bytes = cap + kGroupWidth + 1;
    out.resize(base + 16 + ctrl_bytes + slot_bytes * cap, 0);
    int8_t* ctrl = reinterpret_cast<int8_t*>(&out[base + 16]);
    uint8_t* slots = &out[base + 16 + ctrl_bytes];
    std::memset(ctrl, -128, cap + kGroupWidth);  // kEmpty
    ctrl[cap] = -1;                              // kSentinel
    ctrl[cap + kGroupWidth] = 0;                 // phmap leaves this byte uninitialised
    const uint8_t* kp = static_cast<const uint8_t*>(keys);
    const uint8_t* vp = static_cast<const uint8_t*>(vals);
    for (uint64_t e = 0; e < n; e++) {
        uint64_t key = 0;
        std::memcpy(&key, kp + e * key_bytes, key_bytes);
        const uint64_t h = phmap_mix8(key);
        uint64_t offset = (h >> 7) & cap, index = 0, slot;
        for (;;) {
            int found = -1;
            for (uint64_t i = 0; i < kGroupWidth; i++)
                if (ctrl[offset + i] < -1) { found = (int)i; break; }  // empty or deleted
            if (found >= 0) { slot = (offset + (uint64_t)found) & cap; break; }
            index += kGroupWidth;
            offset = (offset + index) & cap;
        }
        const int8_t h2 = (int8_t)(h & 0x7F);
        ctrl[slot] = h2;
        ctrl[((slot - kGroupWidth) & cap) + 1 + ((kGroupWidth - 1) & cap)] = h2;  // cloned byte
        std::memcpy(slots + slot * slot_bytes, kp + e * key_bytes, key_bytes);
        if (value_bytes) std::memcpy(slots + slot * slot_bytes + key_bytes, vp + e * value_bytes, value_bytes);
    }
}

}  // namespace ksb
