// Host side of the FILE-driven consumers of the pairwise TSV -- what `kSpider cluster` and
// `kSpider pairwise --estimate-ani` read (pykSpider/kSpider2/ks_clustering.py:63-117, ks_pairwise.py:64-73):
//   * ksb_table_read_tsv      multithreaded parse of `P_kSpider_pairwise.tsv` (and, for dist type "ani", of the
//                             one-column `P_kSpider_pairwise.ani_col.tsv`) into columns: int(row[0]), int(row[1]),
//                             float(row[dist_col]) as Python parses them (correctly rounded decimal -> double,
//                             "inf" / "nan" / "-nan" accepted);
//   * ksb_write_sidecar / ksb_table_read_sidecar   the binary side-car `P_kSpider_pairwise.ksb`:
//                             the 32-byte records in the row order of the TSV the drop-in wrote beside it, so the
//                             consumers skip the text.  It names the TSV's size and modification time; a TSV
//                             rewritten later (by any tool) invalidates it and the text is parsed instead.
// The arithmetic on these columns runs on the GPU (table.cu: ksb_cluster_table, ksb_ani_table).
#include <cerrno>
#include <charconv>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include "ksb_common.hpp"
#include "table.hpp"

namespace ksb {

namespace {

struct Mapped {
    const char* p = nullptr;
    size_t n = 0;
    int fd = -1;
    ~Mapped() {
        if (p && n) munmap(const_cast<char*>(p), n);
        if (fd >= 0) close(fd);
    }
    int open_ro(const char* path) {
        fd = ::open(path, O_RDONLY);
        if (fd < 0) { set_error("cannot read %s: %s", path, strerror(errno)); return KSB_ERR_IO; }
        struct stat st;
        if (fstat(fd, &st) != 0) { set_error("cannot stat %s: %s", path, strerror(errno)); return KSB_ERR_IO; }
        n = (size_t)st.st_size;
        if (n == 0) return KSB_OK;
        void* m = mmap(nullptr, n, PROT_READ, MAP_PRIVATE, fd, 0);
        if (m == MAP_FAILED) { p = nullptr; set_error("cannot map %s: %s", path, strerror(errno)); return KSB_ERR_IO; }
        p = static_cast<const char*>(m);
        madvise(m, n, MADV_SEQUENTIAL);
        return KSB_OK;
    }
};

// Python's float(text) on one TSV field: optional surrounding blanks, sign, decimal / exponent form,
// inf / infinity / nan in any case.  std::from_chars is correctly rounded like CPython's dtoa; it does not
// take a leading '+', which float() accepts.
bool parse_double(const char* b, const char* e, double* out) {
    while (b < e && (*b == ' ' || *b == '\r')) b++;
    while (e > b && (e[-1] == ' ' || e[-1] == '\r')) e--;
    if (b < e && *b == '+') { b++; if (b < e && (*b == '-' || *b == '+')) return false; }
    if (b == e) return false;
    auto r = std::from_chars(b, e, *out, std::chars_format::general);
    if (r.ec == std::errc::result_out_of_range) {  // float("1e999") = inf, float("1e-999") = 0.0
        char tmp[64];
        const size_t n = std::min<size_t>((size_t)(e - b), sizeof tmp - 1);
        std::memcpy(tmp, b, n);
        tmp[n] = 0;
        *out = strtod(tmp, nullptr);
        return true;
    }
    return r.ec == std::errc() && r.ptr == e;
}

inline bool parse_u32(const char*& p, const char* e, uint32_t* out) {
    uint64_t v = 0;
    const char* s = p;
    while (p < e && *p >= '0' && *p <= '9') { v = v * 10 + (uint64_t)(*p - '0'); if (v > 0xFFFFFFFFull) return false; p++; }
    if (p == s) return false;
    *out = (uint32_t)v;
    return true;
}

// [b, e) of the data region cut into `threads` pieces at line ends
std::vector<size_t> line_cuts(const char* p, size_t b, size_t e, int threads) {
    std::vector<size_t> cut(threads + 1, e);
    cut[0] = b;
    for (int t = 1; t < threads; t++) {
        size_t at = b + (e - b) * (size_t)t / (size_t)threads;
        if (at < cut[t - 1]) at = cut[t - 1];
        const void* nl = at < e ? memchr(p + at, '\n', e - at) : nullptr;
        cut[t] = nl ? (size_t)(static_cast<const char*>(nl) - p) + 1 : e;
    }
    return cut;
}

void* alloc_host(size_t bytes, bool* pinned) {
    void* m = nullptr;
    *pinned = false;
    if (bytes == 0) bytes = 8;
    if (table_try_pinned(&m, bytes)) { *pinned = true; return m; }
    if (posix_memalign(&m, 4096, bytes) != 0) return nullptr;
    return m;
}

}  // namespace

void table_release(ksb_table* t) {
    if (!t) return;
    for (void* p : {(void*)t->s1, (void*)t->s2, (void*)t->dist, (void*)t->dist2, (void*)t->cont, (void*)t->cont2})
        if (p) { if (t->pinned) table_free_pinned(p); else free(p); }
    delete t;
}

static int alloc_columns(ksb_table* t, uint64_t n, bool doubles, bool second) {
    bool pin = true, p2 = false;
    auto get = [&](size_t bytes) -> void* {
        void* m = nullptr;
        if (pin) { m = alloc_host(bytes, &p2); if (!p2) pin = false; return m; }
        if (posix_memalign(&m, 4096, bytes ? bytes : 8) != 0) return nullptr;
        return m;
    };
    // all columns of one table come from the same allocator: the first decides
    t->s1 = static_cast<uint32_t*>(get(4 * n));
    t->pinned = pin;
    t->s2 = static_cast<uint32_t*>(get(4 * n));
    if (doubles) {
        t->dist = static_cast<double*>(get(8 * n));
        if (second) t->dist2 = static_cast<double*>(get(8 * n));
    } else {
        t->cont = static_cast<float*>(get(4 * n));
        if (second) t->cont2 = static_cast<float*>(get(4 * n));
    }
    if (!t->s1 || !t->s2 || (doubles && (!t->dist || (second && !t->dist2))) || (!doubles && (!t->cont || (second && !t->cont2)))) {
        set_error("out of host memory for %llu table rows", (unsigned long long)n);
        return KSB_ERR_NOMEM;
    }
    if (t->pinned != pin) { set_error("mixed host allocators in one table"); return KSB_ERR_NOMEM; }
    return KSB_OK;
}

}  // namespace ksb

using namespace ksb;

extern "C" {

int ksb_table_read_tsv(const char* pairwise_tsv, int dist_col, int dist_col2, const char* ani_col_tsv, int host_threads,
                       ksb_table** out) {
    if (!pairwise_tsv || !out) { set_error("ksb_table_read_tsv: null argument"); return KSB_ERR_ARG; }
    *out = nullptr;
    if (dist_col < 3 || dist_col > 6 || (dist_col == 6 && !ani_col_tsv) || (dist_col2 != 0 && (dist_col2 < 3 || dist_col2 > 5 || dist_col == 6))) {
        set_error("ksb_table_read_tsv: dist_col must be 3 (min), 4 (avg), 5 (max) or 6 (ani, with the ani_col file); dist_col2 0 or 3..5");
        return KSB_ERR_ARG;
    }
    const int T = std::max(1, host_threads);
    Mapped tsv, ani;
    int rc = tsv.open_ro(pairwise_tsv);
    if (rc) return rc;
    if (dist_col == 6 && (rc = ani.open_ro(ani_col_tsv))) return rc;
    // header line is skipped, whatever it says (next(pairwise_tsv), ks_clustering.py:69)
    const void* h = tsv.n ? memchr(tsv.p, '\n', tsv.n) : nullptr;
    const size_t d0 = h ? (size_t)(static_cast<const char*>(h) - tsv.p) + 1 : tsv.n;
    const std::vector<size_t> cut = line_cuts(tsv.p, d0, tsv.n, T);
    std::vector<uint64_t> first(T + 1, 0);
    parallel_chunks((uint64_t)T, T, [&](int, uint64_t tb, uint64_t te) {
        for (uint64_t t = tb; t < te; t++) {
            uint64_t n = 0;
            const char* p = tsv.p + cut[t];
            const char* e = tsv.p + cut[t + 1];
            while (p < e) {
                const void* nl = memchr(p, '\n', (size_t)(e - p));
                n++;
                p = nl ? static_cast<const char*>(nl) + 1 : e;
            }
            first[t + 1] = n;
        }
    });
    for (int t = 0; t < T; t++) first[t + 1] += first[t];
    const uint64_t n_rows = first[T];
    ksb_table* tb = new (std::nothrow) ksb_table();
    if (!tb) { set_error("out of memory"); return KSB_ERR_NOMEM; }
    tb->n = n_rows;
    tb->dist_col = dist_col;
    tb->dist_col2 = dist_col2;
    rc = alloc_columns(tb, n_rows, true, dist_col2 != 0);
    if (rc) { table_release(tb); return rc; }
    std::vector<uint64_t> bad(T, UINT64_MAX);
    parallel_chunks((uint64_t)T, T, [&](int, uint64_t tb0, uint64_t te) {
        for (uint64_t t = tb0; t < te; t++) {
            uint64_t row = first[t];
            const char* p = tsv.p + cut[t];
            const char* e = tsv.p + cut[t + 1];
            while (p < e) {
                const void* nlv = memchr(p, '\n', (size_t)(e - p));
                const char* le = nlv ? static_cast<const char*>(nlv) : e;
                const char* q = p;
                bool ok = parse_u32(q, le, &tb->s1[row]) && q < le && *q == '\t';
                if (ok) { q++; ok = parse_u32(q, le, &tb->s2[row]) && q < le && *q == '\t'; }
                // fields 2..: walk to dist_col (and dist_col2)
                int col = 1;
                while (ok && col < 6) {
                    q++;  // past the tab
                    col++;
                    const void* tv = memchr(q, '\t', (size_t)(le - q));
                    const char* fe = tv ? static_cast<const char*>(tv) : le;
                    if (col == dist_col) ok = parse_double(q, fe, &tb->dist[row]);
                    if (ok && col == dist_col2) ok = parse_double(q, fe, &tb->dist2[row]);
                    if (col >= std::max(dist_col == 6 ? 0 : dist_col, dist_col2)) break;
                    if (!tv) { ok = false; break; }
                    q = fe;
                }
                if (!ok && bad[t] == UINT64_MAX) bad[t] = row;
                row++;
                p = nlv ? le + 1 : e;
            }
        }
    });
    for (int t = 0; t < T; t++)
        if (bad[t] != UINT64_MAX) {
            set_error("%s: data line %llu is not `int<TAB>int<TAB>...` with a number in column %d", pairwise_tsv, (unsigned long long)(bad[t] + 1), dist_col);
            table_release(tb);
            return KSB_ERR_FORMAT;
        }
    if (dist_col == 6) {  // one float per line after a header line; line k belongs to row k (ks_clustering.py:71-79)
        const void* ah = ani.n ? memchr(ani.p, '\n', ani.n) : nullptr;
        const size_t a0 = ah ? (size_t)(static_cast<const char*>(ah) - ani.p) + 1 : ani.n;
        const std::vector<size_t> acut = line_cuts(ani.p, a0, ani.n, T);
        std::vector<uint64_t> afirst(T + 1, 0);
        parallel_chunks((uint64_t)T, T, [&](int, uint64_t tb0, uint64_t te) {
            for (uint64_t t = tb0; t < te; t++) {
                uint64_t n = 0;
                const char* p = ani.p + acut[t];
                const char* e = ani.p + acut[t + 1];
                while (p < e) { const void* nl = memchr(p, '\n', (size_t)(e - p)); n++; p = nl ? static_cast<const char*>(nl) + 1 : e; }
                afirst[t + 1] = n;
            }
        });
        for (int t = 0; t < T; t++) afirst[t + 1] += afirst[t];
        if (afirst[T] < n_rows) {  // the reference's next(ani_col_file) would raise StopIteration
            set_error("%s holds %llu values for %llu pairwise rows", ani_col_tsv, (unsigned long long)afirst[T], (unsigned long long)n_rows);
            table_release(tb);
            return KSB_ERR_FORMAT;
        }
        std::vector<uint64_t> abad(T, UINT64_MAX);
        parallel_chunks((uint64_t)T, T, [&](int, uint64_t tb0, uint64_t te) {
            for (uint64_t t = tb0; t < te; t++) {
                uint64_t row = afirst[t];
                const char* p = ani.p + acut[t];
                const char* e = ani.p + acut[t + 1];
                while (p < e && row < n_rows) {
                    const void* nlv = memchr(p, '\n', (size_t)(e - p));
                    const char* le = nlv ? static_cast<const char*>(nlv) : e;
                    if (!parse_double(p, le, &tb->dist[row]) && abad[t] == UINT64_MAX) abad[t] = row;
                    row++;
                    p = nlv ? le + 1 : e;
                }
            }
        });
        for (int t = 0; t < T; t++)
            if (abad[t] != UINT64_MAX) {
                set_error("%s: line %llu is not a number", ani_col_tsv, (unsigned long long)(abad[t] + 2));
                table_release(tb);
                return KSB_ERR_FORMAT;
            }
    }
    *out = tb;
    return KSB_OK;
}

// ---- binary side-car ---------------------------------------------------------------------------
namespace {
struct SidecarHeader {
    uint64_t magic;      // "KSBSIDE1"
    uint64_t n_rows;
    uint64_t tsv_bytes;  // size of the TSV written beside it
    uint64_t record_bytes;
    uint64_t tsv_mtime_ns;  // modification time of that TSV: a TSV rewritten later (by any tool) invalidates the side-car
    uint64_t reserved[3];
};
constexpr uint64_t kSideMagic = 0x314544495342534BULL;
}  // namespace

int ksb_write_sidecar_records(const ksb_record* records, uint64_t n_rows, const char* path, uint64_t tsv_bytes, uint64_t tsv_mtime_ns, int append) {
    if (!path || (!records && n_rows)) { set_error("ksb_write_sidecar: null argument"); return KSB_ERR_ARG; }
    SidecarHeader h{};
    FILE* f = nullptr;
    if (append) {
        f = fopen(path, "r+b");
        if (!f || fread(&h, sizeof h, 1, f) != 1 || h.magic != kSideMagic) { if (f) fclose(f); set_error("%s: not a side-car to append to", path); return KSB_ERR_IO; }
    } else {
        f = fopen(path, "wb");
        if (!f) { set_error("cannot write %s: %s", path, strerror(errno)); return KSB_ERR_IO; }
        h.magic = kSideMagic;
        h.record_bytes = sizeof(ksb_record);
    }
    h.n_rows = (append ? h.n_rows : 0) + n_rows;
    h.tsv_bytes = tsv_bytes;
    h.tsv_mtime_ns = tsv_mtime_ns;
    int rc = KSB_OK;
    if (fseek(f, 0, SEEK_SET) != 0 || fwrite(&h, sizeof h, 1, f) != 1) rc = KSB_ERR_IO;
    if (rc == KSB_OK && fseek(f, 0, SEEK_END) != 0) rc = KSB_ERR_IO;
    if (rc == KSB_OK && n_rows && fwrite(records, sizeof(ksb_record), n_rows, f) != n_rows) rc = KSB_ERR_IO;
    if (fclose(f) != 0) rc = KSB_ERR_IO;
    if (rc) set_error("write error on %s", path);
    return rc;
}

int ksb_write_sidecar(const ksb_result* r, const char* path, const char* tsv_path, int append) {
    if (!r || !path || !tsv_path) { set_error("ksb_write_sidecar: null argument"); return KSB_ERR_ARG; }
    uint64_t n = 0;
    const ksb_record* rec = ksb_result_records(r, &n);
    ksb_result_info info;
    ksb_result_get_info(r, &info);
    if (info.n_rows && !rec) { set_error("ksb_write_sidecar: the result holds no host records (run with want_records = 1)"); return KSB_ERR_ARG; }
    struct stat st;
    if (stat(tsv_path, &st) != 0) { set_error("ksb_write_sidecar: cannot stat %s: %s", tsv_path, strerror(errno)); return KSB_ERR_IO; }
    return ksb_write_sidecar_records(rec, n, path, (uint64_t)st.st_size, (uint64_t)st.st_mtim.tv_sec * 1000000000ull + (uint64_t)st.st_mtim.tv_nsec, append);
}

// Re-stamps a side-car with the size and modification time of the TSV it belongs to: the drop-in writes both files at the
// same time (two threads) and can only name the finished TSV afterwards.
int ksb_sidecar_stamp(const char* sidecar, const char* pairwise_tsv) {
    if (!sidecar || !pairwise_tsv) { set_error("ksb_sidecar_stamp: null argument"); return KSB_ERR_ARG; }
    struct stat st;
    if (stat(pairwise_tsv, &st) != 0) { set_error("ksb_sidecar_stamp: cannot stat %s: %s", pairwise_tsv, strerror(errno)); return KSB_ERR_IO; }
    FILE* f = fopen(sidecar, "r+b");
    SidecarHeader h{};
    if (!f || fread(&h, sizeof h, 1, f) != 1 || h.magic != kSideMagic) { if (f) fclose(f); set_error("%s: not a side-car", sidecar); return KSB_ERR_IO; }
    h.tsv_bytes = (uint64_t)st.st_size;
    h.tsv_mtime_ns = (uint64_t)st.st_mtim.tv_sec * 1000000000ull + (uint64_t)st.st_mtim.tv_nsec;
    int rc = KSB_OK;
    if (fseek(f, 0, SEEK_SET) != 0 || fwrite(&h, sizeof h, 1, f) != 1) rc = KSB_ERR_IO;
    if (fclose(f) != 0) rc = KSB_ERR_IO;
    if (rc) set_error("write error on %s", sidecar);
    return rc;
}

int ksb_sidecar_matches(const char* sidecar, const char* pairwise_tsv) {
    if (!sidecar || !pairwise_tsv) return 0;
    struct stat st, ss;
    if (stat(pairwise_tsv, &st) != 0 || stat(sidecar, &ss) != 0) return 0;
    FILE* f = fopen(sidecar, "rb");
    if (!f) return 0;
    SidecarHeader h{};
    const bool ok = fread(&h, sizeof h, 1, f) == 1;
    fclose(f);
    return ok && h.magic == kSideMagic && h.record_bytes == sizeof(ksb_record) && h.tsv_bytes == (uint64_t)st.st_size &&
           h.tsv_mtime_ns == (uint64_t)st.st_mtim.tv_sec * 1000000000ull + (uint64_t)st.st_mtim.tv_nsec &&
           (uint64_t)ss.st_size == sizeof h + h.n_rows * sizeof(ksb_record) ? 1 : 0;
}

int ksb_table_read_sidecar(const char* sidecar, int dist_col, int dist_col2, int host_threads, ksb_table** out) {
    if (!sidecar || !out) { set_error("ksb_table_read_sidecar: null argument"); return KSB_ERR_ARG; }
    *out = nullptr;
    if (dist_col < 3 || dist_col > 5 || (dist_col2 != 0 && (dist_col2 < 3 || dist_col2 > 5))) { set_error("ksb_table_read_sidecar: dist_col must be 3, 4 or 5"); return KSB_ERR_ARG; }
    Mapped m;
    int rc = m.open_ro(sidecar);
    if (rc) return rc;
    SidecarHeader h{};
    if (m.n < sizeof h) { set_error("%s: too short for a side-car", sidecar); return KSB_ERR_FORMAT; }
    std::memcpy(&h, m.p, sizeof h);
    if (h.magic != kSideMagic || h.record_bytes != sizeof(ksb_record) || m.n != sizeof h + h.n_rows * sizeof(ksb_record)) {
        set_error("%s: not a side-car of this library (or truncated)", sidecar);
        return KSB_ERR_FORMAT;
    }
    ksb_table* tb = new (std::nothrow) ksb_table();
    if (!tb) { set_error("out of memory"); return KSB_ERR_NOMEM; }
    tb->n = h.n_rows;
    tb->dist_col = dist_col;
    tb->dist_col2 = dist_col2;
    rc = alloc_columns(tb, h.n_rows, false, dist_col2 != 0);
    if (rc) { table_release(tb); return rc; }
    const ksb_record* rec = reinterpret_cast<const ksb_record*>(m.p + sizeof h);
    parallel_chunks(h.n_rows, std::max(1, host_threads), [&](int, uint64_t b, uint64_t e) {
        for (uint64_t i = b; i < e; i++) {
            const ksb_record& r = rec[i];
            tb->s1[i] = r.source_1;
            tb->s2[i] = r.source_2;
            const float* c = &r.min_containment;
            tb->cont[i] = c[dist_col - 3];
            if (dist_col2) tb->cont2[i] = c[dist_col2 - 3];
        }
    });
    *out = tb;
    return KSB_OK;
}

int ksb_table_get(const ksb_table* t, uint64_t* n_rows, const uint32_t** source_1, const uint32_t** source_2,
                  const double** dist, const float** cont) {
    if (!t) { set_error("ksb_table_get: null table"); return KSB_ERR_ARG; }
    if (n_rows) *n_rows = t->n;
    if (source_1) *source_1 = t->s1;
    if (source_2) *source_2 = t->s2;
    if (dist) *dist = t->dist;
    if (cont) *cont = t->cont;
    return KSB_OK;
}

void ksb_table_free(ksb_table* t) { table_release(t); }

}  // extern "C"
