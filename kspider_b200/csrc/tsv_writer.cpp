// Host text side of the path: the pairwise TSV (src/pairwise.cpp:242-275), "%g" float text
// (ostream<<float with default flags), and the synthetic index-file writer used by tests / bench.
#include <cerrno>
#include <charconv>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include <fcntl.h>
#include <sys/types.h>
#include <unistd.h>

#include "ksb_common.hpp"
#include "phmap_format.hpp"

namespace ksb {

// "00".."99"
static const char kPairs[201] =
    "0001020304050607080910111213141516171819202122232425262728293031323334353637383940414243444546474849"
    "5051525354555657585960616263646566676869707172737475767778798081828384858687888990919293949596979899";

// decimal text of a u32 (two digits per division); returns length
static inline int u32_to_dec(uint32_t v, char* out) {
    const int n = v < 100u ? (v < 10u ? 1 : 2)
                : v < 10000u ? (v < 1000u ? 3 : 4)
                : v < 1000000u ? (v < 100000u ? 5 : 6)
                : v < 100000000u ? (v < 10000000u ? 7 : 8)
                : (v < 1000000000u ? 9 : 10);
    char* p = out + n;
    while (v >= 100u) {
        const uint32_t q = v / 100u;
        std::memcpy(p -= 2, kPairs + 2 * (v - q * 100u), 2);
        v = q;
    }
    if (v >= 10u) std::memcpy(p - 2, kPairs + 2 * v, 2);
    else p[-1] = (char)('0' + v);
    return n;
}

// decimal text of a u64; returns length
static inline int u64_to_dec(uint64_t v, char* out) {
    if (v <= 0xFFFFFFFFull) return u32_to_dec((uint32_t)v, out);
    // at most 20 digits: the top part (< 1.9e11) again by halves, then 9 zero-padded digits
    const uint64_t hi = v / 1000000000ull;
    uint32_t lo = (uint32_t)(v - hi * 1000000000ull);
    const int n = u64_to_dec(hi, out);
    char* p = out + n + 9;
    for (int i = 0; i < 4; i++) {
        const uint32_t q = lo / 100u;
        std::memcpy(p -= 2, kPairs + 2 * (lo - q * 100u), 2);
        lo = q;
    }
    p[-1] = (char)('0' + lo);
    return n + 9;
}

// "%g" of a float (see ksb_format_float below for the why); in this unit so that the row loop inlines it
static inline int format_float(float v, char* buf) {
    uint32_t bits;
    std::memcpy(&bits, &v, 4);
    if (bits == 0) { buf[0] = '0'; return 1; }
    const uint32_t ex = (bits >> 23) & 0xFFu;
    if ((bits >> 31) == 0 && ex >= 113 && ex <= 146) {  // 2^-14 <= v < 2^20, positive, normal
        const uint64_t M = (bits & 0x7FFFFFu) | 0x800000u;
        const int sh = 150 - (int)ex;                    // v = M >> sh, 4 <= sh <= 37
        static const uint64_t P10[12] = {1ull, 10ull, 100ull, 1000ull, 10000ull, 100000ull, 1000000ull, 10000000ull,
                                         100000000ull, 1000000000ull, 10000000000ull, 100000000000ull};
        // decimal exponent k: 10^k <= v < 10^(k+1); start from the binary exponent, fix up exactly
        int k = (int)(((int)ex - 127) * 1233) >> 12;     // floor(log10(2^e)) or one less
        if (k < -5) k = -5;
        if (k > 5) k = 5;
        uint64_t q = 0, rem = 0;
        for (int it = 0; it < 4; it++) {
            const int p = 5 - k;                         // 0 <= p <= 10
            if (p < 0 || p > 10) { k = 99; break; }
            const uint64_t prod = M * P10[p];            // < 2^24 * 10^10 < 2^58
            q = prod >> sh;
            rem = prod & ((1ull << sh) - 1ull);
            if (q < 100000ull) { k--; continue; }
            if (q >= 1000000ull) { k++; continue; }
            break;
        }
        if (k >= -4 && k <= 5 && q >= 100000ull && q < 1000000ull) {
            const uint64_t half = 1ull << (sh - 1);
            if (rem > half || (rem == half && (q & 1ull))) q++;
            if (q == 1000000ull) { q = 100000ull; k++; }
            if (k <= 5) {
                char d[12] = {};
                const uint32_t q32 = (uint32_t)q, hi2 = q32 / 10000u, lo4 = q32 - hi2 * 10000u, mid2 = lo4 / 100u;
                std::memcpy(d, kPairs + 2 * hi2, 2);
                std::memcpy(d + 2, kPairs + 2 * mid2, 2);
                std::memcpy(d + 4, kPairs + 2 * (lo4 - mid2 * 100u), 2);
                int last = 5;                              // last significant digit kept ("%g" strips trailing zeros)
                while (last > 0 && d[last] == '0') last--;
                // (whole digit groups are copied and the length cut afterwards: at most 11 bytes are touched)
                if (k < 0) {
                    std::memcpy(buf, "0.000", 5);
                    const int n = 1 - k;                   // "0." and -k - 1 zeros
                    std::memcpy(buf + n, d, 6);
                    return n + last + 1;
                }
                std::memcpy(buf, d, 6);
                if (last <= k) return k + 1;
                buf[k + 1] = '.';
                std::memcpy(buf + k + 2, d + k + 1, 5);    // k < last <= 5: at most 5 - k of them count
                return last + 2;
            }
        }
    }
    char tmp[32];
    const int n = snprintf(tmp, sizeof tmp, "%g", (double)v);
    std::memcpy(buf, tmp, (size_t)n);
    return n;
}

static const char kTsvHeader[] =
    "source_1\tsource_2\tshared_kmers\tmin_containment\tavg_containment\tmax_containment\n";  // src/pairwise.cpp:244-251

// Rows -> text -> file.  A batch is formatted into per-thread buffers by all threads; a writer thread then puts the buffers
// into the file, in order, while the NEXT batch is being formatted into a second set of buffers: buffered writes to one file
// serialise on its inode lock whoever issues them, so one writer at full speed beside the formatters is what overlaps
// (the copy into the page cache is the other half of the time: 1.64 GB on cfg4).  `base` = file offset of the first row;
// returns the offset behind the last.
struct TextBuf {  // per-thread text of one batch; never value-initialised (std::string::resize would clear 3x what is written)
    char* p = nullptr;
    size_t cap = 0, len = 0;
    TextBuf() = default;
    TextBuf(const TextBuf&) = delete;
    TextBuf& operator=(const TextBuf&) = delete;
    ~TextBuf() { std::free(p); }
    bool reserve(size_t n) {
        if (n <= cap) return true;
        std::free(p);
        p = (char*)std::malloc(n);
        cap = p ? n : 0;
        return p != nullptr;
    }
};

// fmt(first, last, out) writes the text of rows [first, last) at `out` (at most row_max bytes each) and returns its end
template <class Fmt>
static int write_text_batches(uint64_t n, int fd, uint64_t* base, int threads, size_t row_max, const Fmt& fmt) {
    const uint64_t kBatch = 1ull << 22;
    if (threads < 1) threads = 1;
    std::vector<TextBuf> bufs[2];
    bufs[0] = std::vector<TextBuf>(threads);
    bufs[1] = std::vector<TextBuf>(threads);
    std::vector<uint8_t> failed(threads, 0);
    std::thread writer[2];
    int wrc[2] = {KSB_OK, KSB_OK};
    uint64_t at = *base;
    int rc = KSB_OK;
    uint64_t k = 0;
    for (uint64_t b0 = 0; b0 < n && rc == KSB_OK; b0 += kBatch, k++) {
        const int set = (int)(k & 1);
        if (writer[set].joinable()) { writer[set].join(); if (wrc[set] != KSB_OK) rc = wrc[set]; }  // the set's previous batch is in the file
        if (rc != KSB_OK) break;
        const uint64_t bn = std::min(kBatch, n - b0);
        std::vector<TextBuf>& bs = bufs[set];
        for (auto& s : bs) s.len = 0;
        parallel_chunks(bn, threads, [&](int t, uint64_t b, uint64_t e) {
            TextBuf& s = bs[t];
            if (!s.reserve((size_t)(e - b) * row_max + 16)) { failed[t] = 1; return; }
            s.len = (size_t)(fmt(b0 + b, b0 + e, s.p) - s.p);
        });
        for (int t = 0; t < threads; t++)
            if (failed[t]) rc = KSB_ERR_NOMEM;
        if (rc != KSB_OK) break;
        uint64_t bytes = 0;
        for (const auto& s : bs) bytes += s.len;
        const uint64_t off = at;
        at += bytes;
        if (k > 0 && writer[set ^ 1].joinable()) { writer[set ^ 1].join(); if (wrc[set ^ 1] != KSB_OK) rc = wrc[set ^ 1]; }  // file order: batch k - 1 first
        if (rc != KSB_OK) break;
        wrc[set] = KSB_OK;
        writer[set] = std::thread([&bs, &wrc, set, fd, off]() {
            uint64_t pos = off;
            for (const auto& s : bs) {
                const char* p = s.p;
                uint64_t left = s.len;
                while (left) {
                    const ssize_t w = pwrite(fd, p, (size_t)left, (off_t)pos);
                    if (w < 0) { if (errno == EINTR) continue; wrc[set] = KSB_ERR_IO; return; }
                    p += w; pos += (uint64_t)w; left -= (uint64_t)w;
                }
            }
        });
    }
    for (int q = 0; q < 2; q++)
        if (writer[q].joinable()) { writer[q].join(); if (wrc[q] != KSB_OK) rc = wrc[q]; }
    if (rc == KSB_OK) *base = at;
    return rc;
}

static int write_records(const ksb_record* rec, uint64_t n, int fd, uint64_t* base, int threads) {
    const size_t kRowMax = 96;  // 10 + 10 + 20 digits, 3 x 16 for the floats, 6 separators
    return write_text_batches(n, fd, base, threads, kRowMax, [rec](uint64_t first, uint64_t last, char* line) {
        for (uint64_t i = first; i < last; i++) {
            const ksb_record& r = rec[i];
            line += u32_to_dec(r.source_1, line); *line++ = '\t';
            line += u32_to_dec(r.source_2, line); *line++ = '\t';
            line += u64_to_dec(r.shared_kmers, line); *line++ = '\t';
            line += format_float(r.min_containment, line); *line++ = '\t';
            line += format_float(r.avg_containment, line); *line++ = '\t';
            line += format_float(r.max_containment, line); *line++ = '\n';
        }
        return line;
    });
}

static int write_tsv_rows(const ksb_record* rec, uint64_t n, const char* path, int threads, bool append) {
    const int fd = open(path, append ? (O_WRONLY | O_CREAT) : (O_WRONLY | O_CREAT | O_TRUNC), 0644);
    if (fd < 0) { set_error("cannot write %s: %s", path, strerror(errno)); return KSB_ERR_IO; }
    int rc = KSB_OK;
    uint64_t base = 0;
    if (append) {
        const off_t end = lseek(fd, 0, SEEK_END);
        if (end < 0) rc = KSB_ERR_IO; else base = (uint64_t)end;
    } else {
        const size_t hl = sizeof(kTsvHeader) - 1;
        if (pwrite(fd, kTsvHeader, hl, 0) != (ssize_t)hl) rc = KSB_ERR_IO;
        base = hl;
    }
    if (rc == KSB_OK) rc = write_records(rec, n, fd, &base, threads);
    if (close(fd) != 0) rc = KSB_ERR_IO;
    if (rc == KSB_ERR_IO) set_error("write error on %s", path);
    if (rc == KSB_ERR_NOMEM) set_error("out of host memory writing %s", path);
    return rc;
}

static int write_tsv_impl(const ksb_result* r, const char* path, int threads, bool append) {
    if (!r || !path) { set_error("ksb_write_tsv: null argument"); return KSB_ERR_ARG; }
    uint64_t n = 0;
    const ksb_record* rec = ksb_result_records(r, &n);
    ksb_result_info info;
    ksb_result_get_info(r, &info);
    if (info.n_rows && !rec) { set_error("ksb_write_tsv: the result holds no host records (run with want_records = 1)"); return KSB_ERR_ARG; }
    return write_tsv_rows(rec, n, path, threads, append);
}

}  // namespace ksb

extern "C" {

// ostream<<float with default flags == printf("%g") of the promoted value (libstdc++ num_put uses
// __convert_from_v with "%.*g", precision 6).  nan/inf come out as "nan"/"-nan"/"inf" either way.
//
// Fast exact path for 1e-4 <= v < 1e6 (every containment but the tiniest): v = M * 2^E with a 24-bit
// M and -38 <= E <= -4, so v * 10^p (p <= 10) is the exact integer M * 10^p shifted right by -E and
// fits 64 bits; the six significant digits are that quotient rounded half-to-even on the exact
// remainder -- what glibc prints.  Everything else (0 handled inline) goes through snprintf.
int ksb_format_float(float v, char* buf) { return ksb::format_float(v, buf); }

// Python's repr(float) (what ks_pairwise.py:79 writes with f"{avg_ani}"): the shortest digit string that
// round-trips (std::to_chars), laid out by CPython's rule for format code 'r' -- fixed notation when
// 1e-4 <= |v| < 1e16 (with ".0" appended to integers), otherwise d[.ddd]e+XX with at least two exponent
// digits.  buf needs 32 bytes; returns the length.
int ksb_format_double_repr(double v, char* buf) {
    if (std::isnan(v)) { std::memcpy(buf, "nan", 3); return 3; }
    int n = 0;
    if (std::signbit(v)) { buf[n++] = '-'; v = -v; }
    if (std::isinf(v)) { std::memcpy(buf + n, "inf", 3); return n + 3; }
    if (v == 0.0) { std::memcpy(buf + n, "0.0", 3); return n + 3; }
    if (v >= 1e-4 && v < 1e16) {  // repr's fixed range: the shortest round-trip digits in fixed notation are what to_chars(fixed) writes
        char* const b = buf + n;
        char* e = std::to_chars(b, b + 30, v, std::chars_format::fixed).ptr;
        if (!std::memchr(b, '.', (size_t)(e - b))) { *e++ = '.'; *e++ = '0'; }
        return (int)(e - buf);
    }
    char sci[40];
    const auto res = std::to_chars(sci, sci + sizeof sci, v, std::chars_format::scientific);  // d[.ddd]e[+-]XX, shortest
    char digits[24];
    int nd = 0, e10 = 0;
    const char* q = sci;
    for (; q < res.ptr && *q != 'e'; q++)
        if (*q != '.') digits[nd++] = *q;
    if (q < res.ptr) {
        q++;
        const bool neg = *q == '-';
        if (*q == '+' || *q == '-') q++;
        for (; q < res.ptr; q++) e10 = e10 * 10 + (*q - '0');
        if (neg) e10 = -e10;
    }
    if (e10 < -4 || e10 >= 16) {  // CPython: decpt <= -4 || decpt > 16, decpt = e10 + 1
        buf[n++] = digits[0];
        if (nd > 1) { buf[n++] = '.'; std::memcpy(buf + n, digits + 1, (size_t)nd - 1); n += nd - 1; }
        buf[n++] = 'e';
        buf[n++] = e10 < 0 ? '-' : '+';
        const int ae = e10 < 0 ? -e10 : e10;
        if (ae >= 100) buf[n++] = (char)('0' + ae / 100);
        buf[n++] = (char)('0' + (ae / 10) % 10);
        buf[n++] = (char)('0' + ae % 10);
    } else if (e10 < 0) {
        buf[n++] = '0'; buf[n++] = '.';
        for (int i = 0; i < -e10 - 1; i++) buf[n++] = '0';
        std::memcpy(buf + n, digits, (size_t)nd); n += nd;
    } else {
        for (int i = 0; i <= e10; i++) buf[n++] = i < nd ? digits[i] : '0';
        buf[n++] = '.';
        if (nd > e10 + 1) { std::memcpy(buf + n, digits + e10 + 1, (size_t)(nd - e10 - 1)); n += nd - e10 - 1; }
        else buf[n++] = '0';
    }
    return n;
}

// `P_kSpider_pairwise.ani_col.tsv` (ks_pairwise.py:33,64-80): header "avg_ani", one repr(float) per row.
int ksb_write_ani_column(const double* ani, uint64_t n, const char* path, int host_threads) {
    if ((!ani && n) || !path) { ksb::set_error("ksb_write_ani_column: null argument"); return KSB_ERR_ARG; }
    const int fd = open(path, O_WRONLY | O_CREAT | O_TRUNC, 0644);
    if (fd < 0) { ksb::set_error("cannot write %s: %s", path, strerror(errno)); return KSB_ERR_IO; }
    static const char kHeader[] = "avg_ani\n";
    const size_t hl = sizeof(kHeader) - 1;
    int rc = pwrite(fd, kHeader, hl, 0) == (ssize_t)hl ? KSB_OK : KSB_ERR_IO;
    uint64_t base = hl;
    // the pairwise TSV's writer: all threads format a batch while a writer thread puts the previous one into the file
    if (rc == KSB_OK)
        rc = ksb::write_text_batches(n, fd, &base, host_threads, 32, [ani](uint64_t first, uint64_t last, char* line) {
            for (uint64_t i = first; i < last; i++) { line += ksb_format_double_repr(ani[i], line); *line++ = '\n'; }
            return line;
        });
    if (close(fd) != 0) rc = KSB_ERR_IO;
    if (rc == KSB_ERR_IO) ksb::set_error("write error on %s", path);
    if (rc == KSB_ERR_NOMEM) ksb::set_error("out of host memory writing %s", path);
    return rc;
}

int ksb_write_tsv(const ksb_result* r, const char* path, int host_threads) { return ksb::write_tsv_impl(r, path, host_threads, false); }
int ksb_append_tsv(const ksb_result* r, const char* path, int host_threads) { return ksb::write_tsv_impl(r, path, host_threads, true); }
int ksb_write_tsv_records(const ksb_record* records, uint64_t n_rows, const char* path, int host_threads, int append) {
    if ((!records && n_rows) || !path) { ksb::set_error("ksb_write_tsv_records: null argument"); return KSB_ERR_ARG; }
    return ksb::write_tsv_rows(records, n_rows, path, host_threads, append != 0);
}

int ksb_write_index_files(const char* index_prefix, uint64_t n_colors, const uint64_t* color_ids, const uint64_t* offsets,
                          const uint32_t* members, uint64_t n_counts, const uint64_t* count_ids, const uint64_t* count_vals,
                          uint64_t n_totals, const uint32_t* total_ids, const uint32_t* total_vals) {
    if (!index_prefix || (n_colors && (!color_ids || !offsets)) || (n_counts && (!count_ids || !count_vals)) ||
        (n_totals && (!total_ids || !total_vals))) {
        ksb::set_error("ksb_write_index_files: null argument");
        return KSB_ERR_ARG;
    }
    const std::string P(index_prefix);
    auto put = [&](const std::string& path, const std::vector<uint8_t>& bytes) {
        FILE* f = fopen(path.c_str(), "wb");
        if (!f) { ksb::set_error("cannot write %s: %s", path.c_str(), strerror(errno)); return false; }
        const bool ok = bytes.empty() || fwrite(bytes.data(), 1, bytes.size(), f) == bytes.size();
        if (fclose(f) != 0 || !ok) { ksb::set_error("write error on %s", path.c_str()); return false; }
        return true;
    };
    try {
        std::vector<uint8_t> out;
        ksb::append_dump(out, total_ids, total_vals, n_totals, 4, 4);
        if (!put(P + "_groupID_to_kmerCount.bin", out)) return KSB_ERR_IO;
        out.clear();
        ksb::append_dump(out, count_ids, count_vals, n_counts, 8, 8);
        if (!put(P + "_color_count.bin", out)) return KSB_ERR_IO;
        out.clear();
        out.resize(8);
        std::memcpy(out.data(), &n_colors, 8);
        // stream the colour records in blocks so a 1 GB file does not need 2 GB of vector growth
        FILE* f = fopen((P + "_color_to_sources.bin").c_str(), "wb");
        if (!f) { ksb::set_error("cannot write %s_color_to_sources.bin: %s", index_prefix, strerror(errno)); return KSB_ERR_IO; }
        bool ok = true;
        for (uint64_t c = 0; c < n_colors && ok; c++) {
            const size_t at = out.size();
            out.resize(at + 8);
            std::memcpy(&out[at], &color_ids[c], 8);
            ksb::append_dump(out, members + offsets[c], nullptr, offsets[c + 1] - offsets[c], 4, 0);
            if (out.size() >= (1u << 24)) { ok = fwrite(out.data(), 1, out.size(), f) == out.size(); out.clear(); }
        }
        if (ok && !out.empty()) ok = fwrite(out.data(), 1, out.size(), f) == out.size();
        if (fclose(f) != 0 || !ok) { ksb::set_error("write error on %s_color_to_sources.bin", index_prefix); return KSB_ERR_IO; }
    } catch (const std::bad_alloc&) {
        ksb::set_error("out of host memory");
        return KSB_ERR_NOMEM;
    }
    return KSB_OK;
}

}  // extern "C"
