// Host loader: the reference's three index files (or an in-memory CSR) -> sample-compacted,
// member-sorted, size-binned CSR blob in pinned memory.
//
// Replaces, with identical observable semantics, the load phase of kSpider::pairwise:
//   load_colors_to_sources  src/pairwise.cpp:95-111  (colour id narrowed u64->u32, insert_or_assign:
//                                                    the LAST record with a given narrowed id wins)
//   load_colors_count       src/pairwise.cpp:113-121 (key AND value narrowed to u32, slot order,
//                                                    last wins; assert(size) must hold)
//   totals load             src/pairwise.cpp:166-170 (flat_hash_map<u32,u32>; assert(size))
//   colorsCount[colour]     src/pairwise.cpp:221     (missing colour -> weight 0)
// The per-colour work the reference does inside its hot loop that depends only on the index
// (ascending(), Combo enumeration order) is folded in here: members are sorted once.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <memory>
#include <numeric>
#include <thread>

#include <cuda_runtime_api.h>

#include "ksb_common.hpp"
#include "phmap_format.hpp"

namespace ksb {

// Ascending order of one colour's member ids (distinct: they come out of a hash SET, in slot order).  The samples of a colour
// of a clade- or family-structured index lie within a few hundred ids of each other, so the ids are dropped into a bitmap
// over [min, max] (at most 64 words on the stack) and read back in order with ctz -- O(n + range / 64) instead of a
// comparison sort of 40-500 elements, which was 60 % of the member extraction.  Wider ranges take std::sort.
static inline void sort_members(uint32_t* v, uint64_t n) {
    if (n < 2) return;
    uint32_t lo = v[0], hi = v[0];
    for (uint64_t i = 1; i < n; i++) { lo = v[i] < lo ? v[i] : lo; hi = v[i] > hi ? v[i] : hi; }
    const uint64_t range = (uint64_t)hi - lo + 1;
    if (range > 64 * 64 || range > 64 * n) { std::sort(v, v + n); return; }  // sparse in its range: the word scan would dominate
    uint64_t bits[64];
    const uint32_t words = (uint32_t)((range + 63) / 64);
    for (uint32_t w = 0; w < words; w++) bits[w] = 0;
    for (uint64_t i = 0; i < n; i++) { const uint32_t x = v[i] - lo; bits[x >> 6] |= 1ull << (x & 63); }
    uint64_t k = 0;
    for (uint32_t w = 0; w < words; w++)
        for (uint64_t b = bits[w]; b; b &= b - 1) v[k++] = lo + 64 * w + (uint32_t)__builtin_ctzll(b);
    if (k != n) std::sort(v, v + n);  // a repeated id (never from a set): keep every copy, as std::sort would
}

static thread_local std::string g_err;

void set_error(const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
}
const char* last_error() { return g_err.c_str(); }

void parallel_chunks(uint64_t n, int threads, const std::function<void(int, uint64_t, uint64_t)>& fn) {
    if (threads < 1) threads = 1;
    if ((uint64_t)threads > n) threads = n ? (int)n : 1;
    if (threads == 1) { fn(0, 0, n); return; }
    std::vector<std::thread> pool;
    pool.reserve(threads);
    for (int t = 0; t < threads; t++) {
        const uint64_t b = n * (uint64_t)t / (uint64_t)threads, e = n * (uint64_t)(t + 1) / (uint64_t)threads;
        pool.emplace_back(fn, t, b, e);
    }
    for (auto& th : pool) th.join();
}

void row_cuts(const uint64_t* row_updates, uint64_t N, uint64_t U, int W, uint64_t* cut) {
    if (W < 1) W = 1;
    for (int r = 0; r <= W; r++) cut[r] = N;
    cut[0] = 0;
    uint64_t acc = 0;
    int r = 1;
    for (uint64_t i = 0; i < N && r < W; i++) {
        acc += row_updates[i];
        // rank r starts after the first row at which the prefix reaches r/W of the updates
        while (r < W && (unsigned __int128)acc * (unsigned)W >= (unsigned __int128)U * (unsigned)r) cut[r++] = i + 1;
    }
    cut[W] = N;
}

// u32 id -> u32 value lookup; direct array when ids are small, open addressing otherwise.
class IdTable {
public:
    void init(uint64_t n_expected, uint32_t max_id) {
        dense_ = (uint64_t)max_id + 1 <= std::max<uint64_t>(n_expected * 8, 1u << 20);
        if (dense_) {
            val_.assign((uint64_t)max_id + 1, 0);
            used_.assign((uint64_t)max_id + 1, 0);
        } else {
            uint64_t cap = 64;
            while (cap < n_expected * 2 + 2) cap <<= 1;
            mask_ = cap - 1;
            key_.assign(cap, 0);
            val_.assign(cap, 0);
            used_.assign(cap, 0);
        }
    }
    // insert_or_assign; returns true if the key was already present
    bool put(uint32_t k, uint32_t v) {
        const uint64_t s = slot(k);
        const bool had = used_[s];
        used_[s] = 1;
        if (!dense_) key_[s] = k;
        val_[s] = v;
        return had;
    }
    bool get(uint32_t k, uint32_t* v) const {
        if (dense_) {
            if (k >= val_.size() || !used_[k]) return false;
            *v = val_[k];
            return true;
        }
        uint64_t h = mix(k) & mask_;
        while (used_[h]) {
            if (key_[h] == k) { *v = val_[h]; return true; }
            h = (h + 1) & mask_;
        }
        return false;
    }
    bool dense() const { return dense_; }
    void prefetch(uint32_t k) const { if (dense_ && k < val_.size()) { __builtin_prefetch(&val_[k]); __builtin_prefetch(&used_[k]); } }
private:
    static uint64_t mix(uint64_t x) { x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; return x; }
    uint64_t slot(uint32_t k) const {
        if (dense_) return k;
        uint64_t h = mix(k) & mask_;
        while (used_[h] && key_[h] != k) h = (h + 1) & mask_;
        return h;
    }
    bool dense_ = true;
    uint64_t mask_ = 0;
    std::vector<uint32_t> key_, val_;
    std::vector<uint8_t> used_;
};

static inline uint64_t align256(uint64_t x) { return (x + 255) & ~255ULL; }

// K2 items of one mid colour of s members: for every block of kMidCols columns, the rows below
// the block's last column, in chunks of kMidRows.
static inline uint64_t mid_items_of(uint64_t s) {
    uint64_t n = 0;
    for (uint64_t b0 = 0; b0 < s; b0 += kMidCols) {
        const uint64_t rows = std::min<uint64_t>(s - 1, b0 + kMidCols - 1);  // rows a < rows can pair with the block
        n += (rows + kMidRows - 1) / kMidRows;
    }
    return n;
}

// Intermediate form shared by both front ends: every colour of the file / input, members sorted
// ascending and de-duplicated (ORIGINAL sample ids), weight already narrowed.
// resize() without the serial zero-fill (and its page faults on one thread): the member array is written in full, in
// parallel, right after it is sized
template <class T>
struct NoInitAlloc : std::allocator<T> {
    template <class U> struct rebind { using other = NoInitAlloc<U>; };
    NoInitAlloc() = default;
    template <class U> NoInitAlloc(const NoInitAlloc<U>&) {}
    template <class U> void construct(U* p) noexcept { ::new (static_cast<void*>(p)) U; }
    template <class U, class... A> void construct(U* p, A&&... a) { ::new (static_cast<void*>(p)) U(std::forward<A>(a)...); }
};
struct RawIndex {
    std::vector<uint64_t> off;      // [C+1]
    std::vector<uint32_t, NoInitAlloc<uint32_t>> members;  // [M]
    std::vector<uint32_t> weights;  // [C]
    std::vector<uint32_t> total_ids, total_vals;  // slot / input order
    uint64_t n_colors_file = 0;
};

// KSPIDER_LOAD_TIMERS=1: phases of the loader on stderr
struct LoadTimer {
    bool on = std::getenv("KSPIDER_LOAD_TIMERS") && std::atoi(std::getenv("KSPIDER_LOAD_TIMERS")) != 0;
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    void lap(const char* what) {
        if (!on) return;
        const auto t1 = std::chrono::steady_clock::now();
        std::fprintf(stderr, "ksb_load %-28s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(t1 - t0).count());
        t0 = t1;
    }
};

static int build_index(RawIndex& raw, int threads, ksb_index** out) {
    LoadTimer bt;
    std::unique_ptr<ksb_index> ix(new ksb_index());
    IndexDesc& d = ix->desc;
    std::memset(&d, 0, sizeof d);
    d.magic = kIndexMagic;
    d.n_colors_file = raw.n_colors_file;
    const uint64_t C = raw.weights.size();

    // ---- sample compaction: ascending original id -> 0..N-1 (order preserving)
    uint32_t max_id = 0;
    for (uint32_t v : raw.total_ids) max_id = std::max(max_id, v);
    {
        std::vector<uint32_t> tmax(threads > 0 ? threads : 1, 0);
        parallel_chunks(raw.members.size(), threads, [&](int t, uint64_t b, uint64_t e) {
            uint32_t m = 0;
            for (uint64_t i = b; i < e; i++) m = std::max(m, raw.members[i]);
            tmax[t] = m;
        });
        for (uint32_t v : tmax) max_id = std::max(max_id, v);
    }
    std::vector<uint32_t> sample_ids;
    std::vector<uint32_t> remap;  // dense path: original id -> compact
    const bool dense_ids = (uint64_t)max_id + 1 <= (1ull << 28);
    if (dense_ids) {
        std::vector<uint8_t> present((uint64_t)max_id + 1, 0);
        for (uint32_t v : raw.total_ids) present[v] = 1;
        parallel_chunks(raw.members.size(), threads, [&](int, uint64_t b, uint64_t e) {
            // benign same-value race; read before write: with few samples the flags share a handful of cache lines, and
            // unconditional stores from every thread made this loop the slowest phase of the loader
            for (uint64_t i = b; i < e; i++) { const uint32_t m = raw.members[i]; if (!present[m]) present[m] = 1; }
        });
        remap.assign((uint64_t)max_id + 1, 0);
        for (uint64_t v = 0; v <= max_id; v++)
            if (present[v]) { remap[v] = (uint32_t)sample_ids.size(); sample_ids.push_back((uint32_t)v); }
    } else {
        sample_ids.assign(raw.members.begin(), raw.members.end());
        sample_ids.insert(sample_ids.end(), raw.total_ids.begin(), raw.total_ids.end());
        std::sort(sample_ids.begin(), sample_ids.end());
        sample_ids.erase(std::unique(sample_ids.begin(), sample_ids.end()), sample_ids.end());
    }
    const uint64_t N = sample_ids.size();
    auto compact = [&](uint32_t id) -> uint32_t {
        if (dense_ids) return remap[id];
        return (uint32_t)(std::lower_bound(sample_ids.begin(), sample_ids.end(), id) - sample_ids.begin());
    };
    d.n_samples = N;

    const int T = std::max(1, threads);

    bt.lap("  compaction");
    // ---- routing of the int8 Gram path (K4), per sample window.  A colour whose members all lie in a
    // window of `width` samples costs ~width^2 * planes tensor MACs there against s^2/2 reds on K1/K2,
    // so it is a candidate when s >= kGramCrossover * width * sqrt(planes) (and >= kGramMinSize) for
    // the SMALLEST window of the multi-level grid that contains it; a window is used only if its
    // candidates together beat the atomic path including the window's fixed costs.
    // KSPIDER_GRAM = off | auto (default) | <min size> (forces every colour of at least that size).
    const uint64_t NOKEY = ~0ull;
    std::vector<uint64_t> gkey;        // per colour: window key, then (once the windows are fixed) its Gram CSR index
    std::vector<GramWin> gwins;
    std::vector<uint64_t> gram_list;   // colours in Gram CSR order (grouped by window)
    uint64_t gram_min = 0, gram_planes = 1;
    if (N >= 2) {
        const char* env = std::getenv("KSPIDER_GRAM");
        long forced = -1;
        bool off = false;
        if (env && *env) {
            if (!std::strcmp(env, "off")) off = true;
            else if (std::strcmp(env, "auto")) { forced = std::strtol(env, nullptr, 10); if (forced < 2) off = true; }
        }
        if (!off) {
            const uint64_t npad = (N + kGramTile - 1) / kGramTile * kGramTile;
            const uint64_t s_any = forced >= 2 ? (uint64_t)forced : kGramMinSize;
            std::vector<uint32_t> tw(T, 0);
            parallel_chunks(C, T, [&](int t, uint64_t b, uint64_t e) {
                uint32_t m = 0;
                for (uint64_t c = b; c < e; c++)
                    if (raw.off[c + 1] - raw.off[c] >= s_any) m = std::max(m, raw.weights[c]);
                tw[t] = m;
            });
            uint32_t wmax = 0;
            for (uint32_t v : tw) wmax = std::max(wmax, v);
            const uint64_t planes = wmax < (1u << 8) ? 1 : (wmax < (1u << 16) ? 2 : 4);
            struct Level { uint64_t width, stride, smin; bool top; };
            std::vector<Level> levels;
            auto smin_of = [&](uint64_t width) -> uint64_t {
                if (forced >= 2) return (uint64_t)forced;
                return std::max<uint64_t>(kGramMinSize, (uint64_t)std::ceil(kGramCrossover * (double)width * std::sqrt((double)planes)));
            };
            for (uint64_t w = kGramTile;; w <<= 1) {
                if (w >= npad) { if (npad <= kGramMaxSamples) levels.push_back({npad, npad, smin_of(npad), true}); break; }
                if (w > kGramMaxSamples) break;
                levels.push_back({w, w / 2, smin_of(w), false});
            }
            gkey.assign(C, NOKEY);
            std::vector<uint64_t> tn(T, 0);
            std::vector<uint32_t> cfirst(C), clast(C);  // compact first / last member of every candidate: read once, here, by all threads
            parallel_chunks(C, T, [&](int t, uint64_t b, uint64_t e) {
                uint64_t n = 0;
                for (uint64_t c = b; c < e; c++) {
                    const uint64_t sz = raw.off[c + 1] - raw.off[c];
                    if (sz < 2 || sz < s_any || raw.weights[c] == 0) continue;
                    const uint64_t first = compact(raw.members[raw.off[c]]), last = compact(raw.members[raw.off[c + 1] - 1]);
                    cfirst[c] = (uint32_t)first;
                    clast[c] = (uint32_t)last;
                    for (size_t l = 0; l < levels.size(); l++) {
                        const Level& lv = levels[l];
                        const uint64_t x0 = lv.top ? 0 : first / lv.stride * lv.stride;
                        if (last >= x0 + lv.width) continue;
                        if (sz >= lv.smin) { gkey[c] = (x0 << 8) | l; n++; }
                        break;  // larger windows only raise the bar
                    }
                }
                tn[t] = n;
            });
            uint64_t n_cand = 0;
            for (uint64_t v : tn) n_cand += v;
            // (window key, colour) sorted by key, input order inside a window.  The keys are few (one per window) and the
            // colours come in ascending order, so this is a counting sort by key -- a comparison sort of a million pairs
            // was the largest serial piece of the loader.
            std::vector<std::pair<uint64_t, uint64_t>> cand(n_cand);
            {
                std::vector<uint64_t> keys;  // distinct keys, ascending
                {
                    std::vector<std::vector<uint64_t>> tk(T);
                    parallel_chunks(C, T, [&](int t, uint64_t b, uint64_t e) {
                        std::vector<uint64_t>& k = tk[t];
                        uint64_t last = NOKEY;
                        for (uint64_t c = b; c < e; c++)
                            if (gkey[c] != NOKEY && gkey[c] != last) { k.push_back(gkey[c]); last = gkey[c]; }  // runs of one window collapse
                        std::sort(k.begin(), k.end());
                        k.erase(std::unique(k.begin(), k.end()), k.end());
                    });
                    for (auto& k : tk) keys.insert(keys.end(), k.begin(), k.end());
                    std::sort(keys.begin(), keys.end());
                    keys.erase(std::unique(keys.begin(), keys.end()), keys.end());
                }
                const size_t K = keys.size();
                auto slot_of = [&](uint64_t key) { return (size_t)(std::lower_bound(keys.begin(), keys.end(), key) - keys.begin()); };
                // per thread and key: how many candidates; then every thread writes its own at its offsets
                std::vector<std::vector<uint32_t>> cnt(T, std::vector<uint32_t>(K, 0));
                std::vector<uint64_t> cb(T, 0), ce(T, 0);
                parallel_chunks(C, T, [&](int t, uint64_t b, uint64_t e) {
                    cb[t] = b; ce[t] = e;
                    uint64_t last = NOKEY; size_t ls = 0;
                    for (uint64_t c = b; c < e; c++)
                        if (gkey[c] != NOKEY) { if (gkey[c] != last) { last = gkey[c]; ls = slot_of(last); } cnt[t][ls]++; }
                });
                std::vector<std::vector<uint64_t>> at(T, std::vector<uint64_t>(K, 0));
                uint64_t run = 0;
                for (size_t k = 0; k < K; k++)
                    for (int t = 0; t < T; t++) { at[t][k] = run; run += cnt[t][k]; }
                parallel_chunks((uint64_t)T, T, [&](int, uint64_t tb, uint64_t te) {
                    for (uint64_t t = tb; t < te; t++) {
                        uint64_t last = NOKEY; size_t ls = 0;
                        for (uint64_t c = cb[t]; c < ce[t]; c++)
                            if (gkey[c] != NOKEY) { if (gkey[c] != last) { last = gkey[c]; ls = slot_of(last); } cand[at[t][ls]++] = {gkey[c], c}; }
                    }
                });
            }
            // per window: keep it only if the tensor path wins (bitmap build + MMAs + per-item overhead + one pass of reds)
            double gain = 0.0;
            std::vector<std::pair<uint64_t, uint64_t>> kept;  // [begin, end) ranges of cand
            std::vector<std::pair<uint64_t, uint64_t>> kept_span;  // (first sample, width) of the kept windows
            std::vector<uint64_t> kept_extent;                      // samples from the first to the last touched one
            for (uint64_t b = 0; b < cand.size();) {
                uint64_t e = b, upd = 0, mem = 0, lo = ~0ull, hi = 0;
                while (e < cand.size() && cand[e].first == cand[b].first) {
                    const uint64_t c = cand[e].second, sz = raw.off[c + 1] - raw.off[c];
                    upd += sz * (sz - 1) / 2;
                    mem += sz;
                    lo = std::min<uint64_t>(lo, cfirst[c]);
                    hi = std::max<uint64_t>(hi, (uint64_t)clast[c] + 1);
                    e++;
                }
                const Level& lv = levels[cand[b].first & 0xFF];
                // The grid window only decides which colours go together; the window K4 works on is the range
                // of samples its colours really touch, from the first one (tiles may start at any sample) to the
                // last, rounded up to whole tiles.  A 1000-sample family that the grid puts into a 2048-wide window
                // (span <= width / 2) needs 8 x 8 tiles, not 16 x 16: 36 tile pairs instead of 136.
                const uint64_t span_w = std::min<uint64_t>(lv.width, (hi - lo + kGramTile - 1) / kGramTile * kGramTile);
                const double width = (double)span_w, tiles = width / kGramTile;
                const double kpad = (double)((e - b + kGramKBlock - 1) / kGramKBlock * kGramKBlock);
                const double rate = lv.width >= 4 * kGramTile ? kGramMacRate : 0.5 * kGramMacRate;
                const double t_gram = width * width * kpad * (double)planes / rate + (double)mem / 1.0e11 +
                                      tiles * (tiles + 1) / 2 / (double)(4 / planes) * 30e-9 + width * width / 2 / kGramRedRate;
                const double t_red = (double)upd / kGramRedRate;
                if (forced >= 2 || t_gram < t_red) { kept.emplace_back(b, e); kept_span.emplace_back(lo, span_w); kept_extent.push_back(hi - lo); gain += t_red - t_gram; }
                b = e;
            }
            if (forced < 2 && gain < 30e-6) { kept.clear(); kept_span.clear(); kept_extent.clear(); }  // launches + memsets of the path itself
            std::fill(gkey.begin(), gkey.end(), NOKEY);
            uint64_t smallest = ~0ull;
            const bool shrink = !(std::getenv("KSPIDER_GRAM_SHRINK") && std::atoi(std::getenv("KSPIDER_GRAM_SHRINK")) == 0);
            for (size_t ki = 0; ki < kept.size(); ki++) {
                auto& r = kept[ki];
                const Level& lv = levels[cand[r.first].first & 0xFF];
                GramWin w;
                w.c0 = gram_list.size();
                w.n = (uint32_t)(r.second - r.first);
                if (shrink) {
                    w.x0 = (uint32_t)kept_span[ki].first;
                    w.width = (uint32_t)kept_span[ki].second;
                    w.extent = (uint32_t)std::min<uint64_t>(kept_extent[ki], w.width);
                } else {
                    w.x0 = (uint32_t)(cand[r.first].first >> 8);
                    w.width = (uint32_t)std::min<uint64_t>(lv.width, (npad - w.x0 + kGramTile - 1) / kGramTile * kGramTile);
                    w.extent = 0;
                }
                gwins.push_back(w);
                for (uint64_t i = r.first; i < r.second; i++) {
                    gkey[cand[i].second] = gram_list.size();
                    gram_list.push_back(cand[i].second);
                    smallest = std::min<uint64_t>(smallest, raw.off[cand[i].second + 1] - raw.off[cand[i].second]);
                }
            }
            if (!gram_list.empty()) { gram_min = smallest; gram_planes = planes; }
        }
    }
    const bool any_gram = !gram_list.empty();
    d.gram_min_size = gram_min;
    d.gram_planes = gram_planes;

    bt.lap("  gram routing");
    // ---- locality order: when the accumulator cannot be L2-resident, colours are placed in the blob
    // by ascending first member (= the triangle row of their first updates), so the colours the SMs
    // work on at the same time hit the same neighbourhood of rows and the reds stay in L2 instead of
    // each one costing an HBM sector read + write-back.  KSPIDER_SORT = auto (default) | on | off.
    std::vector<uint64_t> order;  // empty = input order
    {
        const char* env = std::getenv("KSPIDER_SORT");
        bool on = N >= 2 && (double)N * (double)(N - 1) * 4.0 > 48.0 * 1048576.0;
        if (env && !std::strcmp(env, "on")) on = true;
        if (env && !std::strcmp(env, "off")) on = false;
        if (on && C > 1) {  // stable counting sort by compact first member (empty colours first)
            std::vector<uint64_t> start(N + 2, 0);
            std::vector<uint32_t> keys(C);  // a cache miss per colour (its first member): taken by all threads, once
            parallel_chunks(C, T, [&](int, uint64_t b, uint64_t e) {
                for (uint64_t c = b; c < e; c++) keys[c] = raw.off[c + 1] > raw.off[c] ? 1u + compact(raw.members[raw.off[c]]) : 0u;
            });
            for (uint64_t c = 0; c < C; c++) start[(uint64_t)keys[c] + 1]++;
            for (uint64_t k = 0; k <= N; k++) start[k + 1] += start[k];
            order.resize(C);
            for (uint64_t c = 0; c < C; c++) order[start[keys[c]]++] = c;
        }
    }
    auto ord = [&](uint64_t ci) -> uint64_t { return order.empty() ? ci : order[ci]; };

    bt.lap("  locality order");
    // ---- classify colours; per-chunk class counts so the fill can run in parallel, in blob order
    const uint32_t K_MID = kSmallMax + 1, K_ZERO = kSmallMax + 2, K_GRAM = kSmallMax + 3, K_N = kSmallMax + 4;
    std::vector<std::vector<uint64_t>> cnt(T, std::vector<uint64_t>(K_N, 0));      // colours per class
    std::vector<std::vector<uint64_t>> mcnt(T, std::vector<uint64_t>(K_N, 0));     // members per class
    std::vector<uint64_t> chunk_b(T + 1, C);
    std::vector<uint64_t> icnt(T, 0), ibase(T, 0);  // K2 items per chunk
    auto klass = [&](uint64_t c) -> uint32_t {
        const uint64_t s = raw.off[c + 1] - raw.off[c];
        if (s < 2) return 0;
        if (raw.weights[c] == 0) return K_ZERO;
        if (any_gram && gkey[c] != NOKEY) return K_GRAM;
        return s <= kSmallMax ? (uint32_t)s : K_MID;
    };
    parallel_chunks(C, T, [&](int t, uint64_t b, uint64_t e) {
        chunk_b[t] = b;
        for (uint64_t ci = b; ci < e; ci++) {
            const uint64_t c = ord(ci);
            const uint32_t k = klass(c);
            if (!k) continue;
            cnt[t][k]++;
            const uint64_t s = raw.off[c + 1] - raw.off[c];
            mcnt[t][k] += (k == K_MID) ? ((s + 7) & ~7ULL) : s;  // mid colours start 16-byte aligned (u16 or u32 members)
            if (k == K_MID) icnt[t] += mid_items_of(s);
        }
    });
    uint64_t n_items = 0;
    for (int t = 0; t < T; t++) { ibase[t] = n_items; n_items += icnt[t]; }
    std::vector<uint64_t> tot(K_N, 0), mtot(K_N, 0);
    std::vector<std::vector<uint64_t>> base(T, std::vector<uint64_t>(K_N, 0)), mbase(T, std::vector<uint64_t>(K_N, 0));
    for (uint32_t k = 0; k < K_N; k++)
        for (int t = 0; t < T; t++) {
            base[t][k] = tot[k]; mbase[t][k] = mtot[k];
            tot[k] += cnt[t][k]; mtot[k] += mcnt[t][k];
        }
    uint64_t moff = 0, woff = 0;
    for (uint32_t s = 2; s <= kSmallMax; s++) {
        SmallClass& sc = d.cls[s];
        sc.s = s;
        sc.pairs = s * (s - 1) / 2;
        sc.magic = (uint32_t)(((1ull << 32) + sc.pairs - 1) / sc.pairs);
        sc.n = tot[s];
        sc.member_off = moff;
        sc.weight_off = woff;
        moff += tot[s] * s;
        woff += tot[s];
        d.n_updates += tot[s] * sc.pairs;
    }
    d.n_small_colors = woff;
    d.n_small_updates = d.n_updates;
    d.n_small_members = moff;
    d.n_mid_colors = tot[K_MID];
    d.n_mid_members_padded = mtot[K_MID];
    d.n_mid_items = n_items;
    d.n_zero_colors = tot[K_ZERO];
    d.n_zero_members = mtot[K_ZERO];
    d.n_gram_colors = tot[K_GRAM];
    d.n_gram_members = mtot[K_GRAM];
    d.n_colors = d.n_small_colors + d.n_mid_colors + d.n_zero_colors + d.n_gram_colors;

    bt.lap("  classify");
    // ---- member width: u16 when every compact id fits (halves the blob = the H2D / broadcast time);
    // KSPIDER_MEMBER_BITS=32 forces u32 (tests)
    {
        const char* mb = std::getenv("KSPIDER_MEMBER_BITS");
        d.member_bits = (N <= 65536 && !(mb && std::atoi(mb) == 32)) ? 16 : 32;
    }
    const uint64_t MB = d.member_bits / 8;
    {   // the small-colour and Gram sections (most of the blob): 10-bit packed members when N <= 1024, 12-bit
        // when N <= 4096, and weights as narrow as the largest one allows.
        // KSPIDER_MEMBER_BITS=12|16|32 / KSPIDER_WEIGHT_BITS=32 override.
        const char* mb = std::getenv("KSPIDER_MEMBER_BITS");
        const int mbv = mb ? std::atoi(mb) : 0;
        d.packed_bits = (N <= 4096 && mbv != 16 && mbv != 32) ? ((N <= 1024 && mbv != 12) ? 10 : 12) : d.member_bits;
        std::vector<uint32_t> tw(T, 0);
        parallel_chunks(C, T, [&](int t, uint64_t b, uint64_t e) {
            uint32_t m = 0;
            for (uint64_t c = b; c < e; c++)
                if (raw.off[c + 1] - raw.off[c] >= 2) m = std::max(m, raw.weights[c]);
            tw[t] = m;
        });
        uint32_t wmax = 0;
        for (uint32_t v : tw) wmax = std::max(wmax, v);
        const char* wb = std::getenv("KSPIDER_WEIGHT_BITS");
        d.weight_bits = (wb && std::atoi(wb) == 32) ? 32 : (wmax < (1u << 8) ? 8 : (wmax < (1u << 16) ? 16 : 32));
    }
    {   // Gram members as offsets inside their window when that is narrower than a compact id (KSPIDER_GRAM_REL=0|10|12|16
        // overrides; a width the widest window does not fit is ignored).  Default: 16 bits when ids need 32 -- the packed
        // widths cost kb_build its 16-byte loads, so they are only taken on request.
        uint64_t maxw = 0;
        for (const GramWin& w : gwins) maxw = std::max<uint64_t>(maxw, w.width);
        const char* gr = std::getenv("KSPIDER_GRAM_REL");
        const int grv = gr ? std::atoi(gr) : -1;
        uint64_t rel = 0;
        if (any_gram) {
            if (grv == 10 || grv == 12 || grv == 16) rel = (uint64_t)grv;
            else if (grv != 0 && d.packed_bits == 32) rel = 16;
            if (rel && (maxw > (1ull << rel) || rel >= d.packed_bits)) rel = 0;
        }
        d.gram_rel_bits = rel;
    }
    {   // Gram colours as ready-made bitmaps when that is the smaller blob (KSPIDER_GRAM_FORM=csr|bitmap overrides)
        uint64_t rows = 0;
        for (const GramWin& w : gwins) rows += gram_win_rows(w);
        const uint64_t csr_bytes = (gram_member_bits(d) * mtot[K_GRAM] + 7) / 8 + 8 * (uint64_t)gram_list.size();
        const char* gf = std::getenv("KSPIDER_GRAM_FORM");
        bool bitmap = any_gram && 16 * rows < csr_bytes;
        if (gf && !std::strcmp(gf, "csr")) bitmap = false;
        if (gf && !std::strcmp(gf, "bitmap")) bitmap = any_gram;
        d.gram_form = bitmap ? 1 : 0;
        d.gram_bitmap_rows = bitmap ? rows : 0;
        if (bitmap) d.gram_rel_bits = 0;
    }
    const uint64_t PB = d.packed_bits, WB = d.weight_bits / 8, GB = gram_member_bits(d);
    auto packed_bytes = [&](uint64_t n) { return (PB * n + 7) / 8 + 8; };
    auto gram_bytes = [&](uint64_t n) { return (GB * n + 7) / 8 + 8; };

    bt.lap("  member width");
    // ---- payload layout
    uint64_t p = 0;
    auto section = [&](uint64_t bytes) { const uint64_t o = p; p = align256(p + bytes); return o; };
    d.off_totals = section(4 * N);
    d.off_sample_ids = section(4 * N);
    d.off_row_updates = section(8 * N);
    d.off_band_lo = section(4 * N);
    d.off_band_hi = section(4 * N);
    d.off_bandstart = section(8 * (N + 1));
    d.off_rowbase = section(8 * N);
    d.off_small_members = section(packed_bytes(d.n_small_members));
    d.off_small_weights = section(WB * d.n_small_colors);
    d.off_mid_offsets = section(8 * (d.n_mid_colors + 1));
    d.off_mid_sizes = section(4 * d.n_mid_colors);
    d.off_mid_members = section(MB * d.n_mid_members_padded + 64);
    d.off_mid_weights = section(4 * d.n_mid_colors);
    d.off_mid_items = section(sizeof(MidItem) * d.n_mid_items);
    d.off_zero_offsets = section(8 * (d.n_zero_colors + 1));
    d.off_zero_members = section(MB * d.n_zero_members);
    d.off_gram_offsets = section(d.gram_form ? 8 * (gwins.size() + 1) : 8 * (d.n_gram_colors + 1));
    d.off_gram_members = section(d.gram_form ? 16 * (d.gram_bitmap_rows + kGramTile) : gram_bytes(d.n_gram_members));
    d.off_gram_weights = section(WB * d.n_gram_colors);
    d.n_gram_wins = gwins.size();
    for (const GramWin& w : gwins) {
        d.n_gram_macs += (uint64_t)w.n * ((uint64_t)w.width * (w.width - 1) / 2) * gram_planes;
        const uint64_t ext = w.extent ? w.extent : w.width;
        d.n_gram_cells += ext * (ext - 1) / 2;
    }
    d.off_gram_wins = section(sizeof(GramWin) * d.n_gram_wins);
    d.payload_bytes = std::max<uint64_t>(p, 256);

    void* mem = nullptr;
    if (cudaMallocHost(&mem, d.payload_bytes) == cudaSuccess) {
        ix->pinned = true;
    } else {
        (void)cudaGetLastError();  // no device / no driver: plain memory is fine for host-only use
        if (posix_memalign(&mem, 4096, d.payload_bytes) != 0) { set_error("out of host memory (%llu bytes)", (unsigned long long)d.payload_bytes); return KSB_ERR_NOMEM; }
        ix->pinned = false;
    }
    ix->payload = static_cast<uint8_t*>(mem);
    std::memset(ix->payload, 0, d.payload_bytes);
    uint8_t* P = ix->payload;
    uint32_t* totals = reinterpret_cast<uint32_t*>(P + d.off_totals);
    uint32_t* sids = reinterpret_cast<uint32_t*>(P + d.off_sample_ids);
    uint64_t* row_updates = reinterpret_cast<uint64_t*>(P + d.off_row_updates);
    uint32_t* band_lo = reinterpret_cast<uint32_t*>(P + d.off_band_lo);
    uint32_t* band_hi = reinterpret_cast<uint32_t*>(P + d.off_band_hi);
    uint8_t* small_m = P + d.off_small_members;  // member sections: u16 or u32 elements
    uint8_t* small_w = P + d.off_small_weights;  // weight sections: u8 / u16 / u32 elements
    auto put_w = [&](uint8_t* sec, uint64_t i, uint32_t w) {
        if (WB == 1) sec[i] = (uint8_t)w;
        else if (WB == 2) reinterpret_cast<uint16_t*>(sec)[i] = (uint16_t)w;
        else reinterpret_cast<uint32_t*>(sec)[i] = w;
    };
    // element i of a small / Gram member section.  12-bit packing: element i = bits [12i, 12i+12); its
    // full byte is stored plainly, its half byte is OR-ed atomically (the other half may belong to a
    // neighbouring colour that another thread is writing; the payload starts zeroed).  10-bit packing:
    // element i = bits [10i, 10i+10) — 4 elements per 5 bytes, so an element always lies in 2 bytes.
    auto put_m = [&](uint8_t* sec, uint64_t bits, uint64_t i, uint32_t v) {
        if (bits == 32) { reinterpret_cast<uint32_t*>(sec)[i] = v; return; }
        if (bits == 16) { reinterpret_cast<uint16_t*>(sec)[i] = (uint16_t)v; return; }
        if (bits == 10) {
            uint8_t* p = sec + ((5 * i) >> 2);
            const uint32_t x = (v & 0x3FFu) << ((i & 3) << 1);
            __atomic_fetch_or(p, (uint8_t)x, __ATOMIC_RELAXED);
            __atomic_fetch_or(p + 1, (uint8_t)(x >> 8), __ATOMIC_RELAXED);
            return;
        }
        uint8_t* p = sec + ((3 * i) >> 1);
        if ((i & 1) == 0) {
            p[0] = (uint8_t)v;
            __atomic_fetch_or(p + 1, (uint8_t)(v >> 8), __ATOMIC_RELAXED);
        } else {
            __atomic_fetch_or(p, (uint8_t)((v & 0xFu) << 4), __ATOMIC_RELAXED);
            p[1] = (uint8_t)(v >> 4);
        }
    };
    uint64_t* mid_off = reinterpret_cast<uint64_t*>(P + d.off_mid_offsets);
    uint8_t* mid_m = P + d.off_mid_members;
    uint32_t* mid_w = reinterpret_cast<uint32_t*>(P + d.off_mid_weights);
    uint32_t* mid_sz = reinterpret_cast<uint32_t*>(P + d.off_mid_sizes);
    MidItem* mid_items = reinterpret_cast<MidItem*>(P + d.off_mid_items);
    uint64_t* zero_off = reinterpret_cast<uint64_t*>(P + d.off_zero_offsets);
    uint8_t* zero_m = P + d.off_zero_members;
    uint64_t* gram_off = reinterpret_cast<uint64_t*>(P + d.off_gram_offsets);
    uint8_t* gram_m = P + d.off_gram_members;
    uint8_t* gram_w = P + d.off_gram_weights;

    std::copy(sample_ids.begin(), sample_ids.end(), sids);
    if (!gwins.empty()) std::memcpy(P + d.off_gram_wins, gwins.data(), sizeof(GramWin) * gwins.size());
    if (!d.gram_form)
        for (uint64_t g = 0, o = 0; g < gram_list.size(); g++) {  // Gram CSR: colours grouped by window
            gram_off[g] = o;
            o += raw.off[gram_list[g] + 1] - raw.off[gram_list[g]];
        }
    std::vector<uint32_t> gram_x0;  // per Gram colour (CSR order): what its stored members are relative to
    if (d.gram_rel_bits) {
        gram_x0.resize(gram_list.size());
        for (const GramWin& w : gwins) std::fill(gram_x0.begin() + w.c0, gram_x0.begin() + w.c0 + w.n, w.x0);
    }
    if (d.gram_form) {  // bitmap form: gram_off = first bitmap row of every window
        uint64_t row = 0;
        for (size_t wi = 0; wi < gwins.size(); wi++) {
            gram_off[wi] = row;
            row += gram_win_rows(gwins[wi]);
        }
        gram_off[gwins.size()] = row;
    }
    uint32_t* gram_bits = reinterpret_cast<uint32_t*>(gram_m);
    // groupID_to_kmerCount[id]: a map, so one value per id; missing -> 0 (operator[] default)
    for (size_t i = 0; i < raw.total_ids.size(); i++) totals[compact(raw.total_ids[i])] = raw.total_vals[i];

    bt.lap("  layout + alloc + memset");
    // ---- fill (parallel, deterministic placement = input order within each class)
    // per thread and sample, in ONE 24-byte record (one cache line touched per member instead of four arrays'): updates whose
    // row it is, sum of the weights of its colours, and the columns it is ever paired with, [lo, hi)
    struct RowAcc { uint64_t rows = 0, wsum = 0; uint32_t lo = 0xFFFFFFFFu, hi = 0; };
    std::vector<std::vector<RowAcc>> t_acc(T);
    std::vector<uint64_t> t_maxs(T, 0), t_mid_updates(T, 0), t_mid_members(T, 0), t_gram_updates(T, 0);
    parallel_chunks(C, T, [&](int t, uint64_t b, uint64_t e) {
        std::vector<RowAcc>& acc = t_acc[t];
        acc.assign(N, RowAcc());
        std::vector<uint64_t> at = base[t], mat = mbase[t];
        uint64_t maxs = 0, midu = 0, midm = 0, gramu = 0, iat = ibase[t];
        std::vector<uint32_t> cm;  // compact members of the current colour
        for (uint64_t ci = b; ci < e; ci++) {
            const uint64_t c = ord(ci);
            const uint32_t k = klass(c);
            if (!k) continue;
            const uint64_t s = raw.off[c + 1] - raw.off[c];
            const uint32_t* src = &raw.members[raw.off[c]];
            const uint32_t w = raw.weights[c];
            uint8_t* sec;        // member section of this colour and its first element there
            uint64_t e0, pbits = PB;
            uint32_t rel0 = 0;   // Gram members may be stored relative to their window
            bool in_bitmap = false;    // bitmap form: the colour's members are not stored as a list (its K-block is filled below)
            const bool packed = (k == K_GRAM) || (k != K_MID && k != K_ZERO);
            if (k == K_MID) {
                sec = mid_m;
                e0 = mat[k];
                mid_off[at[k]] = mat[k];
                mid_sz[at[k]] = (uint32_t)s;
                mid_w[at[k]] = w;
                midu += s * (s - 1) / 2;
                midm += s;
                for (uint64_t b0 = 0; b0 < s; b0 += kMidCols) {
                    const uint64_t rows = std::min<uint64_t>(s - 1, b0 + kMidCols - 1);
                    for (uint64_t a0 = 0; a0 < rows; a0 += kMidRows)
                        mid_items[iat++] = MidItem{(uint32_t)at[k], (uint32_t)b0, (uint32_t)a0, 0};
                }
            } else if (k == K_ZERO) {
                sec = zero_m;
                e0 = mat[k];
                zero_off[at[k]] = mat[k];
            } else if (k == K_GRAM) {
                sec = gram_m;
                if (d.gram_form) {
                    in_bitmap = true;
                    e0 = 0;
                } else {
                    e0 = gram_off[gkey[c]];
                    pbits = GB;
                    if (d.gram_rel_bits) rel0 = gram_x0[gkey[c]];
                }
                put_w(gram_w, gkey[c], w);
                gramu += s * (s - 1) / 2;
            } else {
                const SmallClass& sc = d.cls[k];
                sec = small_m;
                e0 = sc.member_off + at[k] * s;
                put_w(small_w, sc.weight_off + at[k], w);
            }
            cm.resize(s);
            for (uint64_t i = 0; i < s; i++) {
                const uint32_t ci = compact(src[i]);
                cm[i] = ci;
                if (in_bitmap) {}  // the K-blocks are filled below, one owner per K-block, without atomics
                else if (packed) put_m(sec, pbits, e0 + i, ci - rel0);
                else if (MB == 2) reinterpret_cast<uint16_t*>(sec)[e0 + i] = (uint16_t)ci;
                else reinterpret_cast<uint32_t*>(sec)[e0 + i] = ci;
            }
            // members are ascending: row cm[i] pairs with columns cm[i+1] .. cm[s-1]
            const uint32_t last1 = cm[s - 1] + 1;
            for (uint64_t i = 0; i < s; i++) {
                RowAcc& a = acc[cm[i]];
                if (k != K_ZERO) { a.rows += s - 1 - i; a.wsum += w; }
                if (i + 1 < s) { a.lo = std::min(a.lo, cm[i + 1]); a.hi = std::max(a.hi, last1); }
            }
            maxs = std::max(maxs, s);
            at[k]++;
            mat[k] += (k == K_MID) ? ((s + 7) & ~7ULL) : s;
        }
        t_maxs[t] = maxs;
        t_mid_updates[t] = midu;
        t_mid_members[t] = midm;
        t_gram_updates[t] = gramu;
    });
    if (!d.gram_form) gram_off[d.n_gram_colors] = d.n_gram_members;
    if (d.gram_form) {
        // Bitmaps of the Gram windows: a K-block's rows belong to its 128 colours alone, so the K-blocks are dealt to the
        // threads and filled with plain ORs (73 M locked ORs from the colour loop above were a third of its time).
        std::vector<std::pair<uint32_t, uint32_t>> kbs;  // (window, K-block inside it)
        for (size_t wi = 0; wi < gwins.size(); wi++)
            for (uint32_t kb = 0; kb < (gwins[wi].n + kGramKBlock - 1) / kGramKBlock; kb++) kbs.emplace_back((uint32_t)wi, kb);
        parallel_chunks(kbs.size(), T, [&](int, uint64_t b, uint64_t e) {
            for (uint64_t q = b; q < e; q++) {
                const GramWin& gwn = gwins[kbs[q].first];
                const uint32_t kb = kbs[q].second;
                uint32_t* rows = gram_bits + 4 * (gram_off[kbs[q].first] + (uint64_t)kb * gram_win_stride(gwn));
                const uint64_t kn = std::min<uint64_t>(kGramKBlock, gwn.n - (uint64_t)kb * kGramKBlock);
                for (uint64_t k = 0; k < kn; k++) {
                    const uint64_t c = gram_list[gwn.c0 + (uint64_t)kb * kGramKBlock + k];
                    const uint32_t bit = 1u << (k & 31u);
                    uint32_t* col = rows + (k >> 5);
                    for (uint64_t i = raw.off[c]; i < raw.off[c + 1]; i++) col[4 * (uint64_t)(compact(raw.members[i]) - gwn.x0)] |= bit;
                }
            }
        });
    }
    mid_off[d.n_mid_colors] = d.n_mid_members_padded;
    zero_off[d.n_zero_colors] = d.n_zero_members;
    for (int t = 0; t < T; t++) {
        d.max_color_size = std::max(d.max_color_size, t_maxs[t]);
        d.n_updates += t_mid_updates[t];
        d.n_mid_members += t_mid_members[t];
        d.n_gram_updates += t_gram_updates[t];
    }
    d.n_updates += d.n_gram_updates;
    d.n_members = d.n_small_members + d.n_mid_members + d.n_zero_members + d.n_gram_members;
    // ---- per-sample reductions over the threads, in parallel over the samples: updates per row, the cell bound, and the row
    // bands -- cell (i, j) can only be touched for band_lo[i] <= j < band_hi[i]; the accumulator stores exactly these
    // cells, row after row (the full triangle when the colours have no locality)
    std::vector<uint64_t> r_maxsum(T, 0), r_cells(T, 0);
    parallel_chunks(N, T, [&](int tt, uint64_t b, uint64_t e) {
        uint64_t maxsum = 0, cells = 0;
        for (uint64_t i = b; i < e; i++) {
            uint64_t r = 0, ws = 0;
            uint32_t lo = 0xFFFFFFFFu, hi = 0;
            for (int t = 0; t < T; t++)
                if (!t_acc[t].empty()) {
                    const RowAcc& a = t_acc[t][i];
                    r += a.rows; ws += a.wsum;
                    lo = std::min(lo, a.lo); hi = std::max(hi, a.hi);
                }
            if (hi <= lo) lo = hi = 0;
            row_updates[i] = r;
            band_lo[i] = lo;
            band_hi[i] = hi;
            maxsum = std::max(maxsum, ws);
            cells += hi - lo;
        }
        r_maxsum[tt] = maxsum;
        r_cells[tt] = cells;
    });
    d.band_cells = 0;
    for (int t = 0; t < T; t++) { d.max_row_sum = std::max(d.max_row_sum, r_maxsum[t]); d.band_cells += r_cells[t]; }
    bt.lap("  fill");
    {   // the accumulator's row tables travel with the blob: a device handle neither computes nor copies them separately
        uint64_t* bandstart = reinterpret_cast<uint64_t*>(P + d.off_bandstart);
        uint64_t* rowbase = reinterpret_cast<uint64_t*>(P + d.off_rowbase);
        bandstart[0] = 0;
        for (uint64_t i = 0; i < N; i++) {
            bandstart[i + 1] = bandstart[i] + (band_hi[i] - band_lo[i]);
            rowbase[i] = bandstart[i] - band_lo[i];  // wraps for rows whose band starts past their first cells; used modulo 2^64
        }
    }
    ix->totals_file_ids = std::move(raw.total_ids);
    ix->totals_file_vals = std::move(raw.total_vals);
    *out = ix.release();
    return KSB_OK;
}

static int load_files(const char* prefix, int threads, ksb_index** out) {
    const std::string P(prefix);
    LoadTimer lt;
    RawIndex raw;
    MappedFile fs, fc, ft;
    if (!fs.open(P + "_color_to_sources.bin")) return KSB_ERR_IO;
    if (!fc.open(P + "_color_count.bin")) return KSB_ERR_IO;
    if (!ft.open(P + "_groupID_to_kmerCount.bin")) return KSB_ERR_IO;

    // ---- pass 1 over colour->sources: record boundaries (sequential: sizes depend on capacity)
    if (fs.size < 8) { set_error("%s_color_to_sources.bin: shorter than its record count", prefix); return KSB_ERR_FORMAT; }
    uint64_t n_rec;
    std::memcpy(&n_rec, fs.data, 8);
    if (n_rec > fs.size / 24) { set_error("%s_color_to_sources.bin: record count %llu impossible for %llu bytes", prefix, (unsigned long long)n_rec, (unsigned long long)fs.size); return KSB_ERR_FORMAT; }
    struct Rec { uint64_t pos; uint32_t id; uint32_t size; };
    std::vector<Rec> recs(n_rec);
    {
        uint64_t pos = 8;
        uint32_t max_id = 0;
        bool wide = false;
        for (uint64_t r = 0; r < n_rec; r++) {
            if (fs.size - pos < 8) { set_error("%s_color_to_sources.bin: truncated at record %llu", prefix, (unsigned long long)r); return KSB_ERR_FORMAT; }
            uint64_t color;
            std::memcpy(&color, fs.data + pos, 8);
            DumpView v;
            if (!open_dump(fs.data + pos + 8, fs.size - pos - 8, 4, &v)) return KSB_ERR_FORMAT;
            if (v.size > 0xFFFFFFFFull) { set_error("colour with more than 2^32 members"); return KSB_ERR_FORMAT; }
            recs[r] = Rec{pos + 8, (uint32_t)color, (uint32_t)v.size};
            wide |= (color >> 32) != 0;
            max_id = std::max(max_id, (uint32_t)color);
            // the walk is a chain of dependent cache misses (a record's length is in its own header): indexers write runs of
            // colours of one capacity, so the headers a few records ahead are probably where this record's length says
            const uint64_t step = 8 + v.bytes;
            for (uint64_t a = 4; a <= 8; a += 2)
                if (pos + a * step + 24 <= fs.size) __builtin_prefetch(fs.data + pos + a * step, 0, 0);
            pos += step;
        }
        // int_vec_map key is uint32_t: insert_or_assign keeps the last record per narrowed id
        // (the writers iterate a map, so the u64 ids are distinct: two records can only share a NARROWED id when some id
        // does not fit 32 bits -- only then is the table of last records needed)
        IdTable last;
        bool dup = false;
        if (wide) {
            last.init(n_rec, max_id);
            for (uint64_t r = 0; r < n_rec; r++) dup |= last.put(recs[r].id, (uint32_t)std::min<uint64_t>(r, 0xFFFFFFFFull));
        }
        if (dup) {
            if (n_rec > 0xFFFFFFFFull) { set_error("duplicate colour ids with more than 2^32 records are not supported"); return KSB_ERR_FORMAT; }
            std::vector<Rec> kept;
            kept.reserve(n_rec);
            for (uint64_t r = 0; r < n_rec; r++) {
                uint32_t lr = 0;
                last.get(recs[r].id, &lr);
                if (lr == r) kept.push_back(recs[r]);
            }
            recs.swap(kept);
        }
        (void)wide;
    }
    const uint64_t C = recs.size();
    raw.n_colors_file = C;
    lt.lap("record boundaries");

    // ---- colour counts: narrowed key -> narrowed value, slot order, last wins
    IdTable wt;
    {
        DumpView v;
        if (!open_dump(fc.data, fc.size, 16, &v)) return KSB_ERR_FORMAT;
        if (v.size == 0) { set_error("%s_color_count.bin is empty (reference: assert(tmpMap.size()), src/pairwise.cpp:117)", prefix); return KSB_ERR_FORMAT; }
        // two parallel sweeps over the slots: (1) largest narrowed id, live slots, any id above 2^32;
        // (2) the table itself -- distinct u64 keys narrow to distinct u32 keys unless an id is wide, so the direct
        // table takes concurrent puts; otherwise slot order decides which duplicate wins and the sweep stays serial
        const int TC = std::max(1, threads);
        std::vector<uint32_t> tmax(TC, 0);
        std::vector<uint64_t> tlive(TC, 0);
        std::vector<uint8_t> twide(TC, 0);
        parallel_chunks(v.capacity, TC, [&](int t, uint64_t b, uint64_t e) {
            uint32_t mx = 0;
            uint64_t lv = 0;
            bool wd = false;
            for (uint64_t i = b; i < e; i++)
                if (v.ctrl[i] >= 0) { uint64_t k; std::memcpy(&k, v.slots + 16 * i, 8); mx = std::max(mx, (uint32_t)k); wd |= (k >> 32) != 0; lv++; }
            tmax[t] = mx; tlive[t] = lv; twide[t] = wd;
        });
        uint32_t max_id = 0;
        uint64_t live = 0;
        bool wide_ids = false;
        for (int t = 0; t < TC; t++) { max_id = std::max(max_id, tmax[t]); live += tlive[t]; wide_ids |= twide[t] != 0; }
        if (live != v.size) { set_error("%s_color_count.bin: %llu live slots, header says %llu", prefix, (unsigned long long)live, (unsigned long long)v.size); return KSB_ERR_FORMAT; }
        wt.init(v.size, max_id);
        auto sweep = [&](int, uint64_t b, uint64_t e) {
            for (uint64_t i = b; i < e; i++)
                if (v.ctrl[i] >= 0) {
                    uint64_t kv[2];
                    std::memcpy(kv, v.slots + 16 * i, 16);
                    wt.put((uint32_t)kv[0], (uint32_t)kv[1]);
                }
        };
        if (wt.dense() && !wide_ids) parallel_chunks(v.capacity, TC, sweep);
        else sweep(0, 0, v.capacity);
    }

    lt.lap("colour counts");
    // ---- totals, slot order
    {
        DumpView v;
        if (!open_dump(ft.data, ft.size, 8, &v)) return KSB_ERR_FORMAT;
        if (v.size == 0) { set_error("%s_groupID_to_kmerCount.bin is empty (reference: assert, src/pairwise.cpp:170)", prefix); return KSB_ERR_FORMAT; }
        raw.total_ids.reserve(v.size);
        raw.total_vals.reserve(v.size);
        for (uint64_t i = 0; i < v.capacity; i++)
            if (v.ctrl[i] >= 0) {
                uint32_t kv[2];
                std::memcpy(kv, v.slots + 8 * i, 8);
                raw.total_ids.push_back(kv[0]);
                raw.total_vals.push_back(kv[1]);
            }
        if (raw.total_ids.size() != v.size) { set_error("%s_groupID_to_kmerCount.bin: live slots != size", prefix); return KSB_ERR_FORMAT; }
    }

    // ---- pass 2: extract + sort members in parallel
    raw.off.assign(C + 1, 0);
    for (uint64_t c = 0; c < C; c++) raw.off[c + 1] = raw.off[c] + recs[c].size;
    raw.members.resize(raw.off[C]);
    raw.weights.resize(C);
    std::atomic<int> bad{0};
    // chunks of equal FILE BYTES, not of equal colour counts: indexers write colours of very different sizes in runs
    // (500-member family colours after 45-member clade colours), and equal counts left one thread with most of the work
    const int TT = std::max(1, threads);
    std::vector<uint64_t> cstart(TT + 1, C);
    cstart[0] = 0;
    for (int t = 1; t < TT && C; t++) {
        const uint64_t target = 8 + (fs.size - 8) * (uint64_t)t / (uint64_t)TT;
        cstart[t] = (uint64_t)(std::lower_bound(recs.begin(), recs.end(), target, [](const Rec& r, uint64_t pos) { return r.pos < pos; }) - recs.begin());
    }
    parallel_chunks((uint64_t)TT, TT, [&](int, uint64_t tb, uint64_t te) {
      for (uint64_t t = tb; t < te; t++)
        for (uint64_t c = cstart[t]; c < cstart[t + 1]; c++) {
            if (c + 16 < cstart[t + 1]) wt.prefetch(recs[c + 16].id);  // the count lookup is a cache miss per colour
            DumpView v;
            open_dump(fs.data + recs[c].pos, fs.size - recs[c].pos, 4, &v);  // validated in pass 1
            uint32_t* dst = &raw.members[raw.off[c]];
            uint64_t k = 0;
            if (v.size) {
                // eight control bytes at a time: a FULL slot has the top bit of its byte clear
                uint64_t i = 0;
                for (; i + 8 <= v.capacity; i += 8) {
                    uint64_t w8;
                    std::memcpy(&w8, v.ctrl + i, 8);
                    uint64_t full = ~w8 & 0x8080808080808080ull;
                    while (full) {
                        const unsigned b = (unsigned)__builtin_ctzll(full) >> 3;
                        if (k < v.size) std::memcpy(&dst[k], v.slots + 4 * (i + b), 4);
                        k++;
                        full &= full - 1;
                    }
                }
                for (; i < v.capacity; i++)
                    if (v.ctrl[i] >= 0) { if (k < v.size) std::memcpy(&dst[k], v.slots + 4 * i, 4); k++; }
            }
            if (k != v.size) { bad = 1; continue; }
            if (k == 2) { if (dst[0] > dst[1]) std::swap(dst[0], dst[1]); }
            else sort_members(dst, k);
            uint32_t w = 0;
            wt.get(recs[c].id, &w);  // colorsCount[id] default-inserts 0 when absent
            raw.weights[c] = w;
        }
    });
    if (bad) { set_error("%s_color_to_sources.bin: a member set has live slots != size", prefix); return KSB_ERR_FORMAT; }
    lt.lap("members extract + sort");
    const int rc = build_index(raw, threads, out);
    lt.lap("build_index (bin, route, blob)");
    return rc;
}

static int from_csr(uint64_t n_colors, const uint64_t* offsets, const uint32_t* members, const uint32_t* weights,
                    uint64_t n_totals, const uint32_t* total_ids, const uint32_t* total_vals, int threads, ksb_index** out) {
    RawIndex raw;
    raw.n_colors_file = n_colors;
    raw.off.assign(offsets, offsets + n_colors + 1);
    if (n_colors && raw.off[0] != 0) { set_error("offsets[0] must be 0"); return KSB_ERR_ARG; }
    for (uint64_t c = 0; c < n_colors; c++)
        if (raw.off[c + 1] < raw.off[c]) { set_error("offsets not monotone at colour %llu", (unsigned long long)c); return KSB_ERR_ARG; }
    raw.members.assign(members, members + raw.off[n_colors]);
    raw.weights.assign(weights, weights + n_colors);
    // the reference's member lists come from sets: sort, and require distinct members
    std::atomic<int> dup{0};
    parallel_chunks(n_colors, threads, [&](int, uint64_t b, uint64_t e) {
        for (uint64_t c = b; c < e; c++) {
            uint32_t* m = &raw.members[raw.off[c]];
            const uint64_t s = raw.off[c + 1] - raw.off[c];
            std::sort(m, m + s);
            if (std::adjacent_find(m, m + s) != m + s) dup = 1;
        }
    });
    if (dup) { set_error("a colour lists the same sample twice (member lists are sets in the reference)"); return KSB_ERR_ARG; }
    // totals: a map in the reference; last value wins for a repeated id
    raw.total_ids.assign(total_ids, total_ids + n_totals);
    raw.total_vals.assign(total_vals, total_vals + n_totals);
    return build_index(raw, threads, out);
}

int from_csr_owned(uint64_t n_colors, const uint64_t* offsets, const uint32_t* members, const uint32_t* weights,
                   uint64_t n_totals, const uint32_t* total_ids, const uint32_t* total_vals, int threads, ksb_index** out) {
    return from_csr(n_colors, offsets, members, weights, n_totals, total_ids, total_vals, threads, out);
}

// element i of a member section of `bits`-wide compact ids (10 / 12: packed)
static inline uint32_t blob_member(const uint8_t* sec, uint64_t bits, uint64_t i) {
    if (bits == 32) return reinterpret_cast<const uint32_t*>(sec)[i];
    if (bits == 16) return reinterpret_cast<const uint16_t*>(sec)[i];
    if (bits == 10) {
        const uint8_t* q = sec + ((5 * i) >> 2);
        return (((uint32_t)q[0] | ((uint32_t)q[1] << 8)) >> ((i & 3) << 1)) & 0x3FFu;
    }
    const uint8_t* p = sec + ((3 * i) >> 1);
    return (((uint32_t)p[0] | ((uint32_t)p[1] << 8)) >> ((i & 1) << 2)) & 0xFFFu;
}

// Row-block shard: walks every colour of the binned blob, keeps those with a member in rows
// [r0, r1) (as a ROW, i.e. not as the colour's last member), cut down to their members >= r0, and
// bins the result again.  Sample ids / totals are passed through unchanged so the compaction (and
// therefore every cell address of the owned rows) is the one of the whole index.
static int shard_index(const ksb_index* ix, int rank, int world, int threads, ksb_index** out) {
    const IndexDesc& D = ix->desc;
    const uint8_t* P = ix->payload;
    const uint64_t N = D.n_samples;
    std::vector<uint64_t> cut(world + 1, N);
    row_cuts(reinterpret_cast<const uint64_t*>(P + D.off_row_updates), N, D.n_updates, world, cut.data());
    const uint32_t r0 = (uint32_t)cut[rank], r1 = (uint32_t)(rank == world - 1 ? N : cut[rank + 1]);
    const uint32_t* sids = reinterpret_cast<const uint32_t*>(P + D.off_sample_ids);
    const uint32_t* totals = reinterpret_cast<const uint32_t*>(P + D.off_totals);
    // element i of a member section of `bits`-wide compact ids (12 = packed), element i of a weight section
    auto member = [&](const uint8_t* sec, uint64_t bits, uint64_t i) -> uint32_t { return blob_member(sec, bits, i); };
    auto weight = [&](const uint8_t* sec, uint64_t i) -> uint32_t {
        if (D.weight_bits == 8) return sec[i];
        if (D.weight_bits == 16) return reinterpret_cast<const uint16_t*>(sec)[i];
        return reinterpret_cast<const uint32_t*>(sec)[i];
    };
    RawIndex raw;
    raw.n_colors_file = D.n_colors_file;
    raw.off.push_back(0);
    bool corrupt = false;
    // one colour: sorted compact members sec[b .. b+s) of a `bits`-wide section
    auto take = [&](const uint8_t* sec, uint64_t bits, uint64_t b, uint64_t s, uint32_t w, uint32_t add = 0) {
        uint64_t lo = 0, hi = s;  // first member >= r0
        while (lo < hi) { const uint64_t mid = (lo + hi) / 2; if (member(sec, bits, b + mid) + add < r0) lo = mid + 1; else hi = mid; }
        if (s - lo < 2 || member(sec, bits, b + lo) + add >= r1) return;  // no owned row with a later column
        // (a blob adopted through ksb_index_from_blob has had its tables checked, not its member values: a list that is not
        // strictly ascending below N would index past `sids` here and past the Gram bitmaps in build_index)
        uint64_t prev = 0;
        for (uint64_t i = lo; i < s; i++) {
            const uint64_t m = (uint64_t)member(sec, bits, b + i) + add;
            if (m >= N || (i > lo && m <= prev)) { corrupt = true; return; }
            prev = m;
        }
        for (uint64_t i = lo; i < s; i++) raw.members.push_back(sids[member(sec, bits, b + i) + add]);
        raw.off.push_back(raw.members.size());
        raw.weights.push_back(w);
    };
    for (uint32_t sz = 2; sz <= kSmallMax; sz++) {
        const SmallClass& sc = D.cls[sz];
        for (uint64_t c = 0; c < sc.n; c++)
            take(P + D.off_small_members, D.packed_bits, sc.member_off + c * sz, sz, weight(P + D.off_small_weights, sc.weight_off + c));
    }
    {
        const uint64_t* off = reinterpret_cast<const uint64_t*>(P + D.off_mid_offsets);
        const uint32_t* szs = reinterpret_cast<const uint32_t*>(P + D.off_mid_sizes);
        const uint32_t* w = reinterpret_cast<const uint32_t*>(P + D.off_mid_weights);
        for (uint64_t c = 0; c < D.n_mid_colors; c++) take(P + D.off_mid_members, D.member_bits, off[c], szs[c], w[c]);
    }
    {
        const uint64_t* off = reinterpret_cast<const uint64_t*>(P + D.off_zero_offsets);
        for (uint64_t c = 0; c < D.n_zero_colors; c++) take(P + D.off_zero_members, D.member_bits, off[c], off[c + 1] - off[c], 0);
    }
    {
        const uint64_t* off = reinterpret_cast<const uint64_t*>(P + D.off_gram_offsets);
        const GramWin* gw = reinterpret_cast<const GramWin*>(P + D.off_gram_wins);
        if (D.gram_form) {
            // bitmap form: the member lists of a K-block's 128 colours are read back from its rows (samples ascending)
            const uint32_t* bits = reinterpret_cast<const uint32_t*>(P + D.off_gram_members);
            std::vector<uint32_t> lists[kGramKBlock];
            for (uint64_t w = 0; w < D.n_gram_wins; w++) {
                const uint64_t kblocks = ((uint64_t)gw[w].n + kGramKBlock - 1) / kGramKBlock;
                for (uint64_t kb = 0; kb < kblocks; kb++) {
                    const uint64_t kn = std::min<uint64_t>(kGramKBlock, gw[w].n - kb * kGramKBlock);
                    for (uint64_t k = 0; k < kn; k++) lists[k].clear();
                    const uint32_t stride = gram_win_stride(gw[w]);
                    const uint32_t* rows = bits + 4 * (off[w] + kb * stride);
                    for (uint32_t x = 0; x < stride; x++)
                        for (uint32_t wd = 0; wd < 4; wd++)
                            for (uint32_t v = rows[4 * (uint64_t)x + wd]; v; v &= v - 1)
                                lists[32 * wd + (uint32_t)__builtin_ctz(v)].push_back(gw[w].x0 + x);
                    for (uint64_t k = 0; k < kn; k++)
                        take(reinterpret_cast<const uint8_t*>(lists[k].data()), 32, 0, lists[k].size(),
                             weight(P + D.off_gram_weights, gw[w].c0 + kb * kGramKBlock + k));
                }
            }
        }
        uint64_t wi = 0;  // windows partition the Gram colours in ascending c0
        for (uint64_t c = 0; !D.gram_form && c < D.n_gram_colors; c++) {
            while (D.gram_rel_bits && wi + 1 < D.n_gram_wins && c >= gw[wi].c0 + gw[wi].n) wi++;
            take(P + D.off_gram_members, gram_member_bits(D), off[c], off[c + 1] - off[c], weight(P + D.off_gram_weights, c),
                 D.gram_rel_bits ? gw[wi].x0 : 0u);
        }
    }
    if (corrupt) { set_error("ksb_index_shard: the index holds a member list that is not strictly ascending below its sample count"); return KSB_ERR_FORMAT; }
    raw.total_ids.assign(sids, sids + N);
    raw.total_vals.assign(totals, totals + N);
    ksb_index* sh = nullptr;
    const int rc = build_index(raw, threads, &sh);
    if (rc) return rc;
    if (sh->desc.n_samples != N) { ksb_index_free(sh); set_error("shard: sample compaction changed"); return KSB_ERR_ARG; }
    sh->desc.shard_world = (uint64_t)world;
    sh->desc.shard_rank = (uint64_t)rank;
    sh->desc.shard_rb = r0;
    sh->desc.shard_re = r1;
    // pair-updates this shard applies = those of its rows (the cut colours also hold pairs of later rows)
    const uint64_t* ru = reinterpret_cast<const uint64_t*>(sh->payload + sh->desc.off_row_updates);
    uint64_t own = 0;
    for (uint32_t i = r0; i < r1; i++) own += ru[i];
    sh->desc.n_updates = own;
    sh->totals_file_ids = ix->totals_file_ids;
    sh->totals_file_vals = ix->totals_file_vals;
    *out = sh;
    return KSB_OK;
}

}  // namespace ksb

// ------------------------------------------------------------------------------------------- C-ABI

extern "C" {

const char* ksb_last_error(void) { return ksb::last_error(); }
int ksb_abi_version(void) { return KSB_ABI_VERSION; }

int ksb_index_load(const char* index_prefix, int host_threads, ksb_index** out) {
    if (!index_prefix || !out) { ksb::set_error("ksb_index_load: null argument"); return KSB_ERR_ARG; }
    *out = nullptr;
    try {
        return ksb::load_files(index_prefix, host_threads, out);
    } catch (const std::bad_alloc&) {
        ksb::set_error("out of host memory while loading %s", index_prefix);
        return KSB_ERR_NOMEM;
    } catch (const std::exception& e) {
        ksb::set_error("ksb_index_load: %s", e.what());
        return KSB_ERR_FORMAT;
    }
}

int ksb_index_from_csr(uint64_t n_colors, const uint64_t* offsets, const uint32_t* members, const uint32_t* weights,
                       uint64_t n_totals, const uint32_t* total_ids, const uint32_t* total_vals, int host_threads,
                       ksb_index** out) {
    if (!out || !offsets || (!members && n_colors && offsets[n_colors]) || (!weights && n_colors) ||
        (n_totals && (!total_ids || !total_vals))) {
        ksb::set_error("ksb_index_from_csr: null argument");
        return KSB_ERR_ARG;
    }
    *out = nullptr;
    try {
        return ksb::from_csr(n_colors, offsets, members, weights, n_totals, total_ids, total_vals, host_threads, out);
    } catch (const std::bad_alloc&) {
        ksb::set_error("out of host memory");
        return KSB_ERR_NOMEM;
    } catch (const std::exception& e) {
        ksb::set_error("ksb_index_from_csr: %s", e.what());
        return KSB_ERR_ARG;
    }
}

int ksb_index_shard(const ksb_index* ix, int rank, int world, int host_threads, ksb_index** out) {
    if (!ix || !out || world < 1 || rank < 0 || rank >= world) { ksb::set_error("ksb_index_shard: bad argument"); return KSB_ERR_ARG; }
    *out = nullptr;
    if (ix->desc.shard_world) { ksb::set_error("ksb_index_shard: the index is already a shard"); return KSB_ERR_ARG; }
    try {
        return ksb::shard_index(ix, rank, world, host_threads, out);
    } catch (const std::bad_alloc&) {
        ksb::set_error("out of host memory");
        return KSB_ERR_NOMEM;
    } catch (const std::exception& e) {
        ksb::set_error("ksb_index_shard: %s", e.what());
        return KSB_ERR_ARG;
    }
}

namespace ksb {
// A (descriptor, payload) pair that came from another process or rank (NCCL / gloo / a file) is checked before it is
// trusted: the section offsets must be exactly the layout build_index derives from the descriptor's own counts, the
// class table must add up, and the CSR offset arrays and window list inside the payload must stay inside their
// sections.  Then the member VALUES are scanned (validate_members, all host threads): the kernels check a row against the
// owned block but add a column to the row's base as it comes, so a cell address is only as good as the member lists --
// every list strictly ascending below N, every pair of a row inside that row's band, every Gram colour inside its window.
// The member values of a blob whose tables validate_blob has already checked.
static const char* validate_members(const IndexDesc& d, const uint8_t* P) {
    const uint64_t N = d.n_samples;
    const uint32_t* blo = reinterpret_cast<const uint32_t*>(P + d.off_band_lo);
    const uint32_t* bhi = reinterpret_cast<const uint32_t*>(P + d.off_band_hi);
    const uint64_t T = std::min(16u, std::max(1u, std::thread::hardware_concurrency()));
    // (a thread per 2048 colours at most: the 63 size classes of a small index are not worth 63 x 16 thread starts)
    auto par = [&](uint64_t n, const std::function<void(int, uint64_t, uint64_t)>& fn) { parallel_chunks(n, (int)std::min<uint64_t>(T, n / 2048 + 1), fn); };
    std::atomic<int> bad{0};  // 1: a list is not strictly ascending below N, 2: a pair outside its row's band, 3: a Gram colour outside its window
    // one colour of s members, get(i) = compact id: row get(i) pairs with the columns get(i + 1) .. get(s - 1)
    auto check = [&](auto&& get, uint64_t s) -> int {
        if (s == 0) return 0;
        const uint64_t last = get(s - 1);
        if (last >= N) return 1;
        uint64_t prev = get(0);
        if (s > 1 && prev >= last) return 1;
        for (uint64_t i = 1; i < s; i++) {
            const uint64_t m = get(i);
            if (m <= prev || m > last) return 1;           // (so prev < last < N before it indexes the bands)
            if (blo[prev] > m || bhi[prev] <= last) return 2;
            prev = m;
        }
        return 0;
    };
    auto note = [&](int why) { if (why) { int z = 0; bad.compare_exchange_strong(z, why); } };
    for (uint32_t sz = 2; sz <= kSmallMax; sz++) {
        const SmallClass& sc = d.cls[sz];
        const uint8_t* sec = P + d.off_small_members;
        const uint64_t bits = d.packed_bits;
        par(sc.n, [&](int, uint64_t b, uint64_t e) {
            for (uint64_t c = b; c < e && !bad; c++) {
                const uint64_t e0 = sc.member_off + c * sz;
                note(check([&](uint64_t i) { return (uint64_t)blob_member(sec, bits, e0 + i); }, sz));
            }
        });
    }
    {
        const uint64_t* off = reinterpret_cast<const uint64_t*>(P + d.off_mid_offsets);
        const uint32_t* szs = reinterpret_cast<const uint32_t*>(P + d.off_mid_sizes);
        const uint8_t* sec = P + d.off_mid_members;
        par(d.n_mid_colors, [&](int, uint64_t b, uint64_t e) {
            for (uint64_t c = b; c < e && !bad; c++)
                note(check([&](uint64_t i) { return (uint64_t)blob_member(sec, d.member_bits, off[c] + i); }, szs[c]));
        });
    }
    {
        const uint64_t* off = reinterpret_cast<const uint64_t*>(P + d.off_zero_offsets);
        const uint8_t* sec = P + d.off_zero_members;
        par(d.n_zero_colors, [&](int, uint64_t b, uint64_t e) {
            for (uint64_t c = b; c < e && !bad; c++)
                note(check([&](uint64_t i) { return (uint64_t)blob_member(sec, d.member_bits, off[c] + i); }, off[c + 1] - off[c]));
        });
    }
    const GramWin* gw = reinterpret_cast<const GramWin*>(P + d.off_gram_wins);
    const uint64_t* goff = reinterpret_cast<const uint64_t*>(P + d.off_gram_offsets);
    if (!d.gram_form) {
        const uint8_t* sec = P + d.off_gram_members;
        const uint64_t bits = gram_member_bits(d);
        parallel_chunks(d.n_gram_wins, (int)T, [&](int, uint64_t wb, uint64_t we) {
            for (uint64_t w = wb; w < we && !bad; w++) {
                const uint64_t x0 = gw[w].x0, add = d.gram_rel_bits ? x0 : 0, stride = gram_win_stride(gw[w]);
                for (uint64_t c = gw[w].c0; c < gw[w].c0 + gw[w].n && !bad; c++) {
                    const uint64_t sz = goff[c + 1] - goff[c];
                    auto get = [&](uint64_t i) { return (uint64_t)blob_member(sec, bits, goff[c] + i) + add; };
                    note(check(get, sz));
                    if (sz && !bad && (get(0) < x0 || get(sz - 1) - x0 >= stride)) note(3);
                }
            }
        });
        if (!d.gram_rel_bits)  // colours no window names are not run by K4, but the shard builder walks them
            par(d.n_gram_colors, [&](int, uint64_t b, uint64_t e) {
                for (uint64_t c = b; c < e && !bad; c++)
                    note(check([&](uint64_t i) { return (uint64_t)blob_member(sec, bits, goff[c] + i); }, goff[c + 1] - goff[c]));
            });
    } else {
        // bitmap form: a colour's list is read back from the rows of its K-block (ascending by construction); bits of
        // colours past the window's last one would be multiplied with weights that are not theirs
        const uint32_t* bitmap = reinterpret_cast<const uint32_t*>(P + d.off_gram_members);
        std::vector<std::pair<uint32_t, uint32_t>> kbs;
        for (uint64_t w = 0; w < d.n_gram_wins; w++)
            for (uint32_t kb = 0; kb < ((uint64_t)gw[w].n + kGramKBlock - 1) / kGramKBlock; kb++) kbs.emplace_back((uint32_t)w, kb);
        parallel_chunks(kbs.size(), (int)T, [&](int, uint64_t b, uint64_t e) {
            std::vector<uint32_t> lists[kGramKBlock];
            for (uint64_t q = b; q < e && !bad; q++) {
                const GramWin& g = gw[kbs[q].first];
                const uint64_t kb = kbs[q].second, kn = std::min<uint64_t>(kGramKBlock, g.n - kb * kGramKBlock);
                const uint32_t stride = gram_win_stride(g);
                const uint32_t* rows = bitmap + 4 * (goff[kbs[q].first] + kb * stride);
                for (auto& l : lists) l.clear();
                for (uint32_t x = 0; x < stride; x++)
                    for (uint32_t wd = 0; wd < 4; wd++)
                        for (uint32_t v = rows[4 * (uint64_t)x + wd]; v; v &= v - 1) lists[32 * wd + (uint32_t)__builtin_ctz(v)].push_back(g.x0 + x);
                for (uint64_t k = 0; k < kGramKBlock && !bad; k++) {
                    if (k >= kn) { if (!lists[k].empty()) note(3); continue; }
                    const std::vector<uint32_t>& l = lists[k];
                    note(check([&](uint64_t i) { return (uint64_t)l[i]; }, l.size()));
                }
            }
        });
    }
    switch (bad.load()) {
        case 0: return nullptr;
        case 1: return "a member list is not strictly ascending below the sample count";
        case 2: return "a member pair lies outside its row's band";
        default: return "a Gram colour leaves its window";
    }
}

static const char* validate_blob(const IndexDesc& d, const uint8_t* P, uint64_t payload_bytes) {
    if (d.magic != kIndexMagic) return "wrong magic";
    if (d.n_samples > 0x7FFFFFFFull) return "more than 2^31 samples";
    if (d.member_bits != 16 && d.member_bits != 32) return "member_bits";
    if (d.packed_bits != 10 && d.packed_bits != 12 && d.packed_bits != 16 && d.packed_bits != 32) return "packed_bits";
    if (d.weight_bits != 8 && d.weight_bits != 16 && d.weight_bits != 32) return "weight_bits";
    if (d.gram_rel_bits != 0 && d.gram_rel_bits != 10 && d.gram_rel_bits != 12 && d.gram_rel_bits != 16) return "gram_rel_bits";
    const uint64_t lim = 1ull << 40;
    for (uint64_t v : {d.n_small_members, d.n_small_colors, d.n_mid_colors, d.n_mid_members_padded, d.n_mid_items, d.n_zero_colors,
                       d.n_zero_members, d.n_gram_colors, d.n_gram_members, d.n_gram_wins})
        if (v > lim) return "a count is absurd";
    const uint64_t N = d.n_samples, MB = d.member_bits / 8, PB = d.packed_bits, WB = d.weight_bits / 8;
    auto packed_bytes = [&](uint64_t n) { return (PB * n + 7) / 8 + 8; };
    uint64_t p = 0;
    bool ok = true;
    auto section = [&](uint64_t off, uint64_t bytes) { if (off != p) ok = false; p = align256(p + bytes); };
    section(d.off_totals, 4 * N);
    section(d.off_sample_ids, 4 * N);
    section(d.off_row_updates, 8 * N);
    section(d.off_band_lo, 4 * N);
    section(d.off_band_hi, 4 * N);
    section(d.off_bandstart, 8 * (N + 1));
    section(d.off_rowbase, 8 * N);
    section(d.off_small_members, packed_bytes(d.n_small_members));
    section(d.off_small_weights, WB * d.n_small_colors);
    section(d.off_mid_offsets, 8 * (d.n_mid_colors + 1));
    section(d.off_mid_sizes, 4 * d.n_mid_colors);
    section(d.off_mid_members, MB * d.n_mid_members_padded + 64);
    section(d.off_mid_weights, 4 * d.n_mid_colors);
    section(d.off_mid_items, sizeof(MidItem) * d.n_mid_items);
    section(d.off_zero_offsets, 8 * (d.n_zero_colors + 1));
    section(d.off_zero_members, MB * d.n_zero_members);
    if (d.gram_form > 1 || (!d.gram_form && d.gram_bitmap_rows) || d.gram_bitmap_rows > lim) return "gram_form";
    section(d.off_gram_offsets, d.gram_form ? 8 * (d.n_gram_wins + 1) : 8 * (d.n_gram_colors + 1));
    section(d.off_gram_members, d.gram_form ? 16 * (d.gram_bitmap_rows + kGramTile) : (gram_member_bits(d) * d.n_gram_members + 7) / 8 + 8);
    section(d.off_gram_weights, WB * d.n_gram_colors);
    section(d.off_gram_wins, sizeof(GramWin) * d.n_gram_wins);
    if (!ok) return "section offsets do not match the layout of this build";
    if (std::max<uint64_t>(p, 256) != d.payload_bytes || d.payload_bytes != payload_bytes) return "payload size does not match the layout";
    {   // row bands and the tables derived from them
        const uint32_t* blo = reinterpret_cast<const uint32_t*>(P + d.off_band_lo);
        const uint32_t* bhi = reinterpret_cast<const uint32_t*>(P + d.off_band_hi);
        const uint64_t* bs = reinterpret_cast<const uint64_t*>(P + d.off_bandstart);
        const uint64_t* rb = reinterpret_cast<const uint64_t*>(P + d.off_rowbase);
        if (N && bs[0] != 0) return "row band tables";
        for (uint64_t i = 0; i < N; i++)
            if (bhi[i] < blo[i] || bhi[i] > N || bs[i + 1] != bs[i] + (bhi[i] - blo[i]) || rb[i] != bs[i] - blo[i]) return "row band tables";
        if (N && bs[N] != d.band_cells) return "row band tables";
    }
    {   // compact index -> sample id is the ascending order of the ids (build_index's compaction)
        const uint32_t* sid = reinterpret_cast<const uint32_t*>(P + d.off_sample_ids);
        for (uint64_t i = 1; i < N; i++) if (sid[i] <= sid[i - 1]) return "sample ids are not ascending";
    }
    uint64_t nc = 0, nm = 0;
    for (uint32_t s = 2; s <= kSmallMax; s++) {
        const SmallClass& c = d.cls[s];
        if (c.n == 0) continue;
        if (c.s != s || c.n > d.n_small_colors || c.member_off > d.n_small_members || c.member_off + c.n * s > d.n_small_members ||
            c.weight_off + c.n > d.n_small_colors)
            return "a size class leaves its section";
        if (c.pairs != s * (s - 1) / 2 || c.magic != (uint32_t)(((1ull << 32) + c.pairs - 1) / c.pairs)) return "a size class has the wrong pair count";
        nc += c.n;
        nm += c.n * s;
    }
    if (nc != d.n_small_colors || nm != d.n_small_members) return "size classes do not add up";
    auto monotone = [&](uint64_t off, uint64_t n, uint64_t last_max) {
        const uint64_t* a = reinterpret_cast<const uint64_t*>(P + off);
        for (uint64_t i = 0; i < n; i++) if (a[i + 1] < a[i]) return false;
        return a[n] <= last_max;
    };
    if (!monotone(d.off_mid_offsets, d.n_mid_colors, d.n_mid_members_padded)) return "mid offsets";
    if (!monotone(d.off_zero_offsets, d.n_zero_colors, d.n_zero_members)) return "zero-weight offsets";
    if (!d.gram_form && !monotone(d.off_gram_offsets, d.n_gram_colors, d.n_gram_members)) return "Gram offsets";
    {   // a mid colour: its size fits between its offsets, it starts on 8 members (16 bytes: K2 fetches rows with bulk copies)
        const uint64_t* mo = reinterpret_cast<const uint64_t*>(P + d.off_mid_offsets);
        const uint32_t* ms = reinterpret_cast<const uint32_t*>(P + d.off_mid_sizes);
        for (uint64_t c = 0; c < d.n_mid_colors; c++)
            if (ms[c] < 2 || ms[c] > mo[c + 1] - mo[c] || (mo[c] & 7)) return "a mid colour does not fit its offsets";
    }
    const MidItem* mi = reinterpret_cast<const MidItem*>(P + d.off_mid_items);
    {   // a K2 item: a tile of its colour with at least one row (positions; K2 takes min(a0 + kMidRows, s - 1, b0 + kMidCols - 1) - a0 rows)
        const uint32_t* ms = reinterpret_cast<const uint32_t*>(P + d.off_mid_sizes);
        for (uint64_t i = 0; i < d.n_mid_items; i++) {
            if (mi[i].color >= d.n_mid_colors) return "a mid item names a colour that does not exist";
            const uint64_t sz = ms[mi[i].color], b0 = mi[i].b0, a0 = mi[i].a0;
            if (b0 >= sz || b0 % kMidCols || a0 % kMidRows || a0 >= std::min<uint64_t>(sz - 1, b0 + kMidCols - 1)) return "a mid item is not a tile of its colour";
        }
    }
    const GramWin* gw = reinterpret_cast<const GramWin*>(P + d.off_gram_wins);
    for (uint64_t i = 0; i < d.n_gram_wins; i++)
        if (gw[i].c0 > d.n_gram_colors || gw[i].n > d.n_gram_colors - gw[i].c0 || gw[i].width == 0 || gw[i].width % kGramTile || gw[i].width > kGramMaxSamples + kGramTile ||
            gw[i].extent > gw[i].width || (d.gram_rel_bits && gw[i].width > (1ull << d.gram_rel_bits)) ||
            gw[i].c0 != (i ? gw[i - 1].c0 + gw[i - 1].n : 0))  // the windows partition the Gram colours, in order
            return "a Gram window is inconsistent";
    if ((d.n_gram_wins ? gw[d.n_gram_wins - 1].c0 + gw[d.n_gram_wins - 1].n : 0) != d.n_gram_colors) return "a Gram window is inconsistent";
    for (uint64_t i = 0; i < d.n_gram_wins; i++)
        if (gw[i].x0 >= N || (gw[i].extent && (uint64_t)gw[i].x0 + gw[i].extent > N)) return "a Gram window is inconsistent";
    if (d.gram_form) {  // the windows' bitmaps lie back to back, in window order, and colours are numbered in window order
        const uint64_t* wr = reinterpret_cast<const uint64_t*>(P + d.off_gram_offsets);
        uint64_t row = 0, c = 0;
        for (uint64_t i = 0; i < d.n_gram_wins; i++) {
            if (wr[i] != row || gw[i].c0 != c) return "Gram bitmap rows do not match the windows";
            row += gram_win_rows(gw[i]);
            c += gw[i].n;
        }
        if (wr[d.n_gram_wins] != row || row != d.gram_bitmap_rows || c != d.n_gram_colors) return "Gram bitmap rows do not match the windows";
    }
    if (d.shard_world && (d.shard_rank >= d.shard_world || d.shard_rb > d.shard_re || d.shard_re > N)) return "shard fields";
    return validate_members(d, P);
}
}  // namespace ksb

int ksb_index_from_blob(const void* descriptor, uint64_t descriptor_bytes, const void* payload, uint64_t payload_bytes,
                        ksb_index** out) {
    if (!descriptor || !payload || !out) { ksb::set_error("ksb_index_from_blob: null argument"); return KSB_ERR_ARG; }
    *out = nullptr;
    if (descriptor_bytes != sizeof(ksb::IndexDesc)) {
        ksb::set_error("descriptor is %llu bytes, expected %zu", (unsigned long long)descriptor_bytes, sizeof(ksb::IndexDesc));
        return KSB_ERR_ARG;
    }
    ksb::IndexDesc d;
    std::memcpy(&d, descriptor, sizeof d);
    if (d.magic != ksb::kIndexMagic) { ksb::set_error("ksb_index_from_blob: not an index descriptor"); return KSB_ERR_FORMAT; }
    if (d.payload_bytes != payload_bytes) {
        ksb::set_error("payload is %llu bytes, descriptor says %llu", (unsigned long long)payload_bytes, (unsigned long long)d.payload_bytes);
        return KSB_ERR_ARG;
    }
    if (const char* why = ksb::validate_blob(d, static_cast<const uint8_t*>(payload), payload_bytes)) {
        ksb::set_error("ksb_index_from_blob: the blob is not a valid index of this build (%s)", why);
        return KSB_ERR_FORMAT;
    }
    ksb_index* ix = new (std::nothrow) ksb_index();
    if (!ix) { ksb::set_error("out of memory"); return KSB_ERR_NOMEM; }
    ix->desc = d;
    void* mem = nullptr;
    if (cudaMallocHost(&mem, d.payload_bytes) == cudaSuccess) {
        ix->pinned = true;
    } else {
        (void)cudaGetLastError();
        if (posix_memalign(&mem, 4096, d.payload_bytes) != 0) { delete ix; ksb::set_error("out of host memory (%llu bytes)", (unsigned long long)d.payload_bytes); return KSB_ERR_NOMEM; }
    }
    ix->payload = static_cast<uint8_t*>(mem);
    std::memcpy(ix->payload, payload, d.payload_bytes);
    // the totals file order is not part of the blob: slot order (ascending sample id) stands in for it
    const uint32_t* sid = reinterpret_cast<const uint32_t*>(ix->payload + d.off_sample_ids);
    const uint32_t* tot = reinterpret_cast<const uint32_t*>(ix->payload + d.off_totals);
    ix->totals_file_ids.assign(sid, sid + d.n_samples);
    ix->totals_file_vals.assign(tot, tot + d.n_samples);
    *out = ix;
    return KSB_OK;
}

int ksb_index_get_info(const ksb_index* ix, ksb_index_info* info) {
    if (!ix || !info) { ksb::set_error("null argument"); return KSB_ERR_ARG; }
    const ksb::IndexDesc& d = ix->desc;
    std::memset(info, 0, sizeof *info);
    info->n_samples = d.n_samples;
    info->n_colors_file = d.n_colors_file;
    info->n_colors = d.n_colors;
    info->n_members = d.n_members;
    info->n_updates = d.n_updates;
    info->n_small_colors = d.n_small_colors;
    info->n_mid_colors = d.n_mid_colors;
    info->n_gram_colors = d.n_gram_colors;
    info->gram_min_size = d.gram_min_size;
    info->gram_planes = d.gram_planes;
    info->n_gram_updates = d.n_gram_updates;
    info->n_zero_colors = d.n_zero_colors;
    info->max_color_size = d.max_color_size;
    info->payload_bytes = d.payload_bytes;
    info->triangle_cells = d.n_samples ? d.n_samples * (d.n_samples - 1) / 2 : 0;
    info->n_small_members = d.n_small_members;
    info->n_small_updates = d.n_small_updates;
    info->n_gram_members = d.n_gram_members;
    info->n_gram_windows = d.n_gram_wins;
    info->n_gram_macs = d.n_gram_macs;
    info->n_gram_cells = d.n_gram_cells;
    info->pinned = ix->pinned ? 1 : 0;
    info->gram_form = (uint32_t)(d.gram_form | (d.gram_rel_bits << 8));
    return KSB_OK;
}

const uint32_t* ksb_index_sample_ids(const ksb_index* ix) { return ix ? reinterpret_cast<const uint32_t*>(ix->payload + ix->desc.off_sample_ids) : nullptr; }
const uint32_t* ksb_index_totals(const ksb_index* ix) { return ix ? reinterpret_cast<const uint32_t*>(ix->payload + ix->desc.off_totals) : nullptr; }
const void* ksb_index_payload(const ksb_index* ix, uint64_t* bytes) {
    if (!ix) return nullptr;
    if (bytes) *bytes = ix->desc.payload_bytes;
    return ix->payload;
}
const void* ksb_index_descriptor(const ksb_index* ix, uint64_t* bytes) {
    if (!ix) return nullptr;
    if (bytes) *bytes = sizeof(ksb::IndexDesc);
    return &ix->desc;
}

int ksb_index_bands(const ksb_index* ix, const uint32_t** lo, const uint32_t** hi, uint64_t* band_cells) {
    if (!ix) { ksb::set_error("ksb_index_bands: null index"); return KSB_ERR_ARG; }
    if (lo) *lo = reinterpret_cast<const uint32_t*>(ix->payload + ix->desc.off_band_lo);
    if (hi) *hi = reinterpret_cast<const uint32_t*>(ix->payload + ix->desc.off_band_hi);
    if (band_cells) *band_cells = ix->desc.band_cells;
    return KSB_OK;
}

int ksb_index_row_cuts(const ksb_index* ix, int world, uint64_t* cuts_out) {
    if (!ix || !cuts_out || world < 1) { ksb::set_error("ksb_index_row_cuts: bad argument"); return KSB_ERR_ARG; }
    ksb::row_cuts(reinterpret_cast<const uint64_t*>(ix->payload + ix->desc.off_row_updates), ix->desc.n_samples,
                  ix->desc.n_updates, world, cuts_out);
    return KSB_OK;
}

int ksb_blob_row_cuts(const void* descriptor, uint64_t descriptor_bytes, const void* payload, uint64_t payload_bytes,
                      int world, uint64_t* cuts_out) {
    if (!descriptor || !cuts_out || world < 1 || descriptor_bytes != sizeof(ksb::IndexDesc)) { ksb::set_error("ksb_blob_row_cuts: bad argument"); return KSB_ERR_ARG; }
    ksb::IndexDesc d;
    std::memcpy(&d, descriptor, sizeof d);
    if (d.magic != ksb::kIndexMagic || d.payload_bytes != payload_bytes || (!payload && payload_bytes) ||
        d.n_samples > payload_bytes / 8 || d.off_row_updates > payload_bytes - 8 * d.n_samples || (d.off_row_updates & 7)) { /* (no sum that can wrap) */ ksb::set_error("ksb_blob_row_cuts: descriptor does not match the payload"); return KSB_ERR_ARG; }
    ksb::row_cuts(reinterpret_cast<const uint64_t*>(static_cast<const uint8_t*>(payload) + d.off_row_updates), d.n_samples,
                  d.n_updates, world, cuts_out);
    return KSB_OK;
}

int ksb_write_seq_to_kmers(const ksb_index* ix, const char* path) {
    if (!ix || !path) { ksb::set_error("null argument"); return KSB_ERR_ARG; }
    FILE* f = fopen(path, "w");
    if (!f) { ksb::set_error("cannot write %s", path); return KSB_ERR_IO; }
    fputs("ID\tseq\tkmers\n", f);
    for (size_t i = 0; i < ix->totals_file_ids.size(); i++)
        fprintf(f, "%llu\t%u\t%u\n", (unsigned long long)(i + 1), ix->totals_file_ids[i], ix->totals_file_vals[i]);
    if (fclose(f) != 0) { ksb::set_error("write error on %s", path); return KSB_ERR_IO; }
    return KSB_OK;
}

void ksb_index_free(ksb_index* ix) {
    if (!ix) return;
    if (ix->dev_cache && ix->dev_cache_free) ix->dev_cache_free(ix->dev_cache);
    if (ix->payload) {
        if (ix->pinned) cudaFreeHost(ix->payload); else free(ix->payload);
    }
    delete ix;
}

}  // extern "C"
