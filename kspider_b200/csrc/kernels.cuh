// sm_100a kernels of the pairwise hot path.  Everything here is device code + thin launch helpers;
// the C-ABI lives in device_api.cu.
//
// Reference semantics reproduced (dib-lab/kSpider @ 466a267):
//   K1/K2/Kz  src/pairwise.cpp:209-225  every member pair (a<b) of a colour: edges[(min,max)] += (uint32_t)count
//             (Combo::combinations :39-70 enumerates the pairs, ascending() :73-78 orders the ids)
//   K5        src/pairwise.cpp:253-264  float32 containments of every touched pair
//   K6/K7     pykSpider/kSpider2/ks_clustering.py:97-120  threshold + connected components
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "ksb_common.hpp"

namespace ksb {

// Geometry of the owned row block of the banded upper triangle.  Row i stores the cells
// band_lo[i] <= j < band_hi[i] (the only columns row i is ever paired with; the loader computes
// the bands, see index_loader.cpp), rows back to back:
//   cell(i,j) -> bandstart[i] + (j - band_lo[i]) - base = rowbase[i] + j - base   (rows [rb, rb+nrows) are owned)
// Without locality in the colours the bands are [i+1, N) and this is the packed upper triangle.
struct TriGeom {
    unsigned long long* tri;              // owned cells (u64)
    const unsigned long long* rowbase;    // [N]   bandstart[i] - band_lo[i] (wrapping): index of the virtual cell (i, 0)
    const unsigned long long* bandstart;  // [N+1] first cell of row i
    const uint32_t* band_lo;              // [N]
    const uint32_t* band_hi;              // [N]
    unsigned long long base;              // bandstart[rb]
    unsigned long long cells;             // owned cells
    uint32_t n;                           // N
    uint32_t rb;                          // first owned row
    uint32_t nrows;                       // owned rows
    uint32_t cell_bits;                   // 64, or 32: the cells are u32 (tri then points at uint32_t cells) -- taken when the
                                          // loader's bound on any cell (IndexDesc::max_row_sum) fits: every HBM-bound part of
                                          // the step (clear, K4's reds, the post-pass scan) moves half the bytes
};

// Element accessors of the blob's member / weight sections.  Members of the small-colour and Gram
// sections are u32, u16 (N <= 65536), 12-bit packed (N <= 4096: element i = bits [12i, 12i+12) of
// the byte stream) or 10-bit packed (N <= 1024: bits [10i, 10i+10), 4 elements per 5 bytes, so an
// element always lies in 2 bytes); their weights are u8 / u16 / u32, whichever holds the largest.  Narrow sections
// only shorten the H2D / NVLink copy of the end-to-end path; the kernels widen on load.
struct P12 { uint8_t b; };
struct P10 { uint8_t b; };
#ifdef __CUDACC__
template <class MT>
__device__ __forceinline__ uint32_t ldm(const MT* base, unsigned long long i) { return __ldg(base + i); }
template <>
__device__ __forceinline__ uint32_t ldm<P12>(const P12* base, unsigned long long i) {
    const uint8_t* p = reinterpret_cast<const uint8_t*>(base) + ((i * 3ull) >> 1);
    const uint32_t v = (uint32_t)__ldg(p) | ((uint32_t)__ldg(p + 1) << 8);
    return (v >> ((uint32_t)(i & 1ull) << 2)) & 0xFFFu;
}
template <>
__device__ __forceinline__ uint32_t ldm<P10>(const P10* base, unsigned long long i) {
    const uint8_t* p = reinterpret_cast<const uint8_t*>(base) + ((i * 5ull) >> 2);
    const uint32_t v = (uint32_t)__ldg(p) | ((uint32_t)__ldg(p + 1) << 8);
    return (v >> ((uint32_t)(i & 3ull) << 1)) & 0x3FFu;
}
__device__ __forceinline__ uint32_t ldw(const void* base, uint32_t bits, unsigned long long i) {
    if (bits == 8) return __ldg(static_cast<const uint8_t*>(base) + i);
    if (bits == 16) return __ldg(static_cast<const uint16_t*>(base) + i);
    return __ldg(static_cast<const uint32_t*>(base) + i);
}
#endif

struct K1Class {
    unsigned long long task_begin;  // first warp task of this class
    unsigned long long n;           // colours in the class
    unsigned long long member_off;  // u32 units into small-members
    unsigned long long weight_off;  // u32 units into small-weights
    uint32_t s;                     // members per colour
    uint32_t k;                     // colours per warp task (32/s for s<=32, else 1)
};
struct K1Params {
    K1Class cls[kSmallMax];
    unsigned long long n_tasks;
    uint32_t n_cls;
    uint32_t weight_bits;           // 8, 16 or 32: element width of the small-weights section
};

struct LaunchCfg {
    int sms;
    cudaStream_t stream;
    int k1_ctas = 8;  // K1 CTAs (256 threads) per SM; fewer when K4 runs beside it and needs its slot
    // K4 (one CTA per SM, ~200 KB of shared memory) runs beside this kernel: ask for the maximum shared-memory
    // carve-out although the kernel itself uses none.  An SM switches its L1 / shared split only when it is
    // idle, so a K1 grid that got there first with the default (all-L1) split keeps every K4 CTA waiting
    // until K1 has drained -- the two kernels then run back to back instead of side by side.
    bool beside_gram = false;
};

// launchers (return cudaError_t of the launch)
// warp tasks [task_begin, task_end) of the K1 task list
// member sections hold u16 or u32 compact sample ids (IndexDesc::member_bits)
cudaError_t launch_k1_small(const K1Params& p, const void* members, uint32_t member_bits, const void* weights, const TriGeom& g,
                            const LaunchCfg& lc, unsigned long long task_begin, unsigned long long task_end);
cudaError_t launch_k2_mid(const MidItem* items, uint64_t n_items, const unsigned long long* mid_off,
                          const uint32_t* mid_sizes, const void* mid_members, uint32_t member_bits, const uint32_t* mid_weights,
                          const TriGeom& g, const LaunchCfg& lc);
cudaError_t launch_kz_touch(uint64_t n_zero, const unsigned long long* zero_off, const void* zero_members,
                            uint32_t member_bits, uint32_t* touched_bits, const TriGeom& g, const LaunchCfg& lc);

// K3 (spill): chunk buffer the emit kernels append (key = row << shift | column, u32 weight) to
struct SpillBuf {
    unsigned long long* keys;
    uint32_t* vals;
    unsigned long long* cursor;  // entries appended so far
    uint32_t shift;              // bits of a column index
};
cudaError_t launch_k1_spill(const K1Params& p, const void* members, uint32_t member_bits, const void* weights, const TriGeom& g,
                            const SpillBuf& sb, unsigned long long task_begin, unsigned long long task_end,
                            const LaunchCfg& lc);
cudaError_t launch_k2_spill(const MidItem* items, uint64_t item_begin, uint64_t item_end, const unsigned long long* mid_off,
                            const uint32_t* mid_sizes, const void* mid_members, uint32_t member_bits, const uint32_t* mid_weights,
                            const TriGeom& g, const SpillBuf& sb, const LaunchCfg& lc);
cudaError_t launch_k3_emit_csr(uint64_t c_begin, uint64_t c_end, const unsigned long long* off, const void* members,
                               uint32_t member_bits, const void* weights, uint32_t weight_bits, const TriGeom& g, const SpillBuf& sb, const LaunchCfg& lc,
                               const GramWin* wins = nullptr, uint32_t n_wins = 0);  // wins: members are offsets inside the colour's Gram window

// the same for Gram colours that the blob holds as bitmaps (IndexDesc::gram_form 1): colours [c_begin, c_end)
cudaError_t launch_k3_emit_bitmap(uint64_t c_begin, uint64_t c_end, const GramWin* wins, uint32_t n_wins, const unsigned long long* win_rows,
                                  const uint32_t* bits, const void* weights, uint32_t weight_bits, const TriGeom& g, const SpillBuf& sb,
                                  const LaunchCfg& lc);

cudaError_t launch_k5_sparse(const unsigned long long* keys, const unsigned long long* sums, uint64_t n, uint32_t shift,
                             const uint32_t* totals, const uint32_t* sample_ids, ksb_record* records,
                             unsigned long long* pairs, float* cmax, const LaunchCfg& lc);
cudaError_t spill_temp_bytes(uint64_t n, int end_bit, size_t* bytes);
cudaError_t spill_sort_reduce_chunk(void* tmp, size_t tmp_bytes, const unsigned long long* keys_in,
                                    unsigned long long* keys_sorted, const uint32_t* vals_in, uint32_t* vals_sorted,
                                    uint64_t n, int end_bit, unsigned long long* uniq, unsigned long long* sums,
                                    unsigned long long* d_runs, cudaStream_t st);
cudaError_t spill_sort_reduce_runs(void* tmp, size_t tmp_bytes, const unsigned long long* keys_in,
                                   unsigned long long* keys_sorted, const unsigned long long* sums_in,
                                   unsigned long long* sums_sorted, uint64_t n, int end_bit, unsigned long long* uniq,
                                   unsigned long long* sums, unsigned long long* d_runs, cudaStream_t st);

// merge of two sorted runs of unique keys by merge path (duplicates side by side), and the reduce that adds them up
cudaError_t spill_merge_path(const unsigned long long* ak, const unsigned long long* as, uint64_t na, const unsigned long long* bk,
                             const unsigned long long* bs, uint64_t nb, unsigned long long* part, unsigned long long* out_k,
                             unsigned long long* out_s, cudaStream_t st);
uint64_t spill_merge_part_entries(uint64_t total);
cudaError_t cub_reduce_bytes(uint64_t n, size_t* bytes);
cudaError_t spill_reduce_sorted(void* tmp, size_t tmp_bytes, const unsigned long long* keys, const unsigned long long* sums_in, uint64_t n,
                                unsigned long long* uniq, unsigned long long* sums, unsigned long long* d_runs, cudaStream_t st);

// K5: two passes over the owned block.  block_counts has post_blocks(cells)+1 entries.
constexpr uint32_t kPostThreads = 256;
constexpr uint32_t kPostPerThread = 8;
constexpr uint32_t kPostChunk = kPostThreads * kPostPerThread;
inline uint64_t post_blocks(uint64_t cells) { return (cells + kPostChunk - 1) / kPostChunk; }
cudaError_t launch_k5_count(const TriGeom& g, const uint32_t* touched_bits, unsigned long long* block_counts,
                            const LaunchCfg& lc);
// block_rows[b] = row of the first cell of K5 block b (post_blocks(cells) entries): static per pass
cudaError_t launch_k5_block_rows(const TriGeom& g, uint32_t* block_rows, const LaunchCfg& lc);
cudaError_t launch_k5_emit(const TriGeom& g, const uint32_t* touched_bits, const unsigned long long* block_offsets,
                           const uint32_t* block_rows, const uint32_t* totals, const uint32_t* sample_ids, ksb_record* records,
                           unsigned long long* pairs, float* cmax, const LaunchCfg& lc);

// K56: cluster-only scan (containment + threshold + append of the kept pairs only) and the edge-list unions
cudaError_t launch_k56_scan(const TriGeom& g, const uint32_t* touched_bits, const uint32_t* block_rows, const uint32_t* totals,
                            int dist_col, float cstar, unsigned long long* edges, unsigned long long cap,
                            unsigned long long* counters, uint32_t* parent, const LaunchCfg& lc);
// true: the pass is scanned row by row (k56_rows: needs no block_rows table, and hooks the kept edges into `parent` itself)
bool k56_uses_rows(const TriGeom& g);
cudaError_t launch_k6_edges(const unsigned long long* edges, unsigned long long cap, const unsigned long long* counters,
                            uint32_t* parent, uint64_t expect, bool hooked, const LaunchCfg& lc);

// K6+K7
cudaError_t launch_k67_union(const ksb_record* records, const unsigned long long* pairs, const float* colvals, uint64_t n_rows,
                             int dist_col, float cstar, uint32_t* parent, unsigned long long* n_edges, const LaunchCfg& lc);
cudaError_t launch_k6_hook_min(const ksb_record* records, const unsigned long long* pairs, const float* colvals, uint64_t n_rows,
                               int dist_col, float cstar, uint32_t* parent, const LaunchCfg& lc);
cudaError_t launch_k7_init(uint32_t* parent, uint32_t n, const LaunchCfg& lc);
cudaError_t launch_k7_compress(uint32_t* parent, uint32_t n, const LaunchCfg& lc);

// K8: ANI column of the records (ks_pairwise.py:29-82); n_bad counts rows whose distance left [0, 1]
cudaError_t launch_k8_ani(const ksb_record* records, uint64_t n_rows, double inv_k, double* out, unsigned long long* n_bad,
                          const LaunchCfg& lc);

// K4: int8 tcgen05 Gram path (gram.cu).  One launch handles a CHUNK: a list of (sub-)windows whose
// bitmaps share one buffer, and a list of work items over them.
struct GramWinDev {
    unsigned long long c0;        // first Gram colour (index into the Gram CSR) of this (sub-)window
    unsigned long long bits_off;  // first 16-byte bitmap row of its K-block 0 inside the chunk buffer
    uint32_t n;                   // colours
    uint32_t x0;                  // first sample of the window
    uint32_t width;               // samples (multiple of kGramTile)
    uint32_t tiles;               // width / kGramTile
    uint32_t k0;                  // first K-block of the window in the chunk's weight planes
    uint32_t extent;              // samples from x0 to the last one a colour of the window touches (<= width; 0 = unknown):
                                  // rows / columns beyond it are padding of the last tile and hold no update
    uint32_t kstride;             // 16-byte rows from one K-block of the window to the next: width in a chunk buffer kb_build
                                  // fills, the extent in the blob's own bitmaps (gram_win_stride) -- rows a tile reads past a
                                  // K-block's own are then FOREIGN bits, and the epilogue masks every row / column >= extent
    uint32_t pad;
};
constexpr uint32_t kGramSeg = 2048;  // samples per bitmap-build unit: 32 KB of shared memory
struct GramUnit {                    // one bitmap-build unit of a chunk (plan_gram lists them, heaviest first): a K-block
                                     // of <= 128 colours of a window x a segment of <= kGramSeg samples
    unsigned long long c0k;          // first colour of the K-block in the Gram CSR
    unsigned long long bits_row;     // first 16-byte row of the unit in the chunk's bitmap buffer
    uint16_t kn;                     // colours in the K-block
    uint8_t w0, nw;                  // the unit builds the 32-colour words [w0, w0 + nw) of the bitmap rows (nw = 4: whole rows);
                                     // K-blocks of wide windows (long colours) are split into four one-word units
    uint32_t xbase;                  // first sample of the segment
    uint32_t slen;                   // samples in the segment
    uint32_t wq_off;                 // offset of the K-block in a weight byte plane, 0xFFFFFFFF when another unit writes it
};
struct GramItem {
    uint32_t win;                 // index into the chunk's window list
    uint16_t i0, j0;              // first row of the 128-row tile and first column of its (up to NB) 128-column
                                  // tiles, as sample offsets inside the window: j0 = i0 + 128 k, i0 starts at the
                                  // first OWNED row, so a row block never pays for a tile it owns a sliver of
    uint32_t kb0, kb1;            // K-block range inside the window
};
struct GramParams {
    const uint32_t* bits;         // chunk bitmap: per window [kblocks][width][4] u32, bit (colour % 128) of local sample
    const uint8_t* wq;            // [planes][kpad] weight byte planes
    const GramWinDev* wins;
    const GramItem* items;
    uint32_t n_wins;
    uint32_t n_items;
    uint32_t kpad;                // row pitch of wq (multiple of 128)
    uint32_t planes;              // 1, 2 or 4
    uint32_t* counter;            // work counter, zeroed before the launch
};
uint32_t gram_cols_per_item(int planes);
// the variant for chunks of SHORT items (few K-blocks per tile pair): static schedule, double-buffered TMEM
// accumulators, expansion of item n + 1 before the epilogue of item n; planes 1 or 2, NB = 2 / planes
uint32_t gram_cols_per_item_pipe(int planes);
cudaError_t launch_k4_gram_pipe(const GramParams& prm, const TriGeom& g, const LaunchCfg& lc);
cudaError_t launch_kb_build(const unsigned long long* gram_off, const void* gram_members, uint32_t member_bits,
                            const void* gram_weights, uint32_t weight_bits, const GramUnit* units, uint32_t n_units,
                            uint32_t kpad, uint32_t planes, uint32_t* bits, uint8_t* wq, const LaunchCfg& lc);
cudaError_t launch_kb_weights(const GramWinDev* wins, uint32_t n_wins, uint32_t kblocks, const void* gram_weights, uint32_t weight_bits,
                              uint32_t kpad, uint32_t planes, uint8_t* wq, const LaunchCfg& lc);
cudaError_t launch_k4_gram(const GramParams& prm, const TriGeom& g, const LaunchCfg& lc);

// device handle shared by the translation units of the C-ABI (device_api.cu): validates / selects the ordinal
// (-1 = current), returns it or a negative KSB_ERR_*; the per-device service stream results and tables use
int check_device(int device, int* sms);
cudaStream_t service_stream(int dev);

#ifdef __CUDACC__
// lock-free union-find shared by K6/K7 (kernels.cu) and the table kernels (table.cu): smaller root wins
__device__ __forceinline__ uint32_t uf_find(uint32_t* parent, uint32_t x) {
    volatile uint32_t* p = parent;
    for (;;) {
        const uint32_t px = p[x];
        if (px == x) return x;
        const uint32_t ppx = p[px];
        if (ppx != px) p[x] = ppx;  // path halving
        x = px;
    }
}

__device__ __forceinline__ void uf_union(uint32_t* parent, uint32_t a, uint32_t b) {
    for (;;) {
        a = uf_find(parent, a);
        b = uf_find(parent, b);
        if (a == b) return;
        if (a < b) { const uint32_t t = a; a = b; b = t; }
        if (atomicCAS(parent + a, a, b) == a) return;
    }
}
#endif

// tensor-pipe ceiling in K4's instruction shape (gram.cu); *macs = u8 multiply-adds of one launch
cudaError_t launch_imma_peak(uint32_t iters, uint32_t* sink, double* macs, const LaunchCfg& lc);

// red micro-benchmark
cudaError_t launch_bench_red(void* buf, uint64_t cells, uint64_t n_ops, int cell_bits, int pattern, const LaunchCfg& lc);

}  // namespace ksb
