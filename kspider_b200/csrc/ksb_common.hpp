// Shared host-side declarations of libkspider_b200 (not part of the public C-ABI).
#pragma once
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <functional>
#include <string>
#include <vector>

#include "../../include/kspider_b200.h"

namespace ksb {

void set_error(const char* fmt, ...);  // thread-local message for ksb_last_error()

// Simple blocking parallel-for over [0,n) in contiguous chunks, one chunk per thread.
void parallel_chunks(uint64_t n, int threads, const std::function<void(int, uint64_t, uint64_t)>& fn);

// Row-block ownership: cut[r] = first triangle row of rank r (cut[0] = 0, cut[W] = N), chosen so
// that every rank gets ~U/W weighted pair-updates (row_updates[i] = updates whose smaller sample is i).
void row_cuts(const uint64_t* row_updates, uint64_t n_samples, uint64_t n_updates, int world, uint64_t* cut);

constexpr uint32_t kSmallMax = 64;     // colours with 2..64 members go to the warp-per-colour kernel (K1)
constexpr uint32_t kMaxPairsSmall = kSmallMax * (kSmallMax - 1) / 2;  // 2016
// K2 (s > 64) work decomposition: one CTA item = kMidCols columns (threads) x up to kMidRows rows
constexpr uint32_t kMidCols = 256;
constexpr uint32_t kMidRows = 512;

// K4 (Gram) geometry: 128 x 128 sample tiles, K-blocks of 128 colours
constexpr uint32_t kGramTile = 128;
constexpr uint32_t kGramKBlock = 128;
constexpr uint32_t kGramMaxSamples = 32768;  // widest Gram window (bitmap rows / tile grid are O(width), O(width^2))
constexpr uint32_t kGramMinSize = 16;        // colours smaller than this never go to K4
// routing cost model (calibrated on B200): effective Gram MAC rate (u8 MACs/s over the
// full Npad^2 x K x planes box, i.e. including the half that the symmetric tile skips), red rate of
// the atomic kernels, and the resulting per-colour crossover s / Npad = sqrt(2 * red / mac) (x sqrt(planes))
constexpr double kGramMacRate = 1.0e15;
constexpr double kGramRedRate = 2.5e11;
constexpr double kGramCrossover = 0.0224;  // cfg2 sweep in profiles/r01p_crossover.txt: flat between 20 and 23 of 1024; 23 keeps the e2e copy order better

// One Gram window: the colours [c0, c0 + n) of the Gram CSR have all their members inside the sample
// range [x0, x0 + width) (x0 and width multiples of kGramTile), so their contribution is the Gram
// product of a (n x width) incidence block -- the int8 tensor-core work is O(width^2 n) instead of
// O(N^2 n).  Windows come from a fixed multi-level grid (width 128 << level, stride width / 2; the
// top level is the whole padded sample range when it is at most kGramMaxSamples wide).
struct GramWin {
    uint64_t c0;
    uint32_t n;
    uint32_t x0;
    uint32_t width;
    uint32_t extent;  // samples from x0 to the last one touched (<= width); 0 in blobs written before the field existed
};

// One K2 work item: columns [b0, b0+kMidCols) x rows [a0, a0+kMidRows) of mid colour `color`
// (positions in the colour's sorted member list; only pairs with row position < column position count).
struct MidItem {
    uint32_t color;  // index into the mid-colour arrays
    uint32_t b0;
    uint32_t a0;
    uint32_t pad;
};

// One size class of small colours: all colours with exactly s members, stored as a dense
// [n x s] block of (sorted, compact) member ids and an [n] block of weights.
struct SmallClass {
    uint32_t s;        // members per colour
    uint32_t pairs;    // s(s-1)/2
    uint32_t magic;    // ceil(2^32 / pairs): t / pairs == umulhi(t, magic) for t < 2^16
    uint32_t pad;
    uint64_t n;        // colours in the class
    uint64_t member_off;  // in u32 units, relative to the small-members section
    uint64_t weight_off;  // in u32 units, relative to the small-weights section
};

// Fixed-size, trivially copyable description of the binned CSR blob.  Together with the blob it is
// everything a device (or another rank) needs.
struct IndexDesc {
    uint64_t magic;           // 'KSBIDX01'
    uint64_t n_samples;
    uint64_t n_colors_file;
    uint64_t n_colors;        // multi-member, weight != 0 ... plus zero-weight ones
    uint64_t n_members;
    uint64_t n_updates;
    uint64_t n_small_colors, n_small_members;
    uint64_t n_mid_colors, n_mid_members;
    uint64_t n_zero_colors, n_zero_members;
    uint64_t max_color_size;
    uint64_t max_row_sum;     // max over samples of sum of weights of colours containing it (cell bound)
    SmallClass cls[kSmallMax + 1];  // indexed by s; cls[0], cls[1] unused
    // byte offsets of the sections inside the payload (each 256-byte aligned)
    uint64_t off_totals;        // u32[N]
    uint64_t off_sample_ids;    // u32[N]
    uint64_t off_row_updates;   // u64[N]   updates whose row (= smaller sample) is i
    uint64_t off_band_lo;       // u32[N]   first column row i is ever paired with
    uint64_t off_band_hi;       // u32[N]   one past the last such column (lo = hi = 0: row never touched)
    uint64_t band_cells;        // sum of (hi - lo): cells of the banded accumulator
    uint64_t off_bandstart;     // u64[N+1] first cell of row i in the banded accumulator (prefix of hi - lo)
    uint64_t off_rowbase;       // u64[N]   bandstart[i] - band_lo[i] (modulo 2^64): what the kernels add a column to
    uint64_t off_small_members; // member_bits/8 x [n_small_members]
    uint64_t off_small_weights; // u32[n_small_colors]
    uint64_t off_mid_offsets;   // u64[n_mid_colors+1]  start of each colour in the mid-members section,
                                //                       always a multiple of 4 elements (16 B: TMA bulk copies)
    uint64_t off_mid_sizes;     // u32[n_mid_colors]
    uint64_t off_mid_members;   // u32[n_mid_members_padded]
    uint64_t off_mid_weights;   // u32[n_mid_colors]
    uint64_t off_mid_items;     // MidItem[n_mid_items]
    uint64_t n_mid_items;
    uint64_t n_mid_members_padded;
    uint64_t off_zero_offsets;  // u64[n_zero_colors+1]
    uint64_t off_zero_members;  // u32[n_zero_members]
    // colours routed to the int8 Gram path (K4): plain CSR
    uint64_t n_gram_colors, n_gram_members, n_gram_updates;
    uint64_t gram_planes;       // byte planes needed by the largest Gram weight (1..4)
    uint64_t gram_min_size;     // the routing rule that was applied: s >= gram_min_size (0 = path unused)
    uint64_t off_gram_offsets;  // u64[n_gram_colors+1]
    uint64_t off_gram_members;  // u32[n_gram_members]
    uint64_t off_gram_weights;  // u32[n_gram_colors]
    uint64_t off_gram_wins;     // GramWin[n_gram_wins], ascending (x0, width); colours are grouped by window
    uint64_t n_gram_wins;
    uint64_t n_small_updates;   // pair-updates of the K1 colours
    uint64_t n_gram_macs;       // sum over Gram windows of colours x width(width-1)/2 x planes: the u8 MACs K4 must evaluate
    uint64_t n_gram_cells;      // sum over Gram windows of extent(extent-1)/2: cells K4's epilogue can update (one red each)
    uint64_t member_bits;       // 16 or 32: width of the compact sample ids in the mid / zero member sections
                                // (u16 when N <= 65536: halves the blob, i.e. the H2D / broadcast time)
    uint64_t packed_bits;       // 12, 16 or 32: the same for the small-colour and Gram member sections, which can
                                // also be 12-bit packed (N <= 4096); these two sections are most of the blob
    uint64_t weight_bits;       // 8, 16 or 32: element width of the small-colour and Gram weight sections
    uint64_t gram_rel_bits;     // 0: the Gram member section holds compact sample ids, packed_bits wide.  10, 12 or 16: it
                                // holds OFFSETS INSIDE THE COLOUR'S WINDOW (member - GramWin::x0) of that width -- a window
                                // is at most 2^bits samples wide whatever N is, so an index of > 65536 samples ships 2-byte
                                // Gram members instead of 4-byte ones (cfg4: the blob, i.e. the H2D time, 307 -> 163 MB)
    uint64_t gram_form;         // 0: the Gram colours are a CSR (off_gram_offsets = u64[n_gram_colors + 1], off_gram_members =
                                // member lists) and kb_build turns them into bitmaps every step.  1: the blob carries the
                                // BITMAPS K4 reads -- per window, K-block-major, one 16-byte row (= 128 colours) per sample:
                                // bits[(win_row[w] + kblock * width + sample - x0) * 4 + word], bit (colour % 32) -- in
                                // off_gram_members ((gram_bitmap_rows + kGramTile) x 16 B) and off_gram_offsets holds
                                // win_row = u64[n_gram_wins + 1].  Taken when that is SMALLER than the CSR (dense colours:
                                // a clade colour of 45 members in a 128-sample window is 16 B instead of 90): cfg4's
                                // Gram section 147 -> 27 MB, and the step needs no bitmap build.
    uint64_t gram_bitmap_rows;  // 16-byte rows of all windows' bitmaps (form 1)
    // row-block shard (ksb_index_shard): this blob holds only what the owner of triangle rows
    // [shard_rb, shard_re) needs -- colours with a member in that range, cut down to their members
    // >= shard_rb.  shard_world == 0: the whole index.
    uint64_t shard_world, shard_rank, shard_rb, shard_re;
    uint64_t payload_bytes;
};
constexpr uint64_t kIndexMagic = 0x37305844494253ULL | (0x4BULL << 56);
// Bitmap form of the blob: a K-block of a window holds one 16-byte row per sample from x0 up to the LAST one a colour of the
// window touches (`extent`), not up to the tile-padded width -- a 50-sample clade in its 128-sample window is 50 rows per
// K-block, not 128 (cfg4: 27 -> 16 MB of bitmaps).  K4 still fetches whole 128-row tiles: what it reads past a K-block's rows
// are the next K-block's, and every cell of a row or column >= extent is masked in the epilogue.
inline uint32_t gram_win_stride(const GramWin& w) { return w.extent ? w.extent : w.width; }
inline uint64_t gram_win_rows(const GramWin& w) { return (((uint64_t)w.n + kGramKBlock - 1) / kGramKBlock) * gram_win_stride(w); }
inline uint64_t gram_member_bits(const IndexDesc& d) { return d.gram_rel_bits ? d.gram_rel_bits : d.packed_bits; }

}  // namespace ksb

namespace ksb {
// creates the CUDA context of `device` (device_api.cu): the drop-in calls it on a thread of its own while the index files are parsed
int warm_device(int device);
}  // namespace ksb

struct ksb_index {
    ksb::IndexDesc desc;
    uint8_t* payload = nullptr;
    bool pinned = false;
    // ksb_pairwise_cluster keeps the device handle of its last call on this index (allocations, row passes, K4 plan):
    // the next call with the same options only uploads the blob again and runs (device_api.cu; released with the index)
    mutable void* dev_cache = nullptr;
    mutable void (*dev_cache_free)(void*) = nullptr;
    // totals file in slot order, for the seqToKmersNo.tsv writer (src/pairwise.cpp:173-180)
    std::vector<uint32_t> totals_file_ids, totals_file_vals;
};
