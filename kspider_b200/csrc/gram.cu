// K4: exact int8 Gram path for dense colours (tcgen05.mma kind::i8, accumulators in TMEM).
//
// Replaces, for the colours the loader routes here, the same reference lines as K1/K2
// (src/pairwise.cpp:209-225: edges[(i,j)] += (uint32_t)count for every member pair of a colour):
//     S[i][j] = sum_c w_c * B[c][i] * B[c][j]        B = 0/1 incidence (colour x sample)
// computed as D_p = (diag(byte_p(w)) B)^T-tile x B-tile for every byte plane p of the weights
// (u8 x u8 -> s32, 255 * K < 2^31 per accumulation run), combined as sum_p D_p << 8p in u64 and
// added into the same packed upper triangle with red.global.add.u64 (cells i < j only).
//
// Operands never exist in global memory as bytes: a bitmap (one bit per (colour, sample), 16 bytes
// = 128 colours per sample per K-block) is built by kb_build; k4_gram expands it into the
// 128B-swizzled K-major shared-memory tiles the tensor core reads.
#include "kernels.cuh"

namespace ksb {

namespace {

constexpr uint32_t kTile = kGramTile;        // 128 samples per tile side (UMMA M = N = 128)
constexpr uint32_t kKB = kGramKBlock;        // 128 colours per K-block = one 128-byte swizzle atom of u8
constexpr uint32_t kStages = 2;
constexpr uint32_t kProducerWarps = 16;
constexpr uint32_t kProducers = kProducerWarps * 32;   // expansion + epilogue threads
constexpr uint32_t kThreads = kProducers + 64;          // + the MMA-issuing warp + the TMA loader warp
constexpr uint32_t kRing = 4;                           // ring of TMA-loaded bitmap rows + weights, one slot per K-block
constexpr uint32_t kTileBytes = kTile * kKB;  // 16 KB: one operand tile (128 rows x 128 B)
constexpr uint32_t kTmemCols = 512;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok = 0;
    while (!ok) {
        asm volatile(
            "{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
            : "=r"(ok)
            : "r"(bar), "r"(parity)
            : "memory");
    }
}

// Shared-memory matrix descriptor, K-major, SWIZZLE_128B (cute::UMMA::SmemDescriptor):
//   [0,14) start address >> 4 | [16,30) LBO >> 4 (= 1, unused for swizzled K-major) |
//   [32,46) SBO >> 4 (= 1024 B between 8-row groups) | [46,48) version = 1 | [61,64) layout = 2
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
    return (uint64_t)((smem_addr & 0x3FFFFu) >> 4) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}

// Instruction descriptor (cute::UMMA::InstrDescriptor) for kind::i8:
//   c_format [4,6) = 2 (S32), a_format [7,10) = 0 (u8), b_format [10,13) = 0 (u8), a/b major = K (0),
//   n_dim [17,23) = N >> 3, m_dim [24,29) = M >> 4
__device__ __forceinline__ constexpr uint32_t umma_idesc(uint32_t n) { return (2u << 4) | ((n >> 3) << 17) | ((kTile >> 4) << 24); }

__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_1d(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
          "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
          "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// explicit shared-space accesses with 32-bit addresses: the generic pointers derived from the aligned
// dynamic-smem base made the compiler emit LD.E / ST.E with 64-bit address arithmetic in the hot loops
__device__ __forceinline__ void sts128(uint32_t a, uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(x), "r"(y), "r"(z), "r"(w));
}
__device__ __forceinline__ uint4 lds128(uint32_t a) {
    uint4 v;
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
    return v;
}
__device__ __forceinline__ uint2 lds64(uint32_t a) {
    uint2 v;
    asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a));
    return v;
}
__device__ __forceinline__ void sts32(uint32_t a, uint32_t x) { asm volatile("st.shared.b32 [%0], %1;" ::"r"(a), "r"(x) : "memory"); }
__device__ __forceinline__ uint32_t lds32(uint32_t a) {
    uint32_t v;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    return v;
}

// Order of the 128 colours of a K-block along the K axis of the expanded operand tiles.  Colour k is bit (k & 31) of
// word (k >> 5) of a bitmap row; the expansion turns a word into eight 4-byte groups with ONE shift and ONE mask each,
//     group q = (word >> q) & 0x01010101      byte t of group q  <=>  bit q + 8 t,
// so colour k sits at K position (k & ~31) + 4 * (k & 7) + ((k & 31) >> 3).  A Gram product does not care about the
// order of K as long as both operands agree: the weight byte planes (kb_build) are stored in the same order.
__device__ __forceinline__ uint32_t kperm(uint32_t k) { return (k & ~31u) + 4u * (k & 7u) + ((k & 31u) >> 3); }

// address of the virtual cell (x, 0) of the owned block (0: row not owned), u64 or u32 cells (TriGeom::cell_bits)
template <bool C32>
__device__ __forceinline__ unsigned long long gram_row_addr(uint32_t x, const TriGeom& g) {
    if (x - g.rb >= g.nrows) return 0ull;
    return (unsigned long long)g.tri + (C32 ? 4ull : 8ull) * (__ldg(g.rowbase + x) - g.base);
}
template <bool C32>
__device__ __forceinline__ void gram_red(unsigned long long row_addr, uint32_t j, unsigned long long v) {
    if (C32) asm volatile("red.global.add.u32 [%0], %1;" ::"l"(row_addr + 4ull * j), "r"((uint32_t)v) : "memory");
    else asm volatile("red.global.add.u64 [%0], %1;" ::"l"(row_addr + 8ull * j), "l"(v) : "memory");
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// kb_build: one CTA per build unit = (sub-window, K-block of 128 colours, segment of kGramSeg
// samples).  The unit's piece of the K-block-major bitmap
//     bits[(bits_off + kblock * width + sample - x0) * 4 + word]      bit (colour % 128) of the sample
// is assembled in shared memory (shared atomicOr over the K-block's contiguous member array) and written out once,
// coalesced -- zeros included, so the bitmap needs no memset and no global atomics.  Segment 0 also
// writes the byte planes of the weights to wq[plane * kpad + (k0 + kblock) * 128 + kperm(k)].
// ------------------------------------------------------------------------------------------------
template <class MT>
__global__ void __launch_bounds__(256) kb_build(const unsigned long long* __restrict__ gram_off,
                                                const MT* __restrict__ gram_members,
                                                const void* __restrict__ gram_weights, uint32_t weight_bits,
                                                const GramUnit* __restrict__ units, uint32_t n_units,
                                                uint32_t kpad, uint32_t planes, uint32_t* __restrict__ bits,
                                                uint8_t* __restrict__ wq) {
    __shared__ __align__(16) uint32_t sm[kGramSeg * 4];
    __shared__ unsigned long long sm_off[kKB + 1];
    __shared__ uint32_t sm_rel[kKB + 1];
    const uint32_t tid = threadIdx.x;
    if (blockIdx.x >= n_units) return;
    GramUnit nxt = units[blockIdx.x];
    for (uint32_t unit = blockIdx.x; unit < n_units; unit += gridDim.x) {
        // the plan's table says everything about the unit (heaviest units first); the NEXT unit's entry is loaded now
        const GramUnit gu = nxt;
        if (unit + gridDim.x < n_units) nxt = units[unit + gridDim.x];
        const uint32_t slen = gu.slen, kn = gu.kn;
        const uint32_t k0 = min(kn, 32u * gu.w0), k1 = min(kn, 32u * (gu.w0 + gu.nw));  // colours of this unit's words
        // shared layout is WORD-major -- sm[word][sample], word = 32 colours -- so that the atomics of a warp, whose
        // lanes hold consecutive (ascending) members of a colour, fall into different banks; the 16-byte rows of the
        // bitmap are put together when the unit is written out
        for (uint32_t wd = gu.w0; wd < (uint32_t)gu.w0 + gu.nw; wd++)
            for (uint32_t r = tid; r < slen; r += 256) sm[wd * kGramSeg + r] = 0u;
        if (tid <= kn) sm_off[tid] = gram_off[gu.c0k + tid];
        __syncthreads();
        // The colours of a K-block are consecutive in the Gram CSR, so their member lists are ONE contiguous array.
        // Every warp takes a contiguous eighth of it, 32 elements per load instruction (coalesced), sixteen
        // independent loads in flight per lane before the first shared atomic; the colour of a lane's element is
        // followed through the staged offsets (relative, 32-bit) -- consecutive elements of a lane are 32 apart, so
        // the colour index advances by at most one or two per element.
        const unsigned long long o0 = sm_off[k0];
        const uint32_t L = (uint32_t)(sm_off[k1] - o0);
        __syncthreads();
        if (tid >= k0 && tid <= k1) sm_rel[tid] = (uint32_t)(sm_off[tid] - o0);
        __syncthreads();
        const uint32_t xbase = gu.xbase;
        const uint32_t lane = tid & 31u, wrp = tid >> 5;
        if constexpr (sizeof(MT) == 4 || sizeof(MT) == 2) {
            // plain u32 / u16 members: 16-byte loads (4 / 8 members per lane and instruction) over the ALIGNED vectors
            // that cover the array; elements before its start or past its end are masked
            constexpr uint32_t VEC = 16 / sizeof(MT);
            const unsigned long long v_begin = o0 / VEC, v_end = (o0 + L + VEC - 1) / VEC;
            const uint32_t nvec = (uint32_t)(v_end - v_begin);
            const uint32_t per_warp = ((nvec + 8u * 32u - 1u) / (8u * 32u)) * 32u;
            const uint32_t wb = wrp * per_warp, we = min(nvec, wb + per_warp);
            const uint32_t head = (uint32_t)(o0 - v_begin * VEC);      // elements of the first vector that belong to the previous colours
            uint32_t k = 0, next = 0;
            if (wb < we) {  // colour of the first element of the warp's slice: one binary search
                const uint32_t e_first = wb * VEC > head ? wb * VEC - head : 0u;
                uint32_t lo = k0, hi = k1;
                while (hi - lo > 1) {
                    const uint32_t mid = (lo + hi) >> 1;
                    if (sm_rel[mid] <= e_first) lo = mid; else hi = mid;
                }
                k = lo;
                next = sm_rel[k + 1];
            }
            const uint4* vbase = reinterpret_cast<const uint4*>(gram_members) + v_begin;
            for (uint32_t v0 = wb + lane; v0 < we; v0 += 32u * 4u) {
                uint4 q[4];
#pragma unroll
                for (uint32_t u = 0; u < 4; u++) {
                    const uint32_t v = v0 + 32u * u;
                    q[u] = v < we ? __ldg(vbase + v) : make_uint4(0u, 0u, 0u, 0u);
                }
#pragma unroll
                for (uint32_t u = 0; u < 4; u++) {
                    const uint32_t v = v0 + 32u * u;
                    if (v >= we) break;
                    const uint32_t w4[4] = {q[u].x, q[u].y, q[u].z, q[u].w};
#pragma unroll
                    for (uint32_t t = 0; t < VEC; t++) {
                        const uint32_t e = v * VEC + t - head;  // wraps (>= L) for the elements in front of the array
                        if (e >= L) continue;
                        while (next <= e) { k++; next = sm_rel[k + 1]; }
                        const uint32_t m = sizeof(MT) == 4 ? w4[t] : ((w4[t >> 1] >> ((t & 1u) * 16u)) & 0xFFFFu);
                        const uint32_t x = m - xbase;           // members outside the segment wrap to >= slen
                        if (x < slen) atomicOr(sm + (k >> 5) * kGramSeg + x, 1u << (k & 31u));
                    }
                }
            }
        } else {
            // packed members (10 / 12 bits): element-wise loads, sixteen in flight per lane
            const uint32_t per_warp = ((L + 8u * 32u - 1u) / (8u * 32u)) * 32u;   // multiple of 32
            const uint32_t wb = wrp * per_warp, we = min(L, wb + per_warp);
            uint32_t k = 0, next = 0;
            if (wb < we) {
                uint32_t lo = k0, hi = k1;
                while (hi - lo > 1) {
                    const uint32_t mid = (lo + hi) >> 1;
                    if (sm_rel[mid] <= wb) lo = mid; else hi = mid;
                }
                k = lo;
                next = sm_rel[k + 1];
            }
            for (uint32_t e0 = wb + lane; e0 < we; e0 += 32u * 16u) {
                uint32_t xs[16];
#pragma unroll
                for (uint32_t u = 0; u < 16; u++) {
                    const uint32_t e = e0 + 32u * u;
                    xs[u] = e < we ? ldm<MT>(gram_members, o0 + e) - xbase : 0xFFFFFFFFu;
                }
#pragma unroll
                for (uint32_t u = 0; u < 16; u++) {
                    const uint32_t e = e0 + 32u * u;
                    if (e >= we) break;
                    while (next <= e) { k++; next = sm_rel[k + 1]; }
                    if (xs[u] < slen) atomicOr(sm + (k >> 5) * kGramSeg + xs[u], 1u << (k & 31u));
                }
            }
        }
        if (gu.wq_off != 0xFFFFFFFFu && tid < kKB) {
            const uint32_t wt = tid < kn ? ldw(gram_weights, weight_bits, gu.c0k + tid) : 0u;
            for (uint32_t p = 0; p < planes; p++)
                wq[(unsigned long long)p * kpad + (unsigned long long)gu.wq_off + kperm(tid)] = (uint8_t)(wt >> (8u * p));
        }
        __syncthreads();
        uint4* dst = reinterpret_cast<uint4*>(bits) + gu.bits_row;
        if (gu.nw == 4) {
            for (uint32_t r = tid; r < slen; r += 256) dst[r] = make_uint4(sm[r], sm[kGramSeg + r], sm[2u * kGramSeg + r], sm[3u * kGramSeg + r]);
        } else {
            uint32_t* dw = reinterpret_cast<uint32_t*>(dst);
            for (uint32_t wd = gu.w0; wd < (uint32_t)gu.w0 + gu.nw; wd++)
                for (uint32_t r = tid; r < slen; r += 256) dw[4u * r + wd] = sm[wd * kGramSeg + r];
        }
        __syncthreads();
    }
}

// kb_weights: the weight byte planes of a chunk whose bitmaps are part of the blob (IndexDesc::gram_form 1) -- all that
// is left of kb_build there.  One CTA per K-block of the chunk; wq[plane * kpad + (k0 + kblock) * 128 + kperm(k)].
__global__ void __launch_bounds__(kKB) kb_weights(const GramWinDev* __restrict__ wins, uint32_t n_wins,
                                                  const void* __restrict__ gram_weights, uint32_t weight_bits,
                                                  uint32_t kpad, uint32_t planes, uint8_t* __restrict__ wq) {
    __shared__ uint32_t s_win;
    const uint32_t kb = blockIdx.x, tid = threadIdx.x;
    if (tid == 0) {  // the window whose K-blocks [k0, k0 + ceil(n / 128)) hold kb: windows are listed in ascending k0
        uint32_t lo = 0, hi = n_wins;
        while (hi - lo > 1) {
            const uint32_t mid = (lo + hi) >> 1;
            if (wins[mid].k0 <= kb) lo = mid; else hi = mid;
        }
        s_win = lo;
    }
    __syncthreads();
    const GramWinDev w = wins[s_win];
    const uint32_t k = (kb - w.k0) * kKB + tid;
    const uint32_t wt = k < w.n ? ldw(gram_weights, weight_bits, w.c0 + k) : 0u;
    for (uint32_t p = 0; p < planes; p++)
        wq[(unsigned long long)p * kpad + (unsigned long long)kb * kKB + kperm(tid)] = (uint8_t)(wt >> (8u * p));
}

cudaError_t launch_kb_weights(const GramWinDev* wins, uint32_t n_wins, uint32_t kblocks, const void* gram_weights, uint32_t weight_bits,
                              uint32_t kpad, uint32_t planes, uint8_t* wq, const LaunchCfg& lc) {
    if (kblocks == 0) return cudaSuccess;
    kb_weights<<<kblocks, kKB, 0, lc.stream>>>(wins, n_wins, gram_weights, weight_bits, kpad, planes, wq);
    return cudaGetLastError();
}

cudaError_t launch_kb_build(const unsigned long long* gram_off, const void* gram_members, uint32_t member_bits,
                            const void* gram_weights, uint32_t weight_bits, const GramUnit* units, uint32_t n_units,
                            uint32_t kpad, uint32_t planes, uint32_t* bits, uint8_t* wq, const LaunchCfg& lc) {
    if (n_units == 0) return cudaSuccess;
    const unsigned cap = (unsigned)lc.sms * 6u;  // 32 KB of shared memory per CTA
    const unsigned ctas = n_units < cap ? n_units : cap;
    if (member_bits == 10)
        kb_build<P10><<<ctas, 256, 0, lc.stream>>>(gram_off, static_cast<const P10*>(gram_members), gram_weights, weight_bits, units, n_units, kpad, planes, bits, wq);
    else if (member_bits == 12)
        kb_build<P12><<<ctas, 256, 0, lc.stream>>>(gram_off, static_cast<const P12*>(gram_members), gram_weights, weight_bits, units, n_units, kpad, planes, bits, wq);
    else if (member_bits == 16)
        kb_build<uint16_t><<<ctas, 256, 0, lc.stream>>>(gram_off, static_cast<const uint16_t*>(gram_members), gram_weights, weight_bits, units, n_units, kpad, planes, bits, wq);
    else
        kb_build<uint32_t><<<ctas, 256, 0, lc.stream>>>(gram_off, static_cast<const uint32_t*>(gram_members), gram_weights, weight_bits, units, n_units, kpad, planes, bits, wq);
    return cudaGetLastError();
}


// ------------------------------------------------------------------------------------------------
// k4_gram<P>: persistent, one CTA per SM: 16 producer/epilogue warps, 1 MMA-issuing thread, 1 TMA
// loader thread.  Work item = (window, row tile ti, NB = 4/P adjacent column tiles >= ti, K-block
// range), read from the chunk's item list at an index fetched from a global counter.
//   loader    per K-block of 128 colours: cp.async.bulk of the bitmap rows of the A tile (2 KB), of
//             the NB B tiles (NB x 2 KB, contiguous) and of the weight byte planes (P x 128 B) into a
//             4-slot ring, two K-blocks ahead of the producers (`ready` mbarriers, complete_tx).
//   producers expand the bits into the 128B-swizzled K-major u8 tiles of a stage:
//             A_p[row][k] = bit ? byte_p(w_k) : 0 (one tile per weight byte plane), B[row][k] = bit
//             (NB tiles, contiguous, so two column tiles are ONE N = 256 UMMA); every warp arrives on
//             the stage's `full` mbarrier.
//   MMA       waits `full`, issues 4 K-steps x P planes of tcgen05.mma kind::i8 (u8 x u8 -> s32,
//             accumulators in the 512 TMEM columns), commits them to the stage's `empty` mbarrier:
//             expansion of block kb+1 overlaps the MMAs of block kb.
//   epilogue  tcgen05.ld, byte planes combined to u64, 32x16 transposes through shared memory so
//             that one warp instruction adds 2 x 16 consecutive cells of triangle rows
//             (red.global.add.u64).
// ------------------------------------------------------------------------------------------------
template <int P, bool C32>
__global__ void __maxnreg__(P == 4 ? 72 : 64) k4_gram(const __grid_constant__ GramParams prm,
                                                        const __grid_constant__ TriGeom G) {
    constexpr uint32_t NB = 4 / P;
    constexpr uint32_t kStageBytes = (P + NB) * kTileBytes;
    constexpr uint32_t kSlotBytes = (1 + NB) * 2048 + P * 128;  // bitmap rows of (1 + NB) tiles + weight planes
    constexpr uint32_t kWOff = (1 + NB) * 2048;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* ring = smem + kStages * kStageBytes;
    const uint32_t smem_a = smem_u32(smem), ring_a = smem_a + kStages * kStageBytes;
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(ring + kRing * kSlotBytes);
    // bars: full[kStages], empty[kStages], done, ready[kRing]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 1 + kRing);
    uint32_t* item_slot = tmem_slot + 1;
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;

    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(kTmemCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (tid == 32) {
        for (uint32_t s = 0; s < kStages; s++) {
            mbar_init(smem_u32(bars + s), kProducerWarps);
            mbar_init(smem_u32(bars + kStages + s), 1);
        }
        mbar_init(smem_u32(bars + 2 * kStages), 1);
        for (uint32_t s = 0; s < kRing; s++) mbar_init(smem_u32(bars + 2 * kStages + 1 + s), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    const uint32_t bar_done = smem_u32(bars + 2 * kStages);
    const uint32_t bar_ready0 = smem_u32(bars + 2 * kStages + 1);
    uint32_t it = 0, done_phase = 0;  // it: K-blocks processed by this CTA so far, the same in every role

    for (;;) {
        if (tid == 0) *item_slot = atomicAdd(prm.counter, 1u);
        __syncthreads();
        const uint32_t item = *item_slot;
        if (item >= prm.n_items) break;
        const GramItem gi = prm.items[item];
        const GramWinDev win = prm.wins[gi.win];
        const uint32_t i0 = gi.i0, j0 = gi.j0;
        const uint32_t ncol = min(NB, (win.width - j0 + kTile - 1) / kTile);
        const uint32_t kb0 = gi.kb0, kb1 = gi.kb1;

        if (warp < kProducerWarps) {
            // ---------------- producers: expansion
            const uint32_t nunits = (1u + ncol) * 256u;  // (tile, row, half K-block)
            for (uint32_t kb = kb0; kb < kb1; kb++, it++) {
                const uint32_t st = it % kStages, slot = it % kRing;
                mbar_wait(bar_ready0 + 8u * slot, (it / kRing) & 1u);                                      // bits + weights landed
                if (it >= kStages) mbar_wait(smem_u32(bars + kStages + st), ((it / kStages) - 1u) & 1u);  // stage drained by the MMAs
                const uint32_t stage = smem_a + st * kStageBytes;
                const uint32_t rs = ring_a + slot * kSlotBytes;
#pragma unroll
                for (uint32_t k = 0; k < 3; k++) {
                    const uint32_t u = tid + kProducers * k;
                    if (u >= nunits) continue;
                    const uint32_t tsel = u >> 8, r = (u & 255u) >> 1, h = u & 1u, sw = r & 7u;
                    const uint2 bits = lds64(rs + u * 8u);  // slot rows are (tile, row, half) ordered
                    const uint32_t drow = stage + (tsel == 0 ? 0u : (P + tsel - 1) * kTileBytes) + (r >> 3) * 1024u + (r & 7u) * 128u;
#pragma unroll
                    for (uint32_t wd = 0; wd < 2; wd++) {
                        const uint32_t word = wd ? bits.y : bits.x;
                        uint32_t s4[8];  // 32 bits -> 8 x 4 bytes of 0/1
#pragma unroll
                        for (uint32_t n = 0; n < 8; n++) s4[n] = (word >> n) & 0x01010101u;  // K order inside a word: see kperm()
#pragma unroll
                        for (uint32_t c = 0; c < 2; c++) {
                            const uint32_t chunk = (h * 2 + wd) * 2 + c;  // 16-byte chunk inside the 128-byte row
                            const uint32_t dst = drow + ((chunk ^ sw) << 4);
                            if (tsel != 0) {
                                sts128(dst, s4[4 * c], s4[4 * c + 1], s4[4 * c + 2], s4[4 * c + 3]);
                            } else {
#pragma unroll
                                for (int p = 0; p < P; p++) {
                                    const uint4 w = lds128(rs + kWOff + p * 128 + chunk * 16);
                                    sts128(dst + p * kTileBytes, (s4[4 * c] * 0xFFu) & w.x, (s4[4 * c + 1] * 0xFFu) & w.y,
                                           (s4[4 * c + 2] * 0xFFu) & w.z, (s4[4 * c + 3] * 0xFFu) & w.w);
                                }
                            }
                        }
                    }
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to the tensor core
                __syncwarp();
                if (lane == 0) mbar_arrive(smem_u32(bars + st));
            }
            // ---------------- epilogue
            if (kb0 < kb1) {
                mbar_wait(bar_done, done_phase);
                done_phase ^= 1u;
                tc_fence_after();
                const uint32_t tp = smem_a + warp * (32 * 17 * 4);  // stage memory is idle now
                const uint32_t lq = warp & 3u, wg = warp >> 2;
                const uint32_t rsel = lane >> 4, csel = lane & 15u;
                const uint32_t ext = win.extent ? win.extent : win.width;
                for (uint32_t ch = wg; ch < ncol * 8u; ch += 4) {
                    const uint32_t ct = ch >> 3, c16 = ch & 7u;
                    unsigned long long acc[16];
#pragma unroll
                    for (uint32_t y = 0; y < 16; y++) acc[y] = 0ull;
#pragma unroll
                    for (int p = 0; p < P; p++) {
                        uint32_t v[16];
                        tmem_ld16(tmem + ((32u * lq) << 16) + (p * NB + ct) * kTile + c16 * 16u, v);
                        __syncwarp();
#pragma unroll
                        for (uint32_t x = 0; x < 16; x++) sts32(tp + (lane * 17 + x) * 4u, v[x]);  // row = lane, column = x
                        __syncwarp();
#pragma unroll
                        for (uint32_t y = 0; y < 16; y++)
                            acc[y] += (unsigned long long)lds32(tp + ((2 * y + rsel) * 17 + csel) * 4u) << (8 * p);  // row = 2y + rsel, column = csel
                    }
                    const uint32_t jl = j0 + ct * kTile + c16 * 16u + csel;  // tiles may hang over the window's end:
                    const uint32_t j = win.x0 + jl;                           // those rows / columns hold foreign bits
#pragma unroll
                    for (uint32_t y = 0; y < 16; y++) {
                        const uint32_t il = i0 + 32u * lq + 2 * y + rsel;
                        const uint32_t i = win.x0 + il;
                        if (il < jl && jl < ext && j < G.n && acc[y] != 0ull) {  // columns (hence rows) past the extent may hold foreign bits
                            const unsigned long long ra = gram_row_addr<C32>(i, G);
                            if (ra != 0ull) gram_red<C32>(ra, j, acc[y]);
                        }
                    }
                }
            }
        } else if (warp == kProducerWarps) {
            // ---------------- MMA issuer (one thread)
            if (lane == 0) {
                const uint32_t n0 = kTile * min(ncol, 2u);
                const uint32_t n1 = ncol > 2 ? kTile * (ncol - 2) : 0u;
                for (uint32_t kb = kb0; kb < kb1; kb++, it++) {
                    const uint32_t st = it % kStages;
                    mbar_wait(smem_u32(bars + st), (it / kStages) & 1u);
                    tc_fence_after();
                    const uint32_t sbase = smem_u32(smem + st * kStageBytes);
                    const uint64_t bdesc0 = umma_desc_sw128(sbase + P * kTileBytes);
                    const uint64_t bdesc1 = umma_desc_sw128(sbase + (P + 2) * kTileBytes);
#pragma unroll
                    for (int p = 0; p < P; p++) {
                        const uint64_t adesc = umma_desc_sw128(sbase + p * kTileBytes);
#pragma unroll
                        for (uint32_t k = 0; k < 4; k++) {  // 4 x K=32 inside the swizzle atom: +32 B = +2 in the address field
                            const uint32_t acc = (kb > kb0 || k > 0) ? 1u : 0u;
                            umma_i8(tmem + (p * NB) * kTile, adesc + 2 * k, bdesc0 + 2 * k, umma_idesc(n0), acc);
                            if (NB > 2 && n1) umma_i8(tmem + (p * NB + 2) * kTile, adesc + 2 * k, bdesc1 + 2 * k, umma_idesc(n1), acc);
                        }
                    }
                    umma_commit(smem_u32(bars + kStages + st));
                    if (kb + 1 == kb1) umma_commit(bar_done);
                }
            } else {
                it += kb1 - kb0;
            }
        } else {
            // ---------------- TMA loader (one thread): bitmap rows + weights, two K-blocks ahead
            if (lane == 0) {
                const uint32_t bytes = (1u + ncol) * 2048u + P * 128u;
                const uint8_t* gbits = reinterpret_cast<const uint8_t*>(prm.bits);
                for (uint32_t kb = kb0; kb < kb1; kb++, it++) {
                    // slot it % 4 was read for K-block it - 4; its readers finished before `full` of it - 2 completed
                    if (kb >= kb0 + 2) mbar_wait(smem_u32(bars + (it & 1u)), ((it - 2u) / kStages) & 1u);
                    const uint32_t slot = it % kRing;
                    const uint32_t bar = bar_ready0 + 8u * slot;
                    const uint32_t dst = smem_u32(ring + slot * kSlotBytes);
                    const uint8_t* kbase = gbits + (win.bits_off + (unsigned long long)kb * win.kstride) * 16ull;
                    mbar_expect_tx(bar, bytes);
                    tma_load_1d(dst, kbase + (unsigned long long)i0 * 16ull, 2048u, bar);
                    tma_load_1d(dst + 2048u, kbase + (unsigned long long)j0 * 16ull, ncol * 2048u, bar);
#pragma unroll
                    for (int p = 0; p < P; p++)
                        tma_load_1d(dst + kWOff + p * 128, prm.wq + (unsigned long long)p * prm.kpad + (unsigned long long)(win.k0 + kb) * kKB, 128u, bar);
                }
            } else {
                it += kb1 - kb0;
            }
        }
        tc_fence_before();
        __syncthreads();  // TMEM drained, stage + ring memory free, item_slot reusable
        tc_fence_after();
    }
    if (warp == 0) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(kTmemCols) : "memory");
    }
}

// ------------------------------------------------------------------------------------------------
// k_imma_peak: the tensor-pipe ceiling K4 is measured against, in K4's own instruction shape.  One CTA per SM
// fills one A tile and two pairs of B tiles with bytes ONCE and then issues `iters` K-blocks of exactly the
// UMMAs k4_gram<1> issues per K-block (4 K-steps x 2 N = 256 instructions, kind::i8, cta_group::1, M = 128,
// accumulators in the 512 TMEM columns) back to back: no expansion, no loads, no epilogue.  bench.py divides
// K4's u8 MAC rate by this rate ("int8 tensor peak ... must be measured in the same run").
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128, 1) k_imma_peak(uint32_t iters, uint32_t* sink) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem + 5 * kTileBytes);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 1);
    const uint32_t tid = threadIdx.x, warp = tid >> 5;
    for (uint32_t i = tid; i < 5 * kTileBytes / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x01010101u * ((i * 2654435761u) >> 30);
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(kTmemCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (tid == 32) {
        mbar_init(smem_u32(bar), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    if (tid == 0) {
        const uint32_t sbase = smem_u32(smem);
        const uint64_t adesc = umma_desc_sw128(sbase);
        const uint64_t bdesc0 = umma_desc_sw128(sbase + kTileBytes), bdesc1 = umma_desc_sw128(sbase + 3 * kTileBytes);
        for (uint32_t it = 0; it < iters; it++) {
#pragma unroll
            for (uint32_t k = 0; k < 4; k++) {
                const uint32_t acc = (it > 0 || k > 0) ? 1u : 0u;
                umma_i8(tmem, adesc + 2 * k, bdesc0 + 2 * k, umma_idesc(256), acc);
                umma_i8(tmem + 256, adesc + 2 * k, bdesc1 + 2 * k, umma_idesc(256), acc);
            }
        }
        umma_commit(smem_u32(bar));
    }
    mbar_wait(smem_u32(bar), 0);
    tc_fence_after();
    uint32_t v[16];
    tmem_ld16(tmem + ((32u * warp) << 16), v);   // one read so that the accumulators are observable
    if (v[0] == 0xFFFFFFFFu && sink) sink[0] = v[1];
    tc_fence_before();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(kTmemCols) : "memory");
}

cudaError_t launch_imma_peak(uint32_t iters, uint32_t* sink, double* macs, const LaunchCfg& lc) {
    const size_t smem = 1024 + 5 * (size_t)kTileBytes + 64;
    cudaError_t e = cudaFuncSetAttribute(k_imma_peak, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    k_imma_peak<<<(unsigned)lc.sms, 128, smem, lc.stream>>>(iters, sink);
    if (macs) *macs = (double)lc.sms * (double)iters * 8.0 * 128.0 * 256.0 * 32.0;
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// k4_gram_pipe<P> (P = 1, 2): the same Gram product for chunks whose items are SHORT (a handful of K-blocks per
// tile pair: clade / family windows of a FracMinHash index).  There the time of k4_gram goes to the per-item
// serial chain -- fetch the item, fill the pipeline, wait for the last MMA, run the epilogue, CTA-wide barrier --
// not to the tensor pipe.  This variant removes the chain:
//   * static schedule: CTA c takes items c, c + gridDim, ... of the plan's cost-sorted list, so every role walks
//     the same item sequence on its own -- no counter, no CTA-wide barrier between items;
//   * the accumulators are DOUBLE-BUFFERED in TMEM (2 x 256 columns: an item covers NB = 2 / P column tiles), the
//     MMA warp starts item n + 1 in the other buffer as soon as its operands arrive;
//   * the producer warps expand item n + 1 BEFORE they run the epilogue of item n, so the tensor core works on
//     n + 1 (and the loader is already fetching n + 2) while TMEM -> red of item n runs;
//   * three operand stages, the loader's 4-slot ring freed by its own mbarrier, epilogue transposes in their own
//     shared scratch (the stages stay busy).
// Roles and operand expansion as in k4_gram.  `extent` of a window (samples really touched) lets the epilogue skip
// the row / column groups that are padding, and groups entirely below the diagonal are skipped.
// ------------------------------------------------------------------------------------------------
constexpr uint32_t kPipeStages = 3;
template <int P, bool C32>
__global__ void __maxnreg__(64) k4_gram_pipe(const __grid_constant__ GramParams prm, const __grid_constant__ TriGeom G) {
    constexpr uint32_t NB = 2 / P;
    constexpr uint32_t kStageBytes = (P + NB) * kTileBytes;
    constexpr uint32_t kSlotBytes = (1 + NB) * 2048 + P * 128;
    constexpr uint32_t kWOff = (1 + NB) * 2048;
    constexpr uint32_t kScratch = kProducerWarps * 32 * 17 * 4;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* ring = smem + kPipeStages * kStageBytes;
    const uint32_t smem_a = smem_u32(smem), ring_a = smem_a + kPipeStages * kStageBytes;
    const uint32_t scratch_a = ring_a + kRing * kSlotBytes;
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(ring + kRing * kSlotBytes + kScratch);
    // bars: full[3], empty[3], ready[4], ring_free[4], acc_full[2], acc_free[2]
    const uint32_t bar_full = smem_u32(bars), bar_empty = bar_full + 8u * kPipeStages, bar_ready = bar_empty + 8u * kPipeStages;
    const uint32_t bar_rfree = bar_ready + 8u * kRing, bar_afull = bar_rfree + 8u * kRing, bar_afree = bar_afull + 16u;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kPipeStages + 2 * kRing + 4);
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;

    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(kTmemCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (tid == 32) {
        for (uint32_t s = 0; s < kPipeStages; s++) { mbar_init(bar_full + 8u * s, kProducerWarps); mbar_init(bar_empty + 8u * s, 1); }
        for (uint32_t s = 0; s < kRing; s++) { mbar_init(bar_ready + 8u * s, 1); mbar_init(bar_rfree + 8u * s, kProducerWarps); }
        for (uint32_t s = 0; s < 2; s++) { mbar_init(bar_afull + 8u * s, 1); mbar_init(bar_afree + 8u * s, kProducerWarps); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp < kProducerWarps) {
        // ---------------- producers: expand item n + 1, then the epilogue of item n
        uint32_t it = 0, n = 0;
        GramItem pgi{};          // the item whose epilogue is pending
        GramWinDev pwin{};
        bool pending = false;
        auto epilogue = [&](const GramItem& gi, const GramWinDev& win, uint32_t m) {
            const uint32_t b = m & 1u;
            mbar_wait(bar_afull + 8u * b, (m >> 1) & 1u);
            tc_fence_after();
            const uint32_t i0 = gi.i0, j0 = gi.j0;
            const uint32_t ncol = min(NB, (win.width - j0 + kTile - 1) / kTile);
            const uint32_t ext = win.extent ? win.extent : win.width;
            const uint32_t tp = scratch_a + warp * (32 * 17 * 4);
            const uint32_t lq = warp & 3u, wg = warp >> 2;
            const uint32_t rsel = lane >> 4, csel = lane & 15u;
            const uint32_t row0 = i0 + 32u * lq;            // first row of this warp's 32-row group
            if (row0 < ext) {
                // address of the virtual cell (row, 0) of the warp's 32 rows, one per lane: the chunk loop takes
                // them by shuffle instead of a dependent global load in front of every red
                const unsigned long long ra_lane = (row0 + lane < win.width) ? gram_row_addr<C32>(win.x0 + row0 + lane, G) : 0ull;
                const uint32_t ra_lo = (uint32_t)ra_lane, ra_hi = (uint32_t)(ra_lane >> 32);
                for (uint32_t ch = wg; ch < ncol * 8u; ch += 4) {
                    const uint32_t ct = ch >> 3, c16 = ch & 7u;
                    const uint32_t col0 = j0 + ct * kTile + c16 * 16u;
                    if (col0 >= ext || col0 + 15u <= row0) continue;  // padding columns, or every cell on / below the diagonal (warp-uniform)
                    unsigned long long acc[16];
#pragma unroll
                    for (uint32_t y = 0; y < 16; y++) acc[y] = 0ull;
#pragma unroll
                    for (int p = 0; p < P; p++) {
                        uint32_t v[16];
                        tmem_ld16(tmem + ((32u * lq) << 16) + b * 256u + (p * NB + ct) * kTile + c16 * 16u, v);
                        __syncwarp();
#pragma unroll
                        for (uint32_t x = 0; x < 16; x++) sts32(tp + (lane * 17 + x) * 4u, v[x]);
                        __syncwarp();
#pragma unroll
                        for (uint32_t y = 0; y < 16; y++)
                            acc[y] += (unsigned long long)lds32(tp + ((2 * y + rsel) * 17 + csel) * 4u) << (8 * p);
                    }
                    const uint32_t jl = col0 + csel;
                    const uint32_t j = win.x0 + jl;
                    const bool col_ok = jl < ext && j < G.n;  // columns (hence rows) past the extent may hold foreign bits
#pragma unroll
                    for (uint32_t y = 0; y < 16; y++) {
                        const uint32_t src = 2 * y + rsel;
                        const unsigned long long ra = ((unsigned long long)__shfl_sync(0xffffffffu, ra_hi, src) << 32) | __shfl_sync(0xffffffffu, ra_lo, src);
                        if (row0 + src < jl && col_ok && acc[y] != 0ull && ra != 0ull) gram_red<C32>(ra, j, acc[y]);
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_afree + 8u * b);
        };
        for (uint32_t item = blockIdx.x; item < prm.n_items; item += gridDim.x, n++) {
            const GramItem gi = prm.items[item];
            const GramWinDev win = prm.wins[gi.win];
            const uint32_t ncol = min(NB, (win.width - gi.j0 + kTile - 1) / kTile);
            const uint32_t nunits = (1u + ncol) * 256u;  // (tile, row, half K-block)
            for (uint32_t kb = gi.kb0; kb < gi.kb1; kb++, it++) {
                const uint32_t st = it % kPipeStages, slot = it % kRing;
                mbar_wait(bar_ready + 8u * slot, (it / kRing) & 1u);
                if (it >= kPipeStages) mbar_wait(bar_empty + 8u * st, ((it / kPipeStages) - 1u) & 1u);
                const uint32_t stage = smem_a + st * kStageBytes;
                const uint32_t rs = ring_a + slot * kSlotBytes;
#pragma unroll
                for (uint32_t k = 0; k < 2; k++) {
                    const uint32_t u = tid + kProducers * k;
                    if (u >= nunits) continue;
                    const uint32_t tsel = u >> 8, r = (u & 255u) >> 1, h = u & 1u, sw = r & 7u;
                    const uint2 bits = lds64(rs + u * 8u);
                    const uint32_t drow = stage + (tsel == 0 ? 0u : (P + tsel - 1) * kTileBytes) + (r >> 3) * 1024u + (r & 7u) * 128u;
#pragma unroll
                    for (uint32_t wd = 0; wd < 2; wd++) {
                        const uint32_t word = wd ? bits.y : bits.x;
                        uint32_t s4[8];
#pragma unroll
                        for (uint32_t q = 0; q < 8; q++) s4[q] = (word >> q) & 0x01010101u;  // K order inside a word: see kperm()
#pragma unroll
                        for (uint32_t c = 0; c < 2; c++) {
                            const uint32_t chunk = (h * 2 + wd) * 2 + c;
                            const uint32_t dst = drow + ((chunk ^ sw) << 4);
                            if (tsel != 0) {
                                sts128(dst, s4[4 * c], s4[4 * c + 1], s4[4 * c + 2], s4[4 * c + 3]);
                            } else {
#pragma unroll
                                for (int p = 0; p < P; p++) {
                                    const uint4 w = lds128(rs + kWOff + p * 128 + chunk * 16);
                                    sts128(dst + p * kTileBytes, (s4[4 * c] * 0xFFu) & w.x, (s4[4 * c + 1] * 0xFFu) & w.y,
                                           (s4[4 * c + 2] * 0xFFu) & w.z, (s4[4 * c + 3] * 0xFFu) & w.w);
                                }
                            }
                        }
                    }
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncwarp();
                if (lane == 0) { mbar_arrive(bar_full + 8u * st); mbar_arrive(bar_rfree + 8u * slot); }
            }
            if (pending) epilogue(pgi, pwin, n - 1u);
            pgi = gi;
            pwin = win;
            pending = true;
        }
        if (pending) epilogue(pgi, pwin, n - 1u);
    } else if (warp == kProducerWarps) {
        // ---------------- MMA issuer (one thread)
        if (lane == 0) {
            uint32_t it = 0, n = 0;
            for (uint32_t item = blockIdx.x; item < prm.n_items; item += gridDim.x, n++) {
                const GramItem gi = prm.items[item];
                const GramWinDev win = prm.wins[gi.win];
                const uint32_t ncol = min(NB, (win.width - gi.j0 + kTile - 1) / kTile);
                const uint32_t b = n & 1u;
                if (n >= 2u) { mbar_wait(bar_afree + 8u * b, ((n >> 1) - 1u) & 1u); tc_fence_after(); }
                const uint32_t nn = kTile * ncol;
                for (uint32_t kb = gi.kb0; kb < gi.kb1; kb++, it++) {
                    const uint32_t st = it % kPipeStages;
                    mbar_wait(bar_full + 8u * st, (it / kPipeStages) & 1u);
                    tc_fence_after();
                    const uint32_t sbase = smem_a + st * kStageBytes;
                    const uint64_t bdesc = umma_desc_sw128(sbase + P * kTileBytes);
#pragma unroll
                    for (int p = 0; p < P; p++) {
                        const uint64_t adesc = umma_desc_sw128(sbase + p * kTileBytes);
#pragma unroll
                        for (uint32_t k = 0; k < 4; k++)
                            umma_i8(tmem + b * 256u + (p * NB) * kTile, adesc + 2 * k, bdesc + 2 * k, umma_idesc(nn), (kb > gi.kb0 || k > 0) ? 1u : 0u);
                    }
                    umma_commit(bar_empty + 8u * st);
                    if (kb + 1 == gi.kb1) umma_commit(bar_afull + 8u * b);
                }
            }
        }
    } else {
        // ---------------- TMA loader (one thread)
        if (lane == 0) {
            uint32_t it = 0;
            const uint8_t* gbits = reinterpret_cast<const uint8_t*>(prm.bits);
            for (uint32_t item = blockIdx.x; item < prm.n_items; item += gridDim.x) {
                const GramItem gi = prm.items[item];
                const GramWinDev win = prm.wins[gi.win];
                const uint32_t ncol = min(NB, (win.width - gi.j0 + kTile - 1) / kTile);
                const uint32_t bytes = (1u + ncol) * 2048u + P * 128u;
                for (uint32_t kb = gi.kb0; kb < gi.kb1; kb++, it++) {
                    const uint32_t slot = it % kRing;
                    if (it >= kRing) mbar_wait(bar_rfree + 8u * slot, ((it / kRing) - 1u) & 1u);
                    const uint32_t bar = bar_ready + 8u * slot;
                    const uint32_t dst = ring_a + slot * kSlotBytes;
                    const uint8_t* kbase = gbits + (win.bits_off + (unsigned long long)kb * win.kstride) * 16ull;
                    mbar_expect_tx(bar, bytes);
                    tma_load_1d(dst, kbase + (unsigned long long)gi.i0 * 16ull, 2048u, bar);
                    tma_load_1d(dst + 2048u, kbase + (unsigned long long)gi.j0 * 16ull, ncol * 2048u, bar);
#pragma unroll
                    for (int p = 0; p < P; p++)
                        tma_load_1d(dst + kWOff + p * 128, prm.wq + (unsigned long long)p * prm.kpad + (unsigned long long)(win.k0 + kb) * kKB, 128u, bar);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 0) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(kTmemCols) : "memory");
    }
}

static size_t gram_pipe_smem_bytes(int P) {
    const size_t nb = 2 / P;
    return 1024 + kPipeStages * (P + nb) * (size_t)kTileBytes + kRing * ((1 + nb) * 2048 + P * 128) + kProducerWarps * 32 * 17 * 4 +
           8 * (2 * kPipeStages + 2 * kRing + 4) + 16;
}

template <int P>
static cudaError_t launch_gram_pipe_p(const GramParams& prm, const TriGeom& g, const LaunchCfg& lc) {
    const size_t smem = gram_pipe_smem_bytes(P);
    const bool c32 = g.cell_bits == 32;
    cudaError_t e = c32 ? cudaFuncSetAttribute(k4_gram_pipe<P, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)
                        : cudaFuncSetAttribute(k4_gram_pipe<P, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    unsigned long long ctas = prm.n_items < (unsigned long long)lc.sms ? prm.n_items : (unsigned long long)lc.sms;
    if (c32) k4_gram_pipe<P, true><<<(unsigned)ctas, kThreads, smem, lc.stream>>>(prm, g);
    else k4_gram_pipe<P, false><<<(unsigned)ctas, kThreads, smem, lc.stream>>>(prm, g);
    return cudaGetLastError();
}

cudaError_t launch_k4_gram_pipe(const GramParams& prm, const TriGeom& g, const LaunchCfg& lc) {
    if (prm.n_items == 0) return cudaSuccess;
    switch (prm.planes) {
        case 1: return launch_gram_pipe_p<1>(prm, g, lc);
        case 2: return launch_gram_pipe_p<2>(prm, g, lc);
        default: return cudaErrorInvalidValue;
    }
}

uint32_t gram_cols_per_item_pipe(int planes) { return planes <= 1 ? 2u : 1u; }

uint32_t gram_cols_per_item(int planes) { return planes <= 1 ? 4u : (planes == 2 ? 2u : 1u); }

static size_t gram_smem_bytes(int P) {
    const size_t nb = 4 / P;
    return 1024 + kStages * (P + nb) * (size_t)kTileBytes + kRing * ((1 + nb) * 2048 + P * 128) + 8 * (2 * kStages + 1 + kRing) + 16;
}

template <int P>
static cudaError_t launch_gram_p(const GramParams& prm, const TriGeom& g, const LaunchCfg& lc) {
    const size_t smem = gram_smem_bytes(P);
    const bool c32 = g.cell_bits == 32;
    cudaError_t e = c32 ? cudaFuncSetAttribute(k4_gram<P, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)
                        : cudaFuncSetAttribute(k4_gram<P, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    unsigned long long ctas = prm.n_items < (unsigned long long)lc.sms ? prm.n_items : (unsigned long long)lc.sms;
    if (c32) k4_gram<P, true><<<(unsigned)ctas, kThreads, smem, lc.stream>>>(prm, g);
    else k4_gram<P, false><<<(unsigned)ctas, kThreads, smem, lc.stream>>>(prm, g);
    return cudaGetLastError();
}

cudaError_t launch_k4_gram(const GramParams& prm, const TriGeom& g, const LaunchCfg& lc) {
    if (prm.n_items == 0) return cudaSuccess;
    switch (prm.planes) {
        case 1: return launch_gram_p<1>(prm, g, lc);
        case 2: return launch_gram_p<2>(prm, g, lc);
        case 4: return launch_gram_p<4>(prm, g, lc);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace ksb
