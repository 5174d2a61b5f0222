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
constexpr uint32_t kThreads = 256;
constexpr uint32_t kTileBytes = kTile * kKB;  // 16 KB: one operand tile (128 rows x 128 B)

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
constexpr uint32_t kIdesc = (2u << 4) | ((kTile >> 3) << 17) | ((kTile >> 4) << 24);

__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(kIdesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

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

__device__ __forceinline__ unsigned long long gram_row_addr(uint32_t x, const TriGeom& g) {
    if (x - g.rb >= g.nrows) return 0ull;
    const unsigned long long t = ((unsigned long long)x * (unsigned long long)(g.two_n_m1 - x)) >> 1;
    return (unsigned long long)g.tri + 8ull * (t - (unsigned long long)x - 1ull - g.base);
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// kb_build: one warp per Gram colour of the chunk [c0, c0 + n).  Sets bit (colour, sample) in the
// K-block-major bitmap  bits[(kblock * npad + sample) * 4 + word]  and writes the byte planes of the
// colour's weight to wq[plane * kpad + colour].
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) kb_build(const unsigned long long* __restrict__ gram_off,
                                                const uint32_t* __restrict__ gram_members,
                                                const uint32_t* __restrict__ gram_weights, unsigned long long c0,
                                                uint32_t n, uint32_t npad, uint32_t kpad, uint32_t planes,
                                                uint32_t* __restrict__ bits, uint8_t* __restrict__ wq) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t cl = warp; cl < n; cl += nwarps) {
        const unsigned long long off = gram_off[c0 + cl];
        const uint32_t s = (uint32_t)(gram_off[c0 + cl + 1] - off);
        const uint32_t* m = gram_members + off;
        uint32_t* base = bits + (unsigned long long)(cl >> 7) * npad * 4ull + ((cl & 127u) >> 5);
        const uint32_t bit = 1u << (cl & 31u);
        for (uint32_t a = lane; a < s; a += 32) atomicOr(base + 4ull * __ldg(m + a), bit);
        if (lane < planes) wq[(unsigned long long)lane * kpad + cl] = (uint8_t)(__ldg(gram_weights + c0 + cl) >> (8u * lane));
    }
}

cudaError_t launch_kb_build(const unsigned long long* gram_off, const uint32_t* gram_members,
                            const uint32_t* gram_weights, uint64_t c0, uint32_t n, uint32_t npad, uint32_t kpad,
                            uint32_t planes, uint32_t* bits, uint8_t* wq, const LaunchCfg& lc) {
    if (n == 0) return cudaSuccess;
    unsigned ctas = (n + 7) / 8;
    const unsigned cap = (unsigned)lc.sms * 8u;
    if (ctas > cap) ctas = cap;
    kb_build<<<ctas, 256, 0, lc.stream>>>(gram_off, gram_members, gram_weights, c0, n, npad, kpad, planes, bits, wq);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// k4_gram<P>: persistent, one CTA per SM.  Work item = (row tile ti, column tile tj >= ti, K split).
// Per K-block of 128 colours all 256 threads expand bitmap words into the swizzled u8 tiles
// (A_p[row][k] = bit ? byte_p(w_k) : 0, B[row][k] = bit), thread 0 issues 4 x P UMMAs
// (128 x 128 x 32, u8 x u8 -> s32 in TMEM) and commits them to the stage's mbarrier, so the
// expansion of block kb+1 overlaps the MMAs of block kb.  Epilogue: TMEM -> registers
// (tcgen05.ld), planes combined to u64, 32x32 transposes through shared memory so that one warp
// instruction adds 32 consecutive cells of one triangle row.
// ------------------------------------------------------------------------------------------------
template <int P>
__global__ void __launch_bounds__(kThreads, 1) k4_gram(const __grid_constant__ GramParams prm,
                                                        const __grid_constant__ TriGeom G) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // stage layout: [P A tiles][B tile]; after the stages: weights ring, barriers, tmem pointer
    constexpr uint32_t kStageBytes = (P + 1) * kTileBytes;
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint32_t* wsm = reinterpret_cast<uint32_t*>(smem + kStages * kStageBytes);          // [2][P][32] u32 = 128 B per plane
    unsigned long long* bars = reinterpret_cast<unsigned long long*>(wsm + 2 * P * 32);  // empty[kStages], done
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + kStages + 1);
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    constexpr uint32_t kCols = (P == 1) ? 128u : (P == 2 ? 256u : 512u);

    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(kCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (tid == 32) {
        for (uint32_t s = 0; s <= kStages; s++) mbar_init(smem_u32(bars + s), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    const uint32_t bar_done = smem_u32(bars + kStages);

    const uint32_t r = tid & 127u;   // tile row this thread expands
    const uint32_t h = tid >> 7;     // which half of the K-block (words 2h, 2h+1)
    const uint32_t row_off = (r >> 3) * 1024u + (r & 7u) * 128u;
    const uint32_t sw = r & 7u;
    const uint32_t T = prm.tiles;
    const uint4* wq4 = reinterpret_cast<const uint4*>(prm.wq);

    uint32_t it = 0;          // K-blocks processed by this CTA so far (all items): stage / phase bookkeeping
    uint32_t done_phase = 0;
    for (unsigned long long item = blockIdx.x; item < prm.n_items; item += gridDim.x) {
        uint32_t q = (uint32_t)(item / prm.ksplit);
        const uint32_t sp = (uint32_t)(item % prm.ksplit);
        uint32_t ti = prm.tile_lo;
        while (q >= T - ti) { q -= T - ti; ti++; }
        const uint32_t tj = ti + q;
        const uint32_t kb0 = (uint32_t)((unsigned long long)prm.kblocks * sp / prm.ksplit);
        const uint32_t kb1 = (uint32_t)((unsigned long long)prm.kblocks * (sp + 1) / prm.ksplit);
        if (kb0 >= kb1) continue;
        const uint32_t i0 = ti * kTile, j0 = tj * kTile;

        // prologue: weights of kb0 straight to the ring, weights of kb0+1 and bits of kb0 to registers
        uint4 wreg = make_uint4(0, 0, 0, 0);
        if (tid < 8 * P) {
            const uint32_t p = tid >> 3, c = tid & 7u;
            reinterpret_cast<uint4*>(wsm + ((kb0 & 1u) * P + p) * 32)[c] = __ldg(wq4 + ((unsigned long long)p * prm.kpad + (unsigned long long)kb0 * kKB) / 16 + c);
            if (kb0 + 1 < kb1) wreg = __ldg(wq4 + ((unsigned long long)p * prm.kpad + (unsigned long long)(kb0 + 1) * kKB) / 16 + c);
        }
        const uint2* abits = reinterpret_cast<const uint2*>(prm.bits) + ((unsigned long long)i0 + r) * 2 + h;
        const uint2* bbits = reinterpret_cast<const uint2*>(prm.bits) + ((unsigned long long)j0 + r) * 2 + h;
        const unsigned long long kb_stride = (unsigned long long)prm.npad * 2;  // uint2 per K-block
        uint2 an = __ldg(abits + kb_stride * kb0), bn = __ldg(bbits + kb_stride * kb0);
        __syncthreads();

        for (uint32_t kb = kb0; kb < kb1; kb++, it++) {
            const uint32_t st = it % kStages;
            const uint2 a = an, b = bn;
            if (kb + 1 < kb1) { an = __ldg(abits + kb_stride * (kb + 1)); bn = __ldg(bbits + kb_stride * (kb + 1)); }
            // weights: ring slot of kb+1 <- registers, registers <- kb+2
            if (tid < 8 * P) {
                const uint32_t p = tid >> 3, c = tid & 7u;
                if (kb + 1 < kb1) reinterpret_cast<uint4*>(wsm + (((kb + 1) & 1u) * P + p) * 32)[c] = wreg;
                if (kb + 2 < kb1) wreg = __ldg(wq4 + ((unsigned long long)p * prm.kpad + (unsigned long long)(kb + 2) * kKB) / 16 + c);
            }
            if (it >= kStages) mbar_wait(smem_u32(bars + st), ((it / kStages) - 1u) & 1u);  // MMAs that read this stage are done
            uint8_t* stage = smem + st * kStageBytes;
            const uint32_t* wk = wsm + (kb & 1u) * P * 32 + h * 16;
#pragma unroll
            for (uint32_t wd = 0; wd < 2; wd++) {
                const uint32_t av = wd ? a.y : a.x, bv = wd ? b.y : b.x;
                uint32_t sa[8], sb[8];
#pragma unroll
                for (uint32_t n = 0; n < 8; n++) {
                    sa[n] = (((av >> (4 * n)) & 0xFu) * 0x00204081u) & 0x01010101u;
                    sb[n] = (((bv >> (4 * n)) & 0xFu) * 0x00204081u) & 0x01010101u;
                }
#pragma unroll
                for (uint32_t c = 0; c < 2; c++) {
                    const uint32_t chunk = (h * 2 + wd) * 2 + c;            // 16-byte chunk index inside the 128-byte row
                    const uint32_t off = row_off + ((chunk ^ sw) << 4);
                    *reinterpret_cast<uint4*>(stage + P * kTileBytes + off) = make_uint4(sb[4 * c], sb[4 * c + 1], sb[4 * c + 2], sb[4 * c + 3]);
#pragma unroll
                    for (int p = 0; p < P; p++) {
                        const uint32_t* w4 = wk + p * 32 + wd * 8 + c * 4;
                        *reinterpret_cast<uint4*>(stage + p * kTileBytes + off) =
                            make_uint4((sa[4 * c] * 0xFFu) & w4[0], (sa[4 * c + 1] * 0xFFu) & w4[1],
                                       (sa[4 * c + 2] * 0xFFu) & w4[2], (sa[4 * c + 3] * 0xFFu) & w4[3]);
                    }
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to the tensor core
            __syncthreads();
            if (tid == 0) {
                tc_fence_after();
                const uint32_t sbase = smem_u32(stage);
                const uint64_t bdesc = umma_desc_sw128(sbase + P * kTileBytes);
#pragma unroll
                for (int p = 0; p < P; p++) {
                    const uint64_t adesc = umma_desc_sw128(sbase + p * kTileBytes);
#pragma unroll
                    for (uint32_t k = 0; k < 4; k++)  // 4 x K=32 inside the swizzle atom: +32 B = +2 in the address field
                        umma_i8(tmem + p * kTile, adesc + 2 * k, bdesc + 2 * k, (kb > kb0 || k > 0) ? 1u : 0u);
                }
                umma_commit(smem_u32(bars + st));
                if (kb + 1 == kb1) umma_commit(bar_done);
            }
        }
        // ---- epilogue
        mbar_wait(bar_done, done_phase);
        done_phase ^= 1u;
        tc_fence_after();
        uint32_t* tp = reinterpret_cast<uint32_t*>(smem) + warp * (32 * 33);  // stage memory is idle now
        const uint32_t lane_base = 32u * (warp & 3u);
        const uint32_t col_base = 64u * (warp >> 2);
#pragma unroll 1
        for (uint32_t cc = 0; cc < 2; cc++) {
            unsigned long long acc[32];
#pragma unroll
            for (uint32_t x = 0; x < 32; x++) acc[x] = 0ull;
#pragma unroll
            for (int p = 0; p < P; p++) {
                uint32_t v[32];
                tmem_ld32(tmem + (lane_base << 16) + p * kTile + col_base + 32 * cc, v);
                __syncwarp();
#pragma unroll
                for (uint32_t x = 0; x < 32; x++) tp[lane * 33 + x] = v[x];   // row = lane, column = x
                __syncwarp();
#pragma unroll
                for (uint32_t x = 0; x < 32; x++) acc[x] += (unsigned long long)tp[x * 33 + lane] << (8 * p);  // row = x, column = lane
            }
            const uint32_t j = j0 + col_base + 32 * cc + lane;
#pragma unroll
            for (uint32_t x = 0; x < 32; x++) {
                const uint32_t i = i0 + lane_base + x;
                if (i < j && j < G.n && acc[x] != 0ull) {
                    const unsigned long long ra = gram_row_addr(i, G);
                    if (ra != 0ull) asm volatile("red.global.add.u64 [%0], %1;" ::"l"(ra + 8ull * j), "l"(acc[x]) : "memory");
                }
            }
        }
        tc_fence_before();
        __syncthreads();
    }
    if (warp == 0) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(kCols) : "memory");
    }
}

size_t gram_smem_bytes(int planes) {
    return 1024 + kStages * (size_t)(planes + 1) * kTileBytes + 2 * planes * 128 + 8 * (kStages + 1) + 16;
}

template <int P>
static cudaError_t launch_gram_p(const GramParams& prm, const TriGeom& g, const LaunchCfg& lc) {
    const size_t smem = gram_smem_bytes(P);
    cudaError_t e = cudaFuncSetAttribute(k4_gram<P>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    unsigned long long ctas = prm.n_items < (unsigned long long)lc.sms ? prm.n_items : (unsigned long long)lc.sms;
    k4_gram<P><<<(unsigned)ctas, kThreads, smem, lc.stream>>>(prm, g);
    return cudaGetLastError();
}

cudaError_t launch_k4_gram(const GramParams& prm, const TriGeom& g, const LaunchCfg& lc) {
    if (prm.n_items == 0) return cudaSuccess;
    switch (prm.planes) {
        case 1: return launch_gram_p<1>(prm, g, lc);
        case 2: return launch_gram_p<2>(prm, g, lc);
        case 3:
        case 4: return launch_gram_p<4>(prm, g, lc);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace ksb
