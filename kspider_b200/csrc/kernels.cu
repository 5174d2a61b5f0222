// sm_100a kernels of the kSpider pairwise hot path (see kernels.cuh for the reference lines each one replaces).
//
// Common idea of the pair-emission kernels: member lists are sorted, so the smaller member of a
// pair (the triangle ROW) is the one with the smaller position.  Every lane/thread keeps ONE member
// as its COLUMN in a register; rows are broadcast (warp shuffle in K1, shared memory in K2) as a
// ready-made 64-bit "row address" (address of the virtual cell (row, 0) of the owned block, or 0
// when this GPU does not own the row), so one update is: add 8*column, red.global.add.u64.
// Updates of one step land in ONE row of the triangle, in ascending address order.
#include "kernels.cuh"
#include "ddpow.hpp"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include <cub/device/device_scan.cuh>

namespace ksb {

#define KSB_FULL 0xffffffffu

__device__ __forceinline__ unsigned long long row_addr(uint32_t x, const TriGeom& g) {
    if (x - g.rb >= g.nrows) return 0ull;
    return (unsigned long long)g.tri + 8ull * (__ldg(g.rowbase + x) - g.base);
}

__device__ __forceinline__ void red_add_u64(unsigned long long addr, unsigned long long w) {
    asm volatile("red.global.add.u64 [%0], %1;" ::"l"(addr), "l"(w) : "memory");
}

__device__ __forceinline__ unsigned long long shfl_u64(unsigned long long v, uint32_t src) {
    const uint32_t lo = __shfl_sync(KSB_FULL, (uint32_t)v, src);
    const uint32_t hi = __shfl_sync(KSB_FULL, (uint32_t)(v >> 32), src);
    return ((unsigned long long)hi << 32) | lo;
}

// Where an update goes.  Dense: `row` is the address of the virtual cell (row, 0) of the owned block
// and the update is one red.  Spill (K3): `row` is (row index + 1) and the update is appended as
// (key = row_index << shift | column, weight) to a chunk buffer through a warp-aggregated cursor;
// every lane of the warp must call add() (the predicate says whether it has an update).
struct DenseSink {
    static __device__ __forceinline__ unsigned long long row(uint32_t x, const TriGeom& g) { return row_addr(x, g); }
    static __device__ __forceinline__ unsigned long long col(uint32_t x) { return 8ull * x; }
    __device__ __forceinline__ void add(bool pred, unsigned long long r, unsigned long long c, unsigned long long w) const {
        if (pred) red_add_u64(r + c, w);
    }
};
// the same on u32 cells (TriGeom::cell_bits == 32: no cell can exceed 2^32 - 1, so no carry is ever lost)
struct DenseSink32 {
    static __device__ __forceinline__ unsigned long long row(uint32_t x, const TriGeom& g) {
        if (x - g.rb >= g.nrows) return 0ull;
        return (unsigned long long)g.tri + 4ull * (__ldg(g.rowbase + x) - g.base);
    }
    static __device__ __forceinline__ unsigned long long col(uint32_t x) { return 4ull * x; }
    __device__ __forceinline__ void add(bool pred, unsigned long long r, unsigned long long c, unsigned long long w) const {
        if (pred) asm volatile("red.global.add.u32 [%0], %1;" ::"l"(r + c), "r"((uint32_t)w) : "memory");
    }
};
struct SpillSink {
    unsigned long long* keys;
    uint32_t* vals;
    unsigned long long* cursor;
    uint32_t shift;
    static __device__ __forceinline__ unsigned long long row(uint32_t x, const TriGeom& g) {
        return (x - g.rb >= g.nrows) ? 0ull : (unsigned long long)x + 1ull;
    }
    static __device__ __forceinline__ unsigned long long col(uint32_t x) { return x; }
    __device__ __forceinline__ void add(bool pred, unsigned long long r, unsigned long long c, unsigned long long w) const {
        const uint32_t m = __ballot_sync(KSB_FULL, pred);
        if (m == 0u) return;
        const uint32_t lane = threadIdx.x & 31u;
        unsigned long long base = 0;
        if (lane == (uint32_t)(__ffs(m) - 1)) base = atomicAdd(cursor, (unsigned long long)__popc(m));
        base = shfl_u64(base, __ffs(m) - 1);
        if (pred) {
            const unsigned long long o = base + __popc(m & ((1u << lane) - 1u));
            keys[o] = ((r - 1ull) << shift) | c;
            vals[o] = (uint32_t)w;
        }
    }
};

// ------------------------------------------------------------------------------------------------
// K1: one warp per small colour (2 <= s <= 64).  Colours of one size s are a dense [n x s] block, so
// a warp task loads 32/s whole colours (s <= 32) or one colour in two registers (s > 32) with one
// coalesced load, then walks the rows: step a broadcasts row a of every group by shuffle and the
// lanes holding a later member add the colour's weight into cell (row, own member).
// ------------------------------------------------------------------------------------------------
template <class Sink, class MT>
__global__ void __launch_bounds__(256) k1_small(const __grid_constant__ K1Params P,
                                                const MT* __restrict__ members,
                                                const void* __restrict__ weights,
                                                const __grid_constant__ TriGeom G, const Sink sink,
                                                unsigned long long task_begin, unsigned long long task_end) {
    const uint32_t lane = threadIdx.x & 31u;
    const unsigned long long warp = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const unsigned long long nwarps = ((unsigned long long)gridDim.x * blockDim.x) >> 5;
    for (unsigned long long task = task_begin + warp; task < task_end; task += nwarps) {
        uint32_t lo = 0, hi = P.n_cls;
        while (hi - lo > 1) {
            const uint32_t mid = (lo + hi) >> 1;
            if (P.cls[mid].task_begin <= task) lo = mid; else hi = mid;
        }
        const unsigned long long t_in = task - P.cls[lo].task_begin;
        const unsigned long long n = P.cls[lo].n;
        const uint32_t s = P.cls[lo].s;
        const unsigned long long moff = P.cls[lo].member_off, woff = P.cls[lo].weight_off;
        if (s <= 32) {
            const uint32_t k = P.cls[lo].k;
            const uint32_t g = lane / s, a = lane - g * s;
            const unsigned long long col0 = t_in * k;
            const bool valid = g < k && col0 + g < n;
            uint32_t x = 0, w = 0;
            if (valid) {
                x = ldm<MT>(members, moff + col0 * s + lane);
                w = ldw(weights, P.weight_bits, woff + col0 + g);
            }
            const unsigned long long ra = valid ? Sink::row(x, G) : 0ull;
            const uint32_t own = __ballot_sync(KSB_FULL, ra != 0ull);
            if (own == 0u) continue;  // no owned row in this task
            const unsigned long long cadd = Sink::col(x);
            const uint32_t gb = g * s;
            // members are sorted, so the owned rows of a colour are one run of positions: walk only the
            // union of those runs over the task's colours (everything when the triangle is not sharded)
            const uint32_t gm = (own >> gb) & (0xffffffffu >> (32u - s));
            const uint32_t ar0 = __reduce_min_sync(KSB_FULL, gm ? (uint32_t)__ffs(gm) - 1u : s);
            const uint32_t ar1 = min(__reduce_max_sync(KSB_FULL, gm ? 32u - (uint32_t)__clz(gm) : 0u), s - 1u);
            for (uint32_t ar = ar0; ar < ar1; ar++) {
                const unsigned long long r = shfl_u64(ra, gb + ar);
                sink.add(valid && a > ar && r != 0ull, r, cadd, w);
            }
        } else {
            const unsigned long long mi = moff + t_in * s;
            const bool v1 = lane + 32 < s;
            const uint32_t x0 = ldm<MT>(members, mi + lane);
            const uint32_t x1 = v1 ? ldm<MT>(members, mi + 32 + lane) : 0u;
            const unsigned long long w = ldw(weights, P.weight_bits, woff + t_in);
            const unsigned long long ra0 = Sink::row(x0, G);
            const unsigned long long ra1 = v1 ? Sink::row(x1, G) : 0ull;
            if (__ballot_sync(KSB_FULL, (ra0 | ra1) != 0ull) == 0u) continue;
            const unsigned long long c0 = Sink::col(x0), c1 = Sink::col(x1);
#pragma unroll 4
            for (uint32_t ar = 0; ar < 32; ar++) {
                const unsigned long long r = shfl_u64(ra0, ar);
                if (r != 0ull) {  // warp-uniform
                    sink.add(lane > ar, r, c0, w);
                    sink.add(v1, r, c1, w);
                }
            }
            for (uint32_t ar = 32; ar + 1 < s; ar++) {
                const unsigned long long r = shfl_u64(ra1, ar - 32);
                sink.add(r != 0ull && v1 && lane + 32 > ar, r, c1, w);
            }
        }
    }
}

// K1 with 32-bit cell arithmetic (dense sink, owned block < 2^32 cells): a row travels through the
// shuffle as ONE 32-bit word, the virtual cell index of (row, 0) relative to the block -- wrapping,
// because rowbase = bandstart - band_lo may be "negative"; (row word + column) mod 2^32 is the real
// cell index -- and which rows are owned comes from the ballot instead of a 64-bit compare.  The red is
// predicated, not branched around.  About half the instructions per update of the generic kernel:
// alone the kernel is bound by L2's red rate either way, but beside K4 the issue slots it no longer
// takes are K4's.
// tri[cell] += w on the lanes with pred != 0; address arithmetic and predicate stay inside one asm block
// so that the red is predicated instead of branched around
template <bool C32>
__device__ __forceinline__ void red_cell_if(unsigned long long tri, uint32_t cell, unsigned long long w, uint32_t pred) {
    if (C32)
        asm volatile(
            "{\n\t.reg .pred p;\n\t.reg .u64 a;\n\tsetp.ne.u32 p, %3, 0;\n\tmad.wide.u32 a, %1, 4, %0;\n\t@p red.global.add.u32 [a], %2;\n\t}" ::"l"(tri),
            "r"(cell), "r"((uint32_t)w), "r"(pred)
            : "memory");
    else
        asm volatile(
            "{\n\t.reg .pred p;\n\t.reg .u64 a;\n\tsetp.ne.u32 p, %3, 0;\n\tmad.wide.u32 a, %1, 8, %0;\n\t@p red.global.add.u64 [a], %2;\n\t}" ::"l"(tri),
            "r"(cell), "l"(w), "r"(pred)
            : "memory");
}

template <class MT, bool C32 = false>
__global__ void __launch_bounds__(256, 6) k1_small_c32(const __grid_constant__ K1Params P,
                                                    const MT* __restrict__ members,
                                                    const void* __restrict__ weights,
                                                    const __grid_constant__ TriGeom G,
                                                    unsigned long long task_begin, unsigned long long task_end) {
    const uint32_t lane = threadIdx.x & 31u;
    const unsigned long long warp = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const unsigned long long nwarps = ((unsigned long long)gridDim.x * blockDim.x) >> 5;
    const unsigned long long tri = (unsigned long long)G.tri;
    unsigned long long task = task_begin + warp;
    if (task >= task_end) return;
    // The class of a task is looked up once and then followed: a warp's tasks only move forward through
    // the class list.  Per class the lane's group / position and the validity of its group are fixed.
    uint32_t ci = 0;
    {
        uint32_t hi = P.n_cls;
        while (hi - ci > 1) {
            const uint32_t mid = (ci + hi) >> 1;
            if (P.cls[mid].task_begin <= task) ci = mid; else hi = mid;
        }
    }
    unsigned long long c_end = 0;  // first task of the next class; 0 forces the first load below
    unsigned long long c_begin = 0, n = 0, moff = 0, woff = 0;
    uint32_t s = 2, k = 1, g = 0, a = 0, gb = 0, smask = 0, below = 0;
    for (; task < task_end; task += nwarps) {
        if (task >= c_end) {
            while (ci + 1 < P.n_cls && P.cls[ci + 1].task_begin <= task) ci++;
            const K1Class& c = P.cls[ci];
            c_begin = c.task_begin;
            c_end = ci + 1 < P.n_cls ? P.cls[ci + 1].task_begin : ~0ull;
            n = c.n; moff = c.member_off; woff = c.weight_off; s = c.s; k = c.k;
            g = (lane * (65536u / s + 1u)) >> 16;  // = lane / s for lane < 32, s <= 32 (s > 32: unused)
            a = lane - g * s;
            gb = g * s;
            smask = 0xffffffffu >> (32u - min(s, 32u));
            below = (1u << a) - 1u;
        }
        const unsigned long long t_in = task - c_begin;
        if (s <= 32) {
            const unsigned long long col0 = t_in * k;
            const bool valid = g < k && col0 + g < n;
            uint32_t x = 0, w = 0;
            if (valid) {
                x = ldm<MT>(members, moff + col0 * s + lane);
                w = ldw(weights, P.weight_bits, woff + col0 + g);
            }
            const bool owned = valid && (x - G.rb < G.nrows);
            const uint32_t r32 = owned ? (uint32_t)(__ldg(G.rowbase + x) - G.base) : 0u;
            const uint32_t own = __ballot_sync(KSB_FULL, owned);
            if (own == 0u) continue;  // no owned row in this task
            const uint32_t gm = valid ? (own >> gb) & smask : 0u;  // owned positions of this lane's colour
            const uint32_t ar0 = __reduce_min_sync(KSB_FULL, gm ? (uint32_t)__ffs(gm) - 1u : s);
            const uint32_t ar1 = min(__reduce_max_sync(KSB_FULL, gm ? 32u - (uint32_t)__clz(gm) : 0u), s - 1u);
            const uint32_t act = gm & below;  // steps (row positions) in which this lane adds: owned rows before its own position
            const uint32_t sb = valid ? gb : 0u;
            const unsigned long long ww = w;
            // four rows per trip: the four shuffles are issued while the warp is converged, then the four
            // (branched-around) reds; positions past the colour's end have no bit in `act`
            for (uint32_t ar = ar0; ar < ar1; ar += 4) {
                const uint32_t src = sb + ar;
                const uint32_t q0 = __shfl_sync(KSB_FULL, r32, src);
                const uint32_t q1 = __shfl_sync(KSB_FULL, r32, src + 1u);
                const uint32_t q2 = __shfl_sync(KSB_FULL, r32, src + 2u);
                const uint32_t q3 = __shfl_sync(KSB_FULL, r32, src + 3u);
                const uint32_t m = act >> ar;  // ar < ar1 <= 31
                red_cell_if<C32>(tri, q0 + x, ww, m & 1u);
                red_cell_if<C32>(tri, q1 + x, ww, m & 2u);
                red_cell_if<C32>(tri, q2 + x, ww, m & 4u);
                red_cell_if<C32>(tri, q3 + x, ww, m & 8u);
            }
        } else {
            const unsigned long long mi = moff + t_in * s;
            const bool v1 = lane + 32 < s;
            const uint32_t x0 = ldm<MT>(members, mi + lane);
            const uint32_t x1 = v1 ? ldm<MT>(members, mi + 32 + lane) : 0u;
            const unsigned long long w = ldw(weights, P.weight_bits, woff + t_in);
            const bool o0 = x0 - G.rb < G.nrows, o1 = v1 && (x1 - G.rb < G.nrows);
            const uint32_t r0 = o0 ? (uint32_t)(__ldg(G.rowbase + x0) - G.base) : 0u;
            const uint32_t r1 = o1 ? (uint32_t)(__ldg(G.rowbase + x1) - G.base) : 0u;
            const uint32_t own0 = __ballot_sync(KSB_FULL, o0), own1 = __ballot_sync(KSB_FULL, o1);
            if ((own0 | own1) == 0u) continue;
#pragma unroll 4
            for (uint32_t ar = 0; ar < 32; ar++) {
                const uint32_t r = __shfl_sync(KSB_FULL, r0, ar);
                if ((own0 >> ar) & 1u) {  // warp-uniform
                    red_cell_if<C32>(tri, r + x0, w, lane > ar);
                    red_cell_if<C32>(tri, r + x1, w, v1);
                }
            }
            for (uint32_t ar = 32; ar + 1 < s; ar++) {
                const uint32_t r = __shfl_sync(KSB_FULL, r1, ar - 32);
                red_cell_if<C32>(tri, r + x1, w, ((own1 >> (ar - 32)) & 1u) != 0u && v1 && lane + 32 > ar);
            }
        }
    }
}

cudaError_t launch_k1_small(const K1Params& p, const void* members, uint32_t member_bits, const void* weights, const TriGeom& g,
                            const LaunchCfg& lc, unsigned long long task_begin, unsigned long long task_end) {
    if (task_end <= task_begin) return cudaSuccess;
    const unsigned long long warps_per_cta = 8;
    unsigned long long ctas = (task_end - task_begin + warps_per_cta - 1) / warps_per_cta;
    const unsigned long long cap = (unsigned long long)lc.sms * (unsigned long long)lc.k1_ctas;  // 8 x 256 threads = 2048 resident threads / SM
    if (ctas > cap) ctas = cap;
    // 32-bit cell arithmetic whenever the owned block allows it (KSPIDER_K1_CELLS=64 forces the generic kernel)
    const char* k1c = std::getenv("KSPIDER_K1_CELLS");
    const bool force64 = k1c && std::atoi(k1c) == 64;
    const bool c32 = !force64 && g.cells < (1ull << 32);
    const int carve = lc.beside_gram ? (int)cudaSharedmemCarveoutMaxShared : (int)cudaSharedmemCarveoutDefault;
    auto go = [&](auto kern64, auto kern32, auto sink, auto mem) {
        if (c32) {
            cudaFuncSetAttribute(kern32, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
            const unsigned long long cap32 = (unsigned long long)lc.sms * (unsigned long long)(lc.k1_ctas < 6 ? lc.k1_ctas : 6);  // 40 registers: 6 resident CTAs / SM
            if (ctas > cap32) ctas = cap32;  // persistent grid: every CTA must be resident from the start (tasks are strided statically)
            kern32<<<(unsigned)ctas, 256, 0, lc.stream>>>(p, mem, weights, g, task_begin, task_end);
        } else {
            cudaFuncSetAttribute(kern64, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
            kern64<<<(unsigned)ctas, 256, 0, lc.stream>>>(p, mem, weights, g, sink, task_begin, task_end);
        }
    };
    if (g.cell_bits == 32) {  // u32 cells
        if (member_bits == 10) go(k1_small<DenseSink32, P10>, k1_small_c32<P10, true>, DenseSink32{}, static_cast<const P10*>(members));
        else if (member_bits == 12) go(k1_small<DenseSink32, P12>, k1_small_c32<P12, true>, DenseSink32{}, static_cast<const P12*>(members));
        else if (member_bits == 16) go(k1_small<DenseSink32, uint16_t>, k1_small_c32<uint16_t, true>, DenseSink32{}, static_cast<const uint16_t*>(members));
        else go(k1_small<DenseSink32, uint32_t>, k1_small_c32<uint32_t, true>, DenseSink32{}, static_cast<const uint32_t*>(members));
        return cudaGetLastError();
    }
    if (member_bits == 10) go(k1_small<DenseSink, P10>, k1_small_c32<P10>, DenseSink{}, static_cast<const P10*>(members));
    else if (member_bits == 12) go(k1_small<DenseSink, P12>, k1_small_c32<P12>, DenseSink{}, static_cast<const P12*>(members));
    else if (member_bits == 16) go(k1_small<DenseSink, uint16_t>, k1_small_c32<uint16_t>, DenseSink{}, static_cast<const uint16_t*>(members));
    else go(k1_small<DenseSink, uint32_t>, k1_small_c32<uint32_t>, DenseSink{}, static_cast<const uint32_t*>(members));
    return cudaGetLastError();
}

cudaError_t launch_k1_spill(const K1Params& p, const void* members, uint32_t member_bits, const void* weights, const TriGeom& g,
                            const SpillBuf& sb, unsigned long long task_begin, unsigned long long task_end,
                            const LaunchCfg& lc) {
    if (task_end <= task_begin) return cudaSuccess;
    unsigned long long ctas = (task_end - task_begin + 7) / 8;
    const unsigned long long cap = (unsigned long long)lc.sms * 8ull;
    if (ctas > cap) ctas = cap;
    const SpillSink sink{sb.keys, sb.vals, sb.cursor, sb.shift};
    if (member_bits == 10)
        k1_small<SpillSink, P10><<<(unsigned)ctas, 256, 0, lc.stream>>>(p, static_cast<const P10*>(members), weights, g, sink, task_begin, task_end);
    else if (member_bits == 12)
        k1_small<SpillSink, P12><<<(unsigned)ctas, 256, 0, lc.stream>>>(p, static_cast<const P12*>(members), weights, g, sink, task_begin, task_end);
    else if (member_bits == 16)
        k1_small<SpillSink, uint16_t><<<(unsigned)ctas, 256, 0, lc.stream>>>(p, static_cast<const uint16_t*>(members), weights, g, sink, task_begin, task_end);
    else
        k1_small<SpillSink, uint32_t><<<(unsigned)ctas, 256, 0, lc.stream>>>(p, static_cast<const uint32_t*>(members), weights, g, sink, task_begin, task_end);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// K8: ANI column (pykSpider/kSpider2/ks_pairwise.py:29-82, `pairwise --estimate-ani`).  Per TSV row the
// reference parses the PRINTED min / max containment (6 significant digits of the float32) back to a
// double, runs sourmash.distance_utils.containment_to_distance(c, ksize, scale, n_unique_kmers).ani on
// each -- c == 0 -> 0.0, c == 1 -> 1.0, else 1 - (1 - c ** (1.0 / ksize)); a distance outside [0, 1]
// raises -- and writes their mean.  The text round trip is reproduced exactly (text6 below); the power is
// a double-double pow (ddpow.hpp, correctly rounded in practice) where CPython calls glibc's (< 1 ulp, not
// always correctly rounded), so ~99.9 % of the values are bit-identical and the rest differ by one ulp
// (tests state the tolerance).
// ------------------------------------------------------------------------------------------------
__constant__ double kPow10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                  1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};

// strtod(printf("%g", c)) for a finite float 0 <= c < 10: the six significant decimal digits of the
// exact binary value, rounded half-to-even on the exact remainder as glibc does (x * 10^k is split
// exactly into p + err by an fma), then digits / 10^k in ONE correctly rounded division (digits < 2^53
// and 10^k exact for k <= 22: the classic exact strtod path).  NaN for anything else.
__device__ double text6(float c) {
    if (c == 0.0f) return 0.0;
    if (!(c > 0.0f) || !(c < 10.0f)) return __longlong_as_double(0x7ff8000000000000ll);
    const double x = (double)c;
    int e10 = (int)floor(log10(x));
    double P = 1.0, p = 0.0, err = 0.0;
    for (int it = 0; it < 4; it++) {
        const int k = 5 - e10;
        if (k < 5 || k > 22) return __longlong_as_double(0x7ff8000000000000ll);
        P = kPow10[k];
        p = __dmul_rn(x, P);
        err = __fma_rn(x, P, -p);
        if (p < 1e5 || (p == 1e5 && err < 0.0)) { e10--; continue; }
        if (p > 1e6 || (p == 1e6 && err >= 0.0)) { e10++; continue; }
        break;
    }
    const double ip = floor(p);
    const double t = __dsub_rn(__dsub_rn(p, ip), 0.5);  // exact: p < 2^20, so the fraction is a multiple of ulp(p) >= 2^-36
    // |err| <= ulp(p)/2 < |t| unless t == 0: the sign of (t + err) is t's, or err's on a tie of p
    const bool up = t > 0.0 || (t == 0.0 && (err > 0.0 || (err == 0.0 && (((long long)ip) & 1ll))));
    return __ddiv_rn(up ? ip + 1.0 : ip, P);
}

__device__ __forceinline__ double ani_of(double c, double inv_k, bool* bad) {
    if (c == 0.0) return 0.0;   // 1 - 1.0
    if (c == 1.0) return 1.0;   // 1 - 0.0
    // c ** (1.0 / ksize): ksize 1 -> c exactly; otherwise the double-double pow of ddpow.hpp (correctly rounded but for
    // ~2^-45 of the cases; CUDA's pow() is <= 2 ulp, CPython's is glibc's < 1 ulp)
    const double pw = inv_k == 1.0 ? c : ((c > 0.0 && c < 1.0) ? dd_pow(c, inv_k) : pow(c, inv_k));
    const double dist = __dsub_rn(1.0, pw);
    if (!(dist >= 0.0 && dist <= 1.0)) { *bad = true; return dist; }  // sourmash: ValueError (also NaN from text6)
    return __dsub_rn(1.0, dist);
}

__global__ void __launch_bounds__(256) k8_ani(const ksb_record* __restrict__ records, unsigned long long n_rows, double inv_k,
                                              double* __restrict__ out, unsigned long long* n_bad) {
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    uint32_t nb = 0;
    for (unsigned long long r = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += stride) {
        const float cmin = __ldg(&records[r].min_containment), cmax = __ldg(&records[r].max_containment);
        bool bad = false;
        const double a1 = ani_of(text6(cmin), inv_k, &bad);
        const double a2 = ani_of(text6(cmax), inv_k, &bad);
        out[r] = __ddiv_rn(__dadd_rn(a1, a2), 2.0);
        nb += bad ? 1u : 0u;
    }
    nb = __reduce_add_sync(KSB_FULL, nb);
    if ((threadIdx.x & 31u) == 0 && nb) atomicAdd(n_bad, (unsigned long long)nb);
}

cudaError_t launch_k8_ani(const ksb_record* records, uint64_t n_rows, double inv_k, double* out, unsigned long long* n_bad,
                          const LaunchCfg& lc) {
    if (n_rows == 0) return cudaSuccess;
    const uint64_t want = (n_rows + 255) / 256, cap = (uint64_t)lc.sms * 8;
    k8_ani<<<(unsigned)(want < cap ? want : cap), 256, 0, lc.stream>>>(records, n_rows, inv_k, out, n_bad);
    return cudaGetLastError();
}

// The same column from two parsed columns of the TSV on disk (ksb_ani_table): doubles = float(text) as the
// reference parses them (no text6 needed), or float32 from the binary side-car (text6 recovers the text value).
__global__ void __launch_bounds__(256) k8_ani_cols(const double* __restrict__ dmin, const double* __restrict__ dmax,
                                                   const float* __restrict__ fmin, const float* __restrict__ fmax,
                                                   unsigned long long n_rows, double inv_k, double* __restrict__ out,
                                                   unsigned long long* n_bad) {
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    uint32_t nb = 0;
    for (unsigned long long r = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += stride) {
        const double c1 = dmin ? __ldg(dmin + r) : text6(__ldg(fmin + r));
        const double c2 = dmax ? __ldg(dmax + r) : text6(__ldg(fmax + r));
        bool bad = false;
        const double a1 = ani_of(c1, inv_k, &bad);
        const double a2 = ani_of(c2, inv_k, &bad);
        out[r] = __ddiv_rn(__dadd_rn(a1, a2), 2.0);
        nb += bad ? 1u : 0u;
    }
    nb = __reduce_add_sync(KSB_FULL, nb);
    if ((threadIdx.x & 31u) == 0 && nb) atomicAdd(n_bad, (unsigned long long)nb);
}

cudaError_t launch_k8_ani_cols(const double* dmin, const double* dmax, const float* fmin, const float* fmax, uint64_t n_rows,
                               double inv_k, double* out, unsigned long long* n_bad, const LaunchCfg& lc) {
    if (n_rows == 0) return cudaSuccess;
    const uint64_t want = (n_rows + 255) / 256, cap = (uint64_t)lc.sms * 8;
    k8_ani_cols<<<(unsigned)(want < cap ? want : cap), 256, 0, lc.stream>>>(dmin, dmax, fmin, fmax, n_rows, inv_k, out, n_bad);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// K2: block-tiled mid/large colours (s > 64).  One CTA item = 256 columns (one per thread, kept in
// a register) x up to 512 rows.  The rows' member ids are staged into shared memory with one TMA
// bulk copy (cp.async.bulk + mbarrier), converted once to row addresses, then every warp sweeps the
// rows: one shared-memory broadcast + one red per update.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_1d(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
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

template <class Sink, class MT>
__global__ void __launch_bounds__(kMidCols) k2_mid(const MidItem* __restrict__ items, unsigned long long item_begin,
                                                   unsigned long long n_items,
                                                   const unsigned long long* __restrict__ mid_off,
                                                   const uint32_t* __restrict__ mid_sizes,
                                                   const MT* __restrict__ mid_members,
                                                   const uint32_t* __restrict__ mid_weights,
                                                   const __grid_constant__ TriGeom G, const Sink sink) {
    __shared__ __align__(128) MT sm_rows[kMidRows];
    __shared__ __align__(16) unsigned long long sm_ra[kMidRows];
    __shared__ __align__(8) unsigned long long sm_bar;
    const uint32_t tid = threadIdx.x;
    const uint32_t bar = smem_u32(&sm_bar);
    if (tid == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    uint32_t phase = 0;
    const bool sharded = G.nrows < G.n;
    for (unsigned long long it = item_begin + blockIdx.x; it < n_items; it += gridDim.x) {
        const MidItem item = items[it];
        const uint32_t s = __ldg(mid_sizes + item.color);
        const MT* m = mid_members + __ldg(mid_off + item.color);
        const unsigned long long w = __ldg(mid_weights + item.color);
        const uint32_t row_lim = min(s - 1u, item.b0 + kMidCols - 1u);
        const uint32_t a_end = min(item.a0 + kMidRows, row_lim);
        const uint32_t nr = a_end - item.a0;
        if (sharded) {  // rows are sorted ids: reject the item when none can be owned (uniform)
            const uint32_t first = __ldg(m + item.a0), last = __ldg(m + a_end - 1);
            if (last < G.rb || first - G.rb >= G.nrows && first >= G.rb) continue;
        }
        if (tid == 0) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            const uint32_t bytes = (uint32_t)sizeof(MT) * ((nr + 7u) & ~7u);  // colours are padded to 8 members
            mbar_expect_tx(bar, bytes);
            tma_load_1d(smem_u32(sm_rows), m + item.a0, bytes, bar);
        }
        const uint32_t b = item.b0 + tid;
        const uint32_t x = (b < s) ? __ldg(m + b) : 0u;
        mbar_wait(bar, phase);
        phase ^= 1u;
        for (uint32_t r = tid; r < nr; r += kMidCols) sm_ra[r] = Sink::row(sm_rows[r], G);
        __syncthreads();
        // rows r with a0 + r < b pair with this thread's column
        const uint32_t lim = (b < s && b > item.a0) ? min(nr, b - item.a0) : 0u;
        const uint32_t wlim = __reduce_max_sync(KSB_FULL, lim);
        const unsigned long long cadd = Sink::col(x);
#pragma unroll 4
        for (uint32_t r = 0; r < wlim; r++) {
            const unsigned long long ra = sm_ra[r];
            sink.add(ra != 0ull && r < lim, ra, cadd, w);
        }
        __syncthreads();
    }
}

cudaError_t launch_k2_mid(const MidItem* items, uint64_t n_items, const unsigned long long* mid_off,
                          const uint32_t* mid_sizes, const void* mid_members, uint32_t member_bits, const uint32_t* mid_weights,
                          const TriGeom& g, const LaunchCfg& lc) {
    if (n_items == 0) return cudaSuccess;
    unsigned long long ctas = n_items;
    const unsigned long long cap = (unsigned long long)lc.sms * 8ull;
    if (ctas > cap) ctas = cap;
    if (g.cell_bits == 32) {
        if (member_bits == 16)
            k2_mid<DenseSink32, uint16_t><<<(unsigned)ctas, kMidCols, 0, lc.stream>>>(items, 0ull, n_items, mid_off, mid_sizes, static_cast<const uint16_t*>(mid_members), mid_weights, g, DenseSink32{});
        else
            k2_mid<DenseSink32, uint32_t><<<(unsigned)ctas, kMidCols, 0, lc.stream>>>(items, 0ull, n_items, mid_off, mid_sizes, static_cast<const uint32_t*>(mid_members), mid_weights, g, DenseSink32{});
        return cudaGetLastError();
    }
    if (member_bits == 16)
        k2_mid<DenseSink, uint16_t><<<(unsigned)ctas, kMidCols, 0, lc.stream>>>(items, 0ull, n_items, mid_off, mid_sizes, static_cast<const uint16_t*>(mid_members), mid_weights, g, DenseSink{});
    else
        k2_mid<DenseSink, uint32_t><<<(unsigned)ctas, kMidCols, 0, lc.stream>>>(items, 0ull, n_items, mid_off, mid_sizes, static_cast<const uint32_t*>(mid_members), mid_weights, g, DenseSink{});
    return cudaGetLastError();
}

cudaError_t launch_k2_spill(const MidItem* items, uint64_t item_begin, uint64_t item_end, const unsigned long long* mid_off,
                            const uint32_t* mid_sizes, const void* mid_members, uint32_t member_bits, const uint32_t* mid_weights,
                            const TriGeom& g, const SpillBuf& sb, const LaunchCfg& lc) {
    if (item_end <= item_begin) return cudaSuccess;
    unsigned long long ctas = item_end - item_begin;
    const unsigned long long cap = (unsigned long long)lc.sms * 8ull;
    if (ctas > cap) ctas = cap;
    const SpillSink sink{sb.keys, sb.vals, sb.cursor, sb.shift};
    if (member_bits == 16)
        k2_mid<SpillSink, uint16_t><<<(unsigned)ctas, kMidCols, 0, lc.stream>>>(items, item_begin, item_end, mid_off, mid_sizes, static_cast<const uint16_t*>(mid_members), mid_weights, g, sink);
    else
        k2_mid<SpillSink, uint32_t><<<(unsigned)ctas, kMidCols, 0, lc.stream>>>(items, item_begin, item_end, mid_off, mid_sizes, static_cast<const uint32_t*>(mid_members), mid_weights, g, sink);
    return cudaGetLastError();
}

// Generic warp-per-colour emitter over a plain CSR section (Gram-routed and zero-weight colours in
// spill mode; a zero weight appends (key, 0), which is exactly the reference's insert-with-0).
template <class MT>
__global__ void __launch_bounds__(256) k3_emit_csr(unsigned long long c_begin, unsigned long long c_end,
                                                   const unsigned long long* __restrict__ off,
                                                   const MT* __restrict__ members,
                                                   const void* __restrict__ weights, uint32_t weight_bits,  // nullptr = weight 0
                                                   const __grid_constant__ TriGeom G, const SpillSink sink,
                                                   const GramWin* __restrict__ wins, uint32_t n_wins) {  // non-null: members are offsets inside the colour's window
    const uint32_t lane = threadIdx.x & 31u;
    const unsigned long long warp = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const unsigned long long nwarps = ((unsigned long long)gridDim.x * blockDim.x) >> 5;
    for (unsigned long long c = c_begin + warp; c < c_end; c += nwarps) {
        const unsigned long long o = off[c];
        const uint32_t s = (uint32_t)(off[c + 1] - o);
        const unsigned long long w = weights ? ldw(weights, weight_bits, c) : 0ull;
        uint32_t x0 = 0;
        if (wins) {  // the window holding colour c (windows partition the colours in ascending c0): warp-uniform search
            uint32_t lo = 0, hi = n_wins;
            while (hi - lo > 1) {
                const uint32_t mid = (lo + hi) >> 1;
                if (__ldg(&wins[mid].c0) <= c) lo = mid; else hi = mid;
            }
            x0 = __ldg(&wins[lo].x0);
        }
        for (uint32_t a = 0; a + 1 < s; a++) {
            const unsigned long long r = SpillSink::row(ldm<MT>(members, o + a) + x0, G);
            if (r == 0ull) continue;  // warp-uniform
            for (uint32_t b0 = a + 1; b0 < s; b0 += 32) {
                const uint32_t b = b0 + lane;
                sink.add(b < s, r, b < s ? ldm<MT>(members, o + b) + x0 : 0u, w);
            }
        }
    }
}

// Gram colours held as bitmaps (IndexDesc::gram_form 1) in spill mode: a warp per colour walks the colour's bit
// column of its window -- 32 samples per step, members found by ballot -- and for every owned member (row) sweeps the
// rest of the column for the later members.  O(s * width / 32) warp steps per colour instead of O(s^2 / 32); only the
// fallback path pays it.
__global__ void __launch_bounds__(256) k3_emit_bitmap(unsigned long long c_begin, unsigned long long c_end,
                                                      const GramWin* __restrict__ wins, uint32_t n_wins,
                                                      const unsigned long long* __restrict__ win_rows,
                                                      const uint32_t* __restrict__ bits,
                                                      const void* __restrict__ weights, uint32_t weight_bits,
                                                      const __grid_constant__ TriGeom G, const SpillSink sink) {
    const uint32_t lane = threadIdx.x & 31u;
    const unsigned long long warp = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const unsigned long long nwarps = ((unsigned long long)gridDim.x * blockDim.x) >> 5;
    for (unsigned long long c = c_begin + warp; c < c_end; c += nwarps) {
        uint32_t lo = 0, hi = n_wins;
        while (hi - lo > 1) {
            const uint32_t mid = (lo + hi) >> 1;
            if (__ldg(&wins[mid].c0) <= c) lo = mid; else hi = mid;
        }
        const uint32_t x0 = __ldg(&wins[lo].x0), ext = __ldg(&wins[lo].extent);
        const uint32_t width = ext ? ext : __ldg(&wins[lo].width);  // rows of a K-block in the blob (gram_win_stride): the samples the window's colours touch
        const uint32_t lc = (uint32_t)(c - __ldg(&wins[lo].c0));
        // word (lc % 128) / 32 of the 16-byte row of sample x0 in the colour's K-block; the rows follow at 4 words each
        const uint32_t* col = bits + 4ull * (__ldg(win_rows + lo) + (unsigned long long)(lc / kGramKBlock) * width) + ((lc % kGramKBlock) >> 5);
        const uint32_t bit = lc & 31u;
        const unsigned long long w = ldw(weights, weight_bits, c);
        for (uint32_t a0 = 0; a0 < width; a0 += 32) {
            uint32_t rows_m = __ballot_sync(KSB_FULL, a0 + lane < width && ((__ldg(col + 4ull * (a0 + lane)) >> bit) & 1u));
            while (rows_m) {
                const uint32_t a = a0 + (uint32_t)__ffs(rows_m) - 1u;
                rows_m &= rows_m - 1u;
                const unsigned long long r = SpillSink::row(x0 + a, G);
                if (r == 0ull) continue;  // warp-uniform
                for (uint32_t b0 = a & ~31u; b0 < width; b0 += 32) {
                    const uint32_t b = b0 + lane;
                    const bool in = b > a && b < width && ((__ldg(col + 4ull * b) >> bit) & 1u);
                    sink.add(in, r, x0 + b, w);
                }
            }
        }
    }
}

cudaError_t launch_k3_emit_bitmap(uint64_t c_begin, uint64_t c_end, const GramWin* wins, uint32_t n_wins, const unsigned long long* win_rows,
                                  const uint32_t* bits, const void* weights, uint32_t weight_bits, const TriGeom& g, const SpillBuf& sb,
                                  const LaunchCfg& lc) {
    if (c_end <= c_begin) return cudaSuccess;
    unsigned long long ctas = (c_end - c_begin + 7) / 8;
    const unsigned long long cap = (unsigned long long)lc.sms * 8ull;
    if (ctas > cap) ctas = cap;
    const SpillSink sink{sb.keys, sb.vals, sb.cursor, sb.shift};
    k3_emit_bitmap<<<(unsigned)ctas, 256, 0, lc.stream>>>(c_begin, c_end, wins, n_wins, win_rows, bits, weights, weight_bits, g, sink);
    return cudaGetLastError();
}

cudaError_t launch_k3_emit_csr(uint64_t c_begin, uint64_t c_end, const unsigned long long* off, const void* members,
                               uint32_t member_bits, const void* weights, uint32_t weight_bits, const TriGeom& g, const SpillBuf& sb,
                               const LaunchCfg& lc, const GramWin* wins, uint32_t n_wins) {
    if (c_end <= c_begin) return cudaSuccess;
    unsigned long long ctas = (c_end - c_begin + 7) / 8;
    const unsigned long long cap = (unsigned long long)lc.sms * 8ull;
    if (ctas > cap) ctas = cap;
    const SpillSink sink{sb.keys, sb.vals, sb.cursor, sb.shift};
    if (member_bits == 10)
        k3_emit_csr<P10><<<(unsigned)ctas, 256, 0, lc.stream>>>(c_begin, c_end, off, static_cast<const P10*>(members), weights, weight_bits, g, sink, wins, n_wins);
    else if (member_bits == 12)
        k3_emit_csr<P12><<<(unsigned)ctas, 256, 0, lc.stream>>>(c_begin, c_end, off, static_cast<const P12*>(members), weights, weight_bits, g, sink, wins, n_wins);
    else if (member_bits == 16)
        k3_emit_csr<uint16_t><<<(unsigned)ctas, 256, 0, lc.stream>>>(c_begin, c_end, off, static_cast<const uint16_t*>(members), weights, weight_bits, g, sink, wins, n_wins);
    else
        k3_emit_csr<uint32_t><<<(unsigned)ctas, 256, 0, lc.stream>>>(c_begin, c_end, off, static_cast<const uint32_t*>(members), weights, weight_bits, g, sink, wins, n_wins);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// Kz: colours whose narrowed count is 0 still create a row in the reference (edges.try_emplace_l
// inserts the pair with value 0, src/pairwise.cpp:222-225).  They set a bit in a touched-bitmap.
// ------------------------------------------------------------------------------------------------
template <class MT>
__global__ void __launch_bounds__(256) kz_touch(unsigned long long n_zero, const unsigned long long* __restrict__ zero_off,
                                                const MT* __restrict__ zero_members, uint32_t* touched,
                                                const __grid_constant__ TriGeom G) {
    const uint32_t lane = threadIdx.x & 31u;
    const unsigned long long warp = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const unsigned long long nwarps = ((unsigned long long)gridDim.x * blockDim.x) >> 5;
    for (unsigned long long c = warp; c < n_zero; c += nwarps) {
        const unsigned long long off = zero_off[c];
        const uint32_t s = (uint32_t)(zero_off[c + 1] - off);
        const MT* m = zero_members + off;
        for (uint32_t a = 0; a + 1 < s; a++) {
            const uint32_t i = m[a];
            if (i - G.rb >= G.nrows) continue;
            const unsigned long long rowcell = __ldg(G.rowbase + i) - G.base;
            for (uint32_t b = a + 1 + lane; b < s; b += 32) {
                const unsigned long long cell = rowcell + m[b];
                atomicOr(touched + (cell >> 5), 1u << (cell & 31u));
            }
        }
    }
}

cudaError_t launch_kz_touch(uint64_t n_zero, const unsigned long long* zero_off, const void* zero_members,
                            uint32_t member_bits, uint32_t* touched_bits, const TriGeom& g, const LaunchCfg& lc) {
    if (n_zero == 0) return cudaSuccess;
    unsigned long long ctas = (n_zero + 7) / 8;
    const unsigned long long cap = (unsigned long long)lc.sms * 8ull;
    if (ctas > cap) ctas = cap;
    if (member_bits == 16)
        kz_touch<uint16_t><<<(unsigned)ctas, 256, 0, lc.stream>>>(n_zero, zero_off, static_cast<const uint16_t*>(zero_members), touched_bits, g);
    else
        kz_touch<uint32_t><<<(unsigned)ctas, 256, 0, lc.stream>>>(n_zero, zero_off, static_cast<const uint32_t*>(zero_members), touched_bits, g);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// K5: containment + record build over the owned block (two passes: count, then emit in cell order,
// i.e. rows sorted by (source_1, source_2)).
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t k5_flags(const TriGeom& G, const uint32_t* __restrict__ touched,
                                             unsigned long long cell0, unsigned long long v[kPostPerThread]) {
    uint32_t flags = 0;
    if (G.cell_bits == 32) {  // u32 cells (kernel-uniform)
        const uint32_t* t32 = reinterpret_cast<const uint32_t*>(G.tri);
        if (cell0 + kPostPerThread <= G.cells) {
            const uint4* p = reinterpret_cast<const uint4*>(t32 + cell0);
#pragma unroll
            for (uint32_t q = 0; q < kPostPerThread / 4; q++) {
                const uint4 t = p[q];
                v[4 * q] = t.x; v[4 * q + 1] = t.y; v[4 * q + 2] = t.z; v[4 * q + 3] = t.w;
            }
        } else {
#pragma unroll
            for (uint32_t q = 0; q < kPostPerThread; q++) v[q] = (cell0 + q < G.cells) ? t32[cell0 + q] : 0ull;
        }
    } else if (cell0 + kPostPerThread <= G.cells) {
        const ulonglong2* p = reinterpret_cast<const ulonglong2*>(G.tri + cell0);
#pragma unroll
        for (uint32_t q = 0; q < kPostPerThread / 2; q++) {
            const ulonglong2 t = p[q];
            v[2 * q] = t.x;
            v[2 * q + 1] = t.y;
        }
    } else {
#pragma unroll
        for (uint32_t q = 0; q < kPostPerThread; q++) v[q] = (cell0 + q < G.cells) ? G.tri[cell0 + q] : 0ull;
    }
    uint32_t tb = 0;
    if (touched != nullptr && cell0 < G.cells) tb = (touched[cell0 >> 5] >> (cell0 & 31u)) & 0xffu;  // 8 | 32
#pragma unroll
    for (uint32_t q = 0; q < kPostPerThread; q++)
        if (v[q] != 0ull || ((tb >> q) & 1u)) flags |= 1u << q;
    if (cell0 >= G.cells) flags = 0;
    else if (cell0 + kPostPerThread > G.cells) flags &= (1u << (uint32_t)(G.cells - cell0)) - 1u;
    return flags;
}

__global__ void __launch_bounds__(kPostThreads) k5_count(const __grid_constant__ TriGeom G,
                                                         const uint32_t* __restrict__ touched,
                                                         unsigned long long* __restrict__ block_counts) {
    __shared__ uint32_t sm[kPostThreads / 32];
    const unsigned long long cell0 = ((unsigned long long)blockIdx.x * kPostThreads + threadIdx.x) * kPostPerThread;
    unsigned long long v[kPostPerThread];
    uint32_t c = __popc(k5_flags(G, touched, cell0, v));
    c = __reduce_add_sync(KSB_FULL, c);
    if ((threadIdx.x & 31u) == 0) sm[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
#pragma unroll
        for (uint32_t i = 0; i < kPostThreads / 32; i++) t += sm[i];
        block_counts[blockIdx.x] = t;
    }
}

// float32 containment math of src/pairwise.cpp:260-264, IEEE round-to-nearest, no contraction.
// NaNs are canonicalised to the x86 default NaN (0xFFC00000, printed "-nan" like the reference).
__device__ __forceinline__ float canon_nan(float f) { return (f != f) ? __int_as_float(0xFFC00000) : f; }

__device__ __forceinline__ void containment(unsigned long long shared, uint32_t k1, uint32_t k2, float& cmin,
                                            float& cavg, float& cmax) {
    const float sh = __ull2float_rn(shared);
    const float c12 = canon_nan(__fdiv_rn(sh, __uint2float_rn(k2)));  // cont_1_in_2
    const float c21 = canon_nan(__fdiv_rn(sh, __uint2float_rn(k1)));  // cont_2_in_1
    cmin = (c21 < c12) ? c21 : c12;  // std::min(a,b) = (b<a) ? b : a
    cmax = (c12 < c21) ? c21 : c12;  // std::max(a,b) = (a<b) ? b : a
    const float sum = canon_nan(__fadd_rn(c12, c21));
    cavg = canon_nan(__double2float_rn(__ddiv_rn((double)sum, 2.0)));
}

// Row of the first cell of every K5 block (2048 cells) of the owned block: static geometry, computed once
// per pass.  k5_emit starts its per-lane row search from it -- between this block's row and the next
// block's, a handful of rows -- instead of running 17 dependent bandstart loads over all N rows per warp.
__global__ void __launch_bounds__(256) k5_block_rows(const __grid_constant__ TriGeom G, unsigned long long n_blocks,
                                                     uint32_t* __restrict__ block_rows) {
    const unsigned long long b = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n_blocks) return;
    const unsigned long long g0 = G.base + b * kPostChunk;
    uint32_t lo = G.rb, hi = G.rb + G.nrows;  // last row with bandstart <= g0
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (__ldg(G.bandstart + mid) <= g0) lo = mid; else hi = mid;
    }
    block_rows[b] = lo;
}

// Emit pass.  A warp owns 256 consecutive cells and takes them 32 at a time (lane = cell), so the
// flagged lanes of one step write CONSECUTIVE records: the record stores of a warp instruction fill
// whole 32-byte sectors instead of one scattered sector per lane.
__global__ void __launch_bounds__(kPostThreads) k5_emit(const __grid_constant__ TriGeom G,
                                                        const uint32_t* __restrict__ touched,
                                                        const unsigned long long* __restrict__ block_offsets,
                                                        const uint32_t* __restrict__ block_rows,
                                                        const uint32_t* __restrict__ totals,
                                                        const uint32_t* __restrict__ sample_ids,
                                                        ksb_record* __restrict__ records,
                                                        unsigned long long* __restrict__ pairs,
                                                        float* __restrict__ cmax_out) {
    __shared__ uint32_t sm[kPostThreads / 32];
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    const unsigned long long wcell0 = (unsigned long long)blockIdx.x * kPostChunk + wid * (32u * kPostPerThread);
    unsigned long long v[kPostPerThread];
    uint32_t ball[kPostPerThread];
    uint32_t wtotal = 0, mine = 0;
#pragma unroll
    for (uint32_t q = 0; q < kPostPerThread; q++) {
        const unsigned long long cell = wcell0 + 32u * q + lane;
        bool f = false;
        v[q] = 0ull;
        if (cell < G.cells) {
            v[q] = G.cell_bits == 32 ? (unsigned long long)reinterpret_cast<const uint32_t*>(G.tri)[cell] : G.tri[cell];
            f = v[q] != 0ull;
            if (touched != nullptr && !f) f = (touched[cell >> 5] >> (cell & 31u)) & 1u;
        }
        ball[q] = __ballot_sync(KSB_FULL, f);
        wtotal += __popc(ball[q]);
        mine |= (uint32_t)f << q;
    }
    if (lane == 0) sm[wid] = wtotal;
    __syncthreads();
    uint32_t wbase = 0;
#pragma unroll
    for (uint32_t i = 0; i < kPostThreads / 32; i++)
        if (i < wid) wbase += sm[i];
    if (mine == 0) return;
    unsigned long long out = block_offsets[blockIdx.x] + wbase;
    // decode (i,j) of this lane's first cell, then walk 32 cells per step
    const unsigned long long g0 = G.base + wcell0 + lane;
    // last row with bandstart <= g0 (rows with an empty band share their start): it lies between the row of
    // this block's first cell and the row of the next block's first cell (k5_block_rows)
    uint32_t lo = __ldg(block_rows + blockIdx.x);
    uint32_t hi = blockIdx.x + 1u < gridDim.x ? __ldg(block_rows + blockIdx.x + 1u) + 1u : G.rb + G.nrows;
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (__ldg(G.bandstart + mid) <= g0) lo = mid; else hi = mid;
    }
    uint32_t i = lo;
    uint32_t j = __ldg(G.band_lo + i) + (uint32_t)(g0 - __ldg(G.bandstart + i));
    uint32_t jend = __ldg(G.band_hi + i);
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (uint32_t q = 0; q < kPostPerThread; q++) {
        if ((mine >> q) != 0u) {  // only while flagged cells remain for this lane
            while (j >= jend) {   // into the next row(s): cells of consecutive rows are back to back
                const uint32_t over = j - jend;
                i++;
                jend = __ldg(G.band_hi + i);
                j = __ldg(G.band_lo + i) + over;
            }
        }
        if ((mine >> q) & 1u) {
            const unsigned long long o = out + __popc(ball[q] & lt);
            float cmin, cavg, cmax;
            containment(v[q], __ldg(totals + i), __ldg(totals + j), cmin, cavg, cmax);
            uint4 a, b;
            a.x = __ldg(sample_ids + i);
            a.y = __ldg(sample_ids + j);
            a.z = (uint32_t)v[q];
            a.w = (uint32_t)(v[q] >> 32);
            b.x = __float_as_uint(cmin);
            b.y = __float_as_uint(cavg);
            b.z = __float_as_uint(cmax);
            b.w = 0;
            uint4* dst = reinterpret_cast<uint4*>(records + o);
            dst[0] = a;
            dst[1] = b;
            pairs[o] = ((unsigned long long)i << 32) | j;
            cmax_out[o] = cmax;  // the default clustering column, as a dense array for K6
        }
        out += __popc(ball[q]);
        j += 32;
    }
}

// ------------------------------------------------------------------------------------------------
// K56: cluster-only post-pass (north_star 3: "only the edge list or cluster labels return to the host").
// ONE scan of the owned block replaces k5_count + k5_emit + the threshold scan of K6: a warp takes 32
// consecutive cells per step (lane = cell), every touched cell gets its (i, j) and the float32 containment of
// the selected column exactly as K5 computes it (src/pairwise.cpp:253-264), the reference's threshold on the
// printed text is the comparison with c* (ks_clustering.py:97-105, ksb_threshold_float_pct), and only the KEPT
// pairs are appended -- 8 bytes each, (i << 32) | j in compact indices -- to an edge list through a per-CTA
// queue (one global atomicAdd per CTA flush).  No records, no pair array, no column copy, no second read of
// the block.  counters: [0] touched cells (= TSV rows), [1] kept edges (may exceed the list's capacity: the
// host then grows the list and repeats the step), both accumulated over the passes of a step.
// ------------------------------------------------------------------------------------------------
constexpr uint32_t kScanQueue = 1024;  // kept pairs of one block queued in shared memory; a denser block appends the rest directly
constexpr uint32_t kScanCtasPerSm = 6;  // 24 KB of shared memory, 42 registers x 256 threads per CTA
struct ScanCut { uint32_t mode, use_max, m, shift; unsigned long long add; };
__global__ void __launch_bounds__(kPostThreads) k56_scan(const __grid_constant__ TriGeom G,
                                                         const uint32_t* __restrict__ touched,
                                                         const uint32_t* __restrict__ block_rows,
                                                         const uint32_t* __restrict__ totals, int dist_col, float cstar,
                                                         const ScanCut cut,
                                                         unsigned long long* __restrict__ edges, unsigned long long cap,
                                                         unsigned long long* counters, uint32_t n_blocks) {
    // Persistent CTAs over the 2048-cell blocks of the owned block.  A block is staged by ONE TMA bulk copy (16 KB,
    // mbarrier complete_tx) and moved to registers at once; the copy of the NEXT block is issued right after that,
    // so it is in flight while this block is decoded: a stream of 16 KB copies instead of a load -> wait -> compute
    // chain per CTA, with one 16 KB staging buffer (24 KB of shared memory per CTA, 6 CTAs per SM).
    __shared__ __align__(128) unsigned long long buf[1][kPostChunk];
    __shared__ unsigned long long q[kScanQueue];
    __shared__ __align__(8) unsigned long long bars[2];
    __shared__ uint32_t qn, rows_cta;
    __shared__ unsigned long long qbase;
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    const uint32_t bar0 = smem_u32(&bars[0]);
    auto fetch = [&](uint32_t blk, uint32_t slot) {  // thread 0 only
        const unsigned long long c0 = (unsigned long long)blk * kPostChunk;
        const unsigned long long left = G.cells - c0;
        const uint32_t bytes = (uint32_t)(((left < kPostChunk ? left : (unsigned long long)kPostChunk) * 8ull + 15ull) & ~15ull);  // the block is padded by 8 cells
        mbar_expect_tx(bar0 + 8u * slot, bytes);
        tma_load_1d(smem_u32(&buf[slot][0]), G.tri + c0, bytes, bar0 + 8u * slot);
    };
    if (threadIdx.x == 0) {
        mbar_init(bar0, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        qn = 0;
        rows_cta = 0;
        if (blockIdx.x < n_blocks) fetch(blockIdx.x, 0);
    }
    __syncthreads();
    uint32_t it = 0;
    for (uint32_t blk = blockIdx.x; blk < n_blocks; blk += gridDim.x, it++) {
        const uint32_t slot = 0u;
        mbar_wait(bar0, it & 1u);
        const unsigned long long wcell0 = (unsigned long long)blk * kPostChunk + wid * (32u * kPostPerThread);
        unsigned long long v[kPostPerThread];
        uint32_t mine = 0;
        const bool whole = (unsigned long long)(blk + 1u) * kPostChunk <= G.cells && touched == nullptr;  // CTA-uniform
        if (whole) {
#pragma unroll
            for (uint32_t k = 0; k < kPostPerThread; k++) {
                v[k] = buf[slot][wid * (32u * kPostPerThread) + 32u * k + lane];
                mine |= (uint32_t)(v[k] != 0ull) << k;
            }
        } else {
#pragma unroll
            for (uint32_t k = 0; k < kPostPerThread; k++) {
                const unsigned long long cell = wcell0 + 32u * k + lane;
                bool f = false;
                v[k] = 0ull;
                if (cell < G.cells) {
                    v[k] = buf[slot][wid * (32u * kPostPerThread) + 32u * k + lane];
                    f = v[k] != 0ull;
                    if (touched != nullptr && !f) f = (touched[cell >> 5] >> (cell & 31u)) & 1u;
                }
                mine |= (uint32_t)f << k;
            }
        }
        __syncthreads();  // every thread holds its cells: the staging buffer is free
        if (threadIdx.x == 0 && blk + gridDim.x < n_blocks) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy reads before the async-proxy write
            fetch(blk + gridDim.x, 0u);
        }
        const uint32_t nrows = __reduce_add_sync(KSB_FULL, (uint32_t)__popc(mine));
        if (lane == 0 && nrows) atomicAdd(&rows_cta, nrows);
        if (mine != 0u) {
            const unsigned long long g0 = G.base + wcell0 + lane;
            uint32_t lo = __ldg(block_rows + blk);
            uint32_t hi = blk + 1u < n_blocks ? __ldg(block_rows + blk + 1u) + 1u : G.rb + G.nrows;
            while (hi - lo > 1) {
                const uint32_t mid = (lo + hi) >> 1;
                if (__ldg(G.bandstart + mid) <= g0) lo = mid; else hi = mid;
            }
            uint32_t i = lo;
            uint32_t j = __ldg(G.band_lo + i) + (uint32_t)(g0 - __ldg(G.bandstart + i));
            uint32_t jend = __ldg(G.band_hi + i);
            uint32_t ti = __ldg(totals + i);
#pragma unroll
            for (uint32_t k = 0; k < kPostPerThread; k++) {
                if ((mine >> k) == 0u) break;
                while (j >= jend) {
                    const uint32_t over = j - jend;
                    i++;
                    jend = __ldg(G.band_hi + i);
                    j = __ldg(G.band_lo + i) + over;
                    ti = __ldg(totals + i);
                }
                if ((mine >> k) & 1u) {
                    const uint32_t tj = __ldg(totals + j);
                    float c;
                    const uint32_t T = cut.use_max ? max(ti, tj) : min(ti, tj);
                    if (cut.mode == 1u && v[k] < (1ull << 24) && T - 1u < (1u << 24) - 1u && min(ti, tj) != 0u) {
                        // EXACT integer form of `float32 quotient >= c*` (no conversion, no division): the correctly
                        // rounded quotient S / T reaches c* iff the real quotient reaches the midpoint M * 2^-shift
                        // between c* and its predecessor (ties go to the even one of the two), i.e. iff
                        // S * 2^shift >= M * T (> when the tie goes down).  S < 2^24 and T < 2^24 are exact as float32,
                        // M < 2^26, shift <= 39.
                        // as a threshold on S: ceil(M T / 2^shift) when a tie is kept, floor(..) + 1 otherwise (cut.add)
                        const unsigned long long thr = ((unsigned long long)cut.m * T + cut.add) >> cut.shift;
                        if (v[k] >= thr) {
                            const uint32_t pos = atomicAdd(&qn, 1u);
                            if (pos < kScanQueue) q[pos] = ((unsigned long long)i << 32) | j;
                            else {
                                const unsigned long long o = atomicAdd(counters + 1, 1ull);
                                if (o < cap) edges[o] = ((unsigned long long)i << 32) | j;
                            }
                        }
                        j += 32;
                        continue;
                    }
                    if (dist_col != 4 && ti != 0u && tj != 0u) {
                        // min / max of the two quotients = the quotient by the larger / smaller total: correctly rounded
                        // division is monotone in the divisor, so this IS std::min / std::max of the two float32 values
                        // (src/pairwise.cpp:260-264) -- with one division.  A zero total (inf / NaN) takes the full path.
                        c = __fdiv_rn(__ull2float_rn(v[k]), __uint2float_rn(dist_col == 3 ? max(ti, tj) : min(ti, tj)));
                    } else {
                        float cmin, cavg, cmax;
                        containment(v[k], ti, tj, cmin, cavg, cmax);
                        c = dist_col == 3 ? cmin : (dist_col == 4 ? cavg : cmax);
                    }
                    if (!(c < cstar)) {  // NaN is kept, as in `if distance < cutoff: continue`
                        const uint32_t pos = atomicAdd(&qn, 1u);
                        if (pos < kScanQueue) q[pos] = ((unsigned long long)i << 32) | j;
                        else {  // queue full: straight to the list
                            const unsigned long long o = atomicAdd(counters + 1, 1ull);
                            if (o < cap) edges[o] = ((unsigned long long)i << 32) | j;
                        }
                    }
                }
                j += 32;
            }
        }
        __syncthreads();
        const uint32_t n = min(qn, kScanQueue);
        // the block's kept pairs go to the list with ONE global atomicAdd; the row counter once, at the end
        const bool last = blk + gridDim.x >= n_blocks;
        if (n != 0u || last) {
            if (threadIdx.x == 0) {
                qbase = n ? atomicAdd(counters + 1, (unsigned long long)n) : 0ull;
                if (last && rows_cta) atomicAdd(counters, (unsigned long long)rows_cta);
            }
            __syncthreads();
            const unsigned long long base = qbase;
            for (uint32_t k = threadIdx.x; k < n; k += kPostThreads)
                if (base + k < cap) edges[base + k] = q[k];
            __syncthreads();
            if (threadIdx.x == 0) qn = 0;
            __syncthreads();
        }
    }
}

// ------------------------------------------------------------------------------------------------
// k56_rows: the same post-pass walking the block ROW BY ROW.  The banded block stores the cells of a row back to back
// (columns band_lo[i] .. band_hi[i]), so a warp that owns a row knows (i, j) of every cell it loads without any search:
// lane = column, 32 consecutive cells (256 B) and their 32 totals (128 B) per step, four steps in flight; the row's own
// total, band and first cell are loaded once per row.  k56_scan, which cuts the block into 2048-cell chunks regardless
// of rows, spends ~68 instructions per cell on finding and following rows (ncu: issue-bound at 41 % of the HBM peak
// on cfg4); here a cell costs a load, a gather from the (L2-resident) totals, a multiply-add, a shift and a compare.
// Rows are dealt to the warps of the grid round-robin (neighbouring rows of a clade / family differ in length by one
// cell, so every warp gets the same mix).  Kept pairs go to a per-warp queue in shared memory (ballot + popc, no
// atomics) that is flushed to the edge list with one global atomicAdd per ~100 pairs.
// ------------------------------------------------------------------------------------------------
constexpr uint32_t kRowsQueue = 128;      // per warp; flushed when a step could overflow it
constexpr uint32_t kRowsCtasPerSm = 8;
__device__ __noinline__ bool k56_keep_slow(unsigned long long v, uint32_t ti, uint32_t tj, int dist_col, float cstar) {
    float c;
    if (dist_col != 4 && ti != 0u && tj != 0u) {
        // min / max of the two quotients = the quotient by the larger / smaller total (see k56_scan)
        c = __fdiv_rn(__ull2float_rn(v), __uint2float_rn(dist_col == 3 ? max(ti, tj) : min(ti, tj)));
    } else {
        float cmin, cavg, cmax;
        containment(v, ti, tj, cmin, cavg, cmax);
        c = dist_col == 3 ? cmin : (dist_col == 4 ? cavg : cmax);
    }
    return !(c < cstar);  // NaN is kept, as in `if distance < cutoff: continue`
}
template <bool TOUCHED, bool USE_MAX, class CT>
__global__ void __launch_bounds__(256) k56_rows(const __grid_constant__ TriGeom G, const uint32_t* __restrict__ touched,
                                                const uint32_t* __restrict__ totals, int dist_col, float cstar,
                                                const ScanCut cut, unsigned long long* __restrict__ edges,
                                                unsigned long long cap, unsigned long long* counters, uint32_t* parent) {
    __shared__ unsigned long long q[8][kRowsQueue];
    __shared__ uint32_t rows_cta;
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    const uint32_t warp = blockIdx.x * 8u + wid, nwarps = gridDim.x * 8u;
    const uint32_t lane_lt = (1u << lane) - 1u;
    unsigned long long* myq = q[wid];
    uint32_t qn = 0;           // warp-uniform
    uint32_t rows_seen = 0;    // touched cells this lane saw
    if (threadIdx.x == 0) rows_cta = 0;
    const bool fast = cut.mode == 1u;
    const unsigned long long cm = cut.m, cadd = cut.add;
    const uint32_t cshift = cut.shift;
    // Software pipeline.  The meta data of the NEXT row (band, first cell, total: dependent loads) is fetched when the current
    // row starts, and the next 128-cell group -- of this row, or the first one of the next row -- is fetched before the current
    // group is decoded, so neither the per-row prologue nor the per-group loads are waited for in the open.
    struct Meta { uint32_t i, lo, len, ti; unsigned long long c0; };
    struct Group { CT v0, v1, v2, v3; uint32_t t0, t1, t2, t3; };
    auto load_meta = [&](uint32_t r) {
        Meta m;
        m.i = G.rb + r;
        m.lo = __ldg(G.band_lo + m.i);
        const uint32_t hi = __ldg(G.band_hi + m.i);
        m.len = hi > m.lo ? hi - m.lo : 0u;
        m.c0 = __ldg(G.bandstart + m.i) - G.base;  // first cell of the row inside the owned block
        m.ti = __ldg(totals + m.i);
        return m;
    };
    auto load_group = [&](const Meta& m, uint32_t x0) {
        Group g;
        const CT* vp = reinterpret_cast<const CT*>(G.tri) + m.c0 + x0 + lane;
        const uint32_t* tp = totals + m.lo + x0 + lane;
        if (x0 + 128u <= m.len) {  // only the row's last group checks bounds
            g.v0 = vp[0]; g.v1 = vp[32]; g.v2 = vp[64]; g.v3 = vp[96];
            g.t0 = __ldg(tp); g.t1 = __ldg(tp + 32); g.t2 = __ldg(tp + 64); g.t3 = __ldg(tp + 96);
        } else {  // cells past the row's end read as untouched (0) with a harmless total
            const uint32_t x = x0 + lane;
            g.v0 = x < m.len ? vp[0] : (CT)0;        g.t0 = x < m.len ? __ldg(tp) : 1u;
            g.v1 = x + 32u < m.len ? vp[32] : (CT)0; g.t1 = x + 32u < m.len ? __ldg(tp + 32) : 1u;
            g.v2 = x + 64u < m.len ? vp[64] : (CT)0; g.t2 = x + 64u < m.len ? __ldg(tp + 64) : 1u;
            g.v3 = x + 96u < m.len ? vp[96] : (CT)0; g.t3 = x + 96u < m.len ? __ldg(tp + 96) : 1u;
        }
        return g;
    };
    Meta m{};
    Group g{};
    if (warp < G.nrows) {
        m = load_meta(warp);
        if (m.len) g = load_group(m, 0u);
    }
    for (uint32_t r = warp; r < G.nrows; r += nwarps) {
        Meta mn{};
        const bool has_next = r + nwarps < G.nrows;
        if (has_next) mn = load_meta(r + nwarps);
        const uint32_t i = m.i, lo = m.lo, len = m.len, ti = m.ti;
        const unsigned long long c0 = m.c0;
        const unsigned long long ihi = (unsigned long long)i << 32;
        // a kept pair goes to the warp's queue (every lane calls; ballot + popc, no atomics)
        auto enqueue = [&](bool keep, uint32_t x) {
            const uint32_t mk = __ballot_sync(KSB_FULL, keep);
            if (mk != 0u) {
                if (keep) myq[qn + __popc(mk & lane_lt)] = ihi | (lo + x);
                qn += __popc(mk);
                __syncwarp();
                if (qn > kRowsQueue - 32u) {  // flush: one global atomicAdd per ~100 kept pairs
                    unsigned long long base = 0;
                    if (lane == 0) base = atomicAdd(counters + 1, (unsigned long long)qn);
                    base = shfl_u64(base, 0);
                    for (uint32_t k = lane; k < qn; k += 32u) {
                        const unsigned long long ij = myq[k];
                        if (base + k < cap) edges[base + k] = ij;
                        // the hook of K6a (see k6_hook_min) on the way out: fire-and-forget, the unions follow in their own kernel
                        if (parent != nullptr) atomicMin(parent + (uint32_t)ij, (uint32_t)(ij >> 32));
                    }
                    __syncwarp();
                    qn = 0;
                }
            }
        };
        // one cell, any value: (value, total of its column, column)
        auto cell = [&](unsigned long long v, uint32_t tj, uint32_t x, bool inside) {
            bool t = v != 0ull;
            if (TOUCHED) {
                if (!t && inside) {
                    const unsigned long long cidx = c0 + x;
                    t = (__ldg(touched + (cidx >> 5)) >> (cidx & 31u)) & 1u;
                }
            }
            bool keep = false;
            if (t) {
                rows_seen++;
                const uint32_t mn_ = min(ti, tj);
                const uint32_t T = USE_MAX ? max(ti, tj) : mn_;
                // eligible for the exact integer form of `float32 quotient >= c*` (see k56_scan / make_scan_cut):
                // S < 2^24, 0 < T < 2^24, no zero total
                const bool ok = fast && (uint32_t)(v >> 24) == 0u && (uint32_t)(v >> 32) == 0u && T - 1u < (1u << 24) - 1u && (!USE_MAX || mn_ != 0u);
                if (ok) keep = v >= ((cm * T + cadd) >> cshift);
                else keep = k56_keep_slow(v, ti, tj, dist_col, cstar);
            }
            enqueue(keep, x);
        };
        // Whole groups of 128 cells whose sums and totals are all below 2^24 take the integer test without any per-cell
        // eligibility check: a cell is touched iff it is non-zero (no zero-weight colours in this variant), and a zero
        // total needs no special case -- S / 0 = inf >= c* is what S * 2^shift >= M * 0 says too.  c* <= 2^7 makes the
        // threshold on S fit 32 bits (M T < 2^50, shift >= 18).
        const bool row_fast = !TOUCHED && fast && cshift >= 18u && (USE_MAX || ti < (1u << 24));
        for (uint32_t x0 = 0; x0 < len; x0 += 128u) {
            Group gn{};
            if (x0 + 128u < len) gn = load_group(m, x0 + 128u);
            else if (has_next && mn.len) gn = load_group(mn, 0u);   // the next row's first group
            bool small = row_fast;
            if (row_fast) {
                const unsigned long long vo = (unsigned long long)(g.v0 | g.v1 | g.v2 | g.v3);
                small = (vo >> 24) == 0ull;
                if (USE_MAX) small = small && ((g.t0 | g.t1 | g.t2 | g.t3) >> 24) == 0u;
            }
            if (__all_sync(KSB_FULL, small)) {
                const CT vv[4] = {g.v0, g.v1, g.v2, g.v3};
                const uint32_t tt[4] = {g.t0, g.t1, g.t2, g.t3};
                bool kp[4];
#pragma unroll
                for (uint32_t u = 0; u < 4; u++) {
                    const uint32_t S = (uint32_t)vv[u];
                    const uint32_t T = USE_MAX ? max(ti, tt[u]) : min(ti, tt[u]);
                    const uint32_t thr = (uint32_t)((cm * T + cadd) >> cshift);
                    const bool nz = S != 0u;
                    rows_seen += nz ? 1u : 0u;
                    kp[u] = nz && S >= thr;
                }
                if (__any_sync(KSB_FULL, kp[0] || kp[1] || kp[2] || kp[3])) {  // most groups keep nothing: one vote instead of four
#pragma unroll
                    for (uint32_t u = 0; u < 4; u++) enqueue(kp[u], x0 + 32u * u + lane);
                }
            } else {
                const uint32_t x = x0 + lane;
                cell(g.v0, g.t0, x, x < len);
                if (x0 + 32u < len) cell(g.v1, g.t1, x + 32u, x + 32u < len);   // warp-uniform
                if (x0 + 64u < len) cell(g.v2, g.t2, x + 64u, x + 64u < len);
                if (x0 + 96u < len) cell(g.v3, g.t3, x + 96u, x + 96u < len);
            }
            g = gn;
        }
        if (len == 0u && has_next && mn.len) g = load_group(mn, 0u);  // an empty row prefetched nothing
        m = mn;
    }
    if (qn) {
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(counters + 1, (unsigned long long)qn);
        base = shfl_u64(base, 0);
        for (uint32_t k = lane; k < qn; k += 32u) {
            const unsigned long long ij = myq[k];
            if (base + k < cap) edges[base + k] = ij;
            if (parent != nullptr) atomicMin(parent + (uint32_t)ij, (uint32_t)(ij >> 32));
        }
    }
    rows_seen = __reduce_add_sync(KSB_FULL, rows_seen);
    __syncthreads();
    if (lane == 0 && rows_seen) atomicAdd(&rows_cta, rows_seen);  // one counter update per CTA
    __syncthreads();
    if (threadIdx.x == 0 && rows_cta) atomicAdd(counters, (unsigned long long)rows_cta);
}

// hook (atomicMin accelerator, see k6_hook_min) and union over a compact edge list whose length lives on the device
__global__ void __launch_bounds__(256) k6_edges_hook(const unsigned long long* __restrict__ edges, unsigned long long cap,
                                                     const unsigned long long* __restrict__ counters, uint32_t* parent) {
    const unsigned long long n = min(__ldg(counters + 1), cap);
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long r = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += 4ull * stride) {
        unsigned long long ij[4];  // four independent loads in flight per thread
#pragma unroll
        for (uint32_t u = 0; u < 4; u++) ij[u] = r + u * stride < n ? __ldg(edges + r + u * stride) : ~0ull;
#pragma unroll
        for (uint32_t u = 0; u < 4; u++)
            if (ij[u] != ~0ull) atomicMin(parent + (uint32_t)ij[u], (uint32_t)(ij[u] >> 32));
    }
}
__global__ void __launch_bounds__(256) k6_edges_union(const unsigned long long* __restrict__ edges, unsigned long long cap,
                                                      const unsigned long long* __restrict__ counters, uint32_t* parent) {
    const unsigned long long n = min(__ldg(counters + 1), cap);
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long r = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += 4ull * stride) {
        unsigned long long ij[4];
#pragma unroll
        for (uint32_t u = 0; u < 4; u++) ij[u] = r + u * stride < n ? __ldg(edges + r + u * stride) : ~0ull;
#pragma unroll
        for (uint32_t u = 0; u < 4; u++)
            if (ij[u] != ~0ull) uf_union(parent, (uint32_t)(ij[u] >> 32), (uint32_t)ij[u]);
    }
}

// Integer form of the threshold for the min / max column (see k56_scan): c* = m24 * 2^E; the midpoint between c* and its
// predecessor is M * 2^-shift with M < 2^26.  mode 0 = use the float path (avg column, c* of 0 / inf / NaN, or a c* so
// small that S << shift could overflow).
static ScanCut make_scan_cut(int dist_col, float cstar) {
    ScanCut c{};
    c.mode = 0;
    c.use_max = dist_col == 3 ? 1u : 0u;  // min containment = quotient by the LARGER total
    if (dist_col == 4 || !(cstar > 0.0f) || !std::isfinite(cstar)) return c;
    uint32_t bits;
    std::memcpy(&bits, &cstar, 4);
    if ((bits >> 23) == 0u) return c;  // subnormal c*
    uint32_t pb = bits - 1u;           // predecessor (c* > 0 and normal, so this is the next float below)
    float pred;
    std::memcpy(&pred, &pb, 4);
    const double mid = ((double)pred + (double)cstar) * 0.5;  // exact: <= 25 significant bits
    int e = 0;
    const double fr = std::frexp(mid, &e);                    // mid = fr * 2^e, 0.5 <= fr < 1
    const double M = std::ldexp(fr, 26);                      // integer < 2^26
    const int shift = 26 - e;
    if (M != std::floor(M) || shift < 0 || shift > 39) return c;
    c.m = (uint32_t)M;
    c.shift = (uint32_t)shift;
    // round-half-to-even lands on c* iff its mantissa is even: then S 2^shift >= M T keeps (ceil), else > (floor + 1)
    c.add = (bits & 1u) == 0u ? (1ull << shift) - 1ull : (1ull << shift);
    c.mode = 1;
    return c;
}

// row-wise scan unless the rows are too short to feed a warp (fewer than 32 cells per owned row on average) or
// KSPIDER_K56=scan asks for the chunk-wise kernel; KSPIDER_K56=rows forces it; k56_scan stages u64 cells only
bool k56_uses_rows(const TriGeom& g) {
    static const int mode = [] { const char* v = std::getenv("KSPIDER_K56"); return !v ? 0 : (!std::strcmp(v, "scan") ? 1 : (!std::strcmp(v, "rows") ? 2 : 0)); }();
    return mode == 2 || g.cell_bits == 32 || (mode == 0 && g.cells >= 32ull * g.nrows);
}

cudaError_t launch_k56_scan(const TriGeom& g, const uint32_t* touched_bits, const uint32_t* block_rows, const uint32_t* totals,
                            int dist_col, float cstar, unsigned long long* edges, unsigned long long cap,
                            unsigned long long* counters, uint32_t* parent, const LaunchCfg& lc) {
    const uint64_t nb = post_blocks(g.cells);
    if (nb == 0) return cudaSuccess;
    const ScanCut cut = make_scan_cut(dist_col, cstar);
    if (k56_uses_rows(g)) {
        // persistent warps: exactly the CTAs that are resident at once (the variant's registers decide: 4 or 5 per SM)
        const uint64_t want = ((uint64_t)g.nrows + 7) / 8;
        const int variant = (touched_bits ? 4 : 0) | (cut.use_max ? 2 : 0) | (g.cell_bits == 32 ? 1 : 0);
        static int per_sm[8] = {0, 0, 0, 0, 0, 0, 0, 0};  // per kernel variant (the variants share one function-pointer type)
        auto go = [&](auto kern) {
            int& ps = per_sm[variant];
            if (ps == 0) {
                int n = 0;
                if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern, 256, 0) != cudaSuccess || n < 1) { (void)cudaGetLastError(); n = 4; }
                ps = n < (int)kRowsCtasPerSm ? n : (int)kRowsCtasPerSm;
            }
            const uint64_t capr = (uint64_t)lc.sms * (uint64_t)ps;
            kern<<<(unsigned)(want < capr ? want : capr), 256, 0, lc.stream>>>(g, touched_bits, totals, dist_col, cstar, cut, edges, cap, counters, parent);
        };
        switch (variant) {
            case 0: go(k56_rows<false, false, unsigned long long>); break;
            case 1: go(k56_rows<false, false, uint32_t>); break;
            case 2: go(k56_rows<false, true, unsigned long long>); break;
            case 3: go(k56_rows<false, true, uint32_t>); break;
            case 4: go(k56_rows<true, false, unsigned long long>); break;
            case 5: go(k56_rows<true, false, uint32_t>); break;
            case 6: go(k56_rows<true, true, unsigned long long>); break;
            default: go(k56_rows<true, true, uint32_t>); break;
        }
        return cudaGetLastError();
    }
    const uint64_t capc = (uint64_t)lc.sms * kScanCtasPerSm;
    k56_scan<<<(unsigned)(nb < capc ? nb : capc), kPostThreads, 0, lc.stream>>>(g, touched_bits, block_rows, totals, dist_col, cstar, cut, edges,
                                                                                cap, counters, (uint32_t)nb);
    return cudaGetLastError();
}
cudaError_t launch_k6_edges(const unsigned long long* edges, unsigned long long cap, const unsigned long long* counters,
                            uint32_t* parent, uint64_t expect, bool hooked, const LaunchCfg& lc) {
    const unsigned long long want = (std::max<uint64_t>(expect, 1) + 255) / 256, capc = (unsigned long long)lc.sms * 8ull;
    const unsigned ctas = (unsigned)(want < capc ? want : capc);
    if (!hooked) k6_edges_hook<<<ctas, 256, 0, lc.stream>>>(edges, cap, counters, parent);  // k56_rows hooks while it appends
    k6_edges_union<<<ctas, 256, 0, lc.stream>>>(edges, cap, counters, parent);
    return cudaGetLastError();
}

cudaError_t launch_k5_count(const TriGeom& g, const uint32_t* touched_bits, unsigned long long* block_counts,
                            const LaunchCfg& lc) {
    const uint64_t nb = post_blocks(g.cells);
    if (nb == 0) return cudaSuccess;
    k5_count<<<(unsigned)nb, kPostThreads, 0, lc.stream>>>(g, touched_bits, block_counts);
    return cudaGetLastError();
}

cudaError_t launch_k5_block_rows(const TriGeom& g, uint32_t* block_rows, const LaunchCfg& lc) {
    const uint64_t nb = post_blocks(g.cells);
    if (nb == 0) return cudaSuccess;
    k5_block_rows<<<(unsigned)((nb + 255) / 256), 256, 0, lc.stream>>>(g, nb, block_rows);
    return cudaGetLastError();
}

cudaError_t launch_k5_emit(const TriGeom& g, const uint32_t* touched_bits, const unsigned long long* block_offsets,
                           const uint32_t* block_rows, const uint32_t* totals, const uint32_t* sample_ids, ksb_record* records,
                           unsigned long long* pairs, float* cmax, const LaunchCfg& lc) {
    const uint64_t nb = post_blocks(g.cells);
    if (nb == 0) return cudaSuccess;
    k5_emit<<<(unsigned)nb, kPostThreads, 0, lc.stream>>>(g, touched_bits, block_offsets, block_rows, totals, sample_ids, records,
                                                         pairs, cmax);
    return cudaGetLastError();
}

// K5 for the spill path: one thread per (key, sum) of the final sorted run.
__global__ void __launch_bounds__(256) k5_sparse(const unsigned long long* __restrict__ keys,
                                                 const unsigned long long* __restrict__ sums, unsigned long long n,
                                                 uint32_t shift, const uint32_t* __restrict__ totals,
                                                 const uint32_t* __restrict__ sample_ids, ksb_record* __restrict__ records,
                                                 unsigned long long* __restrict__ pairs, float* __restrict__ cmax_out) {
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long r = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += stride) {
        const unsigned long long k = keys[r], v = sums[r];
        const uint32_t i = (uint32_t)(k >> shift), j = (uint32_t)(k & ((1ull << shift) - 1ull));
        float cmin, cavg, cmax;
        containment(v, __ldg(totals + i), __ldg(totals + j), cmin, cavg, cmax);
        uint4 a, b;
        a.x = __ldg(sample_ids + i);
        a.y = __ldg(sample_ids + j);
        a.z = (uint32_t)v;
        a.w = (uint32_t)(v >> 32);
        b.x = __float_as_uint(cmin);
        b.y = __float_as_uint(cavg);
        b.z = __float_as_uint(cmax);
        b.w = 0;
        uint4* dst = reinterpret_cast<uint4*>(records + r);
        dst[0] = a;
        dst[1] = b;
        pairs[r] = ((unsigned long long)i << 32) | j;
        cmax_out[r] = cmax;
    }
}

cudaError_t launch_k5_sparse(const unsigned long long* keys, const unsigned long long* sums, uint64_t n, uint32_t shift,
                             const uint32_t* totals, const uint32_t* sample_ids, ksb_record* records,
                             unsigned long long* pairs, float* cmax, const LaunchCfg& lc) {
    if (n == 0) return cudaSuccess;
    unsigned long long ctas = (n + 255) / 256;
    const unsigned long long cap = (unsigned long long)lc.sms * 8ull;
    if (ctas > cap) ctas = cap;
    k5_sparse<<<(unsigned)ctas, 256, 0, lc.stream>>>(keys, sums, n, shift, totals, sample_ids, records, pairs, cmax);
    return cudaGetLastError();
}

cudaError_t scan_block_counts(void* temp, size_t* temp_bytes, const unsigned long long* in, unsigned long long* out,
                              uint64_t n, cudaStream_t stream) {
    return cub::DeviceScan::ExclusiveSum(temp, *temp_bytes, in, out, (int64_t)n, stream);
}

// ------------------------------------------------------------------------------------------------
// K6 + K7: threshold on the selected containment column and lock-free union-find (smaller root wins,
// so the final label of a component is its smallest compact sample index).
// ------------------------------------------------------------------------------------------------
// K6a: before any union, every kept edge (i < j) lowers parent[j] to i with one fire-and-forget
// atomicMin.  Each link joins two samples that ARE connected and keeps parent[x] <= x, so the result is a
// valid union-find forest whose roots are the minima of their trees; in clique-like components (a clade:
// 50 samples, 1225 kept edges) it already is the final star, and the unions of K6b -- which still process
// EVERY edge, so nothing depends on this pass for correctness -- find equal roots after two loads instead
// of fighting over the same few roots with CAS retries.  Must run over all segments before the first union.
__global__ void __launch_bounds__(256) k6_hook_min(const ksb_record* __restrict__ records,
                                                   const unsigned long long* __restrict__ pairs,
                                                   const float* __restrict__ colvals, unsigned long long n_rows,
                                                   int dist_col, float cstar, uint32_t* parent) {
    // 8 independent coalesced loads per thread: with one load in flight per thread the scan ran at 2.3 TB/s
    constexpr uint32_t R = 8;
    const unsigned long long span = 256ull * R;
    for (unsigned long long base = (unsigned long long)blockIdx.x * span; base < n_rows; base += (unsigned long long)gridDim.x * span) {
        float c[R];
#pragma unroll
        for (uint32_t k = 0; k < R; k++) {
            const unsigned long long r = base + k * 256u + threadIdx.x;
            c[k] = -INFINITY;
            if (r < n_rows) c[k] = colvals ? __ldg(colvals + r) : __ldg(&records[r].min_containment + (dist_col - 3));
        }
#pragma unroll
        for (uint32_t k = 0; k < R; k++) {
            const unsigned long long r = base + k * 256u + threadIdx.x;
            if (r >= n_rows || c[k] < cstar) continue;  // NaN is kept
            const unsigned long long ij = __ldg(pairs + r);
            atomicMin(parent + (uint32_t)ij, (uint32_t)(ij >> 32));
        }
    }
}

// Threshold and union are decoupled inside the CTA: a pass over 4 x 256 rows (four independent coalesced
// loads per thread) appends the kept edges to a shared-memory queue, and the unions are run 256 at a time,
// one per thread, only when the queue holds a full round (the rest is carried to the next pass).  With the
// union done in place (`if (keep) uf_union`) a warp waited ~2.5 us of dependent L2 accesses for the two or
// three lanes that had an edge before it could test its next 32 rows: cfg4 (5e7 rows, 5 % kept) 0.37 ms.
constexpr uint32_t kUnionRows = 8;  // rows per thread and pass
__global__ void __launch_bounds__(256) k67_union(const ksb_record* __restrict__ records,
                                                 const unsigned long long* __restrict__ pairs,
                                                 const float* __restrict__ colvals,  // dense copy of the column, or null
                                                 unsigned long long n_rows, int dist_col, float cstar,
                                                 uint32_t* parent, unsigned long long* n_edges) {
    __shared__ unsigned long long q[256 * (kUnionRows + 1)];
    __shared__ uint32_t qn;
    const uint32_t tid = threadIdx.x, lane = tid & 31u;
    if (tid == 0) qn = 0;
    __syncthreads();
    uint32_t kept = 0;
    const unsigned long long span = 256ull * kUnionRows;
    for (unsigned long long base = (unsigned long long)blockIdx.x * span; base < n_rows; base += (unsigned long long)gridDim.x * span) {
        float c[kUnionRows];
#pragma unroll
        for (uint32_t k = 0; k < kUnionRows; k++) {
            const unsigned long long r = base + k * 256u + tid;
            c[k] = INFINITY;
            if (r < n_rows) c[k] = colvals ? __ldg(colvals + r) : __ldg(&records[r].min_containment + (dist_col - 3));
        }
#pragma unroll
        for (uint32_t k = 0; k < kUnionRows; k++) {
            const unsigned long long r = base + k * 256u + tid;
            const bool keep = r < n_rows && !(c[k] < cstar);  // NaN is kept, as in `if distance < cutoff: continue`
            const uint32_t m = __ballot_sync(KSB_FULL, keep);
            if (m == 0u) continue;
            uint32_t pos = 0;
            if (lane == 0) pos = atomicAdd(&qn, (uint32_t)__popc(m));
            pos = __shfl_sync(KSB_FULL, pos, 0);
            if (keep) {
                q[pos + __popc(m & ((1u << lane) - 1u))] = __ldg(pairs + r);
                kept++;
            }
        }
        __syncthreads();
        uint32_t n = qn;
        while (n >= 256u) {  // full rounds only; edges are taken from the end so nothing has to move
            const unsigned long long ij = q[n - 256u + tid];
            uf_union(parent, (uint32_t)(ij >> 32), (uint32_t)ij);
            n -= 256u;
        }
        __syncthreads();
        if (tid == 0) qn = n;
        __syncthreads();
    }
    const uint32_t n = qn;  // the carried rest: fewer than 256 edges
    if (tid < n) {
        const unsigned long long ij = q[tid];
        uf_union(parent, (uint32_t)(ij >> 32), (uint32_t)ij);
    }
    kept = __reduce_add_sync(KSB_FULL, kept);
    if (lane == 0 && kept) atomicAdd(n_edges, (unsigned long long)kept);
}

__global__ void k7_init(uint32_t* parent, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) parent[i] = i;
}
__global__ void k7_compress(uint32_t* parent, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        uint32_t r = i;
        while (parent[r] != r) r = parent[r];
        parent[i] = r;  // roots never change here; writing a root over an ancestor is safe
    }
}

cudaError_t launch_k67_union(const ksb_record* records, const unsigned long long* pairs, const float* colvals, uint64_t n_rows,
                             int dist_col, float cstar, uint32_t* parent, unsigned long long* n_edges, const LaunchCfg& lc) {
    if (n_rows == 0) return cudaSuccess;
    unsigned long long ctas = (n_rows + 256 * kUnionRows - 1) / (256 * kUnionRows);
    const unsigned long long cap = (unsigned long long)lc.sms * 8ull;
    if (ctas > cap) ctas = cap;
    k67_union<<<(unsigned)ctas, 256, 0, lc.stream>>>(records, pairs, colvals, n_rows, dist_col, cstar, parent, n_edges);
    return cudaGetLastError();
}
cudaError_t launch_k6_hook_min(const ksb_record* records, const unsigned long long* pairs, const float* colvals, uint64_t n_rows,
                               int dist_col, float cstar, uint32_t* parent, const LaunchCfg& lc) {
    if (n_rows == 0) return cudaSuccess;
    unsigned long long ctas = (n_rows + 2047) / 2048;
    const unsigned long long cap = (unsigned long long)lc.sms * 8ull;
    if (ctas > cap) ctas = cap;
    k6_hook_min<<<(unsigned)ctas, 256, 0, lc.stream>>>(records, pairs, colvals, n_rows, dist_col, cstar, parent);
    return cudaGetLastError();
}
cudaError_t launch_k7_init(uint32_t* parent, uint32_t n, const LaunchCfg& lc) {
    if (n == 0) return cudaSuccess;
    k7_init<<<(n + 255) / 256, 256, 0, lc.stream>>>(parent, n);
    return cudaGetLastError();
}
cudaError_t launch_k7_compress(uint32_t* parent, uint32_t n, const LaunchCfg& lc) {
    if (n == 0) return cudaSuccess;
    k7_compress<<<(n + 255) / 256, 256, 0, lc.stream>>>(parent, n);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// red micro-benchmark: the roofline the accumulation kernels are measured against.
// pattern 0: every lane a uniformly random cell; pattern 1: a random 32-cell run per warp.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long mix64(unsigned long long x) {
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
    return x;
}

template <typename T>
__global__ void __launch_bounds__(256) bench_red(T* buf, unsigned long long cells, unsigned long long n_ops, int pattern) {
    const unsigned long long tid = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    const uint32_t lane = threadIdx.x & 31u;
    for (unsigned long long op = tid; op < n_ops; op += stride) {
        unsigned long long cell;
        if (pattern == 0) {
            cell = __umul64hi(mix64(op + 0x9E3779B97F4A7C15ULL), cells);
        } else {
            cell = __umul64hi(mix64((op >> 5) + 0x9E3779B97F4A7C15ULL), cells - 32) + lane;
        }
        atomicAdd(buf + cell, (T)1);
    }
}

cudaError_t launch_bench_red(void* buf, uint64_t cells, uint64_t n_ops, int cell_bits, int pattern, const LaunchCfg& lc) {
    const unsigned ctas = (unsigned)lc.sms * 8u;
    if (cell_bits == 64)
        bench_red<unsigned long long><<<ctas, 256, 0, lc.stream>>>((unsigned long long*)buf, cells, n_ops, pattern);
    else
        bench_red<unsigned int><<<ctas, 256, 0, lc.stream>>>((unsigned int*)buf, cells, n_ops, pattern);
    return cudaGetLastError();
}

}  // namespace ksb

namespace ksb {
__global__ void __launch_bounds__(256) k7_merge(uint32_t* parent, const uint32_t* __restrict__ other, uint32_t n) {
    const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n) {
        const uint32_t o = __ldg(other + v);
        if (o != v) uf_union(parent, v, o);
    }
}
// all the other ranks' label arrays in one launch: others = [n_other][n]
__global__ void __launch_bounds__(256) k7_merge_all(uint32_t* parent, const uint32_t* __restrict__ others, uint32_t n, uint32_t n_other) {
    const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    for (uint32_t w = 0; w < n_other; w++) {
        const uint32_t o = __ldg(others + (unsigned long long)w * n + v);
        if (o != v) uf_union(parent, v, o);
    }
}
cudaError_t launch_k7_merge_all(uint32_t* parent, const uint32_t* others, uint32_t n, uint32_t n_other, const LaunchCfg& lc) {
    if (n == 0 || n_other == 0) return cudaSuccess;
    k7_merge_all<<<(n + 255) / 256, 256, 0, lc.stream>>>(parent, others, n, n_other);
    return cudaGetLastError();
}
cudaError_t launch_k7_merge(uint32_t* parent, const uint32_t* other, uint32_t n, const LaunchCfg& lc) {
    if (n == 0) return cudaSuccess;
    k7_merge<<<(n + 255) / 256, 256, 0, lc.stream>>>(parent, other, n);
    return cudaGetLastError();
}
}  // namespace ksb
