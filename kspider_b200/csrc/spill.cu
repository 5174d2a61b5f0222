// K3: sort-reduce spill path.  Used when the owned block of the dense triangle does not fit the
// budget (or is forced): the pair-emission kernels append (key = row << shift | column, weight)
// to a chunk buffer instead of doing reds; every chunk is radix-sorted and reduced by key
// (cub::DeviceRadixSort / cub::DeviceReduce, library calls like the CUB scan of K5), and the run is
// kept on a stack of sorted runs that are merged pairwise while their sizes are within a factor of two (every entry
// takes part in O(log chunks) merges, not in one re-sort per chunk); a merge is a hand-written merge-path kernel pair
// (k3_merge_partition / k3_merge_tiles) followed by a reduce-by-key that adds up the keys both runs hold.  The
// final run is the result in (source_1, source_2) order.  Replaces the same reference lines as the dense accumulator:
// src/pairwise.cpp:222-225 (edges.try_emplace_l: insert w or add w, also when w == 0).
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_reduce.cuh>
#include <cub/iterator/transform_input_iterator.cuh>

#include "kernels.cuh"

namespace ksb {

struct U32ToU64 {
    __host__ __device__ __forceinline__ unsigned long long operator()(const uint32_t& v) const { return v; }
};
using WideIter = cub::TransformInputIterator<unsigned long long, U32ToU64, const uint32_t*>;

// temp bytes needed for n entries (max over the four CUB calls)
cudaError_t spill_temp_bytes(uint64_t n, int end_bit, size_t* bytes) {
    size_t a = 0, b = 0, c = 0, d = 0;
    unsigned long long* k = nullptr;
    uint32_t* v32 = nullptr;
    unsigned long long* v64 = nullptr;
    cudaError_t e = cub::DeviceRadixSort::SortPairs(nullptr, a, k, k, v32, v32, (int64_t)n, 0, end_bit);
    if (e != cudaSuccess) return e;
    e = cub::DeviceRadixSort::SortPairs(nullptr, b, k, k, v64, v64, (int64_t)n, 0, end_bit);
    if (e != cudaSuccess) return e;
    e = cub::DeviceReduce::ReduceByKey(nullptr, c, k, k, WideIter(v32, U32ToU64()), v64, k, cub::Sum(), (int64_t)n);
    if (e != cudaSuccess) return e;
    e = cub::DeviceReduce::ReduceByKey(nullptr, d, k, k, v64, v64, k, cub::Sum(), (int64_t)n);
    if (e != cudaSuccess) return e;
    *bytes = std::max(std::max(a, b), std::max(c, d));
    return cudaSuccess;
}

// chunk: (keys, u32 weights)[n] -> sorted unique keys + u64 sums; *d_runs = number of unique keys
cudaError_t spill_sort_reduce_chunk(void* tmp, size_t tmp_bytes, const unsigned long long* keys_in,
                                    unsigned long long* keys_sorted, const uint32_t* vals_in, uint32_t* vals_sorted,
                                    uint64_t n, int end_bit, unsigned long long* uniq, unsigned long long* sums,
                                    unsigned long long* d_runs, cudaStream_t st) {
    cudaError_t e = cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, keys_in, keys_sorted, vals_in, vals_sorted, (int64_t)n, 0,
                                                    end_bit, st);
    if (e != cudaSuccess) return e;
    return cub::DeviceReduce::ReduceByKey(tmp, tmp_bytes, keys_sorted, uniq, WideIter(vals_sorted, U32ToU64()), sums, d_runs,
                                          cub::Sum(), (int64_t)n, st);
}

// merge: concatenated runs (keys, u64 sums)[n] -> sorted unique keys + u64 sums
cudaError_t spill_sort_reduce_runs(void* tmp, size_t tmp_bytes, const unsigned long long* keys_in,
                                   unsigned long long* keys_sorted, const unsigned long long* sums_in,
                                   unsigned long long* sums_sorted, uint64_t n, int end_bit, unsigned long long* uniq,
                                   unsigned long long* sums, unsigned long long* d_runs, cudaStream_t st) {
    cudaError_t e = cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, keys_in, keys_sorted, sums_in, sums_sorted, (int64_t)n, 0,
                                                    end_bit, st);
    if (e != cudaSuccess) return e;
    return cub::DeviceReduce::ReduceByKey(tmp, tmp_bytes, keys_sorted, uniq, sums_sorted, sums, d_runs, cub::Sum(), (int64_t)n, st);
}


// ------------------------------------------------------------------------------------------------
// Merge path: two sorted runs of UNIQUE keys (A before B on equal keys) -> one sorted run of na + nb entries in which
// a key held by both runs appears twice, side by side (the reduce-by-key that follows adds them up).
//   k3_merge_partition  per tile of kMergeTile outputs: how many of the outputs before the tile come from A
//                       (binary search along the tile's diagonal);
//   k3_merge_tiles      one CTA per tile: the tile's slices of A and B staged in shared memory, every thread finds
//                       its own diagonal there and merges kMergePerThread outputs serially.
// ------------------------------------------------------------------------------------------------
constexpr uint32_t kMergeThreads = 256, kMergePerThread = 8, kMergeTile = kMergeThreads * kMergePerThread;

__global__ void __launch_bounds__(256) k3_merge_partition(const unsigned long long* __restrict__ ak, unsigned long long na,
                                                          const unsigned long long* __restrict__ bk, unsigned long long nb,
                                                          unsigned long long* __restrict__ part, unsigned long long n_tiles) {
    const unsigned long long t = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t > n_tiles) return;
    const unsigned long long total = na + nb;
    const unsigned long long d = min(t * kMergeTile, total);
    unsigned long long lo = d > nb ? d - nb : 0ull, hi = min(d, na);
    while (lo < hi) {
        const unsigned long long mid = (lo + hi) >> 1;
        if (__ldg(ak + mid) <= __ldg(bk + (d - 1 - mid))) lo = mid + 1; else hi = mid;
    }
    part[t] = lo;
}

__global__ void __launch_bounds__(kMergeThreads) k3_merge_tiles(const unsigned long long* __restrict__ ak, const unsigned long long* __restrict__ as,
                                                                unsigned long long na, const unsigned long long* __restrict__ bk,
                                                                const unsigned long long* __restrict__ bs, unsigned long long nb,
                                                                const unsigned long long* __restrict__ part,
                                                                unsigned long long* __restrict__ ok, unsigned long long* __restrict__ os) {
    __shared__ unsigned long long sk[kMergeTile], sv[kMergeTile];
    const unsigned long long t = blockIdx.x, total = na + nb;
    const unsigned long long d0 = t * kMergeTile, d1 = min(d0 + kMergeTile, total);
    const unsigned long long a0 = part[t], a1 = part[t + 1];
    const unsigned long long b0 = d0 - a0, b1 = d1 - a1;
    const uint32_t la = (uint32_t)(a1 - a0), lb = (uint32_t)(b1 - b0), n = la + lb;
    for (uint32_t i = threadIdx.x; i < n; i += kMergeThreads) {
        if (i < la) { sk[i] = __ldg(ak + a0 + i); sv[i] = __ldg(as + a0 + i); }
        else { sk[i] = __ldg(bk + b0 + (i - la)); sv[i] = __ldg(bs + b0 + (i - la)); }
    }
    __syncthreads();
    const uint32_t d = min(threadIdx.x * kMergePerThread, n);
    uint32_t lo = d > lb ? d - lb : 0u, hi = min(d, la);
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (sk[mid] <= sk[la + (d - 1 - mid)]) lo = mid + 1; else hi = mid;
    }
    uint32_t i = lo, j = d - lo;  // next element of the A slice / of the B slice
#pragma unroll
    for (uint32_t k = 0; k < kMergePerThread; k++) {
        const uint32_t o = d + k;
        if (o >= n) break;
        const bool take_a = j >= lb || (i < la && sk[i] <= sk[la + j]);
        const uint32_t src = take_a ? i : la + j;
        ok[d0 + o] = sk[src];
        os[d0 + o] = sv[src];
        if (take_a) i++; else j++;
    }
}

// out_k / out_s: na + nb entries (duplicates side by side); part: (na + nb) / kMergeTile + 2 entries of scratch
cudaError_t spill_merge_path(const unsigned long long* ak, const unsigned long long* as, uint64_t na, const unsigned long long* bk,
                             const unsigned long long* bs, uint64_t nb, unsigned long long* part, unsigned long long* out_k,
                             unsigned long long* out_s, cudaStream_t st) {
    const uint64_t total = na + nb;
    if (total == 0) return cudaSuccess;
    const uint64_t n_tiles = (total + kMergeTile - 1) / kMergeTile;
    k3_merge_partition<<<(unsigned)((n_tiles + 1 + 255) / 256), 256, 0, st>>>(ak, na, bk, nb, part, n_tiles);
    k3_merge_tiles<<<(unsigned)n_tiles, kMergeThreads, 0, st>>>(ak, as, na, bk, bs, nb, part, out_k, out_s);
    return cudaGetLastError();
}
uint64_t spill_merge_part_entries(uint64_t total) { return (total + kMergeTile - 1) / kMergeTile + 2; }

cudaError_t cub_reduce_bytes(uint64_t n, size_t* bytes) {
    unsigned long long* k = nullptr;
    return cub::DeviceReduce::ReduceByKey(nullptr, *bytes, k, k, k, k, k, cub::Sum(), (int64_t)n);
}

// adjacent equal keys of a merged run added up: (keys, sums)[n] -> unique keys + sums; *d_runs = their number
cudaError_t spill_reduce_sorted(void* tmp, size_t tmp_bytes, const unsigned long long* keys, const unsigned long long* sums_in, uint64_t n,
                                unsigned long long* uniq, unsigned long long* sums, unsigned long long* d_runs, cudaStream_t st) {
    return cub::DeviceReduce::ReduceByKey(tmp, tmp_bytes, keys, uniq, sums_in, sums, d_runs, cub::Sum(), (int64_t)n, st);
}

}  // namespace ksb
