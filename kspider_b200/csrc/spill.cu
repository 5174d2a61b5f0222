// K3: sort-reduce spill path.  Used when the owned block of the dense triangle does not fit the
// budget (or is forced): the pair-emission kernels append (key = row << shift | column, weight)
// to a chunk buffer instead of doing reds; every chunk is radix-sorted and reduced by key
// (cub::DeviceRadixSort / cub::DeviceReduce, library calls like the CUB scan of K5), and the run is
// merged into the accumulated run by one more sort + reduce.  The final run is the result in
// (source_1, source_2) order.  Replaces the same reference lines as the dense accumulator:
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

}  // namespace ksb
