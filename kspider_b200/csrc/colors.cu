// Colour construction on the GPU: from per-sample hash lists to the colour table
// the pairwise pass consumes.
//
// Replaces the END STATE of the reference's incremental colouring loop
//   src/sourmash_indexing.cpp:141-277 (and src/index.cpp:189-331 for k-mers): every hash ends up with
//   the colour whose member set is exactly the set of samples that contain it, and
//   colorsCount[colour] = number of distinct hashes with that set (:262-266, :250-252);
//   groupName_to_kmerCount[sample] = mins.size() (:205).
// The reference reaches that state one hash at a time through string-keyed maps (tagsMap, legend);
// here it is one sort by (hash, sample), a segmentation by hash, a second sort of the segments by a
// 64-bit signature of their sample lists, and an exact run-length grouping (adjacent segments are
// compared member by member, so two different sets are never merged; a signature collision can only
// split one set into two colours, which the pairwise pass adds up again).  Colour ids are not
// reproduced (they depend on the reference's processing order and id recycling, :214-222) and do not
// influence any output of the pairwise path.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>
#include <cub/iterator/counting_input_iterator.cuh>

#include <algorithm>
#include <memory>
#include <new>
#include <vector>

#include "kernels.cuh"

struct ksb_colors {
    std::vector<uint64_t> offsets;  // [C+1]
    std::vector<uint32_t> members;  // sample POSITIONS (0..n_samples-1), ascending inside a colour
    std::vector<uint64_t> counts;   // [C] hashes with exactly this member set
    std::vector<uint64_t> totals;   // [n_samples] hashes listed for the sample (duplicates included, like mins.size())
    uint64_t n_hashes = 0;          // distinct hashes
    float sort_ms = 0.f, total_ms = 0.f;
};

namespace ksb {
namespace {

__device__ __forceinline__ unsigned long long cmix(unsigned long long x) {
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
    return x;
}

// entry i of the concatenated hash lists belongs to the sample whose offset range contains i
__global__ void kc_expand(const unsigned long long* __restrict__ offsets, uint32_t n_samples, unsigned long long n,
                          uint32_t* __restrict__ sample) {
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        uint32_t lo = 0, hi = n_samples;  // last s with offsets[s] <= i
        while (hi - lo > 1) {
            const uint32_t mid = (lo + hi) >> 1;
            if (__ldg(offsets + mid) <= i) lo = mid; else hi = mid;
        }
        sample[i] = lo;
    }
}

// after the sort by hash (stable: samples ascending inside a hash): keep = not a repeat of (hash, sample)
__global__ void kc_keep(const unsigned long long* __restrict__ h, const uint32_t* __restrict__ s, unsigned long long n,
                        uint8_t* __restrict__ keep) {
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        keep[i] = (i == 0 || h[i] != h[i - 1] || s[i] != s[i - 1]) ? 1 : 0;
}

__global__ void kc_heads(const unsigned long long* __restrict__ h, unsigned long long n, uint8_t* __restrict__ head) {
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        head[i] = (i == 0 || h[i] != h[i - 1]) ? 1 : 0;
}

__global__ void kc_iota(uint32_t* __restrict__ v, unsigned long long n) {
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) v[i] = (uint32_t)i;
}

// signature of a segment's sample list (commutative sum of mixes + length)
__global__ void kc_signature(const unsigned long long* __restrict__ seg_start, unsigned long long n_seg,
                             const uint32_t* __restrict__ s, unsigned long long* __restrict__ sig) {
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long g = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; g < n_seg; g += stride) {
        const unsigned long long b = seg_start[g], e = seg_start[g + 1];
        unsigned long long a = cmix(e - b);
        for (unsigned long long i = b; i < e; i++) a += cmix(0x9E3779B97F4A7C15ULL + s[i]);
        sig[g] = a;
    }
}

// order[k] = k-th segment by signature; head2[k] = its sample list differs from the previous one's
__global__ void kc_group(const unsigned long long* __restrict__ seg_start, const unsigned long long* __restrict__ sig_sorted,
                         const uint32_t* __restrict__ order, unsigned long long n_seg, const uint32_t* __restrict__ s,
                         uint8_t* __restrict__ head2) {
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; k < n_seg; k += stride) {
        bool differs = (k == 0) || sig_sorted[k] != sig_sorted[k - 1];
        if (!differs) {
            const unsigned long long b = seg_start[order[k]], e = seg_start[order[k] + 1];
            const unsigned long long pb = seg_start[order[k - 1]], pe = seg_start[order[k - 1] + 1];
            differs = (e - b) != (pe - pb);
            for (unsigned long long i = 0; !differs && i < e - b; i++) differs = s[b + i] != s[pb + i];
        }
        head2[k] = differs ? 1 : 0;
    }
}

// colour c = run [cstart[c], cstart[c+1]) of the sorted segments; its members are those of the run's first segment
__global__ void kc_sizes(const unsigned long long* __restrict__ cstart, unsigned long long n_col,
                         const uint32_t* __restrict__ order, const unsigned long long* __restrict__ seg_start,
                         unsigned long long* __restrict__ size, unsigned long long* __restrict__ count) {
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long c = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; c < n_col; c += stride) {
        const uint32_t g = order[cstart[c]];
        size[c] = seg_start[g + 1] - seg_start[g];
        count[c] = cstart[c + 1] - cstart[c];
    }
}

__global__ void kc_emit(const unsigned long long* __restrict__ cstart, unsigned long long n_col,
                        const uint32_t* __restrict__ order, const unsigned long long* __restrict__ seg_start,
                        const unsigned long long* __restrict__ coff, const uint32_t* __restrict__ s,
                        uint32_t* __restrict__ members) {
    const uint32_t lane = threadIdx.x & 31u;
    const unsigned long long warp = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const unsigned long long nwarps = ((unsigned long long)gridDim.x * blockDim.x) >> 5;
    for (unsigned long long c = warp; c < n_col; c += nwarps) {
        const uint32_t g = order[cstart[c]];
        const unsigned long long b = seg_start[g], n = seg_start[g + 1] - b, o = coff[c];
        for (unsigned long long i = lane; i < n; i += 32) members[o + i] = s[b + i];
    }
}

}  // namespace
}  // namespace ksb

namespace {

// device buffers of one build, freed together
struct Pool {
    std::vector<void*> ptrs;
    ~Pool() { for (void* p : ptrs) cudaFree(p); }
    template <class T>
    cudaError_t get(T** out, uint64_t n) {
        void* p = nullptr;
        const cudaError_t e = cudaMalloc(&p, std::max<uint64_t>(n, 1) * sizeof(T));
        if (e == cudaSuccess) ptrs.push_back(p);
        *out = static_cast<T*>(p);
        return e;
    }
};

inline unsigned grid_for(uint64_t n, int sms) {
    const uint64_t want = (n + 255) / 256, cap = (uint64_t)sms * 8;
    return (unsigned)std::max<uint64_t>(1, std::min(want, cap));
}

}  // namespace

#define KC(call)                                                                                         \
    do {                                                                                                 \
        cudaError_t e_ = (call);                                                                         \
        if (e_ != cudaSuccess) {                                                                         \
            ksb::set_error("ksb_colors_build: %s failed: %s", #call, cudaGetErrorString(e_));            \
            (void)cudaGetLastError();                                                                    \
            return KSB_ERR_CUDA;                                                                         \
        }                                                                                                \
    } while (0)

extern "C" {

int ksb_colors_build(uint64_t n_samples, const uint64_t* offsets, const uint64_t* hashes, int device, ksb_colors** out) {
    using namespace ksb;
    if (!out || !offsets || (n_samples && offsets[n_samples] && !hashes)) { set_error("ksb_colors_build: null argument"); return KSB_ERR_ARG; }
    *out = nullptr;
    if (n_samples > 0x7FFFFFFFull) { set_error("ksb_colors_build: more than 2^31 samples"); return KSB_ERR_ARG; }
    for (uint64_t s = 0; s < n_samples; s++)
        if (offsets[s + 1] < offsets[s]) { set_error("ksb_colors_build: offsets not monotone at sample %llu", (unsigned long long)s); return KSB_ERR_ARG; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { (void)cudaGetLastError(); set_error("ksb_colors_build: no CUDA device (libkspider_b200 has no CPU fallback)"); return KSB_ERR_CUDA; }
    if (device < 0) KC(cudaGetDevice(&device));
    KC(cudaSetDevice(device));
    int sms = 148;
    KC(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    ksb_colors* col = new (std::nothrow) ksb_colors();
    if (!col) { set_error("out of memory"); return KSB_ERR_NOMEM; }
    std::unique_ptr<ksb_colors> guard(col);
    col->totals.resize(n_samples);
    for (uint64_t s = 0; s < n_samples; s++) col->totals[s] = offsets[s + 1] - offsets[s];
    col->offsets.assign(1, 0);
    const uint64_t M = n_samples ? offsets[n_samples] - offsets[0] : 0;
    if (M == 0) { *out = guard.release(); return KSB_OK; }
    if (M > 0xFFFFFFFFull) { set_error("ksb_colors_build: more than 2^32 hash entries in one call are not supported"); return KSB_ERR_ARG; }

    Pool pool;
    cudaEvent_t ev[3];
    for (auto& e : ev) KC(cudaEventCreate(&e));
    struct EvGuard { cudaEvent_t* e; ~EvGuard() { for (int i = 0; i < 3; i++) cudaEventDestroy(e[i]); } } evg{ev};
    unsigned long long *d_off, *d_h, *d_h2;
    uint32_t *d_s, *d_s2;
    uint8_t* d_flag;
    unsigned long long* d_num;
    KC(pool.get(&d_off, n_samples + 1));
    KC(pool.get(&d_h, M));
    KC(pool.get(&d_h2, M + 1));
    KC(pool.get(&d_s, M));
    KC(pool.get(&d_s2, M));
    KC(pool.get(&d_flag, M));
    KC(pool.get(&d_num, 4));
    KC(cudaEventRecord(ev[0]));
    {
        std::vector<unsigned long long> rel(n_samples + 1);
        for (uint64_t s = 0; s <= n_samples; s++) rel[s] = offsets[s] - offsets[0];
        KC(cudaMemcpy(d_off, rel.data(), 8 * (n_samples + 1), cudaMemcpyHostToDevice));
    }
    KC(cudaMemcpy(d_h, hashes + offsets[0], 8 * M, cudaMemcpyHostToDevice));
    kc_expand<<<grid_for(M, sms), 256>>>(d_off, (uint32_t)n_samples, M, d_s);
    KC(cudaGetLastError());

    // ---- sort by hash (stable: samples stay ascending inside a hash)
    size_t tb = 0, tb2 = 0;
    void* d_tmp = nullptr;
    KC(cub::DeviceRadixSort::SortPairs(nullptr, tb, d_h, d_h2, d_s, d_s2, (int64_t)M));
    KC(cub::DeviceSelect::Flagged(nullptr, tb2, d_h2, d_flag, d_h, d_num, (int64_t)M));
    tb = std::max(tb, tb2);
    KC(cub::DeviceSelect::Flagged(nullptr, tb2, cub::CountingInputIterator<unsigned long long>(0ull), d_flag, d_h, d_num, (int64_t)M));
    tb = std::max(tb, tb2);
    KC(cub::DeviceScan::ExclusiveSum(nullptr, tb2, d_h, d_h2, (int64_t)M + 1));
    tb = std::max(tb, tb2);
    KC(pool.get(reinterpret_cast<uint8_t**>(&d_tmp), tb + 256));
    KC(cudaEventRecord(ev[1]));
    KC(cub::DeviceRadixSort::SortPairs(d_tmp, tb, d_h, d_h2, d_s, d_s2, (int64_t)M));
    KC(cudaEventRecord(ev[2]));
    // ---- drop repeated (hash, sample) entries: sorted pairs now in (d_h2, d_s2) -> (d_h, d_s)
    kc_keep<<<grid_for(M, sms), 256>>>(d_h2, d_s2, M, d_flag);
    KC(cudaGetLastError());
    KC(cub::DeviceSelect::Flagged(d_tmp, tb, d_h2, d_flag, d_h, d_num, (int64_t)M));
    KC(cub::DeviceSelect::Flagged(d_tmp, tb, d_s2, d_flag, d_s, d_num, (int64_t)M));
    unsigned long long M2 = 0;
    KC(cudaMemcpy(&M2, d_num, 8, cudaMemcpyDeviceToHost));
    // ---- segments = distinct hashes
    kc_heads<<<grid_for(M2, sms), 256>>>(d_h, M2, d_flag);
    KC(cudaGetLastError());
    unsigned long long* d_seg = d_h2;  // segment starts (positions), reuses the sorted-key buffer
    KC(cub::DeviceSelect::Flagged(d_tmp, tb, cub::CountingInputIterator<unsigned long long>(0ull), d_flag, d_seg, d_num, (int64_t)M2));
    unsigned long long n_seg = 0;
    KC(cudaMemcpy(&n_seg, d_num, 8, cudaMemcpyDeviceToHost));
    KC(cudaMemcpy(d_seg + n_seg, &M2, 8, cudaMemcpyHostToDevice));  // d_h2 has M >= n_seg + 1 entries unless every entry is its own segment
    col->n_hashes = n_seg;

    // ---- signature of every segment's sample list, segments sorted by it
    unsigned long long *d_sig, *d_sig2;
    uint32_t *d_ord, *d_ord2;
    KC(pool.get(&d_sig, n_seg));
    KC(pool.get(&d_sig2, n_seg));
    KC(pool.get(&d_ord, n_seg));
    KC(pool.get(&d_ord2, n_seg));
    kc_signature<<<grid_for(n_seg, sms), 256>>>(d_seg, n_seg, d_s, d_sig);
    KC(cudaGetLastError());
    kc_iota<<<grid_for(n_seg, sms), 256>>>(d_ord, n_seg);
    KC(cudaGetLastError());
    {
        size_t need = 0;
        KC(cub::DeviceRadixSort::SortPairs(nullptr, need, d_sig, d_sig2, d_ord, d_ord2, (int64_t)n_seg));
        void* t2 = d_tmp;
        if (need > tb) { uint8_t* extra = nullptr; KC(pool.get(&extra, need + 256)); t2 = extra; }
        size_t have = std::max(need, tb);
        KC(cub::DeviceRadixSort::SortPairs(t2, have, d_sig, d_sig2, d_ord, d_ord2, (int64_t)n_seg));
    }
    // ---- exact grouping of equal sample lists -> colours
    kc_group<<<grid_for(n_seg, sms), 256>>>(d_seg, d_sig2, d_ord2, n_seg, d_s, d_flag);
    KC(cudaGetLastError());
    unsigned long long* d_cstart = nullptr;  // run starts (positions in the sorted segment order)
    KC(pool.get(&d_cstart, n_seg + 1));
    KC(cub::DeviceSelect::Flagged(d_tmp, tb, cub::CountingInputIterator<unsigned long long>(0ull), d_flag, d_cstart, d_num, (int64_t)n_seg));
    unsigned long long n_col = 0;
    KC(cudaMemcpy(&n_col, d_num, 8, cudaMemcpyDeviceToHost));
    KC(cudaMemcpy(d_cstart + n_col, &n_seg, 8, cudaMemcpyHostToDevice));
    unsigned long long *d_size, *d_coff, *d_count;
    KC(pool.get(&d_size, n_col + 1));
    KC(pool.get(&d_coff, n_col + 1));
    KC(pool.get(&d_count, n_col));
    KC(cudaMemset(d_size + n_col, 0, 8));
    kc_sizes<<<grid_for(n_col, sms), 256>>>(d_cstart, n_col, d_ord2, d_seg, d_size, d_count);
    KC(cudaGetLastError());
    {
        size_t need = 0;
        KC(cub::DeviceScan::ExclusiveSum(nullptr, need, d_size, d_coff, (int64_t)n_col + 1));
        void* t2 = d_tmp;
        if (need > tb) { uint8_t* extra = nullptr; KC(pool.get(&extra, need + 256)); t2 = extra; }
        size_t have = std::max(need, tb);
        KC(cub::DeviceScan::ExclusiveSum(t2, have, d_size, d_coff, (int64_t)n_col + 1));
    }
    uint32_t* d_members = d_s2;  // M2 entries at most (every kept entry belongs to exactly one segment; colours keep one segment each)
    kc_emit<<<grid_for(n_col * 32, sms), 256>>>(d_cstart, n_col, d_ord2, d_seg, d_coff, d_s, d_members);
    KC(cudaGetLastError());
    col->offsets.resize(n_col + 1);
    col->counts.resize(n_col);
    KC(cudaMemcpy(col->offsets.data(), d_coff, 8 * (n_col + 1), cudaMemcpyDeviceToHost));
    KC(cudaMemcpy(col->counts.data(), d_count, 8 * n_col, cudaMemcpyDeviceToHost));
    col->members.resize(col->offsets[n_col]);
    KC(cudaMemcpy(col->members.data(), d_members, 4 * col->offsets[n_col], cudaMemcpyDeviceToHost));
    cudaEvent_t ev_end;
    KC(cudaEventCreate(&ev_end));
    KC(cudaEventRecord(ev_end));
    KC(cudaEventSynchronize(ev_end));
    cudaEventElapsedTime(&col->sort_ms, ev[1], ev[2]);
    cudaEventElapsedTime(&col->total_ms, ev[0], ev_end);
    cudaEventDestroy(ev_end);
    *out = guard.release();
    return KSB_OK;
}

uint64_t ksb_colors_count(const ksb_colors* c) { return c ? c->counts.size() : 0; }
uint64_t ksb_colors_distinct_hashes(const ksb_colors* c) { return c ? c->n_hashes : 0; }
const uint64_t* ksb_colors_offsets(const ksb_colors* c) { return c ? c->offsets.data() : nullptr; }
const uint32_t* ksb_colors_members(const ksb_colors* c) { return c ? c->members.data() : nullptr; }
const uint64_t* ksb_colors_counts(const ksb_colors* c) { return c ? c->counts.data() : nullptr; }
const uint64_t* ksb_colors_totals(const ksb_colors* c) { return c ? c->totals.data() : nullptr; }
float ksb_colors_device_ms(const ksb_colors* c) { return c ? c->total_ms : 0.f; }
void ksb_colors_free(ksb_colors* c) { delete c; }

}  // extern "C"
