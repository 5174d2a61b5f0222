// Device side of the FILE-driven consumers (table_reader.cpp parses the files):
//   ksb_cluster_table  `kSpider cluster` from the pairwise TSV on disk, or from its binary side-car
//                      (pykSpider/kSpider2/ks_clustering.py:63-120): node = id - 1, the edge is dropped iff
//                      float(text) * 100 < cut_off_threshold, connected components over ALL nodes.
//   ksb_ani_table      `kSpider pairwise --estimate-ani` from the TSV on disk (ks_pairwise.py:64-80).
// Same lock-free union-find as K6/K7 (kernels.cuh); label of a component = its smallest node index.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <mutex>

#include "kernels.cuh"
#include "table.hpp"

namespace ksb {

bool table_try_pinned(void** out, size_t bytes) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) { (void)cudaGetLastError(); return false; }
    if (cudaMallocHost(out, bytes) != cudaSuccess) { (void)cudaGetLastError(); return false; }
    return true;
}
void table_free_pinned(void* p) { cudaFreeHost(p); }

namespace {

// TEXT rows: keep iff !(v * 100.0 < cut) -- the Python expression on the parsed text, NaN kept.
// Side-car rows: keep iff !(c < cstar), cstar = smallest float32 whose "%g" text passes the same test
// (ksb_threshold_float_pct): identical decisions, because "%g" and strtod are monotone.
template <bool TEXT>
struct Rows {
    const uint32_t* s1;
    const uint32_t* s2;
    const double* dist;
    const float* cont;
    double cut;
    float cstar;
    __device__ __forceinline__ bool keep(unsigned long long r) const {
        if (TEXT) return !(__dmul_rn(__ldg(dist + r), 100.0) < cut);
        return !(__ldg(cont + r) < cstar);
    }
};

constexpr uint32_t kRowsPerThread = 8;

template <bool TEXT>
__global__ void __launch_bounds__(256) kt_hook_min(const Rows<TEXT> R, unsigned long long n_rows, uint32_t n_nodes,
                                                   uint32_t* parent, uint32_t* bad) {
    const unsigned long long span = 256ull * kRowsPerThread;
    for (unsigned long long base = (unsigned long long)blockIdx.x * span; base < n_rows; base += (unsigned long long)gridDim.x * span) {
        bool k[kRowsPerThread];
#pragma unroll
        for (uint32_t q = 0; q < kRowsPerThread; q++) {
            const unsigned long long r = base + q * 256u + threadIdx.x;
            k[q] = r < n_rows && R.keep(r);
        }
#pragma unroll
        for (uint32_t q = 0; q < kRowsPerThread; q++) {
            if (!k[q]) continue;
            const unsigned long long r = base + q * 256u + threadIdx.x;
            const uint32_t a = __ldg(R.s1 + r) - 1u, b = __ldg(R.s2 + r) - 1u;  // node = id - 1 (:75-76,99-100)
            if (a >= n_nodes || b >= n_nodes) { atomicMin(bad, (uint32_t)min(r, 0xFFFFFFFEull)); continue; }
            if (a != b) atomicMin(parent + max(a, b), min(a, b));
        }
    }
}

template <bool TEXT>
__global__ void __launch_bounds__(256) kt_union(const Rows<TEXT> R, unsigned long long n_rows, uint32_t n_nodes,
                                                uint32_t* parent, unsigned long long* n_edges) {
    __shared__ unsigned long long q[256 * (kRowsPerThread + 1)];
    __shared__ uint32_t qn;
    const uint32_t tid = threadIdx.x, lane = tid & 31u;
    if (tid == 0) qn = 0;
    __syncthreads();
    uint32_t kept = 0;
    const unsigned long long span = 256ull * kRowsPerThread;
    for (unsigned long long base = (unsigned long long)blockIdx.x * span; base < n_rows; base += (unsigned long long)gridDim.x * span) {
        bool k[kRowsPerThread];
#pragma unroll
        for (uint32_t x = 0; x < kRowsPerThread; x++) {
            const unsigned long long r = base + x * 256u + tid;
            k[x] = r < n_rows && R.keep(r);
        }
#pragma unroll
        for (uint32_t x = 0; x < kRowsPerThread; x++) {
            const unsigned long long r = base + x * 256u + tid;
            const uint32_t m = __ballot_sync(0xffffffffu, k[x]);
            if (m == 0u) continue;
            uint32_t pos = 0;
            if (lane == 0) pos = atomicAdd(&qn, (uint32_t)__popc(m));
            pos = __shfl_sync(0xffffffffu, pos, 0);
            if (k[x]) {
                const uint32_t a = __ldg(R.s1 + r) - 1u, b = __ldg(R.s2 + r) - 1u;
                // an endpoint outside the graph was reported by kt_hook_min; keep the queue dense with a no-op edge
                const bool ok = a < n_nodes && b < n_nodes;
                q[pos + __popc(m & ((1u << lane) - 1u))] = ok ? (((unsigned long long)a << 32) | b) : 0ull;
                kept++;
            }
        }
        __syncthreads();
        uint32_t n = qn;
        while (n >= 256u) {
            const unsigned long long ab = q[n - 256u + tid];
            uf_union(parent, (uint32_t)(ab >> 32), (uint32_t)ab);
            n -= 256u;
        }
        __syncthreads();
        if (tid == 0) qn = n;
        __syncthreads();
    }
    const uint32_t n = qn;
    if (tid < n) {
        const unsigned long long ab = q[tid];
        uf_union(parent, (uint32_t)(ab >> 32), (uint32_t)ab);
    }
    kept = __reduce_add_sync(0xffffffffu, kept);
    if (lane == 0 && kept) atomicAdd(n_edges, (unsigned long long)kept);
}

}  // namespace

// kernels.cu (K8 lives there with text6 / ani_of)
cudaError_t launch_k8_ani_cols(const double* dmin, const double* dmax, const float* fmin, const float* fmax, uint64_t n_rows,
                               double inv_k, double* out, unsigned long long* n_bad, const LaunchCfg& lc);

}  // namespace ksb

// ------------------------------------------------------------------------------------------------
// ksb_cluster_edges_noalloc: the threshold + connected-components step of ksb_cluster_table for callers
// that may not allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  No
// counterpart in the reference.  The union-find forest
// and the two counters -- what ksb_cluster_table takes from cudaMallocAsync -- live in ONE file-scope
// device array that every GPU loading the library holds until exit.  All calls on a device share it:
// the part a call uses is re-initialised at its start, and calls that overlap on the GPU race.
// ------------------------------------------------------------------------------------------------
namespace ksb {
namespace {
constexpr uint64_t kNoAllocMaxNodes = KSB_NOALLOC_MAX_NODES;
// [0, 256): counters {u64 n_edges, u32 bad row (0xFFFFFFFF = none)}; [256, ..): parent[kNoAllocMaxNodes]
static __device__ __align__(256) unsigned char g_noalloc_scratch[256 + 4 * kNoAllocMaxNodes];

struct NoAllocCtx {
    bool ready = false;
    int sms = 0;
    unsigned char* base = nullptr;  // DEVICE address of g_noalloc_scratch on this ordinal
};
NoAllocCtx g_noalloc[64];
std::mutex g_noalloc_mu;

__global__ void kt_noalloc_begin(uint32_t* parent, uint32_t n_nodes, unsigned long long* n_edges, uint32_t* bad) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_nodes) parent[i] = i;
    if (i == 0) { *n_edges = 0ull; *bad = 0xFFFFFFFFu; }
}
// compress the forest and hand labels + counters to the caller's buffers in the same pass
__global__ void kt_noalloc_finish(uint32_t* parent, uint32_t n_nodes, const unsigned long long* n_edges, const uint32_t* bad,
                                  uint32_t* labels_out, unsigned long long* counters_out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_nodes) labels_out[i] = uf_find(parent, i);
    if (i == 0 && counters_out) { counters_out[0] = *n_edges; counters_out[1] = (unsigned long long)*bad; }
}
}  // namespace
}  // namespace ksb

using namespace ksb;

extern "C" {

float ksb_threshold_float_pct(double cut_pct);

int ksb_cluster_table(const ksb_table* t, uint64_t n_nodes, double cut_pct, int device, uint32_t* labels_out, uint64_t* n_edges) {
    if (!t || (!labels_out && n_nodes)) { set_error("ksb_cluster_table: null argument"); return KSB_ERR_ARG; }
    if (n_nodes >= 0xFFFFFFFFull) { set_error("ksb_cluster_table: too many nodes"); return KSB_ERR_ARG; }
    if (n_edges) *n_edges = 0;
    int sms = 0;
    const int dev = check_device(device, &sms);
    if (dev < 0) return dev;
    if (n_nodes == 0 && t->n == 0) return KSB_OK;
    const bool text = t->dist != nullptr;
    cudaStream_t st = service_stream(dev);
    const LaunchCfg lc{sms, st};
    uint32_t *d_parent = nullptr, *d_bad = nullptr, *d_s1 = nullptr, *d_s2 = nullptr;
    unsigned long long* d_ne = nullptr;
    void* d_val = nullptr;
    cudaError_t e = cudaSuccess;
    int rc = KSB_OK;
    // rows go through the device in slabs: the table may be larger than what is comfortable to hold twice
    const uint64_t slab = std::min<uint64_t>(std::max<uint64_t>(t->n, 1), 64ull << 20);
    const size_t vbytes = text ? 8 : 4;
    float cstar = text ? 0.f : ksb_threshold_float_pct(cut_pct);
    if (cstar != cstar) cstar = INFINITY;
    do {
        if ((e = cudaMallocAsync(&d_parent, 4 * std::max<uint64_t>(n_nodes, 1), st)) != cudaSuccess) break;
        if ((e = cudaMallocAsync(&d_bad, 4, st)) != cudaSuccess) break;
        if ((e = cudaMallocAsync(&d_ne, 8, st)) != cudaSuccess) break;
        if ((e = cudaMallocAsync(&d_s1, 4 * slab, st)) != cudaSuccess) break;
        if ((e = cudaMallocAsync(&d_s2, 4 * slab, st)) != cudaSuccess) break;
        if ((e = cudaMallocAsync(&d_val, vbytes * slab, st)) != cudaSuccess) break;
        if ((e = cudaMemsetAsync(d_bad, 0xFF, 4, st)) != cudaSuccess) break;
        if ((e = cudaMemsetAsync(d_ne, 0, 8, st)) != cudaSuccess) break;
        if ((e = launch_k7_init(d_parent, (uint32_t)std::max<uint64_t>(n_nodes, 1), lc)) != cudaSuccess) break;
        const unsigned cap = (unsigned)sms * 8u;
        // pass 1 over all slabs: atomicMin hooks (accelerator, see k6_hook_min); pass 2: the unions
        for (int pass = 0; pass < 2 && e == cudaSuccess; pass++) {
            for (uint64_t r0 = 0; r0 < t->n && e == cudaSuccess; r0 += slab) {
                const uint64_t n = std::min(slab, t->n - r0);
                if (pass == 0 || t->n > slab) {  // a table of one slab stays resident between the passes
                    if ((e = cudaMemcpyAsync(d_s1, t->s1 + r0, 4 * n, cudaMemcpyHostToDevice, st)) != cudaSuccess) break;
                    if ((e = cudaMemcpyAsync(d_s2, t->s2 + r0, 4 * n, cudaMemcpyHostToDevice, st)) != cudaSuccess) break;
                    const void* hv = text ? (const void*)(t->dist + r0) : (const void*)(t->cont + r0);
                    if ((e = cudaMemcpyAsync(d_val, hv, vbytes * n, cudaMemcpyHostToDevice, st)) != cudaSuccess) break;
                }
                const unsigned ctas = (unsigned)std::min<uint64_t>(cap, (n + 2047) / 2048);
                if (text) {
                    const Rows<true> R{d_s1, d_s2, static_cast<const double*>(d_val), nullptr, cut_pct, 0.f};
                    if (pass == 0) kt_hook_min<true><<<ctas, 256, 0, st>>>(R, n, (uint32_t)n_nodes, d_parent, d_bad);
                    else kt_union<true><<<ctas, 256, 0, st>>>(R, n, (uint32_t)n_nodes, d_parent, d_ne);
                } else {
                    const Rows<false> R{d_s1, d_s2, nullptr, static_cast<const float*>(d_val), cut_pct, cstar};
                    if (pass == 0) kt_hook_min<false><<<ctas, 256, 0, st>>>(R, n, (uint32_t)n_nodes, d_parent, d_bad);
                    else kt_union<false><<<ctas, 256, 0, st>>>(R, n, (uint32_t)n_nodes, d_parent, d_ne);
                }
                e = cudaGetLastError();
            }
        }
        if (e != cudaSuccess) break;
        if (n_nodes && (e = launch_k7_compress(d_parent, (uint32_t)n_nodes, lc)) != cudaSuccess) break;
        uint32_t bad = 0xFFFFFFFFu;
        unsigned long long ne = 0;
        if (n_nodes && (e = cudaMemcpyAsync(labels_out, d_parent, 4 * n_nodes, cudaMemcpyDeviceToHost, st)) != cudaSuccess) break;
        if ((e = cudaMemcpyAsync(&bad, d_bad, 4, cudaMemcpyDeviceToHost, st)) != cudaSuccess) break;
        if ((e = cudaMemcpyAsync(&ne, d_ne, 8, cudaMemcpyDeviceToHost, st)) != cudaSuccess) break;
        if ((e = cudaStreamSynchronize(st)) != cudaSuccess) break;
        if (bad != 0xFFFFFFFFu) {
            // rustworkx: add_edges_from with a node index that was never added fails the whole command
            const uint64_t br = t->n <= slab ? bad : 0;  // the row index is slab-relative: exact only for one-slab tables
            set_error("ksb_cluster_table: a kept row (e.g. %u %u) names a sample id outside 1..%llu, the nodes the namesMap gives the graph",
                      t->s1[br], t->s2[br], (unsigned long long)n_nodes);
            rc = KSB_ERR_FORMAT;
            break;
        }
        if (n_edges) *n_edges = ne;
    } while (0);
    if (e != cudaSuccess) { set_error("ksb_cluster_table: %s", cudaGetErrorString(e)); (void)cudaGetLastError(); rc = KSB_ERR_CUDA; }
    for (void* p : {(void*)d_parent, (void*)d_bad, (void*)d_ne, (void*)d_s1, (void*)d_s2, d_val})
        if (p) cudaFreeAsync(p, st);
    cudaStreamSynchronize(st);
    return rc;
}

int ksb_ani_table(const ksb_table* t, int ksize, int device, double* ani_out, uint64_t* n_invalid) {
    if (!t || (!ani_out && t->n)) { set_error("ksb_ani_table: null argument"); return KSB_ERR_ARG; }
    if (ksize < 1) { set_error("ksb_ani_table: ksize must be positive"); return KSB_ERR_ARG; }
    const bool text = t->dist != nullptr;
    if ((text && !t->dist2) || (!text && !t->cont2) || t->dist_col != 3 || t->dist_col2 != 5) {
        set_error("ksb_ani_table: the table must hold columns 3 (min) and 5 (max) (ksb_table_read_*(.., 3, 5, ..))");
        return KSB_ERR_ARG;
    }
    if (n_invalid) *n_invalid = 0;
    int sms = 0;
    const int dev = check_device(device, &sms);
    if (dev < 0) return dev;
    if (t->n == 0) return KSB_OK;
    cudaStream_t st = service_stream(dev);
    const LaunchCfg lc{sms, st};
    const uint64_t slab = std::min<uint64_t>(t->n, 64ull << 20);
    const size_t vb = text ? 8 : 4;
    void *d_a = nullptr, *d_b = nullptr;
    double* d_out = nullptr;
    unsigned long long* d_bad = nullptr;
    cudaError_t e = cudaSuccess;
    int rc = KSB_OK;
    do {
        if ((e = cudaMallocAsync(&d_a, vb * slab, st)) != cudaSuccess) break;
        if ((e = cudaMallocAsync(&d_b, vb * slab, st)) != cudaSuccess) break;
        if ((e = cudaMallocAsync(&d_out, 8 * slab, st)) != cudaSuccess) break;
        if ((e = cudaMallocAsync(&d_bad, 8, st)) != cudaSuccess) break;
        if ((e = cudaMemsetAsync(d_bad, 0, 8, st)) != cudaSuccess) break;
        for (uint64_t r0 = 0; r0 < t->n && e == cudaSuccess; r0 += slab) {
            const uint64_t n = std::min(slab, t->n - r0);
            const void* ha = text ? (const void*)(t->dist + r0) : (const void*)(t->cont + r0);
            const void* hb = text ? (const void*)(t->dist2 + r0) : (const void*)(t->cont2 + r0);
            if ((e = cudaMemcpyAsync(d_a, ha, vb * n, cudaMemcpyHostToDevice, st)) != cudaSuccess) break;
            if ((e = cudaMemcpyAsync(d_b, hb, vb * n, cudaMemcpyHostToDevice, st)) != cudaSuccess) break;
            e = text ? launch_k8_ani_cols(static_cast<const double*>(d_a), static_cast<const double*>(d_b), nullptr, nullptr, n, 1.0 / (double)ksize, d_out, d_bad, lc)
                     : launch_k8_ani_cols(nullptr, nullptr, static_cast<const float*>(d_a), static_cast<const float*>(d_b), n, 1.0 / (double)ksize, d_out, d_bad, lc);
            if (e != cudaSuccess) break;
            if ((e = cudaMemcpyAsync(ani_out + r0, d_out, 8 * n, cudaMemcpyDeviceToHost, st)) != cudaSuccess) break;
            if ((e = cudaStreamSynchronize(st)) != cudaSuccess) break;  // ani_out may be pageable: finish before the next slab reuses d_out
        }
        if (e != cudaSuccess) break;
        unsigned long long nb = 0;
        if ((e = cudaMemcpyAsync(&nb, d_bad, 8, cudaMemcpyDeviceToHost, st)) != cudaSuccess) break;
        if ((e = cudaStreamSynchronize(st)) != cudaSuccess) break;
        if (n_invalid) *n_invalid = nb;
    } while (0);
    if (e != cudaSuccess) { set_error("ksb_ani_table: %s", cudaGetErrorString(e)); (void)cudaGetLastError(); rc = KSB_ERR_CUDA; }
    for (void* p : {d_a, d_b, (void*)d_out, (void*)d_bad})
        if (p) cudaFreeAsync(p, st);
    cudaStreamSynchronize(st);
    return rc;
}


int ksb_cluster_edges_noalloc(const uint32_t* d_source_1, const uint32_t* d_source_2, const float* d_cont, uint64_t n_rows,
                              uint64_t n_nodes, double cut_pct, void* stream, uint32_t* d_labels_out, uint64_t* d_counters_out) {
    if (n_nodes > kNoAllocMaxNodes) {
        set_error("ksb_cluster_edges_noalloc: %llu nodes exceed the static scratch of this build (KSB_NOALLOC_MAX_NODES = %llu); "
                  "use ksb_cluster_table", (unsigned long long)n_nodes, (unsigned long long)kNoAllocMaxNodes);
        return KSB_ERR_ARG;
    }
    if ((n_rows && (!d_source_1 || !d_source_2 || !d_cont)) || (n_nodes && !d_labels_out)) {
        set_error("ksb_cluster_edges_noalloc: null argument");
        return KSB_ERR_ARG;
    }
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= 64) {
        set_error("ksb_cluster_edges_noalloc: no usable current device (%s)", cudaGetErrorString(e));
        (void)cudaGetLastError();
        return KSB_ERR_CUDA;
    }
    NoAllocCtx& c = g_noalloc[dev];
    {
        // set-up, once per device, with that device current: in host code the array's name is only a stand-in,
        // the kernels need the address the symbol has on THIS device
        std::lock_guard<std::mutex> lk(g_noalloc_mu);
        if (!c.ready) {
            int major = 0;
            if ((e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev)) == cudaSuccess && major != 10) {
                set_error("ksb_cluster_edges_noalloc: device %d is sm_%dx; this library is built for sm_100a only", dev, major);
                return KSB_ERR_CUDA;
            }
            if (e == cudaSuccess) e = cudaDeviceGetAttribute(&c.sms, cudaDevAttrMultiProcessorCount, dev);
            void* p = nullptr;
            if (e == cudaSuccess) e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
            if (e != cudaSuccess || !p) {
                set_error("ksb_cluster_edges_noalloc: cudaGetSymbolAddress(scratch) on device %d: %s", dev, cudaGetErrorString(e));
                (void)cudaGetLastError();
                return KSB_ERR_CUDA;
            }
            c.base = static_cast<unsigned char*>(p);
            c.ready = true;
        }
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    unsigned long long* d_ne = reinterpret_cast<unsigned long long*>(c.base);
    uint32_t* d_bad = reinterpret_cast<uint32_t*>(c.base + 8);
    uint32_t* d_parent = reinterpret_cast<uint32_t*>(c.base + 256);
    float cstar = ksb_threshold_float_pct(cut_pct);
    if (cstar != cstar) cstar = INFINITY;
    const unsigned init_ctas = (unsigned)std::max<uint64_t>(1, (n_nodes + 255) / 256);
    kt_noalloc_begin<<<init_ctas, 256, 0, st>>>(d_parent, (uint32_t)n_nodes, d_ne, d_bad);
    if (n_rows) {
        const unsigned ctas = (unsigned)std::min<uint64_t>((uint64_t)c.sms * 8u, (n_rows + 2047) / 2048);
        const Rows<false> R{d_source_1, d_source_2, nullptr, d_cont, cut_pct, cstar};
        kt_hook_min<false><<<ctas, 256, 0, st>>>(R, n_rows, (uint32_t)n_nodes, d_parent, d_bad);
        kt_union<false><<<ctas, 256, 0, st>>>(R, n_rows, (uint32_t)n_nodes, d_parent, d_ne);
    }
    kt_noalloc_finish<<<init_ctas, 256, 0, st>>>(d_parent, (uint32_t)n_nodes, d_ne, d_bad, d_labels_out,
                                                  reinterpret_cast<unsigned long long*>(d_counters_out));
    if ((e = cudaGetLastError()) != cudaSuccess) {
        set_error("ksb_cluster_edges_noalloc: %s", cudaGetErrorString(e));
        return KSB_ERR_CUDA;
    }
    return KSB_OK;
}

}  // extern "C"
