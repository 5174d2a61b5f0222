// Device half of the C-ABI (include/kspider_b200.h): resident CSR, owned triangle block(s),
// accumulation, post-pass, clustering.  No CPU fallback: every entry point here needs a CUDA device.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <new>
#include <vector>

#include <sys/mman.h>

#include "kernels.cuh"

namespace ksb {
cudaError_t scan_block_counts(void* temp, size_t* temp_bytes, const unsigned long long* in, unsigned long long* out,
                              uint64_t n, cudaStream_t stream);
}

#define KSB_CUDA(call)                                                                               \
    do {                                                                                             \
        cudaError_t e_ = (call);                                                                     \
        if (e_ != cudaSuccess) {                                                                     \
            ksb::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
            (void)cudaGetLastError();                                                                \
            return KSB_ERR_CUDA;                                                                     \
        }                                                                                            \
    } while (0)

// One dense pass: rows [rb, re) of the triangle are accumulated and scanned together.
struct ksb_pass {
    uint64_t rb, re, base, cells, updates;
};

// One K4 launch: (sub-)windows whose bitmaps share the chunk buffer, and the work items over them.
struct ksb_gram_chunk {
    uint64_t c_begin = 0, n_colors = 0;   // contiguous range of the Gram CSR
    uint64_t bits_rows = 0;               // 16-byte bitmap rows of the chunk
    uint64_t blob_row0 = 0;               // bitmap form of the blob (IndexDesc::gram_form 1): the chunk's first row in the blob's bitmap section
    uint32_t win_begin = 0, n_wins = 0;   // into the plan's window array
    uint32_t item_begin = 0, n_items = 0; // into the plan's item array
    uint32_t kblocks = 0;                 // K-blocks of all windows = pitch of the weight planes / 128
    uint32_t n_units = 0;                 // bitmap-build units (K-block x sample segment)
    uint32_t unit_begin = 0;              // into the plan's unit array
    bool pipe = false;                    // short items: k4_gram_pipe (static schedule, double-buffered accumulators)
};
struct ksb_gram_plan {
    bool built = false;
    std::vector<ksb_gram_chunk> chunks;
    ksb::GramWinDev* d_wins = nullptr;   // the three tables are one device allocation (d_tables), filled by ONE copy kernel
    ksb::GramItem* d_items = nullptr;    // from a pinned staging buffer: a cudaMemcpy from the vectors would queue on the
    ksb::GramUnit* d_units = nullptr;    // copy engine BEHIND the blob that is already on its way (ksb_pairwise)
    void* d_tables = nullptr;
    void* h_tables = nullptr;            // pinned, kept until the device handle is released
    uint64_t h_tables_bytes = 0;
    bool wq_static = false;              // the weight byte planes of the (only) chunk were laid out when the plan was built
    cudaEvent_t ev_tables = nullptr;     // the copy kernel has run (recorded on the handle's stream)
    bool tables_ordered = false;         // the stream the step runs on has been made to wait for it
};

struct ksb_device {
    ksb::IndexDesc desc;
    int device = 0, sms = 148;
    int rank = 0, world = 1;
    uint8_t* d_payload = nullptr;
    bool own_payload = false;
    cudaStream_t stream = nullptr;
    cudaStream_t stream2 = nullptr;        // K4 runs here, beside K1/K2 (tensor/ALU-bound next to red-bound)
    cudaEvent_t ev[12] = {};
    bool overlap = true;                   // KSPIDER_OVERLAP=0 serialises K4 after K1/K2
    // pipelined upload (ksb_pairwise): the blob is copied in pieces on copy_stream and the first
    // accumulation waits for exactly the piece each kernel reads
    cudaStream_t copy_stream = nullptr;
    bool copy_pending = false;
    bool defer_copy_sync = false;          // the cluster-only step: do not stop the host inside ksb_accumulate for the end of the upload
    bool copy_harvest = false;             //   (the stream waits for it; its timer is read when the step is collected)
    cudaEvent_t ev_copy[4] = {};           // copy start, head landed, tail landed, copy end
    struct k1_piece { cudaEvent_t ev; unsigned long long task_begin, task_end; };
    std::vector<k1_piece> k1_pieces;       // K1 task ranges in the order their classes are copied
    std::vector<cudaEvent_t> gram_ev;      // per K4 chunk of pass 0: the copy piece holding its members has landed
    std::vector<cudaEvent_t> gram_ev_own;  // the events of the pieces the Gram section is copied in
    std::vector<uint64_t> gram_piece_end;  // payload offset each of those pieces ends at
    unsigned long long* d_tri = nullptr;   // max-pass cells (+8 pad); u32 cells when cell_bits == 32
    uint32_t cell_bits = 64;               // 32 when no cell can exceed 2^32 - 1 (IndexDesc::max_row_sum; KSPIDER_CELL_BITS=64 forces u64)
    uint32_t* d_touched = nullptr;         // one bit per cell, only when zero-weight colours exist
    uint64_t tri_cells_alloc = 0;
    std::vector<ksb_pass> passes;
    uint32_t cur_pass = 0;
    uint64_t row_begin = 0, row_end = 0, owned_cells = 0, owned_updates = 0;
    std::vector<uint64_t> bandstart;       // host copy, [N+1]: first cell of row i in the banded accumulator
    const unsigned long long* d_bandstart = nullptr;  // [N+1]   both inside the device payload (IndexDesc::off_bandstart / off_rowbase)
    const unsigned long long* d_rowbase = nullptr;    // [N]  bandstart[i] - band_lo[i]
    ksb::K1Params k1;
    ksb_timings t;
    uint64_t launches = 0;
    // scratch of the post pass
    unsigned long long* d_block_counts = nullptr;
    void* d_scan_tmp = nullptr;
    size_t scan_tmp_bytes = 0;
    uint64_t post_blocks_alloc = 0;
    int block_rows_pass = -1;  // pass whose K5 block rows sit behind the two scan arrays of d_block_counts (-1: none)
    // K4 (Gram) scratch: bitmap of one chunk of colours + weight byte planes
    uint32_t* d_gbits = nullptr;
    uint8_t* d_gwq = nullptr;
    uint64_t gbits_bytes = 0, gwq_bytes = 0;
    uint32_t g_planes = 0;
    std::vector<ksb::GramWin> gwins;       // host copy of the blob's Gram windows
    std::vector<ksb_gram_plan> gplan;      // per dense pass, built on first use
    // K3 (spill)
    uint64_t spill_mode = 0, budget = 0;
    bool spill = false;
    // cluster-only step (ksb_device_cluster): union-find forest, kept-edge list, counters, pinned staging
    uint32_t* d_parent = nullptr;
    unsigned long long* d_cl_counters = nullptr;  // [0] touched cells (rows), [1] kept edges appended (may exceed edge_cap)
    unsigned long long* d_edges = nullptr;
    uint64_t edge_cap = 0;
    uint8_t* h_stage = nullptr;                    // pinned: labels[N] then 2 x u64 counters
    uint64_t h_stage_bytes = 0;
    // the step as a CUDA graph: captured once per (column, threshold, outputs), replayed by one launch
    struct graph_key { int dist_col; uint32_t cstar_bits; uint32_t* d_labels_out; bool host_labels, overlap; uint64_t edge_cap; void* d_edges; };
    bool use_graph = true;                         // KSPIDER_GRAPH=0 / ksb_device_set_graph(dev, 0): enqueue every step kernel by kernel
    cudaGraphExec_t cl_graph = nullptr;
    graph_key cl_key{};
    uint64_t cl_graph_launches = 0;
    cudaStream_t cl_stream = nullptr;              // stream of the step enqueued last
    bool cl_host_labels = false, cl_was_graph = false;
};

struct ksb_segment {
    ksb_record* d_records = nullptr;
    unsigned long long* d_pairs = nullptr;  // [n] (i << 32 | j) compact indices, followed by float cmax[n] (K6's default column)
    uint64_t n = 0;
};

struct ksb_result {
    int device = 0;
    uint64_t n_samples = 0;
    std::vector<ksb_segment> seg;
    ksb_record* h_records = nullptr;  // all segments concatenated (only when want_records); pinned unless h_pageable
    uint64_t h_capacity = 0;
    bool h_pageable = false;          // host-only results (want_records == 2: the TSV-writing drop-in) of more than 256 MB live in
                                      // plain 2 MB-aligned memory: page-locking 1.6 GB costs ~0.5 s, a staged copy into it ~0.1 s
    bool have_host = false;
    bool host_only = false;           // want_records == 2: the device copies of a pass are released once its rows are on the host
    ksb_result_info info;
};

namespace ksb {

static inline uint64_t rowstart(const ksb_device* d, uint64_t r) { return d->bandstart.empty() ? 0 : d->bandstart[r]; }

static TriGeom geom_of(const ksb_device* d, const ksb_pass& p) {
    TriGeom g;
    g.tri = d->d_tri;
    g.base = p.base;
    g.cells = p.cells;
    g.rowbase = d->d_rowbase;
    g.bandstart = d->d_bandstart;
    g.band_lo = reinterpret_cast<const uint32_t*>(d->d_payload + d->desc.off_band_lo);
    g.band_hi = reinterpret_cast<const uint32_t*>(d->d_payload + d->desc.off_band_hi);
    g.cell_bits = d->cell_bits;
    g.n = (uint32_t)d->desc.n_samples;
    g.rb = (uint32_t)p.rb;
    g.nrows = (uint32_t)(p.re - p.rb);
    return g;
}

// Per-ordinal state that is expensive to (re)create: properties, a service stream for results and
// clustering, two events, and the stream-ordered memory pool configured to keep its memory.
struct DevCtx {
    bool ready = false;
    int major = 0, sms = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
};
static DevCtx g_ctx[64];
static std::mutex g_ctx_mu;

// Streams and events are recycled per device: creating ~3 streams and ~25 events per ksb_pairwise call
// was a measurable part of the end-to-end time of a millisecond-scale run.
static std::vector<cudaStream_t> g_stream_pool[64];
static std::vector<cudaEvent_t> g_event_pool[64][2];  // [device][timing disabled?]
static cudaError_t pool_stream(int dev, cudaStream_t* out) {
    {
        std::lock_guard<std::mutex> lk(g_ctx_mu);
        auto& v = g_stream_pool[dev];
        if (!v.empty()) { *out = v.back(); v.pop_back(); return cudaSuccess; }
    }
    return cudaStreamCreateWithFlags(out, cudaStreamNonBlocking);
}
static void unpool_stream(int dev, cudaStream_t s) {  // the stream must be idle
    if (!s) return;
    std::lock_guard<std::mutex> lk(g_ctx_mu);
    g_stream_pool[dev].push_back(s);
}
static cudaError_t pool_event(int dev, cudaEvent_t* out, bool timing = true) {
    {
        std::lock_guard<std::mutex> lk(g_ctx_mu);
        auto& v = g_event_pool[dev][timing ? 0 : 1];
        if (!v.empty()) { *out = v.back(); v.pop_back(); return cudaSuccess; }
    }
    return timing ? cudaEventCreate(out) : cudaEventCreateWithFlags(out, cudaEventDisableTiming);
}
static void unpool_event(int dev, cudaEvent_t e, bool timing = true) {
    if (!e) return;
    std::lock_guard<std::mutex> lk(g_ctx_mu);
    g_event_pool[dev][timing ? 0 : 1].push_back(e);
}

int check_device(int device, int* sms) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        set_error("no CUDA device: %s (libkspider_b200 has no CPU fallback)", e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
        (void)cudaGetLastError();
        return KSB_ERR_CUDA;
    }
    if (device < 0) KSB_CUDA(cudaGetDevice(&device));
    if (device >= n || device >= 64) { set_error("device %d out of range (%d devices)", device, n); return KSB_ERR_ARG; }
    KSB_CUDA(cudaSetDevice(device));
    DevCtx& c = g_ctx[device];
    {
        std::lock_guard<std::mutex> lk(g_ctx_mu);
        if (!c.ready) {
            KSB_CUDA(cudaDeviceGetAttribute(&c.major, cudaDevAttrComputeCapabilityMajor, device));
            KSB_CUDA(cudaDeviceGetAttribute(&c.sms, cudaDevAttrMultiProcessorCount, device));
            if (c.major == 10) {
                KSB_CUDA(cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
                KSB_CUDA(cudaEventCreate(&c.e0));
                KSB_CUDA(cudaEventCreate(&c.e1));
                cudaMemPool_t pool;
                KSB_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
                uint64_t keep = UINT64_MAX;  // freed blocks stay in the pool: a step re-uses them
                KSB_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
            }
            c.ready = true;
        }
    }
    if (c.major != 10) {
        set_error("device %d is sm_%dx; this library is built for sm_100a only", device, c.major);
        return KSB_ERR_CUDA;
    }
    if (sms) *sms = c.sms;
    return device;
}

cudaStream_t service_stream(int dev) { return g_ctx[dev].stream; }

int warm_device(int device) {
    const int dev = check_device(device, nullptr);
    if (dev < 0) return dev;
    return cudaFree(nullptr) == cudaSuccess ? KSB_OK : KSB_ERR_CUDA;
}

static void build_k1(ksb_device* d) {
    K1Params& p = d->k1;
    std::memset(&p, 0, sizeof p);
    unsigned long long task = 0;
    uint32_t n = 0;
    // largest colours first: the long tasks start early, the tail is made of short ones
    for (uint32_t s = kSmallMax; s >= 2; s--) {
        const SmallClass& sc = d->desc.cls[s];
        if (sc.n == 0) continue;
        K1Class& c = p.cls[n++];
        c.task_begin = task;
        c.n = sc.n;
        c.member_off = sc.member_off;
        c.weight_off = sc.weight_off;
        c.s = s;
        c.k = s <= 32 ? 32 / s : 1;
        task += (sc.n + c.k - 1) / c.k;
    }
    p.n_cls = n;
    p.n_tasks = task;
    p.weight_bits = (uint32_t)d->desc.weight_bits;
}

// Row block of `rank` (equal weighted pair-updates per rank), then dense passes under the budget.
static int plan_rows(ksb_device* d, const uint64_t* row_updates, uint64_t budget_bytes) {
    const uint64_t N = d->desc.n_samples;
    const uint64_t U = d->desc.n_updates;
    const int W = std::max(1, d->world);
    if (d->desc.shard_world) {  // a row-block shard owns the rows it was cut for, whatever opts say
        d->row_begin = d->desc.shard_rb;
        d->row_end = d->desc.shard_re;
    } else if (W == 1) {
        d->row_begin = 0;
        d->row_end = N;
    } else {
        std::vector<uint64_t> cut(W + 1, N);
        row_cuts(row_updates, N, U, W, cut.data());
        d->row_begin = cut[d->rank];
        d->row_end = cut[d->rank + 1];
        if (d->rank == W - 1) d->row_end = N;
    }
    d->owned_cells = rowstart(d, d->row_end) - rowstart(d, d->row_begin);
    if (d->desc.shard_world || W == 1) {
        d->owned_updates = U;  // a shard's descriptor counts the updates of its own rows
    } else {
        d->owned_updates = 0;
        for (uint64_t i = d->row_begin; i < d->row_end; i++) d->owned_updates += row_updates[i];
    }

    // bytes per cell: 8 (u64) + 1/8 (touched bit) when needed
    const bool need_touch = d->desc.n_zero_colors != 0;
    uint64_t max_cells = budget_bytes / 8;
    if (need_touch) max_cells = (uint64_t)((double)budget_bytes / 8.125);
    if (max_cells < N) max_cells = N;  // at least one row
    d->passes.clear();
    uint64_t rb = d->row_begin;
    if (rb < d->row_end && d->owned_cells <= max_cells) {  // the whole block in one pass: no walk over the rows
        d->passes.push_back(ksb_pass{rb, d->row_end, rowstart(d, rb), d->owned_cells, d->owned_updates});
        rb = d->row_end;
    }
    while (rb < d->row_end) {
        uint64_t re = rb, cells = 0, upd = 0;
        while (re < d->row_end) {
            const uint64_t rc = d->bandstart[re + 1] - d->bandstart[re];
            if (re > rb && cells + rc > max_cells) break;
            cells += rc;
            upd += row_updates[re];
            re++;
        }
        d->passes.push_back(ksb_pass{rb, re, rowstart(d, rb), cells, upd});
        rb = re;
    }
    if (d->passes.empty()) d->passes.push_back(ksb_pass{d->row_begin, d->row_begin, rowstart(d, d->row_begin), 0, 0});
    uint64_t mx = 0;
    for (const ksb_pass& p : d->passes) mx = std::max(mx, p.cells);
    d->tri_cells_alloc = mx;
    return KSB_OK;
}

// KSPIDER_SETUP_TIMERS=1: host wall clock of the set-up phases of a device handle on stderr (where the end-to-end call
// spends its time before the first kernel)
struct SetupTrace {
    bool on;
    std::chrono::steady_clock::time_point tp;
    SetupTrace() : tp(std::chrono::steady_clock::now()) {
        static const bool env = [] { const char* v = std::getenv("KSPIDER_SETUP_TIMERS"); return v && v[0] == '1'; }();
        on = env;
    }
    void lap(const char* what) {
        if (!on) return;
        const auto now = std::chrono::steady_clock::now();
        std::fprintf(stderr, "[ksb setup] %-28s %8.1f us\n", what, std::chrono::duration<double, std::micro>(now - tp).count());
        tp = std::chrono::steady_clock::now();
    }
};

static int finish_device(ksb_device* d, const uint64_t* row_updates, const uint64_t* bandstart, const ksb_opts* opts) {
    SetupTrace tr;
    if (d->desc.magic != kIndexMagic) { set_error("descriptor has a wrong magic (ABI mismatch?)"); return KSB_ERR_ARG; }
    if (d->desc.member_bits != 16 && d->desc.member_bits != 32) { set_error("descriptor: member_bits must be 16 or 32"); return KSB_ERR_ARG; }
    if (d->desc.packed_bits != 10 && d->desc.packed_bits != 12 && d->desc.packed_bits != 16 && d->desc.packed_bits != 32) { set_error("descriptor: packed_bits must be 10, 12, 16 or 32"); return KSB_ERR_ARG; }
    if (d->desc.weight_bits != 8 && d->desc.weight_bits != 16 && d->desc.weight_bits != 32) { set_error("descriptor: weight_bits must be 8, 16 or 32"); return KSB_ERR_ARG; }
    if (d->desc.n_samples > 0x7FFFFFFFull) { set_error("more than 2^31 samples are not supported"); return KSB_ERR_ARG; }
    if (!d->stream) KSB_CUDA(pool_stream(d->device, &d->stream));
    if (!d->stream2) KSB_CUDA(pool_stream(d->device, &d->stream2));
    for (auto& e : d->ev) KSB_CUDA(pool_event(d->device, &e));
    { const char* ov = std::getenv("KSPIDER_OVERLAP"); d->overlap = !(ov && ov[0] == '0'); }
    { const char* gv = std::getenv("KSPIDER_GRAPH"); d->use_graph = !(gv && gv[0] == '0'); }
    uint64_t budget = opts ? opts->dense_budget_bytes : 0;
    if (budget == 0) {
        // default: 80 % of the free memory for a pass -- its cells (8 bytes each) AND, when records are wanted, the rows it
        // can emit (32-byte record + 12 bytes of pair / column per touched cell, every cell in the worst case)
        size_t fr = 0, tot = 0;
        KSB_CUDA(cudaMemGetInfo(&fr, &tot));
        const bool rows = opts && opts->want_records != 0;
        budget = (uint64_t)((double)fr * 0.8 * (rows ? 8.0 / 52.0 : 1.0));
    }
    tr.lap("  streams, events, mem info");
    {   // banded accumulator: the prefix of the band lengths and the per-row base the kernels add columns to are sections of the blob
        const uint64_t N = d->desc.n_samples;
        d->bandstart.assign(bandstart, bandstart + N + 1);
        if (N && (d->bandstart[0] != 0 || d->bandstart[N] != d->desc.band_cells)) { set_error("corrupt row band tables"); return KSB_ERR_ARG; }
        d->d_bandstart = reinterpret_cast<const unsigned long long*>(d->d_payload + d->desc.off_bandstart);
        d->d_rowbase = reinterpret_cast<const unsigned long long*>(d->d_payload + d->desc.off_rowbase);
    }
    tr.lap("  band tables");
    int rc = plan_rows(d, row_updates, budget);
    if (rc) return rc;
    tr.lap("  plan_rows");
    d->budget = budget;
    d->spill = d->spill_mode == 1 || (d->spill_mode == 0 && d->passes.size() > 16);
    if (d->spill) {  // K3: one pseudo-pass over all owned rows, no dense block
        const uint64_t N = d->desc.n_samples;
        d->passes.assign(1, ksb_pass{d->row_begin, d->row_end, rowstart(d, d->row_begin), d->owned_cells, d->owned_updates});
        build_k1(d);
        return KSB_OK;
    }
    const uint64_t cells = d->tri_cells_alloc;
    {   // a cell is a sum of weights of colours holding its row: bounded by the row's own sum of weights (the loader's max_row_sum)
        // u32 cells pay where the block streams through HBM; on a small L2-resident block (cfg2: 4 MB) they change nothing
        // for K1 and cost K4 beside it (measured: cfg2 step 1.86 -> 2.31 ms), so those keep u64.  KSPIDER_CELL_BITS=32|64 forces.
        const char* cb = std::getenv("KSPIDER_CELL_BITS");
        const int cbv = cb ? std::atoi(cb) : 0;
        const bool fits = d->desc.max_row_sum < (1ull << 32);
        d->cell_bits = (fits && cbv != 64 && (cbv == 32 || 8 * cells > (16ull << 20))) ? 32u : 64u;
    }
    KSB_CUDA(cudaMallocAsync(&d->d_tri, (d->cell_bits / 8) * (cells + 16), d->stream));
    if (d->desc.n_zero_colors) KSB_CUDA(cudaMallocAsync(&d->d_touched, 4 * (cells / 32 + 2), d->stream));
    d->post_blocks_alloc = post_blocks(cells) + 1;
    KSB_CUDA(cudaMallocAsync(&d->d_block_counts, 8 * d->post_blocks_alloc * 3, d->stream));  // counts, offsets, u32 block rows
    d->block_rows_pass = -1;
    d->scan_tmp_bytes = 0;
    KSB_CUDA(scan_block_counts(nullptr, &d->scan_tmp_bytes, d->d_block_counts, d->d_block_counts + d->post_blocks_alloc,
                               d->post_blocks_alloc, d->stream));
    KSB_CUDA(cudaMallocAsync(&d->d_scan_tmp, std::max<size_t>(d->scan_tmp_bytes, 256), d->stream));
    tr.lap("  accumulator allocs");
    build_k1(d);
    tr.lap("  build_k1");
    d->g_planes = d->desc.gram_planes <= 1 ? 1 : (d->desc.gram_planes == 2 ? 2 : 4);
    d->gplan.assign(d->passes.size(), ksb_gram_plan());
    return KSB_OK;
}

// One-slot cache of pinned host memory: cudaMallocHost costs ~0.3 ms per MB, a result's record
// buffer is re-used by the next run of the same shape.
static std::mutex g_pin_mu;
static void* g_pin_ptr = nullptr;
static uint64_t g_pin_bytes = 0;

static cudaError_t pinned_acquire(void** out, uint64_t* bytes) {
    {
        std::lock_guard<std::mutex> lk(g_pin_mu);
        if (g_pin_ptr && g_pin_bytes >= *bytes) {
            *out = g_pin_ptr;
            *bytes = g_pin_bytes;
            g_pin_ptr = nullptr;
            g_pin_bytes = 0;
            return cudaSuccess;
        }
    }
    return cudaMallocHost(out, *bytes);
}

static void pinned_release(void* p, uint64_t bytes) {
    void* drop = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_pin_mu);
        if (bytes > g_pin_bytes) { drop = g_pin_ptr; g_pin_ptr = p; g_pin_bytes = bytes; }
        else drop = p;
    }
    if (drop) cudaFreeHost(drop);
}

// the same for the K4 plan tables of a device handle (a few hundred kB; its own slot so that the records' buffer and
// the tables do not evict each other)
static void* g_plan_pin_ptr = nullptr;
static uint64_t g_plan_pin_bytes = 0;
static cudaError_t plan_pinned_acquire(void** out, uint64_t* bytes) {
    {
        std::lock_guard<std::mutex> lk(g_pin_mu);
        if (g_plan_pin_ptr && g_plan_pin_bytes >= *bytes) {
            *out = g_plan_pin_ptr;
            *bytes = g_plan_pin_bytes;
            g_plan_pin_ptr = nullptr;
            g_plan_pin_bytes = 0;
            return cudaSuccess;
        }
    }
    *bytes = std::max<uint64_t>(*bytes + *bytes / 4, 1ull << 20);
    return cudaMallocHost(out, *bytes);
}
static void plan_pinned_release(void* p, uint64_t bytes) {
    void* drop = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_pin_mu);
        if (bytes > g_plan_pin_bytes) { drop = g_plan_pin_ptr; g_plan_pin_ptr = p; g_plan_pin_bytes = bytes; }
        else drop = p;
    }
    if (drop) cudaFreeHost(drop);
}

// pinned host memory -> device memory on the SMs (16 bytes per thread and step): PCIe reads that interleave with a DMA
// copy in flight instead of waiting for it in the copy engine's queue
__global__ void __launch_bounds__(256) k_host_to_device(uint4* __restrict__ dst, const uint4* __restrict__ src, uint32_t n16) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += gridDim.x * blockDim.x) dst[i] = src[i];
}

}  // namespace ksb

using namespace ksb;

static float ev_ms(cudaEvent_t a, cudaEvent_t b) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, a, b) != cudaSuccess) { (void)cudaGetLastError(); return 0.f; }
    return ms;
}

extern "C" {

int ksb_cuda_available(void) {
    int sms = 0;
    const int rc = check_device(-1, &sms);
    return rc < 0 ? rc : KSB_OK;
}

void ksb_device_free(ksb_device* d) {
    if (!d) return;
    cudaSetDevice(d->device);
    if (d->stream) cudaStreamSynchronize(d->stream);
    cudaStream_t fs = d->stream ? d->stream : g_ctx[d->device].stream;
    if (d->own_payload && d->d_payload) cudaFreeAsync(d->d_payload, fs);
    if (d->d_tri) cudaFreeAsync(d->d_tri, fs);
    if (d->d_touched) cudaFreeAsync(d->d_touched, fs);
    if (d->d_block_counts) cudaFreeAsync(d->d_block_counts, fs);
    if (d->d_scan_tmp) cudaFreeAsync(d->d_scan_tmp, fs);
    for (ksb_gram_plan& gp : d->gplan)
        if (gp.d_tables) cudaFreeAsync(gp.d_tables, fs);
    if (d->d_gbits) cudaFreeAsync(d->d_gbits, fs);
    if (d->d_gwq) cudaFreeAsync(d->d_gwq, fs);
    if (d->d_parent) cudaFreeAsync(d->d_parent, fs);
    if (d->d_cl_counters) cudaFreeAsync(d->d_cl_counters, fs);
    if (d->d_edges) cudaFreeAsync(d->d_edges, fs);
    if (fs) cudaStreamSynchronize(fs);
    if (d->h_stage) pinned_release(d->h_stage, d->h_stage_bytes);
    for (ksb_gram_plan& gp : d->gplan) {
        if (gp.h_tables) plan_pinned_release(gp.h_tables, gp.h_tables_bytes);
        if (gp.ev_tables) unpool_event(d->device, gp.ev_tables, false);
    }
    if (d->cl_graph) cudaGraphExecDestroy(d->cl_graph);
    for (auto& e : d->ev) unpool_event(d->device, e);
    if (d->stream) { cudaStreamSynchronize(d->stream); unpool_stream(d->device, d->stream); }
    if (d->stream2) { cudaStreamSynchronize(d->stream2); unpool_stream(d->device, d->stream2); }
    if (d->copy_stream) { cudaStreamSynchronize(d->copy_stream); unpool_stream(d->device, d->copy_stream); }
    for (auto& e : d->ev_copy) unpool_event(d->device, e);
    for (auto& pc : d->k1_pieces) unpool_event(d->device, pc.ev, false);
    for (auto& e : d->gram_ev_own) unpool_event(d->device, e, false);
    (void)cudaGetLastError();
    delete d;
}

static int plan_gram(ksb_device* d, uint32_t pass);

static int opts_common(ksb_device* d, const ksb_opts* opts) {
    d->rank = opts ? opts->rank : 0;
    d->world = (opts && opts->world > 1) ? opts->world : 1;
    if (d->rank < 0 || d->rank >= d->world) { set_error("rank %d outside world %d", d->rank, d->world); return KSB_ERR_ARG; }
    std::memset(&d->t, 0, sizeof d->t);
    d->spill_mode = opts ? opts->spill_mode : 0;
    if (d->spill_mode > 2) { set_error("spill_mode must be 0 (auto), 1 (always) or 2 (never)"); return KSB_ERR_ARG; }
    return KSB_OK;
}

// Enqueues the copy of the WHOLE blob on the copy stream, in the order the kernels consume it, before the device is
// planned -- planning (row passes, K4 chunks and items) then runs under the copy instead of in front of it.
// Head (totals, ids, bands, row tables), the mid / zero sections and the Gram offsets, weights and windows; then the
// Gram section (members or bitmaps: what K4 waits for) in pieces of <= 4 MB, each followed by an event -- once the K4
// chunks of the first pass are known (map_gram_pieces) a chunk waits for the piece its last byte lies in --; then the
// small-colour classes from the largest size down in a few chunks, the order K1 consumes them, an event behind each.
static int enqueue_copy_all(ksb_device* d, const ksb_index* ix) {
    const IndexDesc& D = d->desc;
    // a handle that ksb_pairwise_cluster kept from its last call uploads again with the stream and events it has
    const bool again = d->copy_stream != nullptr;
    if (!again) {
        KSB_CUDA(pool_stream(d->device, &d->copy_stream));
        for (auto& e : d->ev_copy) KSB_CUDA(pool_event(d->device, &e));
    }
    size_t n_gram_ev = 0, n_k1_ev = 0;
    cudaStream_t cs = d->copy_stream;
    cudaEvent_t alloc = nullptr;  // d_payload was allocated in stream order on d->stream
    KSB_CUDA(pool_event(d->device, &alloc, false));
    KSB_CUDA(cudaEventRecord(alloc, d->stream));
    KSB_CUDA(cudaStreamWaitEvent(cs, alloc, 0));
    unpool_event(d->device, alloc, false);
    auto copy = [&](uint64_t b, uint64_t e) -> cudaError_t {
        return e > b ? cudaMemcpyAsync(d->d_payload + b, ix->payload + b, e - b, cudaMemcpyHostToDevice, cs) : cudaSuccess;
    };
    KSB_CUDA(cudaEventRecord(d->ev_copy[0], cs));
    KSB_CUDA(copy(0, D.off_small_members));
    KSB_CUDA(cudaEventRecord(d->ev_copy[1], cs));
    KSB_CUDA(copy(D.off_mid_offsets, D.off_gram_members));
    KSB_CUDA(copy(D.off_gram_weights, D.payload_bytes));
    KSB_CUDA(cudaEventRecord(d->ev_copy[2], cs));
    const uint64_t piece = 4ull << 20;
    for (uint64_t b = D.off_gram_members; b < D.off_gram_weights; b += piece) {
        const uint64_t e = std::min<uint64_t>(b + piece, D.off_gram_weights);
        KSB_CUDA(copy(b, e));
        if (n_gram_ev == d->gram_ev_own.size()) {
            cudaEvent_t ev = nullptr;
            KSB_CUDA(pool_event(d->device, &ev, false));
            d->gram_ev_own.push_back(ev);
            d->gram_piece_end.push_back(e);
        }
        KSB_CUDA(cudaEventRecord(d->gram_ev_own[n_gram_ev++], cs));
    }
    build_k1(d);  // depends on the descriptor only
    const uint64_t mbits = D.packed_bits, wbytes = D.weight_bits / 8;
    const uint64_t small_bytes = mbits * D.n_small_members / 8 + wbytes * D.n_small_colors;
    const uint64_t chunk = std::max<uint64_t>(8ull << 20, small_bytes / 6);
    const K1Params& k1 = d->k1;  // classes in descending size = task order
    for (uint32_t c = 0; c < k1.n_cls;) {
        uint32_t e = c;
        uint64_t bytes = 0;
        while (e < k1.n_cls && (e == c || bytes < chunk)) { bytes += k1.cls[e].n * (mbits * k1.cls[e].s + 8 * wbytes) / 8; e++; }
        // classes c..e-1 have descending s, i.e. DEscending offsets: their members / weights are one run each
        const K1Class &hi = k1.cls[c], &lo = k1.cls[e - 1];
        KSB_CUDA(copy(D.off_small_members + mbits * lo.member_off / 8, D.off_small_members + (mbits * (hi.member_off + hi.n * hi.s) + 7) / 8 + 1));
        KSB_CUDA(copy(D.off_small_weights + wbytes * lo.weight_off, D.off_small_weights + wbytes * (hi.weight_off + hi.n)));
        if (n_k1_ev == d->k1_pieces.size()) {
            ksb_device::k1_piece pc;
            pc.ev = nullptr;
            KSB_CUDA(pool_event(d->device, &pc.ev, false));
            pc.task_begin = hi.task_begin;
            pc.task_end = e < k1.n_cls ? k1.cls[e].task_begin : k1.n_tasks;
            d->k1_pieces.push_back(pc);
        }
        KSB_CUDA(cudaEventRecord(d->k1_pieces[n_k1_ev++].ev, cs));
        c = e;
    }
    KSB_CUDA(cudaEventRecord(d->ev_copy[3], cs));
    d->copy_pending = true;
    return KSB_OK;
}

// the K4 chunks of pass 0 -> the copy piece whose arrival completes their part of the Gram section
static void map_gram_pieces(ksb_device* d, const ksb_index* ix) {
    const IndexDesc& D = d->desc;
    const ksb_gram_plan* plan = (D.n_gram_colors && !d->gplan.empty() && d->gplan[0].built && !d->gplan[0].chunks.empty()) ? &d->gplan[0] : nullptr;
    if (!plan || d->gram_piece_end.empty()) return;  // no K4 launch in pass 0: nobody waits for a piece
    const uint64_t* goff = reinterpret_cast<const uint64_t*>(ix->payload + D.off_gram_offsets);
    const size_t nck = plan->chunks.size();
    d->gram_ev.assign(nck, nullptr);
    for (size_t k = 0; k < nck; k++) {
        const ksb_gram_chunk& c = plan->chunks[k];
        uint64_t end = D.gram_form ? D.off_gram_members + 16 * (c.blob_row0 + c.bits_rows + kGramTile)
                                   : D.off_gram_members + (gram_member_bits(D) * goff[c.c_begin + c.n_colors] + 7) / 8 + 64;
        end = std::min<uint64_t>(end, D.off_gram_weights);
        const size_t pi = std::lower_bound(d->gram_piece_end.begin(), d->gram_piece_end.end(), end) - d->gram_piece_end.begin();
        d->gram_ev[k] = d->gram_ev_own[std::min(pi, d->gram_ev_own.size() - 1)];
    }
}

static int upload_impl(const ksb_index* ix, const ksb_opts* opts, ksb_device** out, bool pipelined) {
    if (!ix || !out) { set_error("ksb_upload: null argument"); return KSB_ERR_ARG; }
    *out = nullptr;
    int sms = 0;
    const int dev = check_device(opts ? opts->device : -1, &sms);
    if (dev < 0) return dev;
    ksb_device* d = new (std::nothrow) ksb_device();
    if (!d) { set_error("out of memory"); return KSB_ERR_NOMEM; }
    d->device = dev;
    d->sms = sms;
    d->desc = ix->desc;
    int rc = opts_common(d, opts);
    if (rc) { ksb_device_free(d); return rc; }
    auto fail = [&](int code) { ksb_device_free(d); return code; };
    cudaError_t e = pool_stream(d->device, &d->stream);
    if (e == cudaSuccess) e = cudaMallocAsync(&d->d_payload, d->desc.payload_bytes, d->stream);
    if (e != cudaSuccess) { set_error("cudaMalloc(%llu) for the index: %s", (unsigned long long)d->desc.payload_bytes, cudaGetErrorString(e)); (void)cudaGetLastError(); return fail(KSB_ERR_CUDA); }
    d->own_payload = true;
    SetupTrace tr;
    auto lap = [&](const char* what) { tr.lap(what); };
    const bool pipe = pipelined && ix->pinned;
    if (pipe) {
        rc = enqueue_copy_all(d, ix);
        if (rc) return fail(rc);
    }
    lap("copy enqueue");
    const uint64_t* row_updates = reinterpret_cast<const uint64_t*>(ix->payload + d->desc.off_row_updates);
    if (d->desc.n_gram_wins) {
        const GramWin* gw = reinterpret_cast<const GramWin*>(ix->payload + d->desc.off_gram_wins);
        d->gwins.assign(gw, gw + d->desc.n_gram_wins);
    }
    rc = finish_device(d, row_updates, reinterpret_cast<const uint64_t*>(ix->payload + d->desc.off_bandstart), opts);
    if (rc) return fail(rc);
    lap("finish_device");
    if (pipe && !d->spill) {
        if (d->desc.n_gram_colors && !d->passes.empty() && d->passes[0].cells) {
            rc = plan_gram(d, 0);  // the copy is cut along the K4 chunks of the first pass
            if (rc) return fail(rc);
        }
        lap("plan_gram");
        map_gram_pieces(d, ix);
        *out = d;
        return KSB_OK;
    }
    if (pipe) {  // the spill path takes the blob whole: wait for the copy that is already under way
        e = cudaStreamSynchronize(d->copy_stream);
        if (e != cudaSuccess) { set_error("index H2D copy: %s", cudaGetErrorString(e)); (void)cudaGetLastError(); return fail(KSB_ERR_CUDA); }
        d->t.h2d_ms = ev_ms(d->ev_copy[0], d->ev_copy[3]);
        d->copy_pending = false;
        *out = d;
        return KSB_OK;
    }
    cudaEventRecord(d->ev[0], d->stream);
    e = cudaMemcpyAsync(d->d_payload, ix->payload, d->desc.payload_bytes, cudaMemcpyHostToDevice, d->stream);
    cudaEventRecord(d->ev[1], d->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(d->stream);
    if (e != cudaSuccess) { set_error("index H2D copy: %s", cudaGetErrorString(e)); (void)cudaGetLastError(); return fail(KSB_ERR_CUDA); }
    d->t.h2d_ms = ev_ms(d->ev[0], d->ev[1]);
    *out = d;
    return KSB_OK;
}

int ksb_upload(const ksb_index* ix, const ksb_opts* opts, ksb_device** out) { return upload_impl(ix, opts, out, false); }

int ksb_adopt_device_payload(const void* descriptor, uint64_t descriptor_bytes, const void* d_payload,
                             uint64_t payload_bytes, const ksb_opts* opts, ksb_device** out) {
    if (!descriptor || !d_payload || !out) { set_error("ksb_adopt_device_payload: null argument"); return KSB_ERR_ARG; }
    *out = nullptr;
    if (descriptor_bytes != sizeof(IndexDesc)) { set_error("descriptor is %llu bytes, expected %zu", (unsigned long long)descriptor_bytes, sizeof(IndexDesc)); return KSB_ERR_ARG; }
    int sms = 0;
    const int dev = check_device(opts ? opts->device : -1, &sms);
    if (dev < 0) return dev;
    ksb_device* d = new (std::nothrow) ksb_device();
    if (!d) { set_error("out of memory"); return KSB_ERR_NOMEM; }
    d->device = dev;
    d->sms = sms;
    std::memcpy(&d->desc, descriptor, sizeof(IndexDesc));
    if (d->desc.payload_bytes != payload_bytes) { set_error("payload is %llu bytes, descriptor says %llu", (unsigned long long)payload_bytes, (unsigned long long)d->desc.payload_bytes); delete d; return KSB_ERR_ARG; }
    int rc = opts_common(d, opts);
    if (rc) { ksb_device_free(d); return rc; }
    d->d_payload = const_cast<uint8_t*>(static_cast<const uint8_t*>(d_payload));
    d->own_payload = false;
    std::vector<uint64_t> row_updates(d->desc.n_samples);
    if (d->desc.n_samples) {
        cudaError_t e = cudaMemcpy(row_updates.data(), d->d_payload + d->desc.off_row_updates, 8 * d->desc.n_samples, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) { set_error("reading row histogram from the device payload: %s", cudaGetErrorString(e)); (void)cudaGetLastError(); ksb_device_free(d); return KSB_ERR_CUDA; }
    }
    std::vector<uint64_t> bst(d->desc.n_samples + 1, 0);
    if (d->desc.n_samples) {
        cudaError_t e = cudaMemcpy(bst.data(), d->d_payload + d->desc.off_bandstart, 8 * (d->desc.n_samples + 1), cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) { set_error("reading row bands from the device payload: %s", cudaGetErrorString(e)); (void)cudaGetLastError(); ksb_device_free(d); return KSB_ERR_CUDA; }
    }
    if (d->desc.n_gram_wins) {
        d->gwins.resize(d->desc.n_gram_wins);
        cudaError_t e = cudaMemcpy(d->gwins.data(), d->d_payload + d->desc.off_gram_wins, sizeof(GramWin) * d->desc.n_gram_wins, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) { set_error("reading Gram windows from the device payload: %s", cudaGetErrorString(e)); (void)cudaGetLastError(); ksb_device_free(d); return KSB_ERR_CUDA; }

    }
    rc = finish_device(d, row_updates.data(), bst.data(), opts);
    if (rc) { ksb_device_free(d); return rc; }
    *out = d;
    return KSB_OK;
}

int ksb_device_n_passes(const ksb_device* d) { return d ? (int)d->passes.size() : 0; }
int ksb_device_set_pass(ksb_device* d, int pass) {
    if (!d || pass < 0 || (size_t)pass >= d->passes.size()) { set_error("pass out of range"); return KSB_ERR_ARG; }
    d->cur_pass = (uint32_t)pass;
    return KSB_OK;
}

int ksb_device_is_spill(const ksb_device* d) { return d && d->spill ? 1 : 0; }

int ksb_device_set_overlap(ksb_device* d, int on) {
    if (!d) { set_error("ksb_device_set_overlap: null device"); return KSB_ERR_ARG; }
    d->overlap = on != 0;
    return KSB_OK;
}

// Plans K4 for one dense pass (rows [p.rb, p.re)): windows that intersect the rows are cut into
// sub-windows of at most 16384 K-blocks (255 * 128 * 16384 < 2^31: an s32 accumulator cannot
// overflow) whose bitmap fits the chunk budget, sub-windows are packed into chunks (one bitmap
// buffer, one launch each), and every chunk gets its work items, largest first.
static int plan_gram(ksb_device* d, uint32_t pass) {
    using namespace ksb;
    ksb_gram_plan& plan = d->gplan[pass];
    if (plan.built) return KSB_OK;
    const ksb_pass& p = d->passes[pass];
    const uint32_t NB_long = gram_cols_per_item((int)d->g_planes);
    uint32_t NB = NB_long;
    // chunks whose windows are all short in K (a clade / family window: a few K-blocks per tile pair) run the
    // pipelined variant; KSPIDER_GRAM_PIPE=0 keeps everything on k4_gram, =<n> moves the K-block limit
    const char* gpe = std::getenv("KSPIDER_GRAM_PIPE");
    const uint32_t pipe_max_kb = d->g_planes > 2 ? 0u : (gpe ? (uint32_t)std::atoi(gpe) : 64u);
    // tuning knob; the default keeps a chunk's bitmap L2-resident (126 MB of L2): 96 MB for a resident index -- cfg2's 162 MB of
    // bitmaps in two chunks instead of three: fewer build / drain phases, step 1.84 -> 1.73 ms --, 64 MB while the blob is still
    // on its way (the end-to-end call: a chunk's K4 starts when ITS members have landed, smaller chunks start earlier)
    const char* cm = std::getenv("KSPIDER_GRAM_CHUNK_MB");
    const uint64_t chunk_rows = ((cm && std::atoi(cm) > 0) ? (uint64_t)std::atoi(cm) : (d->copy_pending ? 64 : 96)) << 16;  // MB -> 16-byte rows
    const char* wv = std::getenv("KSPIDER_GRAM_WAVES");     // tuning knob
    const uint64_t waves = (wv && std::atoi(wv) > 0) ? (uint64_t)std::atoi(wv) : 4;
    std::vector<GramWinDev> wins;
    std::vector<GramItem> items;
    std::vector<GramUnit> units;
    struct Rows { uint32_t ilo, ihi; };  // owned rows of a window, as offsets inside it
    std::vector<Rows> wrows;
    ksb_gram_chunk cur;
    auto close_chunk = [&]() {
        if (cur.n_wins == 0) return;
        uint32_t max_kb = 0;
        for (uint32_t w = cur.win_begin; w < cur.win_begin + cur.n_wins; w++) max_kb = std::max<uint32_t>(max_kb, (wins[w].n + kGramKBlock - 1) / kGramKBlock);
        cur.pipe = max_kb <= pipe_max_kb;
        NB = cur.pipe ? gram_cols_per_item_pipe((int)d->g_planes) : NB_long;
        // base items, then K splits so that there are ~`waves` items per SM
        uint64_t base = 0;
        for (uint32_t w = cur.win_begin; w < cur.win_begin + cur.n_wins; w++)
            for (uint32_t i0 = wrows[w].ilo; i0 < wrows[w].ihi; i0 += kGramTile)
                base += ((wins[w].width - i0 + kGramTile - 1) / kGramTile + NB - 1) / NB;
        const uint64_t ks = base ? (waves * (uint64_t)d->sms + base - 1) / base : 1;
        cur.item_begin = (uint32_t)items.size();
        for (uint32_t w = cur.win_begin; w < cur.win_begin + cur.n_wins; w++) {
            const uint32_t kblocks = (wins[w].n + kGramKBlock - 1) / kGramKBlock;
            const uint32_t split = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(ks, kblocks));
            for (uint32_t i0 = wrows[w].ilo; i0 < wrows[w].ihi; i0 += kGramTile)
                for (uint32_t j0 = i0; j0 < wins[w].width; j0 += NB * kGramTile)
                    for (uint32_t sp = 0; sp < split; sp++) {
                        GramItem it;
                        it.win = w - cur.win_begin;
                        it.i0 = (uint16_t)i0;
                        it.j0 = (uint16_t)j0;
                        it.kb0 = (uint32_t)((uint64_t)kblocks * sp / split);
                        it.kb1 = (uint32_t)((uint64_t)kblocks * (sp + 1) / split);
                        if (it.kb1 > it.kb0) items.push_back(it);
                    }
        }
        cur.n_items = (uint32_t)items.size() - cur.item_begin;
        // bitmap-build units of the chunk, the heaviest (most members to place) first
        cur.unit_begin = (uint32_t)units.size();
        // with few units per CTA slot (a rank's share of a multi-GPU run) the long units of wide windows are the critical
        // path of the build: they are split; with many units the whole-row stores of unsplit units are cheaper
        const bool split_wide = cur.n_units < 3u * 6u * (uint32_t)d->sms;
        if (d->desc.gram_form) cur.n_units = 0;
        for (uint32_t w = cur.win_begin; !d->desc.gram_form && w < cur.win_begin + cur.n_wins; w++) {
            const uint32_t kblocks = (wins[w].n + kGramKBlock - 1) / kGramKBlock, segs = (wins[w].width + kGramSeg - 1) / kGramSeg;
            for (uint32_t kb = 0; kb < kblocks; kb++)
                for (uint32_t sg = 0; sg < segs; sg++) {
                    GramUnit u;
                    u.c0k = wins[w].c0 + (uint64_t)kb * kGramKBlock;
                    u.kn = (uint16_t)std::min<uint64_t>(kGramKBlock, wins[w].n - (uint64_t)kb * kGramKBlock);
                    u.xbase = (d->desc.gram_rel_bits ? 0u : wins[w].x0) + sg * kGramSeg;  // in the coordinates the members are stored in
                    u.slen = std::min<uint32_t>(kGramSeg, wins[w].width - sg * kGramSeg);
                    u.bits_row = wins[w].bits_off + (uint64_t)kb * wins[w].width + (uint64_t)sg * kGramSeg;
                    u.wq_off = sg == 0 ? (wins[w].k0 + kb) * kGramKBlock : 0xFFFFFFFFu;
                    u.w0 = 0;
                    u.nw = 4;
                    if (u.slen >= 512 && split_wide) {  // a wide window holds long colours: four one-word units instead of one long one
                        for (uint8_t wd = 0; wd < 4; wd++) {
                            GramUnit q = u;
                            q.w0 = wd;
                            q.nw = 1;
                            if (wd) q.wq_off = 0xFFFFFFFFu;
                            if (wd < 3) units.push_back(q); else u = q;
                        }
                    }
                    units.push_back(u);
                }
        }
        // proxy for the members a unit places: colours of the K-block x samples of the segment (wide windows hold long colours)
        std::stable_sort(units.begin() + cur.unit_begin, units.end(),
                         [&](const GramUnit& a, const GramUnit& b) { return (uint64_t)a.kn * a.slen * a.nw > (uint64_t)b.kn * b.slen * b.nw; });
        if ((uint32_t)units.size() - cur.unit_begin != cur.n_units) cur.n_units = (uint32_t)units.size() - cur.unit_begin;
        auto cost = [&](const GramItem& it) {
            return (uint64_t)(it.kb1 - it.kb0) * (1 + std::min<uint32_t>(NB, (wins[cur.win_begin + it.win].width - it.j0 + kGramTile - 1) / kGramTile));
        };
        std::stable_sort(items.begin() + cur.item_begin, items.end(), [&](const GramItem& a, const GramItem& b) { return cost(a) > cost(b); });
        if (cur.n_items) plan.chunks.push_back(cur);
    };
    auto new_chunk = [&](uint64_t c_begin) {
        cur = ksb_gram_chunk();
        cur.c_begin = c_begin;
        cur.win_begin = (uint32_t)wins.size();
    };
    new_chunk(0);
    const bool bitmap_form = d->desc.gram_form != 0;  // the chunk's bitmap is a slice of the blob: nothing to build
    uint64_t blob_row = 0;                            // first bitmap row of the current window in the blob
    for (const GramWin& w : d->gwins) {
        const uint64_t x0 = w.x0, x1 = (uint64_t)w.x0 + w.width;
        const uint64_t win_row0 = blob_row;
        blob_row += gram_win_rows(w);
        if (p.re <= x0 || p.rb >= x1 || w.n == 0) {  // no owned row inside the window
            if (cur.n_wins) { close_chunk(); }
            new_chunk(w.c0 + w.n);
            continue;
        }
        const uint32_t tiles = w.width / kGramTile;
        const Rows rows{(uint32_t)(p.rb > x0 ? p.rb - x0 : 0), (uint32_t)(std::min<uint64_t>(p.re, x1) - x0)};
        const uint64_t kb_total = ((uint64_t)w.n + kGramKBlock - 1) / kGramKBlock;
        const uint64_t kb_max = std::max<uint64_t>(1, std::min<uint64_t>(16384, chunk_rows / w.width));
        for (uint64_t kb = 0; kb < kb_total; kb += kb_max) {
            const uint64_t kbs = std::min(kb_max, kb_total - kb);
            const uint64_t kpitch = bitmap_form ? gram_win_stride(w) : w.width;  // rows per K-block where the bitmap lives
            if (cur.n_wins && cur.bits_rows + kbs * kpitch > chunk_rows) { close_chunk(); new_chunk(w.c0 + kb * kGramKBlock); }
            if (cur.n_wins == 0) { cur.c_begin = w.c0 + kb * kGramKBlock; cur.blob_row0 = win_row0 + kb * kpitch; }
            GramWinDev wd;
            wd.c0 = w.c0 + kb * kGramKBlock;
            wd.n = (uint32_t)std::min<uint64_t>(kbs * kGramKBlock, w.n - kb * kGramKBlock);
            wd.bits_off = cur.bits_rows;
            wd.x0 = w.x0;
            wd.width = w.width;
            wd.tiles = tiles;
            wd.k0 = cur.kblocks;
            wd.extent = w.extent ? (uint32_t)std::min<uint64_t>(w.extent, w.width) : 0u;
            wd.kstride = bitmap_form ? gram_win_stride(w) : w.width;
            wd.pad = 0;
            cur.n_units += (uint32_t)kbs * ((w.width + kGramSeg - 1) / kGramSeg);
            wins.push_back(wd);
            wrows.push_back(rows);
            cur.n_wins++;
            cur.n_colors = wd.c0 + wd.n - cur.c_begin;
            cur.bits_rows += kbs * kpitch;
            cur.kblocks += (uint32_t)kbs;
        }
    }
    close_chunk();
    uint64_t need_bits = 0, need_wq = 0;
    for (const ksb_gram_chunk& c : plan.chunks) {
        if (!bitmap_form) need_bits = std::max(need_bits, (c.bits_rows + kGramTile) * 16);  // tiles start at any row: the last may hang over
        need_wq = std::max<uint64_t>(need_wq, (uint64_t)d->g_planes * c.kblocks * kGramKBlock + 256);
    }
    if (!plan.chunks.empty()) {
        KSB_CUDA(cudaStreamSynchronize(d->stream2));  // a grown buffer must not be in use
        if (need_bits > d->gbits_bytes) {
            if (d->d_gbits) KSB_CUDA(cudaFreeAsync(d->d_gbits, d->stream));
            KSB_CUDA(cudaMallocAsync(&d->d_gbits, need_bits, d->stream));
            d->gbits_bytes = need_bits;
        }
        if (need_wq > d->gwq_bytes) {
            if (d->d_gwq) KSB_CUDA(cudaFreeAsync(d->d_gwq, d->stream));
            KSB_CUDA(cudaMallocAsync(&d->d_gwq, need_wq, d->stream));
            d->gwq_bytes = need_wq;
        }
        auto up256 = [](uint64_t b) { return (b + 255) / 256 * 256; };
        const uint64_t o_items = up256(sizeof(GramWinDev) * wins.size());
        const uint64_t o_units = o_items + up256(sizeof(GramItem) * items.size());
        const uint64_t total = o_units + up256(sizeof(GramUnit) * std::max<size_t>(units.size(), 1));
        plan.h_tables_bytes = total;
        KSB_CUDA(plan_pinned_acquire(&plan.h_tables, &plan.h_tables_bytes));
        uint8_t* H = static_cast<uint8_t*>(plan.h_tables);
        std::memcpy(H, wins.data(), sizeof(GramWinDev) * wins.size());
        std::memcpy(H + o_items, items.data(), sizeof(GramItem) * items.size());
        if (!units.empty()) std::memcpy(H + o_units, units.data(), sizeof(GramUnit) * units.size());
        KSB_CUDA(cudaMallocAsync(&plan.d_tables, total, d->stream));
        uint8_t* T = static_cast<uint8_t*>(plan.d_tables);
        plan.d_wins = reinterpret_cast<GramWinDev*>(T);
        plan.d_items = reinterpret_cast<GramItem*>(T + o_items);
        plan.d_units = reinterpret_cast<GramUnit*>(T + o_units);
        const uint32_t n16 = (uint32_t)(total / 16);
        k_host_to_device<<<std::min<uint32_t>((n16 + 255) / 256, 4u * (uint32_t)d->sms), 256, 0, d->stream>>>(
            reinterpret_cast<uint4*>(T), reinterpret_cast<const uint4*>(H), n16);
        KSB_CUDA(cudaGetLastError());
        d->launches++;
        if (bitmap_form && plan.chunks.size() == 1 && d->passes.size() == 1) {
            // One pass, one chunk, bitmaps in the blob: the chunk's weight byte planes are as static as the window and item
            // tables, so they are laid out here, once, instead of in front of K4 in every step (cfg4: 22 us per step).
            const ksb_gram_chunk& c = plan.chunks[0];
            if (d->copy_pending) KSB_CUDA(cudaStreamWaitEvent(d->stream, d->ev_copy[2], 0));  // the weights are on their way
            KSB_CUDA(launch_kb_weights(plan.d_wins + c.win_begin, c.n_wins, c.kblocks, d->d_payload + d->desc.off_gram_weights,
                                       (uint32_t)d->desc.weight_bits, c.kblocks * kGramKBlock, d->g_planes, d->d_gwq, LaunchCfg{d->sms, d->stream}));
            plan.wq_static = true;
            d->launches++;
        }
        KSB_CUDA(pool_event(d->device, &plan.ev_tables, false));
        KSB_CUDA(cudaEventRecord(plan.ev_tables, d->stream));
    }
    plan.built = true;
    return KSB_OK;
}

// K4 over the Gram-routed colours of the current pass, chunk by chunk, on `st`.  The bitmap build of
// the first chunk runs on `st_build` (the main stream): K1 is launched behind it there, so the Gram
// kernel (one big CTA per SM) and K1 become runnable together and the Gram CTAs, submitted first, get
// their slots before K1's CTAs fill the rest of every SM.
static int run_gram(ksb_device* d, const ksb::TriGeom& g, cudaStream_t st_build, cudaStream_t st, cudaEvent_t ev_built) {
    using namespace ksb;
    const int prc = plan_gram(d, d->cur_pass);
    if (prc) return prc;
    ksb_gram_plan& plan = d->gplan[d->cur_pass];
    if (!plan.tables_ordered && plan.ev_tables) {
        // first use of the plan: its tables were written by a kernel on the handle's own stream.  Never reached inside a
        // graph capture: cluster_prepare plans and orders every pass before a capture starts.
        KSB_CUDA(cudaStreamWaitEvent(st_build, plan.ev_tables, 0));
        if (st != st_build) KSB_CUDA(cudaStreamWaitEvent(st, plan.ev_tables, 0));
        plan.tables_ordered = true;
    }
    const IndexDesc& D = d->desc;
    const uint8_t* P = d->d_payload;
    const unsigned long long* goff = reinterpret_cast<const unsigned long long*>(P + D.off_gram_offsets);
    const void* gmem = P + D.off_gram_members;
    const void* gw = P + D.off_gram_weights;
    bool first = true;
    for (const ksb_gram_chunk& c : plan.chunks) {
        GramParams gp;
        gp.bits = D.gram_form ? reinterpret_cast<const uint32_t*>(P + D.off_gram_members) + 4 * c.blob_row0 : d->d_gbits;
        gp.wq = d->d_gwq;
        gp.wins = plan.d_wins + c.win_begin;
        gp.items = plan.d_items + c.item_begin;
        gp.n_wins = c.n_wins;
        gp.n_items = c.n_items;
        gp.kpad = c.kblocks * kGramKBlock;
        gp.planes = d->g_planes;
        gp.counter = reinterpret_cast<uint32_t*>(d->d_gwq + (uint64_t)d->g_planes * gp.kpad);
        cudaStream_t sb = first ? st_build : st;
        if (d->copy_pending) {  // first accumulation after a pipelined upload: wait for this chunk's colours
            const size_t k = &c - plan.chunks.data();
            cudaEvent_t ce = (d->cur_pass == 0 && k < d->gram_ev.size() && d->gram_ev[k]) ? d->gram_ev[k] : d->ev_copy[3];
            KSB_CUDA(cudaStreamWaitEvent(sb, ce, 0));
        }
        KSB_CUDA(cudaMemsetAsync(gp.counter, 0, 256, sb));  // bitmap and weight planes are written in full by kb_build
        if (D.gram_form && plan.wq_static)
            d->launches--;  // nothing to lay out per step (the count below assumes two launches per chunk)
        else if (D.gram_form)
            KSB_CUDA(launch_kb_weights(gp.wins, gp.n_wins, c.kblocks, gw, (uint32_t)D.weight_bits, gp.kpad, gp.planes, d->d_gwq, LaunchCfg{d->sms, sb}));
        else
            KSB_CUDA(launch_kb_build(goff, gmem, (uint32_t)gram_member_bits(D), gw, (uint32_t)D.weight_bits, plan.d_units + c.unit_begin, c.n_units, gp.kpad, gp.planes, d->d_gbits, d->d_gwq,
                                     LaunchCfg{d->sms, sb}));
        if (sb != st) {
            KSB_CUDA(cudaEventRecord(ev_built, sb));
            KSB_CUDA(cudaStreamWaitEvent(st, ev_built, 0));
        }
        KSB_CUDA(c.pipe ? launch_k4_gram_pipe(gp, g, LaunchCfg{d->sms, st}) : launch_k4_gram(gp, g, LaunchCfg{d->sms, st}));
        d->launches += 2;
        first = false;
    }
    return KSB_OK;
}

// phase timers of the accumulation whose events were recorded last; the stream must have been synchronised
static void harvest_accumulate_timers(ksb_device* d) {
    d->t.zero_ms += ev_ms(d->ev[0], d->ev[1]);
    d->t.small_ms += ev_ms(d->ev[1], d->ev[2]);
    d->t.mid_ms += ev_ms(d->ev[2], d->ev[3]);
    d->t.zero_touch_ms += ev_ms(d->ev[3], d->ev[4]);
    d->t.gram_ms += ev_ms(d->ev[9], d->ev[10]);  // overlaps small_ms/mid_ms unless KSPIDER_OVERLAP=0
    d->t.accumulate_ms += ev_ms(d->ev[0], d->ev[8]);
}

int ksb_accumulate(ksb_device* d, void* stream_v, int sync) {
    if (!d) { set_error("ksb_accumulate: null device"); return KSB_ERR_ARG; }
    if (d->spill) { set_error("ksb_accumulate: this device runs the spill path (ksb_device_run_spill)"); return KSB_ERR_ARG; }
    KSB_CUDA(cudaSetDevice(d->device));
    cudaStream_t st = stream_v ? static_cast<cudaStream_t>(stream_v) : d->stream;
    const ksb_pass& p = d->passes[d->cur_pass];
    const TriGeom g = geom_of(d, p);
    const LaunchCfg lc{d->sms, st};
    const IndexDesc& D = d->desc;
    const uint8_t* P = d->d_payload;
    KSB_CUDA(cudaEventRecord(d->ev[0], st));
    KSB_CUDA(cudaMemsetAsync(d->d_tri, 0, (d->cell_bits / 8) * (p.cells + 8), st));
    if (d->d_touched) KSB_CUDA(cudaMemsetAsync(d->d_touched, 0, 4 * (p.cells / 32 + 2), st));
    KSB_CUDA(cudaEventRecord(d->ev[1], st));
    if (d->copy_pending) {  // first accumulation after a pipelined upload
        KSB_CUDA(cudaStreamWaitEvent(st, d->ev_copy[1], 0));
        KSB_CUDA(cudaStreamWaitEvent(st, d->ev_copy[2], 0));
    }
    // K4 goes first, on its own stream: one big CTA per SM (tensor pipe + integer expansion) leaves
    // thread slots for the red-bound K1/K2 CTAs launched right after it on the main stream.
    const bool has_gram = p.cells && D.n_gram_colors;
    if (has_gram && d->overlap) {
        KSB_CUDA(cudaEventRecord(d->ev[9], st));
        const int grc = run_gram(d, g, st, d->stream2, d->ev[11]);
        if (grc) return grc;
        KSB_CUDA(cudaEventRecord(d->ev[10], d->stream2));
    }
    // L2 blocking of the atomic kernels: when the pass's block is far larger than L2 every red costs an
    // HBM sector read + write-back (27 G reds/s measured on 400 MB against 190 G/s on an L2-resident
    // block).  K1/K2 are then launched once per sub-block of rows sized to stay in L2: each launch
    // streams the whole colour list again (cheap: rows outside the sub-block are skipped by warp-uniform
    // tests) but its reds hit L2.  Used when the updates per streamed member make that a win.
    std::vector<TriGeom> subs(1, g);
    {
        const char* env = std::getenv("KSPIDER_L2_BLOCK_MB");  // 0 = off; set = forced (no gain/cost rule)
        const uint64_t blk_mb = env ? (uint64_t)std::atoi(env) : 48;
        const uint64_t atom_upd = D.n_updates - D.n_gram_updates, atom_mem = D.n_members - D.n_gram_members - D.n_zero_members;
        const uint64_t bytes = (d->cell_bits / 8) * p.cells;
        if (blk_mb && bytes > 2 * (blk_mb << 20) && atom_mem) {
            uint64_t n_sub = std::min<uint64_t>(64, (bytes + (blk_mb << 20) - 1) / (blk_mb << 20));
            // gain: the pass's atomic updates at the L2 rate instead of the HBM rate (1/27G - 1/190G s each);
            // cost: every extra launch streams the atomic colours again (~2.8 ps per member) + ~8 us of launches
            const double upd = (double)atom_upd * ((double)p.updates / (double)std::max<uint64_t>(1, D.n_updates));
            const double gain = upd * 32e-12;
            while (!env && n_sub > 1 && gain < 2.0 * (double)(n_sub - 1) * (8e-6 + (double)atom_mem * 2.8e-12)) n_sub /= 2;
            if (n_sub > 1) {
                subs.clear();
                const uint64_t per = (p.cells + n_sub - 1) / n_sub;
                uint64_t rb = p.rb;
                while (rb < p.re) {
                    uint64_t re = rb + 1;
                    while (re < p.re && d->bandstart[re + 1] - d->bandstart[rb] <= per) re++;
                    TriGeom gs = g;
                    gs.rb = (uint32_t)rb;
                    gs.nrows = (uint32_t)(re - rb);
                    subs.push_back(gs);
                    rb = re;
                }
            }
        }
    }
    if (p.cells) {
        LaunchCfg lc1 = lc;
        if (has_gram && d->overlap) {
            // K4 (P = 1: 48 registers x 576 threads = 27.6 K registers, capped by __maxnreg__) leaves room for K1
            // CTAs (32 registers x 256 threads) on every SM: K1 gets 3 per SM (measured on cfg2: 3 -> 2.16 ms/step,
            // 4 -> 2.23, 8 = K4 waits for K1 to drain -> 2.58) and the max-shared carve-out (LaunchCfg::beside_gram).
            // Only when K4 is long enough to be worth slowing K1 for: estimated from the measured rates
            // (5.3e14 u8 MACs/s; 1.9e11 reds/s on an L2-resident block, 2.7e10 otherwise).
            const char* kc = std::getenv("KSPIDER_K1_CTAS");
            const double t4 = (double)D.n_gram_macs / 5.3e14;
            const double t1 = (double)(D.n_updates - D.n_gram_updates) / ((d->cell_bits / 8) * p.cells <= (96ull << 20) ? 1.9e11 : 2.7e10);
            const int fit = t4 >= 0.5 * t1 ? (d->g_planes <= 1 ? 3 : 2) : 8;
            lc1.k1_ctas = (kc && std::atoi(kc) > 0 && std::atoi(kc) <= 8) ? std::atoi(kc) : fit;
            const char* cv = std::getenv("KSPIDER_CARVEOUT");
            lc1.beside_gram = lc1.k1_ctas < 8 && !(cv && std::atoi(cv) == 0);
        }
        const void* sm_m = P + D.off_small_members;
        const uint32_t mb = (uint32_t)D.packed_bits;
        const void* sm_w = P + D.off_small_weights;
        for (const TriGeom& gs : subs) {
            if (d->copy_pending) {
                for (const ksb_device::k1_piece& pc : d->k1_pieces) {
                    KSB_CUDA(cudaStreamWaitEvent(st, pc.ev, 0));
                    KSB_CUDA(launch_k1_small(d->k1, sm_m, mb, sm_w, gs, lc1, pc.task_begin, pc.task_end));
                    d->launches++;
                }
            } else {
                KSB_CUDA(launch_k1_small(d->k1, sm_m, mb, sm_w, gs, lc1, 0ull, d->k1.n_tasks));
                if (d->k1.n_tasks) d->launches++;
            }
        }
    }
    KSB_CUDA(cudaEventRecord(d->ev[2], st));
    if (p.cells && D.n_mid_items) {
        for (const TriGeom& gs : subs) {
            KSB_CUDA(launch_k2_mid(reinterpret_cast<const MidItem*>(P + D.off_mid_items), D.n_mid_items,
                                   reinterpret_cast<const unsigned long long*>(P + D.off_mid_offsets),
                                   reinterpret_cast<const uint32_t*>(P + D.off_mid_sizes),
                                   P + D.off_mid_members, (uint32_t)D.member_bits,
                                   reinterpret_cast<const uint32_t*>(P + D.off_mid_weights), gs, lc));
            d->launches++;
        }
    }
    KSB_CUDA(cudaEventRecord(d->ev[3], st));
    if (p.cells && D.n_zero_colors) {
        KSB_CUDA(launch_kz_touch(D.n_zero_colors, reinterpret_cast<const unsigned long long*>(P + D.off_zero_offsets),
                                 P + D.off_zero_members, (uint32_t)D.member_bits, d->d_touched, g, lc));
        d->launches++;
    }
    KSB_CUDA(cudaEventRecord(d->ev[4], st));
    if (has_gram && d->overlap) {
        KSB_CUDA(cudaStreamWaitEvent(st, d->ev[10], 0));
    } else {
        KSB_CUDA(cudaEventRecord(d->ev[9], st));
        if (has_gram) { const int grc = run_gram(d, g, st, st, d->ev[11]); if (grc) return grc; }
        KSB_CUDA(cudaEventRecord(d->ev[10], st));
    }
    KSB_CUDA(cudaEventRecord(d->ev[8], st));
    if (d->copy_pending) {
        KSB_CUDA(cudaStreamWaitEvent(st, d->ev_copy[3], 0));
        if (d->defer_copy_sync) {
            d->copy_harvest = true;  // the step's scan and unions are enqueued behind this without a host round trip
        } else {
            KSB_CUDA(cudaStreamSynchronize(st));
            d->t.h2d_ms = ev_ms(d->ev_copy[0], d->ev_copy[3]);  // overlaps accumulate_ms
        }
        d->copy_pending = false;
    }
    if (sync) {
        KSB_CUDA(cudaStreamSynchronize(st));
        harvest_accumulate_timers(d);
    }
    return KSB_OK;
}

int ksb_device_read_triangle(const ksb_device* d, uint64_t* host_out, uint64_t cells) {
    if (!d || (!host_out && cells)) { set_error("ksb_device_read_triangle: null argument"); return KSB_ERR_ARG; }
    if (d->spill) { set_error("ksb_device_read_triangle: the spill path has no dense block"); return KSB_ERR_ARG; }
    const ksb_pass& p = d->passes[d->cur_pass];
    if (cells != p.cells) { set_error("ksb_device_read_triangle: the current pass has %llu cells, caller asked for %llu", (unsigned long long)p.cells, (unsigned long long)cells); return KSB_ERR_ARG; }
    KSB_CUDA(cudaSetDevice(d->device));
    KSB_CUDA(cudaStreamSynchronize(d->stream));
    if (cells && d->cell_bits == 32) {  // u32 cells on the device: widened here, in place, from the top down
        KSB_CUDA(cudaMemcpy(host_out, d->d_tri, 4 * cells, cudaMemcpyDeviceToHost));
        const uint32_t* narrow = reinterpret_cast<const uint32_t*>(host_out);
        for (uint64_t i = cells; i-- > 0;) host_out[i] = narrow[i];
    } else if (cells) {
        KSB_CUDA(cudaMemcpy(host_out, d->d_tri, 8 * cells, cudaMemcpyDeviceToHost));
    }
    return KSB_OK;
}

int ksb_device_cell_bits(const ksb_device* d) { return d ? (int)d->cell_bits : 0; }

void* ksb_device_triangle(const ksb_device* d, uint64_t* cells, uint64_t* row_begin, uint64_t* row_end) {
    if (!d) return nullptr;
    const ksb_pass& p = d->passes[d->cur_pass];
    if (cells) *cells = p.cells;
    if (row_begin) *row_begin = p.rb;
    if (row_end) *row_end = p.re;
    return d->d_tri;
}

}  // extern "C"

// ------------------------------------------------------------------------------------ post pass (K5)

extern "C" {

void ksb_result_free(ksb_result* r) {
    if (!r) return;
    cudaSetDevice(r->device);
    for (ksb_segment& s : r->seg) {
        if (s.d_records) cudaFreeAsync(s.d_records, g_ctx[r->device].stream);
        if (s.d_pairs) cudaFreeAsync(s.d_pairs, g_ctx[r->device].stream);
    }
    if (r->h_records) { if (r->h_pageable) free(r->h_records); else pinned_release(r->h_records, sizeof(ksb_record) * r->h_capacity); }
    (void)cudaGetLastError();
    delete r;
}

// Scans the current pass of `d` and appends its rows to *inout (created when *inout == NULL).
int ksb_postpass(ksb_device* d, int want_records, ksb_result** inout) {
    if (!d || !inout) { set_error("ksb_postpass: null argument"); return KSB_ERR_ARG; }
    if (d->spill) { set_error("ksb_postpass: this device runs the spill path (ksb_device_run_spill)"); return KSB_ERR_ARG; }
    KSB_CUDA(cudaSetDevice(d->device));
    ksb_result* r = *inout;
    const bool created = (r == nullptr);
    if (created) {
        r = new (std::nothrow) ksb_result();
        if (!r) { set_error("out of memory"); return KSB_ERR_NOMEM; }
        r->device = d->device;
        r->n_samples = d->desc.n_samples;
        std::memset(&r->info, 0, sizeof r->info);
        r->info.row_begin = d->row_begin;
        r->info.row_end = d->row_end;
        r->info.owned_cells = d->owned_cells;
        r->have_host = want_records != 0;
        r->host_only = want_records == 2;
    }
    auto fail = [&](int code) { if (created) ksb_result_free(r); return code; };
#define KSB_CUDA_R(call)                                                                              \
    do {                                                                                              \
        cudaError_t e_ = (call);                                                                      \
        if (e_ != cudaSuccess) {                                                                      \
            set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__);    \
            (void)cudaGetLastError();                                                                 \
            return fail(KSB_ERR_CUDA);                                                                \
        }                                                                                             \
    } while (0)
    const ksb_pass& p = d->passes[d->cur_pass];
    const TriGeom g = geom_of(d, p);
    const LaunchCfg lc{d->sms, d->stream};
    const uint64_t nb = post_blocks(p.cells);
    unsigned long long* counts = d->d_block_counts;
    unsigned long long* offsets = d->d_block_counts + d->post_blocks_alloc;
    uint64_t n_rows = 0;
    ksb_segment seg;
    KSB_CUDA_R(cudaEventRecord(d->ev[5], d->stream));
    if (nb) {
        KSB_CUDA_R(cudaMemsetAsync(counts + nb, 0, 8, d->stream));
        KSB_CUDA_R(launch_k5_count(g, d->d_touched, counts, lc));
        size_t tb = d->scan_tmp_bytes;
        KSB_CUDA_R(scan_block_counts(d->d_scan_tmp, &tb, counts, offsets, nb + 1, d->stream));
        unsigned long long total = 0;
        KSB_CUDA_R(cudaMemcpyAsync(&total, offsets + nb, 8, cudaMemcpyDeviceToHost, d->stream));
        KSB_CUDA_R(cudaStreamSynchronize(d->stream));
        d->launches += 2;  // k5_count + the scan
        n_rows = total;
        if (n_rows) {
            cudaError_t e = cudaMallocAsync(&seg.d_records, sizeof(ksb_record) * n_rows, d->stream);
            if (e == cudaSuccess) {
                e = cudaMallocAsync(&seg.d_pairs, 12 * n_rows, d->stream);
                if (e != cudaSuccess) { cudaFreeAsync(seg.d_records, d->stream); seg.d_records = nullptr; }
            }
            if (e != cudaSuccess) {
                set_error("device memory for %llu rows of this pass (44 bytes each): %s", (unsigned long long)n_rows, cudaGetErrorString(e));
                (void)cudaGetLastError();
                return fail(e == cudaErrorMemoryAllocation ? KSB_ERR_NOMEM : KSB_ERR_CUDA);
            }
            seg.n = n_rows;
            r->seg.push_back(seg);
            const uint8_t* P = d->d_payload;
            uint32_t* block_rows = reinterpret_cast<uint32_t*>(d->d_block_counts + 2 * d->post_blocks_alloc);
            if (d->block_rows_pass != d->cur_pass) {  // geometry of the pass: computed once, reused by every later post pass
                KSB_CUDA_R(launch_k5_block_rows(g, block_rows, lc));
                d->block_rows_pass = d->cur_pass;
                d->launches++;
            }
            KSB_CUDA_R(launch_k5_emit(g, d->d_touched, offsets, block_rows, reinterpret_cast<const uint32_t*>(P + d->desc.off_totals),
                                      reinterpret_cast<const uint32_t*>(P + d->desc.off_sample_ids), seg.d_records, seg.d_pairs, reinterpret_cast<float*>(seg.d_pairs + n_rows), lc));
            d->launches++;
        }
    }
    KSB_CUDA_R(cudaEventRecord(d->ev[6], d->stream));
    if (r->have_host && n_rows) {
        const uint64_t have = r->info.n_rows, need = have + n_rows;
        if (need > r->h_capacity) {
            const uint64_t cap = std::max<uint64_t>(need, r->h_capacity * 2);
            ksb_record* nh = nullptr;
            uint64_t got = sizeof(ksb_record) * cap;
            static const bool pin_always = [] { const char* v = std::getenv("KSPIDER_RECORDS_PINNED"); return v && v[0] == '1'; }();
            const bool pageable = r->host_only && !pin_always && got > (256ull << 20);
            if (pageable) {
                got = (got + (2ull << 20) - 1) & ~((2ull << 20) - 1);
                void* mem = nullptr;
                if (posix_memalign(&mem, 2ull << 20, got) != 0) { set_error("out of host memory (%llu bytes of records)", (unsigned long long)got); if (created) ksb_result_free(r); return KSB_ERR_NOMEM; }
                madvise(mem, got, MADV_HUGEPAGE);  // 2 MB pages where the system allows: 800 page faults instead of 400 000
                nh = static_cast<ksb_record*>(mem);
            } else {
                KSB_CUDA_R(pinned_acquire(reinterpret_cast<void**>(&nh), &got));
            }
            if (have) std::memcpy(nh, r->h_records, sizeof(ksb_record) * have);
            if (r->h_records) { if (r->h_pageable) free(r->h_records); else pinned_release(r->h_records, sizeof(ksb_record) * r->h_capacity); }
            r->h_pageable = pageable;
            const uint64_t cap_got = got / sizeof(ksb_record);
            r->h_records = nh;
            r->h_capacity = cap_got;
        }
        KSB_CUDA_R(cudaMemcpyAsync(r->h_records + have, seg.d_records, sizeof(ksb_record) * n_rows, cudaMemcpyDeviceToHost, d->stream));
    }
    KSB_CUDA_R(cudaEventRecord(d->ev[7], d->stream));
    KSB_CUDA_R(cudaStreamSynchronize(d->stream));
    if (r->host_only && n_rows) {  // the rows are on the host: 44 bytes per row of device memory go back to the pool
        ksb_segment& sg = r->seg.back();
        cudaFreeAsync(sg.d_records, d->stream);
        cudaFreeAsync(sg.d_pairs, d->stream);
        sg.d_records = nullptr;
        sg.d_pairs = nullptr;
    }
#undef KSB_CUDA_R
    d->t.post_ms += ev_ms(d->ev[5], d->ev[6]);
    d->t.d2h_ms += ev_ms(d->ev[6], d->ev[7]);
    r->info.n_rows += n_rows;
    r->info.n_updates += p.updates;
    r->info.gpu_launches = d->launches;
    r->info.t = d->t;
    *inout = r;
    return KSB_OK;
}

// K3: emission in chunks -> sort + reduce per chunk -> merge into the accumulated run -> records.
int ksb_device_run_spill(ksb_device* d, int want_records, ksb_result** out) {
    if (!d || !out) { set_error("ksb_device_run_spill: null argument"); return KSB_ERR_ARG; }
    if (!d->spill) { set_error("ksb_device_run_spill: this device runs the dense path"); return KSB_ERR_ARG; }
    *out = nullptr;
    KSB_CUDA(cudaSetDevice(d->device));
    cudaStream_t st = d->stream;
    const LaunchCfg lc{d->sms, st};
    const IndexDesc& D = d->desc;
    const uint8_t* P = d->d_payload;
    const uint64_t N = D.n_samples;
    uint32_t shift = 1;
    while ((1ull << shift) < N) shift++;
    const int end_bit = (int)(2 * shift);
    TriGeom g = geom_of(d, d->passes[0]);
    g.tri = nullptr;

    // chunk capacity in updates: 8 + 4 bytes per entry, double-buffered for the sort, + 16 for the reduced run
    const uint64_t biggest = std::max<uint64_t>({(uint64_t)kMidCols * kMidRows, (uint64_t)32 * kMaxPairsSmall,
                                                 D.n_gram_colors || D.n_zero_colors ? D.max_color_size * (D.max_color_size - 1) / 2 : 0});
    uint64_t cap = std::max<uint64_t>(d->budget / 40, 1ull << 20);
    cap = std::min<uint64_t>(cap, std::max<uint64_t>(d->owned_updates + D.n_zero_members * D.max_color_size, 1ull << 20));
    if (cap < biggest) { set_error("dense_budget_bytes too small for the spill path: one work unit emits up to %llu updates", (unsigned long long)biggest); return KSB_ERR_BUDGET; }
    struct Chunk { int kind; uint64_t b, e; };
    std::vector<Chunk> chunks;
    for (uint32_t c = 0; c < d->k1.n_cls; c++) {  // K1 task ranges, class by class
        const K1Class& kc = d->k1.cls[c];
        const uint64_t per_task = (uint64_t)kc.k * kc.s * (kc.s - 1) / 2;
        const uint64_t n_tasks = (kc.n + kc.k - 1) / kc.k, step = std::max<uint64_t>(1, cap / per_task);
        for (uint64_t t = 0; t < n_tasks; t += step) chunks.push_back({0, kc.task_begin + t, kc.task_begin + std::min(n_tasks, t + step)});
    }
    {
        const uint64_t step = std::max<uint64_t>(1, cap / ((uint64_t)kMidCols * kMidRows));
        for (uint64_t t = 0; t < D.n_mid_items; t += step) chunks.push_back({1, t, std::min<uint64_t>(D.n_mid_items, t + step)});
        const uint64_t per = std::max<uint64_t>(1, D.max_color_size * (D.max_color_size - 1) / 2), cstep = std::max<uint64_t>(1, cap / per);
        for (uint64_t t = 0; t < D.n_gram_colors; t += cstep) chunks.push_back({2, t, std::min<uint64_t>(D.n_gram_colors, t + cstep)});
        for (uint64_t t = 0; t < D.n_zero_colors; t += cstep) chunks.push_back({3, t, std::min<uint64_t>(D.n_zero_colors, t + cstep)});
    }

    unsigned long long *keys = nullptr, *keys2 = nullptr, *uniq = nullptr, *sums = nullptr, *d_cnt = nullptr;
    uint32_t *vals = nullptr, *vals2 = nullptr;
    void* tmp = nullptr;
    size_t tmp_bytes = 0;
    // sorted runs of unique keys, oldest first; neighbours are merged while their sizes are within a factor of two, so an
    // entry takes part in O(log chunks) merges (merge path + reduce-by-key) instead of one re-sort of everything per chunk
    struct Run { unsigned long long* k; unsigned long long* s; uint64_t n; };
    std::vector<Run> runs;
    unsigned long long *acc_k = nullptr, *acc_s = nullptr;  // the final run
    uint64_t acc_n = 0;
    ksb_result* r = nullptr;
    ksb_segment seg;
    int rc = KSB_OK;
    cudaError_t e = cudaSuccess;
    cudaEvent_t e0 = g_ctx[d->device].e0, e1 = g_ctx[d->device].e1;
    float emit_ms = 0.f, sort_ms = 0.f;
#define SP(call) do { e = (call); if (e != cudaSuccess) goto cuda_fail; } while (0)
    // merges the two newest runs into one (in place on the stack); false + e set on a CUDA error
    auto merge_top = [&]() -> bool {
        Run b = runs.back(); runs.pop_back();
        Run a = runs.back(); runs.pop_back();
        const uint64_t tot = a.n + b.n;
        unsigned long long *mk = nullptr, *ms = nullptr, *part = nullptr, *nk = nullptr, *ns = nullptr;
        void* mtmp = nullptr;
        size_t mtb = 0;
        bool ok = false;
        unsigned long long mm = 0;
        do {
            if ((e = cub_reduce_bytes(tot, &mtb)) != cudaSuccess) break;
            if ((e = cudaMallocAsync(&mk, 8 * tot, st)) != cudaSuccess) break;
            if ((e = cudaMallocAsync(&ms, 8 * tot, st)) != cudaSuccess) break;
            if ((e = cudaMallocAsync(&part, 8 * spill_merge_part_entries(tot), st)) != cudaSuccess) break;
            if ((e = cudaMallocAsync(&nk, 8 * tot, st)) != cudaSuccess) break;
            if ((e = cudaMallocAsync(&ns, 8 * tot, st)) != cudaSuccess) break;
            if ((e = cudaMallocAsync(&mtmp, std::max<size_t>(mtb, 256), st)) != cudaSuccess) break;
            if ((e = spill_merge_path(a.k, a.s, a.n, b.k, b.s, b.n, part, mk, ms, st)) != cudaSuccess) break;
            if ((e = spill_reduce_sorted(mtmp, mtb, mk, ms, tot, nk, ns, d_cnt + 1, st)) != cudaSuccess) break;
            if ((e = cudaMemcpyAsync(&mm, d_cnt + 1, 8, cudaMemcpyDeviceToHost, st)) != cudaSuccess) break;
            if ((e = cudaStreamSynchronize(st)) != cudaSuccess) break;
            ok = true;
        } while (0);
        d->launches += 3;
        for (void* q : {(void*)mk, (void*)ms, (void*)part, mtmp, (void*)a.k, (void*)a.s, (void*)b.k, (void*)b.s})
            if (q) cudaFreeAsync(q, st);
        if (!ok) { if (nk) cudaFreeAsync(nk, st); if (ns) cudaFreeAsync(ns, st); return false; }
        runs.push_back(Run{nk, ns, mm});
        return true;
    };
    SP(spill_temp_bytes(cap, end_bit, &tmp_bytes));
    SP(cudaMallocAsync(&keys, 8 * cap, st));
    SP(cudaMallocAsync(&keys2, 8 * cap, st));
    SP(cudaMallocAsync(&vals, 4 * cap, st));
    SP(cudaMallocAsync(&vals2, 4 * cap, st));
    SP(cudaMallocAsync(&uniq, 8 * cap, st));
    SP(cudaMallocAsync(&sums, 8 * cap, st));
    SP(cudaMallocAsync(&d_cnt, 16, st));
    SP(cudaMallocAsync(&tmp, std::max<size_t>(tmp_bytes, 256), st));
    for (const Chunk& ch : chunks) {
        const SpillBuf sb{keys, vals, d_cnt, shift};
        SP(cudaMemsetAsync(d_cnt, 0, 16, st));
        SP(cudaEventRecord(e0, st));
        if (ch.kind == 0)
            SP(launch_k1_spill(d->k1, P + D.off_small_members, (uint32_t)D.packed_bits,
                               P + D.off_small_weights, g, sb, ch.b, ch.e, lc));
        else if (ch.kind == 1)
            SP(launch_k2_spill(reinterpret_cast<const MidItem*>(P + D.off_mid_items), ch.b, ch.e,
                               reinterpret_cast<const unsigned long long*>(P + D.off_mid_offsets),
                               reinterpret_cast<const uint32_t*>(P + D.off_mid_sizes),
                               P + D.off_mid_members, (uint32_t)D.member_bits,
                               reinterpret_cast<const uint32_t*>(P + D.off_mid_weights), g, sb, lc));
        else if (ch.kind == 2 && D.gram_form)
            SP(launch_k3_emit_bitmap(ch.b, ch.e, reinterpret_cast<const GramWin*>(P + D.off_gram_wins), (uint32_t)D.n_gram_wins,
                                     reinterpret_cast<const unsigned long long*>(P + D.off_gram_offsets),
                                     reinterpret_cast<const uint32_t*>(P + D.off_gram_members), P + D.off_gram_weights,
                                     (uint32_t)D.weight_bits, g, sb, lc));
        else if (ch.kind == 2)
            SP(launch_k3_emit_csr(ch.b, ch.e, reinterpret_cast<const unsigned long long*>(P + D.off_gram_offsets),
                                  P + D.off_gram_members, (uint32_t)gram_member_bits(D),
                                  P + D.off_gram_weights, (uint32_t)D.weight_bits, g, sb, lc,
                                  D.gram_rel_bits ? reinterpret_cast<const GramWin*>(P + D.off_gram_wins) : nullptr, (uint32_t)D.n_gram_wins));
        else
            SP(launch_k3_emit_csr(ch.b, ch.e, reinterpret_cast<const unsigned long long*>(P + D.off_zero_offsets),
                                  P + D.off_zero_members, (uint32_t)D.member_bits, nullptr, 32u, g, sb, lc));
        SP(cudaEventRecord(e1, st));
        d->launches++;
        unsigned long long n = 0;
        SP(cudaMemcpyAsync(&n, d_cnt, 8, cudaMemcpyDeviceToHost, st));
        SP(cudaStreamSynchronize(st));
        emit_ms += ev_ms(e0, e1);
        if (n > cap) { set_error("spill chunk overflow: %llu > %llu", n, (unsigned long long)cap); rc = KSB_ERR_CUDA; goto done; }
        if (n == 0) continue;
        SP(cudaEventRecord(e0, st));
        SP(spill_sort_reduce_chunk(tmp, tmp_bytes, keys, keys2, vals, vals2, n, end_bit, uniq, sums, d_cnt + 1, st));
        unsigned long long m = 0;
        SP(cudaMemcpyAsync(&m, d_cnt + 1, 8, cudaMemcpyDeviceToHost, st));
        SP(cudaStreamSynchronize(st));
        d->launches += 2;
        {
            Run nr{nullptr, nullptr, m};
            SP(cudaMallocAsync(&nr.k, 8 * std::max<uint64_t>(m, 1), st));
            e = cudaMallocAsync(&nr.s, 8 * std::max<uint64_t>(m, 1), st);
            if (e != cudaSuccess) { cudaFreeAsync(nr.k, st); goto cuda_fail; }
            runs.push_back(nr);
            SP(cudaMemcpyAsync(nr.k, uniq, 8 * m, cudaMemcpyDeviceToDevice, st));
            SP(cudaMemcpyAsync(nr.s, sums, 8 * m, cudaMemcpyDeviceToDevice, st));
        }
        while (runs.size() >= 2 && runs[runs.size() - 2].n <= 2 * runs.back().n)
            if (!merge_top()) goto cuda_fail;
        SP(cudaEventRecord(e1, st));
        SP(cudaStreamSynchronize(st));
        sort_ms += ev_ms(e0, e1);
    }
    SP(cudaEventRecord(e0, st));
    while (runs.size() >= 2)
        if (!merge_top()) goto cuda_fail;
    SP(cudaEventRecord(e1, st));
    SP(cudaStreamSynchronize(st));
    sort_ms += ev_ms(e0, e1);
    if (!runs.empty()) { acc_k = runs[0].k; acc_s = runs[0].s; acc_n = runs[0].n; runs.clear(); }
    // ---- records from the final run
    r = new (std::nothrow) ksb_result();
    if (!r) { set_error("out of memory"); rc = KSB_ERR_NOMEM; goto done; }
    r->device = d->device;
    r->n_samples = N;
    std::memset(&r->info, 0, sizeof r->info);
    r->info.row_begin = d->row_begin;
    r->info.row_end = d->row_end;
    r->info.owned_cells = d->owned_cells;
    r->have_host = want_records != 0;
    SP(cudaEventRecord(e0, st));
    if (acc_n) {
        SP(cudaMallocAsync(&seg.d_records, sizeof(ksb_record) * acc_n, st));
        SP(cudaMallocAsync(&seg.d_pairs, 12 * acc_n, st));
        seg.n = acc_n;
        r->seg.push_back(seg);
        SP(launch_k5_sparse(acc_k, acc_s, acc_n, shift, reinterpret_cast<const uint32_t*>(P + D.off_totals),
                            reinterpret_cast<const uint32_t*>(P + D.off_sample_ids), seg.d_records, seg.d_pairs, reinterpret_cast<float*>(seg.d_pairs + acc_n), lc));
        d->launches++;
        if (r->have_host) {
            uint64_t got = sizeof(ksb_record) * acc_n;
            SP(pinned_acquire(reinterpret_cast<void**>(&r->h_records), &got));
            r->h_capacity = got / sizeof(ksb_record);
            SP(cudaMemcpyAsync(r->h_records, seg.d_records, sizeof(ksb_record) * acc_n, cudaMemcpyDeviceToHost, st));
        }
    }
    SP(cudaEventRecord(e1, st));
    SP(cudaStreamSynchronize(st));
    d->t.post_ms += ev_ms(e0, e1);
    d->t.spill_emit_ms += emit_ms;
    d->t.spill_sort_ms += sort_ms;
    d->t.accumulate_ms += emit_ms + sort_ms;
    r->info.n_rows = acc_n;
    r->info.n_updates = d->owned_updates;
    r->info.gpu_launches = d->launches;
    r->info.t = d->t;
    goto done;
cuda_fail:
    set_error("spill path: %s", cudaGetErrorString(e));
    (void)cudaGetLastError();
    rc = KSB_ERR_CUDA;
done:
#undef SP
    for (void* p : {(void*)keys, (void*)keys2, (void*)vals, (void*)vals2, (void*)uniq, (void*)sums, (void*)d_cnt, tmp, (void*)acc_k, (void*)acc_s})
        if (p) cudaFreeAsync(p, st);
    for (const Run& q : runs) { if (q.k) cudaFreeAsync(q.k, st); if (q.s) cudaFreeAsync(q.s, st); }
    cudaStreamSynchronize(st);
    if (rc != KSB_OK) { if (r) ksb_result_free(r); return rc; }
    *out = r;
    return KSB_OK;
}

float ksb_threshold_float_pct(double cut);
int ksb_cluster_pct(const ksb_result* r, int dist_col, double cut_pct, uint32_t* labels_out, uint64_t* n_edges);
int ksb_cluster_pct_device(const ksb_result* r, int dist_col, double cut_pct, uint32_t* d_labels_out, uint64_t* n_edges);
int ksb_device_get_info(const ksb_device* d, ksb_result_info* info);

// Cluster-only step (north_star 3): for every dense pass, accumulate and scan its block ONCE (K56: containment,
// threshold, append of the kept pairs only), then the unions over the kept-edge list and the label compression --
// everything enqueued on one stream without a host round trip; a single synchronisation at the end brings the
// labels (and the two counters) to the host.  No records are built.  The edge list has a capacity (an eighth of
// the owned cells at first); when a step keeps more edges than fit, the list is grown to the exact need and the
// step is repeated -- the only case with a second synchronisation.
// The enqueued work of a step is captured ONCE into a CUDA graph (both streams of the accumulation, the memsets,
// every kernel, the label / counter copies) and replayed by one cudaGraphLaunch per step: at 8 GPUs the kernels of
// a step are ~20 us each and ~14 launches + ~20 event records of host API time would otherwise BE the step.
static int cluster_prepare(ksb_device* d, cudaStream_t st) {
    const uint32_t N = (uint32_t)d->desc.n_samples;
    if (!d->d_parent) {
        KSB_CUDA(cudaMallocAsync(&d->d_parent, 4ull * N, st));
        KSB_CUDA(cudaMallocAsync(&d->d_cl_counters, 32, st));
        uint64_t got = 4ull * N + 16;
        void* hp = nullptr;
        KSB_CUDA(pinned_acquire(&hp, &got));
        d->h_stage = static_cast<uint8_t*>(hp);
        d->h_stage_bytes = got;
    }
    if (!d->d_edges) {
        if (d->edge_cap == 0) {
            const char* ec = std::getenv("KSPIDER_EDGE_CAP");  // test knob: a short list forces the grow-and-repeat path
            d->edge_cap = (ec && std::atoll(ec) > 0) ? (uint64_t)std::atoll(ec) : std::max<uint64_t>(1ull << 16, d->owned_cells / 8);
        }
        KSB_CUDA(cudaMallocAsync(&d->d_edges, 8 * d->edge_cap, st));
    }
    if (d->desc.n_gram_colors)
        for (uint32_t p = 0; p < d->passes.size(); p++)
            if (d->passes[p].cells) {  // allocates and copies: never inside a capture
                const int rc = plan_gram(d, p);
                if (rc) return rc;
                ksb_gram_plan& gp = d->gplan[p];
                if (!gp.tables_ordered && gp.ev_tables) { KSB_CUDA(cudaStreamWaitEvent(st, gp.ev_tables, 0)); gp.tables_ordered = true; }
            }
    return KSB_OK;
}

static int cluster_body(ksb_device* d, int dist_col, float cstar, cudaStream_t st, uint32_t* d_labels_out, bool host_labels) {
    const uint32_t N = (uint32_t)d->desc.n_samples;
    const LaunchCfg lc{d->sms, st};
    const uint8_t* P = d->d_payload;
    uint32_t* block_rows = reinterpret_cast<uint32_t*>(d->d_block_counts + 2 * d->post_blocks_alloc);
    KSB_CUDA(cudaMemsetAsync(d->d_cl_counters, 0, 32, st));
    KSB_CUDA(launch_k7_init(d->d_parent, N, lc));
    d->launches++;
    bool all_rows = true;
    for (uint32_t p = 0; p < d->passes.size(); p++) {
        d->cur_pass = p;
        d->defer_copy_sync = true;
        const int rc = ksb_accumulate(d, st, 0);
        d->defer_copy_sync = false;
        if (rc) return rc;
        const ksb_pass& ps = d->passes[p];
        if (ps.cells == 0) continue;
        const TriGeom g = geom_of(d, ps);
        const bool rows = k56_uses_rows(g);  // row-wise scan: no block table, and the kept edges are hooked as they are appended
        all_rows = all_rows && rows;
        if (!rows && (d->block_rows_pass != (int)p || d->passes.size() > 1)) {
            KSB_CUDA(launch_k5_block_rows(g, block_rows, lc));
            d->block_rows_pass = (int)p;
            d->launches++;
        }
        KSB_CUDA(cudaEventRecord(d->ev[5], st));
        KSB_CUDA(launch_k56_scan(g, d->d_touched, block_rows, reinterpret_cast<const uint32_t*>(P + d->desc.off_totals), dist_col, cstar,
                                 d->d_edges, d->edge_cap, d->d_cl_counters, d->d_parent, lc));
        KSB_CUDA(cudaEventRecord(d->ev[6], st));
        d->launches++;
    }
    KSB_CUDA(launch_k6_edges(d->d_edges, d->edge_cap, d->d_cl_counters, d->d_parent, d->edge_cap, all_rows, lc));
    KSB_CUDA(launch_k7_compress(d->d_parent, N, lc));
    KSB_CUDA(cudaEventRecord(d->ev[7], st));
    d->launches += all_rows ? 2 : 3;
    if (d_labels_out) KSB_CUDA(cudaMemcpyAsync(d_labels_out, d->d_parent, 4ull * N, cudaMemcpyDeviceToDevice, st));
    if (host_labels) KSB_CUDA(cudaMemcpyAsync(d->h_stage, d->d_parent, 4ull * N, cudaMemcpyDeviceToHost, st));
    KSB_CUDA(cudaMemcpyAsync(d->h_stage + 4ull * N, d->d_cl_counters, 16, cudaMemcpyDeviceToHost, st));
    return KSB_OK;
}

int ksb_device_cluster_enqueue(ksb_device* d, int dist_col, double cut_pct, void* stream_v, uint32_t* d_labels_out, int host_labels) {
    if (!d) { set_error("ksb_device_cluster: null device"); return KSB_ERR_ARG; }
    if (d->spill) { set_error("ksb_device_cluster: this device runs the spill path (ksb_device_run_spill + ksb_cluster)"); return KSB_ERR_ARG; }
    if (dist_col < 3 || dist_col > 5) { set_error("ksb_device_cluster: dist_col must be 3 (min), 4 (avg) or 5 (max)"); return KSB_ERR_ARG; }
    KSB_CUDA(cudaSetDevice(d->device));
    cudaStream_t st = stream_v ? static_cast<cudaStream_t>(stream_v) : d->stream;
    if (d->desc.n_samples == 0) return KSB_OK;
    float cstar = ksb_threshold_float_pct(cut_pct);
    if (cstar != cstar) cstar = INFINITY;
    int rc = cluster_prepare(d, st);
    if (rc) return rc;
    d->cl_stream = st;
    d->cl_host_labels = host_labels != 0;
    d->cl_was_graph = false;
    const bool can_graph = d->use_graph && !d->copy_pending;
    if (!can_graph) return cluster_body(d, dist_col, cstar, st, d_labels_out, host_labels != 0);
    uint32_t cbits;
    std::memcpy(&cbits, &cstar, 4);
    const ksb_device::graph_key key{dist_col, cbits, d_labels_out, host_labels != 0, d->overlap, d->edge_cap, d->d_edges};
    if (!d->cl_graph || std::memcmp(&key, &d->cl_key, sizeof key) != 0) {
        if (d->cl_graph) { cudaGraphExecDestroy(d->cl_graph); d->cl_graph = nullptr; }
        const uint64_t launches_before = d->launches;
        KSB_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
        rc = cluster_body(d, dist_col, cstar, st, d_labels_out, host_labels != 0);
        cudaGraph_t graph = nullptr;
        cudaError_t e = cudaStreamEndCapture(st, &graph);
        if (rc) { if (graph) cudaGraphDestroy(graph); (void)cudaGetLastError(); return rc; }
        if (e != cudaSuccess || !graph) {
            // capture refused (an API call inside the step is not capturable on this driver): run the step eagerly from now on
            (void)cudaGetLastError();
            if (graph) cudaGraphDestroy(graph);
            d->use_graph = false;
            d->launches = launches_before;
            return cluster_body(d, dist_col, cstar, st, d_labels_out, host_labels != 0);
        }
        e = cudaGraphInstantiate(&d->cl_graph, graph, 0);
        cudaGraphDestroy(graph);
        if (e != cudaSuccess) { (void)cudaGetLastError(); d->cl_graph = nullptr; d->use_graph = false; d->launches = launches_before; return cluster_body(d, dist_col, cstar, st, d_labels_out, host_labels != 0); }
        d->cl_key = key;
        d->cl_graph_launches = d->launches - launches_before;
        d->launches = launches_before;
    }
    KSB_CUDA(cudaGraphLaunch(d->cl_graph, st));
    d->launches += d->cl_graph_launches;
    d->cl_was_graph = true;
    return KSB_OK;
}

// Waits for the step enqueued last and hands out its result.  KSB_ERR_BUDGET: the kept-edge list was too short; it
// has been grown to the exact need -- enqueue the step again.
int ksb_device_cluster_collect(ksb_device* d, uint32_t* labels_out, uint64_t* n_edges, uint64_t* n_rows) {
    if (!d) { set_error("ksb_device_cluster_collect: null device"); return KSB_ERR_ARG; }
    const uint32_t N = (uint32_t)d->desc.n_samples;
    if (n_edges) *n_edges = 0;
    if (n_rows) *n_rows = 0;
    if (N == 0) return KSB_OK;
    if (!d->cl_stream || !d->h_stage) { set_error("ksb_device_cluster_collect: no step was enqueued"); return KSB_ERR_ARG; }
    KSB_CUDA(cudaSetDevice(d->device));
    KSB_CUDA(cudaStreamSynchronize(d->cl_stream));
    if (d->copy_harvest) { d->t.h2d_ms = ev_ms(d->ev_copy[0], d->ev_copy[3]); d->copy_harvest = false; }
    unsigned long long cnt[2];
    std::memcpy(cnt, d->h_stage + 4ull * N, 16);
    if (cnt[1] > d->edge_cap) {  // the list was too short: now the need is known exactly
        KSB_CUDA(cudaFreeAsync(d->d_edges, d->cl_stream));
        d->d_edges = nullptr;
        d->edge_cap = cnt[1] + cnt[1] / 16 + 1024;
        set_error("ksb_device_cluster: the kept-edge list was too short (%llu edges); it has been grown, repeat the step", cnt[1]);
        return KSB_ERR_BUDGET;
    }
    if (!d->cl_was_graph) {  // events recorded inside a graph carry no timestamps
        harvest_accumulate_timers(d);  // of the last pass
        d->t.post_ms += ev_ms(d->ev[5], d->ev[6]);
        d->t.cluster_ms += ev_ms(d->ev[6], d->ev[7]);
    }
    if (labels_out) {
        if (!d->cl_host_labels) { set_error("ksb_device_cluster_collect: the step was enqueued without host labels"); return KSB_ERR_ARG; }
        std::memcpy(labels_out, d->h_stage, 4ull * N);
    }
    if (n_rows) *n_rows = cnt[0];
    if (n_edges) *n_edges = cnt[1];
    return KSB_OK;
}

int ksb_device_cluster(ksb_device* d, int dist_col, double cut_pct, void* stream_v, uint32_t* labels_out,
                       uint32_t* d_labels_out, uint64_t* n_edges, uint64_t* n_rows) {
    for (int attempt = 0; attempt < 2; attempt++) {
        int rc = ksb_device_cluster_enqueue(d, dist_col, cut_pct, stream_v, d_labels_out, labels_out != nullptr);
        if (rc) return rc;
        rc = ksb_device_cluster_collect(d, labels_out, n_edges, n_rows);
        if (rc != KSB_ERR_BUDGET) return rc;
    }
    set_error("ksb_device_cluster: edge list overflow after growing");
    return KSB_ERR_CUDA;
}

int ksb_device_set_graph(ksb_device* d, int on) {
    if (!d) { set_error("ksb_device_set_graph: null device"); return KSB_ERR_ARG; }
    d->use_graph = on != 0;
    return KSB_OK;
}

static int pairwise_once(const ksb_index* ix, const ksb_opts* opts, ksb_result** out) {
    using Clock = std::chrono::steady_clock;
    auto ms_since = [](Clock::time_point t) { return std::chrono::duration<float, std::milli>(Clock::now() - t).count(); };
    ksb_device* d = nullptr;
    auto t0 = Clock::now();
    int rc = upload_impl(ix, opts, &d, true);
    if (rc) return rc;
    const float setup_ms = ms_since(t0);
    float acc_ms = 0.f, post_ms = 0.f;
    ksb_result* r = nullptr;
    const int want = opts ? opts->want_records : 1;
    if (d->spill) rc = ksb_device_run_spill(d, want, &r);
    for (int p = 0; !d->spill && p < (int)d->passes.size() && rc == KSB_OK; p++) {
        rc = ksb_device_set_pass(d, p);
        t0 = Clock::now();
        if (rc == KSB_OK) rc = ksb_accumulate(d, nullptr, 1);
        acc_ms += ms_since(t0);
        t0 = Clock::now();
        if (rc == KSB_OK) rc = ksb_postpass(d, want, &r);
        post_ms += ms_since(t0);
    }
    t0 = Clock::now();
    ksb_device_free(d);
    if (rc) { ksb_result_free(r); return rc; }
    r->info.t.host_setup_ms = setup_ms;
    r->info.t.host_accumulate_ms = acc_ms;
    r->info.t.host_post_ms = post_ms;
    r->info.t.host_free_ms = ms_since(t0);
    *out = r;
    return KSB_OK;
}

// One retry policy for the whole call: when a pass runs out of device memory (its rows did not fit beside its cells), the
// run is repeated with a quarter of the dense budget -- more, smaller passes -- instead of failing where the reference,
// which only needs host memory, succeeds.
int ksb_pairwise(const ksb_index* ix, const ksb_opts* opts, ksb_result** out) {
    if (!ix || !out) { set_error("ksb_pairwise: null argument"); return KSB_ERR_ARG; }
    *out = nullptr;
    ksb_opts o{};
    if (opts) o = *opts; else { o.device = -1; o.want_records = 1; }
    int rc = pairwise_once(ix, &o, out);
    for (int attempt = 0; rc == KSB_ERR_NOMEM && attempt < 3; attempt++) {
        if (o.dense_budget_bytes == 0) {
            int dev = o.device;
            if (dev >= 0) cudaSetDevice(dev);
            size_t fr = 0, tot = 0;
            if (cudaMemGetInfo(&fr, &tot) != cudaSuccess) { (void)cudaGetLastError(); break; }
            o.dense_budget_bytes = (uint64_t)((double)fr * 0.8 * 8.0 / 52.0);
        }
        o.dense_budget_bytes = std::max<uint64_t>(o.dense_budget_bytes / 4, 1ull << 20);
        rc = pairwise_once(ix, &o, out);
    }
    return rc;
}

// upload + cluster-only step + release in one call: the end-to-end path when only the labels are wanted
// (host index in, N labels out; H2D = the blob, D2H = 4 N + 16 bytes).  A device that has to spill falls back to
// the record path on the device (sort-reduce run, then K6 + K7 over its records).
// ksb_pairwise_cluster keeps the device handle of its last call on an index -- allocations, row passes, K4 plan and
// its tables, streams and events; nothing of the RESULT -- in the index (ksb_index::dev_cache; released with it).  The next
// call with the same options uploads the blob again (the inputs travel host -> device in every call) and runs: what a caller
// pays per call is the copy and the kernels, not 0.3-0.8 ms of host-side planning that depends on how busy the host is.
// KSPIDER_E2E_CACHE=0 disables.
struct E2eCache {
    std::mutex mu;
    ksb_device* dev = nullptr;
    bool in_use = false;
    ksb_opts key{};
};
static void e2e_cache_free(void* p) {
    E2eCache* c = static_cast<E2eCache*>(p);
    if (!c) return;
    if (c->dev) ksb_device_free(c->dev);
    delete c;
}
static std::mutex g_e2e_mu;  // guards the creation of an index's cache object
static E2eCache* e2e_cache_of(const ksb_index* ix) {
    static const bool on = [] { const char* v = std::getenv("KSPIDER_E2E_CACHE"); return !(v && v[0] == '0'); }();
    if (!on || !ix->pinned) return nullptr;
    std::lock_guard<std::mutex> lk(g_e2e_mu);
    if (!ix->dev_cache) {
        ix->dev_cache = new (std::nothrow) E2eCache();
        ix->dev_cache_free = e2e_cache_free;
    }
    return static_cast<E2eCache*>(ix->dev_cache);
}
static bool same_opts(const ksb_opts& a, const ksb_opts& b) {
    return a.device == b.device && a.rank == b.rank && a.world == b.world && a.dense_budget_bytes == b.dense_budget_bytes && a.spill_mode == b.spill_mode;
}

int ksb_pairwise_cluster(const ksb_index* ix, const ksb_opts* opts, int dist_col, double cut_pct, uint32_t* labels_out,
                         uint32_t* d_labels_out, uint64_t* n_edges, uint64_t* n_rows, ksb_result_info* info_out) {
    if (!ix || (!labels_out && !d_labels_out)) { set_error("ksb_pairwise_cluster: null argument"); return KSB_ERR_ARG; }
    using Clock = std::chrono::steady_clock;
    auto ms_since = [](Clock::time_point t) { return std::chrono::duration<float, std::milli>(Clock::now() - t).count(); };
    ksb_device* d = nullptr;
    auto t0 = Clock::now();
    ksb_opts key{};
    if (opts) key = *opts;
    else key.device = -1;
    if (key.device < 0) { int cur = 0; if (cudaGetDevice(&cur) == cudaSuccess) key.device = cur; else (void)cudaGetLastError(); }
    if (key.world <= 1) { key.world = 1; key.rank = 0; }
    E2eCache* cache = e2e_cache_of(ix);
    bool cached = false, holds_slot = false;
    if (cache) {
        std::lock_guard<std::mutex> lk(cache->mu);
        if (!cache->in_use) {
            cache->in_use = holds_slot = true;
            if (cache->dev && same_opts(cache->key, key)) { d = cache->dev; cache->dev = nullptr; cached = true; }
            else if (cache->dev) { ksb_device_free(cache->dev); cache->dev = nullptr; }
        }
    }
    auto release_slot = [&](ksb_device* keep) {
        if (!holds_slot) { if (keep) ksb_device_free(keep); return; }
        std::lock_guard<std::mutex> lk(cache->mu);
        cache->dev = keep;
        cache->key = key;
        cache->in_use = false;
    };
    int rc = KSB_OK;
    if (cached) {
        // the kept handle: fresh timers, the blob on its way again, the first pass's K4 chunks mapped to the copy pieces
        rc = cudaSetDevice(d->device) == cudaSuccess ? KSB_OK : KSB_ERR_CUDA;
        std::memset(&d->t, 0, sizeof d->t);
        d->launches = 0;
        d->cur_pass = 0;
        if (rc == KSB_OK) rc = enqueue_copy_all(d, ix);
        if (rc == KSB_OK) map_gram_pieces(d, ix);
        if (rc != KSB_OK) { release_slot(nullptr); ksb_device_free(d); return rc; }
    } else {
        rc = upload_impl(ix, opts, &d, true);
        if (rc) { release_slot(nullptr); return rc; }
    }
    const float setup_ms = ms_since(t0);
    t0 = Clock::now();
    ksb_result_info info;
    std::memset(&info, 0, sizeof info);
    if (d->spill) {
        ksb_result* r = nullptr;
        rc = ksb_device_run_spill(d, 0, &r);
        if (rc == KSB_OK && labels_out) rc = ksb_cluster_pct(r, dist_col, cut_pct, labels_out, n_edges);
        if (rc == KSB_OK && d_labels_out) rc = ksb_cluster_pct_device(r, dist_col, cut_pct, d_labels_out, n_edges);
        if (rc == KSB_OK) { info = r->info; if (n_rows) *n_rows = r->info.n_rows; }
        ksb_result_free(r);
    } else {
        uint64_t nr = 0;
        rc = ksb_device_cluster(d, dist_col, cut_pct, nullptr, labels_out, d_labels_out, n_edges, &nr);
        if (rc == KSB_OK) {
            ksb_device_get_info(d, &info);
            info.n_rows = nr;
            if (n_rows) *n_rows = nr;
        }
    }
    const float step_ms = ms_since(t0);
    t0 = Clock::now();
    if (rc == KSB_OK && !d->spill && holds_slot) release_slot(d);  // kept for the next call on this index
    else { release_slot(nullptr); ksb_device_free(d); }
    if (rc) return rc;
    info.t.host_setup_ms = setup_ms;
    info.t.host_accumulate_ms = step_ms;
    info.t.host_free_ms = ms_since(t0);
    if (info_out) *info_out = info;
    return KSB_OK;
}

int ksb_device_get_info(const ksb_device* d, ksb_result_info* info) {
    if (!d || !info) { set_error("null argument"); return KSB_ERR_ARG; }
    std::memset(info, 0, sizeof *info);
    info->n_updates = d->owned_updates;
    info->row_begin = d->row_begin;
    info->row_end = d->row_end;
    info->owned_cells = d->owned_cells;
    info->gpu_launches = d->launches;
    info->t = d->t;
    return KSB_OK;
}

int ksb_result_get_info(const ksb_result* r, ksb_result_info* info) {
    if (!r || !info) { set_error("null argument"); return KSB_ERR_ARG; }
    *info = r->info;
    return KSB_OK;
}

const ksb_record* ksb_result_records(const ksb_result* r, uint64_t* n_rows) {
    if (!r) return nullptr;
    if (n_rows) *n_rows = r->have_host ? r->info.n_rows : 0;
    return r->have_host ? r->h_records : nullptr;
}

}  // extern "C"

// ------------------------------------------------------------------------- threshold + CC (K6 + K7)

extern "C" {

int ksb_format_float(float v, char* buf);

// ks_clustering.py:97-105: the edge is dropped iff float(text) * 100 < float(cutoff) * 100 where
// text is the "%g" rendering of the float32 containment.  Both steps are monotone non-decreasing
// in the float, so the kept set is {c >= c*}: binary search of the bit pattern over [+0, +inf].
float ksb_threshold_float_pct(double cut);
float ksb_threshold_float(double cutoff01) { return ksb_threshold_float_pct(cutoff01 * 100.0); }

// cut = the reference's cut_off_threshold, already in percent (= float(cutoff) * 100)
float ksb_threshold_float_pct(double cut) {
    static thread_local double last_cut = 0.0;   // the search below formats ~31 floats: keep the last answer
    static thread_local float last_val = 0.0f;
    static thread_local bool have_last = false;
    if (have_last && std::memcmp(&last_cut, &cut, sizeof cut) == 0) return last_val;
    auto remember = [&](float v) { last_cut = cut; last_val = v; have_last = true; return v; };
    auto kept = [&](uint32_t bits) {
        float c;
        std::memcpy(&c, &bits, 4);
        char buf[32];
        const int n = ksb_format_float(c, buf);
        buf[n] = 0;
        const double distance = strtod(buf, nullptr) * 100.0;
        return !(distance < cut);
    };
    uint32_t lo = 0, hi = 0x7F800000u;  // +0 .. +inf
    if (kept(lo)) return remember(0.0f);
    if (!kept(hi)) return remember(std::nanf(""));  // cutoff above +inf: nothing but NaN rows survives
    while (hi - lo > 1) {  // invariant: !kept(lo), kept(hi)
        const uint32_t mid = lo + (hi - lo) / 2;
        if (kept(mid)) hi = mid; else lo = mid;
    }
    float c;
    std::memcpy(&c, &hi, 4);
    return remember(c);
}

int ksb_cluster_pct(const ksb_result* r, int dist_col, double cut_pct, uint32_t* labels_out, uint64_t* n_edges);
int ksb_cluster(const ksb_result* r, int dist_col, double cutoff01, uint32_t* labels_out, uint64_t* n_edges) {
    return ksb_cluster_pct(r, dist_col, cutoff01 * 100.0, labels_out, n_edges);
}

static int cluster_core(const ksb_result* r, int dist_col, double cut_pct, uint32_t* labels_out, uint32_t* d_labels_out,
                        uint64_t* n_edges);
int ksb_cluster_pct(const ksb_result* r, int dist_col, double cut_pct, uint32_t* labels_out, uint64_t* n_edges) {
    if (!labels_out) { set_error("ksb_cluster: null argument"); return KSB_ERR_ARG; }
    return cluster_core(r, dist_col, cut_pct, labels_out, nullptr, n_edges);
}
int ksb_cluster_pct_device(const ksb_result* r, int dist_col, double cut_pct, uint32_t* d_labels_out, uint64_t* n_edges) {
    if (!d_labels_out) { set_error("ksb_cluster_pct_device: null argument"); return KSB_ERR_ARG; }
    return cluster_core(r, dist_col, cut_pct, nullptr, d_labels_out, n_edges);
}
// labels land in host memory (labels_out) or stay on the device (d_labels_out, N entries on r's device)
static int cluster_core(const ksb_result* r, int dist_col, double cut_pct, uint32_t* labels_out, uint32_t* d_labels_out,
                        uint64_t* n_edges) {
    if (!r) { set_error("ksb_cluster: null argument"); return KSB_ERR_ARG; }
    if (dist_col < 3 || dist_col > 5) { set_error("ksb_cluster: dist_col must be 3 (min), 4 (avg) or 5 (max)"); return KSB_ERR_ARG; }
    if (r->host_only) { set_error("ksb_cluster: this result was made with want_records = 2 (host records only); its device copies are gone"); return KSB_ERR_ARG; }
    int sms = 0;
    const int dev = check_device(r->device, &sms);
    if (dev < 0) return dev;
    const uint32_t N = (uint32_t)r->n_samples;
    if (n_edges) *n_edges = 0;
    if (N == 0) return KSB_OK;
    float cstar = ksb_threshold_float_pct(cut_pct);
    if (cstar != cstar) cstar = INFINITY;  // keep NaN rows only: c < inf drops every number... except inf itself
    uint32_t* d_parent = nullptr;
    unsigned long long* d_edges = nullptr;
    cudaStream_t st = g_ctx[dev].stream;
    cudaEvent_t e0 = g_ctx[dev].e0, e1 = g_ctx[dev].e1;
    int rc = KSB_OK;
    cudaError_t e = cudaSuccess;
    do {
        if (d_labels_out) d_parent = d_labels_out;
        else if ((e = cudaMallocAsync(&d_parent, 4ull * N, st)) != cudaSuccess) break;
        if ((e = cudaMallocAsync(&d_edges, 8, st)) != cudaSuccess) break;
        const LaunchCfg lc{sms, st};
        cudaEventRecord(e0, st);
        if ((e = cudaMemsetAsync(d_edges, 0, 8, st)) != cudaSuccess) break;
        if ((e = launch_k7_init(d_parent, N, lc)) != cudaSuccess) break;
        const bool hook = !(std::getenv("KSPIDER_HOOK_MIN") && std::atoi(std::getenv("KSPIDER_HOOK_MIN")) == 0);
        if (hook) {
            for (const ksb_segment& s : r->seg)
                if ((e = launch_k6_hook_min(s.d_records, s.d_pairs, dist_col == 5 ? reinterpret_cast<const float*>(s.d_pairs + s.n) : nullptr, s.n, dist_col, cstar, d_parent, lc)) != cudaSuccess) break;
            if (e != cudaSuccess) break;
        }
        for (const ksb_segment& s : r->seg)
            if ((e = launch_k67_union(s.d_records, s.d_pairs, dist_col == 5 ? reinterpret_cast<const float*>(s.d_pairs + s.n) : nullptr, s.n, dist_col, cstar, d_parent, d_edges, lc)) != cudaSuccess) break;
        if (e != cudaSuccess) break;
        if ((e = launch_k7_compress(d_parent, N, lc)) != cudaSuccess) break;
        cudaEventRecord(e1, st);
        unsigned long long ne = 0;
        if (labels_out && (e = cudaMemcpyAsync(labels_out, d_parent, 4ull * N, cudaMemcpyDeviceToHost, st)) != cudaSuccess) break;
        if ((e = cudaMemcpyAsync(&ne, d_edges, 8, cudaMemcpyDeviceToHost, st)) != cudaSuccess) break;
        if ((e = cudaStreamSynchronize(st)) != cudaSuccess) break;
        if (n_edges) *n_edges = ne;
        ksb_result* rw = const_cast<ksb_result*>(r);
        rw->info.t.cluster_ms = ev_ms(e0, e1);
        rw->info.gpu_launches += 2 + (hook ? 2 : 1) * r->seg.size();
    } while (0);
    if (e != cudaSuccess) { set_error("ksb_cluster: %s", cudaGetErrorString(e)); (void)cudaGetLastError(); rc = KSB_ERR_CUDA; }
    if (d_parent && d_parent != d_labels_out) cudaFreeAsync(d_parent, st);
    if (d_edges) cudaFreeAsync(d_edges, st);
    return rc;
}

// ANI column (`kSpider pairwise --estimate-ani`, ks_pairwise.py:29-82): avg_ani of every record, in record
// order, computed on the result's device from the device-resident records (K8) and copied to ani_out.
// n_invalid = rows for which the reference would raise (containment text > 1, inf or nan).
int ksb_ani_column(const ksb_result* r, int ksize, double* ani_out, uint64_t* n_invalid) {
    if (!r || (!ani_out && r->info.n_rows)) { set_error("ksb_ani_column: null argument"); return KSB_ERR_ARG; }
    if (r->host_only) { set_error("ksb_ani_column: this result was made with want_records = 2 (host records only)"); return KSB_ERR_ARG; }
    if (ksize < 1) { set_error("ksb_ani_column: ksize must be positive"); return KSB_ERR_ARG; }
    if (n_invalid) *n_invalid = 0;
    uint64_t total = 0;
    for (const ksb_segment& sg : r->seg) total += sg.n;
    if (total == 0) return KSB_OK;
    int sms = 0;
    const int dev = check_device(r->device, &sms);
    if (dev < 0) return dev;
    cudaStream_t st = g_ctx[dev].stream;
    double* d_out = nullptr;
    unsigned long long* d_bad = nullptr;
    cudaError_t e = cudaSuccess;
    int rc = KSB_OK;
    do {
        if ((e = cudaMallocAsync(&d_out, 8 * total, st)) != cudaSuccess) break;
        if ((e = cudaMallocAsync(&d_bad, 8, st)) != cudaSuccess) break;
        if ((e = cudaMemsetAsync(d_bad, 0, 8, st)) != cudaSuccess) break;
        const LaunchCfg lc{sms, st};
        uint64_t at = 0;
        for (const ksb_segment& sg : r->seg) {
            if (!sg.d_records && sg.n) { e = cudaErrorInvalidValue; break; }
            if ((e = launch_k8_ani(sg.d_records, sg.n, 1.0 / (double)ksize, d_out + at, d_bad, lc)) != cudaSuccess) break;
            at += sg.n;
        }
        if (e != cudaSuccess) break;
        unsigned long long nb = 0;
        if ((e = cudaMemcpyAsync(ani_out, d_out, 8 * total, cudaMemcpyDeviceToHost, st)) != cudaSuccess) break;
        if ((e = cudaMemcpyAsync(&nb, d_bad, 8, cudaMemcpyDeviceToHost, st)) != cudaSuccess) break;
        if ((e = cudaStreamSynchronize(st)) != cudaSuccess) break;
        if (n_invalid) *n_invalid = nb;
        const_cast<ksb_result*>(r)->info.gpu_launches += r->seg.size();
    } while (0);
    if (e != cudaSuccess) { set_error("ksb_ani_column: %s", cudaGetErrorString(e)); (void)cudaGetLastError(); rc = KSB_ERR_CUDA; }
    if (d_out) cudaFreeAsync(d_out, st);
    if (d_bad) cudaFreeAsync(d_bad, st);
    return rc;
}

}  // extern "C"

// ------------------------------------------------------- multi-GPU label merge + red micro-benchmark

namespace ksb {
// union(v, other[v]) for every v: merges the partition `other` into `parent`
cudaError_t launch_k7_merge(uint32_t* parent, const uint32_t* other, uint32_t n, const LaunchCfg& lc);
cudaError_t launch_k7_merge_all(uint32_t* parent, const uint32_t* others, uint32_t n, uint32_t n_other, const LaunchCfg& lc);
}  // namespace ksb

extern "C" {

// Device-side merge for callers that already hold the label arrays on the GPU (e.g. after an NCCL
// all-gather): d_labels_all is [world][n]; the merged labels are written to d_out[n].
int ksb_cluster_merge_device(uint64_t n, int world, const uint32_t* d_labels_all, uint32_t* d_out, void* stream_v) {
    if (!d_labels_all || !d_out || world < 1) { set_error("ksb_cluster_merge_device: bad argument"); return KSB_ERR_ARG; }
    if (n == 0) return KSB_OK;
    int dev = 0, sms = 148;
    KSB_CUDA(cudaGetDevice(&dev));
    KSB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    cudaStream_t st = static_cast<cudaStream_t>(stream_v);
    const LaunchCfg lc{sms, st};
    KSB_CUDA(cudaMemcpyAsync(d_out, d_labels_all, 4 * n, cudaMemcpyDeviceToDevice, st));
    KSB_CUDA(launch_k7_merge_all(d_out, d_labels_all + n, (uint32_t)n, (uint32_t)(world - 1), lc));  // every other rank's labels, one launch
    KSB_CUDA(launch_k7_compress(d_out, (uint32_t)n, lc));
    return KSB_OK;
}

int ksb_cluster_merge(uint64_t n, const uint32_t* labels_a, const uint32_t* labels_b, uint32_t* labels_out) {
    if ((n && (!labels_a || !labels_b || !labels_out))) { set_error("ksb_cluster_merge: null argument"); return KSB_ERR_ARG; }
    if (n == 0) return KSB_OK;
    const int dev = check_device(-1, nullptr);
    if (dev < 0) return dev;
    for (uint64_t i = 0; i < n; i++)
        if (labels_a[i] >= n || labels_b[i] >= n) { set_error("ksb_cluster_merge: label out of range at %llu", (unsigned long long)i); return KSB_ERR_ARG; }
    uint32_t* d_all = nullptr;
    uint32_t* d_out = nullptr;
    KSB_CUDA(cudaMalloc(&d_all, 8 * n));
    cudaError_t e = cudaMalloc(&d_out, 4 * n);
    int rc = KSB_OK;
    if (e == cudaSuccess) e = cudaMemcpy(d_all, labels_a, 4 * n, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_all + n, labels_b, 4 * n, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) rc = ksb_cluster_merge_device(n, 2, d_all, d_out, nullptr);
    if (e == cudaSuccess && rc == KSB_OK) e = cudaMemcpy(labels_out, d_out, 4 * n, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) { set_error("ksb_cluster_merge: %s", cudaGetErrorString(e)); (void)cudaGetLastError(); rc = KSB_ERR_CUDA; }
    cudaFree(d_all);
    if (d_out) cudaFree(d_out);
    return rc;
}

int ksb_bench_imma(uint32_t iters, int reps, float* best_ms, double* macs_per_launch) {
    if (!best_ms || !macs_per_launch || iters == 0 || reps < 1) { set_error("ksb_bench_imma: bad argument"); return KSB_ERR_ARG; }
    int sms = 0;
    const int dev = check_device(-1, &sms);
    if (dev < 0) return dev;
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    const LaunchCfg lc{sms, nullptr};
    float best = 1e30f;
    cudaError_t e = cudaSuccess;
    for (int i = 0; i < reps + 1 && e == cudaSuccess; i++) {  // first launch = warm-up
        cudaEventRecord(a, nullptr);
        e = launch_imma_peak(iters, nullptr, macs_per_launch, lc);
        cudaEventRecord(b, nullptr);
        if (e == cudaSuccess) e = cudaEventSynchronize(b);
        if (i > 0 && e == cudaSuccess) best = std::min(best, ev_ms(a, b));
    }
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    if (e != cudaSuccess) { set_error("ksb_bench_imma: %s", cudaGetErrorString(e)); (void)cudaGetLastError(); return KSB_ERR_CUDA; }
    *best_ms = best;
    return KSB_OK;
}

int ksb_bench_red(uint64_t cells, uint64_t n_ops, int cell_bits, int pattern, int reps, float* best_ms) {
    if (!best_ms || cells < 64 || (cell_bits != 32 && cell_bits != 64) || reps < 1) { set_error("ksb_bench_red: bad argument"); return KSB_ERR_ARG; }
    int sms = 0;
    const int dev = check_device(-1, &sms);
    if (dev < 0) return dev;
    void* buf = nullptr;
    KSB_CUDA(cudaMalloc(&buf, cells * (cell_bits / 8)));
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    cudaMemset(buf, 0, cells * (cell_bits / 8));
    const LaunchCfg lc{sms, nullptr};
    float best = 1e30f;
    cudaError_t e = cudaSuccess;
    for (int i = 0; i < reps + 1 && e == cudaSuccess; i++) {  // first launch = warm-up
        cudaEventRecord(a, nullptr);
        e = launch_bench_red(buf, cells, n_ops, cell_bits, pattern, lc);
        cudaEventRecord(b, nullptr);
        if (e == cudaSuccess) e = cudaEventSynchronize(b);
        if (i > 0 && e == cudaSuccess) best = std::min(best, ev_ms(a, b));
    }
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    cudaFree(buf);
    if (e != cudaSuccess) { set_error("ksb_bench_red: %s", cudaGetErrorString(e)); (void)cudaGetLastError(); return KSB_ERR_CUDA; }
    *best_ms = best;
    return KSB_OK;
}

}  // extern "C"
