// ksb_index_load_colored: a kProcessor `colored_kDataFrame` file set as the input of the pairwise pass -- the
// "kProcessor color table" of the indexers that do not write the three .bin files (index_kmers / index_skipmers /
// index_protein: src/index.cpp:53-80 -> kProcessor::index -> colored_kDataFrame::save).  SURVEY 8 row f4.
//
// Files (lib/kProcessor/src/kDataFrames/colored_kDataFrame.cpp:101-113 writes them for prefix P):
//   P.phmap            kDataFramePHMAP::save (kDataFramePHMAP.cpp:208-224): phmap dump of
//                      parallel_flat_hash_map<u64 kmer, u64 colour, ..., N = 6> (kDataFrame.hpp:14,23-26) =
//                      u64 submap count (64), then per submap one raw_hash_set dump with 16-byte slots
//                      (phmap_dump.h:52-63,93-108)
//   Pcolors.intvectors intVectorsTable::save (colorTable.cpp:256-282): text, first the number of colours, then one
//                      line per colour: `colourID size sample...` -- note: NO separator between P and the name
//   P.namesMap         `count` then `sampleID name` lines (optional here: only its ids are used)
// What the pairwise pass needs from them (src/pairwise.cpp:95-121,166-170 read it from the .bin files):
//   members of a colour   = its line of colors.intvectors
//   count of a colour     = number of k-mers of P.phmap whose value is that colour  -> DEVICE HISTOGRAM below
//   total of a sample     = sum of the counts of the colours that hold it (every k-mer has exactly one colour)
#include <algorithm>
#include <atomic>
#include <cerrno>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "kernels.cuh"
#include "ksb_common.hpp"
#include "phmap_format.hpp"

namespace ksb {

namespace {

// One thread per slot of a submap: a FULL slot (ctrl >= 0) looks its colour up in the ascending id list of the
// colour table and adds 1 to that colour's counter (red.global.add.u64); colours the table does not list are
// counted as orphans.
__global__ void __launch_bounds__(256) kc_colour_histogram(const int8_t* __restrict__ ctrl, const ulonglong2* __restrict__ slots,
                                                           unsigned long long capacity, const unsigned long long* __restrict__ ids,
                                                           unsigned long long n_ids, unsigned long long* counts,
                                                           unsigned long long* orphans, unsigned long long* full) {
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    unsigned int n_full = 0, n_orph = 0;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < capacity; i += stride) {
        if (ctrl[i] < 0) continue;
        n_full++;
        const unsigned long long colour = slots[i].y;
        unsigned long long lo = 0, hi = n_ids;  // first id >= colour
        while (lo < hi) {
            const unsigned long long mid = (lo + hi) >> 1;
            if (__ldg(ids + mid) < colour) lo = mid + 1; else hi = mid;
        }
        if (lo < n_ids && __ldg(ids + lo) == colour) asm volatile("red.global.add.u64 [%0], %1;" ::"l"(counts + lo), "l"(1ull) : "memory");
        else n_orph++;
    }
    n_full = __reduce_add_sync(0xffffffffu, n_full);
    n_orph = __reduce_add_sync(0xffffffffu, n_orph);
    if ((threadIdx.x & 31u) == 0) {
        if (n_full) atomicAdd(full, (unsigned long long)n_full);
        if (n_orph) atomicAdd(orphans, (unsigned long long)n_orph);
    }
}

struct Line { uint64_t pos, end; };

inline bool next_u64(const char*& p, const char* e, uint64_t* out) {
    while (p < e && (*p == ' ' || *p == '\t' || *p == '\r')) p++;
    if (p >= e || *p < '0' || *p > '9') return false;
    uint64_t v = 0;
    while (p < e && *p >= '0' && *p <= '9') { v = v * 10 + (uint64_t)(*p - '0'); p++; }
    *out = v;
    return true;
}

}  // namespace

int from_csr_owned(uint64_t n_colors, const uint64_t* offsets, const uint32_t* members, const uint32_t* weights,
                   uint64_t n_totals, const uint32_t* total_ids, const uint32_t* total_vals, int threads, ksb_index** out);

}  // namespace ksb

using namespace ksb;

extern "C" int ksb_index_load_colored(const char* index_prefix, int host_threads, int device, ksb_index** out) {
    if (!index_prefix || !out) { set_error("ksb_index_load_colored: null argument"); return KSB_ERR_ARG; }
    *out = nullptr;
    const int T = std::max(1, host_threads);
    const std::string prefix(index_prefix);
    try {
        // ---- colour table: colors.intvectors ------------------------------------------------------------
        MappedFile cf;
        if (!cf.open(prefix + "colors.intvectors")) return KSB_ERR_IO;
        const char* cp = reinterpret_cast<const char*>(cf.data);
        const char* ce = cp + cf.size;
        const char* p = cp;
        uint64_t n_decl = 0;
        if (!next_u64(p, ce, &n_decl)) { set_error("%scolors.intvectors: no colour count on the first line", index_prefix); return KSB_ERR_FORMAT; }
        while (p < ce && *p != '\n') p++;
        if (p < ce) p++;
        // line index, then two parallel passes (sizes, fill)
        std::vector<Line> lines;
        lines.reserve(n_decl);
        for (const char* q = p; q < ce;) {
            const void* nl = memchr(q, '\n', (size_t)(ce - q));
            const char* le = nl ? static_cast<const char*>(nl) : ce;
            const char* t = q;
            while (t < le && (*t == ' ' || *t == '\r' || *t == '\t')) t++;
            if (t < le) lines.push_back(Line{(uint64_t)(q - cp), (uint64_t)(le - cp)});
            q = nl ? le + 1 : ce;
        }
        if (lines.size() < n_decl) {  // the reference's `input >>` loop would read garbage past the end
            set_error("%scolors.intvectors: %llu colours declared, %zu lines present", index_prefix, (unsigned long long)n_decl, lines.size());
            return KSB_ERR_FORMAT;
        }
        const uint64_t C = n_decl;  // lines beyond the declared count are never read by the reference (colorTable.cpp:231)
        std::vector<uint64_t> ids(C), off(C + 1, 0);
        std::atomic<uint64_t> bad{UINT64_MAX};
        parallel_chunks(C, T, [&](int, uint64_t b, uint64_t e) {
            for (uint64_t c = b; c < e; c++) {
                const char* q = cp + lines[c].pos;
                const char* le = cp + lines[c].end;
                uint64_t s = 0;
                if (!next_u64(q, le, &ids[c]) || !next_u64(q, le, &s)) { bad = c; continue; }
                off[c + 1] = s;
            }
        });
        if (bad != UINT64_MAX) { set_error("%scolors.intvectors: line %llu is not `colour size samples...`", index_prefix, (unsigned long long)(bad.load() + 2)); return KSB_ERR_FORMAT; }
        for (uint64_t c = 0; c < C; c++) off[c + 1] += off[c];
        std::vector<uint32_t> members(off[C]);
        parallel_chunks(C, T, [&](int, uint64_t b, uint64_t e) {
            for (uint64_t c = b; c < e; c++) {
                const char* q = cp + lines[c].pos;
                const char* le = cp + lines[c].end;
                uint64_t v = 0;
                next_u64(q, le, &v);
                next_u64(q, le, &v);
                for (uint64_t k = off[c]; k < off[c + 1]; k++) {
                    if (!next_u64(q, le, &v) || v > 0xFFFFFFFFull) { bad = c; break; }
                    members[k] = (uint32_t)v;
                }
                if (next_u64(q, le, &v)) bad = c;  // more samples than the declared size
            }
        });
        if (bad != UINT64_MAX) { set_error("%scolors.intvectors: line %llu does not hold exactly `size` sample ids", index_prefix, (unsigned long long)(bad.load() + 2)); return KSB_ERR_FORMAT; }
        // a colour id listed twice: `colors[color] = ...` keeps the LAST one (colorTable.cpp:237)
        std::vector<uint64_t> order(C);
        for (uint64_t c = 0; c < C; c++) order[c] = c;
        std::stable_sort(order.begin(), order.end(), [&](uint64_t a, uint64_t b) { return ids[a] < ids[b]; });
        std::vector<uint64_t> uniq_ids, uniq_src;
        uniq_ids.reserve(C);
        uniq_src.reserve(C);
        for (uint64_t k = 0; k < C; k++) {
            if (k + 1 < C && ids[order[k + 1]] == ids[order[k]]) continue;  // a later line with the same id follows
            uniq_ids.push_back(ids[order[k]]);
            uniq_src.push_back(order[k]);
        }
        const uint64_t U = uniq_ids.size();

        // ---- k-mer -> colour map: histogram of the values on the device -----------------------------------
        MappedFile kf;
        if (!kf.open(prefix + ".phmap")) return KSB_ERR_IO;
        if (kf.size < 8) { set_error("%s.phmap: too short", index_prefix); return KSB_ERR_FORMAT; }
        uint64_t n_sub = 0;
        std::memcpy(&n_sub, kf.data, 8);
        if (n_sub == 0 || n_sub > 4096) { set_error("%s.phmap: %llu submaps (not a parallel_flat_hash_map dump?)", index_prefix, (unsigned long long)n_sub); return KSB_ERR_FORMAT; }
        std::vector<DumpView> subs(n_sub);
        uint64_t at = 8, n_kmers_decl = 0, max_cap = 0;
        for (uint64_t s = 0; s < n_sub; s++) {
            if (!open_dump(kf.data + at, kf.size - at, 16, &subs[s])) return KSB_ERR_FORMAT;
            at += subs[s].bytes;
            n_kmers_decl += subs[s].size;
            if (subs[s].size) max_cap = std::max(max_cap, subs[s].capacity);
        }
        int sms = 0;
        const int dev = check_device(device, &sms);
        if (dev < 0) return dev;
        cudaStream_t st = service_stream(dev);
        std::vector<unsigned long long> counts(U, 0);
        unsigned long long h_stat[2] = {0, 0};
        {
            unsigned long long *d_ids = nullptr, *d_counts = nullptr, *d_stat = nullptr;
            int8_t* d_ctrl = nullptr;
            ulonglong2* d_slots = nullptr;
            cudaError_t e = cudaSuccess;
            do {
                if ((e = cudaMallocAsync(&d_ids, 8 * std::max<uint64_t>(U, 1), st)) != cudaSuccess) break;
                if ((e = cudaMallocAsync(&d_counts, 8 * std::max<uint64_t>(U, 1), st)) != cudaSuccess) break;
                if ((e = cudaMallocAsync(&d_stat, 16, st)) != cudaSuccess) break;
                if ((e = cudaMallocAsync(&d_ctrl, max_cap + 32, st)) != cudaSuccess) break;
                if ((e = cudaMallocAsync(&d_slots, 16 * std::max<uint64_t>(max_cap, 1), st)) != cudaSuccess) break;
                if (U && (e = cudaMemcpyAsync(d_ids, uniq_ids.data(), 8 * U, cudaMemcpyHostToDevice, st)) != cudaSuccess) break;
                if ((e = cudaMemsetAsync(d_counts, 0, 8 * std::max<uint64_t>(U, 1), st)) != cudaSuccess) break;
                if ((e = cudaMemsetAsync(d_stat, 0, 16, st)) != cudaSuccess) break;
                for (uint64_t s = 0; s < n_sub && e == cudaSuccess; s++) {
                    const DumpView& v = subs[s];
                    if (v.size == 0) continue;
                    if ((e = cudaMemcpyAsync(d_ctrl, v.ctrl, v.capacity, cudaMemcpyHostToDevice, st)) != cudaSuccess) break;
                    if ((e = cudaMemcpyAsync(d_slots, v.slots, 16 * v.capacity, cudaMemcpyHostToDevice, st)) != cudaSuccess) break;
                    const unsigned ctas = (unsigned)std::min<uint64_t>((uint64_t)sms * 8, (v.capacity + 255) / 256);
                    kc_colour_histogram<<<ctas, 256, 0, st>>>(d_ctrl, d_slots, v.capacity, d_ids, U, d_counts, d_stat, d_stat + 1);
                    e = cudaGetLastError();
                }
                if (e != cudaSuccess) break;
                if (U && (e = cudaMemcpyAsync(counts.data(), d_counts, 8 * U, cudaMemcpyDeviceToHost, st)) != cudaSuccess) break;
                if ((e = cudaMemcpyAsync(h_stat, d_stat, 16, cudaMemcpyDeviceToHost, st)) != cudaSuccess) break;
                e = cudaStreamSynchronize(st);
            } while (0);
            for (void* q : {(void*)d_ids, (void*)d_counts, (void*)d_stat, (void*)d_ctrl, (void*)d_slots})
                if (q) cudaFreeAsync(q, st);
            cudaStreamSynchronize(st);
            if (e != cudaSuccess) { set_error("ksb_index_load_colored: %s", cudaGetErrorString(e)); (void)cudaGetLastError(); return KSB_ERR_CUDA; }
        }
        if (h_stat[1] != n_kmers_decl) {
            set_error("%s.phmap: %llu live slots, %llu declared", index_prefix, (unsigned long long)h_stat[1], (unsigned long long)n_kmers_decl);
            return KSB_ERR_FORMAT;
        }
        if (h_stat[0]) {  // colored_kDataFrame::getSamplesIDForKmer would fail on these k-mers
            set_error("%s.phmap: %llu k-mers carry a colour that %scolors.intvectors does not list", index_prefix, (unsigned long long)h_stat[0], index_prefix);
            return KSB_ERR_FORMAT;
        }

        // ---- CSR in ascending colour id order, weights = (uint32_t)count like src/pairwise.cpp:118-120, totals
        std::vector<uint64_t> off2(U + 1, 0);
        for (uint64_t k = 0; k < U; k++) off2[k + 1] = off2[k] + (off[uniq_src[k] + 1] - off[uniq_src[k]]);
        std::vector<uint32_t> mem2(off2[U]), weights(U);
        parallel_chunks(U, T, [&](int, uint64_t b, uint64_t e) {
            for (uint64_t k = b; k < e; k++) {
                std::copy(members.begin() + off[uniq_src[k]], members.begin() + off[uniq_src[k] + 1], mem2.begin() + off2[k]);
                weights[k] = (uint32_t)counts[k];
            }
        });
        uint32_t max_id = 0;
        for (uint32_t m : mem2) max_id = std::max(max_id, m);
        std::vector<uint32_t> name_ids;
        {   // every name is a sample, even one that no k-mer reached
            FILE* nf = fopen((prefix + ".namesMap").c_str(), "r");
            if (nf) {
                unsigned long long n = 0, id = 0;
                char name[4096];
                if (fscanf(nf, "%llu", &n) == 1)
                    for (unsigned long long i = 0; i < n && fscanf(nf, "%llu %4095s", &id, name) == 2; i++)
                        if (id <= 0xFFFFFFFFull) name_ids.push_back((uint32_t)id);
                fclose(nf);
            }
        }
        for (uint32_t v : name_ids) max_id = std::max(max_id, v);
        std::vector<uint32_t> total_ids, total_vals;
        if ((uint64_t)max_id + 1 <= (1ull << 28)) {  // dense ids (kProcessor numbers its groups 1..N)
            std::vector<uint64_t> tot((uint64_t)max_id + 1, 0);
            std::vector<uint8_t> present((uint64_t)max_id + 1, 0);
            for (uint64_t k = 0; k < U; k++)
                for (uint64_t a = off2[k]; a < off2[k + 1]; a++) { tot[mem2[a]] += counts[k]; present[mem2[a]] = 1; }
            for (uint32_t v : name_ids) present[v] = 1;
            for (uint64_t s = 0; s <= max_id; s++)
                if (present[s]) { total_ids.push_back((uint32_t)s); total_vals.push_back((uint32_t)tot[s]); }  // u32 like groupID_to_kmerCount's values
        } else {
            total_ids = mem2;
            total_ids.insert(total_ids.end(), name_ids.begin(), name_ids.end());
            std::sort(total_ids.begin(), total_ids.end());
            total_ids.erase(std::unique(total_ids.begin(), total_ids.end()), total_ids.end());
            std::vector<uint64_t> tot(total_ids.size(), 0);
            for (uint64_t k = 0; k < U; k++)
                for (uint64_t a = off2[k]; a < off2[k + 1]; a++)
                    tot[std::lower_bound(total_ids.begin(), total_ids.end(), mem2[a]) - total_ids.begin()] += counts[k];
            total_vals.resize(total_ids.size());
            for (size_t i = 0; i < tot.size(); i++) total_vals[i] = (uint32_t)tot[i];
        }
        return from_csr_owned(U, off2.data(), mem2.data(), weights.data(), total_ids.size(), total_ids.data(), total_vals.data(), T, out);
    } catch (const std::bad_alloc&) {
        set_error("out of host memory");
        return KSB_ERR_NOMEM;
    }
}
