#include "kspider_dropin.hpp"

#include <chrono>
#include <cstdlib>
#include <iostream>
#include <stdexcept>
#include <thread>
#include <vector>

#include "ksb_common.hpp"

namespace {
using Clock = std::chrono::high_resolution_clock;
double secs_since(Clock::time_point t0) { return std::chrono::duration<double, std::milli>(Clock::now() - t0).count() / 1000; }
[[noreturn]] void fail(const char* what) { throw std::runtime_error(std::string("kSpider::pairwise: ") + what + ": " + ksb_last_error()); }
long env_long(const char* name, long dflt) {
    const char* v = std::getenv(name);
    return (v && *v) ? std::strtol(v, nullptr, 10) : dflt;
}
}  // namespace

namespace kSpider {

void pairwise(std::string index_prefix, int user_threads) {
    const int threads = user_threads > 0 ? user_threads : 1;
    const int gpus = (int)std::max(1L, env_long("KSPIDER_GPUS", 1));
    const int dev0 = (int)env_long("KSPIDER_DEVICE", 0);
    const double budget_gb = std::getenv("KSPIDER_DENSE_BUDGET_GB") ? std::atof(std::getenv("KSPIDER_DENSE_BUDGET_GB")) : 0.0;

    // KSPIDER_VERBOSE=1: one extra line on STDERR with the phases of this call (the five stdout lines stay the reference's)
    const bool verbose = env_long("KSPIDER_VERBOSE", 0) != 0;
    double t_load = 0, t_seq = 0, t_gpu = 0, t_tsv = 0, t_side = 0;
    // the reference prints three load timers (src/pairwise.cpp:133,155,181); the loader does the
    // three loads in one call, so the first line carries the time and the other two are ~0
    auto t0 = Clock::now();
    // the CUDA contexts come up (~0.25 s each) while the index files are parsed; an error here is reported by the run itself
    std::vector<std::thread> warm;
    struct Joiner {  // whichever way this function is left
        std::vector<std::thread>& v;
        void join() { for (auto& t : v) if (t.joinable()) t.join(); }
        ~Joiner() { join(); }
    } warm_join{warm};
    for (int g = 0; g < gpus; g++) warm.emplace_back([=] { (void)ksb::warm_device(dev0 + g); });
    ksb_index* ix = nullptr;
    const int lrc = ksb_index_load(index_prefix.c_str(), threads, &ix);
    if (lrc != KSB_OK) {
        const std::string lerr = ksb_last_error();
        warm_join.join();  // (the warm-up may leave its own message behind)
        ksb::set_error("%s", lerr.c_str());
        fail("loading the index");
    }
    t_load = secs_since(t0);
    std::cout << "mapping colors to groups: " << t_load << " secs" << std::endl;
    t0 = Clock::now();
    std::cout << "parsing index colors: " << secs_since(t0) << " secs" << std::endl;
    t0 = Clock::now();
    if (ksb_write_seq_to_kmers(ix, (index_prefix + "_kSpider_seqToKmersNo.tsv").c_str()) != KSB_OK) { ksb_index_free(ix); fail("writing seqToKmersNo.tsv"); }
    t_seq = secs_since(t0);
    std::cout << "kmer counting: " << t_seq << " secs" << std::endl;

    t0 = Clock::now();
    warm_join.join();  // a context that takes longer than the parsing (the first CUDA process on a box) counts as GPU time
    std::vector<ksb_result*> res(gpus, nullptr);
    std::vector<int> rcs(gpus, KSB_OK);
    std::vector<std::string> errs(gpus);
    auto work = [&](int g) {
        ksb_opts o{};
        o.device = dev0 + g;
        o.want_records = 2;  // rows go to the TSV: no device copy is kept once a pass is on the host
        o.dense_budget_bytes = (uint64_t)(budget_gb * 1e9);
        if (gpus == 1) {
            rcs[g] = ksb_pairwise(ix, &o, &res[g]);
        } else {
            // every GPU gets the row-block shard of its own rows (only its colours, cut down and
            // re-binned) over its own PCIe link; owned cells are disjoint, nothing is reduced
            ksb_index* sh = nullptr;
            rcs[g] = ksb_index_shard(ix, g, gpus, std::max(1, threads / gpus), &sh);
            if (rcs[g] == KSB_OK) rcs[g] = ksb_pairwise(sh, &o, &res[g]);
            ksb_index_free(sh);
        }
        if (rcs[g] != KSB_OK) errs[g] = ksb_last_error();
    };
    if (gpus == 1) work(0);
    else {
        std::vector<std::thread> pool;
        for (int g = 0; g < gpus; g++) pool.emplace_back(work, g);
        for (auto& t : pool) t.join();
    }
    for (int g = 0; g < gpus; g++)
        if (rcs[g] != KSB_OK) {
            const std::string msg = "kSpider::pairwise: GPU " + std::to_string(dev0 + g) + ": " + errs[g];
            for (auto* r : res) ksb_result_free(r);
            ksb_index_free(ix);
            throw std::runtime_error(msg);
        }
    t_gpu = secs_since(t0);
    std::cout << "pairwise hashmap construction: " << t_gpu << " secs" << std::endl;
    const std::string out = index_prefix + "_kSpider_pairwise.tsv";
    std::cout << "writing pairwise matrix to " << out << std::endl;
    int rc = KSB_OK;
    t0 = Clock::now();
    // binary side-car for `cluster` / `--estimate-ani` (they then skip the text): the same rows, 32 bytes each, written by a
    // thread of its own while the TSV is formatted, and stamped with the TSV's size and time once both are complete
    const char* sc = std::getenv("KSPIDER_SIDECAR");
    const bool want_side = !(sc && sc[0] == '0');
    const std::string side = index_prefix + "_kSpider_pairwise.ksb";
    int side_rc = KSB_OK;
    std::string side_err;
    double side_secs = 0;
    std::thread side_writer;
    if (want_side)
        side_writer = std::thread([&] {
            const auto ts = Clock::now();
            for (int g = 0; g < gpus && side_rc == KSB_OK; g++) {
                uint64_t n = 0;
                const ksb_record* rec = ksb_result_records(res[g], &n);
                side_rc = ksb_write_sidecar_records(rec, n, side.c_str(), 0, 0, g != 0);
            }
            if (side_rc != KSB_OK) side_err = ksb_last_error();
            side_secs = secs_since(ts);
        });
    for (int g = 0; g < gpus && rc == KSB_OK; g++)
        rc = g == 0 ? ksb_write_tsv(res[g], out.c_str(), threads) : ksb_append_tsv(res[g], out.c_str(), threads);
    t_tsv = secs_since(t0);
    std::string tsv_err = rc == KSB_OK ? "" : ksb_last_error();
    t0 = Clock::now();
    if (want_side) {
        side_writer.join();
        if (rc == KSB_OK && side_rc != KSB_OK) { rc = side_rc; tsv_err = side_err; }
        if (rc == KSB_OK) { rc = ksb_sidecar_stamp(side.c_str(), out.c_str()); if (rc != KSB_OK) tsv_err = ksb_last_error(); }
    }
    t_side = secs_since(t0);  // what the side-car adds to the wall time (its writing overlaps the TSV: side_secs)
    (void)side_secs;
    if (verbose) std::cerr << "ksb_phases load_s=" << t_load << " seq_to_kmers_s=" << t_seq << " gpu_s=" << t_gpu << " tsv_s=" << t_tsv
                           << " sidecar_s=" << t_side << std::endl;
    const std::string err = tsv_err;
    for (auto* r : res) ksb_result_free(r);
    ksb_index_free(ix);
    if (rc != KSB_OK) throw std::runtime_error("kSpider::pairwise: writing " + out + ": " + err);
}

}  // namespace kSpider

// C-ABI door to the drop-in for hosts that cannot pass a std::string by value (ctypes, cgo ...):
// what a SWIG/pybind `_kSpider_internal.pairwise` wrapper would call.  Returns KSB_OK or a negative
// code with the exception text in ksb_last_error().
extern "C" int ksb_dropin_pairwise(const char* index_prefix, int user_threads) {
    if (!index_prefix) { ksb::set_error("ksb_dropin_pairwise: null prefix"); return KSB_ERR_ARG; }
    try {
        kSpider::pairwise(std::string(index_prefix), user_threads);
    } catch (const std::exception& e) {
        const std::string msg = e.what();
        ksb::set_error("%s", msg.c_str());
        return msg.find("CUDA") != std::string::npos || msg.find("GPU") != std::string::npos ? KSB_ERR_CUDA : KSB_ERR_IO;
    }
    return KSB_OK;
}
