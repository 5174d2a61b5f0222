// Drop-in for the reference's public C++ entry point of the hot path
// (include/kSpider.hpp:11, defined src/pairwise.cpp:123): same namespace, name, by-value
// signature and mangled symbol, so pairwise.cpp:3-5 and a SWIG/pybind `_kSpider_internal`
// link against libkspider_b200.so unchanged.
#pragma once
#include <string>

namespace kSpider {
// Reads  <prefix>_color_to_sources.bin, <prefix>_color_count.bin, <prefix>_groupID_to_kmerCount.bin
// Writes <prefix>_kSpider_seqToKmersNo.tsv, <prefix>_kSpider_pairwise.tsv
// user_threads = host loader / formatter threads (the reference's OpenMP thread count).
// Environment: KSPIDER_GPUS (devices to shard the triangle over, default 1), KSPIDER_DEVICE
// (first ordinal, default 0), KSPIDER_DENSE_BUDGET_GB (dense triangle budget per GPU).
// Throws std::runtime_error on any failure (the reference asserts / crashes).
void pairwise(std::string index_prefix, int user_threads);
}  // namespace kSpider
