// Parsed rows of a pairwise TSV (or of its binary side-car): what the file-driven consumers work on.
#pragma once
#include <cstddef>
#include <cstdint>

struct ksb_table {
    uint64_t n = 0;
    uint32_t* s1 = nullptr;   // source_1 as written (ORIGINAL sample ids)
    uint32_t* s2 = nullptr;
    double* dist = nullptr;   // TSV tables: float(text) of column dist_col (or of the ani_col file)
    double* dist2 = nullptr;  // TSV tables: second column (ANI needs min and max)
    float* cont = nullptr;    // side-car tables: the float32 containment whose "%g" text the TSV shows
    float* cont2 = nullptr;
    int dist_col = 0, dist_col2 = 0;
    bool pinned = false;
};

namespace ksb {
bool table_try_pinned(void** out, size_t bytes);  // cudaMallocHost when a driver is there (table.cu)
void table_free_pinned(void* p);
void table_release(ksb_table* t);
}  // namespace ksb
