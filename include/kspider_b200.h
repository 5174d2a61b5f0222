/*
 * kspider_b200.h — C-ABI of libkspider_b200.so: the B200-native replacement of kSpider's pairwise
 * hot path (dib-lab/kSpider @ 466a267).
 *
 * Everything the reference does between "read the three index files" and "write the pairwise TSV /
 * cluster file" sits behind these entry points.  Plain pointers and sizes only; no exceptions cross
 * the boundary; every function returns 0 (KSB_OK) or a negative error code and leaves a message
 * retrievable with ksb_last_error() (thread-local).  The library owns every buffer it hands out
 * until the matching ksb_*_free.
 *
 * Reference interface each entry point replaces (file:line in /root/reference):
 *   ksb_index_load            load_colors_to_sources + load_colors_count + totals load
 *                             (src/pairwise.cpp:95-121,125-137,166-170)
 *   ksb_index_from_csr        the state after src/pairwise.cpp:194 (vec_color_to_ids, colorsCount,
 *                             groupID_to_kmerCount) handed over directly instead of through files
 *   ksb_write_seq_to_kmers    src/pairwise.cpp:173-180
 *   ksb_pairwise              the OpenMP pair loop + PAIRS_COUNTER accumulation
 *                             (src/pairwise.cpp:191-237, Combo :39-70) and the per-edge containment
 *                             math of the writer loop (src/pairwise.cpp:253-264)
 *   ksb_upload/ksb_accumulate/ksb_postpass   the same, split so the accumulation can be timed with
 *                             inputs already resident in HBM
 *   ksb_write_tsv             src/pairwise.cpp:242-275 (header, column order, ostream<<float text)
 *   ksb_cluster               pykSpider/kSpider2/ks_clustering.py:63-120 (threshold on the printed
 *                             containment text, graph build, rx.connected_components)
 *   kSpider::pairwise(std::string,int)  (C++ symbol exported by the same library; declared in
 *                             kspider_b200/csrc/kspider_dropin.hpp) include/kSpider.hpp:11
 *
 * There is NO CPU fallback: every compute entry point fails with KSB_ERR_CUDA when no sm_100
 * device is usable.
 */
#ifndef KSPIDER_B200_H
#define KSPIDER_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KSB_OK 0
#define KSB_ERR_IO (-1)       /* file missing / unreadable / short */
#define KSB_ERR_FORMAT (-2)   /* not a phmap dump of the expected shape, or reference assert would fire */
#define KSB_ERR_NOMEM (-3)
#define KSB_ERR_CUDA (-4)     /* no device, launch or runtime failure */
#define KSB_ERR_ARG (-5)
#define KSB_ERR_BUDGET (-6)   /* owned triangle block does not fit the dense budget and spill is disabled */

#define KSB_ABI_VERSION 1

typedef struct ksb_index ksb_index;    /* host side: parsed, sample-compacted, size-binned CSR (pinned) */
typedef struct ksb_device ksb_device;  /* one GPU: resident CSR + owned triangle block */
typedef struct ksb_result ksb_result;  /* records / edges / labels of one run */

/* Shape of a loaded index.  U = sum over multi-member colours of s(s-1)/2 = weighted pair-updates. */
typedef struct ksb_index_info {
    uint64_t n_samples;        /* N: distinct sample ids in totals U members */
    uint64_t n_colors_file;    /* records in _color_to_sources.bin after u32-key de-duplication */
    uint64_t n_colors;         /* C: colours with >= 2 members */
    uint64_t n_members;        /* M: members of those colours */
    uint64_t n_updates;        /* U */
    uint64_t n_small_colors;   /* 2 <= s <= 64  (flat pair-space kernel) */
    uint64_t n_mid_colors;     /* s > 64        (shared-memory tiled kernel) */
    uint64_t n_gram_colors;    /* colours routed to the int8 tcgen05 Gram kernel (K4); 0 = path unused */
    uint64_t gram_min_size;    /* smallest colour routed to K4 (0 = none); the rule is per Gram window, see DESIGN.md
                                  (KSPIDER_GRAM=off|auto|<min size>) */
    uint64_t gram_planes;      /* weight byte planes the Gram kernel runs (1, 2 or 4) */
    uint64_t n_gram_updates;   /* pair-updates covered by the Gram colours */
    uint64_t n_zero_colors;    /* multi-member colours whose narrowed count is 0 (touched-with-zero rows) */
    uint64_t max_color_size;
    uint64_t payload_bytes;    /* size of the binned CSR blob (what H2D / NCCL broadcast moves) */
    uint64_t triangle_cells;   /* N(N-1)/2 */
    uint64_t n_small_members;  /* members / pair-updates of the K1 colours, members of the K4 colours */
    uint64_t n_small_updates;
    uint64_t n_gram_members;
    uint64_t n_gram_windows;   /* sample windows the K4 colours are grouped into */
    uint64_t n_gram_macs;      /* sum over windows of colours x width(width-1)/2 x planes: u8 MACs K4 must evaluate */
    uint32_t pinned;           /* 1 if the payload lives in cudaMallocHost memory */
    uint32_t gram_form;        /* Gram section of the blob: 0 = member lists (CSR; gram_form >> 8 = bits of a window-relative
                                  member, 0 = absolute ids), 1 = the K-block-major bitmaps K4 reads (no bitmap build per step) */
    uint64_t n_gram_cells;     /* sum over windows of extent(extent-1)/2: cells K4's epilogue can update (one red each) */
} ksb_index_info;

typedef struct ksb_opts {
    int32_t device;              /* CUDA ordinal; -1 = current device */
    int32_t rank;                /* this process owns row block `rank` of `world` (balanced by updates) */
    int32_t world;               /* 0 or 1 = whole triangle */
    int32_t want_records;        /* 1: build the {i,j,shared,min,avg,max} records and bring them to the host;
                                    2: the same, HOST ONLY -- the device copies of a pass (44 bytes per row) are released as
                                       soon as its rows are on the host (what the TSV-writing drop-in wants; such a result
                                       cannot be clustered on the device); 0: no host copy */
    uint64_t dense_budget_bytes; /* 0 = library default: 80% of the free HBM for a pass, i.e. its cells AND (when records
                                    are wanted) the rows it can emit.  When a pass still runs out of device memory,
                                    ksb_pairwise repeats the run with a quarter of the budget (more, smaller passes). */
    uint64_t spill_mode;         /* K3 sort-reduce path: 0 = auto (only when the owned block would need more than
                                    16 dense passes under the budget), 1 = always, 2 = never */
    uint64_t reserved[3];
} ksb_opts;

/* One TSV row before formatting: 32 bytes, ids are the ORIGINAL sample ids of the index files. */
typedef struct ksb_record {
    uint32_t source_1;
    uint32_t source_2;
    uint64_t shared_kmers;
    float min_containment;
    float avg_containment;
    float max_containment;
    uint32_t pad;
} ksb_record;

typedef struct ksb_timings {
    float h2d_ms;         /* CSR blob host->device */
    float zero_ms;        /* clearing the owned triangle block */
    float small_ms;       /* K1 flat pair-space kernel (2 <= s <= 64) */
    float mid_ms;         /* K2 tiled kernel (s > 64) */
    float zero_touch_ms;  /* touched-bitmap pass for zero-weight colours */
    float accumulate_ms;  /* zero + small + mid + zero_touch measured as one interval */
    float post_ms;        /* K5 count + compact */
    float d2h_ms;         /* records device->host */
    float cluster_ms;     /* K6 + K7 */
    float gram_ms;        /* K4: bitmap build + int8 tcgen05 Gram kernel (inside accumulate_ms) */
    float spill_emit_ms;  /* K3: pair emission into chunk buffers */
    float spill_sort_ms;  /* K3: radix sorts + reduce-by-key + run merges */
    float host_setup_ms;  /* ksb_pairwise only, host wall clock: device handle + allocations + copy enqueue */
    float host_accumulate_ms; /* the ksb_accumulate call(s), including the wait for the pipelined copy */
    float host_post_ms;   /* the ksb_postpass call(s) */
    float host_free_ms;   /* releasing the device handle */
} ksb_timings;

typedef struct ksb_result_info {
    uint64_t n_rows;          /* R: touched pairs in the owned row block */
    uint64_t n_updates;       /* updates applied by this device (its row block) */
    uint64_t row_begin;       /* owned rows [row_begin,row_end) in compact sample index space */
    uint64_t row_end;
    uint64_t owned_cells;
    uint64_t gpu_launches;    /* kernels launched by the library for this result so far */
    ksb_timings t;
} ksb_result_info;

const char* ksb_last_error(void);
int ksb_abi_version(void);
/* 0 if a usable sm_100 device exists, KSB_ERR_CUDA otherwise (message says why). */
int ksb_cuda_available(void);

/* ---- host: index ------------------------------------------------------------------------- */
int ksb_index_load(const char* index_prefix, int host_threads, ksb_index** out);
/* members need not be sorted; sample ids are arbitrary u32; colours with < 2 members are ignored. */
int ksb_index_from_csr(uint64_t n_colors, const uint64_t* offsets, const uint32_t* members,
                       const uint32_t* weights, uint64_t n_totals, const uint32_t* total_ids,
                       const uint32_t* total_vals, int host_threads, ksb_index** out);
/* The same index from a kProcessor `colored_kDataFrame` file set -- what index_kmers /
 * index_skipmers / index_protein leave (src/index.cpp:53-80 -> colored_kDataFrame::save,
 * lib/kProcessor/src/kDataFrames/colored_kDataFrame.cpp:101-113): <prefix>.phmap (k-mer -> colour,
 * kDataFramePHMAP.cpp:208-224), <prefix>colors.intvectors (colour -> samples, colorTable.cpp:256-282) and
 * <prefix>.namesMap.  Per-colour counts = histogram of the k-mer map's values, computed on `device` (-1 = current);
 * per-sample totals = sum of the counts of the colours holding the sample; counts / totals narrowed to u32 as
 * src/pairwise.cpp:118-120 does.  KSB_ERR_FORMAT when a k-mer carries a colour the table does not list. */
int ksb_index_load_colored(const char* index_prefix, int host_threads, int device, ksb_index** out);
int ksb_index_get_info(const ksb_index* ix, ksb_index_info* info);
/* compact index -> original sample id (ascending), and per-sample totals in the same order */
const uint32_t* ksb_index_sample_ids(const ksb_index* ix);
const uint32_t* ksb_index_totals(const ksb_index* ix);
/* Row bands of the accumulator: row i (compact index) is only ever paired with columns
 * lo[i] <= j < hi[i]; the device stores exactly these cells, row after row (cell (i,j) of the block
 * ksb_device_triangle returns is at sum_{r<i, r owned}(hi[r]-lo[r]) + j - lo[i]).  Without locality in
 * the colours this is the packed upper triangle. */
int ksb_index_bands(const ksb_index* ix, const uint32_t** lo, const uint32_t** hi, uint64_t* band_cells);
/* the binned CSR blob + its fixed-size descriptor: what another rank needs to adopt the index */
const void* ksb_index_payload(const ksb_index* ix, uint64_t* bytes);
const void* ksb_index_descriptor(const ksb_index* ix, uint64_t* bytes);
/* row-block ownership for `world` GPUs: cuts_out[world+1], rank r owns triangle rows
 * [cuts[r], cuts[r+1]) (compact sample indices), balanced by weighted pair-updates */
int ksb_index_row_cuts(const ksb_index* ix, int world, uint64_t* cuts_out);
/* the same from a (descriptor, HOST payload) pair, for ranks that received the blob by broadcast */
int ksb_blob_row_cuts(const void* descriptor, uint64_t descriptor_bytes, const void* payload, uint64_t payload_bytes,
                      int world, uint64_t* cuts_out);
/* Row-block shard of an index for rank `rank` of `world` (same cuts as ksb_index_row_cuts): a new,
 * smaller index holding only the colours that have a member in the rank's rows, each cut down to its
 * members >= the first owned row, re-binned for the kernels.  A device created from it owns exactly
 * those rows (opts.rank / opts.world are ignored) and produces the same block, bit for bit, as the
 * whole index with opts.rank / opts.world -- but its per-step work is proportional to the rank's
 * share of the pair-updates instead of the whole colour table.  This is what each GPU of a
 * multi-GPU run receives. */
int ksb_index_shard(const ksb_index* ix, int rank, int world, int host_threads, ksb_index** out);
/* Host index from a (descriptor, payload) pair produced by ksb_index_descriptor / ksb_index_payload
 * on another process or rank: the blob is copied into pinned memory of THIS process, so that every
 * rank of a multi-GPU run uploads its own shard over its own PCIe link (ksb_pairwise / ksb_upload).
 * The pair is checked first (KSB_ERR_FORMAT): section layout, row tables, and every member list -- strictly ascending below the
 * sample count, every pair inside its row's band, every Gram colour inside its window -- since the kernels address cells with
 * the member values as they come. */
int ksb_index_from_blob(const void* descriptor, uint64_t descriptor_bytes, const void* payload, uint64_t payload_bytes,
                        ksb_index** out);
int ksb_write_seq_to_kmers(const ksb_index* ix, const char* path);
void ksb_index_free(ksb_index* ix);

/* ---- device: resident CSR + accumulation ---------------------------------------------------- */
/* Copies the blob to the device (timed: t.h2d_ms) and allocates the owned triangle block. */
int ksb_upload(const ksb_index* ix, const ksb_opts* opts, ksb_device** out);
/* Adopts a blob that is ALREADY in device memory (e.g. received by an NCCL broadcast).  The
 * descriptor is host memory.  The device buffer must outlive the ksb_device.  Nothing in device memory can be checked from
 * here: the payload is TRUSTED to be what ksb_index_payload of this build produced (ksb_index_from_blob is the checking door). */
int ksb_adopt_device_payload(const void* descriptor, uint64_t descriptor_bytes, const void* d_payload,
                             uint64_t payload_bytes, const ksb_opts* opts, ksb_device** out);
/* Clears the owned block and runs the pair-emission kernels on `stream` (a cudaStream_t passed as
 * void*; NULL = the library's own stream).  Returns after enqueueing unless sync != 0. */
int ksb_accumulate(ksb_device* dev, void* stream, int sync);
/* device pointer of the owned block of the banded upper triangle: cells of ksb_device_cell_bits(dev) bits -- 32 when the
 * index proves that no cell can exceed 2^32 - 1 (a cell is a sum of weights of colours holding its row; the loader
 * keeps the largest per-sample sum of weights) and the block is large enough to stream through HBM (> 16 MB as u64;
 * KSPIDER_CELL_BITS=32 takes them for any size), 64 otherwise or with KSPIDER_CELL_BITS=64.  The values are those of
 * the reference's u64 PAIRS_COUNTER either way (src/pairwise.cpp:222-225). */
void* ksb_device_triangle(const ksb_device* dev, uint64_t* cells, uint64_t* row_begin, uint64_t* row_end);
int ksb_device_cell_bits(const ksb_device* dev);
/* copies the current pass's block to host memory as u64 cells (cells must equal the pass's cell count) */
int ksb_device_read_triangle(const ksb_device* dev, uint64_t* host_out, uint64_t cells);
/* When the owned row block does not fit dense_budget_bytes it is processed in several dense passes
 * (sub-blocks of rows): select one with ksb_device_set_pass before ksb_accumulate / ksb_postpass. */
int ksb_device_n_passes(const ksb_device* dev);
int ksb_device_set_pass(ksb_device* dev, int pass);
/* K3: 1 if this device handle runs the sort-reduce spill path (no dense block is allocated; the
 * pass API above does not apply), and the spill run itself: emission in chunks of at most
 * dense_budget_bytes / 40 updates, sort + reduce per chunk, merge, records in (i,j) order. */
int ksb_device_is_spill(const ksb_device* dev);
/* K4 normally runs on a second stream beside K1/K2 (default, or KSPIDER_OVERLAP=0 to disable at
 * creation).  on = 0 serialises the kernels so that the per-kernel timings of ksb_timings are those of
 * each kernel running alone (bench.py's roofline leg); results are identical either way. */
int ksb_device_set_overlap(ksb_device* dev, int on);
int ksb_device_run_spill(ksb_device* dev, int want_records, ksb_result** out);
/* owned rows / cells / updates, launches and timings so far */
int ksb_device_get_info(const ksb_device* dev, ksb_result_info* info);
/* K5: containment + record compaction of the current pass (rows sorted by (i,j)).  *inout == NULL
 * creates the result, otherwise the rows are appended to it.  Records also land on the host if
 * want_records (the flag of the creating call decides). */
int ksb_postpass(ksb_device* dev, int want_records, ksb_result** inout);
/* Cluster-only step (north_star (3): "only the edge list or cluster labels return to the host"): every dense pass
 * is accumulated and its block scanned ONCE by K56 -- float32 containment of column dist_col (3 = min, 4 = avg,
 * 5 = max) exactly as ksb_postpass computes it (src/pairwise.cpp:253-264), the reference's threshold on the
 * printed text (ks_clustering.py:97-105; cut_pct = float(cutoff) * 100), and ONLY the kept pairs are appended
 * (8 bytes each) to an edge list -- then lock-free union-find over that list.  No records are built, nothing is
 * read twice.  Everything is enqueued on `stream` (a cudaStream_t as void*; NULL = the library's own) and the
 * call synchronises once, at the end.  labels_out (host, N entries) and / or d_labels_out (device, N entries)
 * receive label = smallest compact sample index of the component, restricted to the edges of the rows this
 * device owns (merge row blocks with ksb_cluster_merge_device).  n_rows = touched pairs (the TSV rows the
 * reference would write), n_edges = kept edges.  Same labels as ksb_postpass + ksb_cluster. */
int ksb_device_cluster(ksb_device* dev, int dist_col, double cut_pct, void* stream, uint32_t* labels_out,
                       uint32_t* d_labels_out, uint64_t* n_edges, uint64_t* n_rows);
/* The same step in two halves, for callers that put more work behind it on the same stream before they wait (the
 * multi-GPU step: label all-gather + merge): _enqueue returns without synchronising -- the step is replayed as ONE
 * CUDA graph launch (captured at the first call for a given column / threshold / output; KSPIDER_GRAPH=0 or
 * ksb_device_set_graph(dev, 0) enqueue it kernel by kernel, which also keeps the per-kernel timers alive) --
 * and _collect waits for it and hands out the labels (if host_labels was set) and the counters.
 * _collect returns KSB_ERR_BUDGET when the kept-edge list was too short: it has been grown, enqueue again. */
int ksb_device_cluster_enqueue(ksb_device* dev, int dist_col, double cut_pct, void* stream, uint32_t* d_labels_out,
                               int host_labels);
int ksb_device_cluster_collect(ksb_device* dev, uint32_t* labels_out, uint64_t* n_edges, uint64_t* n_rows);
int ksb_device_set_graph(ksb_device* dev, int on);
void ksb_device_free(ksb_device* dev);

/* upload + accumulate + postpass in one call (the end-to-end path) */
int ksb_pairwise(const ksb_index* ix, const ksb_opts* opts, ksb_result** out);

/* upload + ksb_device_cluster + release in one call: host index in, cluster labels out -- the end-to-end path
 * when only the labels are wanted (H2D: the blob; D2H: 4 N + 16 bytes).  labels_out: host, d_labels_out: device
 * memory of opts->device (either may be NULL; the multi-GPU path merges the device copies).  info_out (optional): rows, updates,
 * launches, timings of the run.  A device that must spill runs the sort-reduce path and clusters its records.
 * The call keeps the device handle it built (device allocations of the blob and the accumulator, row passes, the K4 plan,
 * streams and events -- never a result) inside `ix` and releases it in ksb_index_free: the next call on the same index
 * with the same device / rank / world / budget / spill mode uploads the blob again and runs, without planning or
 * allocating anything.  KSPIDER_E2E_CACHE=0 builds and releases a handle per call. */
int ksb_pairwise_cluster(const ksb_index* ix, const ksb_opts* opts, int dist_col, double cut_pct, uint32_t* labels_out,
                         uint32_t* d_labels_out, uint64_t* n_edges, uint64_t* n_rows, ksb_result_info* info_out);

/* ---- results ---------------------------------------------------------------------------------- */
int ksb_result_get_info(const ksb_result* r, ksb_result_info* info);
const ksb_record* ksb_result_records(const ksb_result* r, uint64_t* n_rows);
int ksb_write_tsv(const ksb_result* r, const char* path, int host_threads);
/* Appends rows of further results (other row blocks) to an existing TSV without a header. */
int ksb_append_tsv(const ksb_result* r, const char* path, int host_threads);
/* The same writer on host rows that are not inside a result (src/pairwise.cpp:242-275: header unless append, then one line per
 * row in the order given): what ksb_write_tsv / ksb_append_tsv call with the result's rows. */
int ksb_write_tsv_records(const ksb_record* records, uint64_t n_rows, const char* path, int host_threads, int append);
/* K6+K7: threshold column dist_col (3=min,4=avg,5=max) at cutoff01 with the reference's text
 * semantics, union-find over the kept edges.  labels_out[N]: label = smallest compact sample index of
 * the component.  parent_inout (optional, N entries, device-independent host buffer) lets several row
 * blocks be merged: pass the labels of the previous block in, get the merged labels out. */
int ksb_cluster(const ksb_result* r, int dist_col, double cutoff01, uint32_t* labels_out,
                uint64_t* n_edges);
/* same with the threshold already in percent, i.e. the reference's Clusters(cut_off_threshold=...)
 * constructor argument (ks_clustering.py:22,157: float(cutoff) * 100) */
int ksb_cluster_pct(const ksb_result* r, int dist_col, double cut_pct, uint32_t* labels_out, uint64_t* n_edges);
/* same, labels left in DEVICE memory of the result's device (d_labels_out[N]): the multi-GPU path
 * all-gathers and merges them there (ksb_cluster_merge_device) without a host round trip */
int ksb_cluster_pct_device(const ksb_result* r, int dist_col, double cut_pct, uint32_t* d_labels_out, uint64_t* n_edges);
float ksb_threshold_float_pct(double cut_pct);
int ksb_cluster_merge(uint64_t n, const uint32_t* labels_a, const uint32_t* labels_b, uint32_t* labels_out);
/* same on DEVICE buffers of the current device (e.g. the output of an NCCL all-gather):
 * d_labels_all is [world][n], merged labels land in d_out[n]; stream is a cudaStream_t (NULL = default) */
int ksb_cluster_merge_device(uint64_t n, int world, const uint32_t* d_labels_all, uint32_t* d_out, void* stream);
/* smallest float32 c such that the reference keeps an edge whose printed containment is c
 * (strtod("%g" of c) * 100 >= cutoff01 * 100); exposed for tests */
float ksb_threshold_float(double cutoff01);
void ksb_result_free(ksb_result* r);

/* ---- drop-in ------------------------------------------------------------------------------ */
/* kSpider::pairwise(std::string index_prefix, int user_threads) (include/kSpider.hpp:11) behind a
 * C signature: load -> pairwise on KSPIDER_GPUS devices -> both TSVs, same stdout lines. */
int ksb_dropin_pairwise(const char* index_prefix, int user_threads);

/* ---- colour construction on the GPU ---------------------------------------- */
/* Replaces the end state of the reference's incremental colouring loop (src/sourmash_indexing.cpp:
 * 141-277, src/index.cpp:189-331): sample s (s = 0..n_samples-1) lists the u64 hashes
 * hashes[offsets[s] .. offsets[s+1]); colour = the exact set of samples containing a hash,
 * count = number of distinct hashes with that set (colorsCount, :262-266), total[s] = entries listed
 * for s (mins.size(), :205).  One radix sort by hash, segmentation, a sort of the segments by a
 * signature of their sample lists and an exact grouping, all on `device` (-1 = current).  Members
 * are sample POSITIONS, ascending; singleton colours are included (the pairwise pass skips them).
 * Feed the arrays to ksb_index_from_csr / ksb_write_index_files. */
typedef struct ksb_colors ksb_colors;
int ksb_colors_build(uint64_t n_samples, const uint64_t* offsets, const uint64_t* hashes, int device, ksb_colors** out);
uint64_t ksb_colors_count(const ksb_colors* c);
uint64_t ksb_colors_distinct_hashes(const ksb_colors* c);
const uint64_t* ksb_colors_offsets(const ksb_colors* c);   /* [count + 1] */
const uint32_t* ksb_colors_members(const ksb_colors* c);
const uint64_t* ksb_colors_counts(const ksb_colors* c);    /* [count] */
const uint64_t* ksb_colors_totals(const ksb_colors* c);    /* [n_samples] */
float ksb_colors_device_ms(const ksb_colors* c);           /* H2D + sorts + kernels + D2H, CUDA events */
void ksb_colors_free(ksb_colors* c);

/* ---- text ---------------------------------------------------------------------------------- */
/* ostream<<float of the reference = "%g": writes at most 16 bytes, returns length (no NUL). */
int ksb_format_float(float v, char* buf);

/* ---- ANI column: `kSpider pairwise --estimate-ani` (pykSpider/kSpider2/ks_pairwise.py:29-82) ----------
 * ksb_ani_column: avg_ani of every record of `r`, in record order, computed on r's device (K8) from the
 * device-resident records: per row the PRINTED min / max containment (6 significant digits, as the
 * reference re-reads its own TSV) -> sourmash containment_to_distance(...).ani = 1 - (1 - c ** (1/ksize))
 * (0 -> 0.0, 1 -> 1.0) -> mean of the two.  ani_out: host, n_rows doubles.  n_invalid: rows where the
 * reference would raise (printed containment > 1, inf, nan); their value is NaN or out of range.
 * ksb_format_double_repr: Python's repr(float) (the text ks_pairwise.py:79 writes); buf >= 32 bytes.
 * ksb_write_ani_column: `P_kSpider_pairwise.ani_col.tsv` -- "avg_ani" header, one repr per row. */
int ksb_ani_column(const ksb_result* r, int ksize, double* ani_out, uint64_t* n_invalid);
int ksb_format_double_repr(double v, char* buf);
int ksb_write_ani_column(const double* ani, uint64_t n, const char* path, int host_threads);

/* ---- the pairwise TSV ON DISK as input: `kSpider cluster` and `pairwise --estimate-ani` -------------------
 * The reference's two consumers of the pairwise TSV read the FILE (pykSpider/kSpider2/ks_clustering.py:63-117,
 * ks_pairwise.py:64-73).  A ksb_table holds its parsed rows:
 *   ksb_table_read_tsv      multithreaded parse; per data line int(row[0]), int(row[1]) and float(row[dist_col])
 *                           exactly as Python parses them (correctly rounded; "inf", "nan", "-nan" accepted).
 *                           dist_col 3 = min, 4 = avg, 5 = max containment; 6 = "ani": the value of the same line
 *                           of the one-column file ani_col_tsv (ks_clustering.py:70-79).  dist_col2 (0 = none,
 *                           else 3..5) parses a second column (the ANI estimate needs min and max).
 *   ksb_write_sidecar       binary side-car of a result: the 32-byte records in the row order
 *                           of the TSV ksb_write_tsv writes, plus that TSV's size and modification time.  ksb_dropin_pairwise writes it
 *                           as <prefix>_kSpider_pairwise.ksb unless KSPIDER_SIDECAR=0.  tsv_path: the TSV it
 *                           belongs to (its size is recorded); append != 0 adds the rows of a further row block.
 *   ksb_sidecar_matches     1 if `sidecar` is intact and was written for exactly this pairwise_tsv (same size, same
 *                           modification time: a TSV rewritten later by any tool invalidates it), else 0.
 *   ksb_table_read_sidecar  the same table from the side-car: no text is parsed; values stay float32 and the
 *                           consumers apply the text rule through ksb_threshold_float_pct / the exact "%g" round trip.
 *   ksb_cluster_table       ks_clustering.py:45-47,63-120 on the GPU: n_nodes nodes (one per namesMap entry),
 *                           node = id - 1, row dropped iff float(text) * 100 < cut_pct (NaN kept), lock-free
 *                           union-find.  labels_out[n_nodes]: smallest node index of the node's component.
 *                           A KEPT row naming a node >= n_nodes fails with KSB_ERR_FORMAT (rustworkx raises).
 *   ksb_ani_table           ks_pairwise.py:64-80 on the GPU for a table holding columns 3 and 5
 *                           (ksb_table_read_*(.., 3, 5, ..)): ani_out[n_rows] in table (= file) order. */
typedef struct ksb_table ksb_table;
int ksb_table_read_tsv(const char* pairwise_tsv, int dist_col, int dist_col2, const char* ani_col_tsv, int host_threads,
                       ksb_table** out);
int ksb_table_read_sidecar(const char* sidecar, int dist_col, int dist_col2, int host_threads, ksb_table** out);
/* views into the table (valid until ksb_table_free); dist is NULL for side-car tables, cont for TSV tables */
int ksb_table_get(const ksb_table* t, uint64_t* n_rows, const uint32_t** source_1, const uint32_t** source_2,
                  const double** dist, const float** cont);
void ksb_table_free(ksb_table* t);
int ksb_write_sidecar(const ksb_result* r, const char* path, const char* tsv_path, int append);
int ksb_write_sidecar_records(const ksb_record* records, uint64_t n_rows, const char* path, uint64_t tsv_bytes,
                              uint64_t tsv_mtime_ns, int append);
int ksb_sidecar_matches(const char* sidecar, const char* pairwise_tsv);
/* names the (finished) TSV a side-car belongs to: size and modification time into its header.  ksb_dropin_pairwise writes
 * the TSV and the side-car at the same time and stamps the side-car when both are complete. */
int ksb_sidecar_stamp(const char* sidecar, const char* pairwise_tsv);
int ksb_cluster_table(const ksb_table* t, uint64_t n_nodes, double cut_pct, int device, uint32_t* labels_out,
                      uint64_t* n_edges);
int ksb_ani_table(const ksb_table* t, int ksize, int device, double* ani_out, uint64_t* n_invalid);

/* ---- allocation-free variant --------------------------------------------------------------------------------
 * ksb_cluster_edges_noalloc has NO counterpart in the reference; it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  It is the device step of ksb_cluster_table (threshold on the float32 containment
 * through ksb_threshold_float_pct, node = id - 1, lock-free union-find, label = smallest node index of the
 * component) on rows that are ALREADY in device memory, and it calls no cudaMalloc / cudaFree / allocating
 * library routine and does not synchronise: it enqueues four kernels on `stream` (a cudaStream_t as void*, used
 * as given: NULL = the CUDA default stream) of the CURRENT device and returns.
 *   d_source_1, d_source_2, d_cont   device, n_rows entries: sample ids (1-based, as in the TSV) and the float32
 *                                    containment column to threshold
 *   d_labels_out                     device, n_nodes entries
 *   d_counters_out                   device, two u64 (may be NULL): [0] = kept edges, [1] = 0xFFFFFFFF, or the
 *                                    smallest row index (saturated to 0xFFFFFFFE) of a KEPT row naming a node
 *                                    >= n_nodes -- such rows are skipped, where ksb_cluster_table fails
 * The union-find forest and the counters live in one file-scope `static __device__` array of
 * 256 + 4 * KSB_NOALLOC_MAX_NODES bytes that every GPU loading the library holds until exit; n_nodes beyond
 * KSB_NOALLOC_MAX_NODES is refused with KSB_ERR_ARG.  Every call on a device shares that array (the part a
 * call uses is re-initialised at its start): calls that overlap on the GPU -- two streams, two threads -- race;
 * order them.  The first call on a device resolves the array's device address (cudaGetSymbolAddress): make it
 * outside a graph capture (n_rows = 0 is enough). */
#define KSB_NOALLOC_MAX_NODES (1ull << 20)
int ksb_cluster_edges_noalloc(const uint32_t* d_source_1, const uint32_t* d_source_2, const float* d_cont, uint64_t n_rows,
                              uint64_t n_nodes, double cut_pct, void* stream, uint32_t* d_labels_out, uint64_t* d_counters_out);

/* ---- synthetic index files (bench / tests): real phmap dump bytes ------------------------- */
int ksb_write_index_files(const char* index_prefix, uint64_t n_colors, const uint64_t* color_ids,
                          const uint64_t* offsets, const uint32_t* members, uint64_t n_counts,
                          const uint64_t* count_ids, const uint64_t* count_vals, uint64_t n_totals,
                          const uint32_t* total_ids, const uint32_t* total_vals);

/* ---- micro-benchmarks used by bench.py to measure the red/atomic roofline in-run ---------- */
/* n_ops random (pattern 0) / warp-contiguous (pattern 1) red.global.add.{u64,u32} into a buffer of
 * `cells` cells; returns the best-of-`reps` milliseconds. */
int ksb_bench_red(uint64_t cells, uint64_t n_ops, int cell_bits, int pattern, int reps, float* best_ms);
/* int8 tensor-pipe ceiling in K4's own instruction shape (tcgen05.mma kind::i8, cta_group::1, M = 128, N = 256,
 * K = 32; one CTA per SM issuing `iters` K-blocks of 8 instructions on resident tiles): best-of-`reps`
 * milliseconds of one launch and the u8 multiply-adds it performs. */
int ksb_bench_imma(uint32_t iters, int reps, float* best_ms, double* macs_per_launch);

#ifdef __cplusplus
}
#endif
#endif /* KSPIDER_B200_H */
