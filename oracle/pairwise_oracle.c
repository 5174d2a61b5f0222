/*
 * oracle/pairwise_oracle.c — TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C CPU restatement of the reference's pairwise hot path, used ONLY as the checker in
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.  It is
 * never linked into, imported by, or called from the product (kspider_b200/), which must fail
 * loudly when its CUDA library is missing.
 *
 * Parity status: PINNED — this restatement is checked row-for-row against the UNMODIFIED
 * reference (oracle/_ref/pairwise_ref, built from /root/reference/src/pairwise.cpp by
 * oracle/Makefile) on the survey's micro-KAT and on seeded synthetic indexes
 * (tests/test_oracle_pinning.py, fixtures in tests/golden/).
 *
 * What each function follows (reference file:line, dib-lab/kSpider @ 466a267):
 *   ko_dump_*              lib/parallel-hashmap/parallel_hashmap/phmap_dump.h:52-83 (raw_hash_set
 *                          dump = size, capacity, ctrl[capacity+kWidth+1], slots[capacity]) and
 *                          phmap.h:284-296 (ctrl byte >= 0 <=> slot full)
 *   ko_index_load          src/pairwise.cpp:95-121 (load_colors_to_sources / load_colors_count,
 *                          incl. the u64->u32 narrowing of colour ids and counts, and
 *                          insert_or_assign = last one wins) and :166-170 (totals)
 *   ko_pairwise_csr        src/pairwise.cpp:209-225 (all member pairs of each colour, ascending(),
 *                          edges[(i,j)] += (uint32_t)count, inserted even when count == 0)
 *   ko_pairwise_csr_banded the same sums for indexes of BASELINE size (10^9..10^10 updates), where a
 *                          single-threaded hash map takes hours: per smaller sample i the cells (i, j) of
 *                          the columns j it is ever paired with, accumulated by all host threads with
 *                          atomic adds (OpenMP), a byte per cell for "touched" (count 0 still makes a row).
 *                          Another container, the same arithmetic; pinned against ko_pairwise_csr on every
 *                          golden case and on seeded inputs (tests/test_oracle_pinning.py).
 *   ko_containment         src/pairwise.cpp:257-264 (float32 math)
 *   ko_write_tsv           src/pairwise.cpp:242-275 (header, columns, ostream<<float == "%g")
 *   ko_write_seq_to_kmers  src/pairwise.cpp:173-180
 *
 * Row order: the reference iterates a hash map (order unspecified); the oracle emits rows sorted by
 * (source_1, source_2) and every comparison is done on sorted rows.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define KO_OK 0
#define KO_ERR_IO -1
#define KO_ERR_FORMAT -2
#define KO_ERR_NOMEM -3

typedef struct {
    uint64_t n_colors;   /* colours after u32-key de-duplication (ALL sizes, incl. singletons) */
    uint32_t* color_ids; /* [n_colors] narrowed ids */
    uint64_t* offsets;   /* [n_colors+1] */
    uint32_t* members;   /* [offsets[n_colors]] in file slot order */
    uint32_t* weights;   /* [n_colors] (uint32_t)count, 0 when the colour has no count entry */
    uint64_t n_totals;
    uint32_t* total_ids; /* [n_totals] in file slot order (= reference iteration order) */
    uint32_t* total_vals;
} ko_index;

typedef struct {
    uint64_t n_rows;
    uint32_t* src1;
    uint32_t* src2;
    uint64_t* shared;
    float* cmin;
    float* cavg;
    float* cmax;
} ko_result;

/* ------------------------------------------------------------------ file helpers */

static unsigned char* ko_read_file(const char* path, uint64_t* len) {
    FILE* f = fopen(path, "rb");
    if (!f) return NULL;
    fseek(f, 0, SEEK_END);
    long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    unsigned char* buf = (unsigned char*)malloc(n > 0 ? (size_t)n : 1);
    if (!buf) { fclose(f); return NULL; }
    if (n > 0 && fread(buf, 1, (size_t)n, f) != (size_t)n) { free(buf); fclose(f); return NULL; }
    fclose(f);
    *len = (uint64_t)n;
    return buf;
}

/* One raw_hash_set dump starting at p (phmap_dump.h:52-63).  kWidth is 16 on SSE2 builds
 * (phmap.h:395,522-526).  Returns bytes consumed, 0 on a malformed dump. */
typedef struct {
    uint64_t size, capacity;
    const int8_t* ctrl;
    const unsigned char* slots;
} ko_dump;

static uint64_t ko_dump_open(const unsigned char* p, uint64_t avail, uint64_t slot_bytes, ko_dump* d) {
    const uint64_t kWidth = 16;
    if (avail < 16) return 0;
    memcpy(&d->size, p, 8);
    memcpy(&d->capacity, p + 8, 8);
    d->ctrl = NULL;
    d->slots = NULL;
    if (d->size == 0) return 16;
    if (d->capacity == 0 || ((d->capacity + 1) & d->capacity) != 0) return 0; /* must be 2^k-1 */
    if (d->size > d->capacity) return 0;
    uint64_t ctrl_bytes = d->capacity + kWidth + 1;
    uint64_t need = 16 + ctrl_bytes + slot_bytes * d->capacity;
    if (avail < need) return 0;
    d->ctrl = (const int8_t*)(p + 16);
    d->slots = p + 16 + ctrl_bytes;
    return need;
}

/* ------------------------------------------------------------------ tiny u32->u64 hash map */

typedef struct {
    uint64_t cap; /* power of two */
    uint64_t n;
    uint32_t* keys;
    uint64_t* vals;
    uint8_t* used;
} ko_map32;

static int ko_map32_init(ko_map32* m, uint64_t expect) {
    uint64_t cap = 16;
    while (cap < expect * 2 + 2) cap <<= 1;
    m->cap = cap;
    m->n = 0;
    m->keys = (uint32_t*)malloc(cap * 4);
    m->vals = (uint64_t*)malloc(cap * 8);
    m->used = (uint8_t*)calloc(cap, 1);
    return (m->keys && m->vals && m->used) ? 0 : -1;
}
static void ko_map32_free(ko_map32* m) { free(m->keys); free(m->vals); free(m->used); }
static uint64_t ko_mix64(uint64_t x) {
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
    return x;
}
/* returns slot index; *found tells whether key existed */
static uint64_t ko_map32_slot(ko_map32* m, uint32_t key, int* found) {
    uint64_t h = ko_mix64(key) & (m->cap - 1);
    while (m->used[h]) {
        if (m->keys[h] == key) { *found = 1; return h; }
        h = (h + 1) & (m->cap - 1);
    }
    *found = 0;
    return h;
}

/* ------------------------------------------------------------------ index load */

void ko_index_free(ko_index* ix) {
    if (!ix) return;
    free(ix->color_ids); free(ix->offsets); free(ix->members); free(ix->weights);
    free(ix->total_ids); free(ix->total_vals); free(ix);
}

int ko_index_load(const char* prefix, ko_index** out) {
    char path[4096];
    uint64_t len = 0;
    int rc = KO_ERR_FORMAT;
    ko_index* ix = (ko_index*)calloc(1, sizeof(ko_index));
    unsigned char *fs = NULL, *fc = NULL, *ft = NULL;
    ko_map32 cmap; memset(&cmap, 0, sizeof cmap);
    uint64_t* rec_pos = NULL;
    if (!ix) return KO_ERR_NOMEM;

    /* --- colour -> sources (pairwise.cpp:95-111): u64 n, then n x {u64 colour, set<u32> dump}.
     * map key is uint32_t: the colour id is narrowed and insert_or_assign keeps the LAST record. */
    snprintf(path, sizeof path, "%s_color_to_sources.bin", prefix);
    fs = ko_read_file(path, &len);
    if (!fs) { rc = KO_ERR_IO; goto fail; }
    if (len < 8) goto fail;
    uint64_t n_rec; memcpy(&n_rec, fs, 8);
    if (n_rec > len) goto fail;
    rec_pos = (uint64_t*)malloc((n_rec + 1) * 8);
    if (!rec_pos || ko_map32_init(&cmap, n_rec)) { rc = KO_ERR_NOMEM; goto fail; }
    {
        uint64_t pos = 8;
        for (uint64_t r = 0; r < n_rec; r++) {
            if (len - pos < 8) goto fail;
            uint64_t color; memcpy(&color, fs + pos, 8);
            ko_dump d;
            uint64_t used = ko_dump_open(fs + pos + 8, len - pos - 8, 4, &d);
            if (!used) goto fail;
            rec_pos[r] = pos;
            int found;
            uint64_t s = ko_map32_slot(&cmap, (uint32_t)color, &found);
            if (!found) { cmap.used[s] = 1; cmap.keys[s] = (uint32_t)color; cmap.n++; }
            cmap.vals[s] = r; /* last record with this narrowed id wins */
            pos += 8 + used;
        }
    }
    /* surviving records in first-seen order of file position */
    ix->n_colors = cmap.n;
    ix->color_ids = (uint32_t*)malloc((cmap.n + 1) * 4);
    ix->offsets = (uint64_t*)malloc((cmap.n + 1) * 8);
    ix->weights = (uint32_t*)calloc(cmap.n + 1, 4);
    if (!ix->color_ids || !ix->offsets || !ix->weights) { rc = KO_ERR_NOMEM; goto fail; }
    {
        uint64_t total_members = 0, k = 0;
        for (uint64_t r = 0; r < n_rec; r++) {
            uint64_t color; memcpy(&color, fs + rec_pos[r], 8);
            int found; uint64_t s = ko_map32_slot(&cmap, (uint32_t)color, &found);
            if (cmap.vals[s] != r) continue;
            uint64_t sz; memcpy(&sz, fs + rec_pos[r] + 8, 8);
            total_members += sz;
        }
        ix->members = (uint32_t*)malloc((total_members + 1) * 4);
        if (!ix->members) { rc = KO_ERR_NOMEM; goto fail; }
        uint64_t w = 0;
        for (uint64_t r = 0; r < n_rec; r++) {
            uint64_t color; memcpy(&color, fs + rec_pos[r], 8);
            int found; uint64_t s = ko_map32_slot(&cmap, (uint32_t)color, &found);
            if (cmap.vals[s] != r) continue;
            ko_dump d;
            ko_dump_open(fs + rec_pos[r] + 8, len - rec_pos[r] - 8, 4, &d);
            ix->color_ids[k] = (uint32_t)color;
            ix->offsets[k] = w;
            uint64_t live = 0;
            for (uint64_t i = 0; i < d.capacity && d.size; i++)
                if (d.ctrl[i] >= 0) { memcpy(&ix->members[w++], d.slots + 4 * i, 4); live++; }
            if (live != d.size) goto fail;
            cmap.vals[s] = k; /* now: narrowed id -> colour index */
            k++;
        }
        ix->offsets[k] = w;
    }

    /* --- colour counts (pairwise.cpp:113-121): map<u64,u64> dump, key and value narrowed to u32,
     * insert_or_assign in slot order => the last slot with a given narrowed key wins. */
    snprintf(path, sizeof path, "%s_color_count.bin", prefix);
    fc = ko_read_file(path, &len);
    if (!fc) { rc = KO_ERR_IO; goto fail; }
    {
        ko_dump d;
        if (!ko_dump_open(fc, len, 16, &d)) goto fail;
        if (d.size == 0) goto fail; /* reference: assert(tmpMap.size()) pairwise.cpp:117 */
        uint64_t live = 0;
        for (uint64_t i = 0; i < d.capacity; i++) {
            if (d.ctrl[i] < 0) continue;
            uint64_t kv[2]; memcpy(kv, d.slots + 16 * i, 16);
            live++;
            int found; uint64_t s = ko_map32_slot(&cmap, (uint32_t)kv[0], &found);
            if (found) ix->weights[cmap.vals[s]] = (uint32_t)kv[1];
        }
        if (live != d.size) goto fail;
    }

    /* --- per-sample totals (pairwise.cpp:166-170): map<u32,u32> dump, kept in slot order. */
    snprintf(path, sizeof path, "%s_groupID_to_kmerCount.bin", prefix);
    ft = ko_read_file(path, &len);
    if (!ft) { rc = KO_ERR_IO; goto fail; }
    {
        ko_dump d;
        if (!ko_dump_open(ft, len, 8, &d)) goto fail;
        if (d.size == 0) goto fail; /* assert(groupID_to_kmerCount.size()) pairwise.cpp:170 */
        ix->total_ids = (uint32_t*)malloc(d.size * 4);
        ix->total_vals = (uint32_t*)malloc(d.size * 4);
        if (!ix->total_ids || !ix->total_vals) { rc = KO_ERR_NOMEM; goto fail; }
        uint64_t k = 0;
        for (uint64_t i = 0; i < d.capacity; i++) {
            if (d.ctrl[i] < 0) continue;
            if (k >= d.size) goto fail;
            memcpy(&ix->total_ids[k], d.slots + 8 * i, 4);
            memcpy(&ix->total_vals[k], d.slots + 8 * i + 4, 4);
            k++;
        }
        if (k != d.size) goto fail;
        ix->n_totals = k;
    }
    free(fs); free(fc); free(ft); free(rec_pos); ko_map32_free(&cmap);
    *out = ix;
    return KO_OK;
fail:
    free(fs); free(fc); free(ft); free(rec_pos); ko_map32_free(&cmap);
    ko_index_free(ix);
    return rc;
}

/* ------------------------------------------------------------------ pair accumulation */

typedef struct {
    uint64_t cap, n;
    uint64_t* keys; /* (i << 32) | j, i < j ; UINT64_MAX = empty (i<j so never a real key) */
    uint64_t* vals;
} ko_pairmap;

static int ko_pairmap_init(ko_pairmap* m, uint64_t cap) {
    m->cap = cap; m->n = 0;
    m->keys = (uint64_t*)malloc(cap * 8);
    m->vals = (uint64_t*)malloc(cap * 8);
    if (!m->keys || !m->vals) return -1;
    memset(m->keys, 0xFF, cap * 8);
    return 0;
}
static int ko_pairmap_add(ko_pairmap* m, uint64_t key, uint64_t w);
static int ko_pairmap_grow(ko_pairmap* m) {
    ko_pairmap b;
    if (ko_pairmap_init(&b, m->cap * 2)) return -1;
    for (uint64_t i = 0; i < m->cap; i++)
        if (m->keys[i] != UINT64_MAX) ko_pairmap_add(&b, m->keys[i], m->vals[i]);
    free(m->keys); free(m->vals);
    *m = b;
    return 0;
}
static int ko_pairmap_add(ko_pairmap* m, uint64_t key, uint64_t w) {
    if ((m->n + 1) * 10 > m->cap * 7 && ko_pairmap_grow(m)) return -1;
    uint64_t h = ko_mix64(key) & (m->cap - 1);
    while (m->keys[h] != UINT64_MAX) {
        if (m->keys[h] == key) { m->vals[h] += w; return 0; } /* v.second += ccount */
        h = (h + 1) & (m->cap - 1);
    }
    m->keys[h] = key; m->vals[h] = w; m->n++;                  /* inserted with ccount, even 0 */
    return 0;
}

typedef struct { uint64_t key, val; } ko_kv;
static int ko_kv_cmp(const void* a, const void* b) {
    uint64_t x = ((const ko_kv*)a)->key, y = ((const ko_kv*)b)->key;
    return x < y ? -1 : (x > y ? 1 : 0);
}

/* float32 containment math exactly as src/pairwise.cpp:260-264.
 * std::min(a,b) = (b < a) ? b : a ; std::max(a,b) = (a < b) ? b : a. */
void ko_containment(uint64_t shared, uint32_t k1, uint32_t k2, float* cmin, float* cavg, float* cmax) {
    volatile float c12 = (float)shared / (float)k2; /* cont_1_in_2 */
    volatile float c21 = (float)shared / (float)k1; /* cont_2_in_1 */
    float a = c12, b = c21;
    *cmin = (b < a) ? b : a;
    volatile float sum = a + b;
    *cavg = (float)((double)sum / 2.0);
    *cmax = (a < b) ? b : a;
}

void ko_result_free(ko_result* r) {
    if (!r) return;
    free(r->src1); free(r->src2); free(r->shared); free(r->cmin); free(r->cavg); free(r->cmax); free(r);
}

int ko_pairwise_csr(uint64_t n_colors, const uint64_t* offsets, const uint32_t* members,
                    const uint32_t* weights, uint64_t n_totals, const uint32_t* total_ids,
                    const uint32_t* total_vals, ko_result** out) {
    ko_pairmap pm;
    if (ko_pairmap_init(&pm, 1 << 16)) return KO_ERR_NOMEM;
    for (uint64_t c = 0; c < n_colors; c++) {
        const uint32_t* m = members + offsets[c];
        uint64_t s = offsets[c + 1] - offsets[c];
        uint64_t w = weights[c];
        for (uint64_t a = 0; a < s; a++)
            for (uint64_t b = a + 1; b < s; b++) {
                uint32_t i = m[a], j = m[b];
                if (i > j) { uint32_t t = i; i = j; j = t; } /* ascending() */
                if (ko_pairmap_add(&pm, ((uint64_t)i << 32) | j, w)) return KO_ERR_NOMEM;
            }
    }
    ko_kv* kv = (ko_kv*)malloc((pm.n + 1) * sizeof(ko_kv));
    if (!kv) return KO_ERR_NOMEM;
    uint64_t n = 0;
    for (uint64_t i = 0; i < pm.cap; i++)
        if (pm.keys[i] != UINT64_MAX) { kv[n].key = pm.keys[i]; kv[n].val = pm.vals[i]; n++; }
    free(pm.keys); free(pm.vals);
    qsort(kv, n, sizeof(ko_kv), ko_kv_cmp);

    ko_map32 tm;
    if (ko_map32_init(&tm, n_totals)) { free(kv); return KO_ERR_NOMEM; }
    for (uint64_t i = 0; i < n_totals; i++) {
        int found; uint64_t s = ko_map32_slot(&tm, total_ids[i], &found);
        if (!found) { tm.used[s] = 1; tm.keys[s] = total_ids[i]; tm.n++; }
        tm.vals[s] = total_vals[i];
    }
    ko_result* r = (ko_result*)calloc(1, sizeof(ko_result));
    r->n_rows = n;
    r->src1 = (uint32_t*)malloc((n + 1) * 4); r->src2 = (uint32_t*)malloc((n + 1) * 4);
    r->shared = (uint64_t*)malloc((n + 1) * 8);
    r->cmin = (float*)malloc((n + 1) * 4); r->cavg = (float*)malloc((n + 1) * 4); r->cmax = (float*)malloc((n + 1) * 4);
    for (uint64_t x = 0; x < n; x++) {
        uint32_t i = (uint32_t)(kv[x].key >> 32), j = (uint32_t)kv[x].key;
        int f1, f2;
        uint64_t s1 = ko_map32_slot(&tm, i, &f1), s2 = ko_map32_slot(&tm, j, &f2);
        uint32_t k1 = f1 ? (uint32_t)tm.vals[s1] : 0; /* operator[] default 0 (pairwise.cpp:257-258) */
        uint32_t k2 = f2 ? (uint32_t)tm.vals[s2] : 0;
        r->src1[x] = i; r->src2[x] = j; r->shared[x] = kv[x].val;
        ko_containment(kv[x].val, k1, k2, &r->cmin[x], &r->cavg[x], &r->cmax[x]);
    }
    ko_map32_free(&tm);
    free(kv);
    *out = r;
    return KO_OK;
}

/* u32 ascending */
static int ko_u32_cmp(const void* a, const void* b) {
    uint32_t x = *(const uint32_t*)a, y = *(const uint32_t*)b;
    return x < y ? -1 : (x > y ? 1 : 0);
}
static void ko_isort_u32(uint32_t* a, uint64_t n) {
    if (n > 48) { qsort(a, n, 4, ko_u32_cmp); return; }
    for (uint64_t i = 1; i < n; i++) {
        uint32_t v = a[i];
        uint64_t j = i;
        while (j > 0 && a[j - 1] > v) { a[j] = a[j - 1]; j--; }
        a[j] = v;
    }
}

/* Large-index variant of ko_pairwise_csr (see the header): same result, rows in (source_1, source_2) order. */
int ko_pairwise_csr_banded(uint64_t n_colors, const uint64_t* offsets, const uint32_t* members,
                           const uint32_t* weights, uint64_t n_totals, const uint32_t* total_ids,
                           const uint32_t* total_vals, int threads, ko_result** out) {
    const uint64_t M = offsets[n_colors];
    if (threads < 1) threads = 1;
    /* compact sample index = rank among the distinct member ids */
    uint32_t* ids = (uint32_t*)malloc((M + 1) * 4);
    if (!ids) return KO_ERR_NOMEM;
    uint64_t N = 0;
    uint32_t max_id = 0;
    for (uint64_t i = 0; i < M; i++) if (members[i] > max_id) max_id = members[i];
    uint32_t* rank = NULL; /* dense ids: id -> compact index by table instead of by search */
    if (M && (uint64_t)max_id < (1ull << 28)) {
        rank = (uint32_t*)calloc((uint64_t)max_id + 1, 4);
        if (!rank) return KO_ERR_NOMEM;
        for (uint64_t i = 0; i < M; i++) rank[members[i]] = 1;
        for (uint64_t v = 0; v <= max_id; v++)
            if (rank[v]) { rank[v] = (uint32_t)N; ids[N++] = (uint32_t)v; }
    } else {
        memcpy(ids, members, M * 4);
        qsort(ids, M, 4, ko_u32_cmp);
        for (uint64_t i = 0; i < M; i++)
            if (i == 0 || ids[i] != ids[i - 1]) ids[N++] = ids[i];
    }
    /* sorted compact members per colour */
    uint32_t* cm = (uint32_t*)malloc((M + 1) * 4);
    uint32_t* lo = (uint32_t*)malloc((N + 1) * 4);
    uint32_t* hi = (uint32_t*)calloc(N + 1, 4);
    uint64_t* start = (uint64_t*)calloc(N + 2, 8);
    if (!cm || !lo || !hi || !start) return KO_ERR_NOMEM;
    for (uint64_t i = 0; i < N; i++) lo[i] = 0xFFFFFFFFu;
    int zero_w = 0;
#pragma omp parallel for schedule(dynamic, 4096) num_threads(threads) reduction(| : zero_w)
    for (uint64_t c = 0; c < n_colors; c++) {
        const uint64_t b = offsets[c], s = offsets[c + 1] - b;
        for (uint64_t a = 0; a < s; a++) {
            const uint32_t v = members[b + a];
            if (rank) { cm[b + a] = rank[v]; continue; }
            uint64_t l = 0, h = N;
            while (h - l > 1) { const uint64_t m = (l + h) >> 1; if (ids[m] <= v) l = m; else h = m; }
            cm[b + a] = (uint32_t)l;
        }
        ko_isort_u32(cm + b, s);
        if (s >= 2 && weights[c] == 0) zero_w = 1;
    }
    /* bands: row m[p] is paired with the columns m[p+1] .. m[s-1] */
    for (uint64_t c = 0; c < n_colors; c++) {
        const uint64_t b = offsets[c], s = offsets[c + 1] - b;
        for (uint64_t p = 0; p + 1 < s; p++) {
            const uint32_t r = cm[b + p];
            if (cm[b + p + 1] < lo[r]) lo[r] = cm[b + p + 1];
            if (cm[b + s - 1] + 1 > hi[r]) hi[r] = cm[b + s - 1] + 1;
        }
    }
    for (uint64_t i = 0; i < N; i++) {
        if (lo[i] == 0xFFFFFFFFu) lo[i] = hi[i] = 0;
        start[i + 1] = start[i] + (hi[i] - lo[i]);
    }
    const uint64_t cells = start[N];
    uint64_t* acc = (uint64_t*)calloc(cells + 1, 8);
    uint8_t* touched = (uint8_t*)calloc(cells + 1, 1);
    if (!acc || !touched) return KO_ERR_NOMEM;
#pragma omp parallel for schedule(dynamic, 1024) num_threads(threads)
    for (uint64_t c = 0; c < n_colors; c++) {
        const uint64_t b = offsets[c], s = offsets[c + 1] - b;
        const uint64_t w = weights[c];
        for (uint64_t p = 0; p + 1 < s; p++) {
            const uint32_t r = cm[b + p];
            uint64_t* row = acc + start[r] - lo[r];
            uint8_t* trow = touched + start[r] - lo[r];
            for (uint64_t q = p + 1; q < s; q++) {
                const uint32_t j = cm[b + q];
                if (j == r) continue; /* cannot happen: member lists are sets */
                if (w) __atomic_fetch_add(&row[j], w, __ATOMIC_RELAXED);      /* v.second += ccount */
                if (zero_w || !w) trow[j] = 1;                                 /* inserted even when count == 0 */
            }
        }
    }
    /* rows per sample, prefix, emit */
    uint64_t* rfirst = (uint64_t*)calloc(N + 2, 8);
    if (!rfirst) return KO_ERR_NOMEM;
#pragma omp parallel for schedule(static) num_threads(threads)
    for (uint64_t i = 0; i < N; i++) {
        uint64_t n = 0;
        for (uint64_t k = start[i]; k < start[i + 1]; k++) n += (acc[k] != 0 || touched[k]);
        rfirst[i + 1] = n;
    }
    for (uint64_t i = 0; i < N; i++) rfirst[i + 1] += rfirst[i];
    const uint64_t n = rfirst[N];
    ko_map32 tm;
    if (ko_map32_init(&tm, n_totals)) return KO_ERR_NOMEM;
    for (uint64_t i = 0; i < n_totals; i++) {
        int found; uint64_t sl = ko_map32_slot(&tm, total_ids[i], &found);
        if (!found) { tm.used[sl] = 1; tm.keys[sl] = total_ids[i]; tm.n++; }
        tm.vals[sl] = total_vals[i];
    }
    uint32_t* tot = (uint32_t*)calloc(N + 1, 4);
    for (uint64_t i = 0; i < N; i++) {
        int f; uint64_t sl = ko_map32_slot(&tm, ids[i], &f);
        tot[i] = f ? (uint32_t)tm.vals[sl] : 0;  /* operator[] default 0 (pairwise.cpp:257-258) */
    }
    ko_map32_free(&tm);
    ko_result* r = (ko_result*)calloc(1, sizeof(ko_result));
    r->n_rows = n;
    r->src1 = (uint32_t*)malloc((n + 1) * 4); r->src2 = (uint32_t*)malloc((n + 1) * 4);
    r->shared = (uint64_t*)malloc((n + 1) * 8);
    r->cmin = (float*)malloc((n + 1) * 4); r->cavg = (float*)malloc((n + 1) * 4); r->cmax = (float*)malloc((n + 1) * 4);
    if (!r->src1 || !r->src2 || !r->shared || !r->cmin || !r->cavg || !r->cmax) return KO_ERR_NOMEM;
#pragma omp parallel for schedule(dynamic, 64) num_threads(threads)
    for (uint64_t i = 0; i < N; i++) {
        uint64_t x = rfirst[i];
        for (uint64_t k = start[i]; k < start[i + 1]; k++) {
            if (acc[k] == 0 && !touched[k]) continue;
            const uint32_t j = lo[i] + (uint32_t)(k - start[i]);
            r->src1[x] = ids[i]; r->src2[x] = ids[j]; r->shared[x] = acc[k];
            ko_containment(acc[k], tot[i], tot[j], &r->cmin[x], &r->cavg[x], &r->cmax[x]);
            x++;
        }
    }
    free(rank); free(ids); free(cm); free(lo); free(hi); free(start); free(acc); free(touched); free(rfirst); free(tot);
    *out = r;
    return KO_OK;
}

int ko_pairwise(const ko_index* ix, ko_result** out) {
    return ko_pairwise_csr(ix->n_colors, ix->offsets, ix->members, ix->weights, ix->n_totals,
                           ix->total_ids, ix->total_vals, out);
}

/* ostream << float with default flags == printf("%g") of the promoted double (precision 6). */
int ko_format_float(float v, char* buf, size_t n) { return snprintf(buf, n, "%g", (double)v); }

int ko_write_tsv(const ko_result* r, const char* path) {
    FILE* f = fopen(path, "w");
    if (!f) return KO_ERR_IO;
    fputs("source_1\tsource_2\tshared_kmers\tmin_containment\tavg_containment\tmax_containment\n", f);
    for (uint64_t x = 0; x < r->n_rows; x++)
        fprintf(f, "%u\t%u\t%llu\t%g\t%g\t%g\n", r->src1[x], r->src2[x], (unsigned long long)r->shared[x],
                (double)r->cmin[x], (double)r->cavg[x], (double)r->cmax[x]);
    fclose(f);
    return KO_OK;
}

int ko_write_seq_to_kmers(const ko_index* ix, const char* path) {
    FILE* f = fopen(path, "w");
    if (!f) return KO_ERR_IO;
    fputs("ID\tseq\tkmers\n", f);
    for (uint64_t i = 0; i < ix->n_totals; i++)
        fprintf(f, "%llu\t%u\t%u\n", (unsigned long long)(i + 1), ix->total_ids[i], ix->total_vals[i]);
    fclose(f);
    return KO_OK;
}

/* accessors for ctypes */
uint64_t ko_index_n_colors(const ko_index* ix) { return ix->n_colors; }
uint64_t ko_index_n_members(const ko_index* ix) { return ix->offsets[ix->n_colors]; }
uint64_t ko_index_n_totals(const ko_index* ix) { return ix->n_totals; }
const uint32_t* ko_index_color_ids(const ko_index* ix) { return ix->color_ids; }
const uint64_t* ko_index_offsets(const ko_index* ix) { return ix->offsets; }
const uint32_t* ko_index_members(const ko_index* ix) { return ix->members; }
const uint32_t* ko_index_weights(const ko_index* ix) { return ix->weights; }
const uint32_t* ko_index_total_ids(const ko_index* ix) { return ix->total_ids; }
const uint32_t* ko_index_total_vals(const ko_index* ix) { return ix->total_vals; }
uint64_t ko_result_n_rows(const ko_result* r) { return r->n_rows; }
const uint32_t* ko_result_src1(const ko_result* r) { return r->src1; }
const uint32_t* ko_result_src2(const ko_result* r) { return r->src2; }
const uint64_t* ko_result_shared(const ko_result* r) { return r->shared; }
const float* ko_result_cmin(const ko_result* r) { return r->cmin; }
const float* ko_result_cavg(const ko_result* r) { return r->cavg; }
const float* ko_result_cmax(const ko_result* r) { return r->cmax; }
