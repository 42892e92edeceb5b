/* duckhts_ext.c — the DuckDB side of duckhts_b200: a C extension that registers `read_bam`, `read_bcf` and `bgunzip`
 * with the same names, parameters, column schemas and error behaviour as the reference (src/bam_reader.c:1044-1068,
 * src/bcf_reader.c:2055-2081, src/bgzip.c:359-380) and serves them from the GPU through the C-ABI of libduckhts_cuda.so
 * (include/duckhts_cuda.h).  Host code in C, as the reference's; no htslib, no CPU decode path.
 *
 * Entry point: duckhts_init_c_api (reference duckdb_capi/duckdb_extension.h:1179-1195) — the extension is a
 * drop-in for the reference's shared object as far as these table functions go.
 *
 * How a scan maps onto the reference's callbacks (src/bam_reader.c:410-1038):
 *   bind        opens the file on the first GPU (BGZF framing on the host, header inflated on the GPU) in a dh_ctx of its
 *               own (queries never share streams), declares the schema (13 SAM columns [+ 56 SAMtags columns] [+ MAP])
 *   global init projection pushdown -> dh_bam_set_projection; cuts the file into one SOURCE per GPU (contig groups from
 *               the index where the reference scans contig by contig, else virtual-offset ranges; a region query stays on
 *               one GPU), starts every source's pipeline, max_threads = 64
 *   local init  numbers the worker
 *   scan        a worker CLAIMS the next duckdb_vector_size() rows of a source's current batch under that source's lock and
 *               fills its chunk outside the lock: fixed-width columns by memcpy / widening loops, VARCHAR cells of <= 12
 *               bytes by a 16-byte store, longer strings through duckdb_vector_assign_string_element_len, NULLs through
 *               the validity mask.  Batches are reference-counted and go back to the pipeline when their last claim is
 *               filled.  One worker walks the sources in order = file order (the reference's order); several
 *               workers interleave chunks like DuckDB interleaves the reference's threads.
 * Batches cross PCIe in the compact wire layout (include/duckhts_cuda.h, DH_BAM_COMPACT_OUTPUT) unless AUXILIARY_TAGS is
 * projected or DUCKHTS_LAYOUT=duckdb asks for ready-made DuckDB vectors.
 */
#define _GNU_SOURCE
#include <errno.h>
#include <fcntl.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#include "duckdb_ext_abi.h"
#include "duckhts_cuda.h"

static duckdb_ext_api_v1 api;   /* the host's function table, copied by value like DUCKDB_EXTENSION_API_INIT does */

/* ---- GPUs ----------------------------------------------------------------------------------------------------------
 * DUCKHTS_DEVICES = comma-separated CUDA device ordinals a query may use (a device may be named twice: two pipelines on
 * one GPU); DUCKHTS_DEVICE = one ordinal (older spelling); default: every visible device.  The first entry serves bind. */
#define MAX_SOURCES 16
static int device_list(int *devs) {
    int n = 0;
    const char *e = getenv("DUCKHTS_DEVICES");
    if (e && *e) {
        char *dup = strdup(e), *save = NULL;
        for (char *tok = strtok_r(dup, ",", &save); tok && n < MAX_SOURCES; tok = strtok_r(NULL, ",", &save)) devs[n++] = atoi(tok);
        free(dup);
    } else if ((e = getenv("DUCKHTS_DEVICE")) && *e) devs[n++] = atoi(e);
    else {
        const int have = dh_device_count();
        for (int d = 0; d < have && n < MAX_SOURCES; ++d) devs[n++] = d;
    }
    if (n == 0) devs[n++] = 0;
    return n;
}
/* Contexts (streams, events, staging) outlive the query that made them: a session's next query on the same device takes one from the
 * pool instead of creating CUDA streams and re-allocating.  One context serves one query at a time. */
static pthread_mutex_t g_pool_mu = PTHREAD_MUTEX_INITIALIZER;
static struct { dh_ctx *ctx; int device, busy; } g_pool[64];
static dh_ctx *ctx_acquire(int device) {
    pthread_mutex_lock(&g_pool_mu);
    int free_slot = -1;
    for (int i = 0; i < 64; ++i) {
        if (g_pool[i].ctx && !g_pool[i].busy && g_pool[i].device == device) { g_pool[i].busy = 1; dh_ctx *c = g_pool[i].ctx; pthread_mutex_unlock(&g_pool_mu); return c; }
        if (!g_pool[i].ctx && free_slot < 0) free_slot = i;
    }
    pthread_mutex_unlock(&g_pool_mu);
    dh_ctx *c = dh_ctx_create(device);
    if (!c) return NULL;
    pthread_mutex_lock(&g_pool_mu);
    for (int i = 0; i < 64 && free_slot >= 0; ++i)
        if (!g_pool[i].ctx) { g_pool[i].ctx = c; g_pool[i].device = device; g_pool[i].busy = 1; free_slot = -2; break; }
    pthread_mutex_unlock(&g_pool_mu);
    return c;       /* pool full: the context is simply destroyed when released */
}
static void ctx_release(dh_ctx *c) {
    if (!c) return;
    pthread_mutex_lock(&g_pool_mu);
    for (int i = 0; i < 64; ++i)
        if (g_pool[i].ctx == c) { g_pool[i].busy = 0; pthread_mutex_unlock(&g_pool_mu); return; }
    pthread_mutex_unlock(&g_pool_mu);
    dh_ctx_destroy(c);
}

static uint64_t env_u64(const char *name, uint64_t dflt) {
    const char *e = getenv(name);
    return (e && *e) ? strtoull(e, NULL, 0) : dflt;
}

/* whole file -> pinned host memory (the C-ABI copies from it at full PCIe rate); read by a few threads, because one
 * thread copies out of the page cache at a fraction of what the GPU takes */
typedef struct { int fd; uint8_t *dst; uint64_t off, len; int ok; } slurp_part;
static void *slurp_worker(void *arg) {
    slurp_part *p = (slurp_part *)arg;
    uint64_t got = 0;
    p->ok = 1;
    while (got < p->len) {
        const ssize_t k = pread(p->fd, p->dst + got, (size_t)((p->len - got) < (64u << 20) ? (p->len - got) : (64u << 20)), (off_t)(p->off + got));
        if (k <= 0) { if (k < 0 && errno == EINTR) continue; p->ok = 0; break; }
        got += (uint64_t)k;
    }
    return NULL;
}
static uint8_t *slurp_pinned(const char *path, uint64_t *size_out) {
    const int fd = open(path, O_RDONLY);
    if (fd < 0) return NULL;
    struct stat st;
    if (fstat(fd, &st) != 0 || !S_ISREG(st.st_mode)) { close(fd); return NULL; }
    const uint64_t n = (uint64_t)st.st_size;
    uint8_t *p = (uint8_t *)dh_host_alloc(n ? n : 1);
    if (!p) { close(fd); return NULL; }
    int nt = (int)(n >> 26);            /* one reader per 64 MiB, at most 8 */
    if (nt < 1) nt = 1;
    if (nt > 8) nt = 8;
    slurp_part parts[8];
    pthread_t th[8];
    for (int i = 0; i < nt; ++i) {
        parts[i].fd = fd; parts[i].dst = p + n * (uint64_t)i / (uint64_t)nt; parts[i].off = n * (uint64_t)i / (uint64_t)nt;
        parts[i].len = n * (uint64_t)(i + 1) / (uint64_t)nt - parts[i].off; parts[i].ok = 0;
        if (i > 0 && pthread_create(&th[i], NULL, slurp_worker, &parts[i]) != 0) { slurp_worker(&parts[i]); th[i] = 0; }
    }
    slurp_worker(&parts[0]);
    int ok = parts[0].ok;
    for (int i = 1; i < nt; ++i) { if (th[i]) pthread_join(th[i], NULL); ok = ok && parts[i].ok; }
    close(fd);
    if (!ok) { dh_host_free(p); return NULL; }
    *size_out = n;
    return p;
}

static char *named_varchar(duckdb_bind_info info, const char *name) {
    duckdb_value v = api.duckdb_bind_get_named_parameter(info, name);
    char *s = NULL;
    if (v && !api.duckdb_is_null_value(v)) s = api.duckdb_get_varchar(v);
    if (v) api.duckdb_destroy_value(&v);
    return s;
}
static int named_bool(duckdb_bind_info info, const char *name, int dflt) {
    duckdb_value v = api.duckdb_bind_get_named_parameter(info, name);
    int r = dflt;
    if (v && !api.duckdb_is_null_value(v)) r = api.duckdb_get_bool(v) ? 1 : 0;
    if (v) api.duckdb_destroy_value(&v);
    return r;
}
static int64_t named_int(duckdb_bind_info info, const char *name, int64_t dflt) {
    duckdb_value v = api.duckdb_bind_get_named_parameter(info, name);
    int64_t r = dflt;
    if (v && !api.duckdb_is_null_value(v)) r = api.duckdb_get_int64(v);
    if (v) api.duckdb_destroy_value(&v);
    return r;
}
static int file_exists(const char *p) { struct stat st; return p && stat(p, &st) == 0; }

/* index lookup like sam_index_load3 / bcf_index_load3(path, index_path, HTS_IDX_SILENT_FAIL) (htslib/hts.c idx_find_and_load):
 * the explicit path, else PATH.csi, PATH with its suffix replaced by .csi[, PATH.bai, PATH with .bai]; n_ext = 2 for BAM, 1 for BCF */
static dh_index *find_index(dh_ctx *ctx, const char *path, const char *index_path, int n_ext) {
    char cand[5][1100];
    int nc = 0;
    dh_index *ix = NULL;
    if (index_path) snprintf(cand[nc++], sizeof cand[0], "%s", index_path);
    else {
        static const char *const EXT[2] = {".csi", ".bai"};
        const char *dot = strrchr(path, '.'), *slash = strrchr(path, '/');
        for (int e = 0; e < n_ext; ++e) {
            snprintf(cand[nc++], sizeof cand[0], "%s%s", path, EXT[e]);
            if (dot && (!slash || dot > slash)) snprintf(cand[nc++], sizeof cand[0], "%.*s%s", (int)(dot - path), path, EXT[e]);
        }
    }
    for (int i = 0; i < nc && !ix; ++i) {
        if (!file_exists(cand[i])) continue;
        uint64_t isz = 0;
        uint8_t *ibuf = slurp_pinned(cand[i], &isz);
        if (ibuf) { ix = dh_index_load(ctx, ibuf, isz); dh_host_free(ibuf); }
        if (!index_path) break;        /* the first file that exists is the one htslib reads */
    }
    return ix;
}

/* ---- helpers shared by the fillers -------------------------------------------------------------------------------------- */
static inline int bit_valid(const uint64_t *v, uint64_t i) { return !v || ((v[i >> 6] >> (i & 63)) & 1); }
/* 1 if rows [r0, r0 + n) are all valid */
static int range_all_valid(const uint64_t *v, uint64_t r0, uint64_t n) {
    if (!v || !n) return 1;
    const uint64_t r1 = r0 + n;
    uint64_t w = r0 >> 6;
    const uint64_t w1 = (r1 - 1) >> 6;
    for (; w <= w1; ++w) {
        uint64_t m = ~0ull;
        if (w == (r0 >> 6)) m &= ~0ull << (r0 & 63);
        if (w == w1 && (r1 & 63)) m &= ~0ull >> (64 - (r1 & 63));
        if ((v[w] & m) != m) return 0;
    }
    return 1;
}
/* validity bits [r0, r0 + n) of a batch column -> rows [0, n) of the vector's mask (word copies) */
static void copy_validity(duckdb_vector vec, const uint64_t *v, uint64_t r0, idx_t n) {
    if (range_all_valid(v, r0, n)) return;
    api.duckdb_vector_ensure_validity_writable(vec);
    uint64_t *dst = api.duckdb_vector_get_validity(vec);
    const unsigned sh = (unsigned)(r0 & 63);
    const uint64_t w0 = r0 >> 6;
    for (idx_t w = 0; w * 64 < n; ++w) {
        uint64_t bits = v[w0 + w] >> sh;
        if (sh && (w * 64 + (64 - sh)) < n) bits |= v[w0 + w + 1] << (64 - sh);
        const idx_t left = n - w * 64;
        if (left < 64) bits |= ~0ull << left;   /* rows past the chunk stay valid */
        dst[w] = bits;
    }
}
static inline void put_string(duckdb_vector vec, duckdb_string_t *out, idx_t row, const char *p, uint32_t len) {
    if (len <= 12) {
        duckdb_string_t c;
        memset(&c, 0, sizeof c);
        c.value.inlined.length = len;
        memcpy(c.value.inlined.inlined, p, len);
        out[row] = c;
    } else api.duckdb_vector_assign_string_element_len(vec, row, p, len);
}
static inline void put_cell(duckdb_vector vec, duckdb_string_t *out, idx_t row, const duckdb_string_t *cell) {
    if (cell->value.inlined.length <= 12) out[row] = *cell;
    else api.duckdb_vector_assign_string_element_len(vec, row, cell->value.pointer.ptr, cell->value.inlined.length);
}
static duckdb_string_t make_const_cell(const char *s) {   /* cell over memory that outlives the query's chunks only if <= 12 bytes: longer ones are copied per row */
    duckdb_string_t c;
    memset(&c, 0, sizeof c);
    const size_t n = s ? strlen(s) : 0;
    c.value.inlined.length = (uint32_t)n;
    if (n <= 12) memcpy(c.value.inlined.inlined, s, n);
    else { memcpy(c.value.pointer.prefix, s, 4); c.value.pointer.ptr = (char *)s; }
    return c;
}

/* ================================================================================================================
 * read_bam
 * ================================================================================================================ */
/* SAMtags columns of standard_tags := true, in the reference's order (src/bam_reader.c:54-70); type letters as there:
 * i BIGINT, Z VARCHAR, A VARCHAR, B LIST (sub-type in the next letter) */
static const char STD_TAGS[] =
    "AMi ASi BCZ BQZ BZZ CBZ CCZ CGBICMi COZ CPi CQZ CRZ CSZ CTZ CYZ E2Z FIi FSZ FZBSH0i H1i H2i HIi IHi LBZ MCZ MDZ MIZ MLBCMMZ MNi "
    "MQi NHi NMi OAZ OCZ OPi OQZ OXZ PGZ PQi PTZ PUZ Q2Z QTZ QXZ R2Z RGZ RXZ SAZ SMi TCi TSA U2Z UQi ";
#define N_STD_TAGS 56
#define N_CORE 13

typedef struct {
    char *path;
    uint8_t *image;
    uint64_t size;
    dh_ctx *ctx;             /* this query's context on the first device */
    dh_bam *bam;
    dh_index *index;         /* BAI/CSI next to the file (or index_path), NULL if none: sam_index_load3, bam_reader.c:499 */
    char *region;            /* the `region` parameter as given (comma separated), NULL if absent or empty */
    int standard_tags, has_index, n_contigs;
    int n_cols;
    int aux_col;             /* schema index of AUXILIARY_TAGS (auxiliary_tags := true), else -1 */
    uint32_t flag_require, flag_exclude, min_mapq;   /* N3 predicate pushdown (extra named parameters; 0 = none) */
} bam_bind_t;

/* a batch shared by the workers that claimed rows of it */
typedef struct {
    dh_bam_batch b;
    int refs, closed;
} bam_batch_ref;

typedef struct {
    dh_ctx *ctx;             /* NULL for the source that uses the bind's context / handle */
    dh_bam *bam;
    int owns;
    pthread_mutex_t mu;
    bam_batch_ref *cur;
    uint64_t cursor;         /* next row of `cur` to hand out */
    int done;
} bam_source_t;

enum { SC_QNAME, SC_QUAL, SC_STR, SC_LIST };
typedef struct { int abi_col, out, kind, nullable; } slot_col_t;

typedef struct {
    bam_bind_t *bind;
    int n_src;
    bam_source_t src[MAX_SOURCES];
    idx_t n_proj;
    idx_t *col_ids;          /* schema column of output vector i */
    int *abi_col;            /* its column in dh_bam_batch, -2 = AUXILIARY_TAGS, -1 = a column this build always returns NULL for */
    int *out_of_abi;         /* inverse: output vector of a batch column, or -1 */
    int compact;
    int n_slot_cols;
    slot_col_t slot_cols[DH_BAM_MAX_COLS + 1];   /* compact layout: string columns in heap-slot order */
    duckdb_string_t *ref_cells;                  /* n_ref + 1 ("*") */
    int n_ref;
    duckdb_string_t *sample_cells;               /* per @RG entry */
    int n_rg;
    int next_worker;         /* atomic */
    int cancel_from;         /* atomic: sources >= this are not read (an earlier one ended in an error / invalid record) */
    char *deferred_err;      /* a scan that could not start: reported by local init, where the reference finds it (bam_reader.c:601-671) */
    /* DUCKHTS_PROFILE=1: where the scan threads spend their time (nanoseconds, summed over threads; printed when the scan is freed) */
    int profile;
    uint64_t ns_lock, ns_acquire, ns_fill, n_chunks, n_acquire, n_busy;
    double t_init;
    char type_of[DH_BAM_MAX_COLS];               /* per batch column: tag type letter, 0 for core columns */
} bam_global_t;

typedef struct { int worker; } bam_local_t;

static bam_global_t *g_last_bam_global;   /* the scan whose local inits follow (the stable API gives local init no handle on the global state) */

static void bam_bind_free(void *p) {
    bam_bind_t *b = (bam_bind_t *)p;
    if (!b) return;
    if (b->bam) dh_bam_close(b->bam);
    if (b->index) dh_index_free(b->index);
    if (b->ctx) ctx_release(b->ctx);
    if (b->image) dh_host_free(b->image);
    free(b->path);
    free(b->region);
    api.duckdb_free(b);
}
static void bam_release_ref(bam_source_t *s, bam_batch_ref *br) {
    dh_bam_release_batch(s->bam, br->b.slot);
    free(br);
}
static double now_sec(void) { struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec + ts.tv_nsec * 1e-9; }
static inline uint64_t now_ns(void) { struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return (uint64_t)ts.tv_sec * 1000000000ull + (uint64_t)ts.tv_nsec; }
static void bam_global_free(void *p) {
    bam_global_t *g = (bam_global_t *)p;
    if (!g) return;
    bam_global_t *expect = g;
    __atomic_compare_exchange_n(&g_last_bam_global, &expect, NULL, 0, __ATOMIC_ACQ_REL, __ATOMIC_ACQUIRE);
    if (g->profile)
        fprintf(stderr, "[duckhts profile] read_bam: %d source(s), %s layout, %llu chunks, %llu batches (%llu busy waits); thread-seconds: lock wait %.3f, acquire %.3f, fill %.3f; "
                "wall since init %.3f s\n", g->n_src, g->compact ? "compact" : "duckdb", (unsigned long long)g->n_chunks, (unsigned long long)g->n_acquire,
                (unsigned long long)g->n_busy, g->ns_lock * 1e-9, g->ns_acquire * 1e-9, g->ns_fill * 1e-9, now_sec() - g->t_init);
    free(g->deferred_err);
    for (int i = 0; i < g->n_src; ++i) {
        bam_source_t *s = &g->src[i];
        if (s->cur) { free(s->cur); s->cur = NULL; }     /* the handle forgets its held slots at the next scan / close */
        if (s->owns) { if (s->bam) dh_bam_close(s->bam); if (s->ctx) ctx_release(s->ctx); }
        pthread_mutex_destroy(&s->mu);
    }
    free(g->col_ids); free(g->abi_col); free(g->out_of_abi); free(g->ref_cells); free(g->sample_cells);
    api.duckdb_free(g);
}

static void read_bam_bind(duckdb_bind_info info) {
    duckdb_value pv = api.duckdb_bind_get_parameter(info, 0);
    char *path = api.duckdb_get_varchar(pv);
    api.duckdb_destroy_value(&pv);
    if (!path || !path[0]) {
        api.duckdb_bind_set_error(info, "read_bam requires a file path");
        if (path) api.duckdb_free(path);
        return;
    }
    char *region = named_varchar(info, "region"), *index_path = named_varchar(info, "index_path"), *reference = named_varchar(info, "reference");
    const int standard_tags = named_bool(info, "standard_tags", 0), auxiliary_tags = named_bool(info, "auxiliary_tags", 0);
    char err[768];
    err[0] = 0;
    bam_bind_t *b = (bam_bind_t *)api.duckdb_malloc(sizeof *b);
    memset(b, 0, sizeof *b);
    int devs[MAX_SOURCES];
    device_list(devs);
    b->ctx = ctx_acquire(devs[0]);
    if (!b->ctx) snprintf(err, sizeof err, "read_bam: no usable CUDA device (%s)", dh_last_error());
    else {
        /* streamed: only the header blocks are read here; the scan reads the file segment by segment (dh_bam_open_file) */
        struct stat stt;
        if (stat(path, &stt) != 0 || !S_ISREG(stt.st_mode) || access(path, R_OK) != 0) snprintf(err, sizeof err, "Failed to open SAM/BAM/CRAM file: %s", path);
        else if (!(b->bam = dh_bam_open_file(b->ctx, path))) snprintf(err, sizeof err, "Failed to read SAM/BAM/CRAM header");
        else b->size = (uint64_t)stt.st_size;
    }
    if (err[0]) {
        api.duckdb_bind_set_error(info, err);
        bam_bind_free(b);
        api.duckdb_free(path);
        if (region) api.duckdb_free(region);
        if (index_path) api.duckdb_free(index_path);
        if (reference) api.duckdb_free(reference);
        return;
    }
    b->path = strdup(path);
    b->standard_tags = standard_tags;
    b->n_contigs = dh_bam_n_ref(b->bam);
    /* N3 (not in the reference's signature; absent = no filter): rows whose FLAG / MAPQ fail are dropped on the GPU */
    b->flag_require = (uint32_t)named_int(info, "flag_require", 0);
    b->flag_exclude = (uint32_t)named_int(info, "flag_exclude", 0);
    b->min_mapq = (uint32_t)named_int(info, "min_mapq", 0);
    if (region && region[0]) b->region = strdup(region);
    b->index = find_index(b->ctx, path, index_path, 2);
    b->has_index = b->index != NULL;
    duckdb_logical_type t_varchar = api.duckdb_create_logical_type(DUCKDB_TYPE_VARCHAR), t_int = api.duckdb_create_logical_type(DUCKDB_TYPE_INTEGER),
                        t_bigint = api.duckdb_create_logical_type(DUCKDB_TYPE_BIGINT), t_usmall = api.duckdb_create_logical_type(DUCKDB_TYPE_USMALLINT),
                        t_double = api.duckdb_create_logical_type(DUCKDB_TYPE_DOUBLE);
    static const struct { const char *name; int kind; } core[N_CORE] = {
        {"QNAME", 0}, {"FLAG", 3}, {"RNAME", 0}, {"POS", 2}, {"MAPQ", 1}, {"CIGAR", 0}, {"RNEXT", 0}, {"PNEXT", 2}, {"TLEN", 2},
        {"SEQ", 0}, {"QUAL", 0}, {"READ_GROUP_ID", 0}, {"SAMPLE_ID", 0}};
    for (int c = 0; c < N_CORE; ++c)
        api.duckdb_bind_add_result_column(info, core[c].name, core[c].kind == 0 ? t_varchar : core[c].kind == 1 ? t_int : core[c].kind == 2 ? t_bigint : t_usmall);
    b->n_cols = N_CORE;
    if (standard_tags) {
        for (int k = 0; k < N_STD_TAGS; ++k) {
            const char *e = STD_TAGS + 4 * k;
            char name[3] = {e[0], e[1], 0};
            if (e[2] == 'B') {
                duckdb_logical_type lt = api.duckdb_create_list_type(e[3] == 'f' ? t_double : t_bigint);
                api.duckdb_bind_add_result_column(info, name, lt);
                api.duckdb_destroy_logical_type(&lt);
            } else api.duckdb_bind_add_result_column(info, name, e[2] == 'i' ? t_bigint : e[2] == 'f' ? t_double : t_varchar);
        }
        b->n_cols += N_STD_TAGS;
    }
    b->aux_col = -1;
    if (auxiliary_tags) {   /* MAP(VARCHAR, VARCHAR) after the typed tag columns (src/bam_reader.c:539-549) */
        duckdb_logical_type mt = api.duckdb_create_map_type(t_varchar, t_varchar);
        b->aux_col = b->n_cols++;
        api.duckdb_bind_add_result_column(info, "AUXILIARY_TAGS", mt);
        api.duckdb_destroy_logical_type(&mt);
    }
    api.duckdb_destroy_logical_type(&t_varchar); api.duckdb_destroy_logical_type(&t_int); api.duckdb_destroy_logical_type(&t_bigint);
    api.duckdb_destroy_logical_type(&t_usmall); api.duckdb_destroy_logical_type(&t_double);
    api.duckdb_bind_set_bind_data(info, b, bam_bind_free);
    api.duckdb_free(path);
    if (region) api.duckdb_free(region);
    if (index_path) api.duckdb_free(index_path);
    if (reference) api.duckdb_free(reference);
}

/* global init: projection pushdown (src/bam_reader.c:676-679 does it per thread; one GPU pipeline serves all threads here),
 * the sources (one per GPU) and which rows each of them scans (src/bam_reader.c:577-585, 601-671, 689-716) */
static void read_bam_init(duckdb_init_info info) {
    bam_bind_t *b = (bam_bind_t *)api.duckdb_init_get_bind_data(info);
    bam_global_t *g = (bam_global_t *)api.duckdb_malloc(sizeof *g);
    memset(g, 0, sizeof *g);
    g->bind = b;
    g->cancel_from = MAX_SOURCES + 1;
    g->profile = getenv("DUCKHTS_PROFILE") != NULL;
    g->t_init = now_sec();
    g->n_proj = api.duckdb_init_get_column_count(info);
    g->col_ids = (idx_t *)calloc(g->n_proj ? g->n_proj : 1, sizeof(idx_t));
    g->abi_col = (int *)calloc(g->n_proj ? g->n_proj : 1, sizeof(int));
    g->out_of_abi = (int *)malloc(sizeof(int) * DH_BAM_MAX_COLS);
    for (int c = 0; c < DH_BAM_MAX_COLS; ++c) g->out_of_abi[c] = -1;
    /* Tag columns are requested from the C-ABI in projection order; tag k of the request is column 13 + k of every batch. */
    uint64_t mask = 0;
    dh_tag_req tags[DH_BAM_MAX_TAGS];
    int n_tags = 0, want_map = 0;
    for (idx_t i = 0; i < g->n_proj; ++i) {
        const idx_t c = api.duckdb_init_get_column_index(info, i);
        g->col_ids[i] = c;
        if (c < N_CORE) { g->abi_col[i] = (int)c; mask |= 1ull << c; continue; }
        if ((int)c == b->aux_col) { g->abi_col[i] = -2; want_map = 1; continue; }
        const char *e = STD_TAGS + 4 * (c - N_CORE);
        if (c - N_CORE >= N_STD_TAGS || n_tags >= DH_BAM_MAX_TAGS) { g->abi_col[i] = -1; continue; }
        tags[n_tags].tag[0] = e[0]; tags[n_tags].tag[1] = e[1]; tags[n_tags].type = e[2]; tags[n_tags].pad = e[2] == 'B' ? e[3] : 0;
        g->abi_col[i] = N_CORE + n_tags;
        g->type_of[N_CORE + n_tags] = e[2];
        if (N_CORE + n_tags < 64) mask |= 1ull << (N_CORE + n_tags);   /* tag columns past bit 63 are always projected */
        ++n_tags;
    }
    /* a column projected twice is filled once per output vector: keep the first for the slot walk, the rest are copies */
    for (idx_t i = 0; i < g->n_proj; ++i)
        if (g->abi_col[i] >= 0 && g->out_of_abi[g->abi_col[i]] < 0) g->out_of_abi[g->abi_col[i]] = (int)i;
    const char *lay = getenv("DUCKHTS_LAYOUT");
    g->compact = !want_map && !(lay && !strcmp(lay, "duckdb"));
    /* the compact fill walks each row's heap slots once, so every output vector needs its own batch column */
    for (idx_t i = 0; i < g->n_proj && g->compact; ++i)
        if (g->abi_col[i] >= 0 && g->out_of_abi[g->abi_col[i]] != (int)i) g->compact = 0;

    char err[1200];
    err[0] = 0;
    char excl[2 * N_STD_TAGS];
    for (int k = 0; k < N_STD_TAGS; ++k) { excl[2 * k] = STD_TAGS[4 * k]; excl[2 * k + 1] = STD_TAGS[4 * k + 1]; }
    /* how many GPUs: a region query stays on one; otherwise one source per listed device while shards stay above
     * DUCKHTS_MIN_SHARD_BYTES of compressed input (default 256 MiB: below that the second GPU's start-up costs more than it saves) */
    int devs[MAX_SOURCES];
    int n_dev = device_list(devs);
    const uint64_t min_shard = env_u64("DUCKHTS_MIN_SHARD_BYTES", 256ull << 20);
    int n_src = 1;
    const int contig_mode = !b->region && b->index && b->n_contigs > 1;
    if (!b->region) {
        n_src = n_dev;
        while (n_src > 1 && b->size / (uint64_t)n_src < min_shard) --n_src;
        if (contig_mode && n_src > b->n_contigs) n_src = b->n_contigs;
    }
    const int n_slots = (int)env_u64("DUCKHTS_SLOTS", 8);
    const uint32_t batch_rows = (uint32_t)env_u64("DUCKHTS_BATCH_ROWS", 128u << 10);
    int32_t tid_bounds[MAX_SOURCES + 1];
    uint64_t voffs[MAX_SOURCES + 1];
    if (n_src > 1) {
        const int rc = contig_mode ? dh_bam_split_contigs(b->bam, b->index, n_src, tid_bounds) : dh_bam_split(b->bam, n_src, voffs);
        if (rc != DH_OK) n_src = 1;
    }
    for (int i = 0; i < n_src && !err[0]; ++i) {
        bam_source_t *s = &g->src[i];
        pthread_mutex_init(&s->mu, NULL);
        g->n_src = i + 1;
        if (i == 0) { s->bam = b->bam; s->ctx = NULL; s->owns = 0; }
        else {
            s->owns = 1;
            s->ctx = ctx_acquire(devs[i]);
            s->bam = s->ctx ? dh_bam_clone(b->bam, s->ctx) : NULL;
            if (!s->bam) { snprintf(err, sizeof err, "read_bam: cannot start a pipeline on device %d: %s", devs[i], dh_last_error()); break; }
        }
        int rc = dh_bam_set_projection(s->bam, mask, tags, n_tags);
        if (rc == DH_OK) rc = dh_bam_set_aux_map(s->bam, want_map, excl, b->standard_tags ? N_STD_TAGS : 0);   /* the MAP leaves out the typed columns (bam_reader.c:976) */
        if (rc == DH_OK) rc = dh_bam_set_mode(s->bam, g->compact ? DH_BAM_COMPACT_OUTPUT : 0u);
        if (rc == DH_OK) rc = dh_bam_set_filter(s->bam, b->flag_require, b->flag_exclude, b->min_mapq);
        if (rc == DH_OK) rc = dh_bam_configure(s->bam, 0, batch_rows, 0);
        if (rc == DH_OK) rc = dh_bam_set_slots(s->bam, n_slots);
        if (rc != DH_OK) { snprintf(err, sizeof err, "%s", dh_last_error()); break; }
        /* which rows (src/bam_reader.c:601-671): region(s) through the index; else, with an index and more than one contig,
         * contig by contig (reads without a contig are then never returned); else the whole file in order */
        if (b->region) {
            if (!b->index) snprintf(err, sizeof err, "Region query requires an index (.bai/.csi/.crai)");
            else {
                char *dup = strdup(b->region), *save = NULL;
                const char *regs[1024];
                int nreg = 0;
                for (char *tok = strtok_r(dup, ",", &save); tok && nreg < 1024; tok = strtok_r(NULL, ",", &save)) regs[nreg++] = tok;
                rc = dh_bam_scan_regions(s->bam, b->index, regs, nreg);
                if (rc == DH_ENOREGION) snprintf(err, sizeof err, "No reads found for region(s): %s", b->region);
                else if (rc != DH_OK) snprintf(err, sizeof err, "%s", dh_last_error());
                free(dup);
            }
        } else if (contig_mode) {
            const int32_t lo = n_src > 1 ? tid_bounds[i] : 0, hi = n_src > 1 ? tid_bounds[i + 1] : b->n_contigs;
            if (dh_bam_scan_contigs(s->bam, b->index, lo, hi) != DH_OK) snprintf(err, sizeof err, "%s", dh_last_error());
        } else {
            const uint64_t lo = n_src > 1 ? voffs[i] : 0, hi = n_src > 1 ? voffs[i + 1] : UINT64_MAX;
            if (lo == UINT64_MAX) { s->done = 1; continue; }             /* more parts than blocks */
            if (dh_bam_scan_begin(s->bam, lo, hi) != DH_OK) snprintf(err, sizeof err, "%s", dh_last_error());
        }
    }
    /* every GPU starts uploading and inflating now, not when a worker first asks it for rows */
    for (int i = 0; i < g->n_src && !err[0]; ++i)
        if (!g->src[i].done && dh_bam_prefetch(g->src[i].bam) != DH_OK) snprintf(err, sizeof err, "%s", dh_last_error());
    if (err[0]) {
        /* the reference meets these in local init, after DuckDB has the schema: keep the scan state, hand the message to the workers */
        g->deferred_err = strdup(err);
        for (int i = 0; i < g->n_src; ++i) g->src[i].done = 1;
    }
    /* compact layout: what the fillers need besides the batch */
    if (g->compact) {
        g->n_ref = b->n_contigs;
        g->ref_cells = (duckdb_string_t *)calloc((size_t)g->n_ref + 1, sizeof(duckdb_string_t));
        for (int t = 0; t < g->n_ref; ++t) g->ref_cells[t] = make_const_cell(dh_bam_ref_name(b->bam, t));
        g->ref_cells[g->n_ref] = make_const_cell("*");
        g->n_rg = dh_bam_n_rg(b->bam);
        g->sample_cells = (duckdb_string_t *)calloc((size_t)g->n_rg + 1, sizeof(duckdb_string_t));
        for (int r = 0; r < g->n_rg; ++r) g->sample_cells[r] = make_const_cell(dh_bam_rg_sample(b->bam, r));
        /* heap-slot order of a row: QNAME, CIGAR, SEQ, QUAL, 'Z' / 'B' tag columns in request order, READ_GROUP_ID */
        static const int CORE_SLOTS[4] = {0, 5, 9, 10};
        int ns = 0;
        for (int q = 0; q < 4; ++q)
            if (g->out_of_abi[CORE_SLOTS[q]] >= 0) {
                slot_col_t sc = {CORE_SLOTS[q], g->out_of_abi[CORE_SLOTS[q]], q == 0 ? SC_QNAME : q == 3 ? SC_QUAL : SC_STR, 0};
                g->slot_cols[ns++] = sc;
            }
        for (int k = 0; k < n_tags; ++k) {
            const char ty = g->type_of[N_CORE + k];
            if (ty != 'Z' && ty != 'B') continue;
            slot_col_t sc = {N_CORE + k, g->out_of_abi[N_CORE + k], ty == 'B' ? SC_LIST : SC_STR, 1};
            g->slot_cols[ns++] = sc;
        }
        if (g->out_of_abi[11] >= 0) { slot_col_t sc = {11, g->out_of_abi[11], SC_STR, 1}; g->slot_cols[ns++] = sc; }
        g->n_slot_cols = ns;
    }
    api.duckdb_init_set_max_threads(info, 64);
    api.duckdb_init_set_init_data(info, g, bam_global_free);
    __atomic_store_n(&g_last_bam_global, g, __ATOMIC_RELEASE);
}

static void read_bam_local_init(duckdb_init_info info) {
    bam_bind_t *b = (bam_bind_t *)api.duckdb_init_get_bind_data(info);
    bam_global_t *g = __atomic_load_n(&g_last_bam_global, __ATOMIC_ACQUIRE);
    if (g && g->bind == b && g->deferred_err) { api.duckdb_init_set_error(info, g->deferred_err); return; }
    bam_local_t *l = (bam_local_t *)api.duckdb_malloc(sizeof *l);
    l->worker = -1;     /* numbered at the first scan call */
    api.duckdb_init_set_init_data(info, l, api.duckdb_free);
}

static size_t core_width(idx_t c) { return c == 1 ? 2 : c == 4 ? 4 : (c == 3 || c == 7 || c == 8) ? 8 : 16; }

/* rows [r0, r0 + n) of a DuckDB-layout batch -> the output chunk */
static void fill_duckdb_layout(bam_global_t *g, const dh_bam_batch *bt, uint64_t r0, idx_t n, duckdb_data_chunk output) {
    for (idx_t i = 0; i < g->n_proj; ++i) {
        duckdb_vector vec = api.duckdb_data_chunk_get_vector(output, i);
        const int ac = g->abi_col[i];
        if (ac == -2) {   /* AUXILIARY_TAGS: list entries re-based onto the vector's child, key / value cells appended */
            const duckdb_list_entry *cells = (const duckdb_list_entry *)bt->aux_map + r0;
            const uint64_t *valid = bt->aux_map_validity;
            duckdb_list_entry *out = (duckdb_list_entry *)api.duckdb_vector_get_data(vec);
            idx_t total = api.duckdb_list_vector_get_size(vec), need = total;
            for (idx_t r = 0; r < n; ++r) need += cells[r].length;
            api.duckdb_list_vector_reserve(vec, need);
            api.duckdb_list_vector_set_size(vec, need);
            duckdb_vector child = api.duckdb_list_vector_get_child(vec);
            duckdb_vector kv[2] = {api.duckdb_struct_vector_get_child(child, 0), api.duckdb_struct_vector_get_child(child, 1)};
            const duckdb_string_t *src[2] = {(const duckdb_string_t *)bt->aux_keys, (const duckdb_string_t *)bt->aux_values};
            duckdb_string_t *dst[2] = {(duckdb_string_t *)api.duckdb_vector_get_data(kv[0]), (duckdb_string_t *)api.duckdb_vector_get_data(kv[1])};
            for (idx_t r = 0; r < n; ++r) {
                out[r].offset = total;
                out[r].length = cells[r].length;
                if (!bit_valid(valid, r0 + r)) continue;
                for (int q = 0; q < 2; ++q)
                    for (uint64_t e = 0; e < cells[r].length; ++e) put_cell(kv[q], dst[q], total + e, &src[q][cells[r].offset + e]);
                total += cells[r].length;
            }
            copy_validity(vec, valid, r0, n);
            continue;
        }
        if (ac < 0) {   /* column not served: all NULL */
            api.duckdb_vector_ensure_validity_writable(vec);
            uint64_t *val = api.duckdb_vector_get_validity(vec);
            for (idx_t r = 0; r < n; ++r) api.duckdb_validity_set_row_invalid(val, r);
            if (STD_TAGS[4 * (g->col_ids[i] - N_CORE) + 2] == 'B') {
                duckdb_list_entry *le = (duckdb_list_entry *)api.duckdb_vector_get_data(vec);
                for (idx_t r = 0; r < n; ++r) { le[r].offset = 0; le[r].length = 0; }
            }
            continue;
        }
        const idx_t c = g->col_ids[i];
        const char ty = g->type_of[ac];
        const size_t w = c < N_CORE ? core_width(c) : (ty == 'i' || ty == 'f') ? 8 : 16;
        const uint8_t *src = (const uint8_t *)bt->data[ac] + r0 * w;
        void *dst = api.duckdb_vector_get_data(vec);
        const uint64_t *valid = bt->validity[ac];
        if (ty == 'B') {
            /* LIST cells (bam_assign_list_int/_double, src/bam_reader.c:106-138): children are appended to the list child */
            const duckdb_list_entry *cells = (const duckdb_list_entry *)src;
            duckdb_list_entry *out = (duckdb_list_entry *)dst;
            idx_t total = api.duckdb_list_vector_get_size(vec), need = total;
            for (idx_t r = 0; r < n; ++r)
                if (bit_valid(valid, r0 + r)) need += cells[r].length;
            api.duckdb_list_vector_reserve(vec, need);
            api.duckdb_list_vector_set_size(vec, need);
            uint64_t *child = (uint64_t *)api.duckdb_vector_get_data(api.duckdb_list_vector_get_child(vec));
            for (idx_t r = 0; r < n; ++r) {
                if (!bit_valid(valid, r0 + r)) continue;
                memcpy(child + total, (const uint64_t *)bt->heap + cells[r].offset, 8 * cells[r].length);
                out[r].offset = total;
                out[r].length = cells[r].length;
                total += cells[r].length;
            }
        } else if (w != 16) memcpy(dst, src, n * w);
        else {
            const duckdb_string_t *cells = (const duckdb_string_t *)src;
            duckdb_string_t *out = (duckdb_string_t *)dst;
            if (range_all_valid(valid, r0, n)) for (idx_t r = 0; r < n; ++r) put_cell(vec, out, r, &cells[r]);
            else for (idx_t r = 0; r < n; ++r) if (bit_valid(valid, r0 + r)) put_cell(vec, out, r, &cells[r]);
        }
        copy_validity(vec, valid, r0, n);
    }
}

/* rows [r0, r0 + n) of a compact-layout batch -> the output chunk (the expansion the GPU left to the host: widening, dictionary
 * cells, one cell per string; include/duckhts_cuda.h DH_BAM_COMPACT_OUTPUT) */
static void fill_compact_layout(bam_global_t *g, const dh_bam_batch *bt, uint64_t r0, idx_t n, duckdb_data_chunk output) {
    /* fixed-width, dictionary and typed-scalar columns, one column at a time */
    for (idx_t i = 0; i < g->n_proj; ++i) {
        const int ac = g->abi_col[i];
        duckdb_vector vec = api.duckdb_data_chunk_get_vector(output, i);
        if (ac < 0) {   /* column not served: all NULL */
            api.duckdb_vector_ensure_validity_writable(vec);
            uint64_t *val = api.duckdb_vector_get_validity(vec);
            for (idx_t r = 0; r < n; ++r) api.duckdb_validity_set_row_invalid(val, r);
            if (STD_TAGS[4 * (g->col_ids[i] - N_CORE) + 2] == 'B') {
                duckdb_list_entry *le = (duckdb_list_entry *)api.duckdb_vector_get_data(vec);
                for (idx_t r = 0; r < n; ++r) { le[r].offset = 0; le[r].length = 0; }
            }
            continue;
        }
        void *dst = api.duckdb_vector_get_data(vec);
        const uint64_t *valid = bt->validity[ac];
        switch (ac) {
        case 1: memcpy(dst, (const uint16_t *)bt->data[ac] + r0, n * 2); break;                                   /* FLAG */
        case 4: { const uint8_t *s = (const uint8_t *)bt->data[ac] + r0; int32_t *d = (int32_t *)dst; for (idx_t r = 0; r < n; ++r) d[r] = s[r]; break; }
        case 3: case 7: { const int32_t *s = (const int32_t *)bt->data[ac] + r0; int64_t *d = (int64_t *)dst; for (idx_t r = 0; r < n; ++r) d[r] = (int64_t)s[r] + 1; break; }   /* POS = pos + 1 (bam_reader.c:803), PNEXT */
        case 8: { const int32_t *s = (const int32_t *)bt->data[ac] + r0; int64_t *d = (int64_t *)dst; for (idx_t r = 0; r < n; ++r) d[r] = s[r]; break; }
        case 2: case 6: {                                                                                          /* RNAME / RNEXT: sam_hdr_tid2name or "*" */
            const int32_t *s = (const int32_t *)bt->data[ac] + r0;
            duckdb_string_t *d = (duckdb_string_t *)dst;
            for (idx_t r = 0; r < n; ++r) put_cell(vec, d, r, &g->ref_cells[(s[r] >= 0 && s[r] < g->n_ref) ? s[r] : g->n_ref]);
            break;
        }
        case 12: {                                                                                                 /* SAMPLE_ID: SM of the read's @RG line */
            const int32_t *s = (const int32_t *)bt->data[ac] + r0;
            duckdb_string_t *d = (duckdb_string_t *)dst;
            for (idx_t r = 0; r < n; ++r)
                if (bit_valid(valid, r0 + r) && s[r] >= 0 && s[r] < g->n_rg) put_cell(vec, d, r, &g->sample_cells[s[r]]);
            copy_validity(vec, valid, r0, n);
            break;
        }
        default: {
            const char ty = g->type_of[ac];
            if (ty == 'i' || ty == 'f') { memcpy(dst, (const uint64_t *)bt->data[ac] + r0, n * 8); copy_validity(vec, valid, r0, n); }
            else if (ty == 'A') {
                const uint8_t *s = (const uint8_t *)bt->data[ac] + r0;
                duckdb_string_t *d = (duckdb_string_t *)dst;
                for (idx_t r = 0; r < n; ++r) { memset(&d[r], 0, sizeof d[r]); if (s[r]) { d[r].value.inlined.length = 1; d[r].value.inlined.inlined[0] = (char)s[r]; } }
                copy_validity(vec, valid, r0, n);
            }
            break;   /* string columns: the slot walk below */
        }
        }
    }
    /* string columns: one walk over each row's heap slots */
    const int ns = g->n_slot_cols;
    if (!ns) return;
    duckdb_vector vecs[DH_BAM_MAX_COLS + 1];
    void *outp[DH_BAM_MAX_COLS + 1];
    const uint32_t *lens[DH_BAM_MAX_COLS + 1];
    const uint64_t *valids[DH_BAM_MAX_COLS + 1];
    idx_t list_total[DH_BAM_MAX_COLS + 1];
    uint64_t *list_child[DH_BAM_MAX_COLS + 1];
    for (int q = 0; q < ns; ++q) {
        const slot_col_t *sc = &g->slot_cols[q];
        vecs[q] = api.duckdb_data_chunk_get_vector(output, (idx_t)sc->out);
        outp[q] = api.duckdb_vector_get_data(vecs[q]);
        lens[q] = (const uint32_t *)bt->data[sc->abi_col];
        valids[q] = sc->nullable ? bt->validity[sc->abi_col] : NULL;
        list_total[q] = 0;
        list_child[q] = NULL;
        if (sc->kind == SC_LIST) {   /* one reserve per chunk: the children of the chunk's cells */
            const duckdb_list_entry *cells = (const duckdb_list_entry *)bt->data[sc->abi_col] + r0;
            idx_t total = api.duckdb_list_vector_get_size(vecs[q]), need = total;
            for (idx_t r = 0; r < n; ++r) if (bit_valid(valids[q], r0 + r)) need += cells[r].length;
            api.duckdb_list_vector_reserve(vecs[q], need);
            api.duckdb_list_vector_set_size(vecs[q], need);
            list_total[q] = total;
            list_child[q] = (uint64_t *)api.duckdb_vector_get_data(api.duckdb_list_vector_get_child(vecs[q]));
        }
    }
    const uint8_t *heap = bt->heap;
    for (idx_t r = 0; r < n; ++r) {
        const uint64_t row = r0 + r;
        uint64_t off = bt->rec_off[row];
        for (int q = 0; q < ns; ++q) {
            const slot_col_t *sc = &g->slot_cols[q];
            if (valids[q] && !bit_valid(valids[q], row)) {
                if (sc->kind == SC_LIST) { duckdb_list_entry *o = (duckdb_list_entry *)outp[q]; o[r].offset = list_total[q]; o[r].length = 0; }
                continue;
            }
            const char *p = (const char *)heap + off;
            if (sc->kind == SC_LIST) {
                const duckdb_list_entry e = ((const duckdb_list_entry *)bt->data[sc->abi_col])[row];
                duckdb_list_entry *o = (duckdb_list_entry *)outp[q];
                memcpy(list_child[q] + list_total[q], (const uint64_t *)heap + e.offset, 8 * e.length);
                o[r].offset = list_total[q];
                o[r].length = e.length;
                list_total[q] += e.length;
                off += 4 + 8 * e.length;
                continue;
            }
            const uint32_t v = lens[q][row], cap = v & 0x7fffffffu;
            uint32_t len = cap;
            if (sc->kind == SC_QNAME) len = (v >> 31) ? (uint32_t)strnlen(p, cap) : cap - 1;      /* C-string semantics */
            else if (v >> 31) len = (uint32_t)strnlen(p, cap);                                   /* QUAL cut at a byte that wrapped to NUL */
            put_string(vecs[q], (duckdb_string_t *)outp[q], r, p, len);
            off += (cap + 3u) & ~3u;
        }
    }
    for (int q = 0; q < ns; ++q)
        if (valids[q]) copy_validity(vecs[q], valids[q], r0, n);
}

static int pick_source(bam_global_t *g, int worker) {
    const int cancel = __atomic_load_n(&g->cancel_from, __ATOMIC_ACQUIRE);
    const int pref = worker % g->n_src;
    if (pref < cancel && !__atomic_load_n(&g->src[pref].done, __ATOMIC_ACQUIRE)) return pref;
    for (int s = 0; s < g->n_src && s < cancel; ++s)
        if (!__atomic_load_n(&g->src[s].done, __ATOMIC_ACQUIRE)) return s;
    return -1;
}

static void read_bam_scan(duckdb_function_info info, duckdb_data_chunk output) {
    bam_global_t *g = (bam_global_t *)api.duckdb_function_get_init_data(info);
    bam_local_t *l = (bam_local_t *)api.duckdb_function_get_local_init_data(info);
    if (!g || !l) { api.duckdb_data_chunk_set_size(output, 0); return; }
    if (l->worker < 0) l->worker = __atomic_fetch_add(&g->next_worker, 1, __ATOMIC_RELAXED);
    const idx_t vs = api.duckdb_vector_size();
    for (;;) {
        const int si = pick_source(g, l->worker);
        if (si < 0) { api.duckdb_data_chunk_set_size(output, 0); return; }
        bam_source_t *s = &g->src[si];
        const uint64_t t_lock = g->profile ? now_ns() : 0;
        pthread_mutex_lock(&s->mu);
        if (g->profile) __atomic_fetch_add(&g->ns_lock, now_ns() - t_lock, __ATOMIC_RELAXED);
        if (s->done) { pthread_mutex_unlock(&s->mu); continue; }
        if (!s->cur || s->cursor >= s->cur->b.n_rows) {
            if (s->cur) {   /* every row of it is claimed: it goes back when the last claim is filled */
                s->cur->closed = 1;
                if (s->cur->refs == 0) bam_release_ref(s, s->cur);
                s->cur = NULL;
            }
            bam_batch_ref *br = (bam_batch_ref *)calloc(1, sizeof *br);
            const uint64_t t_acq = g->profile ? now_ns() : 0;
            const int rc = dh_bam_acquire_batch(s->bam, &br->b);
            if (g->profile) { __atomic_fetch_add(&g->ns_acquire, now_ns() - t_acq, __ATOMIC_RELAXED); __atomic_fetch_add(rc == DH_EBUSY ? &g->n_busy : &g->n_acquire, 1, __ATOMIC_RELAXED); }
            if (rc == 1) { s->cur = br; s->cursor = 0; }
            else {
                free(br);
                if (rc == DH_EBUSY) { pthread_mutex_unlock(&s->mu); usleep(100); continue; }   /* every slot is being filled from: wait for a release */
                __atomic_store_n(&s->done, 1, __ATOMIC_RELEASE);
                /* rc == 0: end of this source's rows — or an invalid record ended them: the reference stops silently there
                 * (src/bam_reader.c:754-765), and so must everything behind it in file order.  A failed BGZF block (DH_EDATA) is
                 * a read error to htslib = the same silent end.  Anything else (CUDA failure, a pipeline limit) is reported. */
                dh_bam_scan_stats st;
                memset(&st, 0, sizeof st);
                dh_bam_scan_stats_get(s->bam, &st);
                if (rc == DH_EDATA || (rc == 0 && st.status != 0)) {
                    int cur = __atomic_load_n(&g->cancel_from, __ATOMIC_ACQUIRE);
                    while (si + 1 < cur && !__atomic_compare_exchange_n(&g->cancel_from, &cur, si + 1, 0, __ATOMIC_ACQ_REL, __ATOMIC_ACQUIRE)) {}
                }
                pthread_mutex_unlock(&s->mu);
                if (rc < 0 && rc != DH_EDATA) {
                    api.duckdb_function_set_error(info, dh_last_error());
                    api.duckdb_data_chunk_set_size(output, 0);
                    return;
                }
                continue;
            }
        }
        bam_batch_ref *br = s->cur;
        const uint64_t r0 = s->cursor;
        const idx_t n = br->b.n_rows - r0 < vs ? (idx_t)(br->b.n_rows - r0) : vs;
        s->cursor += n;
        br->refs++;
        pthread_mutex_unlock(&s->mu);

        const uint64_t t_fill = g->profile ? now_ns() : 0;
        if (br->b.layout == DH_BAM_COMPACT_OUTPUT) fill_compact_layout(g, &br->b, r0, n, output);
        else fill_duckdb_layout(g, &br->b, r0, n, output);
        if (g->profile) { __atomic_fetch_add(&g->ns_fill, now_ns() - t_fill, __ATOMIC_RELAXED); __atomic_fetch_add(&g->n_chunks, 1, __ATOMIC_RELAXED); }

        pthread_mutex_lock(&s->mu);
        if (--br->refs == 0 && br->closed) bam_release_ref(s, br);
        pthread_mutex_unlock(&s->mu);
        api.duckdb_data_chunk_set_size(output, n);
        return;
    }
}

static void register_read_bam(duckdb_connection con) {
    duckdb_table_function tf = api.duckdb_create_table_function();
    duckdb_logical_type t_varchar = api.duckdb_create_logical_type(DUCKDB_TYPE_VARCHAR), t_bool = api.duckdb_create_logical_type(DUCKDB_TYPE_BOOLEAN),
                        t_int = api.duckdb_create_logical_type(DUCKDB_TYPE_INTEGER);
    api.duckdb_table_function_set_name(tf, "read_bam");
    api.duckdb_table_function_add_parameter(tf, t_varchar);
    api.duckdb_table_function_add_named_parameter(tf, "region", t_varchar);
    api.duckdb_table_function_add_named_parameter(tf, "index_path", t_varchar);
    api.duckdb_table_function_add_named_parameter(tf, "reference", t_varchar);
    api.duckdb_table_function_add_named_parameter(tf, "standard_tags", t_bool);
    api.duckdb_table_function_add_named_parameter(tf, "auxiliary_tags", t_bool);
    /* N3, additions to the reference's signature: predicates evaluated on the GPU (dh_bam_set_filter) */
    api.duckdb_table_function_add_named_parameter(tf, "flag_require", t_int);
    api.duckdb_table_function_add_named_parameter(tf, "flag_exclude", t_int);
    api.duckdb_table_function_add_named_parameter(tf, "min_mapq", t_int);
    api.duckdb_table_function_set_bind(tf, read_bam_bind);
    api.duckdb_table_function_set_init(tf, read_bam_init);
    api.duckdb_table_function_set_local_init(tf, read_bam_local_init);
    api.duckdb_table_function_set_function(tf, read_bam_scan);
    api.duckdb_table_function_supports_projection_pushdown(tf, true);
    api.duckdb_register_table_function(con, tf);
    api.duckdb_destroy_table_function(&tf);
    api.duckdb_destroy_logical_type(&t_varchar);
    api.duckdb_destroy_logical_type(&t_bool);
    api.duckdb_destroy_logical_type(&t_int);
}

/* ================================================================================================================
 * read_bcf — same parameters, schema and NULL rules as the reference (src/bcf_reader.c:451-830, 2055-2081); rows come
 * from dh_bcf batches: site-level cells once per record, FORMAT cells once per (record, sample).  tidy_format := true
 * repeats the site-level cell for each sample row here, on the host, 16 bytes at a time.
 * ================================================================================================================ */
typedef struct {
    uint8_t *image;
    uint64_t size;
    dh_ctx *ctx;             /* this query's context */
    dh_bcf *bcf;
    int tidy, n_samples, n_info, n_fmt, default_gt;
    int info_start, sample_id_col, fmt_start, n_cols;
    dh_bcf_field *info, *fmt;
    dh_index *index;         /* CSI, loaded when a region is given (bcf_reader.c:902-915) */
    char *region;            /* the parameter as given */
    char **regions;          /* comma-separated, blanks trimmed, empty tokens dropped (parse_regions_duckdb, bcf_reader.c:286-329) */
    int n_regions;
    duckdb_string_t *sample_cells;   /* SAMPLE_ID cells of the tidy rows, one per sample */
} bcf_bind_t;

typedef struct {
    dh_bcf_batch b;
    int refs, closed;
} bcf_batch_ref;

/* the scan all workers of a query share: one GPU pipeline, rows handed out chunk by chunk */
typedef struct {
    bcf_bind_t *bind;
    idx_t n_proj;
    idx_t *col_ids;
    pthread_mutex_t mu;
    bcf_batch_ref *cur;
    uint64_t cursor;        /* next row of the batch: records (wide) or records x samples (tidy) */
    int done;
    int next_region;        /* region cursor of a chained region scan (bcf_reader.c:172, 1318-1333) */
    char *deferred_err;     /* a scan that could not start: reported by local init, where the reference finds it */
} bcf_global_t;
static bcf_global_t *g_last_bcf_global;

/* start the scan of the next region that yields an iterator: 1 started, 0 none left, < 0 error */
static int bcf_next_region(bcf_bind_t *b, bcf_global_t *l) {
    while (l->next_region < b->n_regions) {
        const int rc = dh_bcf_scan_region(b->bcf, b->index, b->regions[l->next_region++]);
        if (rc == DH_OK) return 1;
        if (rc != DH_ENOREGION) return rc;      /* unknown contig: "Region query returned no iterator; skipping region" */
    }
    return 0;
}

static void bcf_bind_free(void *p) {
    bcf_bind_t *b = (bcf_bind_t *)p;
    if (!b) return;
    if (b->bcf) dh_bcf_close(b->bcf);
    if (b->index) dh_index_free(b->index);
    if (b->ctx) ctx_release(b->ctx);
    if (b->image) dh_host_free(b->image);
    free(b->region);
    for (int i = 0; i < b->n_regions; ++i) free(b->regions[i]);
    free(b->regions);
    free(b->info);
    free(b->fmt);
    free(b->sample_cells);
    api.duckdb_free(b);
}
static void bcf_global_free(void *p) {
    bcf_global_t *l = (bcf_global_t *)p;
    if (!l) return;
    bcf_global_t *expect = l;
    __atomic_compare_exchange_n(&g_last_bcf_global, &expect, NULL, 0, __ATOMIC_ACQ_REL, __ATOMIC_ACQUIRE);
    free(l->cur);
    free(l->col_ids);
    free(l->deferred_err);
    pthread_mutex_destroy(&l->mu);
    api.duckdb_free(l);
}

static duckdb_logical_type bcf_field_type(int type, int is_list) {   /* create_bcf_field_type, bcf_reader.c:391-417 */
    duckdb_logical_type e = api.duckdb_create_logical_type(type == 0 ? DUCKDB_TYPE_BOOLEAN : type == 1 ? DUCKDB_TYPE_INTEGER : type == 2 ? DUCKDB_TYPE_FLOAT : DUCKDB_TYPE_VARCHAR);
    if (!is_list) return e;
    duckdb_logical_type l = api.duckdb_create_list_type(e);
    api.duckdb_destroy_logical_type(&e);
    return l;
}

static void read_bcf_bind(duckdb_bind_info info) {
    duckdb_value pv = api.duckdb_bind_get_parameter(info, 0);
    char *path = api.duckdb_get_varchar(pv);
    api.duckdb_destroy_value(&pv);
    if (!path || !path[0]) {
        api.duckdb_bind_set_error(info, "read_bcf requires a file path");
        if (path) api.duckdb_free(path);
        return;
    }
    char *region = named_varchar(info, "region"), *index_path = named_varchar(info, "index_path");
    const int tidy = named_bool(info, "tidy_format", 0);
    char err[768];
    err[0] = 0;
    bcf_bind_t *b = (bcf_bind_t *)api.duckdb_malloc(sizeof *b);
    memset(b, 0, sizeof *b);
    int devs[MAX_SOURCES];
    device_list(devs);
    dh_ctx *ctx = b->ctx = ctx_acquire(devs[0]);
    if (!ctx) snprintf(err, sizeof err, "read_bcf: no usable CUDA device (%s)", dh_last_error());
    else {
        if (region && region[0]) {
            b->region = strdup(region);
            b->regions = (char **)calloc(strlen(region) + 1, sizeof(char *));
            char *dup = strdup(region), *save = NULL;
            for (char *tok = strtok_r(dup, ",", &save); tok; tok = strtok_r(NULL, ",", &save)) {
                while (*tok == ' ' || *tok == '\t') ++tok;
                size_t len = strlen(tok);
                while (len > 0 && (tok[len - 1] == ' ' || tok[len - 1] == '\t')) tok[--len] = 0;
                if (len) b->regions[b->n_regions++] = strdup(tok);
            }
            free(dup);
            if (b->n_regions) b->index = find_index(ctx, path, index_path, 1);
        }
        b->image = slurp_pinned(path, &b->size);
        if (!b->image) snprintf(err, sizeof err, "Failed to open BCF/VCF file: %s", path);
        else if (!(b->bcf = dh_bcf_open(ctx, b->image, b->size))) snprintf(err, sizeof err, "Failed to read BCF/VCF header");
        else if (dh_bcf_has_vep(b->bcf)) snprintf(err, sizeof err, "read_bcf: VEP_* columns (CSQ/BCSQ/ANN) are not available in this build");
    }
    if (err[0]) {
        api.duckdb_bind_set_error(info, err);
        bcf_bind_free(b);
    } else {
        b->tidy = tidy;
        b->n_samples = dh_bcf_n_samples(b->bcf);
        b->n_info = dh_bcf_n_info(b->bcf);
        b->n_fmt = dh_bcf_n_format(b->bcf);
        b->info = (dh_bcf_field *)calloc((size_t)b->n_info + 1, sizeof *b->info);
        b->fmt = (dh_bcf_field *)calloc((size_t)b->n_fmt + 1, sizeof *b->fmt);
        for (int i = 0; i < b->n_info; ++i) dh_bcf_info_field(b->bcf, i, &b->info[i]);
        for (int i = 0; i < b->n_fmt; ++i) dh_bcf_format_field(b->bcf, i, &b->fmt[i]);
        if (b->n_samples > 0 && b->n_fmt == 0) {      /* "Add GT as default" (bcf_reader.c:669-678): a VARCHAR column that is always NULL */
            b->default_gt = 1;
            b->n_fmt = 1;
            b->fmt[0].name = "GT"; b->fmt[0].type = 3; b->fmt[0].is_list = 0;
        }
        duckdb_logical_type t_varchar = api.duckdb_create_logical_type(DUCKDB_TYPE_VARCHAR), t_bigint = api.duckdb_create_logical_type(DUCKDB_TYPE_BIGINT),
                            t_double = api.duckdb_create_logical_type(DUCKDB_TYPE_DOUBLE);
        duckdb_logical_type t_vlist = api.duckdb_create_list_type(t_varchar);
        api.duckdb_bind_add_result_column(info, "CHROM", t_varchar);
        api.duckdb_bind_add_result_column(info, "POS", t_bigint);
        api.duckdb_bind_add_result_column(info, "ID", t_varchar);
        api.duckdb_bind_add_result_column(info, "REF", t_varchar);
        api.duckdb_bind_add_result_column(info, "ALT", t_vlist);
        api.duckdb_bind_add_result_column(info, "QUAL", t_double);
        api.duckdb_bind_add_result_column(info, "FILTER", t_vlist);
        int col = 7;
        char name[600];
        b->info_start = col;
        for (int i = 0; i < b->n_info; ++i, ++col) {
            snprintf(name, sizeof name, "INFO_%s", b->info[i].name);
            duckdb_logical_type t = bcf_field_type(b->info[i].type, b->info[i].is_list);
            api.duckdb_bind_add_result_column(info, name, t);
            api.duckdb_destroy_logical_type(&t);
        }
        b->sample_id_col = -1;
        b->fmt_start = col;
        if (b->n_samples > 0) {
            if (tidy) {
                b->sample_id_col = col++;
                api.duckdb_bind_add_result_column(info, "SAMPLE_ID", t_varchar);
                b->fmt_start = col;
                for (int f = 0; f < b->n_fmt; ++f, ++col) {
                    snprintf(name, sizeof name, "FORMAT_%s", b->fmt[f].name);
                    duckdb_logical_type t = bcf_field_type(b->fmt[f].type, b->fmt[f].is_list);
                    api.duckdb_bind_add_result_column(info, name, t);
                    api.duckdb_destroy_logical_type(&t);
                }
            } else {
                for (int sidx = 0; sidx < b->n_samples; ++sidx)
                    for (int f = 0; f < b->n_fmt; ++f, ++col) {
                        snprintf(name, sizeof name, "FORMAT_%s_%s", b->fmt[f].name, dh_bcf_sample_name(b->bcf, sidx));
                        duckdb_logical_type t = bcf_field_type(b->fmt[f].type, b->fmt[f].is_list);
                        api.duckdb_bind_add_result_column(info, name, t);
                        api.duckdb_destroy_logical_type(&t);
                    }
            }
        }
        b->n_cols = col;
        b->sample_cells = (duckdb_string_t *)calloc((size_t)b->n_samples + 1, sizeof(duckdb_string_t));
        for (int sidx = 0; sidx < b->n_samples; ++sidx) b->sample_cells[sidx] = make_const_cell(dh_bcf_sample_name(b->bcf, sidx));
        api.duckdb_destroy_logical_type(&t_varchar); api.duckdb_destroy_logical_type(&t_bigint); api.duckdb_destroy_logical_type(&t_double);
        api.duckdb_destroy_logical_type(&t_vlist);
        api.duckdb_bind_set_bind_data(info, b, bcf_bind_free);
    }
    api.duckdb_free(path);
    if (region) api.duckdb_free(region);
    if (index_path) api.duckdb_free(index_path);
}

/* global init: projection pushdown and the start of the scan; max_threads = 64 workers share the one pipeline */
static void read_bcf_init(duckdb_init_info info) {
    bcf_bind_t *b = (bcf_bind_t *)api.duckdb_init_get_bind_data(info);
    bcf_global_t *l = (bcf_global_t *)api.duckdb_malloc(sizeof *l);
    memset(l, 0, sizeof *l);
    pthread_mutex_init(&l->mu, NULL);
    l->bind = b;
    l->n_proj = api.duckdb_init_get_column_count(info);
    l->col_ids = (idx_t *)calloc(l->n_proj ? l->n_proj : 1, sizeof(idx_t));
    uint32_t core_mask = 0;
    uint8_t *iw = (uint8_t *)calloc((size_t)b->n_info + 1, 1), *fw = (uint8_t *)calloc((size_t)b->n_fmt + 1, 1);
    for (idx_t i = 0; i < l->n_proj; ++i) {
        const idx_t c = api.duckdb_init_get_column_index(info, i);
        l->col_ids[i] = c;
        if (c < 7) core_mask |= 1u << c;
        else if ((int)c < b->info_start + b->n_info) iw[c - (idx_t)b->info_start] = 1;
        else if ((int)c >= b->fmt_start && (int)c != b->sample_id_col) fw[(c - (idx_t)b->fmt_start) % (idx_t)b->n_fmt] = 1;
    }
    int rc = dh_bcf_set_projection(b->bcf, core_mask, iw, b->default_gt ? NULL : fw);
    free(iw);
    free(fw);
    if (rc == DH_OK) rc = dh_bcf_set_slots(b->bcf, (int)env_u64("DUCKHTS_SLOTS", 8));
    {   /* FORMAT cells cross PCIe packed unless DUCKHTS_LAYOUT=duckdb asks for ready-made vectors */
        const char *lay = getenv("DUCKHTS_LAYOUT");
        if (rc == DH_OK) rc = dh_bcf_set_mode(b->bcf, (lay && !strcmp(lay, "duckdb")) ? 0u : DH_BAM_COMPACT_OUTPUT);
    }
    char err[600];
    err[0] = 0;
    if (rc == DH_OK && b->n_regions > 0) {
        if (!b->index) snprintf(err, sizeof err, "Region query requires an index file (.tbi or .csi). Region: %s", b->region);
        else {
            const int started = bcf_next_region(b, l);
            if (started < 0) snprintf(err, sizeof err, "%s", dh_last_error());
            else if (!started) l->done = 1;     /* no region named a contig: an empty result, not an error (bcf_reader.c:952-956) */
        }
    } else if (rc != DH_OK || dh_bcf_scan_begin(b->bcf) != DH_OK) snprintf(err, sizeof err, "%s", dh_last_error());
    if (err[0]) { l->deferred_err = strdup(err); l->done = 1; }   /* the reference meets these in local init (after the schema is known) */
    api.duckdb_init_set_max_threads(info, 64);
    api.duckdb_init_set_init_data(info, l, bcf_global_free);
    __atomic_store_n(&g_last_bcf_global, l, __ATOMIC_RELEASE);
}

static void read_bcf_local_init(duckdb_init_info info) {
    bcf_bind_t *b = (bcf_bind_t *)api.duckdb_init_get_bind_data(info);
    bcf_global_t *l = __atomic_load_n(&g_last_bcf_global, __ATOMIC_ACQUIRE);
    if (l && l->bind == b && l->deferred_err) { api.duckdb_init_set_error(info, l->deferred_err); return; }
    api.duckdb_init_set_init_data(info, api.duckdb_malloc(8), api.duckdb_free);
}

/* ---- column fillers: one batch column -> one output vector, rows [0, n) of the chunk -------------------------------------
 * Cell k of the chunk comes from batch cell idx(k) = (i0 + k) / div * mul + add: site-level columns of tidy rows repeat
 * (div = n_samples), FORMAT columns of tidy rows run on (div = 1), wide FORMAT columns stride by n_samples.
 * type: 0 BOOLEAN 1 INTEGER 2 FLOAT 3 VARCHAR 4 BIGINT 5 DOUBLE */
typedef struct { uint64_t i0, div, mul, add; } cell_map;
static inline uint64_t map_idx(const cell_map *m, idx_t k) { return (m->i0 + k) / m->div * m->mul + m->add; }

static void fill_bcf_column(duckdb_vector vec, idx_t n, int type, int is_list, const void *col, const uint64_t *validity, const cell_map *m, const uint8_t *heap) {
    const int contiguous = m->div == 1 && m->mul == 1;
    uint64_t *val = NULL;
    if (!col) {      /* not produced: NULL everywhere */
        api.duckdb_vector_ensure_validity_writable(vec);
        val = api.duckdb_vector_get_validity(vec);
        for (idx_t k = 0; k < n; ++k) api.duckdb_validity_set_row_invalid(val, k);
        if (is_list) { duckdb_list_entry *out = (duckdb_list_entry *)api.duckdb_vector_get_data(vec); const idx_t t0 = api.duckdb_list_vector_get_size(vec); for (idx_t k = 0; k < n; ++k) { out[k].offset = t0; out[k].length = 0; } }
        return;
    }
    void *dst = api.duckdb_vector_get_data(vec);
    if (is_list) {
        /* one count pass, one reserve / set_size per chunk, one fill pass (the reference does a reserve per cell, bcf_reader.c:1542-1733) */
        const duckdb_list_entry *cells = (const duckdb_list_entry *)col;
        duckdb_list_entry *out = (duckdb_list_entry *)dst;
        idx_t total = api.duckdb_list_vector_get_size(vec), need = total;
        for (idx_t k = 0; k < n; ++k) { const uint64_t i = map_idx(m, k); if (bit_valid(validity, i)) need += cells[i].length; }
        api.duckdb_list_vector_reserve(vec, need);
        api.duckdb_list_vector_set_size(vec, need);
        duckdb_vector child = api.duckdb_list_vector_get_child(vec);
        void *cd = api.duckdb_vector_get_data(child);
        for (idx_t k = 0; k < n; ++k) {
            const uint64_t i = map_idx(m, k);
            out[k].offset = total;
            out[k].length = 0;
            if (!bit_valid(validity, i)) {
                if (!val) { api.duckdb_vector_ensure_validity_writable(vec); val = api.duckdb_vector_get_validity(vec); }
                api.duckdb_validity_set_row_invalid(val, k);
                continue;
            }
            const duckdb_list_entry src = cells[i];
            out[k].length = src.length;
            if (type == 3) {
                const duckdb_string_t *sc = (const duckdb_string_t *)(heap + 16 * src.offset);
                for (uint64_t e = 0; e < src.length; ++e) put_cell(child, (duckdb_string_t *)cd, total + e, &sc[e]);
            } else memcpy((uint8_t *)cd + 4 * total, heap + 4 * src.offset, 4 * src.length);
            total += src.length;
        }
        return;
    }
    const size_t w = type == 0 ? 1 : (type == 1 || type == 2) ? 4 : type == 3 ? 16 : 8;
    if (contiguous) {
        const uint64_t i0 = m->i0 + m->add;
        if (type != 3) memcpy(dst, (const uint8_t *)col + i0 * w, n * w);
        else {
            const duckdb_string_t *cells = (const duckdb_string_t *)col + i0;
            duckdb_string_t *out = (duckdb_string_t *)dst;
            if (range_all_valid(validity, i0, n)) for (idx_t k = 0; k < n; ++k) put_cell(vec, out, k, &cells[k]);
            else for (idx_t k = 0; k < n; ++k) if (bit_valid(validity, i0 + k)) put_cell(vec, out, k, &cells[k]);
        }
        copy_validity(vec, validity, i0, n);
        return;
    }
    /* repeated or strided cells: runs of equal source index are filled from one load; a long string is copied into the
     * vector once per run and its cell repeated (several cells may share one heap string) */
    idx_t k = 0;
    while (k < n) {
        const uint64_t i = map_idx(m, k);
        idx_t e = k + 1;
        if (m->div > 1) { const uint64_t left = m->div - (m->i0 + k) % m->div; e = k + (idx_t)(left < n - k ? left : n - k); }
        if (!bit_valid(validity, i)) {
            if (!val) { api.duckdb_vector_ensure_validity_writable(vec); val = api.duckdb_vector_get_validity(vec); }
            for (idx_t q = k; q < e; ++q) api.duckdb_validity_set_row_invalid(val, q);
            k = e;
            continue;
        }
        switch (type) {
        case 0: memset((uint8_t *)dst + k, ((const uint8_t *)col)[i], e - k); break;
        case 1: case 2: { const uint32_t v = ((const uint32_t *)col)[i]; uint32_t *d = (uint32_t *)dst; for (idx_t q = k; q < e; ++q) d[q] = v; break; }
        case 3: {
            duckdb_string_t *d = (duckdb_string_t *)dst;
            put_cell(vec, d, k, &((const duckdb_string_t *)col)[i]);
            for (idx_t q = k + 1; q < e; ++q) d[q] = d[k];
            break;
        }
        default: { const uint64_t v = ((const uint64_t *)col)[i]; uint64_t *d = (uint64_t *)dst; for (idx_t q = k; q < e; ++q) d[q] = v; break; }
        }
        k = e;
    }
}

/* compact wire layout (dh_bcf_set_mode(DH_BAM_COMPACT_OUTPUT)): the two FORMAT cell kinds the GPU leaves packed */
static void fill_bcf_gt_compact(duckdb_vector vec, idx_t n, const uint32_t *cells, const uint64_t *validity, const cell_map *m, const uint8_t *heap) {
    duckdb_string_t *out = (duckdb_string_t *)api.duckdb_vector_get_data(vec);
    uint64_t *val = NULL;
    for (idx_t k = 0; k < n; ++k) {
        const uint64_t i = map_idx(m, k);
        if (!bit_valid(validity, i)) {
            if (!val) { api.duckdb_vector_ensure_validity_writable(vec); val = api.duckdb_vector_get_validity(vec); }
            api.duckdb_validity_set_row_invalid(val, k);
            continue;
        }
        const uint32_t v = cells[i];
        if (!(v >> 31)) {                         /* <= 3 characters in the cell */
            duckdb_string_t c;
            memset(&c, 0, sizeof c);
            c.value.inlined.length = (v >> 24) & 3u;
            c.value.inlined.inlined[0] = (char)(v & 0xff); c.value.inlined.inlined[1] = (char)((v >> 8) & 0xff); c.value.inlined.inlined[2] = (char)((v >> 16) & 0xff);
            out[k] = c;
        } else {                                  /* {length, bytes} in the heap */
            const uint8_t *p = heap + (v & 0x7fffffffu);
            uint32_t len;
            memcpy(&len, p, 4);
            put_string(vec, out, k, (const char *)p + 4, len);
        }
    }
}
static void fill_bcf_numlist_compact(duckdb_vector vec, idx_t n, const uint16_t *cnt, const uint64_t *validity, const cell_map *m, const uint8_t *heap,
                                     const uint32_t *rec, uint64_t ns) {
    duckdb_list_entry *out = (duckdb_list_entry *)api.duckdb_vector_get_data(vec);
    idx_t total = api.duckdb_list_vector_get_size(vec), need = total;
    for (idx_t k = 0; k < n; ++k) { const uint64_t i = map_idx(m, k); if (bit_valid(validity, i)) need += cnt[i]; }
    api.duckdb_list_vector_reserve(vec, need);
    api.duckdb_list_vector_set_size(vec, need);
    uint32_t *child = (uint32_t *)api.duckdb_vector_get_data(api.duckdb_list_vector_get_child(vec));
    uint64_t *val = NULL;
    for (idx_t k = 0; k < n; ++k) {
        const uint64_t i = map_idx(m, k);
        out[k].offset = total;
        out[k].length = 0;
        if (!bit_valid(validity, i)) {
            if (!val) { api.duckdb_vector_ensure_validity_writable(vec); val = api.duckdb_vector_get_validity(vec); }
            api.duckdb_validity_set_row_invalid(val, k);
            continue;
        }
        const uint64_t r = i / ns, smp = i % ns;
        const uint32_t *src = (const uint32_t *)heap + rec[2 * r] + smp * rec[2 * r + 1];
        const uint32_t c = cnt[i];
        for (uint32_t e = 0; e < c; ++e) child[total + e] = src[e];
        out[k].length = c;
        total += c;
    }
}

static void read_bcf_scan(duckdb_function_info info, duckdb_data_chunk output) {
    bcf_global_t *l = (bcf_global_t *)api.duckdb_function_get_init_data(info);
    if (!l) { api.duckdb_data_chunk_set_size(output, 0); return; }
    bcf_bind_t *b = l->bind;
    const uint64_t ns = (uint64_t)b->n_samples, per_rec = (b->tidy && ns) ? ns : 1;
    const idx_t vs = api.duckdb_vector_size();
    bcf_batch_ref *br = NULL;
    uint64_t r0 = 0;
    idx_t n = 0;
    for (;;) {
        pthread_mutex_lock(&l->mu);
        if (l->done) { pthread_mutex_unlock(&l->mu); api.duckdb_data_chunk_set_size(output, 0); return; }
        if (!l->cur || l->cursor >= l->cur->b.n_records * per_rec) {
            if (l->cur) {
                l->cur->closed = 1;
                if (l->cur->refs == 0) { dh_bcf_release_batch(b->bcf, l->cur->b.slot); free(l->cur); }
                l->cur = NULL;
            }
            bcf_batch_ref *nb = (bcf_batch_ref *)calloc(1, sizeof *nb);
            const int rc = dh_bcf_acquire_batch(b->bcf, &nb->b);
            if (rc == 1) { l->cur = nb; l->cursor = 0; }
            else {
                free(nb);
                if (rc == DH_EBUSY) { pthread_mutex_unlock(&l->mu); usleep(100); continue; }   /* every slot is being filled from: wait for a release */
                if ((rc == 0 || rc == DH_EDATA) && b->n_regions > 0) {   /* end of this region's iterator: chain the next one */
                    const int started = bcf_next_region(b, l);
                    if (started == 1) { pthread_mutex_unlock(&l->mu); continue; }
                    if (started < 0) api.duckdb_function_set_error(info, dh_last_error());
                }
                l->done = 1;
                pthread_mutex_unlock(&l->mu);
                if (rc < 0 && rc != DH_EDATA) api.duckdb_function_set_error(info, dh_last_error());   /* a bad record / block ends the scan silently (bcf_reader.c:1311-1338) */
                api.duckdb_data_chunk_set_size(output, 0);
                return;
            }
        }
        br = l->cur;
        r0 = l->cursor;
        const uint64_t left = br->b.n_records * per_rec - r0;
        n = left < vs ? (idx_t)left : vs;
        l->cursor += n;
        br->refs++;
        pthread_mutex_unlock(&l->mu);
        break;
    }
    static const int CORE_TYPE[7] = {3, 4, 3, 3, 3, 5, 3};
    const dh_bcf_batch *bt = &br->b;
    const cell_map site = {r0, per_rec, 1, 0};            /* record of row k: (r0 + k) / per_rec */
    for (idx_t i = 0; i < l->n_proj; ++i) {
        duckdb_vector vec = api.duckdb_data_chunk_get_vector(output, i);
        const idx_t c = l->col_ids[i];
        if (c < 7) fill_bcf_column(vec, n, CORE_TYPE[c], c == 4 || c == 6, bt->core[c], bt->core_validity[c], &site, bt->heap);
        else if ((int)c < b->info_start + b->n_info) {
            const int k = (int)c - b->info_start;
            fill_bcf_column(vec, n, b->info[k].type, b->info[k].is_list, bt->info[k], bt->info_validity[k], &site, bt->heap);
        } else if ((int)c == b->sample_id_col) {
            duckdb_string_t *out = (duckdb_string_t *)api.duckdb_vector_get_data(vec);
            for (idx_t k = 0; k < n; ++k) put_cell(vec, out, k, &b->sample_cells[(r0 + k) % ns]);
        } else {
            const uint64_t k = c - (idx_t)b->fmt_start, f = k % (uint64_t)b->n_fmt;
            if (b->default_gt) { fill_bcf_column(vec, n, 3, 0, NULL, NULL, &site, bt->heap); continue; }
            const cell_map m = b->tidy ? (cell_map){r0, 1, 1, 0} : (cell_map){r0, 1, ns, k / (uint64_t)b->n_fmt};
            const int is_gt = b->fmt[f].type == 3 && !b->fmt[f].is_list && !strcmp(b->fmt[f].name, "GT");
            const int numlist = b->fmt[f].is_list && (b->fmt[f].type == 1 || b->fmt[f].type == 2);
            if (bt->layout == DH_BAM_COMPACT_OUTPUT && bt->format[f] && is_gt) fill_bcf_gt_compact(vec, n, (const uint32_t *)bt->format[f], bt->format_validity[f], &m, bt->heap);
            else if (bt->layout == DH_BAM_COMPACT_OUTPUT && bt->format[f] && numlist)
                fill_bcf_numlist_compact(vec, n, (const uint16_t *)bt->format[f], bt->format_validity[f], &m, bt->heap, (const uint32_t *)bt->format_rec[f], ns);
            else fill_bcf_column(vec, n, b->fmt[f].type, b->fmt[f].is_list, bt->format[f], bt->format_validity[f], &m, bt->heap);
        }
    }
    pthread_mutex_lock(&l->mu);
    if (--br->refs == 0 && br->closed) { dh_bcf_release_batch(b->bcf, br->b.slot); free(br); }
    pthread_mutex_unlock(&l->mu);
    api.duckdb_data_chunk_set_size(output, n);
}

static void register_read_bcf(duckdb_connection con) {
    duckdb_table_function tf = api.duckdb_create_table_function();
    duckdb_logical_type t_varchar = api.duckdb_create_logical_type(DUCKDB_TYPE_VARCHAR), t_bool = api.duckdb_create_logical_type(DUCKDB_TYPE_BOOLEAN);
    api.duckdb_table_function_set_name(tf, "read_bcf");
    api.duckdb_table_function_add_parameter(tf, t_varchar);
    api.duckdb_table_function_add_named_parameter(tf, "region", t_varchar);
    api.duckdb_table_function_add_named_parameter(tf, "index_path", t_varchar);
    api.duckdb_table_function_add_named_parameter(tf, "tidy_format", t_bool);
    api.duckdb_table_function_set_bind(tf, read_bcf_bind);
    api.duckdb_table_function_set_init(tf, read_bcf_init);
    api.duckdb_table_function_set_local_init(tf, read_bcf_local_init);
    api.duckdb_table_function_set_function(tf, read_bcf_scan);
    api.duckdb_table_function_supports_projection_pushdown(tf, true);
    api.duckdb_register_table_function(con, tf);
    api.duckdb_destroy_table_function(&tf);
    api.duckdb_destroy_logical_type(&t_varchar);
    api.duckdb_destroy_logical_type(&t_bool);
}

/* ================================================================================================================
 * bgunzip — like the reference, the work happens in bind and the scan returns one row (src/bgzip.c:86-200, 300-330)
 * ================================================================================================================ */
typedef struct { char *out_path; int64_t bytes_in, bytes_out; int emitted; } gz_bind_t;
static void gz_bind_free(void *p) { gz_bind_t *g = (gz_bind_t *)p; if (g) { free(g->out_path); api.duckdb_free(g); } }

static void bgunzip_bind(duckdb_bind_info info) {
    duckdb_value pv = api.duckdb_bind_get_parameter(info, 0);
    char *in = api.duckdb_get_varchar(pv);
    api.duckdb_destroy_value(&pv);
    if (!in || !in[0]) { api.duckdb_bind_set_error(info, "bgunzip requires a file path"); if (in) api.duckdb_free(in); return; }
    char *outp = named_varchar(info, "output_path");
    const int keep = named_bool(info, "keep", 1), overwrite = named_bool(info, "overwrite", 0);
    char out_path[4096], err[4608];
    err[0] = 0;
    if (outp) snprintf(out_path, sizeof out_path, "%s", outp);
    else {
        const size_t n = strlen(in);
        if (n > 3 && !strcmp(in + n - 3, ".gz")) snprintf(out_path, sizeof out_path, "%.*s", (int)(n - 3), in);
        else snprintf(out_path, sizeof out_path, "%s.out", in);
    }
    /* one streaming pass, file to file (src/bgzip.c:154-200 is a bgzf_read / fwrite loop): neither file is held in memory, the
     * inflated size is whatever the blocks hold; pinned memory = the three staging slots of dh_bgzf_inflate_fd */
    uint64_t csize = 0, usize = 0;
    dh_ctx *ctx = NULL;
    int in_fd = -1, out_fd = -1;
    if (!overwrite && file_exists(out_path)) snprintf(err, sizeof err, "bgunzip: output '%s' already exists (use overwrite := TRUE to replace)", out_path);
    else {
        int devs[MAX_SOURCES];
        device_list(devs);
        struct stat stt;
        if ((in_fd = open(in, O_RDONLY)) < 0 || fstat(in_fd, &stt) != 0 || !S_ISREG(stt.st_mode)) snprintf(err, sizeof err, "bgunzip: cannot open input %s", in);
        else if (!(ctx = ctx_acquire(devs[0]))) snprintf(err, sizeof err, "bgunzip: no usable CUDA device (%s)", dh_last_error());
        else if ((out_fd = open(out_path, O_WRONLY | O_CREAT | O_TRUNC, 0666)) < 0) snprintf(err, sizeof err, "bgunzip: cannot open output %s: %s", out_path, strerror(errno));
        else {
            dh_inflate_stats st;
            const int rc = dh_bgzf_inflate_fd(ctx, in_fd, out_fd, &st);
            if (rc == DH_EIO) snprintf(err, sizeof err, "bgunzip: write error");
            else if (rc != DH_OK) snprintf(err, sizeof err, "bgunzip: read error");
            csize = (uint64_t)stt.st_size;       /* bytes_in is the size of the input file (src/bgzip.c:208) */
            usize = st.out_bytes;
            if (close(out_fd) != 0 && !err[0]) snprintf(err, sizeof err, "bgunzip: close error");
            out_fd = -1;
        }
    }
    if (in_fd >= 0) close(in_fd);
    if (out_fd >= 0) close(out_fd);
    if (ctx) ctx_release(ctx);
    if (err[0]) api.duckdb_bind_set_error(info, err);
    else {
        if (!keep) unlink(in);
        duckdb_logical_type t_bool = api.duckdb_create_logical_type(DUCKDB_TYPE_BOOLEAN), t_varchar = api.duckdb_create_logical_type(DUCKDB_TYPE_VARCHAR),
                            t_bigint = api.duckdb_create_logical_type(DUCKDB_TYPE_BIGINT);
        api.duckdb_bind_add_result_column(info, "success", t_bool);
        api.duckdb_bind_add_result_column(info, "output_path", t_varchar);
        api.duckdb_bind_add_result_column(info, "bytes_in", t_bigint);
        api.duckdb_bind_add_result_column(info, "bytes_out", t_bigint);
        api.duckdb_destroy_logical_type(&t_bool); api.duckdb_destroy_logical_type(&t_varchar); api.duckdb_destroy_logical_type(&t_bigint);
        gz_bind_t *g = (gz_bind_t *)api.duckdb_malloc(sizeof *g);
        g->out_path = strdup(out_path); g->bytes_in = (int64_t)csize; g->bytes_out = (int64_t)usize; g->emitted = 0;
        api.duckdb_bind_set_bind_data(info, g, gz_bind_free);
    }
    api.duckdb_free(in);
    if (outp) api.duckdb_free(outp);
}
static void bgunzip_init(duckdb_init_info info) { ((gz_bind_t *)api.duckdb_init_get_bind_data(info))->emitted = 0; }
static void bgunzip_scan(duckdb_function_info info, duckdb_data_chunk output) {
    gz_bind_t *g = (gz_bind_t *)api.duckdb_function_get_bind_data(info);
    if (g->emitted) { api.duckdb_data_chunk_set_size(output, 0); return; }
    ((bool *)api.duckdb_vector_get_data(api.duckdb_data_chunk_get_vector(output, 0)))[0] = true;
    api.duckdb_vector_assign_string_element(api.duckdb_data_chunk_get_vector(output, 1), 0, g->out_path ? g->out_path : "");
    ((int64_t *)api.duckdb_vector_get_data(api.duckdb_data_chunk_get_vector(output, 2)))[0] = g->bytes_in;
    ((int64_t *)api.duckdb_vector_get_data(api.duckdb_data_chunk_get_vector(output, 3)))[0] = g->bytes_out;
    g->emitted = 1;
    api.duckdb_data_chunk_set_size(output, 1);
}
static void register_bgunzip(duckdb_connection con) {
    duckdb_table_function tf = api.duckdb_create_table_function();
    duckdb_logical_type t_varchar = api.duckdb_create_logical_type(DUCKDB_TYPE_VARCHAR), t_int = api.duckdb_create_logical_type(DUCKDB_TYPE_INTEGER),
                        t_bool = api.duckdb_create_logical_type(DUCKDB_TYPE_BOOLEAN);
    api.duckdb_table_function_set_name(tf, "bgunzip");
    api.duckdb_table_function_add_parameter(tf, t_varchar);
    api.duckdb_table_function_add_named_parameter(tf, "output_path", t_varchar);
    api.duckdb_table_function_add_named_parameter(tf, "threads", t_int);      /* accepted for compatibility: the GPU does the inflating */
    api.duckdb_table_function_add_named_parameter(tf, "keep", t_bool);
    api.duckdb_table_function_add_named_parameter(tf, "overwrite", t_bool);
    api.duckdb_table_function_set_bind(tf, bgunzip_bind);
    api.duckdb_table_function_set_init(tf, bgunzip_init);
    api.duckdb_table_function_set_function(tf, bgunzip_scan);
    api.duckdb_register_table_function(con, tf);
    api.duckdb_destroy_table_function(&tf);
    api.duckdb_destroy_logical_type(&t_varchar); api.duckdb_destroy_logical_type(&t_int); api.duckdb_destroy_logical_type(&t_bool);
}

/* ================================================================================================================
 * bgzip (N1) — the reference compresses in bind with bgzf_write (src/bgzip.c:209-273); here dh_bgzf_deflate_fd streams the file through
 * the GPU compressor.  Same parameters, same one-row result, same error texts; `threads` is accepted and unused, `level` 0 writes
 * stored blocks and every other level runs the one GPU compressor.
 * ================================================================================================================ */
static void bgzip_bind(duckdb_bind_info info) {
    duckdb_value pv = api.duckdb_bind_get_parameter(info, 0);
    char *in = api.duckdb_get_varchar(pv);
    api.duckdb_destroy_value(&pv);
    if (!in || !in[0]) { api.duckdb_bind_set_error(info, "bgzip requires a file path"); if (in) api.duckdb_free(in); return; }
    char *outp = named_varchar(info, "output_path");
    const int keep = named_bool(info, "keep", 1), overwrite = named_bool(info, "overwrite", 0);
    int level = (int)named_int(info, "level", -1);
    if (level < 0 || level > 9) level = -1;                  /* src/bgzip.c:222-223: anything else opens with mode "w" */
    char out_path[4096], err[4608];
    err[0] = 0;
    if (outp) snprintf(out_path, sizeof out_path, "%s", outp);
    else snprintf(out_path, sizeof out_path, "%s.gz", in);
    uint64_t bytes_in = 0, bytes_out = 0;
    dh_ctx *ctx = NULL;
    int in_fd = -1, out_fd = -1;
    if (!overwrite && file_exists(out_path)) snprintf(err, sizeof err, "bgzip: output '%s' already exists (use overwrite := TRUE to replace)", out_path);
    else {
        int devs[MAX_SOURCES];
        device_list(devs);
        if ((in_fd = open(in, O_RDONLY)) < 0) snprintf(err, sizeof err, "bgzip: cannot open input %s: %s", in, strerror(errno));
        else if (!(ctx = ctx_acquire(devs[0]))) snprintf(err, sizeof err, "bgzip: no usable CUDA device (%s)", dh_last_error());
        else if ((out_fd = open(out_path, O_WRONLY | O_CREAT | O_TRUNC, 0666)) < 0) snprintf(err, sizeof err, "bgzip: cannot open output %s", out_path);
        else {
            dh_deflate_stats st;
            const int rc = dh_bgzf_deflate_fd(ctx, in_fd, out_fd, level, &st);
            if (rc == DH_EIO && strstr(dh_last_error(), "read")) snprintf(err, sizeof err, "bgzip: read error");
            else if (rc != DH_OK) snprintf(err, sizeof err, "bgzip: write error");
            bytes_in = st.in_bytes;
            bytes_out = st.out_bytes;
            if (close(out_fd) != 0 && !err[0]) snprintf(err, sizeof err, "bgzip: close error");
            out_fd = -1;
        }
    }
    if (in_fd >= 0) close(in_fd);
    if (out_fd >= 0) close(out_fd);
    if (ctx) ctx_release(ctx);
    if (err[0]) api.duckdb_bind_set_error(info, err);
    else {
        if (!keep) unlink(in);
        duckdb_logical_type t_bool = api.duckdb_create_logical_type(DUCKDB_TYPE_BOOLEAN), t_varchar = api.duckdb_create_logical_type(DUCKDB_TYPE_VARCHAR),
                            t_bigint = api.duckdb_create_logical_type(DUCKDB_TYPE_BIGINT);
        api.duckdb_bind_add_result_column(info, "success", t_bool);
        api.duckdb_bind_add_result_column(info, "output_path", t_varchar);
        api.duckdb_bind_add_result_column(info, "bytes_in", t_bigint);
        api.duckdb_bind_add_result_column(info, "bytes_out", t_bigint);
        api.duckdb_destroy_logical_type(&t_bool); api.duckdb_destroy_logical_type(&t_varchar); api.duckdb_destroy_logical_type(&t_bigint);
        gz_bind_t *g = (gz_bind_t *)api.duckdb_malloc(sizeof *g);
        g->out_path = strdup(out_path); g->bytes_in = (int64_t)bytes_in; g->bytes_out = (int64_t)bytes_out; g->emitted = 0;
        api.duckdb_bind_set_bind_data(info, g, gz_bind_free);
    }
    api.duckdb_free(in);
    if (outp) api.duckdb_free(outp);
}
static void register_bgzip(duckdb_connection con) {
    duckdb_table_function tf = api.duckdb_create_table_function();
    duckdb_logical_type t_varchar = api.duckdb_create_logical_type(DUCKDB_TYPE_VARCHAR), t_int = api.duckdb_create_logical_type(DUCKDB_TYPE_INTEGER),
                        t_bool = api.duckdb_create_logical_type(DUCKDB_TYPE_BOOLEAN);
    api.duckdb_table_function_set_name(tf, "bgzip");
    api.duckdb_table_function_add_parameter(tf, t_varchar);
    api.duckdb_table_function_add_named_parameter(tf, "output_path", t_varchar);
    api.duckdb_table_function_add_named_parameter(tf, "threads", t_int);
    api.duckdb_table_function_add_named_parameter(tf, "level", t_int);
    api.duckdb_table_function_add_named_parameter(tf, "keep", t_bool);
    api.duckdb_table_function_add_named_parameter(tf, "overwrite", t_bool);
    api.duckdb_table_function_set_bind(tf, bgzip_bind);
    api.duckdb_table_function_set_init(tf, bgunzip_init);
    api.duckdb_table_function_set_function(tf, bgunzip_scan);
    api.duckdb_register_table_function(con, tf);
    api.duckdb_destroy_table_function(&tf);
    api.duckdb_destroy_logical_type(&t_varchar); api.duckdb_destroy_logical_type(&t_int); api.duckdb_destroy_logical_type(&t_bool);
}

/* ================================================================================================================
 * read_tabix (N4) — the generic tab-separated reader of the reference (src/tabix_reader.c:500-778 bind, 852-1030 scan), sequential
 * scan of a bgzip'ed file.  Bind sniffs the first lines like the reference (the index's meta character and line_skip, the header
 * line, the number of fields of the first data line, auto_detect over 100 lines); the scan takes batches of lines from dh_text
 * (every line and field located on the GPU) and applies the reference's per-cell rules: absent, empty and "." are NULL, BIGINT /
 * DOUBLE columns go through strtoll / strtod and are NULL unless the whole field parses.  Not served by this build: region queries
 * (bind error), read_gtf / read_gff.
 * ================================================================================================================ */
#define TBX_MAX_COLS 256
typedef struct {
    char *path;
    dh_ctx *ctx;
    dh_text *text;
    int n_cols, meta_char, line_skip, skip_header_line;
    int types[TBX_MAX_COLS];            /* DUCKDB_TYPE_VARCHAR / BIGINT / INTEGER / DOUBLE */
} tbx_bind_t;
typedef struct {
    tbx_bind_t *bind;
    idx_t n_proj;
    idx_t *col_ids;                     /* cell slot of output vector i, or -1 */
    int *col_types;                     /* its type */
    dh_text_batch b;
    int have;                           /* a batch is open */
    uint64_t next;                      /* next line of it */
    int skip_remaining, skipped_header, finished;
} tbx_global_t;

static void tbx_bind_free(void *p) {
    tbx_bind_t *b = (tbx_bind_t *)p;
    if (!b) return;
    if (b->text) dh_text_close(b->text);
    if (b->ctx) ctx_release(b->ctx);
    free(b->path);
    api.duckdb_free(b);
}
static void tbx_global_free(void *p) { tbx_global_t *g = (tbx_global_t *)p; if (g) { free(g->col_ids); free(g->col_types); api.duckdb_free(g); } }

static int tbx_type_of_name(const char *s) {           /* src/tabix_reader.c:218-232 */
    if (!s) return DUCKDB_TYPE_VARCHAR;
    if (!strcasecmp(s, "INT") || !strcasecmp(s, "INTEGER")) return DUCKDB_TYPE_INTEGER;
    if (!strcasecmp(s, "BIGINT") || !strcasecmp(s, "LONG")) return DUCKDB_TYPE_BIGINT;
    if (!strcasecmp(s, "DOUBLE") || !strcasecmp(s, "FLOAT") || !strcasecmp(s, "REAL")) return DUCKDB_TYPE_DOUBLE;
    return DUCKDB_TYPE_VARCHAR;
}
static int tbx_count_fields(const char *s, size_t n) { int c = 1; for (size_t i = 0; i < n; ++i) c += s[i] == '\t'; return c; }
/* the idx-th field of line s[0, n): NULL if the line has fewer */
static const char *tbx_field(const char *s, size_t n, int idx, int *len) {
    int cur = 0;
    size_t start = 0;
    for (size_t i = 0; i <= n; ++i)
        if (i == n || s[i] == '\t') {
            if (cur == idx) { *len = (int)(i - start); return s + start; }
            ++cur;
            start = i + 1;
        }
    *len = 0;
    return NULL;
}
static int tbx_parse_i64(const char *s, int len, int64_t *out) {      /* parse_int64_span, src/tabix_reader.c:58-68 */
    if (len <= 0 || len >= 128) return 0;
    char buf[128], *end = NULL;
    memcpy(buf, s, (size_t)len);
    buf[len] = 0;
    const long long v = strtoll(buf, &end, 10);
    if (!end || *end) return 0;
    *out = (int64_t)v;
    return 1;
}
static int tbx_parse_f64(const char *s, int len, double *out) {
    if (len <= 0 || len >= 128) return 0;
    char buf[128], *end = NULL;
    memcpy(buf, s, (size_t)len);
    buf[len] = 0;
    const double v = strtod(buf, &end);
    if (!end || *end) return 0;
    *out = v;
    return 1;
}
static int tbx_is_integer(const char *s, int len) {
    int i = 0;
    if (len == 0) return 0;
    if (s[i] == '-' || s[i] == '+') ++i;
    if (i >= len) return 0;
    for (; i < len; ++i) if (s[i] < '0' || s[i] > '9') return 0;
    return 1;
}
static int tbx_is_float(const char *s, int len) {      /* is_float_field: strtod takes the whole field (no length cap there) */
    if (len == 0) return 0;
    char *tmp = (char *)malloc((size_t)len + 1), *end = NULL;
    if (!tmp) return 0;
    memcpy(tmp, s, (size_t)len);
    tmp[len] = 0;
    strtod(tmp, &end);
    const int ok = end && !*end;
    free(tmp);
    return ok;
}
/* meta character and line_skip of the tabix index next to the file (tbx_index_load2 -> tbx->conf, htslib/tbx.c:443-470): a .csi's
 * aux block or a .tbi's header hold {format, col_seq, col_beg, col_end, meta, skip, l_nm}.  Returns 1 if an index was read. */
static int tbx_conf_from_index(dh_ctx *ctx, const char *path, const char *index_path, int *meta, int *skip) {
    char cand[2][1100];
    int nc = 0;
    if (index_path) snprintf(cand[nc++], sizeof cand[0], "%s", index_path);
    else { snprintf(cand[nc++], sizeof cand[0], "%s.csi", path); snprintf(cand[nc++], sizeof cand[0], "%s.tbi", path); }
    for (int i = 0; i < nc; ++i) {
        if (!file_exists(cand[i])) continue;
        uint64_t isz = 0, n = 0;
        uint8_t *ibuf = slurp_pinned(cand[i], &isz);
        if (!ibuf) return 0;
        uint8_t head[4096];
        /* the index is BGZF-compressed: its first block holds the header */
        uint32_t bsize = isz >= 18 ? (uint32_t)(ibuf[16] | (ibuf[17] << 8)) + 1u : 0u;
        int ok = bsize >= 26 && bsize <= isz;
        uint8_t *tmp = ok ? (uint8_t *)malloc(65536) : NULL;
        if (tmp && dh_bgzf_inflate_host(ctx, ibuf, bsize, tmp, 65536, &n, NULL) == DH_OK && n >= 36) memcpy(head, tmp, n < sizeof head ? n : sizeof head);
        else ok = 0;
        free(tmp);
        dh_host_free(ibuf);
        if (!ok) return 0;
        const int32_t *w = NULL;
        if (!memcmp(head, "TBI\1", 4)) w = (const int32_t *)(head + 8);                          /* magic, n_ref, then the conf */
        else if (!memcmp(head, "CSI\1", 4) && *(const int32_t *)(head + 12) >= 28) w = (const int32_t *)(head + 16);   /* magic, min_shift, depth, l_aux, aux */
        if (!w) return 0;
        *meta = w[4];
        *skip = w[5];
        return 1;
    }
    return 0;
}

static void read_tabix_bind(duckdb_bind_info info) {
    duckdb_value pv = api.duckdb_bind_get_parameter(info, 0);
    char *path = api.duckdb_get_varchar(pv);
    api.duckdb_destroy_value(&pv);
    if (!path || !path[0]) { api.duckdb_bind_set_error(info, "read_tabix requires a file path"); if (path) api.duckdb_free(path); return; }
    char *region = named_varchar(info, "region"), *index_path = named_varchar(info, "index_path");
    const int header = named_bool(info, "header", 0), auto_detect = named_bool(info, "auto_detect", 0);
    char err[1400];
    err[0] = 0;
    tbx_bind_t *b = (tbx_bind_t *)api.duckdb_malloc(sizeof *b);
    memset(b, 0, sizeof *b);
    b->meta_char = '#';
    char **names = NULL;
    int n_names = 0, types_given = 0, n_types = 0;
    duckdb_value v = api.duckdb_bind_get_named_parameter(info, "header_names");
    if (v && !api.duckdb_is_null_value(v)) {
        const idx_t n = api.duckdb_get_list_size(v);
        if (n > 0) {
            names = (char **)calloc((size_t)n, sizeof(char *));
            n_names = (int)n;
            for (idx_t i = 0; i < n; ++i) { duckdb_value e = api.duckdb_get_list_child(v, i); char *s = api.duckdb_get_varchar(e); names[i] = strdup(s ? s : ""); if (s) api.duckdb_free(s); api.duckdb_destroy_value(&e); }
        }
    }
    if (v) api.duckdb_destroy_value(&v);
    v = api.duckdb_bind_get_named_parameter(info, "column_types");
    if (v && !api.duckdb_is_null_value(v)) {
        const idx_t n = api.duckdb_get_list_size(v);
        if (n > 0) {
            types_given = 1;
            n_types = (int)(n > TBX_MAX_COLS ? TBX_MAX_COLS : n);
            for (int i = 0; i < n_types; ++i) { duckdb_value e = api.duckdb_get_list_child(v, (idx_t)i); char *s = api.duckdb_get_varchar(e); b->types[i] = tbx_type_of_name(s); if (s) api.duckdb_free(s); api.duckdb_destroy_value(&e); }
            b->n_cols = (int)n;
        }
    }
    if (v) api.duckdb_destroy_value(&v);

    uint8_t *head = NULL;
    uint64_t head_n = 0;
    int devs[MAX_SOURCES];
    device_list(devs);
    if (region && region[0]) snprintf(err, sizeof err, "read_tabix: region queries are not served by this build (sequential scans only)");
    else if (!file_exists(path)) snprintf(err, sizeof err, "Cannot open file");
    else if (!(b->ctx = ctx_acquire(devs[0]))) snprintf(err, sizeof err, "read_tabix: no usable CUDA device (%s)", dh_last_error());
    else if (!(b->text = dh_text_open_file(b->ctx, path))) snprintf(err, sizeof err, "read_tabix: %s", dh_last_error());
    else if (!(head = (uint8_t *)malloc(4u << 20)) || dh_text_head(b->text, head, 4u << 20, &head_n) != DH_OK) snprintf(err, sizeof err, "read_tabix: %s", head ? dh_last_error() : "out of memory");
    if (!err[0]) {
        int skip_lines = 0, from_index_meta = '#';
        if (tbx_conf_from_index(b->ctx, path, index_path, &from_index_meta, &skip_lines)) b->meta_char = from_index_meta ? from_index_meta : '#';
        b->line_skip = skip_lines;
        /* the first lines, as the reference reads them (src/tabix_reader.c:657-680) */
        int n_cols = 0, header_from_skip = 0;
        char *header_candidate = NULL;
        size_t hc_len = 0;
        uint64_t pos = 0;
        while (pos < head_n) {
            uint64_t e = pos;
            while (e < head_n && head[e] != '\n') ++e;
            if (e == head_n && head_n == (4u << 20)) break;          /* a line cut by the end of what was sniffed */
            size_t len = (size_t)(e - pos);
            const char *line = (const char *)head + pos;
            pos = e + 1;
            if (len && line[len - 1] == '\r') --len;
            if (len == 0) continue;
            if (skip_lines > 0) {
                if (header && !names) { free(header_candidate); header_candidate = strndup(line, len); hc_len = len; header_from_skip = 1; }
                --skip_lines;
                continue;
            }
            if (b->meta_char && line[0] == b->meta_char) continue;
            if (header && !names && !header_candidate) { header_candidate = strndup(line, len); hc_len = len; }
            else { n_cols = tbx_count_fields(line, len); break; }
        }
        if (names && n_names > 0) { n_cols = n_names; b->skip_header_line = header ? 1 : 0; }
        else if (header && header_candidate) {
            n_names = tbx_count_fields(header_candidate, hc_len);
            names = (char **)calloc((size_t)n_names, sizeof(char *));
            for (int i = 0; i < n_names; ++i) {
                int flen = 0;
                const char *f = tbx_field(header_candidate, hc_len, i, &flen);
                while (f && flen > 0 && (*f == ' ' || *f == '\t')) { ++f; --flen; }
                while (f && flen > 0 && (f[flen - 1] == ' ' || f[flen - 1] == '\t')) --flen;
                names[i] = f ? strndup(f, (size_t)flen) : strdup("");
            }
            n_cols = n_names;
            b->skip_header_line = header_from_skip ? 0 : 1;
        }
        free(header_candidate);
        if (n_cols == 0) n_cols = 1;
        if (n_cols > TBX_MAX_COLS) n_cols = TBX_MAX_COLS;
        if (b->n_cols == 0) b->n_cols = n_cols;
        if (types_given && b->n_cols != n_cols) snprintf(err, sizeof err, "column_types length does not match detected column count");
        if (!err[0]) {
            if (!types_given) for (int i = 0; i < b->n_cols; ++i) b->types[i] = DUCKDB_TYPE_VARCHAR;
            if (auto_detect && !types_given) {
                /* src/tabix_reader.c:706-757: up to 100 data lines; integers stay BIGINT, anything strtod takes makes DOUBLE, the rest VARCHAR */
                int state[TBX_MAX_COLS];
                for (int i = 0; i < b->n_cols; ++i) state[i] = DUCKDB_TYPE_INTEGER;
                int skip2 = b->line_skip, skip_header = b->skip_header_line, seen = 0;
                pos = 0;
                while (seen < 100 && pos < head_n) {
                    uint64_t e = pos;
                    while (e < head_n && head[e] != '\n') ++e;
                    size_t len = (size_t)(e - pos);
                    const char *line = (const char *)head + pos;
                    pos = e + 1;
                    if (len && line[len - 1] == '\r') --len;
                    if (len == 0) continue;
                    if (skip2 > 0) { --skip2; continue; }
                    if (b->meta_char && line[0] == b->meta_char) continue;
                    if (skip_header) { skip_header = 0; continue; }
                    for (int i = 0; i < b->n_cols; ++i) {
                        int flen = 0;
                        const char *f = tbx_field(line, len, i, &flen);
                        if (!f || flen == 0 || (flen == 1 && f[0] == '.')) continue;
                        if (tbx_is_integer(f, flen)) continue;
                        if (tbx_is_float(f, flen)) { if (state[i] != DUCKDB_TYPE_VARCHAR) state[i] = DUCKDB_TYPE_DOUBLE; }
                        else state[i] = DUCKDB_TYPE_VARCHAR;
                    }
                    ++seen;
                }
                for (int i = 0; i < b->n_cols; ++i) b->types[i] = state[i] == DUCKDB_TYPE_INTEGER ? DUCKDB_TYPE_BIGINT : state[i];
            }
            for (int i = 0; i < b->n_cols; ++i) {
                char col_name[32];
                const char *name = (names && i < n_names && names[i] && names[i][0]) ? names[i] : (snprintf(col_name, sizeof col_name, "column%d", i), col_name);
                duckdb_logical_type t = api.duckdb_create_logical_type((duckdb_type)b->types[i]);
                api.duckdb_bind_add_result_column(info, name, t);
                api.duckdb_destroy_logical_type(&t);
            }
        }
    }
    free(head);
    for (int i = 0; i < n_names; ++i) free(names[i]);
    free(names);
    if (err[0]) { api.duckdb_bind_set_error(info, err); tbx_bind_free(b); }
    else { b->path = strdup(path); api.duckdb_bind_set_bind_data(info, b, tbx_bind_free); }
    api.duckdb_free(path);
    if (region) api.duckdb_free(region);
    if (index_path) api.duckdb_free(index_path);
}

static void read_tabix_init(duckdb_init_info info) {
    tbx_bind_t *b = (tbx_bind_t *)api.duckdb_init_get_bind_data(info);
    tbx_global_t *g = (tbx_global_t *)api.duckdb_malloc(sizeof *g);
    memset(g, 0, sizeof *g);
    g->bind = b;
    g->n_proj = api.duckdb_init_get_column_count(info);
    g->col_ids = (idx_t *)calloc(g->n_proj ? g->n_proj : 1, sizeof(idx_t));
    g->col_types = (int *)calloc(g->n_proj ? g->n_proj : 1, sizeof(int));
    uint32_t proj[TBX_MAX_COLS];
    uint32_t np = 0;
    /* one cell slot per projected schema column; a column asked for twice shares its slot */
    int slot_of[TBX_MAX_COLS];
    for (int i = 0; i < TBX_MAX_COLS; ++i) slot_of[i] = -1;
    for (idx_t i = 0; i < g->n_proj; ++i) {
        const idx_t c = api.duckdb_init_get_column_index(info, i);
        if (c < (idx_t)b->n_cols && slot_of[c] < 0) { slot_of[c] = (int)np; proj[np++] = (uint32_t)c; }
        g->col_ids[i] = c < (idx_t)b->n_cols ? (idx_t)slot_of[c] : (idx_t)-1;
        g->col_types[i] = c < (idx_t)b->n_cols ? b->types[c] : DUCKDB_TYPE_VARCHAR;
    }
    g->skip_remaining = b->line_skip;
    api.duckdb_init_set_init_data(info, g, tbx_global_free);
    api.duckdb_init_set_max_threads(info, 1);
    if (dh_text_configure(b->text, (uint32_t)b->n_cols, proj, np, b->meta_char) != DH_OK || dh_text_scan_begin(b->text) != DH_OK)
        api.duckdb_init_set_error(info, dh_last_error());
}

static void read_tabix_scan(duckdb_function_info info, duckdb_data_chunk output) {
    tbx_global_t *g = (tbx_global_t *)api.duckdb_function_get_init_data(info);
    tbx_bind_t *b = g->bind;
    const idx_t ncol = api.duckdb_data_chunk_get_column_count(output);
    duckdb_vector vecs[TBX_MAX_COLS];
    void *data[TBX_MAX_COLS];
    for (idx_t c = 0; c < ncol && c < TBX_MAX_COLS; ++c) { vecs[c] = api.duckdb_data_chunk_get_vector(output, c); data[c] = api.duckdb_vector_get_data(vecs[c]); }
    idx_t rows = 0;
    while (rows < 2048 && !g->finished) {
        if (!g->have || g->next >= g->b.n_lines) {
            const int rc = dh_text_next_batch(b->text, &g->b);
            if (rc < 0) { api.duckdb_function_set_error(info, dh_last_error()); g->finished = 1; break; }
            if (rc == 0) { g->finished = 1; break; }
            g->have = 1;
            g->next = 0;
        }
        const dh_text_batch *tb = &g->b;
        for (; g->next < tb->n_lines && rows < 2048; ++g->next) {
            const uint8_t fl = tb->flags[g->next];
            /* src/tabix_reader.c:893-902, in that order: empty; the first line_skip lines; meta; the header line */
            if (fl & 1) continue;
            if (g->skip_remaining > 0) { --g->skip_remaining; continue; }
            if (fl & 2) continue;
            if (b->skip_header_line && !g->skipped_header) { g->skipped_header = 1; continue; }
            const uint32_t *cells = tb->cells + g->next * tb->n_proj * 2u;
            for (idx_t c = 0; c < ncol; ++c) {
                const idx_t slot = g->col_ids[c];
                if (slot == (idx_t)-1) continue;
                const uint32_t beg = cells[2 * slot], end = cells[2 * slot + 1];
                const char *f = (const char *)tb->text + beg;
                const int flen = beg == 0xffffffffu ? 0 : (int)(end - beg);
                int null = beg == 0xffffffffu || flen == 0 || (flen == 1 && f[0] == '.');
                const int ty = g->col_types[c];
                if (!null && (ty == DUCKDB_TYPE_BIGINT || ty == DUCKDB_TYPE_INTEGER)) {
                    int64_t val = 0;
                    if (tbx_parse_i64(f, flen, &val)) { if (ty == DUCKDB_TYPE_BIGINT) ((int64_t *)data[c])[rows] = val; else ((int32_t *)data[c])[rows] = (int32_t)val; }
                    else null = 1;
                } else if (!null && ty == DUCKDB_TYPE_DOUBLE) {
                    double val = 0;
                    if (tbx_parse_f64(f, flen, &val)) ((double *)data[c])[rows] = val; else null = 1;
                } else if (!null) put_string(vecs[c], (duckdb_string_t *)data[c], rows, f, (uint32_t)flen);
                if (null) {
                    api.duckdb_vector_ensure_validity_writable(vecs[c]);
                    uint64_t *val = api.duckdb_vector_get_validity(vecs[c]);
                    val[rows >> 6] &= ~(1ull << (rows & 63));
                }
            }
            ++rows;
        }
    }
    api.duckdb_data_chunk_set_size(output, rows);
}

static void register_read_tabix(duckdb_connection con) {
    duckdb_table_function tf = api.duckdb_create_table_function();
    duckdb_logical_type t_varchar = api.duckdb_create_logical_type(DUCKDB_TYPE_VARCHAR), t_bool = api.duckdb_create_logical_type(DUCKDB_TYPE_BOOLEAN);
    duckdb_logical_type t_list = api.duckdb_create_list_type(t_varchar);
    api.duckdb_table_function_set_name(tf, "read_tabix");
    api.duckdb_table_function_add_parameter(tf, t_varchar);
    api.duckdb_table_function_add_named_parameter(tf, "region", t_varchar);
    api.duckdb_table_function_add_named_parameter(tf, "index_path", t_varchar);
    api.duckdb_table_function_add_named_parameter(tf, "attributes_map", t_bool);     /* accepted like the reference's shared registration; read_gtf / read_gff only */
    api.duckdb_table_function_add_named_parameter(tf, "header", t_bool);
    api.duckdb_table_function_add_named_parameter(tf, "header_names", t_list);
    api.duckdb_table_function_add_named_parameter(tf, "auto_detect", t_bool);
    api.duckdb_table_function_add_named_parameter(tf, "column_types", t_list);
    api.duckdb_table_function_set_bind(tf, read_tabix_bind);
    api.duckdb_table_function_set_init(tf, read_tabix_init);
    api.duckdb_table_function_set_function(tf, read_tabix_scan);
    api.duckdb_table_function_supports_projection_pushdown(tf, true);
    api.duckdb_register_table_function(con, tf);
    api.duckdb_destroy_table_function(&tf);
    api.duckdb_destroy_logical_type(&t_list); api.duckdb_destroy_logical_type(&t_varchar); api.duckdb_destroy_logical_type(&t_bool);
}

/* ================================================================================================================
 * bam_index — like the reference (src/hts_index_builder.c:110-160, 326-345): the index is built in bind (sam_index_build3 there,
 * one GPU scan here: dh_bam_index_build) and the scan returns one row (success, index_path, index_format).  BAI only (min_shift 0).
 * ================================================================================================================ */
typedef struct { char *index_path; int emitted; } ix_bind_t;
static void ix_bind_free(void *p) { ix_bind_t *g = (ix_bind_t *)p; if (g) { free(g->index_path); api.duckdb_free(g); } }
static void bam_index_bind(duckdb_bind_info info) {
    duckdb_value pv = api.duckdb_bind_get_parameter(info, 0);
    char *path = api.duckdb_get_varchar(pv);
    api.duckdb_destroy_value(&pv);
    if (!path || !path[0]) { api.duckdb_bind_set_error(info, "bam_index requires a file path"); if (path) api.duckdb_free(path); return; }
    char *index_path = named_varchar(info, "index_path");
    const int64_t min_shift = named_int(info, "min_shift", 0);
    char err[1200], out_path[4200];
    err[0] = 0;
    if (index_path) snprintf(out_path, sizeof out_path, "%s", index_path);
    else snprintf(out_path, sizeof out_path, "%s.bai", path);
    int devs[MAX_SOURCES];
    device_list(devs);
    dh_ctx *ctx = NULL;
    dh_bam *bam = NULL;
    uint8_t *bai = NULL;
    uint64_t n = 0;
    /* every failure reads like the reference's: "bam_index: failed to build index for PATH (error N)" with sam_index_build3's codes
     * (-2 cannot open, -1 could not index, -4 could not save) */
    int code = 0;
    if (min_shift > 0) snprintf(err, sizeof err, "bam_index: CSI output (min_shift > 0) is not available in this build");
    else if (!(ctx = ctx_acquire(devs[0]))) snprintf(err, sizeof err, "bam_index: no usable CUDA device (%s)", dh_last_error());
    else if (!(bam = dh_bam_open_file(ctx, path))) code = -2;
    else if (dh_bam_index_build(bam, &bai, &n) != DH_OK) code = -1;
    else {
        FILE *f = fopen(out_path, "wb");
        if (!f || fwrite(bai, 1, n, f) != n) code = -4;
        if (f && fclose(f) != 0) code = -4;
    }
    if (code) snprintf(err, sizeof err, "bam_index: failed to build index for %s (error %d)", path, code);
    if (bai) dh_free(bai);
    if (bam) dh_bam_close(bam);
    if (ctx) ctx_release(ctx);
    if (err[0]) api.duckdb_bind_set_error(info, err);
    else {
        duckdb_logical_type t_bool = api.duckdb_create_logical_type(DUCKDB_TYPE_BOOLEAN), t_varchar = api.duckdb_create_logical_type(DUCKDB_TYPE_VARCHAR);
        api.duckdb_bind_add_result_column(info, "success", t_bool);
        api.duckdb_bind_add_result_column(info, "index_path", t_varchar);
        api.duckdb_bind_add_result_column(info, "index_format", t_varchar);
        api.duckdb_destroy_logical_type(&t_bool); api.duckdb_destroy_logical_type(&t_varchar);
        ix_bind_t *g = (ix_bind_t *)api.duckdb_malloc(sizeof *g);
        g->index_path = strdup(out_path); g->emitted = 0;
        api.duckdb_bind_set_bind_data(info, g, ix_bind_free);
    }
    api.duckdb_free(path);
    if (index_path) api.duckdb_free(index_path);
}
static void bam_index_init(duckdb_init_info info) { ((ix_bind_t *)api.duckdb_init_get_bind_data(info))->emitted = 0; }
static void bam_index_scan(duckdb_function_info info, duckdb_data_chunk output) {
    ix_bind_t *g = (ix_bind_t *)api.duckdb_function_get_bind_data(info);
    if (g->emitted) { api.duckdb_data_chunk_set_size(output, 0); return; }
    ((bool *)api.duckdb_vector_get_data(api.duckdb_data_chunk_get_vector(output, 0)))[0] = true;
    api.duckdb_vector_assign_string_element(api.duckdb_data_chunk_get_vector(output, 1), 0, g->index_path ? g->index_path : "");
    api.duckdb_vector_assign_string_element(api.duckdb_data_chunk_get_vector(output, 2), 0, "BAI");
    g->emitted = 1;
    api.duckdb_data_chunk_set_size(output, 1);
}
static void register_bam_index(duckdb_connection con) {
    duckdb_table_function tf = api.duckdb_create_table_function();
    duckdb_logical_type t_varchar = api.duckdb_create_logical_type(DUCKDB_TYPE_VARCHAR), t_int = api.duckdb_create_logical_type(DUCKDB_TYPE_INTEGER);
    api.duckdb_table_function_set_name(tf, "bam_index");
    api.duckdb_table_function_add_parameter(tf, t_varchar);
    api.duckdb_table_function_add_named_parameter(tf, "index_path", t_varchar);
    api.duckdb_table_function_add_named_parameter(tf, "min_shift", t_int);
    api.duckdb_table_function_add_named_parameter(tf, "threads", t_int);      /* accepted for compatibility */
    api.duckdb_table_function_set_bind(tf, bam_index_bind);
    api.duckdb_table_function_set_init(tf, bam_index_init);
    api.duckdb_table_function_set_function(tf, bam_index_scan);
    api.duckdb_register_table_function(con, tf);
    api.duckdb_destroy_table_function(&tf);
    api.duckdb_destroy_logical_type(&t_varchar); api.duckdb_destroy_logical_type(&t_int);
}

/* ---- entry point (reference duckdb_capi/duckdb_extension.h:1179-1195) ------------------------------------------------ */
__attribute__((visibility("default"))) bool duckhts_init_c_api(duckdb_extension_info info, struct duckdb_extension_access *access) {
    const duckdb_ext_api_v1 *host = (const duckdb_ext_api_v1 *)access->get_api(info, DUCKDB_EXT_ABI_VERSION);
    if (!host) return false;
    api = *host;
    duckdb_database *db = access->get_database(info);
    duckdb_connection con;
    if (api.duckdb_connect(*db, &con) == DuckDBError) { access->set_error(info, "Failed to open connection to database"); return false; }
    register_read_bam(con);
    register_read_bcf(con);
    register_bgunzip(con);
    register_bgzip(con);
    register_read_tabix(con);
    register_bam_index(con);
    api.duckdb_disconnect(&con);
    return true;
}
