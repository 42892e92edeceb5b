/* duckhts_ext.c — the DuckDB side of duckhts_b200: a C extension that registers `read_bam` and `bgunzip` with the
 * same names, parameters, column schemas and error behaviour as the reference (src/bam_reader.c:1044-1068,
 * src/bgzip.c:359-380) and serves them from the GPU through the C-ABI of libduckhts_cuda.so
 * (include/duckhts_cuda.h).  Host code in C, as the reference's; no htslib, no CPU decode path.
 *
 * Entry point: duckhts_init_c_api (reference duckdb_capi/duckdb_extension.h:1179-1195) — the extension is a
 * drop-in for the reference's shared object as far as these two table functions go.
 *
 * How a scan maps onto the reference's callbacks (src/bam_reader.c:410-1038):
 *   bind        reads the file into pinned memory, dh_bam_open (BGZF framing on the host, header inflated on the GPU),
 *               declares the schema (13 SAM columns [+ 56 SAMtags columns] — same order and types)
 *   global init max_threads = 1: the GPU is the parallelism; rows come back in file order (the reference's sequential
 *               order; in its contig-parallel mode DuckDB interleaves threads nondeterministically)
 *   local init  projection pushdown -> dh_bam_set_projection, dh_bam_scan_begin
 *   scan        copies up to duckdb_vector_size() rows of the current dh_bam_batch into the output chunk: fixed-width
 *               columns by memcpy, VARCHAR cells of <= 12 bytes by a 16-byte store (the kernel wrote duckdb_string_t),
 *               longer strings through duckdb_vector_assign_string_element_len, NULLs through the validity mask.
 */
#define _GNU_SOURCE
#include <errno.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include "duckdb_ext_abi.h"
#include "duckhts_cuda.h"

static duckdb_ext_api_v1 api;   /* the host's function table, copied by value like DUCKDB_EXTENSION_API_INIT does */

/* ---- one GPU context per process (device from DUCKHTS_DEVICE, default 0) ------------------------------------------ */
static pthread_mutex_t g_lock = PTHREAD_MUTEX_INITIALIZER;
static dh_ctx *g_ctx;
static dh_ctx *shared_ctx(void) {
    pthread_mutex_lock(&g_lock);
    if (!g_ctx) {
        const char *d = getenv("DUCKHTS_DEVICE");
        g_ctx = dh_ctx_create(d ? atoi(d) : 0);
    }
    dh_ctx *c = g_ctx;
    pthread_mutex_unlock(&g_lock);
    return c;
}

/* whole file -> pinned host memory (the C-ABI copies from it at full PCIe rate) */
static uint8_t *slurp_pinned(const char *path, uint64_t *size_out) {
    FILE *f = fopen(path, "rb");
    if (!f) return NULL;
    struct stat st;
    if (fstat(fileno(f), &st) != 0 || !S_ISREG(st.st_mode)) { fclose(f); return NULL; }
    uint64_t n = (uint64_t)st.st_size;
    uint8_t *p = (uint8_t *)dh_host_alloc(n ? n : 1);
    if (!p) { fclose(f); return NULL; }
    uint64_t got = 0;
    while (got < n) {
        size_t k = fread(p + got, 1, (size_t)((n - got) < (64u << 20) ? (n - got) : (64u << 20)), f);
        if (k == 0) break;
        got += k;
    }
    fclose(f);
    if (got != n) { dh_host_free(p); return NULL; }
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
    dh_bam *bam;
    dh_index *index;         /* BAI/CSI next to the file (or index_path), NULL if none: sam_index_load3, bam_reader.c:499 */
    char *region;            /* the `region` parameter as given (comma separated), NULL if absent or empty */
    int standard_tags, has_index, n_contigs;
    int n_cols;
    int aux_col;             /* schema index of AUXILIARY_TAGS (auxiliary_tags := true), else -1 */
} bam_bind_t;

typedef struct {
    idx_t n_proj;
    idx_t *col_ids;          /* schema column of output vector i */
    int *abi_col;            /* its column in dh_bam_batch, or -1 for a column this build always returns NULL for */
    dh_bam_batch batch;
    uint64_t cursor;         /* next row of `batch` to hand out */
    int have_batch, done;
} bam_local_t;

static void bam_bind_free(void *p) {
    bam_bind_t *b = (bam_bind_t *)p;
    if (!b) return;
    if (b->bam) dh_bam_close(b->bam);
    if (b->index) dh_index_free(b->index);
    if (b->image) dh_host_free(b->image);
    free(b->path);
    free(b->region);
    api.duckdb_free(b);
}
static void bam_local_free(void *p) {
    bam_local_t *l = (bam_local_t *)p;
    if (!l) return;
    free(l->col_ids);
    free(l->abi_col);
    api.duckdb_free(l);
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
    bam_bind_t *b = NULL;
    dh_ctx *ctx = shared_ctx();
    if (!ctx) snprintf(err, sizeof err, "read_bam: no usable CUDA device (%s)", dh_last_error());
    else {
        b = (bam_bind_t *)api.duckdb_malloc(sizeof *b);
        memset(b, 0, sizeof *b);
        b->image = slurp_pinned(path, &b->size);
        if (!b->image) snprintf(err, sizeof err, "Failed to open SAM/BAM/CRAM file: %s", path);
        else if (!(b->bam = dh_bam_open(ctx, b->image, b->size))) snprintf(err, sizeof err, "Failed to read SAM/BAM/CRAM header");
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
    if (region && region[0]) b->region = strdup(region);
    b->index = find_index(ctx, path, index_path, 2);
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

static void read_bam_init(duckdb_init_info info) {
    api.duckdb_init_set_max_threads(info, 1);
    api.duckdb_init_set_init_data(info, api.duckdb_malloc(8), api.duckdb_free);
}

static void read_bam_local_init(duckdb_init_info info) {
    bam_bind_t *b = (bam_bind_t *)api.duckdb_init_get_bind_data(info);
    bam_local_t *l = (bam_local_t *)api.duckdb_malloc(sizeof *l);
    memset(l, 0, sizeof *l);
    l->n_proj = api.duckdb_init_get_column_count(info);
    l->col_ids = (idx_t *)calloc(l->n_proj ? l->n_proj : 1, sizeof(idx_t));
    l->abi_col = (int *)calloc(l->n_proj ? l->n_proj : 1, sizeof(int));
    /* projection pushdown (src/bam_reader.c:676-679): only projected columns are decoded on the GPU.  Tag columns are
     * requested from the C-ABI in projection order; tag k of the request is column 13 + k of every batch. */
    uint64_t mask = 0;
    dh_tag_req tags[DH_BAM_MAX_TAGS];
    int n_tags = 0, want_map = 0;
    for (idx_t i = 0; i < l->n_proj; ++i) {
        const idx_t c = api.duckdb_init_get_column_index(info, i);
        l->col_ids[i] = c;
        if (c < N_CORE) { l->abi_col[i] = (int)c; mask |= 1ull << c; continue; }
        if ((int)c == b->aux_col) { l->abi_col[i] = -2; want_map = 1; continue; }
        const char *e = STD_TAGS + 4 * (c - N_CORE);
        if (c - N_CORE >= N_STD_TAGS || n_tags >= DH_BAM_MAX_TAGS) { l->abi_col[i] = -1; continue; }
        tags[n_tags].tag[0] = e[0]; tags[n_tags].tag[1] = e[1]; tags[n_tags].type = e[2]; tags[n_tags].pad = e[2] == 'B' ? e[3] : 0;
        l->abi_col[i] = N_CORE + n_tags;
        if (N_CORE + n_tags < 64) mask |= 1ull << (N_CORE + n_tags);   /* tag columns past bit 63 are always projected */
        ++n_tags;
    }
    /* which rows (src/bam_reader.c:601-671): region(s) through the index; else, with an index and more than one contig,
     * contig by contig (reads without a contig are then never returned); else the whole file in order */
    int rc = dh_bam_set_projection(b->bam, mask, tags, n_tags);
    if (rc == DH_OK) {   /* the MAP leaves out the tags served as typed columns (bam_reader.c:976) */
        char excl[2 * N_STD_TAGS];
        for (int k = 0; k < N_STD_TAGS; ++k) { excl[2 * k] = STD_TAGS[4 * k]; excl[2 * k + 1] = STD_TAGS[4 * k + 1]; }
        rc = dh_bam_set_aux_map(b->bam, want_map, excl, b->standard_tags ? N_STD_TAGS : 0);
    }
    char err[1200];
    err[0] = 0;
    if (rc != DH_OK) snprintf(err, sizeof err, "%s", dh_last_error());
    else if (b->region) {
        if (!b->index) snprintf(err, sizeof err, "Region query requires an index (.bai/.csi/.crai)");
        else {
            char *dup = strdup(b->region), *save = NULL;
            const char *regs[1024];
            int nreg = 0;
            for (char *tok = strtok_r(dup, ",", &save); tok && nreg < 1024; tok = strtok_r(NULL, ",", &save)) regs[nreg++] = tok;
            rc = dh_bam_scan_regions(b->bam, b->index, regs, nreg);
            if (rc == DH_ENOREGION) snprintf(err, sizeof err, "No reads found for region(s): %s", b->region);
            else if (rc != DH_OK) snprintf(err, sizeof err, "%s", dh_last_error());
            free(dup);
        }
    } else if (b->index && b->n_contigs > 1) {
        if (dh_bam_scan_contigs(b->bam, b->index, 0, b->n_contigs) != DH_OK) snprintf(err, sizeof err, "%s", dh_last_error());
    } else if (dh_bam_scan_begin(b->bam, 0, UINT64_MAX) != DH_OK) snprintf(err, sizeof err, "%s", dh_last_error());
    if (err[0]) {
        api.duckdb_init_set_error(info, err);
        bam_local_free(l);
        return;
    }
    api.duckdb_init_set_init_data(info, l, bam_local_free);
}

static size_t core_width(idx_t c) { return c == 1 ? 2 : c == 4 ? 4 : (c == 3 || c == 7 || c == 8) ? 8 : 16; }

static void read_bam_scan(duckdb_function_info info, duckdb_data_chunk output) {
    bam_bind_t *b = (bam_bind_t *)api.duckdb_function_get_bind_data(info);
    bam_local_t *l = (bam_local_t *)api.duckdb_function_get_local_init_data(info);
    if (!l || l->done) { api.duckdb_data_chunk_set_size(output, 0); return; }
    while (!l->have_batch || l->cursor >= l->batch.n_rows) {
        const int rc = dh_bam_next_batch(b->bam, &l->batch);
        if (rc == 1) { l->have_batch = 1; l->cursor = 0; continue; }
        l->done = 1;
        /* rc == 0: end of file, or an invalid record ended the scan: the reference stops silently there too
         * (src/bam_reader.c:754-765).  A failed BGZF block (DH_EDATA) is a read error to htslib = the same silent end.
         * Anything else is a CUDA/runtime failure and is reported. */
        if (rc < 0 && rc != DH_EDATA && rc != DH_EFORMAT) api.duckdb_function_set_error(info, dh_last_error());
        api.duckdb_data_chunk_set_size(output, 0);
        return;
    }
    const idx_t vs = api.duckdb_vector_size();
    const uint64_t r0 = l->cursor;
    const idx_t n = l->batch.n_rows - r0 < vs ? (idx_t)(l->batch.n_rows - r0) : vs;
    for (idx_t i = 0; i < l->n_proj; ++i) {
        duckdb_vector vec = api.duckdb_data_chunk_get_vector(output, i);
        const int ac = l->abi_col[i];
        if (ac == -2) {   /* AUXILIARY_TAGS: list entries re-based onto the vector's child, key / value cells appended */
            const duckdb_list_entry *cells = (const duckdb_list_entry *)l->batch.aux_map + r0;
            const uint64_t *valid = l->batch.aux_map_validity;
            duckdb_list_entry *out = (duckdb_list_entry *)api.duckdb_vector_get_data(vec);
            uint64_t *val = NULL;
            idx_t total = api.duckdb_list_vector_get_size(vec), need = total;
            for (idx_t r = 0; r < n; ++r) need += cells[r].length;
            api.duckdb_list_vector_reserve(vec, need);
            api.duckdb_list_vector_set_size(vec, need);
            duckdb_vector child = api.duckdb_list_vector_get_child(vec);
            duckdb_vector kv[2] = {api.duckdb_struct_vector_get_child(child, 0), api.duckdb_struct_vector_get_child(child, 1)};
            const duckdb_string_t *src[2] = {(const duckdb_string_t *)l->batch.aux_keys, (const duckdb_string_t *)l->batch.aux_values};
            for (idx_t r = 0; r < n; ++r) {
                out[r].offset = total;
                out[r].length = cells[r].length;
                if (!((valid[(r0 + r) >> 6] >> ((r0 + r) & 63)) & 1)) {
                    if (!val) { api.duckdb_vector_ensure_validity_writable(vec); val = api.duckdb_vector_get_validity(vec); }
                    api.duckdb_validity_set_row_invalid(val, r);
                    continue;
                }
                for (int q = 0; q < 2; ++q) {
                    duckdb_string_t *dst = (duckdb_string_t *)api.duckdb_vector_get_data(kv[q]);
                    for (uint64_t e = 0; e < cells[r].length; ++e) {
                        const duckdb_string_t *cl = &src[q][cells[r].offset + e];
                        if (cl->value.inlined.length <= 12) dst[total + e] = *cl;
                        else api.duckdb_vector_assign_string_element_len(kv[q], total + e, cl->value.pointer.ptr, cl->value.inlined.length);
                    }
                }
                total += cells[r].length;
            }
            continue;
        }
        if (ac < 0) {   /* column not served: all NULL */
            api.duckdb_vector_ensure_validity_writable(vec);
            uint64_t *val = api.duckdb_vector_get_validity(vec);
            for (idx_t r = 0; r < n; ++r) api.duckdb_validity_set_row_invalid(val, r);
            if (STD_TAGS[4 * (l->col_ids[i] - N_CORE) + 2] == 'B') {
                duckdb_list_entry *le = (duckdb_list_entry *)api.duckdb_vector_get_data(vec);
                for (idx_t r = 0; r < n; ++r) { le[r].offset = 0; le[r].length = 0; }
            }
            continue;
        }
        const idx_t c = l->col_ids[i];
        size_t w = c < N_CORE ? core_width(c) : (STD_TAGS[4 * (c - N_CORE) + 2] == 'i' || STD_TAGS[4 * (c - N_CORE) + 2] == 'f') ? 8 : 16;
        const uint8_t *src = (const uint8_t *)l->batch.data[ac] + r0 * w;
        void *dst = api.duckdb_vector_get_data(vec);
        const uint64_t *valid = l->batch.validity[ac];
        if (c >= N_CORE && STD_TAGS[4 * (c - N_CORE) + 2] == 'B') {
            /* LIST cells (bam_assign_list_int/_double, src/bam_reader.c:106-138): children are appended to the list child */
            const duckdb_list_entry *cells = (const duckdb_list_entry *)src;
            duckdb_list_entry *out = (duckdb_list_entry *)dst;
            idx_t total = api.duckdb_list_vector_get_size(vec), need = total;
            for (idx_t r = 0; r < n; ++r)
                if (!valid || ((valid[(r0 + r) >> 6] >> ((r0 + r) & 63)) & 1)) need += cells[r].length;
            api.duckdb_list_vector_reserve(vec, need);
            api.duckdb_list_vector_set_size(vec, need);
            uint64_t *child = (uint64_t *)api.duckdb_vector_get_data(api.duckdb_list_vector_get_child(vec));
            for (idx_t r = 0; r < n; ++r) {
                if (valid && !((valid[(r0 + r) >> 6] >> ((r0 + r) & 63)) & 1)) continue;
                memcpy(child + total, (const uint64_t *)l->batch.heap + cells[r].offset, 8 * cells[r].length);
                out[r].offset = total;
                out[r].length = cells[r].length;
                total += cells[r].length;
            }
        } else if (w != 16) memcpy(dst, src, n * w);
        else {
            const duckdb_string_t *cells = (const duckdb_string_t *)src;
            duckdb_string_t *out = (duckdb_string_t *)dst;
            for (idx_t r = 0; r < n; ++r) {
                if (valid && !((valid[(r0 + r) >> 6] >> ((r0 + r) & 63)) & 1)) continue;
                if (cells[r].value.inlined.length <= 12) out[r] = cells[r];
                else api.duckdb_vector_assign_string_element_len(vec, r, cells[r].value.pointer.ptr, cells[r].value.inlined.length);
            }
        }
        if (valid) {
            uint64_t *val = NULL;
            for (idx_t r = 0; r < n; ++r)
                if (!((valid[(r0 + r) >> 6] >> ((r0 + r) & 63)) & 1)) {
                    if (!val) { api.duckdb_vector_ensure_validity_writable(vec); val = api.duckdb_vector_get_validity(vec); }
                    api.duckdb_validity_set_row_invalid(val, r);
                }
        }
    }
    l->cursor += n;
    api.duckdb_data_chunk_set_size(output, n);
}

static void register_read_bam(duckdb_connection con) {
    duckdb_table_function tf = api.duckdb_create_table_function();
    duckdb_logical_type t_varchar = api.duckdb_create_logical_type(DUCKDB_TYPE_VARCHAR), t_bool = api.duckdb_create_logical_type(DUCKDB_TYPE_BOOLEAN);
    api.duckdb_table_function_set_name(tf, "read_bam");
    api.duckdb_table_function_add_parameter(tf, t_varchar);
    api.duckdb_table_function_add_named_parameter(tf, "region", t_varchar);
    api.duckdb_table_function_add_named_parameter(tf, "index_path", t_varchar);
    api.duckdb_table_function_add_named_parameter(tf, "reference", t_varchar);
    api.duckdb_table_function_add_named_parameter(tf, "standard_tags", t_bool);
    api.duckdb_table_function_add_named_parameter(tf, "auxiliary_tags", t_bool);
    api.duckdb_table_function_set_bind(tf, read_bam_bind);
    api.duckdb_table_function_set_init(tf, read_bam_init);
    api.duckdb_table_function_set_local_init(tf, read_bam_local_init);
    api.duckdb_table_function_set_function(tf, read_bam_scan);
    api.duckdb_table_function_supports_projection_pushdown(tf, true);
    api.duckdb_register_table_function(con, tf);
    api.duckdb_destroy_table_function(&tf);
    api.duckdb_destroy_logical_type(&t_varchar);
    api.duckdb_destroy_logical_type(&t_bool);
}

/* ================================================================================================================
 * read_bcf — same parameters, schema and NULL rules as the reference (src/bcf_reader.c:451-830, 2055-2081); rows come
 * from dh_bcf batches: site-level cells once per record, FORMAT cells once per (record, sample).  tidy_format := true
 * repeats the site-level cell for each sample row here, on the host, 16 bytes at a time.
 * ================================================================================================================ */
typedef struct {
    uint8_t *image;
    uint64_t size;
    dh_bcf *bcf;
    int tidy, n_samples, n_info, n_fmt, default_gt;
    int info_start, sample_id_col, fmt_start, n_cols;
    dh_bcf_field *info, *fmt;
    dh_index *index;         /* CSI, loaded when a region is given (bcf_reader.c:902-915) */
    char *region;            /* the parameter as given */
    char **regions;          /* comma-separated, blanks trimmed, empty tokens dropped (parse_regions_duckdb, bcf_reader.c:286-329) */
    int n_regions;
} bcf_bind_t;

typedef struct {
    idx_t n_proj;
    idx_t *col_ids;
    dh_bcf_batch batch;
    uint64_t cursor;        /* next row of the batch: records (wide) or records x samples (tidy) */
    int have_batch, done;
    int next_region;        /* region cursor of a chained region scan (bcf_reader.c:172, 1318-1333) */
} bcf_local_t;

/* start the scan of the next region that yields an iterator: 1 started, 0 none left, < 0 error */
static int bcf_next_region(bcf_bind_t *b, bcf_local_t *l) {
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
    if (b->image) dh_host_free(b->image);
    if (b->index) dh_index_free(b->index);
    free(b->region);
    for (int i = 0; i < b->n_regions; ++i) free(b->regions[i]);
    free(b->regions);
    free(b->info);
    free(b->fmt);
    api.duckdb_free(b);
}
static void bcf_local_free(void *p) {
    bcf_local_t *l = (bcf_local_t *)p;
    if (!l) return;
    free(l->col_ids);
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
    bcf_bind_t *b = NULL;
    dh_ctx *ctx = shared_ctx();
    if (!ctx) snprintf(err, sizeof err, "read_bcf: no usable CUDA device (%s)", dh_last_error());
    else {
        b = (bcf_bind_t *)api.duckdb_malloc(sizeof *b);
        memset(b, 0, sizeof *b);
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
        api.duckdb_destroy_logical_type(&t_varchar); api.duckdb_destroy_logical_type(&t_bigint); api.duckdb_destroy_logical_type(&t_double);
        api.duckdb_destroy_logical_type(&t_vlist);
        api.duckdb_bind_set_bind_data(info, b, bcf_bind_free);
    }
    api.duckdb_free(path);
    if (region) api.duckdb_free(region);
    if (index_path) api.duckdb_free(index_path);
}

static void read_bcf_init(duckdb_init_info info) {
    api.duckdb_init_set_max_threads(info, 1);
    api.duckdb_init_set_init_data(info, api.duckdb_malloc(8), api.duckdb_free);
}

static void read_bcf_local_init(duckdb_init_info info) {
    bcf_bind_t *b = (bcf_bind_t *)api.duckdb_init_get_bind_data(info);
    bcf_local_t *l = (bcf_local_t *)api.duckdb_malloc(sizeof *l);
    memset(l, 0, sizeof *l);
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
    const int rc = dh_bcf_set_projection(b->bcf, core_mask, iw, b->default_gt ? NULL : fw);
    free(iw);
    free(fw);
    if (rc == DH_OK && b->n_regions > 0) {
        if (!b->index) {
            char err[600];
            snprintf(err, sizeof err, "Region query requires an index file (.tbi or .csi). Region: %s", b->region);
            api.duckdb_init_set_error(info, err);
            bcf_local_free(l);
            return;
        }
        const int started = bcf_next_region(b, l);
        if (started < 0) {
            api.duckdb_init_set_error(info, dh_last_error());
            bcf_local_free(l);
            return;
        }
        if (!started) l->done = 1;              /* no region named a contig: an empty result, not an error (bcf_reader.c:952-956) */
    } else if (rc != DH_OK || dh_bcf_scan_begin(b->bcf) != DH_OK) {
        api.duckdb_init_set_error(info, dh_last_error());
        bcf_local_free(l);
        return;
    }
    api.duckdb_init_set_init_data(info, l, bcf_local_free);
}

static int cell_valid(const uint64_t *v, uint64_t i) { return !v || ((v[i >> 6] >> (i & 63)) & 1); }
static void set_null(duckdb_vector vec, idx_t row) {
    api.duckdb_vector_ensure_validity_writable(vec);
    api.duckdb_validity_set_row_invalid(api.duckdb_vector_get_validity(vec), row);
}
static void put_varchar(duckdb_vector vec, idx_t row, const duckdb_string_t *cell) {
    if (cell->value.inlined.length <= 12) ((duckdb_string_t *)api.duckdb_vector_get_data(vec))[row] = *cell;
    else api.duckdb_vector_assign_string_element_len(vec, row, cell->value.pointer.ptr, cell->value.inlined.length);
}
/* one cell of a batch column -> row `row` of `vec`.  type: 0 BOOLEAN 1 INTEGER 2 FLOAT 3 VARCHAR 4 BIGINT 5 DOUBLE */
static void emit_cell(duckdb_vector vec, idx_t row, int type, int is_list, const void *col, const uint64_t *validity, uint64_t i, const uint8_t *heap) {
    const int ok = col && cell_valid(validity, i);
    if (is_list) {
        duckdb_list_entry *out = (duckdb_list_entry *)api.duckdb_vector_get_data(vec);
        duckdb_list_entry e;
        e.offset = api.duckdb_list_vector_get_size(vec);
        e.length = 0;
        if (ok) {
            const duckdb_list_entry src = ((const duckdb_list_entry *)col)[i];
            e.length = src.length;
            if (e.length) {
                api.duckdb_list_vector_reserve(vec, e.offset + e.length);
                api.duckdb_list_vector_set_size(vec, e.offset + e.length);
                duckdb_vector child = api.duckdb_list_vector_get_child(vec);
                if (type == 3) {
                    const duckdb_string_t *cells = (const duckdb_string_t *)(heap + 16 * src.offset);
                    for (uint64_t k = 0; k < src.length; ++k) put_varchar(child, e.offset + k, &cells[k]);
                } else memcpy((uint8_t *)api.duckdb_vector_get_data(child) + 4 * e.offset, heap + 4 * src.offset, 4 * src.length);
            }
        } else set_null(vec, row);
        out[row] = e;
        return;
    }
    if (!ok) { set_null(vec, row); return; }
    void *dst = api.duckdb_vector_get_data(vec);
    switch (type) {
    case 0: ((uint8_t *)dst)[row] = ((const uint8_t *)col)[i]; break;
    case 1: case 2: ((uint32_t *)dst)[row] = ((const uint32_t *)col)[i]; break;
    case 3: put_varchar(vec, row, &((const duckdb_string_t *)col)[i]); break;
    default: ((uint64_t *)dst)[row] = ((const uint64_t *)col)[i]; break;
    }
}

static void read_bcf_scan(duckdb_function_info info, duckdb_data_chunk output) {
    bcf_bind_t *b = (bcf_bind_t *)api.duckdb_function_get_bind_data(info);
    bcf_local_t *l = (bcf_local_t *)api.duckdb_function_get_local_init_data(info);
    if (!l || l->done) { api.duckdb_data_chunk_set_size(output, 0); return; }
    const uint64_t ns = (uint64_t)b->n_samples, per_rec = (b->tidy && ns) ? ns : 1;
    while (!l->have_batch || l->cursor >= l->batch.n_records * per_rec) {
        const int rc = dh_bcf_next_batch(b->bcf, &l->batch);
        if (rc == 1) { l->have_batch = 1; l->cursor = 0; continue; }
        l->have_batch = 0;
        if ((rc == 0 || rc == DH_EDATA || rc == DH_EFORMAT) && b->n_regions > 0) {   /* end of this region's iterator: chain the next one */
            const int started = bcf_next_region(b, l);
            if (started == 1) continue;
            if (started < 0) api.duckdb_function_set_error(info, dh_last_error());
        }
        l->done = 1;
        if (rc < 0 && rc != DH_EDATA && rc != DH_EFORMAT) api.duckdb_function_set_error(info, dh_last_error());   /* a bad record / block ends the scan silently (bcf_reader.c:1311-1338) */
        api.duckdb_data_chunk_set_size(output, 0);
        return;
    }
    const idx_t vs = api.duckdb_vector_size();
    const uint64_t r0 = l->cursor, left = l->batch.n_records * per_rec - r0;
    const idx_t n = left < vs ? (idx_t)left : vs;
    static const int CORE_TYPE[7] = {3, 4, 3, 3, 3, 5, 3};
    const dh_bcf_batch *bt = &l->batch;
    for (idx_t i = 0; i < l->n_proj; ++i) {
        duckdb_vector vec = api.duckdb_data_chunk_get_vector(output, i);
        const idx_t c = l->col_ids[i];
        for (idx_t r = 0; r < n; ++r) {
            const uint64_t row = r0 + r, rec = row / per_rec, smp = row % per_rec;
            if (c < 7) emit_cell(vec, r, CORE_TYPE[c], c == 4 || c == 6, bt->core[c], bt->core_validity[c], rec, bt->heap);
            else if ((int)c < b->info_start + b->n_info) {
                const int k = (int)c - b->info_start;
                emit_cell(vec, r, b->info[k].type, b->info[k].is_list, bt->info[k], bt->info_validity[k], rec, bt->heap);
            } else if ((int)c == b->sample_id_col) api.duckdb_vector_assign_string_element(vec, r, dh_bcf_sample_name(b->bcf, (int32_t)smp));
            else {
                const uint64_t k = c - (idx_t)b->fmt_start, f = k % (uint64_t)b->n_fmt, s_idx = b->tidy ? smp : k / (uint64_t)b->n_fmt;
                if (b->default_gt) set_null(vec, r);
                else emit_cell(vec, r, b->fmt[f].type, b->fmt[f].is_list, bt->format[f], bt->format_validity[f], rec * ns + s_idx, bt->heap);
            }
        }
    }
    l->cursor += n;
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
    uint8_t *comp = NULL, *plain = NULL;
    uint64_t csize = 0, usize = 0;
    dh_ctx *ctx = NULL;
    if (!overwrite && file_exists(out_path)) snprintf(err, sizeof err, "bgunzip: output '%s' already exists (use overwrite := TRUE to replace)", out_path);
    else if (!(ctx = shared_ctx())) snprintf(err, sizeof err, "bgunzip: no usable CUDA device (%s)", dh_last_error());
    else if (!(comp = slurp_pinned(in, &csize))) snprintf(err, sizeof err, "bgunzip: cannot open input %s", in);
    else {
        dh_inflate_stats st;
        int rc = dh_bgzf_inflate_host(ctx, comp, csize, NULL, 0, &usize, &st);      /* framing + verdicts, size of the output */
        if (rc == DH_OK) {
            plain = (uint8_t *)dh_host_alloc(usize ? usize : 1);
            if (!plain) rc = DH_ENOMEM;
            else rc = dh_bgzf_inflate_host(ctx, comp, csize, plain, usize, &usize, &st);
        }
        if (rc != DH_OK) snprintf(err, sizeof err, "bgunzip: read error");
        else {
            FILE *f = fopen(out_path, "wb");
            if (!f) snprintf(err, sizeof err, "bgunzip: cannot open output %s: %s", out_path, strerror(errno));
            else {
                if (fwrite(plain, 1, usize, f) != usize) snprintf(err, sizeof err, "bgunzip: write error");
                if (fclose(f) != 0 && !err[0]) snprintf(err, sizeof err, "bgunzip: close error");
            }
        }
    }
    if (comp) dh_host_free(comp);
    if (plain) dh_host_free(plain);
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
    api.duckdb_disconnect(&con);
    return true;
}
