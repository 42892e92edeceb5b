/* fake_duckdb.c — a minimal DuckDB C-API *host* for table-function extensions.  TEST / BENCH INFRASTRUCTURE.
 *
 * DuckDB itself is not available in this environment, so neither the reference extension (compiled unchanged
 * from /root/reference/src by oracle/Makefile -> oracle/_ref/libduckhts_ref.so) nor ours
 * (duckhts_b200/lib/duckhts_b200.duckdb_extension.so) can be LOADed by a real engine.  This program plays the
 * engine's part for exactly the calls those table functions make (last row): it hands the
 * extension a duckdb_ext_api_v1 table (include/duckdb_ext_abi.h, slot order of the stable v1.2.0 API), lets it
 * register its functions, then drives  bind -> global init -> local init -> scan*  the way DuckDB does and
 * prints every cell.  Running both extensions under it on the same file and diffing the output is the
 * cell-for-cell parity check; --time turns it into the CPU baseline "reference extension under the fake host".
 *
 *   fake_duckdb EXT.so FUNCTION PATH [name=value ...] [--cols 0,3,5] [--threads N] [--dump | --hash | --time]
 *
 * Output (--dump): one header line of column names, then one TSV line per row; NULL = \N, lists = [a,b],
 * maps = {k=v,...}.  --hash prints per-column FNV-1a hashes of the same text plus the row count.
 * Rows come out in scan order; with --threads 1 (default) that order is deterministic.
 */
#define _GNU_SOURCE
#include <dlfcn.h>
#include <inttypes.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "duckdb_ext_abi.h"

#define VSIZE 2048

/* ---- engine-side objects behind the opaque handles ------------------------------------------------------------ */
typedef struct LType {
    duckdb_type id;
    struct LType *child;          /* LIST element; MAP: STRUCT(key,value) */
    struct LType **members;       /* STRUCT */
    char **member_names;
    idx_t n_members;
} LType;

typedef struct Vec {
    LType *type;
    void *data;
    uint64_t *validity;           /* NULL = all valid (until ensure_validity_writable) */
    idx_t capacity;
    struct Vec *child;            /* LIST */
    idx_t list_size;
    struct Vec **members;         /* STRUCT */
    /* strings copied by assign_string_element live in an arena, like DuckDB's StringHeap (ArenaAllocator): a bump
     * pointer over a list of blocks, rewound when the chunk is reset */
    char **heap;                  /* arena blocks */
    size_t *heap_cap;
    size_t n_heap, cap_heap;      /* blocks allocated / slots in the block list */
    size_t cur_blk, cur_off;      /* bump position */
} Vec;

typedef struct { Vec **vecs; idx_t ncols, size; } Chunk;

typedef struct Val {
    duckdb_type id;
    int is_null;
    char *s;
    int64_t i;
    struct Val **items;           /* LIST: name=a,b,c on the command line */
    idx_t n_items;
} Val;

typedef struct { char *name; LType *type; } NamedParam;

typedef struct TFunc {
    char *name;
    LType **params; idx_t n_params;
    NamedParam *named; idx_t n_named;
    duckdb_table_function_bind_t bind;
    duckdb_table_function_init_t init, local_init;
    duckdb_table_function_t func;
    int pushdown;
} TFunc;

typedef struct {
    TFunc *tf;
    Val **pos; idx_t n_pos;
    char **named_keys; Val **named_vals; idx_t n_named;
    char **col_names; LType **col_types; idx_t n_cols;
    void *bind_data; duckdb_delete_callback_t bind_free;
    char *error;
} BindInfo;

typedef struct {
    BindInfo *bind;
    void *init_data; duckdb_delete_callback_t init_free;
    idx_t max_threads;
    idx_t *col_ids; idx_t n_col_ids;
    char *error;
} InitInfo;

typedef struct {
    BindInfo *bind;
    InitInfo *global, *local;
    char *error;
} FuncInfo;

static TFunc *g_funcs[64];
static int g_nfuncs;
static char *g_ext_error;

static void *xmalloc(size_t n) { void *p = calloc(1, n ? n : 1); if (!p) { fprintf(stderr, "out of memory\n"); exit(2); } return p; }
static char *xstrdup(const char *s) { char *d = xmalloc(strlen(s) + 1); strcpy(d, s); return d; }

/* ---- logical types ---------------------------------------------------------------------------------------------- */
static LType *ltype_copy(const LType *t) {
    if (!t) return NULL;
    LType *c = xmalloc(sizeof *c);
    c->id = t->id;
    c->child = ltype_copy(t->child);
    c->n_members = t->n_members;
    if (t->n_members) {
        c->members = xmalloc(sizeof(LType *) * t->n_members);
        c->member_names = xmalloc(sizeof(char *) * t->n_members);
        for (idx_t i = 0; i < t->n_members; ++i) { c->members[i] = ltype_copy(t->members[i]); c->member_names[i] = xstrdup(t->member_names[i]); }
    }
    return c;
}
static void ltype_free(LType *t) {
    if (!t) return;
    ltype_free(t->child);
    for (idx_t i = 0; i < t->n_members; ++i) { ltype_free(t->members[i]); free(t->member_names[i]); }
    free(t->members); free(t->member_names); free(t);
}
static duckdb_logical_type api_create_logical_type(duckdb_type id) { LType *t = xmalloc(sizeof *t); t->id = id; return (duckdb_logical_type)t; }
static duckdb_logical_type api_create_list_type(duckdb_logical_type child) {
    LType *t = xmalloc(sizeof *t); t->id = DUCKDB_TYPE_LIST; t->child = ltype_copy((LType *)child); return (duckdb_logical_type)t;
}
static duckdb_logical_type api_create_struct_type(duckdb_logical_type *mt, const char **mn, idx_t n) {
    LType *t = xmalloc(sizeof *t); t->id = DUCKDB_TYPE_STRUCT; t->n_members = n;
    t->members = xmalloc(sizeof(LType *) * n); t->member_names = xmalloc(sizeof(char *) * n);
    for (idx_t i = 0; i < n; ++i) { t->members[i] = ltype_copy((LType *)mt[i]); t->member_names[i] = xstrdup(mn[i]); }
    return (duckdb_logical_type)t;
}
static duckdb_logical_type api_create_map_type(duckdb_logical_type k, duckdb_logical_type v) {
    duckdb_logical_type mt[2] = {k, v}; const char *mn[2] = {"key", "value"};
    LType *st = (LType *)api_create_struct_type(mt, mn, 2);
    LType *t = xmalloc(sizeof *t); t->id = DUCKDB_TYPE_MAP; t->child = st; return (duckdb_logical_type)t;
}
static void api_destroy_logical_type(duckdb_logical_type *t) { if (t && *t) { ltype_free((LType *)*t); *t = NULL; } }
static duckdb_type api_get_type_id(duckdb_logical_type t) { return t ? ((LType *)t)->id : DUCKDB_TYPE_INVALID; }

static size_t type_width(const LType *t) {
    switch (t->id) {
    case DUCKDB_TYPE_BOOLEAN: case DUCKDB_TYPE_TINYINT: case DUCKDB_TYPE_UTINYINT: return 1;
    case DUCKDB_TYPE_SMALLINT: case DUCKDB_TYPE_USMALLINT: return 2;
    case DUCKDB_TYPE_INTEGER: case DUCKDB_TYPE_UINTEGER: case DUCKDB_TYPE_FLOAT: return 4;
    case DUCKDB_TYPE_BIGINT: case DUCKDB_TYPE_UBIGINT: case DUCKDB_TYPE_DOUBLE: return 8;
    case DUCKDB_TYPE_VARCHAR: case DUCKDB_TYPE_BLOB: case DUCKDB_TYPE_LIST: case DUCKDB_TYPE_MAP: return 16;
    default: return 0;
    }
}

/* ---- vectors ----------------------------------------------------------------------------------------------------- */
static Vec *vec_new(LType *t, idx_t cap) {
    Vec *v = xmalloc(sizeof *v);
    v->type = t; v->capacity = cap;
    size_t w = type_width(t);
    if (w) v->data = xmalloc(w * cap);
    if (t->id == DUCKDB_TYPE_LIST || t->id == DUCKDB_TYPE_MAP) v->child = vec_new(t->child, VSIZE);
    if (t->id == DUCKDB_TYPE_STRUCT) {
        v->members = xmalloc(sizeof(Vec *) * t->n_members);
        for (idx_t i = 0; i < t->n_members; ++i) v->members[i] = vec_new(t->members[i], cap);
    }
    return v;
}
static void vec_reserve(Vec *v, idx_t cap) {
    if (cap <= v->capacity) return;
    idx_t nc = v->capacity; while (nc < cap) nc *= 2;
    size_t w = type_width(v->type);
    if (w) { v->data = realloc(v->data, w * nc); memset((char *)v->data + w * v->capacity, 0, w * (nc - v->capacity)); }
    if (v->validity) {
        v->validity = realloc(v->validity, ((nc + 63) / 64) * 8);
        for (idx_t i = (v->capacity + 63) / 64; i < (nc + 63) / 64; ++i) v->validity[i] = ~0ull;
    }
    if (v->type->id == DUCKDB_TYPE_STRUCT) for (idx_t i = 0; i < v->type->n_members; ++i) vec_reserve(v->members[i], nc);
    v->capacity = nc;
}
static void vec_reset(Vec *v) {
    free(v->validity); v->validity = NULL;
    v->cur_blk = 0; v->cur_off = 0;      /* rewind the string arena; the blocks are kept for the next chunk */
    v->list_size = 0;
    if (v->child) vec_reset(v->child);
    if (v->members) for (idx_t i = 0; i < v->type->n_members; ++i) vec_reset(v->members[i]);
}
static void vec_free(Vec *v) {
    if (!v) return;
    vec_reset(v);
    for (size_t i = 0; i < v->n_heap; ++i) free(v->heap[i]);
    free(v->heap); free(v->heap_cap); free(v->data);
    vec_free(v->child);
    if (v->members) { for (idx_t i = 0; i < v->type->n_members; ++i) vec_free(v->members[i]); free(v->members); }
    free(v);
}
static idx_t api_vector_size(void) { return VSIZE; }
static void *api_vector_get_data(duckdb_vector v) { return ((Vec *)v)->data; }
static uint64_t *api_vector_get_validity(duckdb_vector v) { return ((Vec *)v)->validity; }
static void api_vector_ensure_validity_writable(duckdb_vector vv) {
    Vec *v = (Vec *)vv;
    if (v->validity) return;
    idx_t nw = (v->capacity + 63) / 64;
    v->validity = xmalloc(nw * 8);
    for (idx_t i = 0; i < nw; ++i) v->validity[i] = ~0ull;
}
static duckdb_logical_type api_vector_get_column_type(duckdb_vector v) { return (duckdb_logical_type)ltype_copy(((Vec *)v)->type); }
static void api_validity_set_row_invalid(uint64_t *validity, idx_t row) { if (validity) validity[row / 64] &= ~(1ull << (row % 64)); }
static bool api_validity_row_is_valid(uint64_t *validity, idx_t row) { return !validity || ((validity[row / 64] >> (row % 64)) & 1); }
#define ARENA_BLOCK (256u << 10)
static char *arena_alloc(Vec *v, size_t len) {
    while (v->cur_blk < v->n_heap && v->cur_off + len > v->heap_cap[v->cur_blk]) { ++v->cur_blk; v->cur_off = 0; }
    if (v->cur_blk == v->n_heap) {
        if (v->n_heap == v->cap_heap) {
            v->cap_heap = v->cap_heap ? v->cap_heap * 2 : 8;
            v->heap = realloc(v->heap, v->cap_heap * sizeof(char *));
            v->heap_cap = realloc(v->heap_cap, v->cap_heap * sizeof(size_t));
        }
        const size_t cap = len > ARENA_BLOCK ? len : ARENA_BLOCK;
        v->heap[v->n_heap] = malloc(cap);
        if (!v->heap[v->n_heap]) { fprintf(stderr, "out of memory\n"); exit(2); }
        v->heap_cap[v->n_heap++] = cap;
        v->cur_off = 0;
    }
    char *p = v->heap[v->cur_blk] + v->cur_off;
    v->cur_off += len;
    return p;
}
static void api_assign_string_len(duckdb_vector vv, idx_t index, const char *str, idx_t len) {
    Vec *v = (Vec *)vv;
    duckdb_string_t *cell = (duckdb_string_t *)v->data + index;
    memset(cell, 0, sizeof *cell);
    cell->value.inlined.length = (uint32_t)len;
    if (len <= 12) { memcpy(cell->value.inlined.inlined, str, len); return; }
    char *copy = arena_alloc(v, len);
    memcpy(copy, str, len);
    memcpy(cell->value.pointer.prefix, str, 4);
    cell->value.pointer.ptr = copy;
}
static void api_assign_string(duckdb_vector v, idx_t index, const char *str) { api_assign_string_len(v, index, str, strlen(str)); }
static duckdb_vector api_list_get_child(duckdb_vector v) { return (duckdb_vector)((Vec *)v)->child; }
static idx_t api_list_get_size(duckdb_vector v) { return ((Vec *)v)->list_size; }
/* set_size also grows the child: the reference's FILTER writer calls it without duckdb_list_vector_reserve
 * (src/bcf_reader.c:1446,1451), which only works in DuckDB while the child fits its initial capacity */
static duckdb_state api_list_set_size(duckdb_vector v, idx_t n) { vec_reserve(((Vec *)v)->child, n); ((Vec *)v)->list_size = n; return DuckDBSuccess; }
static duckdb_state api_list_reserve(duckdb_vector v, idx_t n) { vec_reserve(((Vec *)v)->child, n); return DuckDBSuccess; }
static duckdb_vector api_struct_get_child(duckdb_vector v, idx_t i) { return (duckdb_vector)((Vec *)v)->members[i]; }
static uint32_t api_string_t_length(duckdb_string_t s) { return s.value.inlined.length; }
static const char *api_string_t_data(duckdb_string_t *s) { return s->value.inlined.length <= 12 ? s->value.inlined.inlined : s->value.pointer.ptr; }

static duckdb_vector api_chunk_get_vector(duckdb_data_chunk c, idx_t i) { return (duckdb_vector)((Chunk *)c)->vecs[i]; }
static void api_chunk_set_size(duckdb_data_chunk c, idx_t n) { ((Chunk *)c)->size = n; }
static idx_t api_chunk_get_size(duckdb_data_chunk c) { return ((Chunk *)c)->size; }
static idx_t api_chunk_get_column_count(duckdb_data_chunk c) { return ((Chunk *)c)->ncols; }

/* ---- values ------------------------------------------------------------------------------------------------------ */
static Val *val_null(void) { Val *v = xmalloc(sizeof *v); v->is_null = 1; v->id = DUCKDB_TYPE_INVALID; return v; }
static Val *val_copy(const Val *s) {
    Val *v = xmalloc(sizeof *v);
    *v = *s;
    if (s->s) v->s = xstrdup(s->s);
    if (s->n_items) { v->items = xmalloc(sizeof(Val *) * s->n_items); for (idx_t i = 0; i < s->n_items; ++i) v->items[i] = val_copy(s->items[i]); }
    return v;
}
static void val_free(Val *x) { if (!x) return; for (idx_t i = 0; i < x->n_items; ++i) val_free(x->items[i]); free(x->items); free(x->s); free(x); }
static void api_destroy_value(duckdb_value *v) { if (v && *v) { val_free((Val *)*v); *v = NULL; } }
static bool api_is_null_value(duckdb_value v) { return !v || ((Val *)v)->is_null; }
static char *api_get_varchar(duckdb_value vv) {      /* caller frees with duckdb_free */
    Val *v = (Val *)vv;
    if (!v || v->is_null) return NULL;
    char buf[64];
    const char *s = v->s;
    if (!s) { if (v->id == DUCKDB_TYPE_BOOLEAN) s = v->i ? "true" : "false"; else { snprintf(buf, sizeof buf, "%" PRId64, v->i); s = buf; } }
    char *d = malloc(strlen(s) + 1);
    strcpy(d, s);
    return d;
}
static bool api_get_bool(duckdb_value vv) {
    Val *v = (Val *)vv;
    if (!v || v->is_null) return false;
    if (v->s) return !strcasecmp(v->s, "true") || !strcmp(v->s, "1") || !strcasecmp(v->s, "t");
    return v->i != 0;
}
static int64_t api_get_int64(duckdb_value vv) { Val *v = (Val *)vv; if (!v || v->is_null) return 0; return v->s ? strtoll(v->s, NULL, 10) : v->i; }
static duckdb_logical_type api_get_value_type(duckdb_value vv) { return api_create_logical_type(vv ? ((Val *)vv)->id : DUCKDB_TYPE_INVALID); }
static idx_t api_get_list_size(duckdb_value v) { return v ? ((Val *)v)->n_items : 0; }
static duckdb_value api_get_list_child(duckdb_value v, idx_t i) {
    Val *x = (Val *)v;
    return (duckdb_value)(x && i < x->n_items ? val_copy(x->items[i]) : val_null());
}

/* ---- table functions ---------------------------------------------------------------------------------------------- */
static duckdb_table_function api_create_table_function(void) { return (duckdb_table_function)xmalloc(sizeof(TFunc)); }
static void api_destroy_table_function(duckdb_table_function *tf) { if (tf) *tf = NULL; /* registered copies stay alive */ }
static void api_tf_set_name(duckdb_table_function tf, const char *name) { ((TFunc *)tf)->name = xstrdup(name); }
static void api_tf_add_parameter(duckdb_table_function tfh, duckdb_logical_type t) {
    TFunc *tf = (TFunc *)tfh;
    tf->params = realloc(tf->params, sizeof(LType *) * (tf->n_params + 1));
    tf->params[tf->n_params++] = ltype_copy((LType *)t);
}
static void api_tf_add_named_parameter(duckdb_table_function tfh, const char *name, duckdb_logical_type t) {
    TFunc *tf = (TFunc *)tfh;
    tf->named = realloc(tf->named, sizeof(NamedParam) * (tf->n_named + 1));
    tf->named[tf->n_named].name = xstrdup(name);
    tf->named[tf->n_named++].type = ltype_copy((LType *)t);
}
static void api_tf_set_bind(duckdb_table_function tf, duckdb_table_function_bind_t f) { ((TFunc *)tf)->bind = f; }
static void api_tf_set_init(duckdb_table_function tf, duckdb_table_function_init_t f) { ((TFunc *)tf)->init = f; }
static void api_tf_set_local_init(duckdb_table_function tf, duckdb_table_function_init_t f) { ((TFunc *)tf)->local_init = f; }
static void api_tf_set_function(duckdb_table_function tf, duckdb_table_function_t f) { ((TFunc *)tf)->func = f; }
static void api_tf_pushdown(duckdb_table_function tf, bool p) { ((TFunc *)tf)->pushdown = p; }
static duckdb_state api_register_table_function(duckdb_connection con, duckdb_table_function tfh) {
    (void)con;
    TFunc *tf = (TFunc *)tfh;
    if (!tf || !tf->name || !tf->bind || !tf->init || !tf->func || g_nfuncs >= 64) return DuckDBError;
    g_funcs[g_nfuncs++] = tf;
    return DuckDBSuccess;
}

static void set_err(char **slot, const char *msg) { free(*slot); *slot = xstrdup(msg ? msg : "(null)"); }
static duckdb_value api_bind_get_parameter(duckdb_bind_info bi, idx_t i) {
    BindInfo *b = (BindInfo *)bi;
    return (duckdb_value)(i < b->n_pos ? val_copy(b->pos[i]) : val_null());
}
static duckdb_value api_bind_get_named_parameter(duckdb_bind_info bi, const char *name) {
    BindInfo *b = (BindInfo *)bi;
    for (idx_t i = 0; i < b->n_named; ++i)
        if (!strcmp(b->named_keys[i], name)) return (duckdb_value)val_copy(b->named_vals[i]);
    return (duckdb_value)val_null();     /* DuckDB hands back a NULL value for a named parameter that was not given */
}
static void api_bind_add_result_column(duckdb_bind_info bi, const char *name, duckdb_logical_type t) {
    BindInfo *b = (BindInfo *)bi;
    b->col_names = realloc(b->col_names, sizeof(char *) * (b->n_cols + 1));
    b->col_types = realloc(b->col_types, sizeof(LType *) * (b->n_cols + 1));
    b->col_names[b->n_cols] = xstrdup(name);
    b->col_types[b->n_cols++] = ltype_copy((LType *)t);
}
static void api_bind_set_bind_data(duckdb_bind_info bi, void *d, duckdb_delete_callback_t f) { ((BindInfo *)bi)->bind_data = d; ((BindInfo *)bi)->bind_free = f; }
static void api_bind_set_cardinality(duckdb_bind_info bi, idx_t c, bool e) { (void)bi; (void)c; (void)e; }
static void api_bind_set_error(duckdb_bind_info bi, const char *e) { set_err(&((BindInfo *)bi)->error, e); }
static void *api_init_get_bind_data(duckdb_init_info ii) { return ((InitInfo *)ii)->bind->bind_data; }
static void api_init_set_init_data(duckdb_init_info ii, void *d, duckdb_delete_callback_t f) { ((InitInfo *)ii)->init_data = d; ((InitInfo *)ii)->init_free = f; }
static idx_t api_init_get_column_count(duckdb_init_info ii) { return ((InitInfo *)ii)->n_col_ids; }
static idx_t api_init_get_column_index(duckdb_init_info ii, idx_t i) { return ((InitInfo *)ii)->col_ids[i]; }
static void api_init_set_max_threads(duckdb_init_info ii, idx_t n) { ((InitInfo *)ii)->max_threads = n; }
static void api_init_set_error(duckdb_init_info ii, const char *e) { set_err(&((InitInfo *)ii)->error, e); }
static void *api_function_get_bind_data(duckdb_function_info fi) { return ((FuncInfo *)fi)->bind->bind_data; }
static void *api_function_get_init_data(duckdb_function_info fi) { return ((FuncInfo *)fi)->global->init_data; }
static void *api_function_get_local_init_data(duckdb_function_info fi) { return ((FuncInfo *)fi)->local ? ((FuncInfo *)fi)->local->init_data : NULL; }
static void api_function_set_error(duckdb_function_info fi, const char *e) { set_err(&((FuncInfo *)fi)->error, e); }

/* ---- connection / misc -------------------------------------------------------------------------------------------- */
static struct _duckdb_database g_db_obj;
static duckdb_database g_db = &g_db_obj;
static struct _duckdb_connection g_con_obj;
static duckdb_state api_connect(duckdb_database db, duckdb_connection *out) { (void)db; *out = &g_con_obj; return DuckDBSuccess; }
static void api_disconnect(duckdb_connection *c) { if (c) *c = NULL; }
static void *api_malloc(size_t n) { return malloc(n ? n : 1); }
static void api_free(void *p) { free(p); }
/* out-of-scope registration calls of the reference's entry point (scalar UDFs, SQL macros) are accepted and ignored */
static void *stub_create(void) { return xmalloc(16); }
static void stub_destroy(void **p) { if (p) { free(*p); *p = NULL; } }
static void stub_void(void) {}
static duckdb_state stub_ok(void) { return DuckDBSuccess; }
static duckdb_state stub_query(duckdb_connection c, const char *sql, void *res) { (void)c; (void)sql; (void)res; return DuckDBError; }

static duckdb_ext_api_v1 g_api;
static void ext_set_error(duckdb_extension_info info, const char *e) { (void)info; set_err(&g_ext_error, e); }
static duckdb_database *ext_get_database(duckdb_extension_info info) { (void)info; return &g_db; }
static const void *ext_get_api(duckdb_extension_info info, const char *version) {
    (void)info;
    if (strcmp(version, DUCKDB_EXT_ABI_VERSION)) { fprintf(stderr, "fake_duckdb: extension asks for API %s, host serves %s\n", version, DUCKDB_EXT_ABI_VERSION); return NULL; }
    return &g_api;
}

static void fill_api(void) {
    duckdb_ext_api_v1 *a = &g_api;
    memset(a, 0, sizeof *a);
    a->duckdb_connect = api_connect; a->duckdb_disconnect = api_disconnect;
    a->duckdb_malloc = api_malloc; a->duckdb_free = api_free; a->duckdb_vector_size = api_vector_size;
    a->duckdb_string_t_length = api_string_t_length; a->duckdb_string_t_data = api_string_t_data;
    a->duckdb_destroy_value = api_destroy_value; a->duckdb_get_bool = api_get_bool; a->duckdb_get_int64 = api_get_int64;
    a->duckdb_get_value_type = api_get_value_type; a->duckdb_get_varchar = api_get_varchar; a->duckdb_is_null_value = api_is_null_value;
    a->duckdb_get_list_size = api_get_list_size; a->duckdb_get_list_child = api_get_list_child;
    a->duckdb_create_logical_type = api_create_logical_type; a->duckdb_create_list_type = api_create_list_type;
    a->duckdb_create_map_type = api_create_map_type; a->duckdb_create_struct_type = api_create_struct_type;
    a->duckdb_get_type_id = api_get_type_id; a->duckdb_destroy_logical_type = api_destroy_logical_type;
    a->duckdb_data_chunk_get_column_count = api_chunk_get_column_count; a->duckdb_data_chunk_get_vector = api_chunk_get_vector;
    a->duckdb_data_chunk_get_size = api_chunk_get_size; a->duckdb_data_chunk_set_size = api_chunk_set_size;
    a->duckdb_vector_get_column_type = api_vector_get_column_type; a->duckdb_vector_get_data = api_vector_get_data;
    a->duckdb_vector_get_validity = api_vector_get_validity; a->duckdb_vector_ensure_validity_writable = api_vector_ensure_validity_writable;
    a->duckdb_vector_assign_string_element = api_assign_string; a->duckdb_vector_assign_string_element_len = api_assign_string_len;
    a->duckdb_list_vector_get_child = api_list_get_child; a->duckdb_list_vector_get_size = api_list_get_size;
    a->duckdb_list_vector_set_size = api_list_set_size; a->duckdb_list_vector_reserve = api_list_reserve;
    a->duckdb_struct_vector_get_child = api_struct_get_child;
    a->duckdb_validity_row_is_valid = api_validity_row_is_valid; a->duckdb_validity_set_row_invalid = api_validity_set_row_invalid;
    a->duckdb_create_table_function = api_create_table_function; a->duckdb_destroy_table_function = api_destroy_table_function;
    a->duckdb_table_function_set_name = api_tf_set_name; a->duckdb_table_function_add_parameter = api_tf_add_parameter;
    a->duckdb_table_function_add_named_parameter = api_tf_add_named_parameter; a->duckdb_table_function_set_bind = api_tf_set_bind;
    a->duckdb_table_function_set_init = api_tf_set_init; a->duckdb_table_function_set_local_init = api_tf_set_local_init;
    a->duckdb_table_function_set_function = api_tf_set_function; a->duckdb_table_function_supports_projection_pushdown = api_tf_pushdown;
    a->duckdb_register_table_function = api_register_table_function;
    a->duckdb_bind_add_result_column = api_bind_add_result_column; a->duckdb_bind_get_parameter = api_bind_get_parameter;
    a->duckdb_bind_get_named_parameter = api_bind_get_named_parameter; a->duckdb_bind_set_bind_data = api_bind_set_bind_data;
    a->duckdb_bind_set_cardinality = api_bind_set_cardinality; a->duckdb_bind_set_error = api_bind_set_error;
    a->duckdb_init_get_bind_data = api_init_get_bind_data; a->duckdb_init_set_init_data = api_init_set_init_data;
    a->duckdb_init_get_column_count = api_init_get_column_count; a->duckdb_init_get_column_index = api_init_get_column_index;
    a->duckdb_init_set_max_threads = api_init_set_max_threads; a->duckdb_init_set_error = api_init_set_error;
    a->duckdb_function_get_bind_data = api_function_get_bind_data; a->duckdb_function_get_init_data = api_function_get_init_data;
    a->duckdb_function_get_local_init_data = api_function_get_local_init_data; a->duckdb_function_set_error = api_function_set_error;
    /* ignored registrations */
    a->duckdb_query = (void *)stub_query; a->duckdb_destroy_result = (void *)stub_void;
    a->duckdb_create_scalar_function = (void *)stub_create; a->duckdb_destroy_scalar_function = (void *)stub_destroy;
    a->duckdb_scalar_function_set_name = (void *)stub_void; a->duckdb_scalar_function_set_varargs = (void *)stub_void;
    a->duckdb_scalar_function_set_special_handling = (void *)stub_void; a->duckdb_scalar_function_set_volatile = (void *)stub_void;
    a->duckdb_scalar_function_add_parameter = (void *)stub_void; a->duckdb_scalar_function_set_return_type = (void *)stub_void;
    a->duckdb_scalar_function_set_extra_info = (void *)stub_void; a->duckdb_scalar_function_set_function = (void *)stub_void;
    a->duckdb_register_scalar_function = (void *)stub_ok; a->duckdb_scalar_function_get_extra_info = (void *)stub_create;
    a->duckdb_scalar_function_set_error = (void *)stub_void;
    a->duckdb_create_scalar_function_set = (void *)stub_create; a->duckdb_destroy_scalar_function_set = (void *)stub_destroy;
    a->duckdb_add_scalar_function_to_set = (void *)stub_ok; a->duckdb_register_scalar_function_set = (void *)stub_ok;
}

/* ---- cell formatting --------------------------------------------------------------------------------------------- */
typedef struct { char *p; size_t n, cap; } Buf;
static void buf_put(Buf *b, const char *s, size_t n) {
    if (b->n + n + 1 > b->cap) { b->cap = (b->n + n + 1) * 2; b->p = realloc(b->p, b->cap); }
    memcpy(b->p + b->n, s, n); b->n += n; b->p[b->n] = 0;
}
static void buf_puts(Buf *b, const char *s) { buf_put(b, s, strlen(s)); }
static void buf_esc(Buf *b, const char *s, size_t n) {      /* keep the TSV one row per line */
    for (size_t i = 0; i < n; ++i) {
        unsigned char c = (unsigned char)s[i];
        if (c == '\t') buf_puts(b, "\\t"); else if (c == '\n') buf_puts(b, "\\n"); else if (c == '\\') buf_puts(b, "\\\\");
        else if (c < 32 || c == 127) { char t[8]; snprintf(t, sizeof t, "\\x%02x", c); buf_puts(b, t); }
        else buf_put(b, (const char *)&c, 1);
    }
}
static void format_cell(Buf *b, Vec *v, idx_t row) {
    if (v->validity && !((v->validity[row / 64] >> (row % 64)) & 1)) { buf_puts(b, "\\N"); return; }
    char t[64];
    switch (v->type->id) {
    case DUCKDB_TYPE_BOOLEAN: buf_puts(b, ((uint8_t *)v->data)[row] ? "true" : "false"); break;
    case DUCKDB_TYPE_TINYINT: snprintf(t, sizeof t, "%d", ((int8_t *)v->data)[row]); buf_puts(b, t); break;
    case DUCKDB_TYPE_UTINYINT: snprintf(t, sizeof t, "%u", ((uint8_t *)v->data)[row]); buf_puts(b, t); break;
    case DUCKDB_TYPE_SMALLINT: snprintf(t, sizeof t, "%d", ((int16_t *)v->data)[row]); buf_puts(b, t); break;
    case DUCKDB_TYPE_USMALLINT: snprintf(t, sizeof t, "%u", ((uint16_t *)v->data)[row]); buf_puts(b, t); break;
    case DUCKDB_TYPE_INTEGER: snprintf(t, sizeof t, "%d", ((int32_t *)v->data)[row]); buf_puts(b, t); break;
    case DUCKDB_TYPE_UINTEGER: snprintf(t, sizeof t, "%u", ((uint32_t *)v->data)[row]); buf_puts(b, t); break;
    case DUCKDB_TYPE_BIGINT: snprintf(t, sizeof t, "%" PRId64, ((int64_t *)v->data)[row]); buf_puts(b, t); break;
    case DUCKDB_TYPE_UBIGINT: snprintf(t, sizeof t, "%" PRIu64, ((uint64_t *)v->data)[row]); buf_puts(b, t); break;
    case DUCKDB_TYPE_FLOAT: snprintf(t, sizeof t, "%.9g", ((float *)v->data)[row]); buf_puts(b, t); break;
    case DUCKDB_TYPE_DOUBLE: snprintf(t, sizeof t, "%.17g", ((double *)v->data)[row]); buf_puts(b, t); break;
    case DUCKDB_TYPE_VARCHAR: case DUCKDB_TYPE_BLOB: {
        duckdb_string_t *s = (duckdb_string_t *)v->data + row;
        buf_esc(b, api_string_t_data(s), s->value.inlined.length);
        break;
    }
    case DUCKDB_TYPE_LIST: case DUCKDB_TYPE_MAP: {
        duckdb_list_entry e = ((duckdb_list_entry *)v->data)[row];
        buf_puts(b, v->type->id == DUCKDB_TYPE_MAP ? "{" : "[");
        for (uint64_t i = 0; i < e.length; ++i) {
            if (i) buf_puts(b, ",");
            if (v->type->id == DUCKDB_TYPE_MAP) {
                format_cell(b, v->child->members[0], e.offset + i); buf_puts(b, "="); format_cell(b, v->child->members[1], e.offset + i);
            } else format_cell(b, v->child, e.offset + i);
        }
        buf_puts(b, v->type->id == DUCKDB_TYPE_MAP ? "}" : "]");
        break;
    }
    case DUCKDB_TYPE_STRUCT:
        buf_puts(b, "(");
        for (idx_t i = 0; i < v->type->n_members; ++i) { if (i) buf_puts(b, ","); format_cell(b, v->members[i], row); }
        buf_puts(b, ")");
        break;
    default: buf_puts(b, "?"); break;
    }
}

/* ---- driver -------------------------------------------------------------------------------------------------------- */
enum { MODE_DUMP, MODE_HASH, MODE_TIME };
typedef struct {
    TFunc *tf; BindInfo *bind; InitInfo *global;
    idx_t *col_ids; idx_t n_col_ids;
    int mode;
    pthread_mutex_t *out_lock;
    uint64_t rows; uint64_t *hash, *mhash; char *error;
} Worker;

static double now_s(void) { struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec + ts.tv_nsec * 1e-9; }

static void *worker_main(void *arg) {
    Worker *w = (Worker *)arg;
    InitInfo local = {0};
    local.bind = w->bind; local.col_ids = w->col_ids; local.n_col_ids = w->n_col_ids;
    if (w->tf->local_init) {
        w->tf->local_init((duckdb_init_info)&local);
        if (local.error) { w->error = local.error; return NULL; }
    }
    Chunk ch = {0};
    ch.ncols = w->n_col_ids;
    ch.vecs = xmalloc(sizeof(Vec *) * (ch.ncols ? ch.ncols : 1));
    for (idx_t i = 0; i < ch.ncols; ++i) ch.vecs[i] = vec_new(w->bind->col_types[w->col_ids[i]], VSIZE);
    FuncInfo fi = {0};
    fi.bind = w->bind; fi.global = w->global; fi.local = &local;
    Buf line = {0};
    for (;;) {
        ch.size = 0;
        for (idx_t i = 0; i < ch.ncols; ++i) vec_reset(ch.vecs[i]);
        w->tf->func((duckdb_function_info)&fi, (duckdb_data_chunk)&ch);
        if (fi.error) { w->error = fi.error; break; }
        if (ch.size == 0) break;
        w->rows += ch.size;
        if (w->mode == MODE_TIME) continue;
        if (w->mode == MODE_DUMP) pthread_mutex_lock(w->out_lock);
        for (idx_t r = 0; r < ch.size; ++r) {
            line.n = 0;
            for (idx_t c = 0; c < ch.ncols; ++c) {
                size_t start = line.n;
                format_cell(&line, ch.vecs[c], r);
                if (w->mode == MODE_HASH) {
                    uint64_t h = w->hash[c], m = 1469598103934665603ull;
                    for (size_t k = start; k < line.n; ++k) {
                        h ^= (unsigned char)line.p[k]; h *= 1099511628211ull;
                        m ^= (unsigned char)line.p[k]; m *= 1099511628211ull;
                    }
                    h ^= 0xff; h *= 1099511628211ull;       /* cell separator */
                    w->hash[c] = h;
                    m ^= m >> 30; m *= 0xbf58476d1ce4e5b9ull; m ^= m >> 27; m *= 0x94d049bb133111ebull; m ^= m >> 31;
                    w->mhash[c] += m;                        /* multiset hash: independent of row order and of the thread that saw the row */
                } else if (c + 1 < ch.ncols) buf_puts(&line, "\t");
            }
            if (w->mode == MODE_DUMP) { fwrite(line.p, 1, line.n, stdout); fputc('\n', stdout); }
        }
        if (w->mode == MODE_DUMP) pthread_mutex_unlock(w->out_lock);
    }
    for (idx_t i = 0; i < ch.ncols; ++i) vec_free(ch.vecs[i]);
    free(ch.vecs); free(line.p);
    if (local.init_free && local.init_data) local.init_free(local.init_data);
    return NULL;
}

static Val *parse_value(const char *text, const LType *t) {
    Val *v = xmalloc(sizeof *v);
    v->id = t ? t->id : DUCKDB_TYPE_VARCHAR;
    if (!strcmp(text, "NULL")) { v->is_null = 1; return v; }
    if (v->id == DUCKDB_TYPE_LIST) {
        char *dup = xstrdup(text), *save = NULL;
        idx_t cap = 1;
        for (const char *q = text; *q; ++q) cap += *q == ',';
        v->items = xmalloc(sizeof(Val *) * cap);
        for (char *tok = strtok_r(dup, ",", &save); tok; tok = strtok_r(NULL, ",", &save)) v->items[v->n_items++] = parse_value(tok, t->child);
        free(dup);
        return v;
    }
    if (v->id == DUCKDB_TYPE_BOOLEAN) v->i = !strcasecmp(text, "true") || !strcmp(text, "1");
    else if (v->id == DUCKDB_TYPE_VARCHAR) v->s = xstrdup(text);
    else v->i = strtoll(text, NULL, 10);
    return v;
}

int main(int argc, char **argv) {
    if (argc < 4) {
        fprintf(stderr, "usage: %s EXT.so FUNCTION PATH [name=value ...] [--cols i,j,..] [--threads N] [--dump|--hash|--time] [--repeat N] [--entry SYMBOL]\n", argv[0]);
        return 2;
    }
    const char *ext = argv[1], *fname = argv[2], *entry = NULL;
    int mode = MODE_DUMP, threads = 1;
    const char *cols = NULL;
    fill_api();
    void *h = dlopen(ext, RTLD_NOW | RTLD_LOCAL);
    if (!h) { fprintf(stderr, "dlopen: %s\n", dlerror()); return 2; }
    for (int i = 4; i < argc; ++i) if (!strcmp(argv[i], "--entry") && i + 1 < argc) entry = argv[i + 1];
    bool (*init)(duckdb_extension_info, struct duckdb_extension_access *) = NULL;
    const char *cands[] = {entry, "duckhts_init_c_api", "duckhts_b200_init_c_api", NULL};
    for (int i = 0; i < 3 && !init; ++i) if (cands[i]) *(void **)&init = dlsym(h, cands[i]);
    if (!init) { fprintf(stderr, "no *_init_c_api entry point in %s\n", ext); return 2; }
    struct duckdb_extension_access access = {ext_set_error, ext_get_database, ext_get_api};
    struct _duckdb_extension_info info_obj = {0};
    if (!init(&info_obj, &access)) { fprintf(stderr, "extension init failed: %s\n", g_ext_error ? g_ext_error : "(no message)"); return 2; }
    TFunc *tf = NULL;
    for (int i = 0; i < g_nfuncs; ++i) if (!strcmp(g_funcs[i]->name, fname)) tf = g_funcs[i];
    if (!tf) { fprintf(stderr, "table function %s not registered (have:", fname); for (int i = 0; i < g_nfuncs; ++i) fprintf(stderr, " %s", g_funcs[i]->name); fprintf(stderr, ")\n"); return 2; }

    int repeat = 1, want_time = 0, want_types = 0;
    for (int i = 4; i < argc; ++i) if (!strcmp(argv[i], "--repeat") && i + 1 < argc) repeat = atoi(argv[i + 1]);
    if (repeat < 1) repeat = 1;
    int rc = 0;
    double first_s = 0, best_s = 0, first_bind = 0, best_bind = 0;
    for (int run = 0; run < repeat; ++run) {
    BindInfo bind = {0};
    bind.tf = tf;
    bind.pos = xmalloc(sizeof(Val *) * 4);
    bind.pos[bind.n_pos++] = parse_value(argv[3], tf->n_params ? tf->params[0] : NULL);
    bind.named_keys = xmalloc(sizeof(char *) * argc);
    bind.named_vals = xmalloc(sizeof(Val *) * argc);
    for (int i = 4; i < argc; ++i) {
        if (!strcmp(argv[i], "--dump")) mode = MODE_DUMP;
        else if (!strcmp(argv[i], "--hash")) mode = MODE_HASH;
        else if (!strcmp(argv[i], "--time")) { want_time = 1; if (mode != MODE_HASH) mode = MODE_TIME; }
        else if (!strcmp(argv[i], "--cols") && i + 1 < argc) cols = argv[++i];
        else if (!strcmp(argv[i], "--threads") && i + 1 < argc) threads = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--repeat") && i + 1 < argc) ++i;
        else if (!strcmp(argv[i], "--entry") && i + 1 < argc) ++i;
        else if (!strcmp(argv[i], "--types")) want_types = 1;
        else {
            char *eq = strchr(argv[i], '=');
            if (!eq) { fprintf(stderr, "bad argument %s\n", argv[i]); return 2; }
            char *key = strndup(argv[i], (size_t)(eq - argv[i]));
            const LType *t = NULL;
            for (idx_t k = 0; k < tf->n_named; ++k) if (!strcmp(tf->named[k].name, key)) t = tf->named[k].type;
            if (!t) { printf("BIND ERROR\tInvalid named parameter \"%s\" for function %s\n", key, fname); return 1; }
            bind.named_keys[bind.n_named] = key;
            bind.named_vals[bind.n_named++] = parse_value(eq + 1, t);
        }
    }
    double t0 = now_s();
    tf->bind((duckdb_bind_info)&bind);
    const double t_bind = now_s() - t0;
    if (bind.error) { printf("BIND ERROR\t%s\n", bind.error); return 1; }
    idx_t *col_ids = xmalloc(sizeof(idx_t) * (bind.n_cols + 1));
    idx_t n_col_ids = 0;
    if (cols && tf->pushdown) {
        char *dup = xstrdup(cols);
        for (char *tok = strtok(dup, ","); tok; tok = strtok(NULL, ",")) {
            idx_t c = (idx_t)strtoull(tok, NULL, 10);
            if (c >= bind.n_cols) { fprintf(stderr, "column %s out of range (%" PRIu64 " columns)\n", tok, bind.n_cols); return 2; }
            col_ids[n_col_ids++] = c;
        }
    } else for (idx_t c = 0; c < bind.n_cols; ++c) col_ids[n_col_ids++] = c;
    InitInfo global = {0};
    global.bind = &bind; global.col_ids = col_ids; global.n_col_ids = n_col_ids; global.max_threads = 1;
    tf->init((duckdb_init_info)&global);
    if (global.error) { printf("INIT ERROR\t%s\n", global.error); return 1; }
    if (threads < 1) threads = 1;
    if ((idx_t)threads > global.max_threads) threads = (int)global.max_threads;
    if (mode == MODE_DUMP) {
        for (idx_t c = 0; c < n_col_ids; ++c) printf("%s%s", bind.col_names[col_ids[c]], c + 1 < n_col_ids ? "\t" : "\n");
        /* --types: the duckdb_type id of every column (of a LIST's element behind a slash) */
        if (want_types) for (idx_t c = 0; c < n_col_ids; ++c) {
            const LType *t = bind.col_types[col_ids[c]];
            printf("#%d", (int)t->id);
            if (t->child) printf("/%d", (int)t->child->id);
            printf("%s", c + 1 < n_col_ids ? "\t" : "\n");
        }
    }
    pthread_mutex_t lock = PTHREAD_MUTEX_INITIALIZER;
    Worker *ws = xmalloc(sizeof(Worker) * threads);
    pthread_t *th = xmalloc(sizeof(pthread_t) * threads);
    for (int i = 0; i < threads; ++i) {
        ws[i].tf = tf; ws[i].bind = &bind; ws[i].global = &global; ws[i].col_ids = col_ids; ws[i].n_col_ids = n_col_ids;
        ws[i].mode = mode; ws[i].out_lock = &lock;
        ws[i].hash = xmalloc(sizeof(uint64_t) * (n_col_ids + 1));
        ws[i].mhash = xmalloc(sizeof(uint64_t) * (n_col_ids + 1));
        for (idx_t c = 0; c < n_col_ids; ++c) ws[i].hash[c] = 1469598103934665603ull;
        if (threads == 1) worker_main(&ws[i]); else pthread_create(&th[i], NULL, worker_main, &ws[i]);
    }
    uint64_t rows = 0;
    for (int i = 0; i < threads; ++i) {
        if (threads > 1) pthread_join(th[i], NULL);
        rows += ws[i].rows;
        if (ws[i].error) { printf("SCAN ERROR\t%s\n", ws[i].error); rc = 1; }
    }
    double dt = now_s() - t0;
    if (mode == MODE_HASH) {
        printf("rows\t%" PRIu64 "\n", rows);
        /* per column: the chained hash (row order matters; printed for one thread only, where the order is defined) and the
         * multiset hash (sum of per-cell hashes: the same for any row order and any split of the rows over threads) */
        for (idx_t c = 0; c < n_col_ids; ++c) {
            uint64_t msum = 0;
            for (int i = 0; i < threads; ++i) msum += ws[i].mhash[c];
            if (threads == 1) printf("%s\t%016" PRIx64 "\t%016" PRIx64 "\n", bind.col_names[col_ids[c]], ws[0].hash[c], msum);
            else printf("%s\t-\t%016" PRIx64 "\n", bind.col_names[col_ids[c]], msum);
        }
    }
    if (mode == MODE_TIME || want_time) {
        if (run == 0) { first_s = best_s = dt; first_bind = best_bind = t_bind; }
        else if (dt < best_s) { best_s = dt; best_bind = t_bind; }
        if (run + 1 == repeat)
            /* seconds = the first run of this process (cold: what a new session pays); best_seconds = the fastest of --repeat runs in
             * the same process (warm: a session that has run the function before) */
            printf("{\"function\":\"%s\",\"threads\":%d,\"max_threads\":%" PRIu64 ",\"rows\":%" PRIu64 ",\"seconds\":%.6f,\"bind_seconds\":%.6f,"
                   "\"runs\":%d,\"best_seconds\":%.6f,\"best_bind_seconds\":%.6f}\n", fname, threads, global.max_threads, rows, first_s, first_bind, repeat, best_s, best_bind);
    }
    if (global.init_free && global.init_data) global.init_free(global.init_data);
    if (bind.bind_free && bind.bind_data) bind.bind_free(bind.bind_data);
    if (rc) break;
    }
    fflush(stdout);
    return rc;
}
