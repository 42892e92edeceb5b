/* duckhts_cuda.h — C-ABI of libduckhts_cuda.so: the drop-in boundary between host C code (the DuckDB
 * table functions read_bam / bgunzip, or any other FFI caller) and the sm_100a CUDA kernels.
 *
 * Plain C: opaque handles, plain pointers and sizes, no C++/torch types.  Every entry point names the
 * reference interface it replaces (paths relative to the reference tree; "htslib/" =
 * third_party/htslib/).  INTEGRATION.md shows how the reference's src/bam_reader.c and src/bgzip.c
 * would bind these.
 *
 * Threading: a dh_ctx is bound to one CUDA device.  Scan handles (dh_bam, dh_bcf) lock internally, so any number of
 * host threads may call dh_*_acquire_batch / dh_*_release_batch / dh_*_next_batch on ONE handle (the workers of one
 * query); handles that may be driven at the same time (two queries, one handle per GPU) need a dh_ctx each, because
 * streams, ticket counters and staging slots belong to the context.  Several contexts may share a device.
 * Errors: functions return 0 (DH_OK) or a negative DH_E* code; dh_last_error() gives the message of
 * the calling thread's last failure.  There is NO CPU fallback: without a usable CUDA device
 * dh_ctx_create fails and nothing else can be called.
 */
#ifndef DUCKHTS_CUDA_H
#define DUCKHTS_CUDA_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define DH_OK 0
#define DH_EINVAL (-1)  /* bad argument */
#define DH_ECUDA (-2)   /* CUDA runtime error (message in dh_last_error) */
#define DH_EFORMAT (-3) /* malformed BGZF/BAM framing (htslib BGZF_ERR_HEADER / bam_read1 < -1) */
#define DH_EDATA (-4)   /* a block failed inflate / CRC / ISIZE (htslib BGZF_ERR_ZLIB / BGZF_ERR_CRC) */
#define DH_ENOMEM (-5)
#define DH_EIO (-6)
#define DH_ENOREGION (-7) /* no requested region names a reference sequence (sam_itr_regarray returned NULL) */
#define DH_EBUSY (-8)     /* every batch slot of the scan is held by the caller: release one and call again */

typedef struct dh_ctx dh_ctx;

const char *dh_version(void);
const char *dh_last_error(void);
int dh_device_count(void);

/* One context per GPU.  Creates the streams, the pinned staging and the device arenas lazily. */
dh_ctx *dh_ctx_create(int device);
void dh_ctx_destroy(dh_ctx *ctx);
int dh_ctx_sm_count(const dh_ctx *ctx);

/* The library keeps the device and pinned buffers of finished scans (a session runs many; allocating them costs more than a small
 * scan) up to 48 GiB per device and 8 GiB pinned; dh_trim() gives them back.  DH_ALLOC_CACHE=0 in the environment disables it. */
void dh_trim(void);

/* Pinned host memory: buffers handed to the *_host entry points copy at full PCIe rate only when they
 * come from here (pageable memory works, through the driver's bounce buffers). */
void *dh_host_alloc(size_t bytes);
void dh_host_free(void *p);

/* ---- A1: BGZF block framing -----------------------------------------------------------------------
 * Replaces check_header() + the BSIZE walk of bgzf_read_block()/bgzf_mt_read_block()
 * (htslib/bgzf.c:896-903, 1147-1216, 1485-1539).  Host job: the chain is serial (18-byte header per
 * block).  Walks `data[start..size)`; stops at end of data, at `cap` blocks, or at the first invalid
 * header.  Returns the number of blocks written (>= 0) or DH_EFORMAT; *end_off = offset after the
 * last block accepted.  A trailing partial block is reported as DH_EFORMAT only if strict != 0. */
typedef struct {
    uint64_t coffset; /* file offset of the block's first header byte (the BGZF "block address") */
    uint32_t csize;   /* BSIZE: total block size, header and footer included */
    uint32_t isize;   /* ISIZE footer */
    uint32_t crc;     /* CRC32 footer */
    uint32_t reserved;
} dh_bgzf_block;
int64_t dh_bgzf_walk(const uint8_t *data, uint64_t size, uint64_t start, dh_bgzf_block *out, int64_t cap,
                     uint64_t *end_off, int strict);

/* ---- A2/A3: BGZF inflate (+ fused CRC32/ISIZE check) -------------------------------------------------
 * Replaces the bgzf_read() loop of bgzip_bind_common(decompress=1) (src/bgzip.c:154-200) and beneath
 * it bgzf_read_block -> inflate_block -> bgzf_uncompress (htslib/bgzf.c:1004-1239, 808-824, 762-804).
 *
 * dh_bgzf_inflate_host: whole BGZF image in host memory -> inflated bytes in host memory
 * (framing on the host, H2D, inflate kernel, D2H; pipelined over segments).  *out_size receives the
 * inflated length; if out == NULL only the length and per-block verdicts are computed (still inflates).
 * Returns DH_OK, DH_EFORMAT (bad header; *bad_block = index) or DH_EDATA (*bad_block = first failing
 * block, *bad_code = dh::INF_ERR_* code). */
typedef struct {
    uint64_t n_blocks;
    uint64_t comp_bytes;   /* BGZF bytes consumed */
    uint64_t out_bytes;    /* inflated bytes produced */
    int64_t bad_block;     /* -1 if none */
    int32_t bad_code;
    int32_t eof_marker;    /* 1 if the image ends with the 28-byte EOF block (bgzf_check_EOF) */
    float kernel_ms;       /* sum of inflate-kernel durations (CUDA events on the launching streams) */
    float total_ms;        /* first H2D enqueued -> last D2H complete (CUDA events) */
} dh_inflate_stats;
int dh_bgzf_inflate_host(dh_ctx *ctx, const uint8_t *comp, uint64_t comp_size, uint8_t *out, uint64_t out_cap,
                         uint64_t *out_size, dh_inflate_stats *stats);

/* Streaming form, file descriptor to file descriptor — what bgunzip does (src/bgzip.c:154-200: a bgzf_read / write loop
 * over 64 KB buffers; here segments of ~192 MB compressed travel through three pinned staging slots: pread -> H2D ->
 * inflate -> D2H -> pwrite, the reads and writes of one segment overlapping the GPU work of the others).  Neither the
 * input nor the output is ever held whole: pinned memory = the slots.  in_fd must be seekable (pread).  Returns DH_OK,
 * DH_EFORMAT / DH_EDATA (read errors; stats->bad_block), DH_EIO (errno of the failing read / write in the message). */
int dh_bgzf_inflate_fd(dh_ctx *ctx, int in_fd, int out_fd, dh_inflate_stats *stats);

/* ---- N1: BGZF compression (the reference's bgzip(): src/bgzip.c:209-273; bgzf_write / bgzf_close, htslib/bgzf.c:624-683,
 * 1805-1850, 2138-2150).  The input is cut into 0xff00-byte blocks like bgzf_write cuts it; each becomes one BGZF block (LZ77 +
 * dynamic Huffman on the GPU, a stored block where that would not shrink); the 28-byte EOF marker ends the stream.  The
 * DEFLATE bits are this library's own; inflated bytes, block boundaries in inflated space, CRC-32 and ISIZE are what htslib
 * writes.  level: 0 = stored blocks, 1..9 and -1 (default) = the one GPU compressor (bgzip's level is a CPU time knob). */
typedef struct dh_deflate_stats {
    uint64_t in_bytes, out_bytes;   /* out_bytes includes the EOF marker */
    uint64_t n_blocks;              /* data blocks written (without the EOF marker) */
    uint64_t h2d_bytes, d2h_bytes;
    float kernel_ms;                /* sum over launches: compress + pack kernels (CUDA events) */
    float total_ms;
} dh_deflate_stats;
uint64_t dh_bgzf_deflate_bound(uint64_t in_size);      /* an output capacity that always suffices */
int dh_bgzf_deflate_host(dh_ctx *ctx, const uint8_t *in, uint64_t in_size, int level, uint8_t *out, uint64_t out_cap,
                         uint64_t *out_size, dh_deflate_stats *stats);
/* file descriptor to file descriptor, streamed through three staging slots (in_fd seekable); what bgzip() runs */
int dh_bgzf_deflate_fd(dh_ctx *ctx, int in_fd, int out_fd, int level, dh_deflate_stats *stats);

/* ---- N4: the text path (read_tabix's sequential scan, src/tabix_reader.c:852-1030: hts_getline per line, get_field per cell).
 * A bgzip'ed tab-separated text file comes back as batches of whole lines: the inflated text and, for every line and every
 * projected column, the [begin, end) offsets of the field (a '\r' in front of the '\n' is not part of the line, like
 * bgzf_getline).  flags[i] marks the lines the reference skips wherever they stand: 1 = an empty line, 2 = one that starts with the
 * meta character.  What depends on the line NUMBER (line_skip, the header line) and the typed conversions stay with the caller. */
typedef struct dh_text dh_text;
typedef struct dh_text_batch {
    uint64_t n_lines, line0;        /* lines in this batch; number of the first one in the file */
    const uint8_t *text;            /* pinned host memory, valid until the next call on the handle */
    uint64_t text_bytes;
    const uint32_t *cells;          /* [n_lines][n_proj][2]: begin, end; begin = 0xffffffff: the line has fewer fields */
    const uint8_t *flags;           /* [n_lines] */
    uint32_t n_proj;
    int32_t status;                 /* with return value 0: 0 = end of file, -2 = a cut or corrupt block ended the scan */
} dh_text_batch;
dh_text *dh_text_open_file(dh_ctx *ctx, const char *path);          /* NULL: unreadable, or not BGZF (dh_last_error) */
void dh_text_close(dh_text *t);
int dh_text_head(dh_text *t, uint8_t *buf, uint64_t cap, uint64_t *n);      /* the first inflated bytes (whole blocks, <= cap) */
int dh_text_configure(dh_text *t, uint32_t n_cols, const uint32_t *proj, uint32_t n_proj, int meta_char);
int dh_text_scan_begin(dh_text *t);
int dh_text_next_batch(dh_text *t, dh_text_batch *out);             /* 1 = a batch, 0 = done (out->status), < 0 = error */

/* Device-resident variant used for kernel-level measurement and by the BAM scan: upload once, inflate
 * any number of times.  dh_bgzf_dev_load frames `comp` on the host, uploads it and the block table;
 * dh_bgzf_dev_inflate enqueues the kernel(s) on the context's stream WITHOUT synchronising;
 * dh_ctx_sync waits.  dh_bgzf_dev_check downloads the per-block verdicts (first failure, if any);
 * dh_bgzf_dev_read copies a slice of the inflated bytes back. */
typedef struct dh_bgzf_dev dh_bgzf_dev;
dh_bgzf_dev *dh_bgzf_dev_load(dh_ctx *ctx, const uint8_t *comp, uint64_t comp_size);
void dh_bgzf_dev_free(dh_bgzf_dev *d);
uint64_t dh_bgzf_dev_nblocks(const dh_bgzf_dev *d);
uint64_t dh_bgzf_dev_comp_bytes(const dh_bgzf_dev *d);
uint64_t dh_bgzf_dev_out_bytes(const dh_bgzf_dev *d);
int dh_bgzf_dev_inflate(dh_bgzf_dev *d);
/* One inflate pass with CUDA events between the kernels (synchronises after every launch group): ms[0] = phase 1
 * (Huffman decode, one stream per lane), ms[1] = phase 2 (literal/match placement + CRC-32), ms[2] = fallback filter. */
int dh_bgzf_dev_inflate_timed(dh_bgzf_dev *d, float *ms);
int dh_bgzf_dev_check(dh_bgzf_dev *d, int64_t *bad_block, int32_t *bad_code);
int dh_bgzf_dev_read(dh_bgzf_dev *d, uint64_t offset, uint8_t *dst, uint64_t n);

/* ---- No-allocation inflate.  No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The lane-group inflate
 * kernel (inflate + CRC-32 + ISIZE per block, as dh_bgzf_inflate_host runs it) with its scratch — block descriptors, per-block
 * verdicts, work ticket — in a fixed file-scope __device__ array of DH_NOALLOC_SCRATCH_BYTES that every GPU loading the library
 * holds until exit; the call makes no cudaMalloc / cudaFree / cudaMallocHost and needs no dh_ctx.
 *   d_comp   DEVICE buffer holding the BGZF image from file offset 0: block i starts at d_comp + blocks[i].coffset.  16-byte
 *            aligned, and comp_cap (its size) must leave DH_NOALLOC_COMP_PAD readable bytes after every block (the kernel's bulk
 *            copies run ahead of the payload).
 *   blocks   HOST array from dh_bgzf_walk, n_blocks <= DH_NOALLOC_MAX_BLOCKS (more: DH_EINVAL, split the call).
 *   d_out    DEVICE buffer; block i's bytes go to d_out + (sum of ISIZE of the blocks before it); *out_size = the total.
 *   stream   a cudaStream_t (NULL = the default stream); the call returns after the stream has drained.
 * Returns DH_OK, DH_EINVAL (limit or argument), DH_ECUDA, or DH_EDATA with *bad_block / *bad_code = the first failing block.
 * Every call on a device shares the one array: calls of one process are serialised inside the library; do not rely on its
 * contents between calls, and calls from different processes use their own copies. */
#define DH_NOALLOC_MAX_BLOCKS 16384u
#define DH_NOALLOC_SCRATCH_BYTES 590080u
#define DH_NOALLOC_COMP_PAD 8192u
int dh_bgzf_inflate_noalloc(int device, const uint8_t *d_comp, uint64_t comp_cap, const dh_bgzf_block *blocks, uint32_t n_blocks,
                            uint8_t *d_out, uint64_t out_cap, uint64_t *out_size, int64_t *bad_block, int32_t *bad_code, void *stream);

/* ---- A4-A15: read_bam scan -----------------------------------------------------------------------------
 * Replaces, for one BAM image in host memory, what bam_read_bind/local_init/bam_read_function get from htslib
 * (src/bam_reader.c:410-1038): sam_open + sam_hdr_read (bam_hdr_read, htslib/sam.c:229-342), then per record
 * sam_read1 -> bam_read1 (sam.c:779-855, 4124-4134) and the per-row column conversions (bam_reader.c:779-966),
 * here done by batches on the GPU: BGZF inflate -> record-boundary walk -> columnar decode straight into
 * DuckDB vector layouts.  Rows come back in file order, exactly the rows the reference's sequential scan
 * returns; an invalid record ends the scan silently after the rows before it (bam_reader.c:754-765) with the
 * htslib error code reported in dh_bam_batch.status / dh_bam_scan_status. */
typedef struct dh_bam dh_bam;

/* Frames the image (A1), inflates the header blocks on the GPU and parses the BAM header on the host.
 * `data` must stay valid (and should be pinned, dh_host_alloc) until dh_bam_close.  NULL on error. */
dh_bam *dh_bam_open(dh_ctx *ctx, const uint8_t *data, uint64_t size);
/* The same on a FILE, streamed (north-star item (a); htslib never holds more than a block or two either, bgzf.c:1004-1239):
 * only the header blocks are read here.  Scans read the file segment by segment with parallel pread into two pinned staging
 * buffers (dh_bam_configure's seg_comp_bytes each, default 512 MiB), frame the BGZF blocks as the bytes arrive, and copy to the
 * device from there; the first batch leaves before the file has been read, and pinned memory does not grow with the file.
 * Everything else (projection, ranges, index scans, split, clone, batches) works as on a memory image; DH_BAM_DEVICE_INPUT
 * does not apply. */
dh_bam *dh_bam_open_file(dh_ctx *ctx, const char *path);
void dh_bam_close(dh_bam *b);
int32_t dh_bam_n_ref(const dh_bam *b);                       /* sam_hdr_nref */
const char *dh_bam_ref_name(const dh_bam *b, int32_t tid);   /* sam_hdr_tid2name */
uint32_t dh_bam_ref_len(const dh_bam *b, int32_t tid);
const char *dh_bam_header_text(const dh_bam *b, uint32_t *len);
uint64_t dh_bam_first_voffset(const dh_bam *b);              /* virtual offset of the first alignment record */
uint64_t dh_bam_n_blocks(const dh_bam *b);

/* Column schema = the reference's (bam_reader.c:189-204): 0 QNAME 1 FLAG 2 RNAME 3 POS 4 MAPQ 5 CIGAR 6 RNEXT
 * 7 PNEXT 8 TLEN 9 SEQ 10 QUAL 11 READ_GROUP_ID 12 SAMPLE_ID, then one column per requested typed tag
 * (standard_tags, bam_reader.c:54-70).  col_mask bit c = column c is projected (projection pushdown,
 * bam_reader.c:676-679); tag k is column 13+k (tag columns past bit 63, i.e. k >= 51, are always projected).  Tag types: 'i' BIGINT, 'f' DOUBLE, 'Z' VARCHAR (SAM Z or H),
 * 'A' VARCHAR, 'c' BIGINT CIGAR metric (N3 below), 'B' LIST of 8-byte children (BIGINT or DOUBLE by the column's declared sub-type; the cell writer
 * follows the value's RUNTIME sub-type as bam_assign_list_int/_double do, bam_reader.c:106-138, 954-962). */
#define DH_BAM_CORE_COLS 13
#define DH_BAM_MAX_TAGS 63
#define DH_BAM_MAX_COLS (DH_BAM_CORE_COLS + DH_BAM_MAX_TAGS)
typedef struct {
    char tag[2];
    char type;
    char pad;
} dh_tag_req;
int dh_bam_set_projection(dh_bam *b, uint64_t col_mask, const dh_tag_req *tags, int n_tags);

/* A16: the auxiliary_tags := true column (bam_reader.c:539-549, 967-1031): every aux entry of the record, in record
 * order, as key = tag / value = bam_aux_to_string text (bam_reader.c:140-183; integers "%" PRId64, floats "%g", B
 * arrays "sub,v,v,..").  `exclude` = n_exclude two-byte tags left out of the map (the reference leaves out the
 * tags that standard_tags := true serves as typed columns).  Call before a scan begins. */
int dh_bam_set_aux_map(dh_bam *b, int enable, const char *exclude, int n_exclude);

/* Pipeline knobs (0 = keep): compressed bytes per inflate segment, rows per output batch, carry buffer for
 * records that straddle segments.  Small values exercise the multi-segment paths in tests. */
int dh_bam_configure(dh_bam *b, uint64_t seg_comp_bytes, uint32_t batch_rows, uint64_t carry_bytes);

/* Measurement modes (bench.py `value`): DH_BAM_DEVICE_INPUT stages the whole compressed image in HBM once, so
 * scans read it from there instead of copying segments host->device; DH_BAM_DEVICE_OUTPUT leaves each batch in
 * the device arena (no device->host copy; the batch's host pointers are then NOT filled).  Default 0: host in,
 * host out. */
#define DH_BAM_DEVICE_INPUT 1u
#define DH_BAM_DEVICE_OUTPUT 2u
/* DH_BAM_COMPACT_OUTPUT: batches leave the GPU in the COMPACT WIRE LAYOUT instead of DuckDB vector layout (fewer bytes
 * cross PCIe: 16-byte duckdb_string_t cells become 4-byte lengths, dictionary columns become codes, BAM's own 32-bit
 * fields stay 32-bit).  The consumer expands it while it fills its vectors (duckhts_b200/host/duckhts_ext.c does).
 *   data[c] per row:  FLAG uint16 | MAPQ uint8 | POS, PNEXT int32 = BAM pos / next_pos (0-based: add 1) | TLEN int32
 *                     RNAME, RNEXT int32 = tid (-1 = "*"; names: dh_bam_ref_name) | SAMPLE_ID int32 = @RG entry
 *                     (-1 = NULL; dh_bam_rg_sample) | 'i' tags int64, 'f' tags double | 'A' tags uint8 (0 = "") |
 *                     'B' tags duckdb_list_entry as in the DuckDB layout | every other VARCHAR column uint32 = length.
 *   String bytes: rec_off[row] = offset in `heap` of the row's slots, which follow one another in the order QNAME,
 *   CIGAR, SEQ, QUAL, then the 'Z' and 'B' tag columns in request order, then READ_GROUP_ID; projected columns only;
 *   a slot is the string padded to a multiple of 4 bytes ('B': 4 + 8 * length bytes, children 8-byte aligned; NULL
 *   cells have no slot).  Two length words carry a flag in bit 31 for values htslib cuts at an embedded NUL (strlen
 *   semantics of duckdb_vector_assign_string_element): QNAME = l_read_name (NUL included; the string is
 *   l_read_name - 1 bytes, or, flagged, strnlen(slot, l_read_name)), QUAL = l_seq (flagged: strnlen(slot, l_seq)).
 * validity[c] as in the DuckDB layout. */
#define DH_BAM_COMPACT_OUTPUT 4u
int dh_bam_set_mode(dh_bam *b, uint32_t flags);
/* @RG dictionary of the header, in header order (entries with an ID): SAMPLE_ID codes of the compact layout index it. */
int32_t dh_bam_n_rg(const dh_bam *b);
const char *dh_bam_rg_id(const dh_bam *b, int32_t i, uint32_t *len);
const char *dh_bam_rg_sample(const dh_bam *b, int32_t i);   /* SM value, NULL if the line has none (-> NULL cell) */

/* Start a scan over the records that START at virtual offsets in [voff_beg, voff_end) (bgzf virtual offset =
 * coffset << 16 | uoffset, htslib/bgzf.h:261).  voff_beg = 0 means "first record"; voff_end = UINT64_MAX means
 * "to EOF".  voff_beg must be a record boundary (first record, a BAI chunk start, or a dh_bam_split boundary).
 * This is the unit of multi-GPU sharding: each GPU scans one contiguous range with its own dh_ctx/dh_bam. */
int dh_bam_scan_begin(dh_bam *b, uint64_t voff_beg, uint64_t voff_end);

/* VARCHAR cells are DuckDB duckdb_string_t (16 bytes): <= 12 bytes inlined, else prefix + pointer into `heap`
 * (or, for RNAME/RNEXT/SAMPLE_ID, into a dictionary owned by the dh_bam handle).  Fixed-width columns are plain
 * arrays: FLAG uint16, POS/PNEXT/TLEN int64, MAPQ int32, 'i' tags int64, 'f' tags double.  'B' tag cells are
 * duckdb_list_entry {offset, length} with the children at ((const uint64_t *)heap)[offset .. offset+length)
 * (int64 or double bit patterns; the host appends them to the DuckDB list child vector).  validity[c] is a
 * DuckDB validity mask (bit r = row r valid) or NULL when the column cannot be NULL. */
typedef struct {
    uint64_t n_rows;
    uint64_t row0;        /* scan-relative index of the first row of this batch */
    int32_t status;       /* 0, or the htslib error code that ended the scan right after this batch */
    int32_t n_cols;
    const void *data[DH_BAM_MAX_COLS];
    const uint64_t *validity[DH_BAM_MAX_COLS];
    const uint8_t *heap;
    uint64_t heap_bytes;
    uint64_t bytes;       /* bytes copied device -> host for this batch */
    /* AUXILIARY_TAGS MAP(VARCHAR, VARCHAR) column, when dh_bam_set_aux_map enabled it (else NULL / 0): one
     * duckdb_list_entry {offset, length} per row into the n_aux key / value cells (duckdb_string_t, long values point
     * into `heap`); rows without a pair are NULL (validity bit 0) as at bam_reader.c:982-991. */
    const void *aux_map;
    const uint64_t *aux_map_validity;
    const void *aux_keys;
    const void *aux_values;
    uint64_t n_aux;
    int32_t slot;               /* pipeline slot holding this batch: the argument of dh_bam_release_batch */
    uint32_t layout;            /* 0 = DuckDB vector layout, DH_BAM_COMPACT_OUTPUT = compact wire layout */
    const uint32_t *rec_off;    /* compact layout: heap offset of every row's string slots; else NULL */
} dh_bam_batch;
/* Returns 1 and fills *out (buffers stay valid until the next call on this handle), 0 at end of scan, < 0 on
 * error. */
int dh_bam_next_batch(dh_bam *b, dh_bam_batch *out);
/* The same with explicit lifetime, for consumers that work on several batches at once (the worker threads of one
 * DuckDB query): an acquired batch stays valid until dh_bam_release_batch(b, out->slot).  The pipeline has
 * dh_bam_set_slots() slots (default 6, at least 3, at most 32; set before the scan begins); slots that are neither
 * held nor in flight are refilled ahead of the consumer.  Returns like dh_bam_next_batch; if every slot is held it
 * returns DH_EBUSY rather than wait for a release that only the caller can make. */
int dh_bam_set_slots(dh_bam *b, int n_slots);
/* Starts the pipeline without waiting for anything: the host-to-device copies and the inflate of the first segments
 * are enqueued and the call returns.  A consumer that drives several handles (one per GPU) calls this on each right
 * after dh_bam_scan_*, so that every GPU is at work before the first batch is asked for. */
int dh_bam_prefetch(dh_bam *b);
int dh_bam_acquire_batch(dh_bam *b, dh_bam_batch *out);
int dh_bam_release_batch(dh_bam *b, int32_t slot);

/* Per-column 64-bit checksums of one batch: sums[c] += sum over rows of a hash of the cell's VALUE (strings by their
 * bytes, NULL as a constant), so the result is independent of the layout (DuckDB / compact), of where the batch lives
 * (host: computed by host threads; device output: by a kernel on the batch arena) and of how a scan is cut into
 * batches, ranges or GPUs: checksums of disjoint row sets add up modulo 2^64.  Used for parity at sizes no cell-by-cell
 * comparison reaches (bench.py: tiled input = tiles x the unit's checksums).  sums has DH_BAM_MAX_COLS entries; columns
 * not projected are left alone.  The batch must be one this handle returned and still holds. */
int dh_bam_batch_checksum(dh_bam *b, const dh_bam_batch *batch, uint64_t *sums);

typedef struct {
    uint64_t rows;
    uint64_t h2d_bytes, d2h_bytes;
    uint64_t inflated_bytes;
    uint64_t n_blocks;
    uint64_t rewalk_rounds;   /* boundary speculation misses that needed a re-walk */
    int32_t status;           /* final scan status (0 = clean EOF / range end) */
    int32_t pad;
    uint64_t end_voffset;     /* virtual offset of the first record NOT returned (for range merging) */
    float inflate_ms, walk_ms, decode_ms;  /* sums of CUDA-event kernel times */
} dh_bam_scan_stats;
int dh_bam_scan_stats_get(const dh_bam *b, dh_bam_scan_stats *out);

/* n_parts + 1 virtual offsets cutting the alignment section into contiguous ranges of about equal compressed
 * size on BGZF block boundaries (htslib-written BAM starts every block on a record boundary, htslib/sam.c:883).
 * voffs[0] = first record, voffs[n_parts] = UINT64_MAX. */
int dh_bam_split(const dh_bam *b, int n_parts, uint64_t *voffs);

/* ---- A8: BAI/CSI index, region queries, contig-parallel scans -----------------------------------------------
 * Replaces hts_idx_load3 -> idx_read / idx_read_core (htslib/hts.c:2920-3055) and, per query, sam_itr_regarray /
 * sam_itr_queryi -> hts_itr_query / hts_itr_multi_bam (hts.c:3421-3745) plus the record filter of hts_itr_next /
 * hts_itr_multi_next (hts.c:4266-4604) that src/bam_reader.c:642-671,689-716,749-752 drives.  The index is parsed on
 * the host (a CSI file is BGZF: it is inflated by the GPU kernel first); a query becomes a few virtual-offset ranges
 * and an overlap filter evaluated inside the record-boundary kernel, so only matching records produce rows.  Rows:
 * the records, in file order, that overlap any requested interval (each once) — the iterator's result on a
 * coordinate-sorted file.  `data` is the index file image (BAI: "BAI\1"; CSI: BGZF-wrapped "CSI\1"). */
typedef struct dh_index dh_index;
dh_index *dh_index_load(dh_ctx *ctx, const uint8_t *data, uint64_t size);
void dh_index_free(dh_index *ix);
int32_t dh_index_n_ref(const dh_index *ix);
int dh_index_is_csi(const dh_index *ix);
uint64_t dh_index_n_no_coor(const dh_index *ix);
/* Region strings as read_bam's `region` parameter carries them after splitting at commas (src/bam_reader.c:318-348):
 * "chr", "chr:beg-end", "chr:beg", "chr:-end", "{chr}:beg-end"; 1-based inclusive; thousands separators allowed.
 * Unknown names are skipped like htslib does; if none is left the call returns DH_ENOREGION (the reference reports
 * "No reads found for region(s): ..." at init).  Starts the scan: follow with dh_bam_next_batch. */
int dh_bam_scan_regions(dh_bam *b, const dh_index *idx, const char *const *regions, int n_regions);
/* The reference's index-driven scan without a region (src/bam_reader.c:577-585,689-716): contigs tid_lo..tid_hi-1, one
 * whole-contig query each, in contig order.  Reads without a contig (tid = -1) are not returned.  This is
 * also the unit of region-sharded multi-GPU scans: rank r scans the contig group dh_bam_split_contigs gives it. */
int dh_bam_scan_contigs(dh_bam *b, const dh_index *idx, int32_t tid_lo, int32_t tid_hi);
/* n_parts + 1 contig bounds cutting [0, n_ref) into contiguous groups of about equal compressed size (from the
 * index's chunk offsets). */
int dh_bam_split_contigs(const dh_bam *b, const dh_index *idx, int n_parts, int32_t *tid_bounds);
/* The same from the index alone (host code, no GPU): contigs [0, dh_index_n_ref). */
int dh_index_split_contigs(const dh_index *idx, int n_parts, int32_t *tid_bounds);
/* A second handle on the same image for another context (= another GPU): block table, header and dictionaries are
 * copied from `src` instead of being walked, inflated and parsed again.  This is how one query is spread over the
 * GPUs of a box: one handle per GPU, each scanning the range / contig group dh_bam_split(_contigs) gives it. */
dh_bam *dh_bam_clone(const dh_bam *src, dh_ctx *ctx);

/* ---- N2: index build ---------------------------------------------------------------------------------------------------
 * Replaces bam_index(path) = sam_index_build3 -> sam_index (htslib/sam.c:989-1027: sam_read1 + hts_idx_push per record) ->
 * hts_idx_finish -> hts_idx_save (htslib/hts.c:2400-2642, 2754-2860), src/hts_index_builder.c:110-160, for BAI (min_shift 0).
 * The file is scanned once on the GPU (inflate + boundary walk; no column is decoded): a kernel emits the runs of records with
 * equal (contig, bin) and the linear index, the host closes the chunks, applies hts_idx_finish's rules (bins whose chunks span
 * less than 64 KB of file move to their parent, adjacent chunks of one BGZF block merge) and serialises the BAI.  The result has
 * the bins, chunks, linear index, per-contig mapped / unmapped counts and n_no_coor htslib writes (bins in ascending order; htslib
 * writes them in its hash table's order).  *bai is malloc'ed: free it with dh_free.  DH_EFORMAT: the file is not coordinate-sorted
 * or a coordinate does not fit a BAI (2^29). */
int dh_bam_index_build(dh_bam *b, uint8_t **bai, uint64_t *bai_size);
void dh_free(void *p);

/* ---- N3: predicate pushdown ------------------------------------------------------------------------------------------
 * Record filter evaluated in the boundary kernel next to the region test, so rejected records cost no row, no heap byte
 * and no PCIe byte.  A record passes iff (FLAG & flag_require) == flag_require, (FLAG & flag_exclude) == 0 and
 * MAPQ >= min_mapq: the conjunctions of the reference's flag predicates (is_paired, is_unmapped, is_reverse, ...,
 * src/kmer_udf.c:1082-1183) and a mapping-quality floor.  All zero = no filter.  Set before the scan begins. */
int dh_bam_set_filter(dh_bam *b, uint32_t flag_require, uint32_t flag_exclude, uint32_t min_mapq);
/* CIGAR metrics as derived columns: the reference's scalar UDFs over read_bam's CIGAR text (src/kmer_udf.c:197-293, registered
 * at :1231-1238) computed by the decode kernel from the binary CIGAR.  Request them through dh_bam_set_projection as tag columns of
 * type 'c' (a BIGINT column, 8 bytes per row in both layouts, NULL where the UDF returns NULL: "*", an operation of length 0, an
 * operation the text parser rejects):  tag[0] = 'S' cigar_has_soft_clip, 'H' cigar_has_hard_clip (0 / 1), 'l' cigar_left_soft_clip,
 * 'r' cigar_right_soft_clip, 'q' cigar_query_length, 'a' cigar_aligned_query_length, 'R' cigar_reference_length,
 * 'o' cigar_has_op(CIGAR, tag[1]) (0 / 1; tag[1] one of MIDNSHP=X, lower case accepted).  The SAM-flag predicates
 * (is_paired ... is_supplementary, sam_flag_has) are single AND tests of the FLAG column; as row filters they are dh_bam_set_filter. */

/* ---- B1-B10: read_bcf scan --------------------------------------------------------------------------------------
 * Replaces, for one BCF image in host memory, what bcf_read_bind / bcf_read_local_init / bcf_read_function get from
 * htslib (src/bcf_reader.c:451-2053): bcf_hdr_read (htslib/vcf.c:1710+) -> dictionaries and column schema; per record
 * bcf_read -> bcf_read1_core + bcf_record_check (vcf.c:1874-1911, 2040-2212), bcf_unpack (vcf.c:4178-4302),
 * bcf_get_info_* / bcf_get_format_* / bcf_get_genotypes (vcf.c:6056-6248) and the column writers incl. the GT string
 * builder (bcf_reader.c:1381-1462, 1542-1733, 1738-1998) — done by batches on the GPU.  Records come back in file
 * order; an invalid record ends the scan silently after the records before it (bcf_reader.c:1311-1338) with htslib's
 * code (-2) in the batch / scan status.
 *
 * Layout of a batch: site-level columns (CHROM POS ID REF ALT QUAL FILTER, INFO_*) hold ONE cell per record; FORMAT
 * columns hold one cell per (record, sample), sample fastest — exactly the row order of tidy_format := true, where the
 * site-level cell of row r is cell r / n_samples (the reference repeats it per sample row, bcf_reader.c:1283-1290);
 * in wide mode column FORMAT_<f>_<sample s> is cell r * n_samples + s of FORMAT column f.  Cell types: VARCHAR =
 * duckdb_string_t (long strings point into `heap`), BOOLEAN uint8, INTEGER int32, FLOAT float (bit copies), POS int64,
 * QUAL double, LIST = duckdb_list_entry {offset, length} with the children at `heap` + offset * child size (4 for
 * INTEGER / FLOAT children, 16 for VARCHAR children).  validity = DuckDB mask (bit = 1 valid) or NULL when the column
 * cannot be NULL; the cells of NULL rows are unspecified. */
typedef struct dh_bcf dh_bcf;
typedef struct {
    const char *name;      /* header ID */
    int32_t type;          /* 0 Flag -> BOOLEAN, 1 Integer -> INTEGER, 2 Float -> FLOAT, 3 String -> VARCHAR (create_bcf_field_type) */
    int32_t is_list;       /* LIST column (Number, corrected by the VCF-spec table of src/include/vcf_types.h, is not fixed) */
} dh_bcf_field;
dh_bcf *dh_bcf_open(dh_ctx *ctx, const uint8_t *data, uint64_t size);
void dh_bcf_close(dh_bcf *b);
int32_t dh_bcf_n_samples(const dh_bcf *b);
const char *dh_bcf_sample_name(const dh_bcf *b, int32_t i);
int32_t dh_bcf_n_contigs(const dh_bcf *b);
const char *dh_bcf_contig_name(const dh_bcf *b, int32_t rid);
int32_t dh_bcf_n_info(const dh_bcf *b);              /* INFO_<id> columns, in header-dictionary order */
int32_t dh_bcf_n_format(const dh_bcf *b);            /* FORMAT_<id> columns, in header-dictionary order */
int dh_bcf_info_field(const dh_bcf *b, int32_t i, dh_bcf_field *out);
int dh_bcf_format_field(const dh_bcf *b, int32_t i, dh_bcf_field *out);
int dh_bcf_has_vep(const dh_bcf *b);                 /* header declares CSQ / BCSQ / ANN (VEP_* columns: not served) */
const char *dh_bcf_header_text(const dh_bcf *b, uint32_t *len);
/* Projection pushdown: bit c of core_mask = core column c (0 CHROM .. 6 FILTER); info_want[i] / format_want[i] != 0 =
 * INFO / FORMAT column i.  NULL arrays = all. */
#define DH_BCF_CORE_COLS 7
int dh_bcf_set_projection(dh_bcf *b, uint32_t core_mask, const uint8_t *info_want, const uint8_t *format_want);
/* Pipeline knobs (0 = keep): compressed bytes per inflate segment, records per batch, bytes of the per-batch region
 * for strings longer than 12 bytes and string-list cells. */
int dh_bcf_configure(dh_bcf *b, uint64_t seg_comp_bytes, uint32_t batch_records, uint64_t string_bytes);
/* DH_BAM_DEVICE_INPUT / DH_BAM_DEVICE_OUTPUT as for dh_bam.  DH_BAM_COMPACT_OUTPUT: the FORMAT columns -- 2,504 cells per record
 * and column, where the bytes of a batch are -- leave the GPU compact (site-level columns keep the DuckDB layout):
 *   GT            uint32 per cell: bit 31 clear: [25:24] = length (<= 3), [23:0] = the characters ("0|1"); bit 31 set: [30:0] = offset
 *                 in `heap` of {uint32 length, bytes} (alleles >= 10, polyploid calls)
 *   INTEGER[] / FLOAT[]   uint16 per cell = number of elements kept (missing and vector_end dropped); the cell of sample s of record r
 *                 has its children at ((uint32 *)heap)[first + s * per], {first, per} = ((uint32 *)format_rec[f])[2 r .. 2 r + 1]
 *   everything else as in the DuckDB layout.  48 -> 21 bytes per (record, sample) for GT, DP, GQ, AD. */
int dh_bcf_set_mode(dh_bcf *b, uint32_t flags);
int dh_bcf_scan_begin(dh_bcf *b);
/* Region query through the CSI index (bcf_itr_querys + bcf_itr_next, htslib/vcf.h:1321-1330; used by
 * src/bcf_reader.c:917-951): the records with rid == contig, pos < end and pos + rlen > beg, in file order.  One
 * region per scan; the reference chains several regions as a union of single-region iterators (duplicates kept), the
 * caller does the same with one scan per region.  DH_ENOREGION: the region names no contig (the reference warns and
 * skips it). */
int dh_bcf_scan_region(dh_bcf *b, const dh_index *idx, const char *region);
typedef struct {
    uint64_t n_records;
    uint64_t n_samples;
    uint64_t rec0;         /* scan-relative index of the first record of this batch */
    int32_t status;        /* 0, or -2 when an invalid record ended the scan right after this batch */
    int32_t pad;
    const void *core[DH_BCF_CORE_COLS];
    const uint64_t *core_validity[DH_BCF_CORE_COLS];
    const void *const *info;                 /* dh_bcf_n_info entries; NULL where not projected */
    const uint64_t *const *info_validity;
    const void *const *format;               /* dh_bcf_n_format entries */
    const uint64_t *const *format_validity;
    const uint8_t *heap;
    uint64_t heap_bytes;
    uint64_t bytes;        /* bytes copied device -> host for this batch */
    int32_t slot;          /* pipeline slot: the argument of dh_bcf_release_batch */
    uint32_t layout;       /* 0 = DuckDB vector layout, DH_BAM_COMPACT_OUTPUT = compact FORMAT cells (dh_bcf_set_mode) */
    const void *const *format_rec;   /* compact layout, numeric LIST FORMAT columns: per record {uint32 first child, uint32 values per sample}; else NULL */
} dh_bcf_batch;
/* Returns 1 and fills *out (buffers stay valid until the next call on this handle), 0 at end of scan, < 0 on error. */
int dh_bcf_next_batch(dh_bcf *b, dh_bcf_batch *out);
/* Explicit lifetime, as for dh_bam: valid until dh_bcf_release_batch(b, out->slot). */
int dh_bcf_set_slots(dh_bcf *b, int n_slots);
int dh_bcf_acquire_batch(dh_bcf *b, dh_bcf_batch *out);
int dh_bcf_release_batch(dh_bcf *b, int32_t slot);
int dh_bcf_scan_stats_get(const dh_bcf *b, dh_bam_scan_stats *out);

/* CUDA-event timer on the context's stream (the stream every kernel of this library is launched on). */
int dh_ctx_sync(dh_ctx *ctx);
int dh_timer_start(dh_ctx *ctx);
int dh_timer_stop(dh_ctx *ctx, float *ms);
/* Flush L2 by overwriting a scratch buffer larger than L2 (timing hygiene between iterations). */
int dh_flush_l2(dh_ctx *ctx);
/* Host<->device copy bandwidth probe (pinned, `bytes` per direction); GB/s. */
int dh_measure_pcie(dh_ctx *ctx, uint64_t bytes, float *h2d_gbs, float *d2h_gbs);
/* Same, both directions running at the same time (what the read_bam pipeline sees). */
int dh_measure_pcie_bidir(dh_ctx *ctx, uint64_t bytes, float *h2d_gbs, float *d2h_gbs);
/* Number of kernels this library has launched on this context so far (for bench.py's gpu_launches). */
uint64_t dh_ctx_launch_count(const dh_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif /* DUCKHTS_CUDA_H */
