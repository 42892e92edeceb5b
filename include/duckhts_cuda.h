/* duckhts_cuda.h — C-ABI of libduckhts_cuda.so: the drop-in boundary between host C code (the DuckDB
 * table functions read_bam / bgunzip, or any other FFI caller) and the sm_100a CUDA kernels.
 *
 * Plain C: opaque handles, plain pointers and sizes, no C++/torch types.  Every entry point names the
 * reference interface it replaces (paths relative to the reference tree; "htslib/" =
 * third_party/htslib/).  INTEGRATION.md shows how the reference's src/bam_reader.c and src/bgzip.c
 * would bind these.
 *
 * Threading: a dh_ctx is bound to one CUDA device and may be used by one host thread at a time.
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

typedef struct dh_ctx dh_ctx;

const char *dh_version(void);
const char *dh_last_error(void);
int dh_device_count(void);

/* One context per GPU.  Creates the streams, the pinned staging and the device arenas lazily. */
dh_ctx *dh_ctx_create(int device);
void dh_ctx_destroy(dh_ctx *ctx);
int dh_ctx_sm_count(const dh_ctx *ctx);

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
int dh_bgzf_dev_check(dh_bgzf_dev *d, int64_t *bad_block, int32_t *bad_code);
int dh_bgzf_dev_read(dh_bgzf_dev *d, uint64_t offset, uint8_t *dst, uint64_t n);

/* CUDA-event timer on the context's stream (the stream every kernel of this library is launched on). */
int dh_ctx_sync(dh_ctx *ctx);
int dh_timer_start(dh_ctx *ctx);
int dh_timer_stop(dh_ctx *ctx, float *ms);
/* Flush L2 by overwriting a scratch buffer larger than L2 (timing hygiene between iterations). */
int dh_flush_l2(dh_ctx *ctx);
/* Host<->device copy bandwidth probe (pinned, `bytes` per direction); GB/s. */
int dh_measure_pcie(dh_ctx *ctx, uint64_t bytes, float *h2d_gbs, float *d2h_gbs);
/* Number of kernels this library has launched on this context so far (for bench.py's gpu_launches). */
uint64_t dh_ctx_launch_count(const dh_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif /* DUCKHTS_CUDA_H */
