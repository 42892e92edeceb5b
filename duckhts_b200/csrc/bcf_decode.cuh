// bcf_decode.cuh — device-side types + launchers of the BCF record-boundary / validation kernel (B2, B3) and the
// columnar decode kernel (B4-B10).  See bcf_decode.cu for the reference functions each kernel replaces.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "bam_decode.cuh"   // StrCell, BlockWalk, WalkParams

namespace dh {

// header value types (htslib/vcf.h BCF_HT_*) and record value types (BCF_BT_*)
enum { HT_FLAG = 0, HT_INT = 1, HT_REAL = 2, HT_STR = 3 };
enum { BT_NULL = 0, BT_INT8 = 1, BT_INT16 = 2, BT_INT32 = 3, BT_FLOAT = 5, BT_CHAR = 7 };
// per-column meta byte: low 2 bits = HT_*, bit 4 = list column (Number != fixed), bit 5 = the GT column
constexpr uint8_t META_LIST = 16, META_GT = 32;

constexpr int BCF_CORE_COLS = 7;   // CHROM POS ID REF ALT QUAL FILTER (src/bcf_reader.c:552-577)
enum { BC_CHROM = 0, BC_POS, BC_ID, BC_REF, BC_ALT, BC_QUAL, BC_FILTER };
constexpr int BCF_MAX_FMT = 64;    // FORMAT fields one record may carry (n_fmt is a byte; more than 64 distinct is unheard of)

// DuckDB LIST parent cell (duckdb_list_entry, reference duckdb_capi/duckdb.h:393-399)
struct ListEntry { uint64_t offset, length; };

struct BcfTables {
    const StrCell *ctg_cells;    // n_ctg cells (contig names); len = 0xffffffff marks a hole in the dictionary
    const StrCell *id_cells;     // n_ids cells (FILTER/INFO/FORMAT names); len = 0xffffffff marks a hole
    const int16_t *info_col;     // dictionary id -> INFO column, -1 if the id is not an INFO field
    const int16_t *fmt_col;      // dictionary id -> FORMAT column, -1 if not
    const uint8_t *info_meta;    // per INFO column
    const uint8_t *fmt_meta;     // per FORMAT column
    const uint8_t *info_want;    // per INFO column: projected
    const uint8_t *fmt_want;     // per FORMAT column: projected
    int32_t n_ctg, n_ids, n_info, n_fmt;
    uint32_t n_samples;
    uint32_t core_mask;          // bit c = core column c projected
    int32_t gt_id;               // dictionary id of FORMAT/GT when the header is older than VCFv4.4 (updatephasing), else -1
    uint32_t any_fmt;            // any FORMAT column projected
};

struct BcfDecodeParams {
    const uint8_t *u;
    const BlockWalk *walk;
    const int64_t *blk_start;
    const uint16_t *rec;
    int64_t rec_bias;
    const uint32_t *row_base;       // records of the batch before this block
    const uint64_t *heap_base;      // fixed-region bytes of the batch before this block
    uint32_t b_lo, b_hi;
    uint8_t *core[BCF_CORE_COLS];   // site-level columns, one cell per RECORD (nullptr = not projected)
    unsigned long long *core_valid[BCF_CORE_COLS];
    uint8_t *const *info;           // n_info device pointers (host-mapped table), one cell per record
    unsigned long long *const *info_valid;
    uint8_t *const *fmt;            // n_fmt device pointers, one cell per (record, sample), sample fastest
    unsigned long long *const *fmt_valid;
    uint8_t *heap;                  // [fixed region | bump region]; LIST offsets count child elements from here
    uint64_t heap_host;             // host address the heap is copied to (long VARCHAR cells point into it)
    uint64_t fixed_bytes;           // size of the fixed region (list children at structural capacity)
    uint64_t bump_cap;              // size of the bump region (strings > 12 bytes, string-list cells)
    unsigned long long *bump_used;  // bytes requested from the bump region (may exceed bump_cap: the host then retries)
    // compact wire layout (dh_bcf_set_mode(DH_BAM_COMPACT_OUTPUT), include/duckhts_cuda.h): the GT column is 4 bytes per cell, numeric
    // LIST FORMAT columns are a 2-byte element count per cell + {first child, values per sample} per RECORD
    uint32_t compact, pad;
    uint8_t *const *fmt_rec;        // n_fmt device pointers: per record {uint32 first child (4-byte units from `heap`), uint32 values per sample}
};

cudaError_t launch_bcf_walk(const WalkParams &p, const BcfTables &t, cudaStream_t stream);
cudaError_t launch_bcf_decode(const BcfDecodeParams &p, const BcfTables &t, cudaStream_t stream);

}  // namespace dh
