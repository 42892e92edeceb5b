// bam_decode.cuh — device-side types + launchers of the BAM record-boundary (A5-A7) and columnar decode
// (A9-A15) kernels.  See bam_decode.cu for the reference functions each kernel replaces.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace dh {

// DuckDB VARCHAR cell (duckdb_string_t, reference duckdb_capi/duckdb.h:376-391): <= 12 bytes inlined
// (zero padded), else 4-byte prefix + pointer.  The pointer the kernel stores is a HOST address inside
// the pinned heap the batch is copied into, so the cells are valid DuckDB cells after the D2H copy.
struct alignas(16) StrCell {
    uint32_t len;
    uint32_t w[3];  // inlined bytes 0..11 | prefix, ptr_lo, ptr_hi
};

struct TagReq {
    uint8_t t0, t1;
    uint8_t type;  // 'i' BIGINT, 'f' DOUBLE, 'Z' VARCHAR (Z or H), 'A' VARCHAR(1), 'B' LIST (duckdb_list_entry + 8-byte children in the heap)
    uint8_t want;  // column projected
};

// @RG dictionary entry: ID bytes live in a side heap; sample = prebuilt cell (SM value) or absent
struct alignas(16) RgEntry {
    StrCell sample;
    uint32_t id_off, id_len, has_sample, pad;
};

constexpr int BAM_CORE_COLS = 13;  // QNAME FLAG RNAME POS MAPQ CIGAR RNEXT PNEXT TLEN SEQ QUAL READ_GROUP_ID SAMPLE_ID
constexpr int BAM_MAX_TAGS = 63;   // + 1 internal RG request = 64 = two per lane
constexpr int BAM_MAX_COLS = BAM_CORE_COLS + BAM_MAX_TAGS;
enum { COL_QNAME = 0, COL_FLAG, COL_RNAME, COL_POS, COL_MAPQ, COL_CIGAR, COL_RNEXT, COL_PNEXT, COL_TLEN, COL_SEQ,
       COL_QUAL, COL_RG, COL_SAMPLE };

// what both kernels need to know about the file and the projection
struct ScanTables {
    const StrCell *ref_cells;  // n_ref + 1 cells; [n_ref] = "*"
    const RgEntry *rg;
    const uint8_t *rg_ids;
    const TagReq *tags;        // n_tags user requests
    int32_t n_ref;
    uint32_t n_rg;
    uint32_t n_tags;
    uint32_t any_tag_cols;     // any tag column projected
    uint32_t aux_map;          // the AUXILIARY_TAGS MAP column is projected (A16)
    uint32_t compact;          // compact wire layout (DH_BAM_COMPACT_OUTPUT): heap slots for every string, 4-byte aligned
    const uint32_t *aux_excl;  // 65,536-bit set of tags (t0 | t1 << 8) left out of the MAP (the standard_tags columns), or nullptr
    uint64_t col_mask;         // bit c set = core column c projected (tag columns: TagReq.want)
    // N3 predicate pushdown (dh_bam_set_filter): FLAG bits that must be set / clear, MAPQ floor; all 0 = none
    uint32_t flag_require, flag_exclude, min_mapq, pad2;
};

// per BGZF block result of the boundary walk (device -> host)
struct BlockWalk {
    int64_t entry;        // offset (relative to the segment's inflated base) of the first record starting in the block
    int64_t exit_off;     // offset of the first record NOT consumed by this block's walk
    uint64_t heap_bytes;  // heap bytes the block's records need (4-byte aligned slots)
    uint32_t count;       // records that start in this block and were accepted
    int32_t stop;         // 0; 1 = record at exit_off needs bytes beyond the segment; <0 = bam_read1 error code there
    uint32_t map_entries; // AUXILIARY_TAGS: key/value pairs of the block's records
    uint32_t map_heap;    // AUXILIARY_TAGS: heap bytes of their long values (the tail of the block's heap_bytes)
};

// one interval of a region query: [beg, end) 0-based on contig tid (hts_reglist_t intervals after reg_compact)
struct RegIntv {
    int64_t beg, end;
    int32_t tid, pad;
};

struct WalkParams {
    const uint8_t *u;          // segment inflated base (offset 0); bytes [-carry, seg_end) are valid
    const int64_t *blk_start;  // n+1 block start offsets (blk_start[0] may be negative: carry pseudo-block)
    const int64_t *entry_in;   // per block: fixed entry, or INT64_MIN = speculate
    const uint8_t *todo;       // per block: 1 = walk it in this round (nullptr = all)
    BlockWalk *out;
    uint32_t n;
    int32_t is_final;          // 1: the data ends at seg_end (EOF): an incomplete record is an error, not a carry
    int64_t seg_end;
    int64_t limit_end;         // records starting at or after this offset are outside the scan range
    // in-block offset of every record the walk accepts: block i's records sit at rec[(blk_start[i] - rec_bias) / 36 + i ...]
    // (a record is at least 36 bytes, so the ranges of consecutive blocks cannot overlap)
    uint16_t *rec;
    int64_t rec_bias;          // = blk_start[0]
    // region filter: 0 none; 1 = every placed record (tid >= 0, end > 0: the reference's per-contig index scan);
    // 2 = records overlapping one of the filt_n intervals (device array, sorted by tid then beg, disjoint)
    int32_t filt_mode;
    uint32_t filt_n;
    const RegIntv *filt;
};

struct DecodeParams {
    const uint8_t *u;
    const BlockWalk *walk;        // count per block
    const int64_t *blk_start;     // block start offsets (as given to the walk)
    const uint16_t *rec;          // record offsets written by the walk
    int64_t rec_bias;
    const uint32_t *row_base;     // per block, relative to the batch
    const uint64_t *heap_base;    // per block, relative to the batch heap
    uint32_t b_lo, b_hi;          // blocks of this batch
    uint8_t *col[BAM_MAX_COLS];   // device column data (nullptr = not projected)
    unsigned long long *valid[BAM_MAX_COLS];  // validity words (bit = 1 valid); nullptr = column has no NULLs
    uint8_t *heap;                // device heap
    uint64_t heap_host;           // host address the heap will be copied to
    uint32_t *rec_off;            // compact layout: heap offset of each row's slots (nullptr in the DuckDB layout)
};

// AUXILIARY_TAGS MAP(VARCHAR, VARCHAR) column (bam_reader.c:967-1031): one list entry per row, key / value cells per pair
struct AuxMapParams {
    const uint8_t *u;
    const BlockWalk *walk;
    const int64_t *blk_start;
    const uint16_t *rec;
    int64_t rec_bias;
    const uint32_t *row_base;
    const uint64_t *heap_base;
    const uint32_t *map_base;     // per block: pairs of the batch before this block
    uint32_t b_lo, b_hi;
    uint8_t *entries;             // duckdb_list_entry per row
    unsigned long long *valid;    // row validity (NULL row = no pair)
    uint8_t *keys, *values;       // duckdb_string_t cells, one per pair
    uint8_t *heap;
    uint64_t heap_host;
};

// ---- N2: index build (bam_index) -----------------------------------------------------------------------------------------
// hts_idx_push (htslib/hts.c:2553-2642) closes a chunk whenever the bin changes: the file is a sequence of RUNS of records with the
// same (tid, bin).  One thread per BGZF block emits the runs of its records (the host merges runs that continue across blocks) and
// feeds the linear index (first record start per 16 kb window = atomicMin over inflated offsets, which order like virtual offsets).
struct IdxRun {
    uint64_t start_u;             // inflated offset (relative to the scan's block table) of the run's first record
    int32_t tid;
    uint32_t bin;                 // hts_reg2bin(beg, end, 14, 5); 4680 for reads without a contig
    uint32_t n_mapped, n_unmapped;
};
struct IdxBlock {                 // per walk block
    uint32_t base, n_runs;        // its runs: runs[base .. base + n_runs)
    int32_t first_tid, last_tid;  // sortedness across blocks (hts_idx_push :2597-2616)
    int64_t first_beg, last_beg;
    int32_t err;                  // 0; 1 unsorted positions; 2 coordinate beyond 2^29 (hts_idx_check_range); 3 linear index too short
    uint32_t count;
};
struct IndexParams {
    const uint8_t *u;
    const BlockWalk *walk;
    const int64_t *blk_start;
    const uint16_t *rec;
    int64_t rec_bias;
    uint32_t n;                   // walk blocks (incl. the carry pseudo-block 0)
    int64_t u_base;               // table-relative inflated offset of segment offset 0
    IdxRun *runs;
    uint32_t run_cap;
    unsigned int *counter;        // runs claimed so far
    IdxBlock *blocks;
    unsigned long long *lin;      // linear index: per contig lin_n[tid] windows at lin_base[tid], initialised to ~0
    const uint32_t *lin_base, *lin_n;
    int32_t n_ref;
};
cudaError_t launch_bam_index(const IndexParams &p, const ScanTables &t, cudaStream_t stream);
cudaError_t launch_bam_auxmap(const AuxMapParams &p, const ScanTables &t, cudaStream_t stream);
cudaError_t launch_bam_walk(const WalkParams &p, const ScanTables &t, cudaStream_t stream);
cudaError_t launch_bam_decode(const DecodeParams &p, const ScanTables &t, cudaStream_t stream);
cudaError_t launch_bam_decode_compact(const DecodeParams &p, const ScanTables &t, cudaStream_t stream);

// ---- per-column checksums of a batch (dh_bam_batch_checksum) ----------------------------------------------------------
// One definition for the device kernel, the host threads and the test-side restatement: cell hash = mix64 of the value
// (fixed width), of a Horner polynomial over the bytes (strings) or over the 8-byte children (lists); NULL = a constant.
constexpr uint64_t CK_NULL = 0x9e3779b97f4a7c15ull, CK_K = 0x100000001b3ull;
__host__ __device__ inline uint64_t ck_mix(uint64_t x) {
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull; x ^= x >> 27; x *= 0x94d049bb133111ebull; x ^= x >> 31;
    return x;
}
__host__ __device__ inline uint64_t ck_bytes(const uint8_t *p, uint32_t n) {
    uint64_t h = 0;
    for (uint32_t i = 0; i < n; ++i) h = h * CK_K + (uint64_t)p[i] + 1ull;
    return ck_mix(h ^ ((uint64_t)n * CK_NULL));
}
__host__ __device__ inline uint64_t ck_words(const uint64_t *p, uint64_t n) {
    uint64_t h = 0;
    for (uint64_t i = 0; i < n; ++i) h = h * CK_K + p[i] + 1ull;
    return ck_mix(h ^ (n * CK_NULL));
}
enum { CK_SKIP = 0, CK_U16, CK_MAPQ, CK_POS, CK_TLEN, CK_STR, CK_DICT, CK_SAMPLE, CK_TAG_I, CK_TAG_F, CK_TAG_A, CK_TAG_B };
struct ChecksumParams {
    DecodeParams dp;              // column pointers of the batch (device or host addresses, all on one side)
    uint64_t n_rows;
    uint64_t heap_bytes;
    uint32_t compact;
    uint32_t n_cols;
    uint8_t kind[BAM_MAX_COLS];   // CK_* per column
    const uint64_t *ref_hash;     // n_ref + 1 hashes of the contig names ([n_ref] = "*"): compact RNAME / RNEXT
    const uint64_t *sample_hash;  // per @RG entry: hash of its SM value: compact SAMPLE_ID
    const uint64_t *dict_ptr;     // DuckDB layout: host addresses of dictionary strings longer than 12 bytes, sorted ...
    const uint64_t *dict_hash;    // ... and their hashes
    uint32_t n_dict;
    int32_t n_ref;
    uint64_t *sums;               // BAM_MAX_COLS accumulators
};
cudaError_t launch_bam_checksum(const ChecksumParams &p, cudaStream_t stream);
void bam_checksum_host(const ChecksumParams &p, int n_threads);

}  // namespace dh
