// bcf_decode.cu — sm_100a kernels for the BCF half of the scan hot path.
//
//   bcf_walk_kernel    replaces bcf_read1_core (htslib/vcf.c:1874-1911: l_shared / l_indiv framing, the 24-byte fixed
//                      part) and bcf_record_check (vcf.c:2040-2212: structural validation of every typed field of the
//                      shared and individual sections) for every BGZF block in parallel.  BCF records straddle BGZF
//                      blocks routinely (bcf_write never flushes on record boundaries, vcf.c:2554-2556), so a block's
//                      first record offset is speculated — the first offset whose 32-byte fixed part is plausible
//                      (n_sample equal to the header's, rid in range, sane lengths) and whose successor is plausible
//                      too — and the host validates the chain (entry[i] == exit[i-1]) exactly as for BAM.
//   bcf_decode_kernel  replaces bcf_unpack (vcf.c:4178-4302), bcf_get_info_values / bcf_get_format_values /
//                      bcf_get_format_string / bcf_get_genotypes (vcf.c:6056-6248) and the per-row column writers of
//                      bcf_read_function (src/bcf_reader.c:1381-1462 core columns, 1542-1733 INFO, 1738-1998 FORMAT
//                      incl. the GT string builder 1904-1957): one CTA per BGZF block, one thread serialises the
//                      site-level fields of a record, all threads spread over its samples.
//
// Output layout: site-level columns hold ONE cell per record and FORMAT columns one cell per (record, sample); the
// tidy expansion (one row per sample, site-level cells repeated: bcf_reader.c:1283-1290) is a stride, not a copy —
// the host shim repeats the 16-byte cells when it fills DuckDB chunks, so neither HBM nor PCIe carries 2,504 copies
// of CHROM/POS/INFO.  LIST columns are duckdb_list_entry{offset,length} cells + children in the batch heap, child
// slots at structural capacity (n values per sample), missing / vector_end values dropped inside the slot.
#include "bcf_decode.cuh"

namespace dh {

namespace {

__device__ __forceinline__ uint32_t b8(const uint8_t *u, int64_t o) { return u[o]; }
__device__ __forceinline__ uint32_t b16(const uint8_t *u, int64_t o) { return (uint32_t)u[o] | ((uint32_t)u[o + 1] << 8); }
__device__ __forceinline__ uint32_t b32(const uint8_t *u, int64_t o) {   // unaligned LE32
    uintptr_t a = (uintptr_t)(u + o);
    const uint32_t *q = reinterpret_cast<const uint32_t *>(a & ~(uintptr_t)3);
    uint32_t sh = (uint32_t)(a & 3u) * 8u;
    uint32_t lo = q[0];
    if (sh == 0) return lo;
    return __funnelshift_r(lo, q[1], sh);
}
__device__ __forceinline__ uint32_t type_shift(uint32_t t) { return t > 7u ? 0u : (0x00232100u >> (4u * t)) & 15u; }   // bcf_type_shift, vcf.c:91
__device__ __forceinline__ uint64_t align16(uint64_t n) { return (n + 15ull) & ~15ull; }

// bcf_dec_typed_int1_safe (vcf.c:1916-1948)
__device__ __forceinline__ bool typed_int1(const uint8_t *u, int64_t &p, int64_t end, int32_t &val) {
    if (end - p < 2) return false;
    const uint32_t t = b8(u, p++) & 15u;
    if (t == BT_INT8) { val = (int32_t)(int8_t)b8(u, p++); return true; }
    if (end - p < (int64_t)(1u << type_shift(t))) return false;
    if (t == BT_INT16) { val = (int32_t)(int16_t)b16(u, p); p += 2; return true; }
    if (t == BT_INT32) { val = (int32_t)b32(u, p); p += 4; return true; }
    return false;
}
// bcf_dec_size_safe (vcf.c:1950-1963)
__device__ __forceinline__ bool dec_size(const uint8_t *u, int64_t &p, int64_t end, int32_t &num, uint32_t &type) {
    if (p >= end) return false;
    const uint32_t b = b8(u, p);
    type = b & 15u;
    if ((b >> 4) != 15u) { ++p; num = (int32_t)(b >> 4); return true; }
    ++p;
    if (!typed_int1(u, p, end, num)) return false;
    return num >= 0;
}
__device__ __forceinline__ int32_t dec_int1(const uint8_t *u, int64_t p, uint32_t t) {   // bcf_dec_int1
    if (t == BT_INT8) return (int32_t)(int8_t)b8(u, p);
    if (t == BT_INT16) return (int32_t)(int16_t)b16(u, p);
    return (int32_t)b32(u, p);
}

struct FmtDesc {
    int64_t data;      // offset of the first sample's values
    int32_t key, n;    // dictionary id, values per sample
    uint32_t type, size;   // BT_*, bytes per sample
    uint64_t slot;     // fixed-region offset (relative to the record's base) of its list children, if it has any
};

struct BRec {
    int64_t off, end, shared_end, indiv, indiv_end;
    int64_t id_p, als_p, flt_p, info_p;
    int64_t pos;
    int32_t rid, rlen, id_n, flt_n;
    uint32_t qual_bits, n_info, n_allele, n_sample, n_fmt, flt_type;
    uint64_t alt_slot, flt_slot, info_slot0, need;
};

constexpr uint32_t IS_INT = (1u << BT_INT8) | (1u << BT_INT16) | (1u << BT_INT32);
constexpr uint32_t IS_VALID = IS_INT | (1u << BT_NULL) | (1u << BT_FLOAT) | (1u << BT_CHAR);

__device__ __forceinline__ bool id_ok(const BcfTables &t, int32_t key) { return key >= 0 && key < t.n_ids && t.id_cells[key].len != 0xffffffffu; }

// Frame + validate the record at `off` as bcf_read1_core + bcf_record_check do, and lay out its fixed-region needs.
// Returns 0, 1 (needs bytes beyond seg_end and the stream continues) or -2 (every failure of bcf_read is -2).
// fd: optional array of BCF_MAX_FMT descriptors filled for the decode kernel.
__device__ int bcf_parse(const uint8_t *u, int64_t off, int64_t seg_end, bool is_final, const BcfTables &t, BRec &r, FmtDesc *fd) {
    const int64_t avail = seg_end - off;
    if (avail < 32) return is_final ? -2 : 1;                         // vcf.c:1879-1882 (avail == 0 never gets here)
    uint32_t l_shared = b32(u, off);
    const uint32_t l_indiv = b32(u, off + 4);
    if (l_shared < 24) return -2;                                     // :1885
    l_shared -= 24;
    const int64_t total = 32 + (int64_t)l_shared + (int64_t)l_indiv;
    if (avail < total) return is_final ? -2 : 1;                      // :1907-1908 short bgzf_read
    r.off = off;
    r.end = off + total;
    r.rid = (int32_t)b32(u, off + 8);
    const uint32_t pos32 = b32(u, off + 12);
    r.pos = pos32 == 0xffffffffu ? -1 : (int64_t)pos32;               // :1895-1896
    r.rlen = (int32_t)b32(u, off + 16);
    r.qual_bits = b32(u, off + 20);
    r.n_info = b16(u, off + 24);
    r.n_allele = b16(u, off + 26);
    const uint32_t w = b32(u, off + 28);
    r.n_sample = w & 0xffffffu;
    r.n_fmt = w >> 24;
    if ((!l_indiv || !r.n_sample) && r.n_fmt) r.n_fmt = 0;            // :1905
    int64_t p = off + 32;
    r.shared_end = p + l_shared;
    r.indiv = r.shared_end;
    r.indiv_end = r.end;
    bool err = false;
    // ---- bcf_record_check (vcf.c:2040-2212) ----
    if (r.rid < 0 || r.rid >= t.n_ctg || t.ctg_cells[r.rid].len == 0xffffffffu) err = true;
    int32_t num;
    uint32_t type;
    const int64_t end = r.shared_end;
    if (!dec_size(u, p, end, num, type)) return -2;                   // ID
    if (type != BT_CHAR) err = true;
    int64_t bytes = (int64_t)num << type_shift(type);
    if (end - p < bytes) return -2;
    r.id_p = p;
    r.id_n = num;
    p += bytes;
    if (r.n_allele < 1) err = true;
    r.als_p = p;
    for (uint32_t i = 0; i < r.n_allele; ++i) {                        // REF and ALT
        if (!dec_size(u, p, end, num, type)) return -2;
        if (type != BT_CHAR) err = true;
        bytes = (int64_t)num << type_shift(type);
        if (end - p < bytes) return -2;
        p += bytes;
    }
    if (!dec_size(u, p, end, num, type)) return -2;                   // FILTER
    r.flt_p = p;
    r.flt_n = 0;
    r.flt_type = type;
    if (num > 0) {
        bytes = (int64_t)num << type_shift(type);
        if (!((1u << type) & IS_INT)) {
            err = true;
            if (end - p < bytes) return -2;
        } else {
            if (end - p < bytes) return -2;
            for (int32_t i = 0; i < num; ++i)
                if (!id_ok(t, dec_int1(u, p + ((int64_t)i << type_shift(type)), type))) err = true;
            r.flt_n = num;
        }
        p += bytes;
    }
    uint64_t need = 0;
    r.alt_slot = need;
    if ((t.core_mask >> BC_ALT) & 1u) need += 16ull * (r.n_allele > 1 ? r.n_allele - 1 : 0);
    r.flt_slot = need;
    if ((t.core_mask >> BC_FILTER) & 1u) need += 16ull * (uint64_t)(r.flt_n > 0 ? r.flt_n : 1);
    r.info_p = p;
    r.info_slot0 = need;
    for (uint32_t i = 0; i < r.n_info; ++i) {                          // INFO
        int32_t key = -1;
        if (!typed_int1(u, p, end, key)) return -2;
        const bool kok = id_ok(t, key);
        if (!kok) err = true;
        if (!dec_size(u, p, end, num, type)) return -2;
        if (!((1u << type) & IS_VALID) || (type == BT_NULL && num > 0)) err = true;
        bytes = (int64_t)num << type_shift(type);
        if (end - p < bytes) return -2;
        p += bytes;
        if (kok) {
            const int c = t.info_col[key];
            if (c >= 0 && t.info_want[c] && (t.info_meta[c] & META_LIST) && (t.info_meta[c] & 3u) != HT_STR) need += align16(4ull * (uint64_t)num);
        }
    }
    p = r.indiv;
    const int64_t iend = r.indiv_end;
    if (r.n_fmt > (uint32_t)BCF_MAX_FMT) return -2;                    // (limit of this implementation)
    for (uint32_t i = 0; i < r.n_fmt; ++i) {                           // FORMAT
        int32_t key = -1;
        if (!typed_int1(u, p, iend, key)) return -2;
        const bool kok = id_ok(t, key);
        if (!kok) err = true;
        if (!dec_size(u, p, iend, num, type)) return -2;
        if (!((1u << type) & IS_VALID) || (type == BT_NULL && num > 0)) err = true;
        bytes = ((int64_t)num << type_shift(type)) * (int64_t)r.n_sample;
        if (iend - p < bytes) { if (t.gt_id >= 0 && key == t.gt_id) err = true; else return -2; bytes = iend - p; }   // updatephasing :1993-1994 / bad_indiv
        if (fd) {
            fd[i].data = p;
            fd[i].key = kok ? key : -1;
            fd[i].n = num;
            fd[i].type = type;
            fd[i].size = (uint32_t)num << type_shift(type);
            fd[i].slot = need;
        }
        p += bytes;
        if (kok) {
            const int c = t.fmt_col[key];
            if (c >= 0 && t.fmt_want[c] && (t.fmt_meta[c] & META_LIST) && (t.fmt_meta[c] & 3u) != HT_STR)
                need += align16(4ull * (uint64_t)num * (uint64_t)t.n_samples);
        }
    }
    // records whose sample count differs from the header's cannot be addressed per sample (htslib would read past the
    // record): treated as invalid here
    if (r.n_fmt && r.n_sample != t.n_samples) err = true;
    r.need = need;
    return err ? -2 : 0;
}

// cheap "does a record start here" test for speculative block entries; never decides correctness
__device__ __forceinline__ bool bcf_plausible(const uint8_t *u, int64_t off, int64_t seg_end, const BcfTables &t) {
    if (seg_end - off < 32) return false;
    const uint32_t w = b32(u, off + 28);
    if ((w & 0xffffffu) != t.n_samples && (w & 0xffffffu) != 0u) return false;
    const uint32_t l_shared = b32(u, off), l_indiv = b32(u, off + 4);
    if (l_shared < 24u + 3u || l_shared > (1u << 26) || l_indiv > (1u << 30)) return false;
    const int32_t rid = (int32_t)b32(u, off + 8);
    if (rid < 0 || rid >= t.n_ctg) return false;
    const uint32_t n_allele = b16(u, off + 26);
    if (n_allele < 1) return false;
    if ((w >> 24) > (uint32_t)BCF_MAX_FMT) return false;
    if ((w & 0xffffffu) && (w >> 24) && (uint64_t)l_indiv < (uint64_t)(w >> 24) * 2ull) return false;
    // the ID descriptor must be a character vector
    const uint32_t d = b8(u, off + 32);
    if ((d & 15u) != BT_CHAR) return false;
    return true;
}

}  // namespace

// =======================================================================================================
// boundary walk + validation: one thread per BGZF block
// =======================================================================================================
__global__ void __launch_bounds__(128) bcf_walk_kernel(const WalkParams p, const BcfTables t) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.n) return;
    if (p.todo && !p.todo[i]) return;
    const uint8_t *u = p.u;
    const int64_t start = p.blk_start[i], bend = p.blk_start[i + 1];
    const int64_t lim = bend < p.limit_end ? bend : p.limit_end;
    int64_t off = p.entry_in[i];
    BRec r;
    if (off == INT64_MIN) {
        // speculate: the first offset in the block where a structurally valid record starts
        int64_t s = start;
        for (; s < bend; ++s) {
            if (!bcf_plausible(u, s, p.seg_end, t)) continue;
            if (bcf_parse(u, s, p.seg_end, false, t, r, nullptr) < 0) continue;
            break;
        }
        off = s;   // == bend if nothing starts here (a long record spans the block)
    }
    uint16_t *rec = p.rec + ((start - p.rec_bias) / 32 + (int64_t)i);
    BlockWalk w;
    w.entry = off;
    w.count = 0;
    w.map_entries = w.map_heap = 0;
    w.heap_bytes = 0;
    w.stop = 0;
    while (off < lim) {
        if (p.is_final && off == p.seg_end) break;
        const int rc = bcf_parse(u, off, p.seg_end, p.is_final != 0, t, r, nullptr);
        if (rc != 0) { w.stop = rc; break; }
        if (p.filt_mode) {
            // region query: bcf_readrec hands hts_itr_next (tid, beg, end) = (rid, pos, pos + rlen) (htslib/vcf.c:2267-2276); the
            // iterator keeps `end > beg_q && end_q > beg` on the queried contig (hts.c:4304-4313)
            bool keep = false;
            for (uint32_t q = 0; q < p.filt_n && !keep; ++q)
                keep = p.filt[q].tid == r.rid && r.pos + (int64_t)r.rlen > p.filt[q].beg && p.filt[q].end > r.pos;
            if (!keep) { off = r.end; continue; }
        }
        w.heap_bytes += r.need;
        rec[w.count] = (uint16_t)(off - start);
        w.count++;
        off = r.end;
    }
    w.exit_off = off;
    p.out[i] = w;
}

// =======================================================================================================
// columnar decode: one CTA per BGZF block
// =======================================================================================================
namespace {

constexpr uint32_t I32_MISSING = 0x80000000u, I32_VEND = 0x80000001u, F32_MISSING = 0x7F800001u, F32_VEND = 0x7F800002u;

// value j of a typed vector as the 32-bit word bcf_get_info_values / bcf_get_format_values store (vcf.c:6098-6131,
// 6219-6244): narrow integers widened, their missing / vector_end sentinels mapped to the 32-bit ones; floats by bits
__device__ __forceinline__ uint32_t word_at(const uint8_t *u, int64_t base, int32_t j, uint32_t type) {
    if (type == BT_INT8) { const uint32_t v = b8(u, base + j); return v == 0x80u ? I32_MISSING : v == 0x81u ? I32_VEND : (uint32_t)(int32_t)(int8_t)v; }
    if (type == BT_INT16) { const uint32_t v = b16(u, base + 2ll * j); return v == 0x8000u ? I32_MISSING : v == 0x8001u ? I32_VEND : (uint32_t)(int32_t)(int16_t)v; }
    return b32(u, base + 4ll * j);
}
__device__ __forceinline__ bool word_is_vend(uint32_t w, uint32_t type) { return type == BT_FLOAT ? w == F32_VEND : w == I32_VEND; }
__device__ __forceinline__ bool numeric_type(uint32_t type) { return type == BT_INT8 || type == BT_INT16 || type == BT_INT32 || type == BT_FLOAT; }

__device__ __forceinline__ void put_cell(uint8_t *col, uint64_t row, uint32_t len, uint32_t w0, uint32_t w1, uint32_t w2) {
    reinterpret_cast<uint4 *>(col)[row] = make_uint4(len, w0, w1, w2);
}
// bytes from the bump region (16-byte granules); UINT64_MAX when it is exhausted (the host re-runs the batch with more)
__device__ __forceinline__ uint64_t bump_alloc(const BcfDecodeParams &p, uint64_t bytes) {
    const uint64_t a = align16(bytes);
    const uint64_t o = atomicAdd(p.bump_used, (unsigned long long)a);
    return o + a <= p.bump_cap ? o : UINT64_MAX;
}
// VARCHAR cell for the bytes u[src, src+len): <= 12 inlined, else copied into the bump region
__device__ void put_str(const BcfDecodeParams &p, uint8_t *col, uint64_t row, const uint8_t *u, int64_t src, uint32_t len) {
    if (len <= 12u) {
        uint32_t w[3] = {0u, 0u, 0u};
        for (uint32_t k = 0; k < len; ++k) w[k >> 2] |= b8(u, src + k) << (8u * (k & 3u));
        put_cell(col, row, len, w[0], w[1], w[2]);
        return;
    }
    const uint64_t o = bump_alloc(p, len);
    if (o == UINT64_MAX) { put_cell(col, row, 0u, 0u, 0u, 0u); return; }
    uint8_t *dst = p.heap + p.fixed_bytes + o;
    for (uint32_t k = 0; k < len; ++k) dst[k] = (uint8_t)b8(u, src + k);
    const uint64_t hp = p.heap_host + p.fixed_bytes + o;
    put_cell(col, row, len, b32(u, src), (uint32_t)hp, (uint32_t)(hp >> 32));
}
// length of a character vector as bcf_fmt_array prints it (vcf.c: up to the first NUL)
__device__ __forceinline__ uint32_t cstr_len(const uint8_t *u, int64_t src, uint32_t n) {
    uint32_t k = 0;
    while (k < n && b8(u, src + k) != 0u) ++k;
    return k;
}
__device__ __forceinline__ void set_bit(unsigned long long *v, uint64_t row) {
    atomicOr(reinterpret_cast<unsigned int *>(v) + (row >> 5), 1u << (row & 31u));
}
// validity bits of 32 consecutive rows starting at row0 (any alignment): two 32-bit atomics per warp
__device__ __forceinline__ void set_bits32(unsigned long long *v, uint64_t row0, uint32_t mask) {
    if (!mask) return;
    unsigned int *v32 = reinterpret_cast<unsigned int *>(v);
    const uint32_t sh = (uint32_t)(row0 & 31u);
    atomicOr(&v32[row0 >> 5], mask << sh);
    if (sh && (mask >> (32u - sh))) atomicOr(&v32[(row0 >> 5) + 1u], mask >> (32u - sh));
}
__device__ __forceinline__ uint32_t dec_len(int32_t v) {   // characters kputw prints
    uint32_t a = v < 0 ? (uint32_t)(-(int64_t)v) : (uint32_t)v, d = 1;
    while (a >= 10u) { a /= 10u; ++d; }
    return d + (v < 0 ? 1u : 0u);
}
__device__ __forceinline__ void dec_put(uint8_t *dst, int32_t v, uint32_t n) {   // n = dec_len(v)
    uint32_t a = v < 0 ? (uint32_t)(-(int64_t)v) : (uint32_t)v;
    for (uint32_t k = n; k-- > (v < 0 ? 1u : 0u);) { dst[k] = (uint8_t)('0' + a % 10u); a /= 10u; }
    if (v < 0) dst[0] = '-';
}

}  // namespace

constexpr int BCF_MAX_FMT_COLS = 256;   // FORMAT columns a header may declare (host refuses more)
constexpr int BCF_MAX_INFO_COLS = 1024;

__global__ void __launch_bounds__(256) bcf_decode_kernel(const __grid_constant__ BcfDecodeParams p, const BcfTables t) {
    // Two copies of the per-record state: thread 0 parses record k + 1 and writes its site-level cells while warps 1..7 write the
    // FORMAT cells of record k (r1: thread 0 did this with 255 threads waiting at the barrier).
    __shared__ BRec r_s[2];
    __shared__ FmtDesc fd_s[2][BCF_MAX_FMT];
    __shared__ int16_t col2desc_s[2][BCF_MAX_FMT_COLS];
    __shared__ uint32_t seen[BCF_MAX_INFO_COLS / 32];
    __shared__ uint64_t rec_base_s[2];
    const uint32_t bi = p.b_lo + blockIdx.x;
    if (bi >= p.b_hi) return;
    const uint32_t count = p.walk[bi].count;
    if (!count) return;
    const uint8_t *u = p.u;
    const int64_t bstart = p.blk_start[bi];
    const uint16_t *rec = p.rec + ((bstart - p.rec_bias) / 32 + (int64_t)bi);
    const uint32_t ns = t.n_samples, tid = threadIdx.x, lane = tid & 31u;
    uint64_t hp = p.heap_base[bi];   // thread 0 only

    // site-level part of record k (thread 0): parse, core and INFO columns, FORMAT descriptors for the other warps
    auto site_level = [&](const uint32_t k, const int q) {
        BRec &r = r_s[q];
        FmtDesc *fd = fd_s[q];
        int16_t *col2desc = col2desc_s[q];
        const uint64_t R = (uint64_t)p.row_base[bi] + k;
        {
            bcf_parse(u, bstart + rec[k], INT64_MAX / 2, true, t, r, fd);   // validated by the walk: cannot fail
            const uint64_t base = hp;
            rec_base_s[q] = base;
            hp += r.need;
            for (int c = 0; c < t.n_fmt; ++c) col2desc[c] = -1;
            for (uint32_t i = 0; i < r.n_fmt; ++i) {
                const int c = fd[i].key >= 0 ? t.fmt_col[fd[i].key] : -1;
                if (c >= 0 && col2desc[c] < 0) col2desc[c] = (int16_t)i;   // first entry with this key wins (vcf.c:6193-6195)
            }
            // ---------------- core columns (src/bcf_reader.c:1381-1462) ----------------
            if (p.core[BC_CHROM]) reinterpret_cast<StrCell *>(p.core[BC_CHROM])[R] = t.ctg_cells[r.rid];
            if (p.core[BC_POS]) reinterpret_cast<int64_t *>(p.core[BC_POS])[R] = r.pos + 1;
            if (p.core[BC_ID]) {
                const uint32_t len = cstr_len(u, r.id_p, (uint32_t)r.id_n);
                const bool dot = r.id_n == 0 || (len == 1u && b8(u, r.id_p) == '.');
                if (!dot) { put_str(p, p.core[BC_ID], R, u, r.id_p, len); set_bit(p.core_valid[BC_ID], R); }
            }
            {
                int64_t ap = r.als_p;
                uint8_t *alt_cells = p.heap + base + r.alt_slot;
                for (uint32_t a = 0; a < r.n_allele; ++a) {
                    int32_t num;
                    uint32_t type;
                    dec_size(u, ap, r.shared_end, num, type);
                    const uint32_t len = cstr_len(u, ap, (uint32_t)num);
                    uint8_t *col = a == 0 ? p.core[BC_REF] : (p.core[BC_ALT] ? alt_cells : nullptr);
                    const uint64_t row = a == 0 ? R : (uint64_t)(a - 1);
                    if (col) {
                        if (num == 0) put_cell(col, row, 1u, (uint32_t)'.', 0u, 0u);   // bcf_fmt_array: n == 0 prints "."
                        else put_str(p, col, row, u, ap, len);
                    }
                    ap += (int64_t)num << type_shift(type);
                }
                if (p.core[BC_ALT]) {
                    ListEntry e;
                    e.offset = (base + r.alt_slot) / 16;
                    e.length = r.n_allele > 1 ? r.n_allele - 1 : 0;
                    reinterpret_cast<ListEntry *>(p.core[BC_ALT])[R] = e;
                }
            }
            if (p.core[BC_QUAL]) {
                double q = 0.0;
                if (r.qual_bits != F32_MISSING) { q = (double)__uint_as_float(r.qual_bits); set_bit(p.core_valid[BC_QUAL], R); }
                reinterpret_cast<double *>(p.core[BC_QUAL])[R] = q;
            }
            if (p.core[BC_FILTER]) {
                StrCell *cells = reinterpret_cast<StrCell *>(p.heap + base + r.flt_slot);
                ListEntry e;
                e.offset = (base + r.flt_slot) / 16;
                if (r.flt_n == 0) {
                    StrCell pass;
                    pass.len = 4u; pass.w[0] = 0x53534150u; pass.w[1] = 0u; pass.w[2] = 0u;   // "PASS"
                    cells[0] = pass;
                    e.length = 1;
                } else {
                    for (int32_t f = 0; f < r.flt_n; ++f)
                        cells[f] = t.id_cells[dec_int1(u, r.flt_p + ((int64_t)f << type_shift(r.flt_type)), r.flt_type)];
                    e.length = (uint64_t)r.flt_n;
                }
                reinterpret_cast<ListEntry *>(p.core[BC_FILTER])[R] = e;
            }
            // ---------------- INFO columns (src/bcf_reader.c:1542-1733; vcf.c:6056-6138) ----------------
            for (int c = 0; c < (t.n_info + 31) / 32; ++c) seen[c] = 0u;
            int64_t ip = r.info_p;
            uint64_t slot = r.info_slot0;
            for (uint32_t i = 0; i < r.n_info; ++i) {
                int32_t key = -1, num;
                uint32_t type;
                typed_int1(u, ip, r.shared_end, key);
                dec_size(u, ip, r.shared_end, num, type);
                const int64_t vp = ip;
                ip += (int64_t)num << type_shift(type);
                const int c = t.info_col[key];
                if (c < 0 || !t.info_want[c]) continue;
                const uint32_t meta = t.info_meta[c], ht = meta & 3u;
                const bool is_list = (meta & META_LIST) != 0;
                uint64_t my_slot = slot;
                if (is_list && ht != HT_STR) slot += align16(4ull * (uint64_t)num);
                if ((seen[c >> 5] >> (c & 31)) & 1u) continue;           // first entry with this key wins (vcf.c:6063-6065)
                seen[c >> 5] |= 1u << (c & 31);
                uint8_t *col = p.info[c];
                unsigned long long *val = p.info_valid[c];
                if (ht == HT_FLAG) { col[R] = 1; continue; }
                if (ht == HT_STR) {
                    if (num <= 0) continue;                                // ret_info = info->len must be > 0
                    const uint32_t len = cstr_len(u, vp, (uint32_t)num);
                    if (len == 1u && b8(u, vp) == '.') continue;           // "." -> NULL
                    if (!is_list) { put_str(p, col, R, u, vp, len); set_bit(val, R); continue; }
                    // process_comma_separated_list (bcf_reader.c:996-1061)
                    uint32_t tokens = 0, ts = 0;
                    for (uint32_t q = 0; q < len; ++q) if (b8(u, vp + q) == ',') { ++tokens; ts = q + 1; }
                    if (len > ts) ++tokens;
                    ListEntry e;
                    e.offset = 0;
                    e.length = tokens;
                    if (tokens) {
                        const uint64_t o = bump_alloc(p, 16ull * tokens);
                        if (o == UINT64_MAX) e.length = 0;
                        else {
                            uint8_t *cells = p.heap + p.fixed_bytes + o;
                            e.offset = (p.fixed_bytes + o) / 16;
                            uint32_t wi = 0;
                            ts = 0;
                            for (uint32_t q = 0; q < len; ++q)
                                if (b8(u, vp + q) == ',') { put_str(p, cells, wi++, u, vp + ts, q - ts); ts = q + 1; }
                            if (len > ts) put_str(p, cells, wi, u, vp + ts, len - ts);
                        }
                    }
                    reinterpret_cast<ListEntry *>(col)[R] = e;
                    set_bit(val, R);
                    continue;
                }
                // Integer / Float: values up to the first vector_end (vcf.c:6098-6110), by the RECORD's type; the column
                // reads the 32-bit words by the HEADER's type
                if (!numeric_type(type)) continue;
                const uint32_t miss = ht == HT_INT ? I32_MISSING : F32_MISSING, vend = ht == HT_INT ? I32_VEND : F32_VEND;
                int32_t ret = 0;
                while (ret < num && !word_is_vend(word_at(u, vp, ret, type), type)) ++ret;
                if (ret <= 0) continue;
                if (!is_list) {
                    const uint32_t w0 = word_at(u, vp, 0, type);
                    if (w0 != miss) { reinterpret_cast<uint32_t *>(col)[R] = w0; set_bit(val, R); }
                    continue;
                }
                uint32_t *child = reinterpret_cast<uint32_t *>(p.heap + base + my_slot);
                uint32_t n_out = 0;
                for (int32_t j = 0; j < ret; ++j) {
                    const uint32_t w = word_at(u, vp, j, type);
                    if (w != miss && w != vend) child[n_out++] = w;
                }
                ListEntry e;
                e.offset = (base + my_slot) / 4;
                e.length = n_out;
                reinterpret_cast<ListEntry *>(col)[R] = e;
                set_bit(val, R);
            }
        }
    };

    constexpr uint32_t WORKERS = 256u - 32u;       // warps 1..7
    if (tid == 0) site_level(0, 0);
    __syncthreads();
    for (uint32_t k = 0; k < count; ++k) {
        const uint64_t R = (uint64_t)p.row_base[bi] + k;
        const int q = (int)(k & 1u);
        const BRec &r = r_s[q];
        const FmtDesc *fd = fd_s[q];
        const int16_t *col2desc = col2desc_s[q];
        if (tid < 32u) {
            if (tid == 0 && k + 1 < count) site_level(k + 1, q ^ 1);
        } else if (t.any_fmt && r.n_fmt) {
            // ---------------- FORMAT columns: warps 1..7 over the samples (bcf_reader.c:1738-1998) ----------------
            const uint32_t wtid = tid - 32u;
            const uint64_t base = rec_base_s[q];
            for (int c = 0; c < t.n_fmt; ++c) {
                if (!t.fmt_want[c]) continue;
                const int di = col2desc[c];
                if (di < 0) continue;
                const FmtDesc d = fd[di];
                const uint32_t meta = t.fmt_meta[c], ht = meta & 3u;
                const bool is_list = (meta & META_LIST) != 0, is_gt = (meta & META_GT) != 0;
                uint8_t *col = p.fmt[c];
                unsigned long long *val = p.fmt_valid[c];
                for (uint32_t s0 = 0; s0 < ns; s0 += WORKERS) {
                    const uint32_t s = s0 + wtid;
                    const uint64_t row = R * ns + s;
                    bool valid = false;
                    if (s < ns) {
                        const int64_t vp = d.data + (int64_t)s * d.size;
                        if (ht == HT_STR && !is_gt) {
                            if (!is_list) {   // bcf_get_format_string (vcf.c:6140-6177): n bytes per sample, C-string semantics
                                const int64_t sp = d.data + (int64_t)s * d.n;
                                put_str(p, col, row, u, sp, cstr_len(u, sp, (uint32_t)d.n));
                                valid = true;
                            }
                        } else if (d.n > 0 && numeric_type(d.type)) {
                            if (is_gt) {
                                // GT string builder (bcf_reader.c:1904-1957; bcf_gt_* macros htslib/vcf.h:1033-1038)
                                uint32_t len = 0;
                                for (int32_t a = 0; a < d.n && d.type != BT_FLOAT; ++a) {   // (a float GT vector is not a genotype: NULL)
                                    const uint32_t w = word_at(u, vp, a, d.type);
                                    if (w == I32_VEND) break;
                                    const int32_t v = (int32_t)w;
                                    len += (a > 0 ? 1u : 0u) + ((v >> 1) ? dec_len((v >> 1) - 1) : 1u);
                                }
                                if (len > 0 && p.compact) {
                                    // compact GT cell: <= 3 characters ("0|1") ride in the cell, [25:24] = length, [23:0] = the bytes; longer
                                    // strings (alleles >= 10, polyploid calls) go to the bump region as {uint32 length, bytes}, cell = 1 << 31 | heap offset
                                    uint8_t tmp[12];
                                    uint8_t *dst = tmp;
                                    uint64_t o = 0;
                                    if (len > 3u) { o = bump_alloc(p, 4ull + len); dst = o == UINT64_MAX ? nullptr : p.heap + p.fixed_bytes + o + 4; }
                                    uint32_t cell = 0u;
                                    if (dst) {
                                        uint32_t q = 0;
                                        for (int32_t a = 0; a < d.n; ++a) {
                                            const uint32_t w = word_at(u, vp, a, d.type);
                                            if (w == I32_VEND) break;
                                            const int32_t v = (int32_t)w;
                                            if (a > 0) dst[q++] = (v & 1) ? '|' : '/';
                                            if (!(v >> 1)) dst[q++] = '.';
                                            else { const uint32_t n = dec_len((v >> 1) - 1); dec_put(dst + q, (v >> 1) - 1, n); q += n; }
                                        }
                                        if (len <= 3u) cell = (len << 24) | (uint32_t)tmp[0] | (len > 1u ? (uint32_t)tmp[1] << 8 : 0u) | (len > 2u ? (uint32_t)tmp[2] << 16 : 0u);
                                        else {
                                            *reinterpret_cast<uint32_t *>(dst - 4) = len;
                                            cell = 0x80000000u | (uint32_t)(p.fixed_bytes + o);
                                        }
                                    }
                                    reinterpret_cast<uint32_t *>(col)[row] = cell;
                                    valid = true;
                                } else if (len > 0) {
                                    uint8_t tmp[12];
                                    uint8_t *dst = tmp;
                                    uint64_t o = 0;
                                    if (len > 12u) { o = bump_alloc(p, len); dst = o == UINT64_MAX ? nullptr : p.heap + p.fixed_bytes + o; }
                                    if (dst) {
                                        uint32_t q = 0;
                                        for (int32_t a = 0; a < d.n; ++a) {
                                            const uint32_t w = word_at(u, vp, a, d.type);
                                            if (w == I32_VEND) break;
                                            const int32_t v = (int32_t)w;
                                            if (a > 0) dst[q++] = (v & 1) ? '|' : '/';
                                            if (!(v >> 1)) dst[q++] = '.';
                                            else { const uint32_t n = dec_len((v >> 1) - 1); dec_put(dst + q, (v >> 1) - 1, n); q += n; }
                                        }
                                        if (len <= 12u) {
                                            uint32_t w3[3] = {0u, 0u, 0u};
                                            for (uint32_t b = 0; b < len; ++b) w3[b >> 2] |= (uint32_t)tmp[b] << (8u * (b & 3u));
                                            put_cell(col, row, len, w3[0], w3[1], w3[2]);
                                        } else {
                                            const uint64_t hpn = p.heap_host + p.fixed_bytes + o;
                                            put_cell(col, row, len, (uint32_t)dst[0] | ((uint32_t)dst[1] << 8) | ((uint32_t)dst[2] << 16) | ((uint32_t)dst[3] << 24),
                                                     (uint32_t)hpn, (uint32_t)(hpn >> 32));
                                        }
                                        valid = true;
                                    } else { put_cell(col, row, 0u, 0u, 0u, 0u); valid = true; }
                                }
                            } else if (ht == HT_INT || ht == HT_REAL) {
                                const uint32_t miss = ht == HT_INT ? I32_MISSING : F32_MISSING, vend = ht == HT_INT ? I32_VEND : F32_VEND;
                                if (!is_list) {
                                    const uint32_t w0 = word_at(u, vp, 0, d.type);
                                    reinterpret_cast<uint32_t *>(col)[row] = w0;
                                    valid = w0 != miss;
                                } else {
                                    uint32_t *child = reinterpret_cast<uint32_t *>(p.heap + base + d.slot) + (uint64_t)s * (uint32_t)d.n;
                                    uint32_t n_out = 0;
                                    for (int32_t j = 0; j < d.n; ++j) {
                                        const uint32_t w = word_at(u, vp, j, d.type);
                                        if (word_is_vend(w, d.type)) break;     // the rest of the vector is vector_end (vcf.c:6233-6236)
                                        if (w != miss && w != vend) child[n_out++] = w;
                                    }
                                    if (p.compact) {
                                        reinterpret_cast<uint16_t *>(col)[row] = (uint16_t)(n_out < 65535u ? n_out : 65535u);
                                        if (s == 0u) reinterpret_cast<uint2 *>(p.fmt_rec[c])[R] = make_uint2((uint32_t)((base + d.slot) / 4), (uint32_t)d.n);
                                    } else {
                                        ListEntry e;
                                        e.offset = (base + d.slot) / 4 + (uint64_t)s * (uint32_t)d.n;
                                        e.length = n_out;
                                        reinterpret_cast<ListEntry *>(col)[row] = e;
                                    }
                                    valid = true;
                                }
                            }
                        }
                    }
                    const uint32_t mask = __ballot_sync(0xffffffffu, valid);
                    if (lane == 0) set_bits32(val, R * ns + (s0 + (wtid & ~31u)), mask);
                }
            }
        }
        __syncthreads();
    }
}

cudaError_t launch_bcf_walk(const WalkParams &p, const BcfTables &t, cudaStream_t stream) {
    if (p.n == 0) return cudaSuccess;
    bcf_walk_kernel<<<(p.n + 127) / 128, 128, 0, stream>>>(p, t);
    return cudaGetLastError();
}
cudaError_t launch_bcf_decode(const BcfDecodeParams &p, const BcfTables &t, cudaStream_t stream) {
    if (p.b_hi <= p.b_lo) return cudaSuccess;
    bcf_decode_kernel<<<p.b_hi - p.b_lo, 256, 0, stream>>>(p, t);
    return cudaGetLastError();
}

}  // namespace dh
