// bam_decode.cu — sm_100a kernels for the BAM half of the read_bam hot path.
//
//   bam_walk_kernel    replaces the record FRAMING + VALIDATION of bam_read1 (htslib/sam.c:779-855, incl. the
//                      bam_tag2cigar fast path :675-730 and the CIGAR/qlen check :841-852) and sam_read1_bam's
//                      tid range check (:4124-4134), run for every BGZF block in parallel: one thread per block
//                      chains `off += 4 + block_size` from a speculative first-record offset (block start for
//                      htslib-written files, else the first offset that looks like a record header), counts the
//                      records that start in the block and the string-heap bytes they need.  The host validates
//                      the speculation (entry[i] == exit[i-1], the pointer-jumping check) and re-launches the walk
//                      for the few blocks that guessed wrong.
//   bam_decode_kernel  replaces the per-row, per-column append loop of bam_read_function
//                      (src/bam_reader.c:747-1031): one warp per BGZF block walks its records and writes DuckDB
//                      flat vectors (FLAG/POS/MAPQ/PNEXT/TLEN), duckdb_string_t cells + string heap (QNAME, CIGAR
//                      text = cigar_to_kstring :375-383, SEQ = seq_to_string :390-394, QUAL = qual_to_string
//                      :400-404, RNAME/RNEXT dictionary cells, READ_GROUP_ID, SAMPLE_ID :879-918) and the typed
//                      SAM tag columns (:920-966; aux walk = bam_aux_get/skip_aux htslib/sam.c:4785-4855,
//                      conversions :5093-5171) with validity bitmasks.
//
// Integer/byte work only: HBM-bound, no tensor cores.  Layout: records of one block are consecutive rows, so a
// warp's column stores are contiguous per column; the long strings of a record (QNAME, SEQ, QUAL, ...) are
// consecutive 4-byte-aligned slots of ONE heap, so a warp writes ~320 contiguous bytes per record.
#include "bam_decode.cuh"

namespace dh {

// ---- byte-stream loads -------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t ld8(const uint8_t *u, int64_t o) { return u[o]; }
__device__ __forceinline__ uint32_t ld32u(const uint8_t *u, int64_t o) {  // unaligned LE32 (may touch 3 bytes beyond)
    uintptr_t a = (uintptr_t)(u + o);
    const uint32_t *q = reinterpret_cast<const uint32_t *>(a & ~(uintptr_t)3);
    uint32_t sh = (uint32_t)(a & 3u) * 8u;
    uint32_t lo = q[0];
    if (sh == 0) return lo;
    return __funnelshift_r(lo, q[1], sh);
}
__device__ __forceinline__ uint32_t mask_bytes(uint32_t v, uint32_t nvalid) {  // keep the low nvalid (0..4) bytes
    return nvalid >= 4 ? v : (v & ((1u << (8u * nvalid)) - 1u));
}
__device__ __forceinline__ uint32_t align4(uint32_t n) { return (n + 3u) & ~3u; }

__device__ __forceinline__ int aux_type2size(uint32_t t) {  // htslib/sam.c:732-748
    switch (t) {
    case 'A': case 'c': case 'C': return 1;
    case 's': case 'S': return 2;
    case 'i': case 'I': case 'f': return 4;
    case 'd': return 8;
    case 'Z': case 'H': case 'B': return (int)t;
    default: return 0;
    }
}

struct Rec {
    int64_t off, end;        // record start (block_size word) and end
    int64_t qname, cigar, seq, qual, aux;
    int64_t cg_lo, cg_hi;    // aux byte range of a consumed CG tag (cut out of the record by bam_tag2cigar), or INT64_MIN
    uint32_t block_len, l_qname, n_cigar, cig_n, flag, mapq;
    int32_t tid, pos, l_qseq, mtid, mpos, isize;
    uint32_t cigar_text_len; // exact length of the CIGAR text (0 if cig_n == 0)
};

// Offsets may be negative (bytes carried over from the previous segment), so "none/corrupt" is NONE, never -1.
constexpr int64_t NONE = INT64_MIN;

// scalar skip_aux (sam.c:4785-4809): s = offset of the type byte.  Returns next offset, or NONE if corrupt.
// *complete: for Z/H, whether a NUL terminator was found.
__device__ __forceinline__ int64_t skip_aux_scalar(const uint8_t *u, int64_t s, int64_t end, bool *complete) {
    *complete = true;
    if (s >= end) return end;
    int size = aux_type2size(ld8(u, s));
    ++s;
    if (size == 'Z' || size == 'H') {
        while (s < end && ld8(u, s) != 0) ++s;
        if (s < end) return s + 1;
        *complete = false;
        return end;
    }
    if (size == 'B') {
        if (end - s < 5) return NONE;
        size = aux_type2size(ld8(u, s));
        ++s;
        uint32_t n = ld32u(u, s);
        s += 4;
        if (size == 0 || (uint64_t)(end - s) < (uint64_t)size * n) return NONE;
        return s + (int64_t)((uint64_t)size * n);
    }
    if (size == 0) return NONE;
    if (end - s < size) return NONE;
    return s + size;
}

// scalar bam_aux_get (sam.c:4832-4855) over aux bytes [a0, end): returns offset of the type byte or -1
__device__ __forceinline__ int64_t aux_get_scalar(const uint8_t *u, int64_t a0, int64_t end, uint32_t t0, uint32_t t1) {
    if (end - a0 <= 2) return NONE;
    int64_t s = a0 + 2;
    for (;;) {
        bool complete;
        int64_t nx = skip_aux_scalar(u, s, end, &complete);
        if (ld8(u, s - 2) == t0 && ld8(u, s - 1) == t1) {
            if (nx == NONE) return NONE;
            uint32_t ty = ld8(u, s);
            if ((ty == 'Z' || ty == 'H') && !complete) return NONE;
            return s;
        }
        if (nx == NONE || end - nx <= 2) return NONE;
        s = nx + 2;
    }
}

__device__ __forceinline__ uint32_t dec_digits(uint32_t v) {
    uint32_t d = 1;
    d += v >= 10u; d += v >= 100u; d += v >= 1000u; d += v >= 10000u; d += v >= 100000u;
    d += v >= 1000000u; d += v >= 10000000u; d += v >= 100000000u; d += v >= 1000000000u;
    return d;
}

// Frame + validate the record at `off` exactly as bam_read1 + sam_read1_bam do.
// Returns 0 (ok), 1 (needs bytes beyond seg_end and the stream continues), or the negative htslib error code.
__device__ __forceinline__ int parse_record(const uint8_t *u, int64_t off, int64_t seg_end, bool is_final, int32_t n_ref,
                                            Rec &r) {
    int64_t avail = seg_end - off;
    if (avail < 4) return is_final ? -2 : 1;                 // sam.c:789-792 truncated
    int32_t block_len = (int32_t)ld32u(u, off);
    if (block_len < 32) return -4;                           // :794
    if (avail - 4 < 32) return is_final ? -3 : 1;            // :801
    int64_t x = off + 4;
    r.off = off;
    r.block_len = (uint32_t)block_len;
    r.end = x + block_len;
    r.tid = (int32_t)ld32u(u, x);
    r.pos = (int32_t)ld32u(u, x + 4);
    uint32_t x2 = ld32u(u, x + 8), x3 = ld32u(u, x + 12);
    r.l_qname = x2 & 0xffu;
    r.mapq = (x2 >> 8) & 0xffu;
    r.flag = x3 >> 16;
    r.n_cigar = x3 & 0xffffu;
    r.l_qseq = (int32_t)ld32u(u, x + 16);
    r.mtid = (int32_t)ld32u(u, x + 20);
    r.mpos = (int32_t)ld32u(u, x + 24);
    r.isize = (int32_t)ld32u(u, x + 28);
    uint32_t l_extranul = (r.l_qname & 3u) ? 4u - (r.l_qname & 3u) : 0u;
    uint32_t new_l_data = (uint32_t)block_len - 32u + l_extranul;
    if (new_l_data > 2147483647u || r.l_qseq < 0 || r.l_qname < 1) return -4;   // :819
    if (((uint64_t)r.n_cigar << 2) + r.l_qname + l_extranul + (((uint64_t)r.l_qseq + 1) >> 1) + (uint64_t)r.l_qseq >
        (uint64_t)new_l_data)
        return -4;                                           // :820-823
    if (avail - 4 < (int64_t)block_len) return is_final ? -4 : 1;  // :827-835 short read
    r.qname = x + 32;
    r.cigar = r.qname + r.l_qname;
    r.seq = r.cigar + 4ll * r.n_cigar;
    r.qual = r.seq + (((int64_t)r.l_qseq + 1) >> 1);
    r.aux = r.qual + r.l_qseq;
    r.cig_n = r.n_cigar;
    r.cg_lo = r.cg_hi = NONE;
    // bam_tag2cigar (sam.c:675-730): fast-path test first
    if (r.n_cigar && ld32u(u, r.cigar) == (4u | ((uint32_t)r.l_qseq << 4)) && r.tid >= 0 && r.pos >= 0) {
        int64_t CG = aux_get_scalar(u, r.aux, r.end, 'C', 'G');
        if (CG != NONE && ld8(u, CG) == 'B' && (ld8(u, CG + 1) == 'I' || ld8(u, CG + 1) == 'i')) {
            uint32_t CG_len = ld32u(u, CG + 2);
            if (CG_len >= r.n_cigar && CG_len < (1u << 29)) {
                r.cig_n = CG_len;
                r.cigar = CG + 6;
                r.cg_lo = CG - 2;
                r.cg_hi = CG + 6 + 4ll * CG_len;
            }
        }
    }
    r.cigar_text_len = 0;
    if (r.cig_n) {                                           // :841-852
        int64_t qlen = 0;
        uint32_t tl = 0;
        for (uint32_t i = 0; i < r.cig_n; ++i) {
            uint32_t c = ld32u(u, r.cigar + 4ll * i);
            if ((0x3C1A7u >> ((c & 15u) << 1)) & 1u) qlen += c >> 4;  // bam_cigar_type bit 0: consumes query
            tl += dec_digits(c >> 4) + 1u;
        }
        r.cigar_text_len = tl;
        if (r.l_qseq > 0 && !(r.flag & 4u) && qlen != r.l_qseq) return -4;
    }
    if (r.tid >= n_ref || r.tid < -1 || r.mtid >= n_ref || r.mtid < -1) return -3;  // sam.c:4127-4131
    return 0;
}

// cheap "does a record header start here" test for speculative block entries (never decides correctness:
// the host checks every speculation against the predecessor's exit)
__device__ __forceinline__ bool plausible_header(const uint8_t *u, int64_t off, int64_t seg_end, int32_t n_ref) {
    if (seg_end - off < 36) return false;
    uint32_t bl = ld32u(u, off);
    if (bl < 32u || bl >= (1u << 28)) return false;
    int32_t tid = (int32_t)ld32u(u, off + 4), pos = (int32_t)ld32u(u, off + 8);
    uint32_t x2 = ld32u(u, off + 12), x3 = ld32u(u, off + 16);
    int32_t l_qseq = (int32_t)ld32u(u, off + 20), mtid = (int32_t)ld32u(u, off + 24), mpos = (int32_t)ld32u(u, off + 28);
    uint32_t l_qname = x2 & 0xffu, n_cigar = x3 & 0xffffu;
    if (tid < -1 || tid >= n_ref || mtid < -1 || mtid >= n_ref || pos < -1 || mpos < -1) return false;
    if (l_qname < 1 || l_qseq < 0 || l_qseq >= (1 << 28)) return false;
    if (32ull + l_qname + 4ull * n_cigar + (((uint64_t)l_qseq + 1) >> 1) + (uint64_t)l_qseq > bl) return false;
    int64_t qend = off + 36 + l_qname - 1;
    if (qend < seg_end && ld8(u, qend) != 0) return false;
    return true;
}

// string-heap slots (4-byte aligned) a record needs for the projected columns; both kernels use these rules
__device__ __forceinline__ uint32_t slot_bytes(uint32_t bound) { return bound > 12u ? align4(bound) : 0u; }
__device__ __forceinline__ bool qual_absent(const uint8_t *u, const Rec &r) { return r.l_qseq <= 0 || ld8(u, r.qual) == 0xffu; }

// =======================================================================================================
// boundary walk: one thread per BGZF block
// =======================================================================================================
__global__ void __launch_bounds__(128) bam_walk_kernel(const WalkParams p, const ScanTables t) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.n) return;
    if (p.todo && !p.todo[i]) return;
    const uint8_t *u = p.u;
    const int64_t start = p.blk_start[i], bend = p.blk_start[i + 1];
    const int64_t lim = bend < p.limit_end ? bend : p.limit_end;
    int64_t off = p.entry_in[i];
    if (off == INT64_MIN) {
        // speculate: htslib-written BAM starts every block on a record boundary (sam.c:883 bgzf_flush_try)
        off = start;
        if (!(plausible_header(u, off, p.seg_end, t.n_ref))) {
            int64_t s = start + 1;
            for (; s < bend; ++s) {
                if (!plausible_header(u, s, p.seg_end, t.n_ref)) continue;
                int64_t nx = s + 4 + (int64_t)ld32u(u, s);  // double check with the following header when in range
                if (nx + 36 <= p.seg_end && !plausible_header(u, nx, p.seg_end, t.n_ref)) continue;
                break;
            }
            off = s;  // == bend if nothing starts here (a long record spans the block)
        }
    }
    BlockWalk w;
    w.entry = off;
    w.count = 0;
    w.heap_bytes = 0;
    w.stop = 0;
    bool want_aux_strings = ((t.col_mask >> COL_RG) & 1ull) != 0;
    for (uint32_t k = 0; k < t.n_tags; ++k)
        want_aux_strings |= t.tags[k].type == 'Z' && ((t.col_mask >> (BAM_CORE_COLS + k)) & 1ull);
    while (off < lim) {
        Rec r;
        int rc = parse_record(u, off, p.seg_end, p.is_final != 0, t.n_ref, r);
        if (rc != 0) { w.stop = rc; break; }
        uint64_t need = 0;
        if ((t.col_mask >> COL_QNAME) & 1ull) need += slot_bytes(r.l_qname);
        if ((t.col_mask >> COL_CIGAR) & 1ull) need += slot_bytes(r.cigar_text_len);
        if ((t.col_mask >> COL_SEQ) & 1ull) need += slot_bytes((uint32_t)r.l_qseq);
        if (((t.col_mask >> COL_QUAL) & 1ull) && !qual_absent(u, r)) need += slot_bytes((uint32_t)r.l_qseq);
        if (want_aux_strings && r.end - r.aux > 2) {
            // every Z/H value longer than 12 bytes whose tag is a projected string column gets a slot
            int64_t s = r.aux + 2;
            for (;;) {
                if (s - 2 == r.cg_lo) {  // tag consumed by bam_tag2cigar
                    if (r.end - r.cg_hi <= 2) break;
                    s = r.cg_hi + 2;
                    continue;
                }
                bool complete;
                int64_t nx = skip_aux_scalar(u, s, r.end, &complete);
                uint32_t ty = ld8(u, s);
                if (nx != NONE && complete && (ty == 'Z' || ty == 'H')) {
                    uint32_t len = (uint32_t)(nx - 1 - (s + 1));
                    if (len > 12u) {
                        uint32_t t0 = ld8(u, s - 2), t1 = ld8(u, s - 1);
                        bool wanted = (t0 == 'R' && t1 == 'G' && ((t.col_mask >> COL_RG) & 1ull));
                        for (uint32_t k = 0; k < t.n_tags && !wanted; ++k)
                            wanted = t.tags[k].t0 == t0 && t.tags[k].t1 == t1 && t.tags[k].type == 'Z' &&
                                     ((t.col_mask >> (BAM_CORE_COLS + k)) & 1ull);
                        if (wanted) need += align4(len);
                    }
                }
                if (nx == NONE || r.end - nx <= 2) break;
                s = nx + 2;
            }
        }
        w.heap_bytes += need;
        w.count++;
        off = r.end;
    }
    w.exit_off = off;
    p.out[i] = w;
}

// =======================================================================================================
// columnar decode: one warp per BGZF block
// =======================================================================================================
constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ void set_valid(unsigned long long *v, uint32_t row) {
    atomicOr(&v[row >> 6], 1ull << (row & 63u));
}
__device__ __forceinline__ void store_cell(uint8_t *col, uint32_t row, uint32_t len, uint32_t w0, uint32_t w1, uint32_t w2) {
    reinterpret_cast<uint4 *>(col)[row] = make_uint4(len, w0, w1, w2);
}

// Warp-cooperative VARCHAR cell from `len` bytes at u[src..]; `bound` is the slot size rule shared with the walk.
__device__ __forceinline__ void emit_str(const DecodeParams &p, int c, uint32_t row, const uint8_t *u, int64_t src,
                                         uint32_t len, uint32_t bound, uint64_t &hp, uint32_t lane) {
    uint64_t slot = hp;
    hp += slot_bytes(bound);
    if (len <= 12u) {
        if (lane == 0) {
            uint32_t w0 = len > 0 ? mask_bytes(ld32u(u, src), len) : 0u;
            uint32_t w1 = len > 4 ? mask_bytes(ld32u(u, src + 4), len - 4u) : 0u;
            uint32_t w2 = len > 8 ? mask_bytes(ld32u(u, src + 8), len - 8u) : 0u;
            store_cell(p.col[c], row, len, w0, w1, w2);
        }
        return;
    }
    uint32_t *dst = reinterpret_cast<uint32_t *>(p.heap + slot);
    uint32_t words = (len + 3u) >> 2;
    for (uint32_t k = lane; k < words; k += 32u) dst[k] = mask_bytes(ld32u(u, src + 4ll * k), len - 4u * k);
    if (lane == 0) {
        uint64_t hptr = p.heap_host + slot;
        store_cell(p.col[c], row, len, ld32u(u, src), (uint32_t)hptr, (uint32_t)(hptr >> 32));
    }
}

// lane-parallel strlen bounded by [s, end): returns offset of the NUL or `end`
__device__ __forceinline__ int64_t find_nul(const uint8_t *u, int64_t s, int64_t end, uint32_t lane) {
    for (int64_t b = s; b < end; b += 32) {
        int64_t q = b + lane;
        bool z = q < end && ld8(u, q) == 0;
        unsigned m = __ballot_sync(FULL, z);
        if (m) return b + (__ffs(m) - 1);
    }
    return end;
}

__global__ void __launch_bounds__(128) bam_decode_kernel(const __grid_constant__ DecodeParams p, const ScanTables t) {
    __shared__ uint32_t stage_all[4][4];  // per warp: first 16 chars of the CIGAR text
    const uint32_t lane = threadIdx.x & 31u, wib = threadIdx.x >> 5;
    const uint32_t bi = p.b_lo + blockIdx.x * 4u + wib;
    if (bi >= p.b_hi) return;
    uint8_t *stage = reinterpret_cast<uint8_t *>(stage_all[wib]);
    const uint8_t *u = p.u;
    const BlockWalk bw = p.walk[bi];
    int64_t off = bw.entry;
    uint32_t row = p.row_base[bi];
    uint64_t hp = p.heap_base[bi];
    const uint64_t mask = t.col_mask;
    // tag requests owned by this lane: k0 = lane, k1 = lane + 32; request n_tags is the internal RG lookup
    const uint32_t n_req = t.n_tags + 1u;
    uint32_t rq_tag[2], rq_type[2];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        uint32_t k = lane + 32u * j;
        if (k < t.n_tags) { rq_tag[j] = t.tags[k].t0 | (t.tags[k].t1 << 8); rq_type[j] = t.tags[k].type; }
        else if (k == t.n_tags) { rq_tag[j] = 'R' | ('G' << 8); rq_type[j] = 'Z'; }
        else { rq_tag[j] = 0xffffffffu; rq_type[j] = 0; }
    }
    const bool any_tag_cols = (mask >> COL_RG) != 0;  // RG, SAMPLE or any tag column

    for (uint32_t rec_i = 0; rec_i < bw.count; ++rec_i, ++row) {
        Rec r;
        // counts/entries come from the validated walk, so this cannot fail; is_final=false keeps it cheap
        parse_record(u, off, INT64_MAX / 2, false, t.n_ref, r);  // (off may be negative: carried bytes)
        // ---- fixed-width columns (bam_reader.c:790-855): one lane each ----
        if (lane == 0 && ((mask >> COL_FLAG) & 1)) reinterpret_cast<uint16_t *>(p.col[COL_FLAG])[row] = (uint16_t)r.flag;
        if (lane == 1 && ((mask >> COL_POS) & 1)) reinterpret_cast<int64_t *>(p.col[COL_POS])[row] = (int64_t)r.pos + 1;
        if (lane == 2 && ((mask >> COL_MAPQ) & 1)) reinterpret_cast<int32_t *>(p.col[COL_MAPQ])[row] = (int32_t)r.mapq;
        if (lane == 3 && ((mask >> COL_PNEXT) & 1)) reinterpret_cast<int64_t *>(p.col[COL_PNEXT])[row] = (int64_t)r.mpos + 1;
        if (lane == 4 && ((mask >> COL_TLEN) & 1)) reinterpret_cast<int64_t *>(p.col[COL_TLEN])[row] = (int64_t)r.isize;
        if (lane == 5 && ((mask >> COL_RNAME) & 1))
            reinterpret_cast<uint4 *>(p.col[COL_RNAME])[row] = reinterpret_cast<const uint4 *>(t.ref_cells)[r.tid >= 0 ? r.tid : t.n_ref];
        if (lane == 6 && ((mask >> COL_RNEXT) & 1))
            reinterpret_cast<uint4 *>(p.col[COL_RNEXT])[row] = reinterpret_cast<const uint4 *>(t.ref_cells)[r.mtid >= 0 ? r.mtid : t.n_ref];
        // ---- QNAME: C string, at most l_qname bytes (missing-NUL fix-up, sam.c:828-832) ----
        if ((mask >> COL_QNAME) & 1) {
            int64_t z = find_nul(u, r.qname, r.qname + r.l_qname, lane);
            emit_str(p, COL_QNAME, row, u, r.qname, (uint32_t)(z - r.qname), r.l_qname, hp, lane);
        }
        // ---- CIGAR text (cigar_to_kstring, bam_reader.c:375-383; "*" when n_cigar == 0, :819-834) ----
        if ((mask >> COL_CIGAR) & 1) {
            const uint32_t L = r.cigar_text_len;
            if (r.cig_n == 0) {
                if (lane == 0) store_cell(p.col[COL_CIGAR], row, 1, '*', 0, 0);
            } else {
                uint64_t slot = hp;
                hp += slot_bytes(L);
                uint8_t *dst = L > 12u ? p.heap + slot : nullptr;
                if (lane < 4) reinterpret_cast<uint32_t *>(stage)[lane] = 0;
                __syncwarp();
                uint32_t base = 0;
                for (uint32_t k0 = 0; k0 < r.cig_n; k0 += 32u) {
                    uint32_t k = k0 + lane, c = 0, d = 0;
                    if (k < r.cig_n) { c = ld32u(u, r.cigar + 4ll * k); d = dec_digits(c >> 4) + 1u; }
                    uint32_t incl = d;  // inclusive warp scan of d
#pragma unroll
                    for (int s = 1; s < 32; s <<= 1) { uint32_t v = __shfl_up_sync(FULL, incl, s); if (lane >= (uint32_t)s) incl += v; }
                    uint32_t o = base + incl - d;
                    if (k < r.cig_n) {
                        uint32_t v = c >> 4, q = o + d - 1;
                        const char opc = "MIDNSHP=XB??????"[c & 15u];  // BAM_CIGAR_STR, sam.h:112-130
                        if (dst) dst[q] = (uint8_t)opc;
                        if (q < 16u) stage[q] = (uint8_t)opc;
                        for (uint32_t j = d - 1; j-- > 0;) {
                            uint32_t dg = '0' + v % 10u;
                            v /= 10u;
                            if (dst) dst[o + j] = (uint8_t)dg;
                            if (o + j < 16u) stage[o + j] = (uint8_t)dg;
                        }
                    }
                    base += __shfl_sync(FULL, incl, 31);
                }
                __syncwarp();
                if (lane == 0) {
                    const uint32_t *sw = reinterpret_cast<const uint32_t *>(stage);
                    if (L <= 12u) store_cell(p.col[COL_CIGAR], row, L, sw[0], sw[1], sw[2]);
                    else { uint64_t hptr = p.heap_host + slot; store_cell(p.col[COL_CIGAR], row, L, sw[0], (uint32_t)hptr, (uint32_t)(hptr >> 32)); }
                }
                __syncwarp();
            }
        }
        // ---- SEQ: 4-bit nibbles -> ASCII (seq_to_string, bam_reader.c:390-394; "*" when l_qseq == 0) ----
        if ((mask >> COL_SEQ) & 1) {
            const uint32_t L = (uint32_t)r.l_qseq;
            const uint64_t LUT_LO = 0x565352474d43413dull, LUT_HI = 0x4e42444b48595754ull;  // "=ACMGRSV" "TWYHKDBN"
            auto nt = [&](uint32_t c) -> uint32_t { return (uint32_t)(((c & 8u) ? LUT_HI : LUT_LO) >> (8u * (c & 7u))) & 0xffu; };
            auto word = [&](uint32_t k) -> uint32_t {  // chars 4k..4k+3
                uint32_t b0 = ld8(u, r.seq + 2ll * k), b1 = (4u * k + 2u < L) ? ld8(u, r.seq + 2ll * k + 1) : 0u;
                uint32_t v = nt(b0 >> 4) | (nt(b0 & 15u) << 8) | (nt(b1 >> 4) << 16) | (nt(b1 & 15u) << 24);
                return mask_bytes(v, L - 4u * k);
            };
            if (L == 0) { if (lane == 0) store_cell(p.col[COL_SEQ], row, 1, '*', 0, 0); }
            else if (L <= 12u) {
                if (lane == 0) store_cell(p.col[COL_SEQ], row, L, word(0), L > 4 ? word(1) : 0u, L > 8 ? word(2) : 0u);
            } else {
                uint64_t slot = hp;
                hp += align4(L);
                uint32_t *dst = reinterpret_cast<uint32_t *>(p.heap + slot);
                uint32_t words = (L + 3u) >> 2;
                for (uint32_t k = lane; k < words; k += 32u) dst[k] = word(k);
                if (lane == 0) { uint64_t hptr = p.heap_host + slot; store_cell(p.col[COL_SEQ], row, L, word(0), (uint32_t)hptr, (uint32_t)(hptr >> 32)); }
            }
        }
        // ---- QUAL + 33 (qual_to_string, bam_reader.c:400-404; "*" if absent; strlen-truncated) ----
        if ((mask >> COL_QUAL) & 1) {
            if (qual_absent(u, r)) { if (lane == 0) store_cell(p.col[COL_QUAL], row, 1, '*', 0, 0); }
            else {
                const uint32_t L = (uint32_t)r.l_qseq;
                uint64_t slot = hp;
                hp += slot_bytes(L);
                auto word = [&](uint32_t k) -> uint32_t {
                    uint32_t w = mask_bytes(ld32u(u, r.qual + 4ll * k), L - 4u * k);
                    uint32_t s = ((w & 0x7f7f7f7fu) + 0x21212121u) ^ (w & 0x80808080u);  // per-byte +33 mod 256
                    return mask_bytes(s, L - 4u * k);
                };
                uint32_t words = (L + 3u) >> 2, cut = L;  // cut = position of the first produced NUL (q == 223)
                uint32_t *dst = L > 12u ? reinterpret_cast<uint32_t *>(p.heap + slot) : nullptr;
                for (uint32_t k0 = 0; k0 < words; k0 += 32u) {
                    uint32_t k = k0 + lane, w = 0x01010101u, zpos = 0xffffffffu;
                    if (k < words) {
                        w = word(k);
                        if (dst) dst[k] = w;
                        uint32_t nv = L - 4u * k < 4u ? L - 4u * k : 4u;
                        for (uint32_t j = 0; j < nv; ++j)
                            if (((w >> (8u * j)) & 0xffu) == 0u) { zpos = 4u * k + j; break; }
                    }
#pragma unroll
                    for (int s = 16; s > 0; s >>= 1) { uint32_t o = __shfl_xor_sync(FULL, zpos, s); zpos = o < zpos ? o : zpos; }
                    if (zpos < cut) cut = zpos;
                }
                if (lane == 0) {
                    if (cut <= 12u) store_cell(p.col[COL_QUAL], row, cut, cut > 0 ? mask_bytes(word(0), cut) : 0u,
                                               cut > 4 ? mask_bytes(word(1), cut - 4u) : 0u, cut > 8 ? mask_bytes(word(2), cut - 8u) : 0u);
                    else { uint64_t hptr = p.heap_host + slot; store_cell(p.col[COL_QUAL], row, cut, word(0), (uint32_t)hptr, (uint32_t)(hptr >> 32)); }
                }
            }
        }
        // ---- aux tags: READ_GROUP_ID, SAMPLE_ID, typed standard tags ----
        if (any_tag_cols) {
            // hit*: 0 = absent, -1 = decided NULL (matched but invalid), else offset of the type byte relative to the record
            int32_t hit0 = 0, hit1 = 0;
            if (r.end - r.aux > 2) {
                int64_t s = r.aux + 2;
                for (;;) {
                    if (s - 2 == r.cg_lo) {
                        if (r.end - r.cg_hi <= 2) break;
                        s = r.cg_hi + 2;
                        continue;
                    }
                    // skip_aux, with the Z/H scan done lane-parallel
                    uint32_t ty = ld8(u, s);
                    int size = aux_type2size(ty);
                    int64_t nx;
                    bool complete = true;
                    if (size == 'Z' || size == 'H') {
                        int64_t z = find_nul(u, s + 1, r.end, lane);
                        complete = z < r.end;
                        nx = complete ? z + 1 : r.end;
                    } else nx = skip_aux_scalar(u, s, r.end, &complete);
                    uint32_t tg = ld8(u, s - 2) | (ld8(u, s - 1) << 8);
                    const int32_t hv = (nx == NONE || !complete) ? -1 : (int32_t)(s - r.off);
                    if (hit0 == 0 && rq_tag[0] == tg) hit0 = hv;
                    if (hit1 == 0 && rq_tag[1] == tg) hit1 = hv;
                    if (nx == NONE || r.end - nx <= 2) break;
                    s = nx + 2;
                }
            }
            // numeric / char columns: the owning lane converts (bam_aux2i / bam_aux2f / bam_aux2A, sam.c:5093-5141)
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                uint32_t k = lane + 32u * j;
                if (k >= t.n_tags || !((mask >> (BAM_CORE_COLS + k)) & 1ull)) continue;
                int c = BAM_CORE_COLS + (int)k;
                if (rq_type[j] == 'Z') continue;
                const int32_t hj = j ? hit1 : hit0;
                if (hj <= 0) { if (rq_type[j] == 'A') store_cell(p.col[c], row, 0, 0, 0, 0); continue; }
                int64_t a = r.off + hj;
                uint32_t ty = ld8(u, a);
                int64_t iv = 0;
                bool is_int = true;
                switch (ty) {
                case 'c': iv = (int8_t)ld8(u, a + 1); break;
                case 'C': iv = ld8(u, a + 1); break;
                case 's': iv = (int16_t)(ld8(u, a + 1) | (ld8(u, a + 2) << 8)); break;
                case 'S': iv = ld8(u, a + 1) | (ld8(u, a + 2) << 8); break;
                case 'i': iv = (int32_t)ld32u(u, a + 1); break;
                case 'I': iv = ld32u(u, a + 1); break;
                default: is_int = false; break;
                }
                if (rq_type[j] == 'i') reinterpret_cast<int64_t *>(p.col[c])[row] = iv;  // non-integer type -> 0
                else if (rq_type[j] == 'f') {
                    double dv;
                    if (ty == 'd') dv = __longlong_as_double(((long long)ld32u(u, a + 5) << 32) | ld32u(u, a + 1));
                    else if (ty == 'f') dv = (double)__uint_as_float(ld32u(u, a + 1));
                    else dv = is_int ? (double)iv : 0.0;
                    reinterpret_cast<double *>(p.col[c])[row] = dv;
                } else {  // 'A': one character, or "" on a type mismatch (bam_aux2A returns 0)
                    uint32_t ch = ty == 'A' ? ld8(u, a + 1) : 0u;
                    store_cell(p.col[c], row, ch ? 1u : 0u, ch, 0, 0);
                }
                set_valid(p.valid[c], row);
            }
            // string columns, one request at a time (warp-cooperative emit)
            for (uint32_t k = 0; k < n_req; ++k) {
                const bool is_rg = k == t.n_tags;
                uint32_t ktype = __shfl_sync(FULL, (k >> 5) ? rq_type[1] : rq_type[0], k & 31u);
                if (ktype != 'Z') continue;
                int c = is_rg ? COL_RG : BAM_CORE_COLS + (int)k;
                bool want_cell = is_rg ? ((mask >> COL_RG) & 1ull) : ((mask >> c) & 1ull);
                bool want_sample = is_rg && ((mask >> COL_SAMPLE) & 1ull);
                if (!want_cell && !want_sample) continue;
                int32_t hrel = __shfl_sync(FULL, (k >> 5) ? hit1 : hit0, k & 31u);
                int64_t h = r.off + hrel;
                uint32_t ty = hrel > 0 ? ld8(u, h) : 0u;
                bool ok = hrel > 0 && (ty == 'Z' || ty == 'H');   // bam_aux2Z: NULL for any other type
                int64_t z = 0;
                if (ok) z = find_nul(u, h + 1, r.end, lane);
                uint32_t len = ok ? (uint32_t)(z - (h + 1)) : 0u;
                if (want_cell) {
                    if (ok) { emit_str(p, c, row, u, h + 1, len, len, hp, lane); if (lane == 0) set_valid(p.valid[c], row); }
                    else if (lane == 0) store_cell(p.col[c], row, 0, 0, 0, 0);
                }
                if (want_sample) {
                    // @RG ID -> SM dictionary (sam_hdr_find_tag_id, bam_reader.c:892-918): first @RG line with this ID
                    int found = -1;
                    if (ok) {
                        for (uint32_t e0 = 0; e0 < t.n_rg && found < 0; e0 += 32u) {
                            uint32_t e = e0 + lane;
                            bool m = false;
                            if (e < t.n_rg && t.rg[e].id_len == len) {
                                m = true;
                                const uint8_t *id = t.rg_ids + t.rg[e].id_off;
                                for (uint32_t q = 0; q < len; ++q)
                                    if (id[q] != ld8(u, h + 1 + q)) { m = false; break; }
                            }
                            unsigned bm = __ballot_sync(FULL, m);
                            if (bm) found = (int)(e0 + __ffs(bm) - 1);
                        }
                    }
                    if (lane == 0) {
                        if (found >= 0 && t.rg[found].has_sample) {
                            reinterpret_cast<uint4 *>(p.col[COL_SAMPLE])[row] = *reinterpret_cast<const uint4 *>(&t.rg[found].sample);
                            set_valid(p.valid[COL_SAMPLE], row);
                        } else store_cell(p.col[COL_SAMPLE], row, 0, 0, 0, 0);
                    }
                }
            }
        }
        off = r.end;
    }
}

cudaError_t launch_bam_walk(const WalkParams &p, const ScanTables &t, cudaStream_t stream) {
    if (p.n == 0) return cudaSuccess;
    bam_walk_kernel<<<(p.n + 127u) / 128u, 128, 0, stream>>>(p, t);
    return cudaGetLastError();
}
cudaError_t launch_bam_decode(const DecodeParams &p, const ScanTables &t, cudaStream_t stream) {
    if (p.b_hi <= p.b_lo) return cudaSuccess;
    bam_decode_kernel<<<(p.b_hi - p.b_lo + 3u) / 4u, 128, 0, stream>>>(p, t);
    return cudaGetLastError();
}

}  // namespace dh
