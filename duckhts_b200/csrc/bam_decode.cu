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
#include <stdlib.h>

#include <mutex>

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
    int64_t rlen;            // bam_cigar2rlen: reference bases the CIGAR consumes (0 if cig_n == 0)
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

// scalar bam_aux_get (sam.c:4832-4855) over aux bytes [a0, end): offset of the type byte, NONE (no such tag: errno ENOENT) or
// CORRUPT (errno EINVAL: a broken entry in front of the tag, or the tag's own value incomplete; bam_aux_next :4823, bam_aux_get :4841-4842)
constexpr int64_t CORRUPT = INT64_MIN + 1;
__device__ __forceinline__ int64_t aux_get_scalar(const uint8_t *u, int64_t a0, int64_t end, uint32_t t0, uint32_t t1) {
    if (end - a0 <= 2) return NONE;
    int64_t s = a0 + 2;
    for (;;) {
        bool complete;
        int64_t nx = skip_aux_scalar(u, s, end, &complete);
        if (ld8(u, s - 2) == t0 && ld8(u, s - 1) == t1) {
            if (nx == NONE) return CORRUPT;
            uint32_t ty = ld8(u, s);
            if ((ty == 'Z' || ty == 'H') && !complete) return CORRUPT;
            return s;
        }
        if (nx == NONE) return CORRUPT;
        if (end - nx <= 2) return NONE;
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
        if (CG == CORRUPT) return -4;       // bam_tag2cigar: bad aux data (errno != ENOENT, sam.c:690-694) -> bam_read1 returns -4 (:837-838)
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
    r.rlen = 0;
    if (r.cig_n) {                                           // :841-852
        int64_t qlen = 0, rlen = 0;
        uint32_t tl = 0;
        for (uint32_t i = 0; i < r.cig_n; ++i) {
            uint32_t c = ld32u(u, r.cigar + 4ll * i);
            if ((0x3C1A7u >> ((c & 15u) << 1)) & 1u) qlen += c >> 4;  // bam_cigar_type bit 0: consumes query
            if ((0x3C1A7u >> ((c & 15u) << 1)) & 2u) rlen += c >> 4;  // bit 1: consumes reference (bam_cigar2rlen)
            tl += dec_digits(c >> 4) + 1u;
        }
        r.cigar_text_len = tl;
        r.rlen = rlen;
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

// ---- AUXILIARY_TAGS (A16) ---------------------------------------------------------------------------------------------
// Upper bound of the text bam_aux_to_string (bam_reader.c:140-183) writes for the value whose type byte is at s; exact for
// A / Z / H.  Bounds above 12 reserve a heap slot (slot_bytes); both kernels use this rule.
__device__ __forceinline__ uint32_t aux_text_bound(const uint8_t *u, int64_t s, int64_t nx, int64_t end) {
    switch (ld8(u, s)) {
    case 'A': return 1u;
    case 'c': case 'C': case 's': case 'S': case 'i': case 'I': return 11u;   // "-2147483648"
    case 'f': case 'd': return 13u;                                           // "-1.79769e+308"
    case 'Z': case 'H': return (uint32_t)((nx == end && ld8(u, end - 1) != 0 ? end : nx - 1) - (s + 1));
    case 'B': {
        const uint32_t sub = ld8(u, s + 1), n = ld32u(u, s + 2);
        const uint32_t per = sub == 'c' ? 5u : sub == 'C' ? 4u : sub == 's' ? 7u : sub == 'S' ? 6u : sub == 'i' ? 12u : sub == 'I' ? 11u : sub == 'f' ? 14u : 2u;
        return 1u + per * n;
    }
    default: return 0u;
    }
}
__device__ __forceinline__ bool aux_excluded(const ScanTables &t, uint32_t key) { return t.aux_excl && ((t.aux_excl[key >> 5] >> (key & 31u)) & 1u); }
// pairs and heap bytes of one record's MAP cell: the bam_aux_first / bam_aux_next walk of bam_reader.c:969-980.  An entry
// whose value does not fit the record (skip_aux fails) is still returned by bam_aux_next and counted; the walk ends after it.
__device__ __forceinline__ void aux_map_need(const uint8_t *u, const Rec &r, const ScanTables &t, uint32_t &cnt, uint32_t &heap) {
    cnt = heap = 0;
    if (r.end - r.aux <= 2) return;
    int64_t s = r.aux + 2;
    for (;;) {
        if (s - 2 == r.cg_lo) {   // tag consumed by bam_tag2cigar
            if (r.end - r.cg_hi <= 2) return;
            s = r.cg_hi + 2;
            continue;
        }
        bool complete;
        const int64_t nx = skip_aux_scalar(u, s, r.end, &complete);
        if (!aux_excluded(t, ld8(u, s - 2) | (ld8(u, s - 1) << 8))) {
            ++cnt;
            if (nx != NONE) heap += slot_bytes(aux_text_bound(u, s, nx, r.end));
        }
        if (nx == NONE || r.end - nx <= 2) return;
        s = nx + 2;
    }
}

// Region filter of indexed scans (A8).  htslib reads the records of the index's chunks and keeps those with
// `end > beg_q && end_q > beg` on the queried contig (hts_itr_next, htslib/hts.c:4301-4311; multi-region form
// hts_itr_multi_next :4565-4597), end = bam_endpos (sam.c:519-524: pos + rlen, rlen = 1 for unmapped reads or a CIGAR
// that consumes no reference).  Here the host turns the chunk list into one virtual-offset range and the overlap
// test runs in the walk, so rejected records never cost a row, a heap byte or a PCIe byte.
__device__ __forceinline__ bool region_pass(const WalkParams &p, const Rec &r) {
    if (r.tid < 0) return false;
    int64_t rl = (r.flag & 4u) ? 0 : r.rlen;
    if (rl == 0) rl = 1;
    const int64_t beg = r.pos, end = (int64_t)r.pos + rl;
    if (p.filt_mode == 1) return end > 0;          // one whole-contig query per contig: [0, HTS_POS_MAX)
    uint32_t lo = 0, hi = p.filt_n;                // first interval (sorted by tid, beg; disjoint) with (tid, end) > (r.tid, beg)
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        const RegIntv v = p.filt[mid];
        if (v.tid < r.tid || (v.tid == r.tid && v.end <= beg)) lo = mid + 1; else hi = mid;
    }
    if (lo >= p.filt_n) return false;
    const RegIntv v = p.filt[lo];
    return v.tid == r.tid && end > v.beg;
}

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
    uint16_t *rec = p.rec + ((start - p.rec_bias) / 36 + (int64_t)i);
    BlockWalk w;
    w.entry = off;
    w.count = 0;
    w.map_entries = w.map_heap = 0;
    w.heap_bytes = 0;
    w.stop = 0;
    bool want_aux_strings = ((t.col_mask >> COL_RG) & 1ull) != 0;
    for (uint32_t k = 0; k < t.n_tags; ++k)
        want_aux_strings |= (t.tags[k].type == 'Z' || t.tags[k].type == 'B') && t.tags[k].want;
    while (off < lim) {
        Rec r;
        int rc = parse_record(u, off, p.seg_end, p.is_final != 0, t.n_ref, r);
        if (rc != 0) { w.stop = rc; break; }
        if (p.filt_mode && !region_pass(p, r)) { off = r.end; continue; }   // hts_itr_next / hts_itr_multi_next overlap test
        // N3 predicate pushdown (dh_bam_set_filter): the conjunction of flag tests and the MAPQ floor
        if ((r.flag & t.flag_require) != t.flag_require || (r.flag & t.flag_exclude) != 0u || r.mapq < t.min_mapq) { off = r.end; continue; }
        uint64_t need = 0;
        const bool compact = t.compact != 0u;
        if (compact) {   // every string has a slot (bytes padded to 4), "*" placeholders included: compact_slots() below
            const uint32_t L = (uint32_t)r.l_qseq;
            if ((t.col_mask >> COL_QNAME) & 1ull) need += align4(r.l_qname);
            if ((t.col_mask >> COL_CIGAR) & 1ull) need += align4(r.cig_n ? r.cigar_text_len : 1u);
            if ((t.col_mask >> COL_SEQ) & 1ull) need += align4(L ? L : 1u);
            if ((t.col_mask >> COL_QUAL) & 1ull) need += align4(qual_absent(u, r) ? 1u : L);
        } else {
            if ((t.col_mask >> COL_QNAME) & 1ull) need += slot_bytes(r.l_qname);
            if ((t.col_mask >> COL_CIGAR) & 1ull) need += slot_bytes(r.cigar_text_len);
            if ((t.col_mask >> COL_SEQ) & 1ull) need += slot_bytes((uint32_t)r.l_qseq);
            if (((t.col_mask >> COL_QUAL) & 1ull) && !qual_absent(u, r)) need += slot_bytes((uint32_t)r.l_qseq);
        }
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
                    if (len > (compact ? 0u : 12u)) {
                        uint32_t t0 = ld8(u, s - 2), t1 = ld8(u, s - 1);
                        // one slot per destination column: READ_GROUP_ID and an RG tag column each get their own copy
                        uint32_t wanted = (t0 == 'R' && t1 == 'G' && ((t.col_mask >> COL_RG) & 1ull)) ? 1u : 0u;
                        for (uint32_t k = 0; k < t.n_tags; ++k)
                            wanted += (t.tags[k].t0 == t0 && t.tags[k].t1 == t1 && t.tags[k].type == 'Z' && t.tags[k].want) ? 1u : 0u;
                        need += (uint64_t)wanted * align4(len);
                    }
                }
                if (nx != NONE && ty == 'B') {   // children of a projected B-array LIST column: 8 bytes each (+4 to reach 8-byte alignment)
                    const uint32_t t0 = ld8(u, s - 2), t1 = ld8(u, s - 1);
                    for (uint32_t k = 0; k < t.n_tags; ++k)
                        if (t.tags[k].t0 == t0 && t.tags[k].t1 == t1 && t.tags[k].type == 'B' && t.tags[k].want) {
                            need += 4ull + 8ull * ld32u(u, s + 2);
                            break;
                        }
                }
                if (nx == NONE || r.end - nx <= 2) break;
                s = nx + 2;
            }
        }
        if (t.aux_map) {
            uint32_t mc, mh;
            aux_map_need(u, r, t, mc, mh);
            w.map_entries += mc;
            w.map_heap += mh;
            need += mh;
        }
        w.heap_bytes += need;
        rec[w.count] = (uint16_t)(off - start);   // where the decode kernel finds this record (relative to the block start)
        w.count++;
        off = r.end;
    }
    w.exit_off = off;
    p.out[i] = w;
}

// ---- N3: CIGAR metrics as derived columns (tag request type 'c') ---------------------------------------------------------
// What the reference's scalar UDFs compute from read_bam's CIGAR text (src/kmer_udf.c:197-269 parse_cigar_metrics, :271-293
// cigar_has_operator_text, registered at :1231-1238), here from the binary CIGAR next to the decode that holds it in registers:
//   t0 = 'S' cigar_has_soft_clip  'H' cigar_has_hard_clip (0 / 1)   'l' cigar_left_soft_clip   'r' cigar_right_soft_clip
//        'q' cigar_query_length   'a' cigar_aligned_query_length    'R' cigar_reference_length  'o' cigar_has_op(CIGAR, t1) (0 / 1)
// NULL (false returned) where the UDF returns NULL: "*" (no CIGAR), an operation of length 0 ("0M": op_len <= 0), an operation the
// text parser does not know ('B', '?'); cigar_has_op answers false for "*" and stops at its operator before it sees later flaws.
__device__ __forceinline__ bool cigar_metric(const uint8_t *u, const Rec &r, uint32_t metric, uint32_t opch, long long &out) {
    out = 0;
    if (r.cig_n == 0u) return metric == 'o';
    long long ql = 0, al = 0, rl = 0, first_len = 0, last_len = 0;
    uint32_t first_op = 15u, last_op = 15u;
    bool soft = false, hard = false;
    for (uint32_t k = 0; k < r.cig_n; ++k) {
        const uint32_t c = ld32u(u, r.cigar + 4ll * k), ln = c >> 4, op = c & 15u;
        if (ln == 0u) return false;
        if (metric == 'o') {
            if ((uint32_t)"MIDNSHP=XB??????"[op] == opch) { out = 1; return true; }
            continue;
        }
        switch (op) {
        case 0: case 7: case 8: ql += ln; al += ln; rl += ln; break;   // M = X
        case 1: ql += ln; break;                                        // I
        case 4: ql += ln; soft = true; break;                           // S
        case 5: hard = true; break;                                     // H
        case 2: case 3: rl += ln; break;                                // D N
        case 6: break;                                                  // P
        default: return false;
        }
        if (k == 0u) { first_op = op; first_len = ln; }
        last_op = op;
        last_len = ln;
    }
    switch (metric) {
    case 'S': out = soft; break;
    case 'H': out = hard; break;
    case 'l': out = first_op == 4u ? first_len : 0; break;
    case 'r': out = last_op == 4u ? last_len : 0; break;
    case 'q': out = ql; break;
    case 'a': out = al; break;
    case 'R': out = rl; break;
    default: break;   // 'o': operator not found
    }
    return true;
}

// =======================================================================================================
// columnar decode: one warp per BGZF block, 32 records per step
// =======================================================================================================
// Step layout (ncu on the first version, which walked ONE record per warp step, showed ~6.5k cycles per record):
//   A  every lane parses ITS OWN record (offsets come from the walk: rec[] is read coalesced) and stores the
//      fixed-width columns -> 32 consecutive rows per store instruction;
//   B  a warp scan of the per-record heap needs gives every record its string slots;
//   C  the long strings (QNAME, SEQ nibbles -> ASCII, QUAL+33, long CIGAR text) are produced record by record
//      by the whole warp: 8 output bytes per lane, contiguous stores into the heap;
//   D  every lane builds the 16-byte duckdb_string_t cells of its own row and the typed tag values ->
//      512 contiguous bytes per column per store instruction; validity bits go out as one word per 32 rows.
constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ void store_cell(uint8_t *col, uint32_t row, uint32_t len, uint32_t w0, uint32_t w1, uint32_t w2) {
    reinterpret_cast<uint4 *>(col)[row] = make_uint4(len, w0, w1, w2);
}
// validity bits of the 32 rows [row0, row0 + 32): two 32-bit atomics per warp instead of one per row
__device__ __forceinline__ void set_valid_group(unsigned long long *v, uint32_t row0, uint32_t mask, uint32_t lane) {
    if (lane == 0 && mask) {
        unsigned int *v32 = reinterpret_cast<unsigned int *>(v);
        const uint32_t sh = row0 & 31u;
        atomicOr(&v32[row0 >> 5], mask << sh);
        if (sh && (mask >> (32u - sh))) atomicOr(&v32[(row0 >> 5) + 1u], mask >> (32u - sh));
    }
}
// <= 12 bytes at u[src..src+len) as the three payload words of an inlined cell
__device__ __forceinline__ void inline_words(const uint8_t *u, int64_t src, uint32_t len, uint32_t &w0, uint32_t &w1, uint32_t &w2) {
    w0 = len > 0 ? mask_bytes(ld32u(u, src), len) : 0u;
    w1 = len > 4 ? mask_bytes(ld32u(u, src + 4), len - 4u) : 0u;
    w2 = len > 8 ? mask_bytes(ld32u(u, src + 8), len - 8u) : 0u;
}
__device__ __forceinline__ void store_str_cell(uint8_t *col, uint32_t row, const uint8_t *u, int64_t src, uint32_t len, uint64_t hptr) {
    if (len <= 12u) {
        uint32_t w0, w1, w2;
        inline_words(u, src, len, w0, w1, w2);
        store_cell(col, row, len, w0, w1, w2);
    } else store_cell(col, row, len, ld32u(u, src), (uint32_t)hptr, (uint32_t)(hptr >> 32));
}
// two packed bytes (b0 | b1 << 8) of 4-bit bases, first base in the high nibble (bam_seqi, htslib/sam.h:325) ->
// four characters of "=ACMGRSVTWYHKDBN" (seq_nt16_str, htslib/hts.c:260) by byte permutes
__device__ __forceinline__ uint32_t seq4(uint32_t h) {
    const uint32_t s = ((h & 0x0f0fu) << 4) | ((h >> 4) & 0x0f0fu);   // nibbles in base order
    const uint32_t s7 = s & 0x7777u;
    const uint32_t a = __byte_perm(0x4d43413du, 0x56535247u, s7);     // "=ACM" "GRSV"
    const uint32_t b = __byte_perm(0x48595754u, 0x4e42444bu, s7);     // "TWYH" "KDBN"
    return __byte_perm(a, b, 0x3210u | ((s & 0x8888u) >> 1));
}
// per-byte +33 mod 256 (qual_to_string, bam_reader.c:400-404)
__device__ __forceinline__ uint32_t qual4(uint32_t w) { return ((w & 0x7f7f7f7fu) + 0x21212121u) ^ (w & 0x80808080u); }
// index of the first zero byte of w (0..3), 4 if none
__device__ __forceinline__ uint32_t first_zero_byte(uint32_t w) {
    const uint32_t z = (w - 0x01010101u) & ~w & 0x80808080u;
    return z ? (uint32_t)(__ffs(z) - 1) >> 3 : 4u;
}
// bam_aux_get (sam.c:4832-4855) for one lane's record, skipping a CG tag consumed by bam_tag2cigar:
// offset of the type byte of the first entry with this tag, or NONE (absent, or present but malformed)
__device__ __forceinline__ int64_t aux_find(const uint8_t *u, const Rec &r, uint32_t tg) {
    if (r.end - r.aux <= 2) return NONE;
    int64_t s = r.aux + 2;
    for (;;) {
        if (s - 2 == r.cg_lo) {
            if (r.end - r.cg_hi <= 2) return NONE;
            s = r.cg_hi + 2;
            continue;
        }
        bool complete;
        const int64_t nx = skip_aux_scalar(u, s, r.end, &complete);
        if ((ld8(u, s - 2) | (ld8(u, s - 1) << 8)) == tg) return (nx == NONE || !complete) ? NONE : s;
        if (nx == NONE || r.end - nx <= 2) return NONE;
        s = nx + 2;
    }
}
// bam_aux2Z view of a hit: length of the NUL-terminated value, or -1 if the entry is not Z/H (-> NULL)
__device__ __forceinline__ int32_t aux_zlen(const uint8_t *u, const Rec &r, int64_t h) {
    if (h == NONE) return -1;
    const uint32_t ty = ld8(u, h);
    if (ty != 'Z' && ty != 'H') return -1;
    int64_t z = h + 1;
    while (z < r.end && ld8(u, z) != 0) ++z;
    return (int32_t)(z - (h + 1));
}
// per-lane copy of a long aux string into its heap slot (rare: Z values longer than 12 bytes)
__device__ __forceinline__ void copy_bytes(uint8_t *dst, const uint8_t *u, int64_t src, uint32_t len) {
    uint32_t *d = reinterpret_cast<uint32_t *>(dst);
    for (uint32_t k = 0; 4u * k < len; ++k) d[k] = mask_bytes(ld32u(u, src + 4ll * k), len - 4u * k);
}
// <= 12 characters of CIGAR text in three words (cigar_to_kstring, bam_reader.c:375-383)
__device__ __forceinline__ void cigar_inline(const uint8_t *u, const Rec &r, uint32_t &w0, uint32_t &w1, uint32_t &w2) {
    w0 = w1 = w2 = 0u;
    uint32_t pos = 0;
    for (uint32_t i = 0; i < r.cig_n; ++i) {
        const uint32_t c = ld32u(u, r.cigar + 4ll * i);
        uint32_t v = c >> 4;
        const uint32_t nd = dec_digits(v);
        for (uint32_t j = nd + 1u; j-- > 0;) {                       // op char last, digits right to left
            uint32_t ch;
            if (j == nd) ch = (uint32_t)"MIDNSHP=XB??????"[c & 15u];   // BAM_CIGAR_STR, sam.h:112-130
            else { ch = '0' + v % 10u; v /= 10u; }
            const uint32_t q = pos + j, sh = 8u * (q & 3u);
            w0 |= (q >> 2) == 0u ? ch << sh : 0u;
            w1 |= (q >> 2) == 1u ? ch << sh : 0u;
            w2 |= (q >> 2) == 2u ? ch << sh : 0u;
        }
        pos += nd + 1u;
    }
}

__global__ void __launch_bounds__(128) bam_decode_kernel(const __grid_constant__ DecodeParams p, const ScanTables t) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t bi = p.b_lo + blockIdx.x * 4u + (threadIdx.x >> 5);
    if (bi >= p.b_hi) return;
    const uint32_t count = p.walk[bi].count;
    if (!count) return;
    const uint8_t *u = p.u;
    const int64_t bstart = p.blk_start[bi];
    const uint16_t *rec = p.rec + ((bstart - p.rec_bias) / 36 + (int64_t)bi);
    const uint64_t mask = t.col_mask;
    const bool cQ = (mask >> COL_QNAME) & 1, cC = (mask >> COL_CIGAR) & 1, cS = (mask >> COL_SEQ) & 1, cU = (mask >> COL_QUAL) & 1,
               cRG = (mask >> COL_RG) & 1, cSM = (mask >> COL_SAMPLE) & 1;
    const bool any_tag_cols = t.any_tag_cols != 0;
    uint32_t row0 = p.row_base[bi];
    uint64_t hp = p.heap_base[bi];

    for (uint32_t g = 0; g < count; g += 32u, row0 += 32u) {
        const uint32_t cnt = count - g < 32u ? count - g : 32u;
        const bool act = lane < cnt;
        const uint32_t row = row0 + lane;
        // ---------------- A: own record ----------------
        Rec r;
        r.off = r.end = r.qname = r.cigar = r.seq = r.qual = r.aux = 0;
        r.cg_lo = r.cg_hi = NONE;
        r.block_len = r.n_cigar = r.cig_n = r.flag = r.mapq = r.cigar_text_len = 0u;
        r.l_qname = 1u;
        r.tid = r.pos = r.mtid = r.mpos = r.isize = -1;
        r.l_qseq = 0;
        uint32_t roff = 0;
        if (act) {
            roff = rec[g + lane];
            // entries and counts come from the validated walk, so this cannot fail; is_final = false keeps it cheap
            parse_record(u, bstart + roff, INT64_MAX / 2, false, t.n_ref, r);
            if ((mask >> COL_FLAG) & 1) reinterpret_cast<uint16_t *>(p.col[COL_FLAG])[row] = (uint16_t)r.flag;
            if ((mask >> COL_POS) & 1) reinterpret_cast<int64_t *>(p.col[COL_POS])[row] = (int64_t)r.pos + 1;
            if ((mask >> COL_MAPQ) & 1) reinterpret_cast<int32_t *>(p.col[COL_MAPQ])[row] = (int32_t)r.mapq;
            if ((mask >> COL_PNEXT) & 1) reinterpret_cast<int64_t *>(p.col[COL_PNEXT])[row] = (int64_t)r.mpos + 1;
            if ((mask >> COL_TLEN) & 1) reinterpret_cast<int64_t *>(p.col[COL_TLEN])[row] = (int64_t)r.isize;
            if ((mask >> COL_RNAME) & 1)
                reinterpret_cast<uint4 *>(p.col[COL_RNAME])[row] = reinterpret_cast<const uint4 *>(t.ref_cells)[r.tid >= 0 ? r.tid : t.n_ref];
            if ((mask >> COL_RNEXT) & 1)
                reinterpret_cast<uint4 *>(p.col[COL_RNEXT])[row] = reinterpret_cast<const uint4 *>(t.ref_cells)[r.mtid >= 0 ? r.mtid : t.n_ref];
        }
        const uint32_t L = (uint32_t)r.l_qseq;
        const bool absent = !act || qual_absent(u, r);
        // QNAME is a C string of at most l_qname bytes (missing-NUL fix-up, sam.c:828-832)
        uint32_t qn = r.l_qname;
        if (act && cQ) {
            for (uint32_t k = 0; k < r.l_qname; k += 4u) {
                const uint32_t z = first_zero_byte(ld32u(u, r.qname + k));
                if (z < 4u) { if (k + z < qn) qn = k + z; break; }
            }
        }
        // ---------------- B: heap slots (the same rules as the walk; 4-byte aligned, in this order) ----------------
        uint32_t need = 0;
        if (act) {
            if (cQ) need += slot_bytes(r.l_qname);
            if (cC) need += slot_bytes(r.cigar_text_len);
            if (cS) need += slot_bytes(L);
            if (cU && !absent) need += slot_bytes(L);
        }
        const uint32_t aux_slot0 = need;   // long Z values follow, in request order (tags, then RG)
        int64_t h_rg = NONE;
        int32_t z_rg = -1;
        if (act && (cRG || cSM)) {
            h_rg = aux_find(u, r, 'R' | ('G' << 8));
            z_rg = aux_zlen(u, r, h_rg);
        }
        if (act && any_tag_cols)
            for (uint32_t k = 0; k < t.n_tags; ++k) {
                if ((t.tags[k].type != 'Z' && t.tags[k].type != 'B') || !t.tags[k].want) continue;
                const int64_t a = aux_find(u, r, t.tags[k].t0 | (t.tags[k].t1 << 8));
                if (t.tags[k].type == 'B') {
                    if (a != NONE && ld8(u, a) == 'B') need += 4u + 8u * ld32u(u, a + 2);
                    continue;
                }
                const int32_t zl = aux_zlen(u, r, a);
                if (zl > 12) need += align4((uint32_t)zl);
            }
        if (act && cRG && z_rg > 12) need += align4((uint32_t)z_rg);
        uint32_t incl = need;
#pragma unroll
        for (int sft = 1; sft < 32; sft <<= 1) { const uint32_t v = __shfl_up_sync(FULL, incl, sft); if (lane >= (uint32_t)sft) incl += v; }
        const uint32_t rel = incl - need;                       // this record's first slot, relative to hp
        const uint32_t total = __shfl_sync(FULL, incl, 31);
        // ---------------- C: long strings, record by record, whole warp ----------------
        uint32_t qcut = L;                                      // QUAL is strlen-truncated: first produced NUL
        const uint32_t lqnc = r.l_qname | (r.n_cigar << 8), lab = L | (absent ? 0x80000000u : 0u);
        for (uint32_t q = 0; q < cnt; ++q) {
            const uint32_t o_ = __shfl_sync(FULL, roff, q), lq_ = __shfl_sync(FULL, lqnc, q), la_ = __shfl_sync(FULL, lab, q),
                           ct_ = __shfl_sync(FULL, r.cigar_text_len, q), rel_ = __shfl_sync(FULL, rel, q);
            const uint32_t lqn = lq_ & 0xffu, ncg = lq_ >> 8, Lq = la_ & 0x7fffffffu;
            const int64_t qname = bstart + o_ + 36, seq = qname + lqn + 4ll * ncg, qual = seq + (((int64_t)Lq + 1) >> 1);
            uint8_t *slot = p.heap + hp + rel_;
            if (cQ && lqn > 12u) {
                uint32_t *d = reinterpret_cast<uint32_t *>(slot);
                for (uint32_t k = lane; 4u * k < lqn; k += 32u) d[k] = mask_bytes(ld32u(u, qname + 4ll * k), lqn - 4u * k);
                slot += align4(lqn);
            }
            if (cC && ct_ > 12u) {
                // long CIGAR text: lanes take operations, a warp scan of the text widths places them
                const int64_t cig = bstart + __shfl_sync(FULL, (int32_t)(r.cigar - bstart), q);
                const uint32_t cn = __shfl_sync(FULL, r.cig_n, q);
                uint32_t base = 0;
                for (uint32_t k0 = 0; k0 < cn; k0 += 32u) {
                    const uint32_t k = k0 + lane;
                    uint32_t c = 0, dgt = 0;
                    if (k < cn) { c = ld32u(u, cig + 4ll * k); dgt = dec_digits(c >> 4) + 1u; }
                    uint32_t inc = dgt;
#pragma unroll
                    for (int sft = 1; sft < 32; sft <<= 1) { const uint32_t v = __shfl_up_sync(FULL, inc, sft); if (lane >= (uint32_t)sft) inc += v; }
                    if (k < cn) {
                        uint32_t v = c >> 4, at = base + inc - 1u;
                        slot[at] = (uint8_t)"MIDNSHP=XB??????"[c & 15u];
                        for (uint32_t j = dgt - 1u; j-- > 0;) { slot[--at] = (uint8_t)('0' + v % 10u); v /= 10u; }
                    }
                    base += __shfl_sync(FULL, inc, 31);
                }
                slot += align4(ct_);
            }
            if (cS && Lq > 12u) {
                uint32_t *d = reinterpret_cast<uint32_t *>(slot);
                for (uint32_t k = lane; 8u * k < Lq; k += 32u) {             // 4 packed bytes -> 8 characters per lane
                    const uint32_t pk = ld32u(u, seq + 4ll * k);
                    d[2u * k] = mask_bytes(seq4(pk & 0xffffu), Lq - 8u * k);
                    if (8u * k + 4u < Lq) d[2u * k + 1u] = mask_bytes(seq4(pk >> 16), Lq - 8u * k - 4u);
                }
                slot += align4(Lq);
            }
            if (cU && !(la_ >> 31) && Lq > 12u) {
                uint32_t *d = reinterpret_cast<uint32_t *>(slot);
                uint32_t zpos = 0xffffffffu;
                for (uint32_t k = lane; 8u * k < Lq; k += 32u) {
                    const uint32_t a = mask_bytes(qual4(ld32u(u, qual + 8ll * k)), Lq - 8u * k) | (Lq - 8u * k < 4u ? 0xffffffffu << (8u * (Lq - 8u * k)) : 0u);
                    d[2u * k] = mask_bytes(a, Lq - 8u * k);
                    uint32_t z = first_zero_byte(a);
                    if (z < 4u && zpos == 0xffffffffu) zpos = 8u * k + z;
                    if (8u * k + 4u < Lq) {
                        const uint32_t rem = Lq - 8u * k - 4u;
                        const uint32_t b = mask_bytes(qual4(ld32u(u, qual + 8ll * k + 4)), rem) | (rem < 4u ? 0xffffffffu << (8u * rem) : 0u);
                        d[2u * k + 1u] = mask_bytes(b, rem);
                        z = first_zero_byte(b);
                        if (z < 4u && zpos == 0xffffffffu) zpos = 8u * k + 4u + z;
                    }
                }
                zpos = __reduce_min_sync(FULL, zpos);
                if (lane == q && zpos < qcut) qcut = zpos;
            }
        }
        __syncwarp();   // the heap of this step is complete (cells below read CIGAR prefixes back)
        // ---------------- D: cells and tag values of the lane's own row ----------------
        const uint64_t hbase = p.heap_host + hp + rel;          // host address of this record's first slot
        if (act) {
            uint64_t hs = hbase;
            if (cQ) { store_str_cell(p.col[COL_QNAME], row, u, r.qname, qn, hs); hs += slot_bytes(r.l_qname); }
            if (cC) {
                // cigar_to_kstring (bam_reader.c:375-383); "*" when n_cigar == 0 (:819-834)
                const uint32_t ctl = r.cigar_text_len;
                if (r.cig_n == 0u) store_cell(p.col[COL_CIGAR], row, 1, '*', 0, 0);
                else if (ctl <= 12u) { uint32_t w0, w1, w2; cigar_inline(u, r, w0, w1, w2); store_cell(p.col[COL_CIGAR], row, ctl, w0, w1, w2); }
                else store_cell(p.col[COL_CIGAR], row, ctl, *reinterpret_cast<const uint32_t *>(p.heap + (hs - p.heap_host)), (uint32_t)hs, (uint32_t)(hs >> 32));
                hs += slot_bytes(ctl);
            }
            if (cS) {
                // seq_to_string (bam_reader.c:390-394); "*" when l_qseq == 0 (:857-866)
                if (L == 0u) store_cell(p.col[COL_SEQ], row, 1, '*', 0, 0);
                else if (L <= 12u) {
                    const uint32_t p0 = ld32u(u, r.seq), p1 = L > 8u ? ld32u(u, r.seq + 4) : 0u;
                    store_cell(p.col[COL_SEQ], row, L, mask_bytes(seq4(p0 & 0xffffu), L), L > 4u ? mask_bytes(seq4(p0 >> 16), L - 4u) : 0u,
                               L > 8u ? mask_bytes(seq4(p1 & 0xffffu), L - 8u) : 0u);
                } else store_cell(p.col[COL_SEQ], row, L, seq4(ld32u(u, r.seq) & 0xffffu), (uint32_t)hs, (uint32_t)(hs >> 32));
                hs += slot_bytes(L);
            }
            if (cU) {
                // qual_to_string (bam_reader.c:400-404); "*" if absent (:868-877); strlen semantics
                if (absent) store_cell(p.col[COL_QUAL], row, 1, '*', 0, 0);
                else {
                    uint32_t w0 = mask_bytes(qual4(ld32u(u, r.qual)), L), w1 = L > 4u ? mask_bytes(qual4(ld32u(u, r.qual + 4)), L - 4u) : 0u,
                             w2 = L > 8u ? mask_bytes(qual4(ld32u(u, r.qual + 8)), L - 8u) : 0u;
                    uint32_t cut = qcut;
                    if (L <= 12u) {                               // short reads never went through step C
                        uint32_t z = first_zero_byte(w0 | (L < 4u ? 0xffffffffu << (8u * L) : 0u));
                        if (z < 4u) cut = z;
                        else if (L > 4u && (z = first_zero_byte(w1 | (L < 8u ? 0xffffffffu << (8u * (L - 4u)) : 0u))) < 4u) cut = 4u + z;
                        else if (L > 8u && (z = first_zero_byte(w2 | (L < 12u ? 0xffffffffu << (8u * (L - 8u)) : 0u))) < 4u) cut = 8u + z;
                    }
                    if (cut <= 12u) store_cell(p.col[COL_QUAL], row, cut, mask_bytes(w0, cut), cut > 4u ? mask_bytes(w1, cut - 4u) : 0u,
                                               cut > 8u ? mask_bytes(w2, cut - 8u) : 0u);
                    else store_cell(p.col[COL_QUAL], row, cut, w0, (uint32_t)hs, (uint32_t)(hs >> 32));
                    hs += slot_bytes(L);
                }
            }
        }
        // typed standard tags (bam_reader.c:920-966): 'i' bam_aux2i, 'f' bam_aux2f, 'A' bam_aux2A, 'Z' bam_aux2Z
        uint64_t hz = hbase + aux_slot0;                        // long Z values, in request order
        if (any_tag_cols)
            for (uint32_t k = 0; k < t.n_tags; ++k) {
                const int c = BAM_CORE_COLS + (int)k;
                if (!t.tags[k].want) continue;
                const uint32_t ty_req = t.tags[k].type;
                bool valid = false;
                if (act && ty_req == 'c') {
                    long long mv;
                    valid = cigar_metric(u, r, t.tags[k].t0, t.tags[k].t1, mv);
                    reinterpret_cast<int64_t *>(p.col[c])[row] = mv;
                } else if (act) {
                    const int64_t a = aux_find(u, r, t.tags[k].t0 | (t.tags[k].t1 << 8));
                    if (ty_req == 'B') {
                        // bam_assign_list_int / _double (bam_reader.c:106-138): the writer follows the value's RUNTIME subtype
                        // (f, d -> bam_auxB2f, else bam_auxB2i), 8 bytes per element whatever the column's child type
                        uint64_t off8 = 0, len = 0;
                        if (a != NONE && ld8(u, a) == 'B') {
                            valid = true;
                            const uint32_t sub = ld8(u, a + 1);
                            len = ld32u(u, a + 2);
                            const uint64_t h8 = (hz + 7ull) & ~7ull;
                            off8 = (h8 - p.heap_host) >> 3;
                            unsigned long long *dst = reinterpret_cast<unsigned long long *>(p.heap + (h8 - p.heap_host));
                            for (uint32_t e = 0; e < len; ++e) {
                                long long iv = 0;
                                bool is_int = true;
                                switch (sub) {
                                case 'c': iv = (int8_t)ld8(u, a + 6 + e); break;
                                case 'C': iv = ld8(u, a + 6 + e); break;
                                case 's': iv = (int16_t)(ld8(u, a + 6 + 2ll * e) | (ld8(u, a + 7 + 2ll * e) << 8)); break;
                                case 'S': iv = ld8(u, a + 6 + 2ll * e) | (ld8(u, a + 7 + 2ll * e) << 8); break;
                                case 'i': iv = (int32_t)ld32u(u, a + 6 + 4ll * e); break;
                                case 'I': iv = ld32u(u, a + 6 + 4ll * e); break;
                                default: is_int = false; break;
                                }
                                if (sub == 'f') dst[e] = (unsigned long long)__double_as_longlong((double)__uint_as_float(ld32u(u, a + 6 + 4ll * e)));
                                else if (sub == 'd') dst[e] = 0ull;                   // bam_auxB2f: not 'f' -> integer getter -> EINVAL, 0
                                else dst[e] = is_int ? (unsigned long long)iv : 0ull;
                            }
                            hz += 4ull + 8ull * len;
                        }
                        reinterpret_cast<ulonglong2 *>(p.col[c])[row] = make_ulonglong2(off8, len);
                    } else if (ty_req == 'Z') {
                        const int32_t zl = aux_zlen(u, r, a);        // NULL for any non-Z/H type (bam_aux2Z)
                        if (zl >= 0) {
                            valid = true;
                            if (zl > 12) copy_bytes(p.heap + (hz - p.heap_host), u, a + 1, (uint32_t)zl);
                            store_str_cell(p.col[c], row, u, a + 1, (uint32_t)zl, hz);
                            if (zl > 12) hz += align4((uint32_t)zl);
                        } else store_cell(p.col[c], row, 0, 0, 0, 0);
                    } else if (a == NONE) {
                        if (ty_req == 'A') store_cell(p.col[c], row, 0, 0, 0, 0);
                    } else {
                        valid = true;
                        const uint32_t ty = ld8(u, a);
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
                        if (ty_req == 'i') reinterpret_cast<int64_t *>(p.col[c])[row] = iv;   // non-integer type -> 0
                        else if (ty_req == 'f') {
                            double dv;
                            if (ty == 'd') dv = __longlong_as_double(((long long)ld32u(u, a + 5) << 32) | ld32u(u, a + 1));
                            else if (ty == 'f') dv = (double)__uint_as_float(ld32u(u, a + 1));
                            else dv = is_int ? (double)iv : 0.0;
                            reinterpret_cast<double *>(p.col[c])[row] = dv;
                        } else {  // 'A': one character, or "" on a type mismatch (bam_aux2A returns 0)
                            const uint32_t ch = ty == 'A' ? ld8(u, a + 1) : 0u;
                            store_cell(p.col[c], row, ch ? 1u : 0u, ch, 0, 0);
                        }
                    }
                }
                set_valid_group(p.valid[c], row0, __ballot_sync(FULL, valid), lane);
            }
        // READ_GROUP_ID / SAMPLE_ID (bam_reader.c:879-918): RG:Z value; SM of the first @RG line with that ID
        if (cRG) {
            if (act) {
                if (z_rg >= 0) {
                    if (z_rg > 12) copy_bytes(p.heap + (hz - p.heap_host), u, h_rg + 1, (uint32_t)z_rg);
                    store_str_cell(p.col[COL_RG], row, u, h_rg + 1, (uint32_t)z_rg, hz);
                } else store_cell(p.col[COL_RG], row, 0, 0, 0, 0);
            }
            set_valid_group(p.valid[COL_RG], row0, __ballot_sync(FULL, act && z_rg >= 0), lane);
        }
        if (cSM) {
            int found = -1;
            if (act && z_rg >= 0)
                for (uint32_t e = 0; e < t.n_rg && found < 0; ++e) {
                    if (t.rg[e].id_len != (uint32_t)z_rg) continue;
                    const uint8_t *id = t.rg_ids + t.rg[e].id_off;
                    bool m = true;
                    for (int32_t q = 0; q < z_rg && m; ++q) m = id[q] == ld8(u, h_rg + 1 + q);
                    if (m) found = (int)e;
                }
            const bool ok = found >= 0 && t.rg[found].has_sample;
            if (act) {
                if (ok) reinterpret_cast<uint4 *>(p.col[COL_SAMPLE])[row] = *reinterpret_cast<const uint4 *>(&t.rg[found].sample);
                else store_cell(p.col[COL_SAMPLE], row, 0, 0, 0, 0);
            }
            set_valid_group(p.valid[COL_SAMPLE], row0, __ballot_sync(FULL, ok), lane);
        }
        hp += total;
    }
}


// =======================================================================================================
// columnar decode, COMPACT WIRE LAYOUT (DH_BAM_COMPACT_OUTPUT, include/duckhts_cuda.h): same conversions, fewer bytes
// =======================================================================================================
// Same step structure as bam_decode_kernel.  What leaves differs: BAM's 32-bit fields stay 32-bit, dictionary columns are
// codes, and a VARCHAR cell is a 4-byte length with the bytes in the row's heap slots (every string, short ones too;
// slots 4-byte aligned, rec_off[row] = first slot).  514 -> ~410 bytes per record on the WGS-like bench file: the
// device-to-host link is the binding resource of the scan, and the host touches every cell anyway when it fills DuckDB's vectors.
constexpr uint32_t LEN_FLAG = 0x80000000u;   // QNAME / QUAL length words: the value ends at an embedded NUL (host: strnlen)
__device__ __forceinline__ void put_short(uint8_t *dst, uint32_t len, uint32_t w0, uint32_t w1, uint32_t w2) {
    uint32_t *d = reinterpret_cast<uint32_t *>(dst);
    if (len > 0u) d[0] = w0;
    if (len > 4u) d[1] = w1;
    if (len > 8u) d[2] = w2;
}

__global__ void __launch_bounds__(128) bam_decode_compact_kernel(const __grid_constant__ DecodeParams p, const ScanTables t) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t bi = p.b_lo + blockIdx.x * 4u + (threadIdx.x >> 5);
    if (bi >= p.b_hi) return;
    const uint32_t count = p.walk[bi].count;
    if (!count) return;
    const uint8_t *u = p.u;
    const int64_t bstart = p.blk_start[bi];
    const uint16_t *rec = p.rec + ((bstart - p.rec_bias) / 36 + (int64_t)bi);
    const uint64_t mask = t.col_mask;
    const bool cQ = (mask >> COL_QNAME) & 1, cC = (mask >> COL_CIGAR) & 1, cS = (mask >> COL_SEQ) & 1, cU = (mask >> COL_QUAL) & 1,
               cRG = (mask >> COL_RG) & 1, cSM = (mask >> COL_SAMPLE) & 1;
    const bool any_tag_cols = t.any_tag_cols != 0;
    uint32_t row0 = p.row_base[bi];
    uint64_t hp = p.heap_base[bi];

    for (uint32_t g = 0; g < count; g += 32u, row0 += 32u) {
        const uint32_t cnt = count - g < 32u ? count - g : 32u;
        const bool act = lane < cnt;
        const uint32_t row = row0 + lane;
        // ---------------- A: own record, fixed-width columns at BAM's own widths ----------------
        Rec r;
        r.off = r.end = r.qname = r.cigar = r.seq = r.qual = r.aux = 0;
        r.cg_lo = r.cg_hi = NONE;
        r.block_len = r.n_cigar = r.cig_n = r.flag = r.mapq = r.cigar_text_len = 0u;
        r.l_qname = 1u;
        r.tid = r.pos = r.mtid = r.mpos = r.isize = -1;
        r.l_qseq = 0;
        uint32_t roff = 0;
        if (act) {
            roff = rec[g + lane];
            parse_record(u, bstart + roff, INT64_MAX / 2, false, t.n_ref, r);
            if ((mask >> COL_FLAG) & 1) reinterpret_cast<uint16_t *>(p.col[COL_FLAG])[row] = (uint16_t)r.flag;
            if ((mask >> COL_POS) & 1) reinterpret_cast<int32_t *>(p.col[COL_POS])[row] = r.pos;
            if ((mask >> COL_MAPQ) & 1) p.col[COL_MAPQ][row] = (uint8_t)r.mapq;
            if ((mask >> COL_PNEXT) & 1) reinterpret_cast<int32_t *>(p.col[COL_PNEXT])[row] = r.mpos;
            if ((mask >> COL_TLEN) & 1) reinterpret_cast<int32_t *>(p.col[COL_TLEN])[row] = r.isize;
            if ((mask >> COL_RNAME) & 1) reinterpret_cast<int32_t *>(p.col[COL_RNAME])[row] = r.tid >= 0 ? r.tid : -1;
            if ((mask >> COL_RNEXT) & 1) reinterpret_cast<int32_t *>(p.col[COL_RNEXT])[row] = r.mtid >= 0 ? r.mtid : -1;
        }
        const uint32_t L = (uint32_t)r.l_qseq;
        const bool absent = !act || qual_absent(u, r);
        uint32_t qn = r.l_qname;   // strlen of the name within l_qname bytes
        if (act && cQ) {
            for (uint32_t k = 0; k < r.l_qname; k += 4u) {
                const uint32_t z = first_zero_byte(ld32u(u, r.qname + k));
                if (z < 4u) { if (k + z < qn) qn = k + z; break; }
            }
        }
        // ---------------- B: heap slots (the rules of the walk; every string has one) ----------------
        const uint32_t ce = r.cig_n ? r.cigar_text_len : 1u, se = L ? L : 1u, ue = absent ? 1u : L;
        const uint32_t s_q = cQ ? align4(r.l_qname) : 0u, s_c = cC ? align4(ce) : 0u, s_s = cS ? align4(se) : 0u, s_u = cU ? align4(ue) : 0u;
        uint32_t need = act ? s_q + s_c + s_s + s_u : 0u;
        const uint32_t aux_slot0 = need;
        int64_t h_rg = NONE;
        int32_t z_rg = -1;
        if (act && (cRG || cSM)) {
            h_rg = aux_find(u, r, 'R' | ('G' << 8));
            z_rg = aux_zlen(u, r, h_rg);
        }
        if (act && any_tag_cols)
            for (uint32_t k = 0; k < t.n_tags; ++k) {
                if ((t.tags[k].type != 'Z' && t.tags[k].type != 'B') || !t.tags[k].want) continue;
                const int64_t a = aux_find(u, r, t.tags[k].t0 | (t.tags[k].t1 << 8));
                if (t.tags[k].type == 'B') {
                    if (a != NONE && ld8(u, a) == 'B') need += 4u + 8u * ld32u(u, a + 2);
                    continue;
                }
                const int32_t zl = aux_zlen(u, r, a);
                if (zl > 0) need += align4((uint32_t)zl);
            }
        if (act && cRG && z_rg > 0) need += align4((uint32_t)z_rg);
        uint32_t incl = need;
#pragma unroll
        for (int sft = 1; sft < 32; sft <<= 1) { const uint32_t v = __shfl_up_sync(FULL, incl, sft); if (lane >= (uint32_t)sft) incl += v; }
        const uint32_t rel = incl - need;
        const uint32_t total = __shfl_sync(FULL, incl, 31);
        if (act) p.rec_off[row] = (uint32_t)(hp + rel);
        // ---------------- C: strings longer than 12 bytes, record by record, whole warp ----------------
        uint32_t qcut = L;
        const uint32_t lqnc = r.l_qname | (r.n_cigar << 8), lab = L | (absent ? 0x80000000u : 0u);
        for (uint32_t q = 0; q < cnt; ++q) {
            const uint32_t o_ = __shfl_sync(FULL, roff, q), lq_ = __shfl_sync(FULL, lqnc, q), la_ = __shfl_sync(FULL, lab, q),
                           ct_ = __shfl_sync(FULL, r.cigar_text_len, q), rel_ = __shfl_sync(FULL, rel, q), cn = __shfl_sync(FULL, r.cig_n, q);
            const uint32_t lqn = lq_ & 0xffu, ncg = lq_ >> 8, Lq = la_ & 0x7fffffffu;
            const bool abs_ = (la_ >> 31) != 0u;
            const int64_t qname = bstart + o_ + 36, seq = qname + lqn + 4ll * ncg, qual = seq + (((int64_t)Lq + 1) >> 1);
            uint8_t *slot = p.heap + hp + rel_;
            if (cQ) {
                if (lqn > 12u) {
                    uint32_t *d = reinterpret_cast<uint32_t *>(slot);
                    for (uint32_t k = lane; 4u * k < lqn; k += 32u) d[k] = mask_bytes(ld32u(u, qname + 4ll * k), lqn - 4u * k);
                }
                slot += align4(lqn);
            }
            if (cC) {
                if (cn && ct_ > 12u) {
                    const int64_t cig = bstart + __shfl_sync(FULL, (int32_t)(r.cigar - bstart), q);
                    uint32_t base = 0;
                    for (uint32_t k0 = 0; k0 < cn; k0 += 32u) {
                        const uint32_t k = k0 + lane;
                        uint32_t c = 0, dgt = 0;
                        if (k < cn) { c = ld32u(u, cig + 4ll * k); dgt = dec_digits(c >> 4) + 1u; }
                        uint32_t inc = dgt;
#pragma unroll
                        for (int sft = 1; sft < 32; sft <<= 1) { const uint32_t v = __shfl_up_sync(FULL, inc, sft); if (lane >= (uint32_t)sft) inc += v; }
                        if (k < cn) {
                            uint32_t v = c >> 4, at = base + inc - 1u;
                            slot[at] = (uint8_t)"MIDNSHP=XB??????"[c & 15u];
                            for (uint32_t j = dgt - 1u; j-- > 0;) { slot[--at] = (uint8_t)('0' + v % 10u); v /= 10u; }
                        }
                        base += __shfl_sync(FULL, inc, 31);
                    }
                }
                slot += align4(cn ? ct_ : 1u);
            }
            if (cS) {
                if (Lq > 12u) {
                    uint32_t *d = reinterpret_cast<uint32_t *>(slot);
                    for (uint32_t k = lane; 8u * k < Lq; k += 32u) {
                        const uint32_t pk = ld32u(u, seq + 4ll * k);
                        d[2u * k] = mask_bytes(seq4(pk & 0xffffu), Lq - 8u * k);
                        if (8u * k + 4u < Lq) d[2u * k + 1u] = mask_bytes(seq4(pk >> 16), Lq - 8u * k - 4u);
                    }
                }
                slot += align4(Lq ? Lq : 1u);
            }
            if (cU) {
                if (!abs_ && Lq > 12u) {
                    uint32_t *d = reinterpret_cast<uint32_t *>(slot);
                    uint32_t zpos = 0xffffffffu;
                    for (uint32_t k = lane; 8u * k < Lq; k += 32u) {
                        const uint32_t a = mask_bytes(qual4(ld32u(u, qual + 8ll * k)), Lq - 8u * k) | (Lq - 8u * k < 4u ? 0xffffffffu << (8u * (Lq - 8u * k)) : 0u);
                        d[2u * k] = mask_bytes(a, Lq - 8u * k);
                        uint32_t z = first_zero_byte(a);
                        if (z < 4u && zpos == 0xffffffffu) zpos = 8u * k + z;
                        if (8u * k + 4u < Lq) {
                            const uint32_t rem = Lq - 8u * k - 4u;
                            const uint32_t b = mask_bytes(qual4(ld32u(u, qual + 8ll * k + 4)), rem) | (rem < 4u ? 0xffffffffu << (8u * rem) : 0u);
                            d[2u * k + 1u] = mask_bytes(b, rem);
                            z = first_zero_byte(b);
                            if (z < 4u && zpos == 0xffffffffu) zpos = 8u * k + 4u + z;
                        }
                    }
                    zpos = __reduce_min_sync(FULL, zpos);
                    if (lane == q && zpos < qcut) qcut = zpos;
                }
                slot += align4(abs_ ? 1u : Lq);
            }
        }
        __syncwarp();
        // ---------------- D: short strings and length words of the lane's own row ----------------
        uint8_t *const slot0 = p.heap + hp + rel;
        if (act) {
            uint8_t *hs = slot0;
            if (cQ) {
                if (r.l_qname <= 12u) { uint32_t w0, w1, w2; inline_words(u, r.qname, r.l_qname, w0, w1, w2); put_short(hs, r.l_qname, w0, w1, w2); }
                reinterpret_cast<uint32_t *>(p.col[COL_QNAME])[row] = r.l_qname | (qn != r.l_qname - 1u ? LEN_FLAG : 0u);
                hs += s_q;
            }
            if (cC) {
                if (r.cig_n == 0u) put_short(hs, 1u, '*', 0u, 0u);
                else if (ce <= 12u) { uint32_t w0, w1, w2; cigar_inline(u, r, w0, w1, w2); put_short(hs, ce, w0, w1, w2); }
                reinterpret_cast<uint32_t *>(p.col[COL_CIGAR])[row] = ce;
                hs += s_c;
            }
            if (cS) {
                if (L == 0u) put_short(hs, 1u, '*', 0u, 0u);
                else if (L <= 12u) {
                    const uint32_t p0 = ld32u(u, r.seq), p1 = L > 8u ? ld32u(u, r.seq + 4) : 0u;
                    put_short(hs, L, mask_bytes(seq4(p0 & 0xffffu), L), L > 4u ? mask_bytes(seq4(p0 >> 16), L - 4u) : 0u,
                              L > 8u ? mask_bytes(seq4(p1 & 0xffffu), L - 8u) : 0u);
                }
                reinterpret_cast<uint32_t *>(p.col[COL_SEQ])[row] = se;
                hs += s_s;
            }
            if (cU) {
                uint32_t cut = qcut;
                if (absent) put_short(hs, 1u, '*', 0u, 0u);
                else if (L <= 12u) {
                    const uint32_t w0 = mask_bytes(qual4(ld32u(u, r.qual)), L), w1 = L > 4u ? mask_bytes(qual4(ld32u(u, r.qual + 4)), L - 4u) : 0u,
                                   w2 = L > 8u ? mask_bytes(qual4(ld32u(u, r.qual + 8)), L - 8u) : 0u;
                    put_short(hs, L, w0, w1, w2);
                    uint32_t z = first_zero_byte(w0 | (L < 4u ? 0xffffffffu << (8u * L) : 0u));
                    if (z < 4u) cut = z;
                    else if (L > 4u && (z = first_zero_byte(w1 | (L < 8u ? 0xffffffffu << (8u * (L - 4u)) : 0u))) < 4u) cut = 4u + z;
                    else if (L > 8u && (z = first_zero_byte(w2 | (L < 12u ? 0xffffffffu << (8u * (L - 8u)) : 0u))) < 4u) cut = 8u + z;
                }
                reinterpret_cast<uint32_t *>(p.col[COL_QUAL])[row] = absent ? 1u : (L | (cut != L ? LEN_FLAG : 0u));
                hs += s_u;
            }
        }
        // typed standard tags
        uint8_t *hz = slot0 + aux_slot0;
        if (any_tag_cols)
            for (uint32_t k = 0; k < t.n_tags; ++k) {
                const int c = BAM_CORE_COLS + (int)k;
                if (!t.tags[k].want) continue;
                const uint32_t ty_req = t.tags[k].type;
                bool valid = false;
                if (act && ty_req == 'c') {
                    long long mv;
                    valid = cigar_metric(u, r, t.tags[k].t0, t.tags[k].t1, mv);
                    reinterpret_cast<int64_t *>(p.col[c])[row] = mv;
                } else if (act) {
                    const int64_t a = aux_find(u, r, t.tags[k].t0 | (t.tags[k].t1 << 8));
                    if (ty_req == 'B') {
                        uint64_t off8 = 0, len = 0;
                        if (a != NONE && ld8(u, a) == 'B') {
                            valid = true;
                            const uint32_t sub = ld8(u, a + 1);
                            len = ld32u(u, a + 2);
                            const uint64_t ho = (uint64_t)(hz - p.heap), h8 = (ho + 7ull) & ~7ull;
                            off8 = h8 >> 3;
                            unsigned long long *dst = reinterpret_cast<unsigned long long *>(p.heap + h8);
                            for (uint32_t e = 0; e < len; ++e) {
                                long long iv = 0;
                                bool is_int = true;
                                switch (sub) {
                                case 'c': iv = (int8_t)ld8(u, a + 6 + e); break;
                                case 'C': iv = ld8(u, a + 6 + e); break;
                                case 's': iv = (int16_t)(ld8(u, a + 6 + 2ll * e) | (ld8(u, a + 7 + 2ll * e) << 8)); break;
                                case 'S': iv = ld8(u, a + 6 + 2ll * e) | (ld8(u, a + 7 + 2ll * e) << 8); break;
                                case 'i': iv = (int32_t)ld32u(u, a + 6 + 4ll * e); break;
                                case 'I': iv = ld32u(u, a + 6 + 4ll * e); break;
                                default: is_int = false; break;
                                }
                                if (sub == 'f') dst[e] = (unsigned long long)__double_as_longlong((double)__uint_as_float(ld32u(u, a + 6 + 4ll * e)));
                                else if (sub == 'd') dst[e] = 0ull;
                                else dst[e] = is_int ? (unsigned long long)iv : 0ull;
                            }
                            hz += 4ull + 8ull * len;
                        }
                        reinterpret_cast<ulonglong2 *>(p.col[c])[row] = make_ulonglong2(off8, len);
                    } else if (ty_req == 'Z') {
                        const int32_t zl = aux_zlen(u, r, a);
                        if (zl >= 0) {
                            valid = true;
                            if (zl > 0) copy_bytes(hz, u, a + 1, (uint32_t)zl);
                            hz += align4((uint32_t)zl);
                        }
                        reinterpret_cast<uint32_t *>(p.col[c])[row] = zl >= 0 ? (uint32_t)zl : 0u;
                    } else if (a == NONE) {
                        if (ty_req == 'A') p.col[c][row] = 0;
                    } else {
                        valid = true;
                        const uint32_t ty = ld8(u, a);
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
                        if (ty_req == 'i') reinterpret_cast<int64_t *>(p.col[c])[row] = iv;
                        else if (ty_req == 'f') {
                            double dv;
                            if (ty == 'd') dv = __longlong_as_double(((long long)ld32u(u, a + 5) << 32) | ld32u(u, a + 1));
                            else if (ty == 'f') dv = (double)__uint_as_float(ld32u(u, a + 1));
                            else dv = is_int ? (double)iv : 0.0;
                            reinterpret_cast<double *>(p.col[c])[row] = dv;
                        } else p.col[c][row] = ty == 'A' ? (uint8_t)ld8(u, a + 1) : (uint8_t)0;
                    }
                }
                set_valid_group(p.valid[c], row0, __ballot_sync(FULL, valid), lane);
            }
        if (cRG) {
            if (act) {
                if (z_rg > 0) copy_bytes(hz, u, h_rg + 1, (uint32_t)z_rg);
                reinterpret_cast<uint32_t *>(p.col[COL_RG])[row] = z_rg >= 0 ? (uint32_t)z_rg : 0u;
            }
            set_valid_group(p.valid[COL_RG], row0, __ballot_sync(FULL, act && z_rg >= 0), lane);
        }
        if (cSM) {
            int found = -1;
            if (act && z_rg >= 0)
                for (uint32_t e = 0; e < t.n_rg && found < 0; ++e) {
                    if (t.rg[e].id_len != (uint32_t)z_rg) continue;
                    const uint8_t *id = t.rg_ids + t.rg[e].id_off;
                    bool m = true;
                    for (int32_t q = 0; q < z_rg && m; ++q) m = id[q] == ld8(u, h_rg + 1 + q);
                    if (m) found = (int)e;
                }
            const bool ok = found >= 0 && t.rg[found].has_sample;
            if (act) reinterpret_cast<int32_t *>(p.col[COL_SAMPLE])[row] = ok ? found : -1;
            set_valid_group(p.valid[COL_SAMPLE], row0, __ballot_sync(FULL, ok), lane);
        }
        hp += total;
    }
}


// =======================================================================================================
// AUXILIARY_TAGS MAP column (A16)
// =======================================================================================================
// printf("%g") of a double, exactly as glibc rounds it (correct rounding of the exact binary value, ties to even): the six
// significant digits come from an exact big-integer division R / S (the "fixed format" branch of Steele & White's dragon4).
// 36 words cover the largest cases (2^1024 numerators, 10^324 * 2^53 for subnormals).  Values are rare in the aux area
// (types f / d and B:f only), so the routine is kept out of line.
struct Big {
    uint32_t w[36];
    int n;   // words in use
};
__device__ __forceinline__ void big_set(Big &a, uint64_t v) {
    a.w[0] = (uint32_t)v;
    a.w[1] = (uint32_t)(v >> 32);
    a.n = a.w[1] ? 2 : 1;
}
__device__ __forceinline__ void big_mul(Big &a, uint32_t m) {
    uint64_t c = 0;
    for (int i = 0; i < a.n; ++i) {
        c += (uint64_t)a.w[i] * m;
        a.w[i] = (uint32_t)c;
        c >>= 32;
    }
    if (c && a.n < 36) a.w[a.n++] = (uint32_t)c;
}
__device__ __forceinline__ void big_shl(Big &a, int bits) {
    const int ws = bits >> 5, bs = bits & 31;
    int n = a.n + ws + 1;
    if (n > 36) n = 36;
    for (int i = n - 1; i >= 0; --i) {
        const int j = i - ws;
        uint32_t v = 0;
        if (j >= 0 && j < a.n) v = a.w[j] << bs;
        if (bs && j - 1 >= 0 && j - 1 < a.n) v |= a.w[j - 1] >> (32 - bs);
        a.w[i] = v;
    }
    a.n = n;
    while (a.n > 1 && a.w[a.n - 1] == 0) --a.n;
}
__device__ __forceinline__ int big_cmp(const Big &a, const Big &b) {
    if (a.n != b.n) return a.n > b.n ? 1 : -1;
    for (int i = a.n - 1; i >= 0; --i)
        if (a.w[i] != b.w[i]) return a.w[i] > b.w[i] ? 1 : -1;
    return 0;
}
__device__ __forceinline__ void big_sub(Big &a, const Big &b) {   // a -= b, a >= b
    uint32_t borrow = 0;
    for (int i = 0; i < a.n; ++i) {
        const uint64_t d = (uint64_t)a.w[i] - (i < b.n ? b.w[i] : 0u) - borrow;
        a.w[i] = (uint32_t)d;
        borrow = (uint32_t)(d >> 63);
    }
    while (a.n > 1 && a.w[a.n - 1] == 0) --a.n;
}
__device__ __forceinline__ void big_pow10(Big &a, int e) {   // a *= 10^e
    for (; e >= 9; e -= 9) big_mul(a, 1000000000u);
    uint32_t m = 1;
    for (; e > 0; --e) m *= 10u;
    if (m > 1) big_mul(a, m);
}
// six significant decimal digits of |x| (finite, non-zero): digits in [100000, 999999], x ~ digits * 10^(exp10 - 5)
__device__ __noinline__ void g6_digits(double ax, uint32_t &digits, int &exp10) {
    const uint64_t bits = (uint64_t)__double_as_longlong(ax);
    int be = (int)((bits >> 52) & 0x7ffu);
    uint64_t m = bits & 0xfffffffffffffull;
    if (be) m |= 1ull << 52; else be = 1;
    const int k = be - 1075;   // ax = m * 2^k
    Big R, S;
    big_set(R, m);
    big_set(S, 1);
    if (k >= 0) big_shl(R, k); else big_shl(S, -k);
    int e = (int)floor(log10(ax));
    if (e >= 0) big_pow10(S, e); else big_pow10(R, -e);
    while (big_cmp(R, S) < 0) { big_mul(R, 10u); --e; }          // the estimate may be one off either way
    for (;;) {
        Big T = S;
        big_mul(T, 10u);
        if (big_cmp(R, T) < 0) break;
        S = T;
        ++e;
    }
    uint32_t d6 = 0;
    for (int i = 0; i < 6; ++i) {
        uint32_t d = 0;
        while (big_cmp(R, S) >= 0) { big_sub(R, S); ++d; }
        d6 = d6 * 10u + d;
        if (i < 5) big_mul(R, 10u);
    }
    big_shl(R, 1);                                               // remainder against one half: ties go to the even digit
    const int c = big_cmp(R, S);
    if (c > 0 || (c == 0 && (d6 & 1u))) ++d6;
    if (d6 == 1000000u) { d6 = 100000u; ++e; }
    digits = d6;
    exp10 = e;
}

// text sink of one MAP value: a heap slot when the bound needs one, else 16 bytes of local memory
struct Sink {
    uint8_t *dst;
    uint32_t len;
};
__device__ __forceinline__ void put(Sink &k, uint32_t c) { k.dst[k.len++] = (uint8_t)c; }
__device__ __forceinline__ void put_i64(Sink &k, long long v) {   // "%" PRId64
    unsigned long long a = v < 0 ? 0ull - (unsigned long long)v : (unsigned long long)v;
    if (v < 0) put(k, '-');
    char tmp[20];
    int n = 0;
    do { tmp[n++] = (char)('0' + a % 10ull); a /= 10ull; } while (a);
    while (n) put(k, tmp[--n]);
}
__device__ __noinline__ void put_g(Sink &k, double x) {   // "%g"
    const uint64_t bits = (uint64_t)__double_as_longlong(x);
    if (bits >> 63) put(k, '-');
    const uint64_t mag = bits & 0x7fffffffffffffffull;
    if (mag > 0x7ff0000000000000ull) { put(k, 'n'); put(k, 'a'); put(k, 'n'); return; }
    if (mag == 0x7ff0000000000000ull) { put(k, 'i'); put(k, 'n'); put(k, 'f'); return; }
    if (mag == 0) { put(k, '0'); return; }
    uint32_t d6;
    int X;
    g6_digits(__longlong_as_double((long long)mag), d6, X);
    int nd = 6;
    while (nd > 1 && d6 % 10u == 0u) { d6 /= 10u; --nd; }   // %g drops trailing zeros
    char ds[6];
    for (int i = nd - 1; i >= 0; --i) { ds[i] = (char)('0' + d6 % 10u); d6 /= 10u; }
    if (X < -4 || X >= 6) {                                  // exponent form, at least two exponent digits
        put(k, ds[0]);
        if (nd > 1) { put(k, '.'); for (int i = 1; i < nd; ++i) put(k, ds[i]); }
        put(k, 'e');
        put(k, X < 0 ? '-' : '+');
        const uint32_t ax = (uint32_t)(X < 0 ? -X : X);
        if (ax >= 100u) put(k, '0' + ax / 100u);
        put(k, '0' + (ax / 10u) % 10u);
        put(k, '0' + ax % 10u);
    } else if (X >= 0) {
        for (int i = 0; i <= X; ++i) put(k, i < nd ? ds[i] : '0');
        if (nd > X + 1) { put(k, '.'); for (int i = X + 1; i < nd; ++i) put(k, ds[i]); }
    } else {
        put(k, '0');
        put(k, '.');
        for (int i = -1; i > X; --i) put(k, '0');
        for (int i = 0; i < nd; ++i) put(k, ds[i]);
    }
}
__device__ __forceinline__ long long aux_int_at(const uint8_t *u, uint32_t ty, int64_t o) {   // get_int_aux_val, sam.c:5126-5141
    switch (ty) {
    case 'c': return (int8_t)ld8(u, o);
    case 'C': return ld8(u, o);
    case 's': return (int16_t)(ld8(u, o) | (ld8(u, o + 1) << 8));
    case 'S': return ld8(u, o) | (ld8(u, o + 1) << 8);
    case 'i': return (int32_t)ld32u(u, o);
    case 'I': return ld32u(u, o);
    default: return 0;
    }
}
// bam_aux_to_string (bam_reader.c:140-183) of the complete entry whose type byte is at s
__device__ __forceinline__ void aux_value_text(Sink &k, const uint8_t *u, int64_t s, uint32_t zlen) {
    const uint32_t ty = ld8(u, s);
    switch (ty) {
    case 'A': { const uint32_t c = ld8(u, s + 1); if (c) put(k, c); break; }      // the cell is assigned with strlen semantics
    case 'c': case 'C': case 's': case 'S': case 'i': case 'I': put_i64(k, aux_int_at(u, ty, s + 1)); break;
    case 'f': put_g(k, (double)__uint_as_float(ld32u(u, s + 1))); break;
    case 'd': put_g(k, __longlong_as_double((long long)(ld32u(u, s + 1) | ((uint64_t)ld32u(u, s + 5) << 32)))); break;
    case 'Z': case 'H': for (uint32_t i = 0; i < zlen; ++i) put(k, ld8(u, s + 1 + i)); break;
    case 'B': {
        const uint32_t sub = ld8(u, s + 1), n = ld32u(u, s + 2);
        const uint32_t esz = (sub == 'c' || sub == 'C') ? 1u : (sub == 's' || sub == 'S') ? 2u : 4u;
        put(k, sub);
        for (uint32_t e = 0; e < n; ++e) {
            put(k, ',');
            if (sub == 'f') put_g(k, (double)__uint_as_float(ld32u(u, s + 6 + 4ll * e)));
            else if (sub == 'd') put(k, '0');                                     // bam_auxB2f of a non-float sub-type: 0
            else put_i64(k, aux_int_at(u, sub, s + 6 + (int64_t)esz * e));
        }
        break;
    }
    default: break;
    }
}

// One warp per BGZF block, one lane per record (as the decode kernel): the pairs of a row sit at consecutive child indices,
// rows in order, so the host appends keys / values to the DuckDB child vectors with two copies per vector chunk.
__global__ void __launch_bounds__(128) bam_auxmap_kernel(const __grid_constant__ AuxMapParams p, const ScanTables t) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t bi = p.b_lo + blockIdx.x * 4u + (threadIdx.x >> 5);
    if (bi >= p.b_hi) return;
    const BlockWalk bw = p.walk[bi];
    if (!bw.count) return;
    const uint8_t *u = p.u;
    const int64_t bstart = p.blk_start[bi];
    const uint16_t *rec = p.rec + ((bstart - p.rec_bias) / 36 + (int64_t)bi);
    uint32_t row0 = p.row_base[bi];
    uint32_t child = p.map_base[bi];
    uint64_t hp = p.heap_base[bi] + bw.heap_bytes - bw.map_heap;   // the MAP's slots are the tail of the block's heap range
    for (uint32_t g = 0; g < bw.count; g += 32u, row0 += 32u) {
        const bool act = lane < bw.count - g;
        Rec r;
        uint32_t cnt = 0, need = 0;
        if (act) {
            parse_record(u, bstart + rec[g + lane], INT64_MAX / 2, false, t.n_ref, r);
            aux_map_need(u, r, t, cnt, need);
        }
        uint32_t ci = cnt, hi = need;   // inclusive warp scans
#pragma unroll
        for (int sft = 1; sft < 32; sft <<= 1) {
            const uint32_t a = __shfl_up_sync(FULL, ci, sft), b = __shfl_up_sync(FULL, hi, sft);
            if (lane >= (uint32_t)sft) { ci += a; hi += b; }
        }
        uint32_t at = child + ci - cnt;
        uint64_t hs = hp + hi - need;
        if (act) {
            reinterpret_cast<ulonglong2 *>(p.entries)[row0 + lane] = make_ulonglong2(at, cnt);
            if (cnt) {
                int64_t s = r.aux + 2;
                for (;;) {
                    if (s - 2 == r.cg_lo) {
                        if (r.end - r.cg_hi <= 2) break;
                        s = r.cg_hi + 2;
                        continue;
                    }
                    bool complete;
                    const int64_t nx = skip_aux_scalar(u, s, r.end, &complete);
                    const uint32_t t0 = ld8(u, s - 2), t1 = ld8(u, s - 1);
                    if (!aux_excluded(t, t0 | (t1 << 8))) {
                        // key: tagbuf[3] assigned with strlen semantics
                        const uint32_t kl = t0 == 0 ? 0u : t1 == 0 ? 1u : 2u;
                        store_cell(p.keys, at, kl, kl == 2 ? (t0 | (t1 << 8)) : kl == 1 ? t0 : 0u, 0u, 0u);
                        uint8_t local[16];
                        Sink k;
                        k.len = 0;
                        k.dst = local;
                        uint32_t bound = 0;
                        if (nx != NONE) {
                            bound = aux_text_bound(u, s, nx, r.end);
                            if (bound > 12u) k.dst = p.heap + hs;
                            const uint32_t ty = ld8(u, s);
                            aux_value_text(k, u, s, (ty == 'Z' || ty == 'H') ? bound : 0u);
                        }
                        if (k.len <= 12u) {
                            uint32_t w[3] = {0u, 0u, 0u};
                            for (uint32_t i = 0; i < k.len; ++i) w[i >> 2] |= (uint32_t)k.dst[i] << (8u * (i & 3u));
                            store_cell(p.values, at, k.len, w[0], w[1], w[2]);
                        } else {
                            const uint64_t hptr = p.heap_host + hs;
                            store_cell(p.values, at, k.len, k.dst[0] | (k.dst[1] << 8) | (k.dst[2] << 16) | ((uint32_t)k.dst[3] << 24),
                                       (uint32_t)hptr, (uint32_t)(hptr >> 32));
                        }
                        hs += slot_bytes(bound);
                        ++at;
                    }
                    if (nx == NONE || r.end - nx <= 2) break;
                    s = nx + 2;
                }
            }
        }
        set_valid_group(p.valid, row0, __ballot_sync(FULL, act && cnt > 0), lane);
        child += __shfl_sync(FULL, ci, 31);
        hp += __shfl_sync(FULL, hi, 31);
    }
}

// =======================================================================================================
// N2: runs of equal (tid, bin) + linear index for bam_index (htslib/hts.c hts_idx_push :2553-2642; sam_index, sam.c:989-1027)
// =======================================================================================================
__device__ __forceinline__ uint32_t reg2bin_bai(int64_t beg, int64_t end) {   // hts_reg2bin(beg, end, 14, 5), htslib/hts.h
    --end;
    if (beg >> 14 == end >> 14) return (uint32_t)(4681 + (beg >> 14));
    if (beg >> 17 == end >> 17) return (uint32_t)(585 + (beg >> 17));
    if (beg >> 20 == end >> 20) return (uint32_t)(73 + (beg >> 20));
    if (beg >> 23 == end >> 23) return (uint32_t)(9 + (beg >> 23));
    if (beg >> 26 == end >> 26) return (uint32_t)(1 + (beg >> 26));
    return 0u;
}
__global__ void __launch_bounds__(128) bam_index_kernel(const IndexParams p, const ScanTables t) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.n) return;
    const uint32_t count = p.walk[i].count;
    const int64_t bstart = p.blk_start[i];
    const uint16_t *rec = p.rec + ((bstart - p.rec_bias) / 36 + (int64_t)i);
    IdxBlock ob;
    ob.base = ob.n_runs = 0;
    ob.first_tid = ob.last_tid = -2;
    ob.first_beg = ob.last_beg = 0;
    ob.err = 0;
    ob.count = count;
    // pass 0 counts the runs, pass 1 writes them (one atomicAdd per block keeps a block's runs together and in order)
    for (int pass = 0; pass < 2 && count; ++pass) {
        uint32_t nruns = 0, cur_bin = 0xffffffffu, nm = 0, nu = 0;
        int32_t cur_tid = -2;
        int64_t last_beg = 0;
        IdxRun *out = p.runs + ob.base;
        for (uint32_t j = 0; j < count; ++j) {
            Rec r;
            const int64_t off = bstart + rec[j];
            parse_record(p.u, off, INT64_MAX / 2, false, t.n_ref, r);
            int64_t beg = r.pos, rl = (r.flag & 4u) ? 0 : r.rlen;          // bam_endpos (sam.c:519-524)
            if (rl == 0) rl = 1;
            int64_t end = beg + rl;
            if (r.tid < 0) { beg = -1; end = 0; }
            if (pass == 0) {
                if (j == 0) { ob.first_tid = r.tid; ob.first_beg = beg; }
                else if (r.tid == cur_tid && r.tid >= 0 && last_beg > beg) ob.err = ob.err ? ob.err : 1;   // unsorted positions
                if (r.tid >= 0 && (beg > (1ll << 29) || end > (1ll << 29))) ob.err = ob.err ? ob.err : 2;
                ob.last_tid = r.tid;
            }
            if (r.tid >= 0) {                                                  // shoehorn [-1, 0) into the leftmost window (hts.c:2625-2627)
                if (beg < 0) beg = 0;
                if (end <= 0) end = 1;
            }
            last_beg = beg;                                                    // z.last_coor is the clamped start (hts.c:2641)
            if (pass == 0) ob.last_beg = beg;
            const uint32_t bin = reg2bin_bai(beg, end);
            if (r.tid != cur_tid || bin != cur_bin) {
                if (pass == 1 && nruns) { IdxRun &q = out[nruns - 1]; q.n_mapped = nm; q.n_unmapped = nu; }
                if (pass == 1) { IdxRun &q = out[nruns]; q.start_u = (uint64_t)(p.u_base + off); q.tid = r.tid; q.bin = bin; q.n_mapped = q.n_unmapped = 0; }
                ++nruns;
                cur_tid = r.tid;
                cur_bin = bin;
                nm = nu = 0;
            }
            if (r.flag & 4u) ++nu; else ++nm;
            if (pass == 1 && r.tid >= 0) {
                // linear index (insert_to_l, hts.c:2338-2362): the first record that touches a window, in file order
                const int64_t w0 = beg >> 14, w1 = (end - 1) >> 14;
                const uint64_t me = (uint64_t)(p.u_base + off);
                for (int64_t w = w0; w <= w1; ++w) {
                    if ((uint64_t)w >= p.lin_n[r.tid]) { p.blocks[i].err = 3; break; }
                    atomicMin(&p.lin[p.lin_base[r.tid] + w], (unsigned long long)me);
                }
            }
        }
        if (pass == 0) {
            ob.n_runs = nruns;
            ob.base = atomicAdd(p.counter, nruns);
            if (ob.base + nruns > p.run_cap) { ob.err = 4; ob.n_runs = 0; p.blocks[i] = ob; return; }
            p.blocks[i] = ob;
        } else if (nruns) { IdxRun &q = out[nruns - 1]; q.n_mapped = nm; q.n_unmapped = nu; }
    }
    if (!count) p.blocks[i] = ob;
}

cudaError_t launch_bam_index(const IndexParams &p, const ScanTables &t, cudaStream_t stream) {
    if (p.n == 0) return cudaSuccess;
    bam_index_kernel<<<(p.n + 127u) / 128u, 128, 0, stream>>>(p, t);
    return cudaGetLastError();
}

// DH_CARVEOUT_MAX=1 (A/B, profiles/r2i): ask for the largest shared-memory carve-out although these kernels use none, so that an SM
// configured for the inflate's 222 KB phase-1 CTA does not have to drain before one of their CTAs may join it.
static void column_kernels_carveout() {
    static std::once_flag once[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return;
    std::call_once(once[dev], [] {
        const char *e = getenv("DH_CARVEOUT_MAX");
        if (!e || !atoi(e)) return;
        cudaFuncSetAttribute(bam_walk_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        cudaFuncSetAttribute(bam_decode_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        cudaFuncSetAttribute(bam_decode_compact_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    });
}

cudaError_t launch_bam_walk(const WalkParams &p, const ScanTables &t, cudaStream_t stream) {
    column_kernels_carveout();
    if (p.n == 0) return cudaSuccess;
    bam_walk_kernel<<<(p.n + 127u) / 128u, 128, 0, stream>>>(p, t);
    return cudaGetLastError();
}
cudaError_t launch_bam_auxmap(const AuxMapParams &p, const ScanTables &t, cudaStream_t stream) {
    if (p.b_hi <= p.b_lo) return cudaSuccess;
    bam_auxmap_kernel<<<(p.b_hi - p.b_lo + 3u) / 4u, 128, 0, stream>>>(p, t);
    return cudaGetLastError();
}
cudaError_t launch_bam_decode(const DecodeParams &p, const ScanTables &t, cudaStream_t stream) {
    if (p.b_hi <= p.b_lo) return cudaSuccess;
    bam_decode_kernel<<<(p.b_hi - p.b_lo + 3u) / 4u, 128, 0, stream>>>(p, t);
    return cudaGetLastError();
}
cudaError_t launch_bam_decode_compact(const DecodeParams &p, const ScanTables &t, cudaStream_t stream) {
    if (p.b_hi <= p.b_lo) return cudaSuccess;
    bam_decode_compact_kernel<<<(p.b_hi - p.b_lo + 3u) / 4u, 128, 0, stream>>>(p, t);
    return cudaGetLastError();
}

}  // namespace dh
