// inflate_core.cuh — raw DEFLATE (RFC 1951) decoder core for one BGZF block.
//
// Replaces, for the read_bam / bgunzip hot path, what the reference gets from zlib's
// inflate() at htslib/bgzf.c:762-804 (bgzf_uncompress: inflateInit2(-15) + inflate(Z_FINISH)).
// This is our own decoder, not a port of zlib: table layout, build and decode loop are designed
// for a GPU lane-group that keeps its tables in shared memory.
//
// The code is written against an "IO policy" so that the SAME source is
//   * run by every lane of a lane-group on the device (IO = TMA-fed shared-memory ring in,
//     cooperative global-memory stores out; see bgzf_inflate.cu), and
//   * compiled for the host with a trivial IO (tests/: table builder + decode loop are unit-tested
//     on the CPU box against zlib streams; the host build is NEVER used by the product path).
//
// Acceptance rules follow zlib's so that valid/invalid verdicts match the reference's inflate():
// over-subscribed or incomplete code sets are errors (incomplete allowed only when the longest code
// is 1 bit), HLIT>286 / HDIST>30 rejected, missing end-of-block code rejected, litlen symbols
// 286/287 and distance symbols 30/31 rejected, distance beyond the start of the block rejected,
// output beyond 65536 bytes rejected (bgzf.c:810 gives zlib a 64 KiB output buffer).
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define DH_HD __host__ __device__ __forceinline__
#define DH_HDM __host__ __device__ __forceinline__
#define DH_HD_NOINLINE static __host__ __device__ __noinline__
#else
#define DH_HD static inline
#define DH_HDM inline
#define DH_HD_NOINLINE static __attribute__((noinline))
#endif

namespace dh {

// ---- table geometry -------------------------------------------------------------------------
// Root index widths; worst-case table sizes are the classic "enough" bounds for canonical codes of
// max length 15 built with remaining-code-space sized subtables: (286 syms, root 9) -> 852,
// (32 syms, root 8) -> 402.  tests/test_inflate_core.py re-derives both bounds by dynamic programming.
constexpr int LIT_ROOT = 9;
constexpr int DIST_ROOT = 8;
constexpr int LIT_ENOUGH = 852;
constexpr int DIST_ENOUGH = 402;
constexpr int PRE_ROOT = 7;  // code-length code: 19 symbols, max 7 bits -> 128 entries, no subtables
constexpr int MAX_LENS = 320;  // 288 litlen + 32 dist code lengths
constexpr uint32_t BGZF_MAX_OUT = 65536;  // htslib/bgzf.h:50 BGZF_MAX_BLOCK_SIZE

// ---- table entry ----------------------------------------------------------------------------
// bits  0..3  : total code length in bits to drop from the bit buffer (1..15)
// bits  4..7  : number of extra bits (length/distance symbols) | subtable index width (KIND_SUB)
// bits  8..10 : kind
// bits 16..31 : payload — literal byte | length base | distance base | subtable start index
enum : uint32_t { KIND_LIT = 0, KIND_BASE = 1, KIND_EOB = 2, KIND_SUB = 3, KIND_BAD = 4 };
DH_HD uint32_t mk_entry(uint32_t payload, uint32_t kind, uint32_t extra, uint32_t len) {
    return (payload << 16) | (kind << 8) | (extra << 4) | len;
}
DH_HD uint32_t e_len(uint32_t e) { return e & 15u; }
DH_HD uint32_t e_extra(uint32_t e) { return (e >> 4) & 15u; }
DH_HD uint32_t e_kind(uint32_t e) { return (e >> 8) & 7u; }
DH_HD uint32_t e_payload(uint32_t e) { return e >> 16; }

// error codes written to the per-block status word (0 = OK)
enum : int {
    INF_OK = 0,
    INF_ERR_BTYPE = 1,        // reserved block type 3
    INF_ERR_STORED_LEN = 2,   // LEN != ~NLEN
    INF_ERR_TOO_MANY_SYMS = 3,
    INF_ERR_PRECODE = 4,      // bad code-length code set
    INF_ERR_LENS_REPEAT = 5,  // repeat with no previous length / repeat overflows
    INF_ERR_NO_EOB = 6,       // code length of symbol 256 is zero
    INF_ERR_LITLEN_SET = 7,   // over-subscribed / incomplete litlen code
    INF_ERR_DIST_SET = 8,     // over-subscribed / incomplete distance code
    INF_ERR_BAD_LITLEN = 9,   // invalid literal/length code in data
    INF_ERR_BAD_DIST = 10,    // invalid distance code in data
    INF_ERR_DIST_TOO_FAR = 11,
    INF_ERR_OUT_OVERFLOW = 12,  // more than 64 KiB of output
    INF_ERR_IN_OVERRUN = 13,    // stream runs past the compressed payload (zlib: Z_BUF_ERROR)
    INF_ERR_ISIZE = 14,         // inflated length != ISIZE footer (stricter than the reference)
    INF_ERR_CRC = 15,           // CRC-32 of output != footer (htslib BGZF_ERR_CRC)
    INF_ERR_HEADER = 16,        // bad BGZF header (htslib BGZF_ERR_HEADER); set by host framing
    INF_NEEDS_V1 = 100          // internal: lockstep kernel's compact tables too small; redone by the lane-group kernel
};

// ---- small helpers ---------------------------------------------------------------------------
DH_HD uint32_t bitrev(uint32_t v, int nbits) {
#if defined(__CUDA_ARCH__)
    return __brev(v) >> (32 - nbits);
#else
    uint32_t r = 0;
    for (int i = 0; i < nbits; ++i) r |= ((v >> i) & 1u) << (nbits - 1 - i);
    return r;
#endif
}

// RFC 1951 §3.2.5 length / distance symbol tables, computed arithmetically (no constant-memory reads).
// length symbol s (257..285) -> (base, extra bits)
DH_HD void len_sym(uint32_t s, uint32_t &base, uint32_t &extra) {
    uint32_t i = s - 257;
    if (i < 8) { base = 3 + i; extra = 0; }
    else if (i == 28) { base = 258; extra = 0; }
    else { extra = (i >> 2) - 1; base = 3 + ((4 + (i & 3)) << extra); }
}
// distance symbol s (0..29) -> (base, extra bits)
DH_HD void dist_sym(uint32_t s, uint32_t &base, uint32_t &extra) {
    if (s < 4) { base = 1 + s; extra = 0; }
    else { extra = (s >> 1) - 1; base = 1 + ((2 + (s & 1)) << extra); }
}

// ---- canonical Huffman table build -------------------------------------------------------------
// lens[0..n): code length per symbol (0 = unused).  table: root table of 2^root entries followed by
// subtables.  sorted: scratch for n uint16.  mode: 0 = litlen, 1 = distance, 2 = code-length code.
// Returns 0 on success, 1 on over-subscribed / disallowed-incomplete code.
// Serial version (one lane does all the work; every lane of the group calls it with identical
// arguments on the device, where `leader` gates the shared-memory stores).
// Out of line on purpose: five call sites, cold relative to the symbol loop, and it only touches the
// group's shared-memory tables (no decoder registers escape).
DH_HD_NOINLINE int build_table(bool leader, const uint8_t *lens, int n, int root, uint32_t *table, int enough,
                               uint16_t *sorted, int mode) {
    uint32_t count[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) count[i] = 0;
    for (int s = 0; s < n; ++s) {
        uint32_t l = lens[s];
        // count[l]++ without dynamic register indexing
#pragma unroll
        for (int i = 0; i < 16; ++i) count[i] += (l == (uint32_t)i);
    }
    int left = 1, maxlen = 0;
    uint32_t offs[16];
    offs[0] = 0;
    offs[1] = 0;
#pragma unroll
    for (int l = 1; l < 16; ++l) {
        left = (left << 1) - (int)count[l];
        if (left < 0) return 1;  // over-subscribed
        if (count[l]) maxlen = l;
        if (l < 15) offs[l + 1] = offs[l] + count[l];
    }
    const uint32_t nroot = 1u << root;
    if (maxlen == 0) {
        // no codes at all: legal (e.g. a block with only literals has an empty distance code);
        // any use of the table hits a BAD entry.
        if (leader)
            for (uint32_t i = 0; i < nroot; ++i) table[i] = mk_entry(0, KIND_BAD, 0, 1);
        return 0;
    }
    if (left > 0 && (mode == 2 || maxlen != 1)) return 1;  // incomplete set (zlib inftrees.c rule)

    // symbols sorted by (length, symbol)
    if (leader) {
        uint32_t cur[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) cur[i] = offs[i];
        for (int s = 0; s < n; ++s) {
            uint32_t l = lens[s];
            if (l) {
                uint32_t pos = 0;
#pragma unroll
                for (int i = 1; i < 16; ++i)
                    if (l == (uint32_t)i) { pos = cur[i]; cur[i]++; }
                sorted[pos] = (uint16_t)s;
            }
        }
        // pre-fill root with BAD so that incomplete (single 1-bit code) sets reject the unused half
        if (left > 0)
            for (uint32_t i = 0; i < nroot; ++i) table[i] = mk_entry(0, KIND_BAD, 0, 1);
    }
    if (!leader) return 0;  // the group meets again at the caller's io.sync()

    uint32_t code = 0;        // next canonical code (MSB-first value) at the current length
    uint32_t idx = 0;         // index into sorted[]
    uint32_t next_free = nroot;  // next unused entry for subtables
    uint32_t sub_prefix = 0xffffffffu, sub_start = 0, sub_bits = 0;
    for (int len = 1; len <= maxlen; ++len) {
        uint32_t cnt = 0;
#pragma unroll
        for (int i = 1; i < 16; ++i) cnt = (len == i) ? count[i] : cnt;
        for (uint32_t k = 0; k < cnt; ++k, ++idx, ++code) {
            uint32_t sym = sorted[idx];
            uint32_t ent;
            if (mode == 0) {
                if (sym < 256) ent = mk_entry(sym, KIND_LIT, 0, len);
                else if (sym == 256) ent = mk_entry(0, KIND_EOB, 0, len);
                else if (sym < 286) { uint32_t b, x; len_sym(sym, b, x); ent = mk_entry(b, KIND_BASE, x, len); }
                else ent = mk_entry(0, KIND_BAD, 0, len);
            } else if (mode == 1) {
                if (sym < 30) { uint32_t b, x; dist_sym(sym, b, x); ent = mk_entry(b, KIND_BASE, x, len); }
                else ent = mk_entry(0, KIND_BAD, 0, len);
            } else {
                ent = mk_entry(sym, KIND_LIT, 0, len);
            }
            uint32_t rev = bitrev(code, len);  // LSB-first bit order as it appears in the stream
            if (len <= root) {
                for (uint32_t i = rev; i < nroot; i += (1u << len)) table[i] = ent;
            } else {
                uint32_t prefix = rev & (nroot - 1);
                if (prefix != sub_prefix) {
                    // open a new subtable; size it by the remaining code space (so "enough" holds)
                    sub_prefix = prefix;
                    uint32_t curr = len - root;
                    int l2 = 1 << curr;
                    // codes of this length still to place under this prefix start at k
                    l2 -= (int)(cnt - k);
                    while (l2 > 0 && (int)(curr + root) < maxlen) {
                        ++curr;
                        uint32_t c2 = 0;
#pragma unroll
                        for (int i = 1; i < 16; ++i) c2 = ((int)(curr + root) == i) ? count[i] : c2;
                        l2 = (l2 << 1) - (int)c2;
                    }
                    sub_bits = curr;
                    sub_start = next_free;
                    next_free += 1u << sub_bits;
                    if ((int)next_free > enough) {  // cannot happen (see "enough" bounds); never write out of range
                        next_free = sub_start;
                        sub_start = nroot;
                        sub_bits = 0;
                    }
                    table[prefix] = mk_entry(sub_start, KIND_SUB, sub_bits, root);
                }
                uint32_t hi = rev >> root;
                for (uint32_t i = hi; i < (1u << sub_bits); i += (1u << (len - root)))
                    if ((int)(sub_start + i) < enough) table[sub_start + i] = ent;
            }
        }
        code <<= 1;
    }
    return 0;
}

// Order in which code-length-code lengths are stored (RFC 1951 §3.2.7)
DH_HD uint32_t precode_order(uint32_t i) {
    // 16,17,18,0,8,7,9,6,10,5,11,4,12,3,13,2,14,1,15 packed 5 bits each into two 64-bit words
    const uint64_t lo = 16ull | (17ull << 5) | (18ull << 10) | (0ull << 15) | (8ull << 20) | (7ull << 25) |
                        (9ull << 30) | (6ull << 35) | (10ull << 40) | (5ull << 45) | (11ull << 50) | (4ull << 55);
    const uint64_t hi = 12ull | (3ull << 5) | (13ull << 10) | (2ull << 15) | (14ull << 20) | (1ull << 25) |
                        (15ull << 30);
    return (i < 12) ? (uint32_t)((lo >> (5 * i)) & 31) : (uint32_t)((hi >> (5 * (i - 12))) & 31);
}

// ---- per-stream decoder state that lives in (shared) memory -----------------------------------
struct InflateTables {
    uint32_t lit[LIT_ENOUGH];
    uint32_t dist[DIST_ENOUGH];
    uint8_t lens[MAX_LENS];
};

// ---- the decode loop ---------------------------------------------------------------------------
// IO policy requirements:
//   uint32_t load32(uint32_t bytepos)        4 stream bytes (little endian) at payload byte offset
//   void     literal(uint32_t byte)          append one byte
//   bool     match(uint32_t len, uint32_t dist)   append a back-reference; false if dist too far / overflow
//   bool     stored(uint32_t bytepos, uint32_t len) append raw stream bytes; false on overflow
//   uint32_t produced()                      bytes appended so far (literals may still be buffered)
//   bool     leader()                        lane that performs table stores
//   void     sync()                          make table stores visible to the group
template <typename IO>
struct Inflater {
    IO &io;
    InflateTables &t;
    uint64_t bitbuf = 0;
    uint32_t bitcnt = 0;   // valid bits in bitbuf
    uint32_t inpos = 0;    // next payload byte to load
    uint32_t in_len;       // payload length in bytes

    DH_HDM Inflater(IO &io_, InflateTables &t_, uint32_t in_len_) : io(io_), t(t_), in_len(in_len_) {}

    DH_HDM void refill() {  // guarantee >= 32 valid bits (actually 33..64)
        if (bitcnt <= 32) {
            bitbuf |= (uint64_t)io.load32(inpos) << bitcnt;
            inpos += 4;
            bitcnt += 32;
        }
    }
    DH_HDM uint32_t peek(uint32_t n) const { return (uint32_t)bitbuf & ((1u << n) - 1u); }
    DH_HDM void drop(uint32_t n) { bitbuf >>= n; bitcnt -= n; }
    DH_HDM uint32_t take(uint32_t n) { uint32_t v = peek(n); drop(n); return v; }
    // payload bits consumed so far
    DH_HDM uint32_t bits_used() const { return inpos * 8u - bitcnt; }

    DH_HDM int read_dynamic_header() {
        refill();
        uint32_t hlit = take(5) + 257, hdist = take(5) + 1, hclen = take(4) + 4;
        if (hlit > 286 || hdist > 30) return INF_ERR_TOO_MANY_SYMS;
        // code-length code lengths -> t.lens[0..19)
        if (io.leader())
            for (int i = 0; i < 19; ++i) t.lens[i] = 0;
        io.sync();
        for (uint32_t i = 0; i < hclen; ++i) {
            refill();
            uint32_t v = take(3);
            if (io.leader()) t.lens[precode_order(i)] = (uint8_t)v;
        }
        io.sync();
        // the precode table borrows the (not yet built) distance table area; its sort scratch sits behind it
        uint32_t *pre = t.dist;
        uint16_t *scratch = reinterpret_cast<uint16_t *>(t.dist + (1 << PRE_ROOT));
        if (build_table(io.leader(), t.lens, 19, PRE_ROOT, pre, 1 << PRE_ROOT, scratch, 2)) return INF_ERR_PRECODE;
        io.sync();
        // litlen + distance code lengths
        uint32_t n = hlit + hdist, i = 0, prev = 0;
        while (i < n) {
            refill();
            uint32_t e = pre[peek(PRE_ROOT)];
            if (e_kind(e) != KIND_LIT) return INF_ERR_PRECODE;
            drop(e_len(e));
            uint32_t sym = e_payload(e);
            if (sym < 16) {
                if (io.leader()) t.lens[i] = (uint8_t)sym;
                prev = sym;
                ++i;
            } else {
                uint32_t rep, val;
                if (sym == 16) {
                    if (i == 0) return INF_ERR_LENS_REPEAT;
                    rep = 3 + take(2);
                    val = prev;
                } else if (sym == 17) {
                    rep = 3 + take(3);
                    val = 0;
                } else {
                    rep = 11 + take(7);
                    val = 0;
                }
                if (i + rep > n) return INF_ERR_LENS_REPEAT;
                if (io.leader())
                    for (uint32_t k = 0; k < rep; ++k) t.lens[i + k] = (uint8_t)val;
                i += rep;
                prev = val;
            }
        }
        io.sync();
        if (t.lens[256] == 0) return INF_ERR_NO_EOB;
        // distance lengths must move out of the way of nothing: lit build uses t.dist as sort scratch AFTER
        // the precode is dead, so stash the distance lengths at the top of lens[] (288..319) first.
        if (io.leader()) {
            // copy backwards-safe: destination index >= source index because hlit <= 288
            for (int k = (int)hdist - 1; k >= 0; --k) t.lens[288 + k] = t.lens[hlit + k];
            for (uint32_t k = hdist; k < 32; ++k) t.lens[288 + k] = 0;
            for (uint32_t k = hlit; k < 288; ++k) t.lens[k] = 0;
        }
        io.sync();
        if (build_table(io.leader(), t.lens, 288, LIT_ROOT, t.lit, LIT_ENOUGH, reinterpret_cast<uint16_t *>(t.dist), 0))
            return INF_ERR_LITLEN_SET;
        io.sync();
        uint16_t dscratch[32];
        if (build_table(io.leader(), t.lens + 288, 32, DIST_ROOT, t.dist, DIST_ENOUGH, dscratch, 1))
            return INF_ERR_DIST_SET;
        io.sync();
        return INF_OK;
    }

    DH_HDM int build_fixed() {
        if (io.leader()) {
            for (int i = 0; i < 144; ++i) t.lens[i] = 8;
            for (int i = 144; i < 256; ++i) t.lens[i] = 9;
            for (int i = 256; i < 280; ++i) t.lens[i] = 7;
            for (int i = 280; i < 288; ++i) t.lens[i] = 8;
            for (int i = 288; i < 320; ++i) t.lens[i] = 5;
        }
        io.sync();
        if (build_table(io.leader(), t.lens, 288, LIT_ROOT, t.lit, LIT_ENOUGH, reinterpret_cast<uint16_t *>(t.dist), 0))
            return INF_ERR_LITLEN_SET;
        io.sync();
        uint16_t dscratch[32];
        if (build_table(io.leader(), t.lens + 288, 32, DIST_ROOT, t.dist, DIST_ENOUGH, dscratch, 1))
            return INF_ERR_DIST_SET;
        io.sync();
        return INF_OK;
    }

    // decode symbols of one Huffman-coded deflate block until end-of-block
    DH_HDM int decode_block() {
        const uint32_t *lit = t.lit;
        const uint32_t *dist = t.dist;
        const uint32_t in_limit = in_len + 8;
        for (;;) {
            if (inpos > in_limit) return INF_ERR_IN_OVERRUN;  // bounds how far a corrupt stream can read ahead
            refill();  // >= 32 bits: enough for litlen code (15) + length extra (5) ...
            uint32_t e = lit[peek(LIT_ROOT)];
            if (e_kind(e) == KIND_SUB) e = lit[e_payload(e) + (((uint32_t)(bitbuf >> LIT_ROOT)) & ((1u << e_extra(e)) - 1u))];
            drop(e_len(e));
            uint32_t kind = e_kind(e);
            if (kind == KIND_LIT) {
                io.literal(e_payload(e));
                continue;
            }
            if (kind == KIND_BASE) {
                uint32_t len = e_payload(e) + take(e_extra(e));
                refill();  // ... then distance code (15) + distance extra (13)
                uint32_t d = dist[peek(DIST_ROOT)];
                if (e_kind(d) == KIND_SUB)
                    d = dist[e_payload(d) + (((uint32_t)(bitbuf >> DIST_ROOT)) & ((1u << e_extra(d)) - 1u))];
                if (e_kind(d) != KIND_BASE) return INF_ERR_BAD_DIST;
                drop(e_len(d));
                uint32_t dd = e_payload(d) + take(e_extra(d));
                if (!io.match(len, dd)) return (dd > io.produced()) ? INF_ERR_DIST_TOO_FAR : INF_ERR_OUT_OVERFLOW;
                continue;
            }
            if (kind == KIND_EOB) return INF_OK;
            return INF_ERR_BAD_LITLEN;
        }
    }

    // whole stream: returns INF_OK or an error code; output length is io.produced()
    DH_HDM int run() {
        for (;;) {
            refill();
            uint32_t bfinal = take(1), btype = take(2);
            int rc;
            if (btype == 0) {
                drop(bitcnt & 7u);  // to byte boundary
                refill();
                uint32_t len = take(16), nlen = take(16);
                if ((len ^ 0xffffu) != nlen) return INF_ERR_STORED_LEN;
                uint32_t bytepos = bits_used() >> 3;  // bitcnt is a multiple of 8 here
                if (bytepos + len > in_len) return INF_ERR_IN_OVERRUN;
                if (!io.stored(bytepos, len)) return INF_ERR_OUT_OVERFLOW;
                // restart the bit reader after the stored bytes
                inpos = bytepos + len;
                bitbuf = 0;
                bitcnt = 0;
                rc = INF_OK;
            } else if (btype == 1) {
                rc = build_fixed();
                if (rc == INF_OK) rc = decode_block();
            } else if (btype == 2) {
                rc = read_dynamic_header();
                if (rc == INF_OK) rc = decode_block();
            } else {
                return INF_ERR_BTYPE;
            }
            if (rc != INF_OK) return rc;
            if (io.produced() > BGZF_MAX_OUT) return INF_ERR_OUT_OVERFLOW;
            if (bits_used() > in_len * 8u) return INF_ERR_IN_OVERRUN;
            if (bfinal) return INF_OK;
            if (inpos > in_len + 8) return INF_ERR_IN_OVERRUN;
        }
    }
};

}  // namespace dh
