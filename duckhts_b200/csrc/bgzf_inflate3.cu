// bgzf_inflate3.cu — "lockstep" BGZF inflate for sm_100a, third generation.
//
// Replaces bgzf_uncompress / inflate_block / bgzf_decode_func (htslib/bgzf.c:762-804, 808-824, 1373-1384):
// raw DEFLATE inflate of every BGZF block + CRC-32 and ISIZE footer checks.
//
// What ncu said about generation 2 (profiles/r1c_*): phase 1 ran 3 warps per SM at 233 warp-instructions and
// ~1000 cycles per symbol step (one warp per scheduler, every dependent instruction exposed, 13 % of the samples
// waiting on the next compressed word from L2); phase 2 (one THREAD per block) read 8x the bytes it produced
// from DRAM because 300k threads with 64 KB working sets thrash L2.  Generation 3:
//
//   phase 1  bgzf_decode3_kernel — still ONE DEFLATE STREAM PER LANE (Huffman decoding is bit-serial), but
//     * 6 warps per SM: one pool of 589 u16 entries per stream (litlen root 7 + distance root 6 and their sub-tables,
//       lane-interleaved) fills the 227 KB of an SM; the few streams whose tables need more are handed, at the header
//       that overflowed, to 4 retry CTAs of the same grid that run 4 warps on 892-entry pools (see P1<> below);
//     * the header pass (code-length decode, table builds) keeps the lanes converged: nothing returns early, the
//       symbol loops have constant trip counts and end in __syncwarp (r1j: 3-7 of 32 threads per instruction before);
//     * compressed bytes reach each lane through a private 2 x 16-byte shared-memory ring filled by cp.async
//       (one 16-byte chunk in flight while the other is consumed), so the symbol loop never waits on L2/DRAM;
//     * the bit reader is a 64-bit window (w0,w1) + one funnel shift per lookup, no 64-bit shifts;
//     * table entries carry length base/extra and distance mantissa/extra, so no RFC-1951 arithmetic in the loop;
//     * the loop is straight-line (both lookups always issued, literal path predicated, one vote per step); a lane that
//       decodes a literal uses the step's second lookup (the distance lookup of the match lanes) on the NEXT literal/length
//       code and takes it too if it is a literal: two symbols per step on literal-heavy streams (r1h: -6 % BAM, -14 % BCF);
//     * output = a LITERAL STREAM (bytes, 4 per store) + a COMMAND STREAM (one word per match), not final bytes.
//   phase 2  ONE WARP PER BLOCK places the literals, replays the matches and checks CRC-32 / ISIZE against the BGZF footer (CRC in a
//     lane-interleaved Horner form: H = H*x^1024 + w through four 256-entry tables).  bgzf_resolve5_kernel (round 2, the default):
//     lanes own output BYTES -- rows of 32 bytes over a shared-memory window of the block's recent output, one descriptor lookup and
//     one generic load per byte, 16-byte stores; see its own header below.  bgzf_resolve3_kernel (round 1, DH_RESOLVE5=0): lanes own
//     COMMANDS -- warp scans give every lane its destination, ragged byte loops in global memory, matches copied in dependency
//     rounds.  bgzf_resolve4_kernel (DH_RESOLVE4=1): one CTA per block with the whole window in shared memory; measured slowest.
// Streams whose tables do not fit the 892-entry pool either (litlen > 640 or litlen + distance > 892 entries; never
// seen from zlib/libdeflate) stay flagged INF_NEEDS_V1 and are redone by the lane-group kernel (bgzf_inflate.cu).
#include <stdlib.h>

#include <mutex>

#include "bgzf_inflate.cuh"
#include "crc32_core.cuh"
#include "inflate_core.cuh"
#include "bgzf_resolve5.cuh"

namespace dh {
namespace v3 {

#ifndef DH_P1_WARPS
#define DH_P1_WARPS 6
#endif
#ifndef DH_LROOT
#define DH_LROOT 7
#endif
#ifndef DH_DROOT
#define DH_DROOT 6
#endif
constexpr int LROOT = DH_LROOT, DROOT = DH_DROOT;
constexpr unsigned FULL = 0xffffffffu;
// One pool of TAB_N u16 entries per stream: the literal/length table (root 7 + sub-tables) at [0, L), the distance table
// (root 6 + sub-tables) right behind it at [L, L + D).  The second-level lookup of a step is always issued, so small roots cost
// no instructions; they make the tables smaller and the root fill of the header pass shorter.  zlib level-6 BAM blocks measure
// L <= 470 and L + D <= 602 (median 546, p99 578) at roots 7/6, BCF genotype blocks L + D <= 508 (worst cases: L = 660 at root 8,
// D = 400 at root 7).  The kernel is bound by the issue latency of one dependent chain per warp (ncu: 25 % issue slots busy with
// 4 warps per SM), so more warps per SM is the lever and the pool size decides how many fit into the 227 KB of an SM:
//   W = 4: 892 entries (every BAM/BCF stream seen fits), W = 6: 589 entries (~1 % of the BAM blocks do not fit).
// The main CTAs run P1<DH_P1_WARPS>; a lane whose tables do not fit queues its stream (as it stood at that deflate block header),
// and the retry CTAs of the same grid resume it as P1<4> (decode_pass<P1Retry, true> below).  What does not fit there either
// is flagged INF_NEEDS_V1 and redone by the lane-group kernel (bgzf_inflate.cu), as before.
// BYTES = the CTA's shared-memory budget.  232448 (227 KB) + the 1 KB the system reserves per CTA is ALL of an SM's 228 KB: while
// such a CTA is resident no other CTA fits on its SM, whatever its size (every CTA costs the reserved 1 KB) -- the column kernels
// and phase 2 of the previous segment wait.  The "small" variant (DH_P1_SMALL_BYTES, default 222 KB: 576 entries, the p99 of the
// BAM blocks is 578) leaves room for five CTAs without shared memory of their own next to it (DH_P1_SMALL=1, see launch below).
template <int W, int BYTES = 232448>
struct P1 {
    static constexpr int WARPS = W, THREADS = W * 32;
    static constexpr int TAB_N = (BYTES / THREADS - 32) / 2;
    static constexpr int SCR_N = 96;                 // build scratch at the end of the pool (free again before the distance table is built)
    static constexpr int SCR_LENS = TAB_N - SCR_N;   // litlen + distance code lengths, FOUR per entry (<= 316 lengths = 79 entries)
    static constexpr int SCR_NC = TAB_N - 16;        // next_code[16]
    // the literal/length table may not reach the scratch while it is being built; link entries carry a 9-bit sub-table offset
    static constexpr int LIT_CAP = TAB_N - SCR_N < (1 << LROOT) + 512 ? TAB_N - SCR_N : (1 << LROOT) + 512;
    // per warp: tables [TAB_N][32] u16, then the input ring [2 chunks][32 lanes][16 bytes]
    static constexpr int WARP_SMEM = TAB_N * 64 + 2 * 512;
    static_assert(WARP_SMEM * WARPS <= BYTES && BYTES <= 232448, "phase 1 is sized to fill (most of) the 227 KB of one SM");
};
#ifndef DH_TAIL_TAKE
#define DH_TAIL_TAKE 1
#endif
#ifndef DH_CONVERT
#define DH_CONVERT 1
#endif
#ifndef DH_RETRY_CTAS
#define DH_RETRY_CTAS 4
#endif
#ifndef DH_P1_SMALL_BYTES
#define DH_P1_SMALL_BYTES (222 * 1024)
#endif
using P1Main = P1<DH_P1_WARPS>;
using P1Retry = P1<4>;
using P1MainS = P1<DH_P1_WARPS, DH_P1_SMALL_BYTES>;
using P1RetryS = P1<4, DH_P1_SMALL_BYTES>;
template <class M, class R>
constexpr int p1_smem() { return M::WARP_SMEM * M::WARPS > R::WARP_SMEM * R::WARPS ? M::WARP_SMEM * M::WARPS : R::WARP_SMEM * R::WARPS; }

// ---- table entries (u16) ---------------------------------------------------------------------------------
// litlen:  bit15 = 1: length code   [14:12] extra bits  [11:4] base-3  [3:0] code length
//          bit15 = 0: [14:13] 0 literal / 1 end-of-block / 2 link / 3 invalid   [12:4] value  [3:0] code length
//          (link: value = sub-table offset behind the root, code-length field = sub-table index bits)
// dist:    [15:14] 0 distance code / 2 link / 3 invalid
//          code: [13:10] extra bits  [9:8] mantissa m (distance = 1 + (m << extra) + extra value)  [3:0] code length
//          link: [12:4] sub-table offset behind the root  [3:0] sub-table index bits
__device__ __forceinline__ uint32_t lit_entry(uint32_t sym, uint32_t l) {
    if (sym < 256u) return (sym << 4) | l;
    if (sym == 256u) return (1u << 13) | l;
    if (sym < 286u) { uint32_t b, x; len_sym(sym, b, x); return 0x8000u | (x << 12) | ((b - 3u) << 4) | l; }
    return (3u << 13) | l;
}
__device__ __forceinline__ uint32_t dist_entry(uint32_t sym, uint32_t l) {
    if (sym >= 30u) return (3u << 14) | l;
    uint32_t x = sym < 4u ? 0u : (sym >> 1) - 1u, m = sym < 4u ? sym : 2u + (sym & 1u);
    return (x << 10) | (m << 8) | l;
}
constexpr uint32_t LIT_BAD = (3u << 13) | 1u, DIST_BAD = (3u << 14) | 1u;

// ---- shared-memory access by 32-bit shared address ----------------------------------------------------------
__device__ __forceinline__ uint32_t lds16(uint32_t a) { uint16_t v; asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a)); return v; }
__device__ __forceinline__ void sts16(uint32_t a, uint32_t v) { asm volatile("st.shared.u16 [%0], %1;" ::"r"(a), "h"((uint16_t)v)); }
__device__ __forceinline__ uint32_t lds32(uint32_t a) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ void cp_async16(uint32_t saddr, const void *g) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(saddr), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait1() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait0() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// ---- bit reader: 64-bit window (w0,w1) over the stream, w2 = next word, fed from the lane's cp.async ring -------
struct Bits {
    uint32_t w0, w1, w2;
    uint32_t bitpos;      // bits of w0 already consumed; < 32 after normalize()
    uint32_t wi;          // stream word (relative to the 16-byte aligned base) that the next advance() loads into w2
    uint32_t ring;        // shared address of this lane's 16 bytes in ring chunk 0 (chunk 1 is 512 bytes further)
    const uint8_t *g;     // 16-byte aligned global address of stream word 0

    __device__ __forceinline__ void advance() {
        w0 = w1;
        w1 = w2;
        bitpos -= 32u;
        const uint32_t c = wi >> 2, k = wi & 3u;
        if (k == 0u) cp_async_wait1();                       // chunk c has landed (chunk c+1 may still be in flight)
        w2 = lds32(ring + ((c & 1u) << 9) + (k << 2));
        if (k == 3u) { cp_async16(ring + ((c & 1u) << 9), g + ((size_t)(c + 2u) << 4)); cp_async_commit(); }
        ++wi;
    }
    // start reading `bit` bits behind the 16-byte aligned address g0 (bit < 128: the start of a payload; more: a resumed stream)
    __device__ __forceinline__ void open(const uint8_t *g0, uint32_t bit, uint32_t ring_addr) {
        ring = ring_addr;
        g = g0 + ((size_t)(bit >> 7) << 4);
        cp_async_wait0();                                    // nothing of the previous stream may still be landing
        cp_async16(ring, g);
        cp_async_commit();
        cp_async16(ring + 512u, g + 16);
        cp_async_commit();
        cp_async_wait1();
        w0 = lds32(ring);
        w1 = lds32(ring + 4u);
        w2 = lds32(ring + 8u);
        wi = 3u;
        bitpos = bit & 127u;
        while (bitpos >= 32u) advance();
    }
    // Branch-free form of `if (bitpos >= 32) advance()` for the symbol loop.  ncu (r1h) showed the two refill blocks of a step
    // as 27 % of its instructions with 5-9 of 32 lanes active: some lane crosses a word in almost every step, so the warp
    // paid for the divergent block anyway.  Here every lane loads its next ring word and selects; the refill of the ring
    // (one 16-byte cp.async per four words) and its commit are predicated instructions, and wait_group 1 runs every time
    // (it only ever has to wait for the chunk that is about to be read).
    __device__ __forceinline__ void normalize() {
        const bool adv = bitpos >= 32u;
        const uint32_t c = wi >> 2, k = wi & 3u;
        const uint32_t slot = ring + ((c & 1u) << 9);
        cp_async_wait1();
        const uint32_t nw = lds32(slot + (k << 2));
        const uint32_t issue = (adv && k == 3u) ? 1u : 0u;
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %0, 0;\n\t@p cp.async.cg.shared.global [%1], [%2], 16;\n\t@p cp.async.commit_group;\n\t}"
                     ::"r"(issue), "r"(slot), "l"(g + ((size_t)(c + 2u) << 4)) : "memory");
        w0 = adv ? w1 : w0;
        w1 = adv ? w2 : w1;
        w2 = adv ? nw : w2;
        bitpos -= adv ? 32u : 0u;
        wi += adv ? 1u : 0u;
    }
    __device__ __forceinline__ uint32_t window() const { return __funnelshift_r(w0, w1, bitpos); }  // next 32 bits
    // header-path helpers (n <= 25)
    __device__ __forceinline__ uint32_t take(uint32_t n) {
        normalize();
        uint32_t v = window() & ((1u << n) - 1u);
        bitpos += n;
        return v;
    }
    // bits consumed since the 16-byte aligned base
    __device__ __forceinline__ uint32_t consumed() const { return (wi - 3u) * 32u + bitpos; }
};

// per-lane decoder state
struct Lane {
    Bits in;
    uint32_t tab;         // shared address of this lane's column: entry i at tab + i * 64
    uint32_t dbase;       // entry index of the distance root table (= size of the literal/length table)
    uint8_t *lit;         // literal stream (grows upwards, 4 bytes per store)
    uint32_t *cmd_top;    // command k is stored at cmd_top[-k]
    uint32_t lit_acc, nlit, ncmd, lit_run, pos, isize;
    uint32_t bit_base;    // bits in front of the 16-byte chunk the reader was opened at (0 unless the stream was resumed)
    uint32_t end_bits;    // consumed() value at the end of the payload
    uint32_t wi_limit;    // a corrupt stream may not fetch words beyond this
    int status;
};

__device__ __forceinline__ uint32_t tget(const Lane &s, uint32_t i) { return lds16(s.tab + (i << 6)); }
__device__ __forceinline__ void tput(const Lane &s, uint32_t i, uint32_t v) { sts16(s.tab + (i << 6), v); }

// cold-path literal (stored blocks): same stream format as the hot loop
__device__ __forceinline__ void put_literal(Lane &s, uint32_t byte) {
    s.lit_acc |= byte << (8u * (s.nlit & 3u));
    ++s.nlit;
    if ((s.nlit & 3u) == 0u) { *reinterpret_cast<uint32_t *>(s.lit + s.nlit - 4u) = s.lit_acc; s.lit_acc = 0u; }
    ++s.pos;
    if (++s.lit_run == 255u) { s.cmd_top[-(int)s.ncmd] = 0x80000000u | 255u; ++s.ncmd; s.lit_run = 0u; }
}

// ---- canonical Huffman table build, one lane's private column -------------------------------------------------
// counts of code lengths 0..15 in 16-bit fields of four 64-bit registers (no dynamically indexed local array)
struct Counts {
    uint64_t c0 = 0, c1 = 0, c2 = 0, c3 = 0;
    __device__ __forceinline__ void add(uint32_t l) {
        const uint64_t one = 1ull << ((l & 3u) * 16u);
        const uint32_t q = l >> 2;
        c0 += q == 0 ? one : 0ull;
        c1 += q == 1 ? one : 0ull;
        c2 += q == 2 ? one : 0ull;
        c3 += q == 3 ? one : 0ull;
    }
    __device__ __forceinline__ void addv(uint32_t l, uint32_t v) {
        const uint64_t val = (uint64_t)v << ((l & 3u) * 16u);
        const uint32_t q = l >> 2;
        c0 += q == 0 ? val : 0ull;
        c1 += q == 1 ? val : 0ull;
        c2 += q == 2 ? val : 0ull;
        c3 += q == 3 ? val : 0ull;
    }
    __device__ __forceinline__ uint32_t get(int l) const {
        const int q = l >> 2;
        const uint64_t w = q == 0 ? c0 : q == 1 ? c1 : q == 2 ? c2 : c3;
        return (uint32_t)(w >> ((l & 3) * 16)) & 0xffffu;
    }
};

// Kraft check shared by both builds: returns INF_OK / err; sets maxlen and left (unused code space at length 15)
__device__ __forceinline__ int check_lengths(const Counts &c, int err, int &maxlen, int &left) {
    left = 1;
    maxlen = 0;
#pragma unroll
    for (int l = 1; l <= 15; ++l) {
        const uint32_t cl = c.get(l);
        left = (left << 1) - (int)cl;
        if (left < 0) return err;                   // over-subscribed
        if (cl) maxlen = l;
    }
    if (left > 0 && maxlen > 1) return err;         // incomplete set: zlib allows it only for a single 1-bit code
    return INF_OK;
}

// Root prefixes that carry codes longer than `root` bits get a sub-table sized by the longest code under them.
// Canonical codes grow with length, so these prefixes are the top end [p_first, nroot) of the prefix space.
// mark(p, bits) / finish(p) -> stores; shared by both builds through lambdas on the code ranges.
template <typename FirstFn>
__device__ __forceinline__ int layout_subtables(const Lane &s, const Counts &c, const int base, const int root, const int cap,
                                                const int maxlen, const uint32_t link_kind, int &used, FirstFn first_code) {
    const int nroot = 1 << root;
    int p_first = nroot;
    for (int l = root + 1; l <= maxlen; ++l) {
        const uint32_t cl = c.get(l);
        if (!cl) continue;
        const uint32_t first = first_code(l);
        const int p0 = (int)(first >> (l - root)), p1 = (int)((first + cl - 1u) >> (l - root));
        if (p0 < p_first) p_first = p0;
        for (int p = p0; p <= p1; ++p) tput(s, base + (__brev((uint32_t)p) >> (32 - root)), (uint32_t)(l - root));
    }
    int next_free = nroot;
    for (int p = p_first; p < nroot; ++p) {
        const uint32_t ri = base + (__brev((uint32_t)p) >> (32 - root));
        const uint32_t bits = tget(s, ri);
        const int off = next_free;
        next_free += 1 << bits;
        if (next_free > cap) return INF_NEEDS_V1;
        tput(s, ri, link_kind | ((uint32_t)(off - nroot) << 4) | bits);
    }
    used = next_free;
    return INF_OK;
}

// store entry `ent` for the code `cd` (MSB-first value) of length l
__device__ __forceinline__ void place_code(const Lane &s, const int base, const int root, const uint32_t cd, const uint32_t l,
                                           const uint32_t ent) {
    const uint32_t rev = __brev(cd) >> (32u - l);
    const uint32_t nroot = 1u << root;
    if ((int)l <= root) {
        for (uint32_t k = rev; k < nroot; k += 1u << l) tput(s, base + k, ent);
    } else {
        const uint32_t link = tget(s, base + (rev & (nroot - 1u)));
        const uint32_t sub = base + nroot + ((link >> 4) & 511u), bits = link & 15u;
        for (uint32_t k = rev >> root; k < (1u << bits); k += 1u << (l - (uint32_t)root)) tput(s, sub + k, ent);
    }
}

// ---- table builds: every lane of `hm` (the lanes that are at a deflate block header) goes through them together ----
// ncu source page (profiles/r1j): the header code ran with 3-7 of 32 threads per instruction.  The lanes enter it together but
// drift apart in the code-length loop and, with error returns inside the loops, the compiler only reconverged them at the end
// of the header.  Now nothing returns early: a lane that failed (or holds a stored block) carries n = 0 through the builds,
// the symbol loops have constant trip counts, and __syncwarp(hm) closes every iteration, so the replication loops of
// place_code run side by side.
template <class K>
__device__ __forceinline__ uint32_t parked_len(const Lane &s, uint32_t j) { return (tget(s, K::SCR_LENS + (j >> 2)) >> (4u * (j & 3u))) & 15u; }

// literal/length table from the n code lengths parked (four per entry) at K::SCR_LENS; `used` = entries it occupies
template <class K>
__device__ __forceinline__ int build_litlen(const Lane &s, const unsigned hm, const int n, int &used) {
    Counts c;
    for (int j = 0; j < 288; j += 4) {
        const uint32_t w = tget(s, K::SCR_LENS + (j >> 2));
#pragma unroll
        for (int q = 0; q < 4; ++q) c.add(j + q < n ? (w >> (4 * q)) & 15u : 0u);
    }
    c.c0 &= ~0xffffull;  // length 0 = unused symbol
    int maxlen, left;
    int rc = check_lengths(c, INF_ERR_LITLEN_SET, maxlen, left);
    used = 1 << LROOT;
    bool ok = rc == INF_OK && n > 0;
    if (ok) {
        uint32_t code = 0;
#pragma unroll
        for (int l = 1; l <= 15; ++l) {                 // next_code[l]: first canonical code of length l
            tput(s, K::SCR_NC + l, code);
            code = (code + c.get(l)) << 1;
        }
        if (left > 0 || maxlen == 0)
            for (int k = 0; k < (1 << LROOT); ++k) tput(s, k, LIT_BAD);
        if (maxlen > LROOT) {
            rc = layout_subtables(s, c, 0, LROOT, K::LIT_CAP, maxlen, 2u << 13, used, [&](int l) -> uint32_t { return tget(s, K::SCR_NC + l); });
            ok = rc == INF_OK;
        }
    }
    __syncwarp(hm);
    const int m = ok ? n : 0;
    for (int j = 0; j < 288; j += 4) {
        const uint32_t w = tget(s, K::SCR_LENS + (j >> 2));
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const uint32_t l = j + q < m ? (w >> (4 * q)) & 15u : 0u;
            if (l) {
                const uint32_t cd = tget(s, K::SCR_NC + l);
                tput(s, K::SCR_NC + l, cd + 1u);
                place_code(s, 0, LROOT, cd, l, lit_entry((uint32_t)(j + q), l));
            }
            __syncwarp(hm);
        }
    }
    return rc;
}

// distance table at [base, K::TAB_N) from n <= 32 code lengths held as nibbles in four registers; needs no scratch
template <class K>
__device__ __forceinline__ int build_dist(const Lane &s, const unsigned hm, const int base, const int n, const uint32_t d0,
                                       const uint32_t d1, const uint32_t d2, const uint32_t d3) {
    auto nib = [&](int j) -> uint32_t { const uint32_t w = j < 8 ? d0 : j < 16 ? d1 : j < 24 ? d2 : d3; return (w >> (4 * (j & 7))) & 15u; };
    int rc = (n > 0 && base + (1 << DROOT) > K::TAB_N) ? INF_NEEDS_V1 : INF_OK;
    Counts c;
#pragma unroll
    for (int j = 0; j < 32; ++j) c.add(j < n ? nib(j) : 0u);
    c.c0 &= ~0xffffull;
    int maxlen = 0, left = 0;
    if (rc == INF_OK) rc = check_lengths(c, INF_ERR_DIST_SET, maxlen, left);
    bool ok = rc == INF_OK && n > 0;
    Counts nc;                                          // next canonical code of every length, 16-bit fields like the counts
    if (ok) {
        if (left > 0 || maxlen == 0)
            for (int k = 0; k < (1 << DROOT); ++k) tput(s, base + k, DIST_BAD);
        if (maxlen > DROOT) {
            int used;
            const int room = K::TAB_N - base, cap = room < (1 << DROOT) + 512 ? room : (1 << DROOT) + 512;
            rc = layout_subtables(s, c, base, DROOT, cap, maxlen, 2u << 14, used, [&](int l) -> uint32_t {
                uint32_t code = 0;
                for (int q = 1; q < l; ++q) code = (code + c.get(q)) << 1;
                return code;
            });
            ok = rc == INF_OK;
        }
        uint32_t code = 0;
#pragma unroll
        for (int l = 1; l <= 15; ++l) {
            nc.addv((uint32_t)l, code);
            code = (code + c.get(l)) << 1;
        }
    }
    __syncwarp(hm);
    const int m = ok ? n : 0;
#pragma unroll 4
    for (int j = 0; j < 32; ++j) {
        const uint32_t l = j < m ? nib(j) : 0u;
        if (l) {
            const uint32_t cd = nc.get((int)l);
            nc.add(l);
            place_code(s, base, DROOT, cd, l, dist_entry((uint32_t)j, l));
        }
        __syncwarp(hm);
    }
    return rc;
}

// ---- deflate block header (RFC 1951 3.2.3-3.2.7) ------------------------------------------------------------------------
// stored block: LEN/NLEN, then the bytes go through the literal path
__device__ __forceinline__ int copy_stored(Lane &s) {
    s.in.bitpos = (s.in.bitpos + 7u) & ~7u;            // LEN/NLEN start at the next byte boundary
    const uint32_t len = s.in.take(16), nlen = s.in.take(16);
    if ((len ^ 0xffffu) != nlen) return INF_ERR_STORED_LEN;
    if (s.pos + len > s.isize) return INF_ERR_ISIZE;
    if (s.in.consumed() + 8u * len > s.end_bits) return INF_ERR_IN_OVERRUN;
    for (uint32_t i = 0; i < len; ++i) put_literal(s, s.in.take(8));
    return INF_OK;
}

// dynamic block: HLIT/HDIST/HCLEN, the code-length code, then the litlen + distance code lengths parked at K::SCR_LENS
// (distance lengths also returned as nibbles in d0..d3).  Per-lane control flow; the lanes meet again behind it.
template <class K>
__device__ __forceinline__ int read_code_lengths(Lane &s, uint32_t &hlit, uint32_t &hdist, uint32_t &d0, uint32_t &d1, uint32_t &d2,
                                                 uint32_t &d3) {
    hlit = s.in.take(5) + 257u;
    hdist = s.in.take(5) + 1u;
    const uint32_t hclen = s.in.take(4) + 4u;
    if (hlit > 286u || hdist > 30u) return INF_ERR_TOO_MANY_SYMS;
    // code-length code: 19 x 3 bits, kept in one register
    uint64_t pl = 0;
    for (uint32_t i = 0; i < hclen; ++i) pl |= (uint64_t)s.in.take(3) << (3u * precode_order(i));
    {
        Counts c;
        for (int sym = 0; sym < 19; ++sym) c.add((uint32_t)(pl >> (3 * sym)) & 7u);
        c.c0 &= ~0xffffull;
        int left = 1;
#pragma unroll
        for (int q = 1; q < 8; ++q) { left = (left << 1) - (int)c.get(q); if (left < 0) return INF_ERR_PRECODE; }
        if (left > 0) return INF_ERR_PRECODE;   // an incomplete code-length code is never allowed (zlib CODES rule)
    }
    // 7-bit direct table in litlen-region slots [0,128): entry = sym << 4 | len
    uint32_t code = 0;
    for (uint32_t l = 1; l < 8u; ++l) {
        for (uint32_t sym = 0; sym < 19u; ++sym) {
            if (((uint32_t)(pl >> (3u * sym)) & 7u) != l) continue;
            const uint32_t rev = __brev(code) >> (32u - l);
            for (uint32_t k = rev; k < 128u; k += 1u << l) tput(s, k, (sym << 4) | l);
            ++code;
        }
        code <<= 1;
    }
    // litlen + distance code lengths, parked four per entry
    const uint32_t n = hlit + hdist;
    uint32_t i = 0, prev = 0, pend = 0;
    auto put_len = [&](uint32_t idx, uint32_t v) {
        pend |= v << (4u * (idx & 3u));
        if ((idx & 3u) == 3u) { tput(s, K::SCR_LENS + (idx >> 2), pend); pend = 0u; }
    };
    while (i < n) {
        s.in.normalize();
        const uint32_t e = tget(s, s.in.window() & 127u);
        s.in.bitpos += e & 15u;
        const uint32_t sym = e >> 4;
        if (sym < 16u) { put_len(i, sym); prev = sym; ++i; continue; }
        uint32_t rep, val;
        if (sym == 16u) { if (i == 0u) return INF_ERR_LENS_REPEAT; rep = 3u + s.in.take(2); val = prev; }
        else if (sym == 17u) { rep = 3u + s.in.take(3); val = 0u; }
        else { rep = 11u + s.in.take(7); val = 0u; }
        if (i + rep > n) return INF_ERR_LENS_REPEAT;
        for (uint32_t k = 0; k < rep; ++k) put_len(i + k, val);
        i += rep;
        prev = val;
    }
    if (n & 3u) tput(s, K::SCR_LENS + (n >> 2), pend);
    if (s.in.consumed() > s.end_bits) return INF_ERR_IN_OVERRUN;
    if (parked_len<K>(s, 256u) == 0u) return INF_ERR_NO_EOB;
    d0 = d1 = d2 = d3 = 0u;
    for (uint32_t k = 0; k < hdist; ++k) {
        const uint32_t v = parked_len<K>(s, hlit + k) << (4u * (k & 7u)), q = k >> 3;
        d0 |= q == 0u ? v : 0u; d1 |= q == 1u ? v : 0u; d2 |= q == 2u ? v : 0u; d3 |= q == 3u ? v : 0u;
    }
    return INF_OK;
}

// returns INF_OK with tables built (or, *stored = true, with a stored block copied), or an error / INF_NEEDS_V1.
// Called by all lanes of `hm` together; nothing in it returns before the last line.
template <class K>
__device__ __forceinline__ int read_block_header(Lane &s, const unsigned hm, uint32_t &bfinal, bool &stored) {
    int rc = INF_OK;
    stored = false;
    bfinal = s.in.take(1);
    const uint32_t btype = s.in.take(2);
    uint32_t hlit = 288u, hdist = 32u;
    uint32_t d0 = 0x55555555u, d1 = 0x55555555u, d2 = 0x55555555u, d3 = 0x55555555u;  // fixed code: 32 distance codes of 5 bits
    if (btype == 3u) rc = INF_ERR_BTYPE;
    else if (btype == 0u) { stored = true; rc = copy_stored(s); }
    else if (btype == 1u) {
        // fixed Huffman code (RFC 1951 3.2.6): lengths parked where the dynamic path would put them
        for (uint32_t j = 0; j < 288u; j += 4u) {
            const uint32_t l = j < 144u ? 8u : j < 256u ? 9u : j < 280u ? 7u : 8u;
            tput(s, K::SCR_LENS + (j >> 2), l * 0x1111u);
        }
    } else rc = read_code_lengths<K>(s, hlit, hdist, d0, d1, d2, d3);
    __syncwarp(hm);
    const bool want = rc == INF_OK && !stored;
    int lit_used;
    const int rl = build_litlen<K>(s, hm, want ? (int)hlit : 0, lit_used);
    if (want && rl != INF_OK) rc = rl;
    const bool want_d = want && rl == INF_OK;
    s.dbase = (uint32_t)lit_used;
    const int rd = build_dist<K>(s, hm, lit_used, want_d ? (int)hdist : 0, d0, d1, d2, d3);
    if (want_d && rd != INF_OK) rc = rd;
    return rc;
}

}  // namespace v3

using namespace v3;

// =========================================================================================================
// phase 1: Huffman decode, one stream per lane -> literal stream + command stream
// =========================================================================================================
// commands (32 bits):  [31]=0 match:   [30:23] literals since the previous command (0..254)  [22:15] length-3  [14:0] distance-1
//                      [31]=1 literals: [30:0] run of literals with no match behind it (255 inside a run, or the tail)
// The retry queue: `retry` = {count, ticket, finished main warps, -} + 8 words per entry {block, bit position of the deflate block
// header that did not fit, nlit, ncmd, lit_run, pos, lit_acc, -}; word 0 of every entry is 0xffffffff until the entry is complete.
// The LAST `retry_ctas` CTAs of the grid run decode_pass<P1Retry, true> (4 warps, the larger pool) next to the main CTAs: they take
// queued streams in groups of 32 as they arrive and RESUME them at the header that overflowed, so almost all of the retry work
// hides behind the main pass (a separate retry launch cost one generation, ~6.5 ms, per launch).  Every wait is bounded: a retry
// warp that gives up leaves its blocks flagged INF_NEEDS_V1, and the lane-group kernel redoes them.
__device__ __forceinline__ uint32_t ld_vol(const unsigned int *p) { return *reinterpret_cast<const volatile unsigned int *>(p); }
constexpr uint32_t RETRY_HDR = 4, RETRY_ENTRY = 8, SPIN_LIMIT = 1u << 21;

template <class K, bool RETRY>
__device__ __forceinline__ void decode_pass(const uint8_t *__restrict__ comp, const BlockDesc *__restrict__ blocks, const uint32_t nblocks,
                                            int *__restrict__ status, uint32_t *__restrict__ ncmd_out, uint8_t *__restrict__ work,
                                            unsigned int *__restrict__ counter, unsigned int *retry, const uint32_t n_main_warps,
                                            const bool linger = true) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t warp_base = (uint32_t)__cvta_generic_to_shared(smem_raw) + warp * (uint32_t)K::WARP_SMEM;
    enum { ST_HEADER = 0, ST_RUN = 1, ST_DONE = 2 };

    for (;;) {
        // ---- a generation: 32 consecutive blocks, one per lane ----
        uint32_t base = 0, b = nblocks, start_bit = 0xffffffffu;
        Lane s;
        s.lit_acc = s.nlit = s.ncmd = s.lit_run = s.pos = 0u;
        if (!RETRY) {
            if (lane == 0) base = atomicAdd(counter, 32u);
            base = __shfl_sync(FULL, base, 0);
            if (base >= nblocks) break;
            b = base + lane;
        } else {
            uint32_t take = 0;
            if (lane == 0) {
                for (uint32_t spins = 0; spins < SPIN_LIMIT; ++spins) {
                    const bool all_in = ld_vol(retry + 2) == n_main_warps;      // read before the count: all_in => the count is final
                    __threadfence();
                    const uint32_t tk = ld_vol(retry + 1), avail = ld_vol(retry) - tk;
                    // full groups while the main pass is in full swing; once its last generations have been handed out (and the main
                    // CTAs that finish turn into retry CTAs, see the kernel), whatever is there, so that no stream waits for company
                    const bool last_round = DH_TAIL_TAKE && ld_vol(counter) >= nblocks;
                    if (avail >= 32u || ((all_in || last_round) && avail)) {
                        const uint32_t want = avail < 32u ? avail : 32u;
                        if (atomicCAS(retry + 1, tk, tk + want) == tk) { base = tk; take = want; break; }
                        continue;
                    }
                    if (all_in || !linger) break;     // a converted main CTA does not wait for work: it would hold its SM
                    __nanosleep(2000);
                }
            }
            take = __shfl_sync(FULL, take, 0);
            base = __shfl_sync(FULL, base, 0);
            if (!take) break;
            if (lane < take) {
                const unsigned int *q = retry + RETRY_HDR + (size_t)(base + lane) * RETRY_ENTRY;
                uint32_t qb = 0xffffffffu;
                for (uint32_t spins = 0; spins < SPIN_LIMIT && (qb = ld_vol(q)) == 0xffffffffu; ++spins) __nanosleep(500);
                __threadfence();
                if (qb < nblocks) {
                    b = qb;
                    start_bit = ld_vol(q + 1);
                    s.nlit = ld_vol(q + 2); s.ncmd = ld_vol(q + 3); s.lit_run = ld_vol(q + 4); s.pos = ld_vol(q + 5); s.lit_acc = ld_vol(q + 6);
                }
            }
        }
        bool queued = false;
        s.tab = warp_base + lane * 2u;
        s.dbase = 1u << LROOT;
        s.status = INF_OK;
        s.isize = 0u;
        s.end_bits = 0u;
        s.wi_limit = 0u;
        s.lit = work;
        s.cmd_top = reinterpret_cast<uint32_t *>(work);
        int st = ST_DONE;
        uint32_t bfinal = 0;
        if (b < nblocks) {
            const BlockDesc d = blocks[b];
            s.isize = d.isize;
            s.lit = work + d.cmd_off;
            s.cmd_top = reinterpret_cast<uint32_t *>(work + d.cmd_off + inflate_work_cap(d.isize)) - 1;
            const uint32_t mis = (uint32_t)((uintptr_t)(comp + d.in_off) & 15u);
            if (!RETRY) start_bit = mis * 8u;
            s.in.open(comp + d.in_off - mis, start_bit, warp_base + (uint32_t)K::TAB_N * 64u + lane * 16u);
            s.bit_base = start_bit & ~127u;                  // consumed() counts from the 16-byte chunk the reader was opened at
            s.end_bits = (mis + d.in_len) * 8u - s.bit_base;
            s.wi_limit = ((mis + d.in_len + 3u) >> 2) + 8u - (s.bit_base >> 5);
            st = ST_HEADER;
        }
        auto fail = [&](int code) { s.status = code; st = ST_DONE; };
        // end of the final deflate block: flush the literal word, announce the trailing literals
        auto finish = [&]() {
            if (s.nlit & 3u) *reinterpret_cast<uint32_t *>(s.lit + (s.nlit & ~3u)) = s.lit_acc;
            if (s.lit_run) { s.cmd_top[-(int)s.ncmd] = 0x80000000u | s.lit_run; ++s.ncmd; s.lit_run = 0u; }
            if (s.in.consumed() > s.end_bits) s.status = INF_ERR_IN_OVERRUN;
            else if (s.pos != s.isize) s.status = INF_ERR_ISIZE;
            st = ST_DONE;
        };

        while (__any_sync(FULL, st != ST_DONE)) {
            // ---------------- deflate block headers (in lock-step for zlib-made streams) ----------------
            const unsigned hm = __ballot_sync(FULL, st == ST_HEADER);
            if (st == ST_HEADER) {
                bool stored;
                const uint32_t hdr_bit = s.in.consumed() + s.bit_base;
                const int rc = read_block_header<K>(s, hm, bfinal, stored);
                if (!RETRY && rc == INF_NEEDS_V1 && n_main_warps) {
                    // the tables do not fit this pool: hand the stream, as it stood at this header, to the retry CTAs
                    status[b] = INF_NEEDS_V1;
                    unsigned int *q = retry + RETRY_HDR + (size_t)atomicAdd(retry, 1u) * RETRY_ENTRY;
                    q[1] = hdr_bit; q[2] = s.nlit; q[3] = s.ncmd; q[4] = s.lit_run; q[5] = s.pos; q[6] = s.lit_acc;
                    __threadfence();
                    *reinterpret_cast<volatile unsigned int *>(q) = b;
                    queued = true;
                }
                if (rc != INF_OK) fail(rc);
                else if (!stored) st = ST_RUN;
                else if (bfinal) finish();
            }
            // ---------------- symbols: one per lane per step, until every running lane has hit something rare ----------------
            while (__any_sync(FULL, st == ST_RUN)) {
                uint32_t rare;
                uint32_t e = 0;
                do {
                    rare = 0u;
                    if (st == ST_RUN) {
                        s.in.normalize();
                        uint32_t bits = s.in.window();
                        e = tget(s, bits & ((1u << LROOT) - 1u));
                        {   // second level: always issued, so that the step is straight-line code
                            const bool link = (e >> 13) == 2u;
                            const uint32_t sub = (1u << LROOT) + ((e >> 4) & 511u) + ((bits >> LROOT) & ((1u << (e & 15u)) - 1u));
                            const uint32_t e2 = tget(s, link ? sub : 0u);
                            e = link ? e2 : e;
                        }
                        const uint32_t cl = e & 15u;
                        bits >>= cl;
                        s.in.bitpos += cl;
                        const bool is_match = (e >> 15) != 0u, is_lit = (e >> 13) == 0u;
                        if (is_lit) {                               // literal
                            s.lit_acc |= ((e >> 4) & 255u) << (8u * (s.nlit & 3u));
                            ++s.nlit;
                            ++s.lit_run;
                            ++s.pos;
                            if ((s.nlit & 3u) == 0u) { *reinterpret_cast<uint32_t *>(s.lit + s.nlit - 4u) = s.lit_acc; s.lit_acc = 0u; }
                            rare = (s.lit_run == 255u) | (s.pos > s.isize);
                        }
                        // ONE more table lookup for every lane: the distance code of a match, or -- for a lane that has just
                        // taken a literal -- the next literal/length code, consumed only if it is a literal too.  Literal-heavy
                        // streams (base qualities, genotype bytes) then move two symbols per step for the instruction count of one.
                        const uint32_t xb = is_match ? (e >> 12) & 7u : 0u;
                        const uint32_t len = 3u + ((e >> 4) & 255u) + (bits & ((1u << xb) - 1u));
                        s.in.bitpos += xb;
                        s.in.normalize();
                        bits = s.in.window();
                        const uint32_t tb = is_match ? s.dbase : 0u, rb = is_match ? (uint32_t)DROOT : (uint32_t)LROOT;
                        uint32_t d = tget(s, tb + (bits & ((1u << rb) - 1u)));
                        {
                            const bool link = is_match ? (d >> 14) == 2u : (d >> 13) == 2u;
                            const uint32_t sub = tb + (1u << rb) + ((d >> 4) & 511u) + ((bits >> rb) & ((1u << (d & 15u)) - 1u));
                            const uint32_t d2 = tget(s, link ? sub : tb);
                            d = link ? d2 : d;
                        }
                        if (is_match) {                             // length + distance
                            const uint32_t dl = d & 15u, dx = (d >> 10) & 15u;
                            bits >>= dl;
                            const uint32_t dist = 1u + (((d >> 8) & 3u) << dx) + (bits & ((1u << dx) - 1u));
                            s.in.bitpos += dl + dx;
                            s.cmd_top[-(int)s.ncmd] = (s.lit_run << 23) | ((len - 3u) << 15) | (dist - 1u);
                            ++s.ncmd;
                            s.lit_run = 0u;
                            rare = ((d >> 14) != 0u) | (dist > s.pos);
                            s.pos += len;
                            rare |= s.pos > s.isize;
                            e = 0x8000u | (d & 0xc000u) >> 2;       // keep the distance verdict for the slow path
                        } else if (is_lit) {
                            if (!rare && (d >> 13) == 0u) {         // a second literal
                                s.in.bitpos += d & 15u;
                                s.lit_acc |= ((d >> 4) & 255u) << (8u * (s.nlit & 3u));
                                ++s.nlit;
                                ++s.lit_run;
                                ++s.pos;
                                if ((s.nlit & 3u) == 0u) { *reinterpret_cast<uint32_t *>(s.lit + s.nlit - 4u) = s.lit_acc; s.lit_acc = 0u; }
                                rare = (s.lit_run == 255u) | (s.pos > s.isize);
                            }
                        } else {
                            rare = 1u;                              // end of block / invalid code
                        }
                        rare |= s.in.wi > s.wi_limit;
                    }
                } while (!__any_sync(FULL, rare));
                // ---------------- rare events ----------------
                if (rare) {
                    if (s.in.wi > s.wi_limit) fail(INF_ERR_IN_OVERRUN);
                    else if ((e >> 13) == 0u) {                     // literal
                        if (s.pos > s.isize) fail(INF_ERR_ISIZE);
                        else { s.cmd_top[-(int)s.ncmd] = 0x80000000u | 255u; ++s.ncmd; s.lit_run = 0u; }
                    } else if (e >> 15) {                           // match
                        if (e & 0x3000u) fail(INF_ERR_BAD_DIST);
                        else if (s.pos > s.isize) fail(INF_ERR_ISIZE);
                        else fail(INF_ERR_DIST_TOO_FAR);
                    } else if ((e >> 13) == 1u) {                   // end of block
                        if (s.in.consumed() > s.end_bits) fail(INF_ERR_IN_OVERRUN);
                        else if (bfinal) finish();
                        else st = ST_HEADER;
                    } else fail(INF_ERR_BAD_LITLEN);
                }
            }
        }
        if (b < nblocks && !queued) {
            status[b] = s.status;
            ncmd_out[b] = s.ncmd;
        }
    }
    cp_async_wait0();
    if (!RETRY && lane == 0) { __threadfence(); atomicAdd(retry + 2, 1u); }
}

// n_main_warps = 0: no retry CTAs in this grid (every lane keeps its INF_NEEDS_V1 verdict)
template <class P1Main, class P1Retry>
__global__ void __launch_bounds__(P1Main::THREADS, 1)
bgzf_decode3_kernel(const uint8_t *__restrict__ comp, const BlockDesc *__restrict__ blocks, uint32_t nblocks,
                    int *__restrict__ status, uint32_t *__restrict__ ncmd_out, uint8_t *__restrict__ work,
                    unsigned int *__restrict__ counter, unsigned int *retry, uint32_t retry_ctas) {
    static_assert(P1Retry::THREADS <= P1Main::THREADS, "the retry role uses a subset of the CTA's warps");
    const uint32_t main_ctas = gridDim.x - retry_ctas;
    const uint32_t n_main_warps = retry_ctas ? main_ctas * (uint32_t)P1Main::WARPS : 0u;
    if (blockIdx.x < main_ctas) {
        decode_pass<P1Main, false>(comp, blocks, nblocks, status, ncmd_out, work, counter, retry, n_main_warps);
        if (!retry_ctas || !DH_CONVERT) return;
        __syncthreads();          // all warps of this CTA are out of the main pass: its shared memory can be laid out for the retry role
    }
    // retry role: the dedicated CTAs from the start and until the queue is final and empty; every main CTA once it has run out of
    // main work, for as long as it finds queued streams (the backlog would otherwise keep 16 warps busy for most of a generation
    // while 140 SMs idle: r1k capture, 11.4 ms instead of 7.1).  A converted CTA that finds the queue empty exits: inside the
    // scan pipeline the column kernels of the other stream are waiting for its SM (lingering there cost 25 % of the scan rate).
    if (threadIdx.x < P1Retry::THREADS)
        decode_pass<P1Retry, true>(comp, blocks, nblocks, status, ncmd_out, work, counter, retry, n_main_warps, blockIdx.x >= main_ctas);
}

// =========================================================================================================
// phase 2: one warp per block — place literals, replay matches in dependency rounds, CRC-32 + footer check
// =========================================================================================================
// CRC tables (host-built, uploaded once per device): [k*256 + i], k = 0..3: byte i at byte position k of a 32-bit
// register advanced by 128 bytes (multiplication by x^1024 mod P, split by byte); k = 4: the classic byte table;
// [1280 + l]: x^(8 * (128 - 4l)) for lane l.
// (CRC3_WORDS and build_crc3_tables: bgzf_resolve5.cuh, shared with the host-side tests)
__device__ uint32_t g_crc3[CRC3_WORDS];

#ifndef DH_LONG_COPY
#define DH_LONG_COPY 16
#endif
constexpr uint32_t LONG_COPY = DH_LONG_COPY;   // copies of at least this many bytes are done by the whole warp

// SMEM_TABS = false: the CRC tables are read through the read-only path (L1) instead of a shared-memory copy, so that the CTA needs no
// shared memory of its own and fits next to a (small-variant) phase-1 CTA of the next segment.
template <bool SMEM_TABS>
__global__ void __launch_bounds__(256, 4)
bgzf_resolve3_kernel(const BlockDesc *__restrict__ blocks, uint32_t nblocks, uint8_t *__restrict__ outbuf,
                     int *__restrict__ status, const uint32_t *__restrict__ ncmd_in, const uint8_t *__restrict__ work) {
    __shared__ uint32_t tabs_s[SMEM_TABS ? CRC3_WORDS : 1];
    if (SMEM_TABS) {
        for (int i = threadIdx.x; i < CRC3_WORDS; i += 256) tabs_s[i] = g_crc3[i];
        __syncthreads();
    }
    auto tab = [&](uint32_t i) -> uint32_t { return SMEM_TABS ? tabs_s[i] : __ldg(g_crc3 + i); };
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t b = blockIdx.x * 8u + (threadIdx.x >> 5);
    if (b >= nblocks) return;
    if (status[b] != INF_OK) return;
    const BlockDesc d = blocks[b];
    uint8_t *o = outbuf + d.out_off;
    const uint8_t *lit = work + d.cmd_off;
    const uint32_t *cmd_top = reinterpret_cast<const uint32_t *>(work + d.cmd_off + inflate_work_cap(d.isize)) - 1;
    const uint32_t n = ncmd_in[b];
    uint32_t pos = 0, lpos = 0;

    for (uint32_t k0 = 0; k0 < n; k0 += 32u) {
        const uint32_t k = k0 + lane;
        const uint32_t c = k < n ? cmd_top[-(int)k] : 0x80000000u;
        const bool isrun = (c >> 31) != 0u;
        const uint32_t lr = isrun ? (c & 0x7fffffffu) : (c >> 23);
        const uint32_t len = isrun ? 0u : ((c >> 15) & 255u) + 3u;
        const uint32_t dist = (c & 0x7fffu) + 1u;
        // inclusive warp scans: output bytes and literal bytes up to and including this command
        uint32_t ae = lr + len, le = lr;
#pragma unroll
        for (int sft = 1; sft < 32; sft <<= 1) {
            const uint32_t va = __shfl_up_sync(FULL, ae, sft), vl = __shfl_up_sync(FULL, le, sft);
            if (lane >= (uint32_t)sft) { ae += va; le += vl; }
        }
        const uint32_t dl = pos + ae - lr - len;     // where this command's literals go
        const uint32_t dm = dl + lr;                 // where its match goes
        const uint32_t ls = lpos + le - lr;          // its literals in the literal stream
        // ---- literals ----
        {
            const uint32_t small = lr < LONG_COPY ? lr : 0u;
            const uint32_t mx = __reduce_max_sync(FULL, small);
            for (uint32_t i = 0; i < mx; ++i)
                if (i < small) o[dl + i] = lit[ls + i];
            uint32_t big = __ballot_sync(FULL, lr >= LONG_COPY);
            while (big) {
                const int L = __ffs(big) - 1;
                big &= big - 1u;
                const uint32_t d_ = __shfl_sync(FULL, dl, L), s_ = __shfl_sync(FULL, ls, L), n_ = __shfl_sync(FULL, lr, L);
                for (uint32_t i = lane; i < n_; i += 32u) o[d_ + i] = lit[s_ + i];
            }
        }
        // ---- matches: which earlier lanes of this step write bytes that this lane reads? ----
        uint32_t pend = __ballot_sync(FULL, len > 0u);
        if (pend) {
            const uint32_t src = dm - dist;
            const uint32_t src_end = dist < len ? dm : src + len;    // source bytes outside the lane's own destination
            const uint32_t me = dm + len;
            uint32_t lo = 0, hi = 0;                                 // lo = #lanes with me <= src, hi = #lanes with dm < src_end
#pragma unroll
            for (int step = 16; step; step >>= 1) {
                const uint32_t vme = __shfl_sync(FULL, me, (int)lo + step - 1), vms = __shfl_sync(FULL, dm, (int)hi + step - 1);
                if (vme <= src) lo += (uint32_t)step;
                if (vms < src_end) hi += (uint32_t)step;
            }
            const uint32_t depmask = hi > lo ? (((1u << hi) - 1u) & ~((1u << lo) - 1u)) : 0u;
            __syncwarp();                                            // literals of this step are in place
            while (pend) {
                const bool ready = ((pend >> lane) & 1u) && !(depmask & pend);
                uint32_t big = __ballot_sync(FULL, ready && len >= LONG_COPY);
                while (big) {
                    const int L = __ffs(big) - 1;
                    big &= big - 1u;
                    const uint32_t d_ = __shfl_sync(FULL, dm, L), s_ = __shfl_sync(FULL, src, L), n_ = __shfl_sync(FULL, len, L),
                                   p_ = __shfl_sync(FULL, dist, L);
                    for (uint32_t i = lane; i < n_; i += 32u) o[d_ + i] = o[s_ + (p_ < n_ ? i % p_ : i)];
                }
                const uint32_t nb = (ready && len < LONG_COPY) ? len : 0u;
                const uint32_t mx = __reduce_max_sync(FULL, nb);
                for (uint32_t i = 0; i < mx; ++i)
                    if (i < nb) o[dm + i] = o[src + i];
                __syncwarp();
                pend &= ~__ballot_sync(FULL, ready);
            }
        }
        pos += __shfl_sync(FULL, ae, 31);
        lpos += __shfl_sync(FULL, le, 31);
        __syncwarp();
    }

    // ---- CRC-32 (htslib/bgzf.c:793) of the finished block ----
    const uint8_t *p = o, *end = o + d.isize;
    const uint8_t *a0 = reinterpret_cast<const uint8_t *>(((uintptr_t)o + 127u) & ~(uintptr_t)127);
    if (a0 > end) a0 = end;
    uint32_t r = 0xffffffffu;
    if (lane == 0)
        for (; p < a0; ++p) r = tab(4 * 256 + ((r ^ *p) & 0xffu)) ^ (r >> 8);
    r = __shfl_sync(FULL, r, 0);
    const uint32_t rows = (uint32_t)(end - a0) >> 7;
    if (rows) {
        const uint32_t *wp = reinterpret_cast<const uint32_t *>(a0) + lane;
        uint32_t H = 0;
        for (uint32_t row = 0; row < rows; ++row) {
            uint32_t w = wp[(size_t)row * 32u];
            if (row == 0 && lane == 0) w ^= r;
            H = tab(H & 0xffu) ^ tab(256 + ((H >> 8) & 0xffu)) ^ tab(512 + ((H >> 16) & 0xffu)) ^ tab(768 + (H >> 24)) ^ w;
        }
        uint32_t v = crc_mulmod(tab(5 * 256 + lane), H);
#pragma unroll
        for (int sft = 16; sft; sft >>= 1) v ^= __shfl_xor_sync(FULL, v, sft);
        r = v;
    }
    if (lane == 0) {
        for (p = a0 + ((size_t)rows << 7); p < end; ++p) r = tab(4 * 256 + ((r ^ *p) & 0xffu)) ^ (r >> 8);
        if ((r ^ 0xffffffffu) != d.crc) status[b] = INF_ERR_CRC;
        else if (pos != d.isize) status[b] = INF_ERR_ISIZE;
    }
}

// =========================================================================================================
// phase 2, fourth generation: one CTA per block, the output window in SHARED MEMORY
// =========================================================================================================
// ncu on bgzf_resolve3_kernel (profiles/r1k): 4.59 GB read + 1.20 GB written to produce 1.22 GB — every back-reference byte went
// through L2/DRAM as a 1-byte LDG -> STG pair (40 % of the samples on that chain, 15 cycles per issued instruction).  A BGZF block
// never refers outside itself (its window is its own <= 64 KB of output), so here the block is put together on chip:
//   * the CTA's 128 threads take the block's commands 128 at a time: one block-wide scan gives every command its destination;
//     the literals of the window's commands go from the literal stream into shared memory;
//   * a match waits only for the earlier matches OF ITS WINDOW whose destination overlaps its source (two binary searches over
//     the window's destination bounds in shared memory give that range; everything in front of the window is final), and copies
//     shared -> shared (~30 cycles per dependent byte instead of an L2 round trip);
//   * the CRC-32 is computed from shared memory by all four warps (the lane-interleaved Horner form of phase 2, one piece per
//     warp, pieces combined by multiplication with x^(8 * bytes behind the piece));
//   * the finished block leaves with coalesced 16-byte stores: the window is laid out in shared memory at the block's own
//     alignment modulo 16, so both sides of the copy are aligned (north-star item (b): LZ77 output written as 16-byte stores).
// DRAM traffic of phase 2 = literal + command streams in, inflated bytes out.  Three CTAs (3 x ~72 KB) share an SM.
constexpr int R4_THREADS = 128, R4_WARPS = R4_THREADS / 32;
constexpr int R4_WIN = 65536 + 32;                 // the window + alignment slack
__device__ uint32_t g_crc4_x2n[32];
struct R4Smem {
    uint8_t win[R4_WIN];
    uint32_t tabs[CRC3_WORDS];
    uint32_t x2n[32];
    uint32_t dm[R4_THREADS], me[R4_THREADS];       // destination bounds of the window's matches (ascending)
    uint32_t part_a[R4_WARPS], part_l[R4_WARPS];   // scan partials
    uint32_t pend[R4_WARPS];                       // matches of the window not copied yet
    uint32_t piece[R4_WARPS];                      // CRC registers of the four pieces
    uint32_t r0;                                   // CRC register behind the bytes in front of the first 128-byte row
};

__global__ void __launch_bounds__(R4_THREADS, 3)
bgzf_resolve4_kernel(const BlockDesc *__restrict__ blocks, uint32_t nblocks, uint8_t *__restrict__ outbuf,
                     int *__restrict__ status, const uint32_t *__restrict__ ncmd_in, const uint8_t *__restrict__ work) {
    extern __shared__ __align__(16) uint8_t r4_raw[];
    R4Smem &sm = *reinterpret_cast<R4Smem *>(r4_raw);
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    for (int i = tid; i < CRC3_WORDS; i += R4_THREADS) sm.tabs[i] = g_crc3[i];
    if (tid < 32) sm.x2n[tid] = g_crc4_x2n[tid];
    const uint32_t b = blockIdx.x;
    if (b >= nblocks) return;
    if (status[b] != INF_OK) return;                 // uniform for the CTA
    const BlockDesc d = blocks[b];
    uint8_t *o = outbuf + d.out_off;
    uint8_t *win = sm.win + ((uintptr_t)o & 15u);    // same alignment modulo 16 as the destination
    const uint8_t *lit = work + d.cmd_off;
    const uint32_t *cmd_top = reinterpret_cast<const uint32_t *>(work + d.cmd_off + inflate_work_cap(d.isize)) - 1;
    const uint32_t n = ncmd_in[b];
    uint32_t pos = 0, lpos = 0;
    __syncthreads();

    for (uint32_t k0 = 0; k0 < n; k0 += R4_THREADS) {
        const uint32_t k = k0 + tid;
        const uint32_t c = k < n ? cmd_top[-(int)k] : 0x80000000u;
        const bool isrun = (c >> 31) != 0u;
        const uint32_t lr = isrun ? (c & 0x7fffffffu) : (c >> 23);
        const uint32_t len = isrun ? 0u : ((c >> 15) & 255u) + 3u;
        const uint32_t dist = (c & 0x7fffu) + 1u;
        // block-wide inclusive scans: output bytes and literal bytes up to and including this command
        uint32_t ae = lr + len, le = lr;
#pragma unroll
        for (int sft = 1; sft < 32; sft <<= 1) {
            const uint32_t va = __shfl_up_sync(FULL, ae, sft), vl = __shfl_up_sync(FULL, le, sft);
            if (lane >= (uint32_t)sft) { ae += va; le += vl; }
        }
        if (lane == 31u) { sm.part_a[warp] = ae; sm.part_l[warp] = le; }
        __syncthreads();
        uint32_t tot_a = 0, tot_l = 0;
#pragma unroll
        for (int w = 0; w < R4_WARPS; ++w) {
            const uint32_t pa = sm.part_a[w], pl = sm.part_l[w];
            if ((uint32_t)w < warp) { ae += pa; le += pl; }
            tot_a += pa;
            tot_l += pl;
        }
        const uint32_t dl = pos + ae - lr - len;     // where this command's literals go
        const uint32_t dm = dl + lr;                 // where its match goes
        const uint32_t ls = lpos + le - lr;          // its literals in the literal stream
        const bool sane = pos + tot_a <= d.isize;    // phase 1 checked every command against ISIZE: this only fences shared memory
        // ---- literals: stream -> window ----
        if (sane) {
            const uint32_t small = lr < 32u ? lr : 0u;
            for (uint32_t i = 0; i < small; ++i) win[dl + i] = lit[ls + i];
            uint32_t big = __ballot_sync(FULL, lr >= 32u);
            while (big) {
                const int L = __ffs(big) - 1;
                big &= big - 1u;
                const uint32_t d_ = __shfl_sync(FULL, dl, L), s_ = __shfl_sync(FULL, ls, L), n_ = __shfl_sync(FULL, lr, L);
                for (uint32_t i = lane; i < n_; i += 32u) win[d_ + i] = lit[s_ + i];
            }
        }
        sm.dm[tid] = dm;
        sm.me[tid] = dm + len;
        const uint32_t mine = __ballot_sync(FULL, len > 0u && sane);
        if (lane == 0u) sm.pend[warp] = mine;
        __syncthreads();                             // literals of the window are in place, bounds and pending bits visible
        // ---- matches: which earlier matches of this window write bytes that this one reads? ----
        const uint32_t src = dm - dist;
        const uint32_t src_end = dist < len ? dm : src + len;    // source bytes outside the match's own destination
        uint32_t lo = 0, hi = 0;                                 // lo = #commands with me <= src, hi = #commands with dm < src_end
        if (len > 0u) {
#pragma unroll
            for (int step = R4_THREADS / 2; step; step >>= 1) {
                if (sm.me[lo + step - 1] <= src) lo += (uint32_t)step;
                if (sm.dm[hi + step - 1] < src_end) hi += (uint32_t)step;
            }
            if (hi > tid) hi = tid;                              // only commands in front of this one can be waited for
        }
        bool todo = len > 0u && sane;
        for (;;) {
            bool ready = false;
            if (todo) {
                uint32_t blocked = 0;
                if (hi > lo) {
#pragma unroll
                    for (int w = 0; w < R4_WARPS; ++w) {
                        const uint32_t w_lo = (uint32_t)w * 32u, w_hi = w_lo + 32u;
                        if (hi <= w_lo || lo >= w_hi) continue;
                        const uint32_t a = lo > w_lo ? lo - w_lo : 0u, e = hi < w_hi ? hi - w_lo : 32u;
                        const uint32_t m = (e >= 32u ? 0xffffffffu : ((1u << e) - 1u)) & ~((1u << a) - 1u);
                        blocked |= sm.pend[w] & m;
                    }
                }
                ready = blocked == 0u;
                if (ready) {
                    if (dist >= len && len >= 8u && ((((uintptr_t)(win + dm)) | ((uintptr_t)(win + src))) & 3u) == 0u) {   // aligned, not overlapping: words
                        uint32_t i = 0;
                        for (; i + 4u <= len; i += 4u) *reinterpret_cast<uint32_t *>(win + dm + i) = *reinterpret_cast<const uint32_t *>(win + src + i);
                        for (; i < len; ++i) win[dm + i] = win[src + i];
                    } else {
                        for (uint32_t i = 0; i < len; ++i) win[dm + i] = win[src + i];   // dist < len: the copy feeds itself, byte by byte
                    }
                }
            }
            __syncthreads();                         // copies of this round are visible
            const uint32_t done = __ballot_sync(FULL, ready);
            if (lane == 0u && done) sm.pend[warp] &= ~done;
            todo = todo && !ready;
            __syncthreads();
            uint32_t any = 0;
#pragma unroll
            for (int w = 0; w < R4_WARPS; ++w) any |= sm.pend[w];
            if (!any) break;
        }
        pos += tot_a;
        lpos += tot_l;
        if (!sane) break;                            // uniform: tot_a is the same for every thread
    }
    __syncthreads();

    // ---- CRC-32 (htslib/bgzf.c:793) of the finished block, from shared memory ----
    // rows of 128 bytes from the first 128-byte-aligned address of the window; lane-interleaved Horner per warp over its piece
    const uint32_t isize = d.isize;
    const uint32_t head = (uint32_t)(((uintptr_t)win + 127u) & ~(uintptr_t)127) - (uint32_t)(uintptr_t)win;   // bytes in front of the rows
    const uint32_t h0 = head < isize ? head : isize;
    const uint32_t rows = (isize - h0) >> 7;
    const uint32_t *T0 = sm.tabs + 4 * 256;
    if (tid == 0) {
        uint32_t r = 0xffffffffu;
        for (uint32_t i = 0; i < h0; ++i) r = T0[(r ^ win[i]) & 0xffu] ^ (r >> 8);
        sm.r0 = r;
    }
    __syncthreads();
    const uint32_t r0 = sm.r0;                       // the register coming out of the head bytes: goes into row 0, whichever warp owns it
    {
        const uint32_t r_lo = rows * warp / R4_WARPS, r_hi = rows * (warp + 1u) / R4_WARPS;
        uint32_t reg = 0;
        if (r_hi > r_lo) {
            const uint32_t *wp = reinterpret_cast<const uint32_t *>(win + h0) + (size_t)r_lo * 32u + lane;
            uint32_t H = 0;
            for (uint32_t row = r_lo; row < r_hi; ++row, wp += 32) {
                uint32_t w = *wp;
                if (row == 0u && lane == 0u) w ^= r0;
                H = sm.tabs[H & 0xffu] ^ sm.tabs[256 + ((H >> 8) & 0xffu)] ^ sm.tabs[512 + ((H >> 16) & 0xffu)] ^ sm.tabs[768 + (H >> 24)] ^ w;
            }
            uint32_t v = crc_mulmod(sm.tabs[5 * 256 + lane], H);
#pragma unroll
            for (int sft = 16; sft; sft >>= 1) v ^= __shfl_xor_sync(FULL, v, sft);
            reg = v;
            // advance the piece's register over the rows behind it: multiplication by x^(8 * bytes)
            if (lane == 0u) reg = crc_shift(sm.x2n, reg, (rows - r_hi) << 7);
        }
        if (lane == 0u) sm.piece[warp] = reg;
    }
    __syncthreads();
    if (tid == 0) {
        uint32_t r = rows ? (sm.piece[0] ^ sm.piece[1] ^ sm.piece[2] ^ sm.piece[3]) : r0;   // no row: r0 never went into a word
        for (uint32_t i = h0 + (rows << 7); i < isize; ++i) r = T0[(r ^ win[i]) & 0xffu] ^ (r >> 8);
        if ((r ^ 0xffffffffu) != d.crc) status[b] = INF_ERR_CRC;
        else if (pos != isize) status[b] = INF_ERR_ISIZE;
    }
    // ---- the block leaves: coalesced 16-byte stores (window and destination share their alignment modulo 16) ----
    {
        const uint32_t mis = (uint32_t)((uintptr_t)o & 15u), pre = mis ? 16u - mis : 0u;
        const uint32_t p0 = pre < isize ? pre : isize;
        if (tid < p0) o[tid] = win[tid];
        const uint32_t vecs = (isize - p0) >> 4;
        const uint4 *sv = reinterpret_cast<const uint4 *>(win + p0);
        uint4 *gv = reinterpret_cast<uint4 *>(o + p0);
        for (uint32_t i = tid; i < vecs; i += R4_THREADS) gv[i] = sv[i];
        const uint32_t t0 = p0 + (vecs << 4);
        if (t0 + tid < isize && tid < 16u) o[t0 + tid] = win[t0 + tid];
    }
}

// phase 2, fifth generation (bgzf_resolve5.cuh holds the per-warp body, shared with the host-side SIMT emulation of the tests)
__global__ void __launch_bounds__(256, 4)
bgzf_resolve5_kernel(const BlockDesc *__restrict__ blocks, uint32_t nblocks, uint8_t *__restrict__ outbuf,
                     int *__restrict__ status, const uint32_t *__restrict__ ncmd_in, const uint8_t *__restrict__ work, const int pf_mode) {
    __shared__ uint32_t tabs_s[CRC3_WORDS];
    __shared__ __align__(16) uint8_t warp_s[8][R5_WARP_SMEM];
    for (int i = threadIdx.x; i < CRC3_WORDS; i += 256) tabs_s[i] = g_crc3[i];
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t b = blockIdx.x * 8u + (threadIdx.x >> 5);
    if (b >= nblocks) return;
    if (status[b] != INF_OK) return;
    resolve5_warp(blocks[b], ncmd_in[b], outbuf, work, status + b, tabs_s, warp_s[threadIdx.x >> 5], lane, pf_mode);
}

// ---- host side ---------------------------------------------------------------------------------------------------

// How the two phases are placed (measured A/B in profiles/r2i_*):
//   DH_P1_SMALL=1     phase 1 runs with the 222 KB pools, so other CTAs can share its SMs,
//   DH_RESOLVE_LDG=1  phase 2 keeps no shared-memory copy of the CRC tables (needed to fit next to phase 1),
//   stream2 != null   phase 2 (and whatever the caller enqueues behind it) runs on stream2 behind `mid`, so that phase 1 of the
//                     caller's NEXT launch on `stream` overlaps it.
struct V3Config { bool p1_small, resolve_ldg, resolve4, resolve5; int r5_pf; };
static const V3Config &v3_config() {
    static const V3Config c = {getenv("DH_P1_SMALL") && atoi(getenv("DH_P1_SMALL")) != 0, getenv("DH_RESOLVE_LDG") && atoi(getenv("DH_RESOLVE_LDG")) != 0,
                               getenv("DH_RESOLVE4") != nullptr, !getenv("DH_RESOLVE5") || atoi(getenv("DH_RESOLVE5")) != 0,
                               getenv("DH_R5_PF") ? atoi(getenv("DH_R5_PF")) : 1};
    return c;
}

static uint32_t v3_retry_ctas(int sm_count) {
    return P1Main::TAB_N < P1Retry::TAB_N ? (sm_count >= 64 ? (uint32_t)DH_RETRY_CTAS : sm_count >= 8 ? 1u : 0u) : 0u;
}
// Streams that phase 1 decodes at a time on `sm_count` SMs: every main CTA holds THREADS streams, one per lane, and a warp takes its
// next 32 only when all of its lanes are done.  A launch of k * round blocks ends all its warps together; anything else leaves the
// SMs to idle through the last generation (100 M-record bench, 468,651 blocks: launches of 2.06 rounds ran phase 1 in 153 ms, launches
// of 3 rounds in 120 ms).
uint32_t bgzf_inflate_v3_round(int sm_count) {
    const uint32_t r = v3_retry_ctas(sm_count);
    return ((uint32_t)sm_count > r ? (uint32_t)sm_count - r : 1u) * (uint32_t)P1Main::THREADS;
}

cudaError_t launch_bgzf_inflate_v3(const uint8_t *comp, const BlockDesc *blocks, uint32_t nblocks, uint8_t *out, int *status,
                                   uint8_t *work, unsigned int *counter, int sm_count, cudaStream_t stream, cudaEvent_t between_phases,
                                   cudaStream_t stream2, cudaEvent_t mid) {
    static std::mutex mu;
    static bool ready[64] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    {
        std::lock_guard<std::mutex> lk(mu);   // several contexts (threads, devices) may arrive here at once
        if (!ready[dev]) {
            static uint32_t host_tabs[CRC3_WORDS];
            build_crc3_tables(host_tabs);
            e = cudaMemcpyToSymbol(g_crc3, host_tabs, sizeof host_tabs);
            if (e != cudaSuccess) return e;
            e = cudaFuncSetAttribute(bgzf_decode3_kernel<P1Main, P1Retry>, cudaFuncAttributeMaxDynamicSharedMemorySize, p1_smem<P1Main, P1Retry>());
            if (e != cudaSuccess) return e;
            e = cudaFuncSetAttribute(bgzf_decode3_kernel<P1MainS, P1RetryS>, cudaFuncAttributeMaxDynamicSharedMemorySize, p1_smem<P1MainS, P1RetryS>());
            if (e != cudaSuccess) return e;
            uint32_t x2n[32];
            crc_x2n_init(x2n);
            e = cudaMemcpyToSymbol(g_crc4_x2n, x2n, sizeof x2n);
            if (e != cudaSuccess) return e;
            e = cudaFuncSetAttribute(bgzf_resolve4_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(R4Smem));
            if (e != cudaSuccess) return e;
            e = cudaFuncSetAttribute(bgzf_resolve4_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
            if (e != cudaSuccess) return e;
            if (getenv("DH_CARVEOUT_MAX") && atoi(getenv("DH_CARVEOUT_MAX"))) {       // see bam_decode.cu
                cudaFuncSetAttribute(bgzf_resolve3_kernel<false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
                cudaFuncSetAttribute(bgzf_resolve3_kernel<true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
            }
            ready[dev] = true;
        }
    }
    const V3Config &cfg = v3_config();
    e = cudaMemsetAsync(counter, 0, sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    if (nblocks == 0) {
        if (stream2 && mid) { if ((e = cudaEventRecord(mid, stream)) != cudaSuccess) return e; return cudaStreamWaitEvent(stream2, mid, 0); }
        return cudaSuccess;
    }
    uint32_t *ncmd = reinterpret_cast<uint32_t *>(work);
    unsigned int *retry = reinterpret_cast<unsigned int *>(work + inflate_work_retry(nblocks));   // queue header + 8 words per block
    uint8_t *areas = work + inflate_work_header(nblocks);
    e = cudaMemsetAsync(retry, 0xff, inflate_work_header(nblocks) - inflate_work_retry(nblocks), stream);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(retry, 0, RETRY_HDR * sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    // main CTAs: one per SM at most, 6 warps x 32 streams each; retry CTAs (same pool as the main pass => none needed)
    const uint32_t retry_ctas = v3_retry_ctas(sm_count);
    uint32_t ctas = (nblocks + P1Main::THREADS - 1) / P1Main::THREADS;
    if (ctas > (uint32_t)sm_count - retry_ctas) ctas = (uint32_t)sm_count - retry_ctas;
    if (cfg.p1_small)
        bgzf_decode3_kernel<P1MainS, P1RetryS><<<ctas + retry_ctas, P1MainS::THREADS, p1_smem<P1MainS, P1RetryS>(), stream>>>(
            comp, blocks, nblocks, status, ncmd, areas, counter, retry, retry_ctas);
    else
        bgzf_decode3_kernel<P1Main, P1Retry><<<ctas + retry_ctas, P1Main::THREADS, p1_smem<P1Main, P1Retry>(), stream>>>(
            comp, blocks, nblocks, status, ncmd, areas, counter, retry, retry_ctas);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    if (between_phases && (e = cudaEventRecord(between_phases, stream)) != cudaSuccess) return e;
    cudaStream_t s2 = stream;
    if (stream2 && mid) {
        if ((e = cudaEventRecord(mid, stream)) != cudaSuccess) return e;
        if ((e = cudaStreamWaitEvent(stream2, mid, 0)) != cudaSuccess) return e;
        s2 = stream2;
    }
    // phase 2: bgzf_resolve5_kernel (warp per block, byte-parallel rows, shared-memory window, 16-byte stores) is the default since
    // r2v: 94.9 ms against 104.7 ms of bgzf_resolve3_kernel (DH_RESOLVE5=0: warp per block, command-parallel, in global memory) on the
    // 100 M-record bench (468,651 blocks), and half its DRAM reads.  DH_RESOLVE4=1 selects the CTA-per-block shared-memory kernel,
    // measured r2b: 295.5 ms -- three 72 KB CTAs per SM put 3 blocks in flight where the warp-per-block kernels have 32.
    if (cfg.resolve4) bgzf_resolve4_kernel<<<nblocks, R4_THREADS, sizeof(R4Smem), s2>>>(blocks, nblocks, out, status, ncmd, areas);
    else if (cfg.resolve5) bgzf_resolve5_kernel<<<(nblocks + 7u) / 8u, 256, 0, s2>>>(blocks, nblocks, out, status, ncmd, areas, cfg.r5_pf);
    else if (cfg.resolve_ldg) bgzf_resolve3_kernel<false><<<(nblocks + 7u) / 8u, 256, 0, s2>>>(blocks, nblocks, out, status, ncmd, areas);
    else bgzf_resolve3_kernel<true><<<(nblocks + 7u) / 8u, 256, 0, s2>>>(blocks, nblocks, out, status, ncmd, areas);
    return cudaGetLastError();
}


// first failing block of a launch, written straight to (pinned host) memory: the scan pipeline learns the verdict
// without a device-to-host copy that would queue behind the column batches on the copy engine
__global__ void __launch_bounds__(256) status_first_bad_kernel(const int *__restrict__ status, uint32_t nblocks, int *__restrict__ out) {
    __shared__ unsigned int first;
    if (threadIdx.x == 0) first = 0xffffffffu;
    __syncthreads();
    unsigned int mine = 0xffffffffu;
    for (uint32_t i = threadIdx.x; i < nblocks; i += 256u)
        if (status[i] != 0) { mine = i; break; }
    if (mine != 0xffffffffu) atomicMin(&first, mine);
    __syncthreads();
    if (threadIdx.x == 0) {
        out[1] = first == 0xffffffffu ? 0 : status[first];
        out[0] = first == 0xffffffffu ? -1 : (int)first;
    }
}
cudaError_t launch_status_first_bad(const int *status, uint32_t nblocks, int *out, cudaStream_t stream) {
    status_first_bad_kernel<<<1, 256, 0, stream>>>(status, nblocks, out);
    return cudaGetLastError();
}

}  // namespace dh
