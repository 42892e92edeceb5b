// bgzf_inflate2.cu — "lockstep" BGZF inflate for sm_100a, second generation.
//
// Replaces bgzf_uncompress / inflate_block / bgzf_decode_func (htslib/bgzf.c:762-804, 808-824, 1373-1384).
//
// Why a second kernel: ncu on the first (lane-group) kernel showed 7.9 of 32 lanes active per warp instruction
// and ~25 warp-instructions per output byte (profiles/r1a_inflate_v1_ncu_summary.csv).  DEFLATE is bit-serial per
// stream, so the only way to fill a warp is ONE STREAM PER LANE; the two things that normally make that
// impossible are handled like this:
//   * Huffman tables of 32 streams do not fit next to each other as u32 tables -> u16 entries, lane-interleaved
//     ([entry][lane], so a warp-wide lookup at 32 different indices is at most a 2-way bank conflict), 852 litlen +
//     332 distance entries per stream = 75.8 KB per warp, 3 warps (96 streams) per SM in 227 KB of shared memory.
//   * LZ77 copies diverge and need the stream's own earlier output -> phase 1 does NOT copy.  It writes literals
//     to their final position (whole aligned 8-byte words; bytes that belong to matches are don't-care) and
//     emits one 32-bit command per match.  Huffman decoding never looks at output bytes, so phase 1 is a pure,
//     branch-light decode loop that all 32 lanes run in lock-step: zlib ends a deflate block every 16,383
//     symbols (lit_bufsize - 1, memLevel 8), so the lanes of a warp reach end-of-block, parse the next dynamic
//     header and build their tables at the same step.
//   * Phase 2 (a second kernel, thread per block, no shared tables, 2048 threads/SM to hide L2 latency) replays
//     the match commands as byte copies inside the block's output and then computes the block's CRC-32
//     (slice-by-8) and compares CRC and ISIZE with the BGZF footer.
// Streams whose distance code needs more than 332 table entries (worst case is 400; zlib/libdeflate output never
// comes close) are flagged INF_NEEDS_V1 and re-done by the first-generation kernel in filter mode.
#include "bgzf_inflate.cuh"
#include "crc32_core.cuh"
#include "inflate_core.cuh"

namespace dh {

namespace v2 {

constexpr int WARPS = 3;
constexpr int THREADS = WARPS * 32;
constexpr int LROOT = 9, DROOT = 7;
constexpr int LIT_N = 852;        // worst case for 286 symbols, root 9 (zlib's ENOUGH_LENS)
constexpr int DIST_N = 352;       // region size; the distance TABLE may use DIST_CAP of it
constexpr int DIST_CAP = 332;     // worst case for 30 symbols, root 7 is 400 (enumerated in DESIGN.md) -> fallback above this
constexpr int NC_LIT = 316;       // dist-region slots [316,332): next_code[] while the litlen table is built
constexpr int NC_DIST = 332;      // dist-region slots [332,348): next_code[] while the distance table is built
constexpr int TAB_N = LIT_N + DIST_N;
constexpr unsigned FULL = 0xffffffffu;

enum : uint32_t { K_LIT = 0, K_LEN = 1, K_EOB = 2, K_LINK = 3, K_BAD = 4 };
__device__ __forceinline__ uint32_t mk(uint32_t kind, uint32_t val, uint32_t len) { return (kind << 13) | (val << 4) | len; }
__device__ __forceinline__ uint32_t te_len(uint32_t e) { return e & 15u; }
__device__ __forceinline__ uint32_t te_val(uint32_t e) { return (e >> 4) & 511u; }
__device__ __forceinline__ uint32_t te_kind(uint32_t e) { return e >> 13; }

struct WarpTab { uint16_t e[TAB_N][32]; };

// ---- bit reader (one per lane; aligned 32-bit loads, one word of load-ahead) ---------------------------
struct Bits {
    const uint32_t *p;   // next word to load into w_next
    uint64_t buf;
    uint32_t cnt;        // valid bits in buf
    uint32_t w_next;
    uint32_t loaded;     // payload bits loaded so far (to detect running past the payload)
    __device__ __forceinline__ void open(const uint8_t *payload) {
        uintptr_t a = (uintptr_t)payload;
        const uint32_t *q = reinterpret_cast<const uint32_t *>(a & ~(uintptr_t)3);
        uint32_t mis = (uint32_t)(a & 3u);
        buf = (uint64_t)(q[0] >> (8u * mis));
        cnt = 32u - 8u * mis;
        loaded = cnt;
        w_next = q[1];
        p = q + 2;
    }
    __device__ __forceinline__ void refill() {  // >= 33 valid bits afterwards
        if (cnt <= 32u) {
            buf |= (uint64_t)w_next << cnt;
            cnt += 32u;
            loaded += 32u;
            w_next = *p++;
        }
    }
    __device__ __forceinline__ uint32_t peek(uint32_t n) const { return (uint32_t)buf & ((1u << n) - 1u); }
    __device__ __forceinline__ void drop(uint32_t n) { buf >>= n; cnt -= n; }
    __device__ __forceinline__ uint32_t take(uint32_t n) { uint32_t v = peek(n); drop(n); return v; }
    __device__ __forceinline__ uint32_t used() const { return loaded - cnt; }
};

// ---- canonical table build, one lane's private column of the warp table ---------------------------------
// lens come from len_of(j), j in [0,n).  base/root/cap describe the destination region; nc = first of 16 scratch slots
// for next_code[].  Returns INF_OK, an INF_ERR_* code, or INF_NEEDS_V1 when the table does not fit.
template <bool LIT, typename LenFn>
__device__ __forceinline__ int build_table(WarpTab *wt, const uint32_t lane, const int base, const int root, const int cap,
                                           const int n, const int nc, LenFn len_of) {
    // counts per length in 16-bit fields of four 64-bit registers
    uint64_t c0 = 0, c1 = 0, c2 = 0, c3 = 0;
    for (int j = 0; j < n; ++j) {
        uint32_t l = len_of(j);
        uint64_t one = 1ull << ((l & 3u) * 16u);
        uint32_t s = l >> 2;
        c0 += s == 0 ? one : 0ull;
        c1 += s == 1 ? one : 0ull;
        c2 += s == 2 ? one : 0ull;
        c3 += s == 3 ? one : 0ull;
    }
    c0 &= ~0xffffull;  // length 0 = unused symbol
    auto count = [&](int l) -> uint32_t {
        int s = l >> 2;
        uint64_t w = s == 0 ? c0 : s == 1 ? c1 : s == 2 ? c2 : c3;
        return (uint32_t)(w >> ((l & 3) * 16)) & 0xffffu;
    };
    int left = 1, maxlen = 0;
    uint32_t code = 0;
    const int BR = base;  // root entries start here
#pragma unroll
    for (int l = 1; l <= 15; ++l) {
        uint32_t cl = count(l);
        left = (left << 1) - (int)cl;
        if (left < 0) return LIT ? INF_ERR_LITLEN_SET : INF_ERR_DIST_SET;   // over-subscribed
        if (cl) maxlen = l;
        wt->e[DIST_N * 0 + LIT_N + nc + l][lane] = (uint16_t)code;          // next_code[l] (first canonical code of length l)
        code = (code + cl) << 1;
    }
    const int nroot = 1 << root;
    if (maxlen == 0) {  // no codes: legal (e.g. literals only -> empty distance code); every lookup is BAD
        for (int k = 0; k < nroot; ++k) wt->e[BR + k][lane] = (uint16_t)mk(K_BAD, 0, 1);
        return INF_OK;
    }
    if (left > 0) {
        if (maxlen != 1) return LIT ? INF_ERR_LITLEN_SET : INF_ERR_DIST_SET;  // incomplete set (zlib inftrees.c rule)
        for (int k = 0; k < nroot; ++k) wt->e[BR + k][lane] = (uint16_t)mk(K_BAD, 0, 1);
    }
    // sub-table layout: every root prefix that carries codes longer than root gets a sub-table sized by the longest
    // code under it.  Canonical codes grow left to right, so these prefixes are the top end of the prefix space.
    if (maxlen > root) {
        int p_first = nroot;
        for (int l = root + 1; l <= maxlen; ++l) {
            uint32_t cl = count(l);
            if (!cl) continue;
            uint32_t first = wt->e[LIT_N + nc + l][lane];
            int p0 = (int)(first >> (l - root)), p1 = (int)((first + cl - 1u) >> (l - root));
            if (p0 < p_first) p_first = p0;
            for (int p = p0; p <= p1; ++p) wt->e[BR + (int)(__brev((uint32_t)p) >> (32 - root))][lane] = (uint16_t)(l - root);
        }
        int next_free = nroot;
        for (int p = p_first; p < nroot; ++p) {
            int ri = BR + (int)(__brev((uint32_t)p) >> (32 - root));
            uint32_t bits = wt->e[ri][lane];
            int off = next_free;
            next_free += 1 << bits;
            if (next_free > cap) return INF_NEEDS_V1;
            wt->e[ri][lane] = (uint16_t)mk(K_LINK, (uint32_t)(off - nroot), bits);
        }
    }
    // place the symbols
    for (int j = 0; j < n; ++j) {
        uint32_t l = len_of(j);
        if (!l) continue;
        uint32_t cd = wt->e[LIT_N + nc + (int)l][lane];
        wt->e[LIT_N + nc + (int)l][lane] = (uint16_t)(cd + 1u);
        uint32_t rev = __brev(cd) >> (32u - l);
        uint32_t ent;
        if (LIT) ent = j < 256 ? mk(K_LIT, (uint32_t)j, l) : j == 256 ? mk(K_EOB, 0, l) : j < 286 ? mk(K_LEN, (uint32_t)j - 257u, l) : mk(K_BAD, 0, l);
        else ent = j < 30 ? mk(K_LEN, (uint32_t)j, l) : mk(K_BAD, 0, l);
        if ((int)l <= root) {
            for (uint32_t k = rev; k < (uint32_t)nroot; k += 1u << l) wt->e[BR + (int)k][lane] = (uint16_t)ent;
        } else {
            uint32_t link = wt->e[BR + (int)(rev & (uint32_t)(nroot - 1))][lane];
            int sub = BR + nroot + (int)te_val(link);
            uint32_t bits = te_len(link);
            for (uint32_t k = rev >> root; k < (1u << bits); k += 1u << (l - (uint32_t)root)) wt->e[sub + (int)k][lane] = (uint16_t)ent;
        }
    }
    return INF_OK;
}

// per-lane decoder state
struct Lane {
    Bits in;
    uint8_t *out;        // block's first output byte
    uint32_t *cmd;       // block's command buffer
    uint64_t acc;        // literal bytes of the current aligned 8-byte output word
    uint32_t accmask;    // which bytes of acc are literals
    uint32_t pos;        // output position
    uint32_t ncmd, lit_run;
    uint32_t isize, in_bits;
    int status;
};

__device__ __forceinline__ void flush_word(Lane &s) {
    if (s.accmask) {
        // word address in global memory: the block's output is not 8-byte aligned in general
        uintptr_t a = (uintptr_t)(s.out + s.pos - 1) & ~(uintptr_t)7;   // word of the LAST byte written (pos > 0 here)
        uint8_t *w = reinterpret_cast<uint8_t *>(a);
        if (w >= s.out && w + 8 <= s.out + s.isize) *reinterpret_cast<uint64_t *>(w) = s.acc;
        else {
#pragma unroll
            for (int k = 0; k < 8; ++k)
                if ((s.accmask >> k) & 1u) w[k] = (uint8_t)(s.acc >> (8 * k));
        }
    }
    s.acc = 0;
    s.accmask = 0;
}
__device__ __forceinline__ void put_literal(Lane &s, uint32_t byte) {
    uint32_t k = (uint32_t)((uintptr_t)(s.out + s.pos) & 7u);
    s.acc |= (uint64_t)byte << (8u * k);
    s.accmask |= 1u << k;
    s.pos++;
    s.lit_run++;
    if (k == 7u) flush_word(s);
}
__device__ __forceinline__ void skip_bytes(Lane &s, uint32_t len) {  // a match: leave the bytes to phase 2
    uintptr_t a0 = (uintptr_t)(s.out + s.pos) >> 3, a1 = (uintptr_t)(s.out + s.pos + len) >> 3;
    if (a0 != a1 && s.accmask) flush_word(s);
    s.pos += len;
}

}  // namespace v2

using namespace v2;

// =========================================================================================================
// phase 1: Huffman decode, one stream per lane
// =========================================================================================================
__global__ void __launch_bounds__(THREADS, 1)
bgzf_decode_kernel(const uint8_t *__restrict__ comp, const BlockDesc *__restrict__ blocks, uint32_t nblocks,
                   uint8_t *__restrict__ outbuf, int *__restrict__ status, uint32_t *__restrict__ ncmd_out,
                   uint32_t *__restrict__ cmdbuf, unsigned int *__restrict__ counter) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    WarpTab *wt = reinterpret_cast<WarpTab *>(smem_raw) + warp;
    enum { ST_HEADER = 0, ST_RUN = 1, ST_DONE = 2 };

    for (;;) {
        // ---- a generation: 32 consecutive blocks, one per lane ----
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(counter, 32u);
        base = __shfl_sync(FULL, base, 0);
        if (base >= nblocks) break;
        const uint32_t b = base + lane;
        Lane s;
        int st = ST_DONE;
        uint32_t bfinal = 0;
        s.status = INF_OK;
        s.pos = s.ncmd = s.lit_run = 0;
        s.acc = 0;
        s.accmask = 0;
        s.isize = 0;
        s.in_bits = 0;
        s.out = outbuf;
        s.cmd = cmdbuf;
        if (b < nblocks) {
            const BlockDesc d = blocks[b];
            s.out = outbuf + d.out_off;
            s.cmd = cmdbuf + d.cmd_off;
            s.isize = d.isize;
            s.in_bits = d.in_len * 8u;
            s.in.open(comp + d.in_off);
            st = ST_HEADER;
        } else {
            s.in.open(comp);
        }
        auto fail = [&](int code) { s.status = code; st = ST_DONE; };

        while (__any_sync(FULL, st != ST_DONE)) {
            // ---------------- deflate block header (lanes in ST_HEADER; in lock-step for zlib-made streams) ----------------
            if (st == ST_HEADER) {
                s.in.refill();
                bfinal = s.in.take(1);
                const uint32_t btype = s.in.take(2);
                if (btype == 0) {
                    // stored: LEN bytes follow at the next byte boundary; they go through the literal path
                    s.in.drop(s.in.cnt & 7u);
                    s.in.refill();
                    uint32_t len = s.in.take(16), nlen = s.in.take(16);
                    if ((len ^ 0xffffu) != nlen) fail(INF_ERR_STORED_LEN);
                    else if (s.pos + len > s.isize) fail(INF_ERR_ISIZE);
                    else if (s.in.used() + 8u * len > s.in_bits) fail(INF_ERR_IN_OVERRUN);
                    else {
                        for (uint32_t i = 0; i < len; ++i) { s.in.refill(); put_literal(s, s.in.take(8)); }
                        st = bfinal ? ST_DONE : ST_HEADER;
                    }
                } else if (btype == 3) {
                    fail(INF_ERR_BTYPE);
                } else {
                    int rc = INF_OK;
                    uint32_t hlit = 288, hdist = 32;
                    uint32_t dl0 = 0x55555555u, dl1 = 0x55555555u, dl2 = 0x55555555u, dl3 = 0x55555555u;  // fixed code: 32 distance codes of 5 bits
                    if (btype == 1) {
                        // fixed Huffman code (RFC 1951 3.2.6): lengths written where the dynamic path would put them
                        for (int j = 0; j < 288; ++j)
                            wt->e[LIT_N + j][lane] = (uint16_t)(j < 144 ? 8 : j < 256 ? 9 : j < 280 ? 7 : 8);
                    } else {
                        hlit = s.in.take(5) + 257u;
                        hdist = s.in.take(5) + 1u;
                        const uint32_t hclen = s.in.take(4) + 4u;
                        if (hlit > 286u || hdist > 30u) rc = INF_ERR_TOO_MANY_SYMS;
                        // code-length code: 19 x 3 bits, kept in one register
                        uint64_t pl = 0;
                        for (uint32_t i = 0; i < hclen; ++i) { s.in.refill(); pl |= (uint64_t)s.in.take(3) << (3u * precode_order(i)); }
                        // 7-bit direct table in litlen-region slots [0,128): entry = sym << 4 | len, 0 = invalid
                        if (rc == INF_OK) {
                            int left = 1;
                            uint32_t cnt7[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                            for (int sym = 0; sym < 19; ++sym) { uint32_t l = (uint32_t)(pl >> (3 * sym)) & 7u;
#pragma unroll
                                for (int q = 1; q < 8; ++q) cnt7[q] += l == (uint32_t)q; }
#pragma unroll
                            for (int q = 1; q < 8; ++q) { left = (left << 1) - (int)cnt7[q]; if (left < 0) rc = INF_ERR_PRECODE; }
                            if (left > 0) rc = INF_ERR_PRECODE;   // incomplete code-length code is never allowed (zlib CODES rule)
                        }
                        if (rc == INF_OK) {
                            uint32_t code = 0;
                            for (uint32_t l = 1; l < 8; ++l) {
                                for (uint32_t sym = 0; sym < 19; ++sym) {
                                    if (((uint32_t)(pl >> (3u * sym)) & 7u) != l) continue;
                                    uint32_t rev = __brev(code) >> (32u - l);
                                    for (uint32_t k = rev; k < 128u; k += 1u << l) wt->e[k][lane] = (uint16_t)((sym << 4) | l);
                                    ++code;
                                }
                                code <<= 1;
                            }
                            // litlen + distance code lengths -> dist-region slots [0, hlit + hdist)
                            const uint32_t n = hlit + hdist;
                            uint32_t i = 0, prev = 0;
                            while (i < n) {
                                s.in.refill();
                                uint32_t e = wt->e[s.in.peek(7)][lane];
                                s.in.drop(e & 15u);
                                uint32_t sym = e >> 4;
                                if (sym < 16u) { wt->e[LIT_N + i][lane] = (uint16_t)sym; prev = sym; ++i; }
                                else {
                                    uint32_t rep, val;
                                    if (sym == 16u) { if (i == 0) { rc = INF_ERR_LENS_REPEAT; break; } rep = 3u + s.in.take(2); val = prev; }
                                    else if (sym == 17u) { rep = 3u + s.in.take(3); val = 0; }
                                    else { rep = 11u + s.in.take(7); val = 0; }
                                    if (i + rep > n) { rc = INF_ERR_LENS_REPEAT; break; }
                                    for (uint32_t k = 0; k < rep; ++k) wt->e[LIT_N + i + k][lane] = (uint16_t)val;
                                    i += rep;
                                    prev = val;
                                }
                            }
                            if (rc == INF_OK && wt->e[LIT_N + 256][lane] == 0) rc = INF_ERR_NO_EOB;
                            if (rc == INF_OK) {
                                dl0 = dl1 = dl2 = dl3 = 0;
                                for (uint32_t k = 0; k < hdist; ++k) {
                                    uint32_t v = (uint32_t)wt->e[LIT_N + hlit + k][lane] << (4u * (k & 7u)), q = k >> 3;
                                    dl0 |= q == 0 ? v : 0u; dl1 |= q == 1 ? v : 0u; dl2 |= q == 2 ? v : 0u; dl3 |= q == 3 ? v : 0u;
                                }
                            }
                        }
                    }
                    if (rc == INF_OK)
                        rc = build_table<true>(wt, lane, 0, LROOT, LIT_N, (int)hlit, NC_LIT, [&](int j) -> uint32_t { return wt->e[LIT_N + j][lane]; });
                    if (rc == INF_OK)
                        rc = build_table<false>(wt, lane, LIT_N, DROOT, DIST_CAP, (int)hdist, NC_DIST,
                                                [&](int j) -> uint32_t { uint32_t w = j < 8 ? dl0 : j < 16 ? dl1 : j < 24 ? dl2 : dl3; return (w >> (4 * (j & 7))) & 15u; });
                    if (rc != INF_OK) fail(rc);
                    else st = ST_RUN;
                }
            }
            // ---------------- symbols until end-of-block: one symbol per lane per step ----------------
            while (__any_sync(FULL, st == ST_RUN)) {
                if (st != ST_RUN) continue;
                s.in.refill();
                uint32_t e = wt->e[s.in.peek(LROOT)][lane];
                if (te_kind(e) == K_LINK)
                    e = wt->e[(1 << LROOT) + te_val(e) + (((uint32_t)(s.in.buf >> LROOT)) & ((1u << te_len(e)) - 1u))][lane];
                s.in.drop(te_len(e));
                const uint32_t kind = te_kind(e);
                if (kind == K_LIT) {
                    if (s.pos >= s.isize) { fail(INF_ERR_ISIZE); continue; }
                    put_literal(s, te_val(e));
                } else if (kind == K_LEN) {
                    uint32_t lbase, lextra;
                    len_sym(257u + te_val(e), lbase, lextra);
                    const uint32_t len = lbase + s.in.take(lextra);
                    s.in.refill();
                    uint32_t d = wt->e[LIT_N + s.in.peek(DROOT)][lane];
                    if (te_kind(d) == K_LINK)
                        d = wt->e[LIT_N + (1 << DROOT) + te_val(d) + (((uint32_t)(s.in.buf >> DROOT)) & ((1u << te_len(d)) - 1u))][lane];
                    if (te_kind(d) != K_LEN) { fail(INF_ERR_BAD_DIST); continue; }
                    s.in.drop(te_len(d));
                    uint32_t dbase, dextra;
                    dist_sym(te_val(d), dbase, dextra);
                    const uint32_t dist = dbase + s.in.take(dextra);
                    if (dist > s.pos) { fail(INF_ERR_DIST_TOO_FAR); continue; }
                    if (s.pos + len > s.isize) { fail(INF_ERR_ISIZE); continue; }
                    // command: [31]=0 [30:23] literals since the previous command (< 256) [22:15] len-3 [14:0] dist-1;
                    // a longer literal run is announced by a skip command [31]=1 [30:0] run
                    if (s.lit_run > 255u) { s.cmd[s.ncmd++] = 0x80000000u | s.lit_run; s.lit_run = 0; }
                    s.cmd[s.ncmd++] = (s.lit_run << 23) | ((len - 3u) << 15) | (dist - 1u);
                    s.lit_run = 0;
                    skip_bytes(s, len);
                } else if (kind == K_EOB) {
                    if (s.in.used() > s.in_bits) fail(INF_ERR_IN_OVERRUN);
                    else st = bfinal ? ST_DONE : ST_HEADER;
                } else {
                    fail(INF_ERR_BAD_LITLEN);
                }
                if (st == ST_RUN && s.in.used() > s.in_bits + 64u) fail(INF_ERR_IN_OVERRUN);  // bounds a corrupt stream's read-ahead
            }
        }
        if (b < nblocks) {
            if (s.accmask) flush_word(s);
            if (s.status == INF_OK && s.pos != s.isize) s.status = INF_ERR_ISIZE;
            status[b] = s.status;
            ncmd_out[b] = s.ncmd;
        }
    }
}

// =========================================================================================================
// phase 2: replay the match commands, then CRC-32 + footer check; one thread per block
// =========================================================================================================
__global__ void __launch_bounds__(256)
bgzf_resolve_kernel(const BlockDesc *__restrict__ blocks, uint32_t nblocks, uint8_t *__restrict__ outbuf,
                    int *__restrict__ status, const uint32_t *__restrict__ ncmd_in, const uint32_t *__restrict__ cmdbuf) {
    __shared__ uint32_t crc_tab[8 * 256];
    crc_table_column(crc_tab, (int)threadIdx.x);
    __syncthreads();
    const uint32_t b = blockIdx.x * 256u + threadIdx.x;
    if (b >= nblocks) return;
    if (status[b] != INF_OK) return;
    const BlockDesc d = blocks[b];
    uint8_t *o = outbuf + d.out_off;
    const uint32_t *c = cmdbuf + d.cmd_off;
    const uint32_t n = ncmd_in[b];
    uint32_t pos = 0;
    for (uint32_t i = 0; i < n; ++i) {
        const uint32_t cmd = c[i];
        if (cmd >> 31) { pos += cmd & 0x7fffffffu; continue; }
        pos += cmd >> 23;
        uint32_t len = ((cmd >> 15) & 0xffu) + 3u;
        const uint32_t dist = (cmd & 0x7fffu) + 1u;
        uint8_t *dst = o + pos;
        const uint8_t *src = dst - dist;
        pos += len;
        if (dist >= 4u) {
            // 4 bytes per step: the source word is complete because dist >= 4
            while (len >= 4u) {
                uintptr_t a = (uintptr_t)src;
                const uint32_t *q = reinterpret_cast<const uint32_t *>(a & ~(uintptr_t)3);
                uint32_t sh = (uint32_t)(a & 3u) * 8u;
                uint32_t v = sh ? __funnelshift_r(q[0], q[1], sh) : q[0];
                if (((uintptr_t)dst & 3u) == 0) *reinterpret_cast<uint32_t *>(dst) = v;
                else { dst[0] = (uint8_t)v; dst[1] = (uint8_t)(v >> 8); dst[2] = (uint8_t)(v >> 16); dst[3] = (uint8_t)(v >> 24); }
                src += 4; dst += 4; len -= 4u;
            }
        }
        while (len) { *dst++ = *src++; --len; }
    }
    // CRC-32 (htslib/bgzf.c:793) + ISIZE already checked in phase 1
    uint32_t crc = crc_update(crc_tab, 0xffffffffu, o, d.isize) ^ 0xffffffffu;
    if (crc != d.crc) status[b] = INF_ERR_CRC;
}

cudaError_t launch_bgzf_inflate_v2(const uint8_t *comp, const BlockDesc *blocks, uint32_t nblocks, uint8_t *out, int *status,
                                   uint32_t *work, unsigned int *counter, int sm_count, cudaStream_t stream) {
    static bool configured = false;
    const size_t smem = sizeof(WarpTab) * WARPS;
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(bgzf_decode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        configured = true;
    }
    cudaError_t e = cudaMemsetAsync(counter, 0, sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    if (nblocks == 0) return cudaSuccess;
    uint32_t *ncmd = work;
    uint32_t *cmds = work + ((nblocks + 63u) & ~63u);
    uint32_t ctas = (nblocks + THREADS - 1) / THREADS;
    if (ctas > (uint32_t)sm_count) ctas = (uint32_t)sm_count;
    bgzf_decode_kernel<<<ctas, THREADS, smem, stream>>>(comp, blocks, nblocks, out, status, ncmd, cmds, counter);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    bgzf_resolve_kernel<<<(nblocks + 255u) / 256u, 256, 0, stream>>>(blocks, nblocks, out, status, ncmd, cmds);
    return cudaGetLastError();
}

}  // namespace dh
