// bgzf_deflate.cu — BGZF compression on sm_100a.
//
// Replaces bgzf_write / deflate_block / bgzf_compress (htslib/bgzf.c:624-683, 1805-1850, 2068-2110) behind the reference's
// bgzip() table function (src/bgzip.c:209-273): the input is cut into BGZF_BLOCK_SIZE = 0xff00-byte blocks exactly like
// bgzf_write cuts it, every block becomes one gzip member (18-byte header with the BC extra field, a raw DEFLATE stream,
// CRC-32, ISIZE), the file ends with the 28-byte EOF marker.  The DEFLATE streams are NOT zlib's bit for bit (no two
// compressors agree); what is identical to the reference's output is everything a reader can observe: the inflated bytes,
// the block boundaries in inflated space (hence every virtual offset's low 16 bits), CRC and ISIZE.
//
// One WARP per block, two sweeps over the block's bytes:
//   sweep 1  LZ77, 32 positions per step: a 4-byte hash per lane, the candidate is the nearest earlier lane of the step with the
//            same hash (__match_any_sync: distances < 32, the homopolymer / quality runs of sequencing data) or else the hash
//            table's (shared memory, 4096 x u16, last position per hash); matches are extended by word compares and taken
//            greedily left to right with ballots.  Per position one role byte (literal / match start / covered) and per match
//            one word go to a scratch area; literal/length and distance histograms accumulate in shared memory.
//   build    length-limited Huffman codes (15 bits; 7 for the code-length alphabet): rank sort by the warp, two-queue tree and
//            the overflow repair by lane 0, canonical codes bit-reversed for LSB-first packing; the dynamic-block header.
//   sweep 2  every lane looks up the bits of its position (a literal's code, or length code + extra + distance code + extra,
//            <= 48 bits), a warp scan gives the bit offsets, the bits are OR-ed into a 512-byte shared-memory window and the
//            finished words leave with coalesced 32-bit stores.  CRC-32 of the input by the same warp (lane-interleaved Horner
//            form, as in the inflate's phase 2).
// A block that would not shrink is written as a stored block (BTYPE 0), like zlib does.  Blocks land in fixed 64 KB slots; a scan
// + gather pass packs them back to back.
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "bgzf_deflate.cuh"
#include "crc32_core.cuh"

namespace dh {
namespace dfl {

constexpr unsigned FULL = 0xffffffffu;
#ifndef DH_DFL_HASH_BITS
#define DH_DFL_HASH_BITS 11
#endif
#ifndef DH_DFL_CTAS
#define DH_DFL_CTAS 4
#endif
constexpr int WARPS = 8;                      // per CTA; DH_DFL_CTAS CTAs per SM: the kernel waits on L1/L2 loads (ncu r2o: 4.0 of 10.75 cycles
                                              // per instruction on the long scoreboard at 16 warps per SM), so shared memory is kept under 7 KB per warp
constexpr int HASH_BITS = DH_DFL_HASH_BITS;
constexpr uint32_t MIN_MATCH = 4, MAX_MATCH = 258;
constexpr int NLL = 288, ND = 32, NCL = 19;
constexpr int CRCW = 5 * 256 + 32;            // the same tables as the inflate's phase 2 (bgzf_inflate3.cu)
__device__ uint32_t g_crc_d[CRCW];

// per-warp shared memory
struct WarpSmem {
    union {
        uint16_t htab[1 << HASH_BITS];        // sweep 1
        struct {                              // Huffman build (after sweep 1)
            uint32_t f[2 * NLL];              // weights: sorted leaves [0, n), internal nodes [n, 2n - 1)
            uint16_t s[NLL];                  // symbol of sorted leaf i
            uint16_t parent[2 * NLL];
            uint8_t depth[2 * NLL];
            uint16_t rle[NLL + ND + 4];       // code-length symbols: sym | extra << 5 (alive until the header is written)
        } hb;
        uint32_t win[128];                    // sweep 2: output window (ring of 32-bit words); lies over hb.f, which is dead by then
    };
    // the histograms turn into the code tables once the lengths are known and the block's size has been summed up
    union { uint32_t hist_ll[NLL]; uint32_t code_ll[NLL]; };      // code tables: reversed code | length << 16
    union { uint32_t hist_d[ND]; uint32_t code_d[ND]; };
    union { uint32_t hist_cl[NCL + 1]; uint32_t code_cl[NCL + 1]; };
    uint8_t len_ll[NLL], len_d[ND], len_cl[NCL + 1];
};
static_assert(sizeof(WarpSmem) * WARPS * DH_DFL_CTAS <= 227 * 1024, "shared memory of the resident CTAs");

__device__ __forceinline__ uint32_t load32u(const uint8_t *p) {     // unaligned 32-bit load (may touch up to 7 bytes behind p)
    const uint32_t *a = reinterpret_cast<const uint32_t *>(reinterpret_cast<uintptr_t>(p) & ~(uintptr_t)3);
    const uint32_t sh = (uint32_t)(reinterpret_cast<uintptr_t>(p) & 3u) * 8u;
    const uint32_t lo = a[0];
    if (sh == 0) return lo;
    return __funnelshift_r(lo, a[1], sh);
}
__device__ __forceinline__ uint32_t bitrev(uint32_t code, uint32_t len) { return __brev(code) >> (32u - len); }

// RFC 1951 length / distance codes, arithmetically (l = length - 3, d = distance - 1)
__device__ __forceinline__ void len_code(uint32_t l, uint32_t &sym, uint32_t &nextra, uint32_t &extra) {
    if (l < 8u) { sym = 257u + l; nextra = 0; extra = 0; return; }
    if (l == 255u) { sym = 285u; nextra = 0; extra = 0; return; }
    const uint32_t nb = 31u - (uint32_t)__clz(l);
    sym = 257u + 4u * (nb - 1u) + ((l >> (nb - 2u)) & 3u);
    nextra = nb - 2u;
    extra = l & ((1u << nextra) - 1u);
}
__device__ __forceinline__ void dist_code(uint32_t d, uint32_t &sym, uint32_t &nextra, uint32_t &extra) {
    if (d < 4u) { sym = d; nextra = 0; extra = 0; return; }
    const uint32_t nb = 31u - (uint32_t)__clz(d);
    sym = 2u * nb + ((d >> (nb - 1u)) & 1u);
    nextra = nb - 1u;
    extra = d & ((1u << nextra) - 1u);
}

// ---------------------------------------------------------------------------------------------------------------------
// Length-limited Huffman code lengths for hist[0..nsym) -> lens[0..nsym).  At least two symbols get a code (zlib does the same:
// a decoder may refuse a code set with a single code of length 0).  Warp-cooperative; sm.hb is the scratch.
// ---------------------------------------------------------------------------------------------------------------------
__device__ void build_lengths(WarpSmem &sm, uint32_t *hist, uint8_t *lens, int nsym, int maxbits, uint32_t lane) {
    // force two used symbols
    uint32_t used = 0;
    for (int i = (int)lane; i < nsym; i += 32) used += hist[i] != 0u;
#pragma unroll
    for (int sft = 16; sft; sft >>= 1) used += __shfl_xor_sync(FULL, used, sft);
    if (used < 2u && lane == 0) {
        for (int i = 0; i < nsym && used < 2u; ++i)
            if (hist[i] == 0u) { hist[i] = 1u; ++used; }
    }
    __syncwarp();
    // compact + rank sort by (weight, symbol)
    int n = 0;
    for (int i0 = 0; i0 < nsym; i0 += 32) {
        const int i = i0 + (int)lane;
        const bool nz = i < nsym && hist[i] != 0u;
        const uint32_t m = __ballot_sync(FULL, nz);
        if (nz) {
            const uint32_t w = hist[i];
            // rank = number of used symbols that sort in front of (w, i)
            int rank = 0;
            for (int j = 0; j < nsym; ++j) {
                const uint32_t wj = hist[j];
                rank += (wj != 0u) && (wj < w || (wj == w && j < i));
            }
            sm.hb.f[rank] = w;
            sm.hb.s[rank] = (uint16_t)i;
        }
        n += __popc(m);
    }
    __syncwarp();
    for (int i = (int)lane; i < nsym; i += 32) lens[i] = 0;
    __syncwarp();
    if (lane == 0) {
        // two-queue Huffman: leaves 0..n-1 (ascending weight), internal nodes n..2n-2 appear in ascending weight too
        int li = 0, ij = n, k = n;
        while (k < 2 * n - 1) {
            int a, b;
            if (li < n && (ij >= k || sm.hb.f[li] <= sm.hb.f[ij])) a = li++; else a = ij++;
            if (li < n && (ij >= k || sm.hb.f[li] <= sm.hb.f[ij])) b = li++; else b = ij++;
            sm.hb.f[k] = sm.hb.f[a] + sm.hb.f[b];
            sm.hb.parent[a] = (uint16_t)k;
            sm.hb.parent[b] = (uint16_t)k;
            ++k;
        }
        sm.hb.depth[2 * n - 2] = 0;
        for (int v = 2 * n - 3; v >= 0; --v) sm.hb.depth[v] = (uint8_t)(sm.hb.depth[sm.hb.parent[v]] + 1u);
        // limit to maxbits: move overflowing leaves up and pay with the deepest leaf that still has room (zlib's gen_bitlen repair)
        int bl[17];
        for (int b = 0; b <= 16; ++b) bl[b] = 0;
        int overflow = 0;
        for (int i = 0; i < n; ++i) {
            int d = sm.hb.depth[i];
            if (d > maxbits) { d = maxbits; ++overflow; }
            ++bl[d];
        }
        while (overflow > 0) {
            int bits = maxbits - 1;
            while (bl[bits] == 0) --bits;
            --bl[bits];
            bl[bits + 1] += 2;
            --bl[maxbits];
            overflow -= 2;
        }
        // hand the lengths out again, longest to the rarest symbol
        int i = 0;
        for (int bits = maxbits; bits >= 1; --bits)
            for (int c = bl[bits]; c > 0; --c) lens[sm.hb.s[i++]] = (uint8_t)bits;
    }
    __syncwarp();
}

// canonical codes (RFC 1951 3.2.2), stored bit-reversed with the length in the high half
__device__ void assign_codes(const uint8_t *lens, uint32_t *codes, int nsym, uint32_t lane) {
    if (lane == 0) {
        uint32_t cnt[16], next[16];
        for (int b = 0; b < 16; ++b) cnt[b] = 0;
        for (int i = 0; i < nsym; ++i) ++cnt[lens[i]];
        cnt[0] = 0;
        uint32_t code = 0;
        next[0] = 0;
        for (int b = 1; b < 16; ++b) { code = (code + cnt[b - 1]) << 1; next[b] = code; }
        for (int i = 0; i < nsym; ++i) {
            const uint32_t l = lens[i];
            codes[i] = l ? (bitrev(next[l]++, l) | (l << 16)) : 0u;
        }
    }
    __syncwarp();
}

// lane 0's bit writer for the block header: whole words go straight to the slot, the unfinished word is handed to the window
struct BitWriter {
    uint32_t *out32;
    uint64_t acc;
    uint32_t nacc, widx;
    __device__ __forceinline__ void put(uint32_t v, uint32_t nb) {
        acc |= (uint64_t)v << nacc;
        nacc += nb;
        if (nacc >= 32u) { out32[widx++] = (uint32_t)acc; acc >>= 32; nacc -= 32u; }
    }
};

// CRC-32 of p[0..n) by one warp (zlib's crc32, htslib/bgzf.c:662): lane 0 eats the bytes in front of the first 128-byte boundary
// and behind the last whole 128-byte row, the rows go through the lane-interleaved Horner form H = H * x^1024 + w.
__device__ uint32_t warp_crc32(const uint8_t *p, uint32_t n, uint32_t lane) {
    auto tab = [&](uint32_t i) -> uint32_t { return __ldg(g_crc_d + i); };
    const uint8_t *end = p + n;
    const uint8_t *a0 = reinterpret_cast<const uint8_t *>((reinterpret_cast<uintptr_t>(p) + 127u) & ~(uintptr_t)127);
    if (a0 > end) a0 = end;
    uint32_t r = 0xffffffffu;
    if (lane == 0)
        for (const uint8_t *q = p; q < a0; ++q) r = tab(4 * 256 + ((r ^ *q) & 0xffu)) ^ (r >> 8);
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
    if (lane == 0)
        for (const uint8_t *q = a0 + ((size_t)rows << 7); q < end; ++q) r = tab(4 * 256 + ((r ^ *q) & 0xffu)) ^ (r >> 8);
    return __shfl_sync(FULL, r, 0) ^ 0xffffffffu;
}

// Scratch of one block: role bytes [0, 0xff00), then the match words (<= 0xff00 / 4 + 1)
constexpr uint32_t ROLE_BYTES = 0xff00u;
constexpr uint32_t SCRATCH_STRIDE = ROLE_BYTES + (0xff00u / MIN_MATCH + 64u) * 4u;

// =====================================================================================================================
// the kernel: warp w of the grid compresses block w of the launch
// =====================================================================================================================
__global__ void __launch_bounds__(WARPS * 32, DH_DFL_CTAS)
bgzf_deflate_kernel(const uint8_t *__restrict__ in, uint64_t in_bytes, uint32_t nblocks, uint8_t *__restrict__ slots,
                    uint8_t *__restrict__ scratch, uint32_t *__restrict__ csize, int level) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    const uint32_t b = blockIdx.x * WARPS + wid;
    if (b >= nblocks) return;
    WarpSmem &sm = reinterpret_cast<WarpSmem *>(smem_raw)[wid];
    const uint64_t start = (uint64_t)b * DEFLATE_BLOCK_IN;
    const uint32_t n = (uint32_t)(in_bytes - start < (uint64_t)DEFLATE_BLOCK_IN ? in_bytes - start : (uint64_t)DEFLATE_BLOCK_IN);
    const uint8_t *src = in + start;
    uint8_t *slot = slots + (size_t)b * DEFLATE_SLOT_STRIDE;
    uint8_t *roles = scratch + (size_t)b * SCRATCH_STRIDE;
    uint32_t *matches = reinterpret_cast<uint32_t *>(roles + ROLE_BYTES);
    const uint32_t lt = (1u << lane) - 1u;

    const uint32_t crc = warp_crc32(src, n, lane);

    // ---------------- sweep 1: LZ77 ----------------
    for (uint32_t i = lane; i < (1u << HASH_BITS); i += 32u) sm.htab[i] = 0xffffu;
    for (uint32_t i = lane; i < (uint32_t)NLL; i += 32u) sm.hist_ll[i] = 0u;
    sm.hist_d[lane] = 0u;
    if (lane < (uint32_t)NCL + 1u) sm.hist_cl[lane] = 0u;
    __syncwarp();
    uint32_t cov = 0, nmatch = 0, extra_bits = 0;
    const bool lz = level > 0;
    for (uint32_t base = 0; base < n && lz; base += 32u) {
        const uint32_t p = base + lane;
        const bool valid = p < n;
        const bool can = p + MIN_MATCH <= n;
        const uint32_t w = valid ? load32u(src + p) : 0u;
        const uint32_t h = (w * 0x9E3779B1u) >> (32 - HASH_BITS);
        uint32_t cand = can ? (uint32_t)sm.htab[h] : 0xffffu;
        const uint32_t grp = __match_any_sync(FULL, can ? h : (0x10000u + lane));
        if (can && (grp & lt)) cand = base + (31u - (uint32_t)__clz(grp & lt));      // nearest earlier lane of this step
        uint32_t len = 0, dist = 0;
        if (cand != 0xffffu) {
            const uint32_t maxlen = n - p < MAX_MATCH ? n - p : MAX_MATCH;
            const uint8_t *a = src + cand, *c = src + p;
            while (len < maxlen) {
                const uint32_t x = load32u(a + len) ^ load32u(c + len);
                if (x) { len += (uint32_t)(__ffs((int)x) - 1) >> 3; break; }
                len += 4u;
            }
            if (len > maxlen) len = maxlen;
            dist = p - cand;
            if (len < MIN_MATCH || dist > 32768u) len = 0;     // DEFLATE's window
        }
        __syncwarp();
        if (can && !(grp & ~lt & ~(1u << lane))) sm.htab[h] = (uint16_t)p;              // the last lane of a hash group records it
        // greedy choice, left to right
        uint32_t role = !valid ? 3u : (p < cov ? 1u : 0u);
        uint32_t avail = __ballot_sync(FULL, len != 0u && p >= cov);
        while (avail) {
            const int L = __ffs((int)avail) - 1;
            const uint32_t pL = base + (uint32_t)L, lenL = __shfl_sync(FULL, len, L);
            if ((int)lane == L) role = 2u;
            else if (valid && p > pL && p < pL + lenL) role = 1u;
            cov = pL + lenL;
            const uint32_t span = cov - base;
            avail &= span >= 32u ? 0u : ~((1u << span) - 1u);
        }
        if (valid) roles[p] = (uint8_t)role;
        const uint32_t mm = __ballot_sync(FULL, role == 2u);
        if (role == 2u) {
            matches[nmatch + __popc(mm & lt)] = ((len - 3u) << 16) | (dist - 1u);
            uint32_t ls, le, lx, ds, de, dx;
            len_code(len - 3u, ls, le, lx);
            dist_code(dist - 1u, ds, de, dx);
            atomicAdd(&sm.hist_ll[ls], 1u);
            atomicAdd(&sm.hist_d[ds], 1u);
            extra_bits += le + de;
        } else if (role == 0u) atomicAdd(&sm.hist_ll[w & 0xffu], 1u);
        nmatch += (uint32_t)__popc(mm);
        __syncwarp();
    }
#pragma unroll
    for (int sft = 16; sft; sft >>= 1) extra_bits += __shfl_xor_sync(FULL, extra_bits, sft);
    if (lane == 0) sm.hist_ll[256] = 1u;      // end of block
    __syncwarp();

    // ---------------- Huffman codes + the size of a dynamic block ----------------
    uint32_t nll = 257, nd = 1, ncl = 4, nrle = 0;
    uint64_t total_bits = ~0ull;
    if (lz) {
        build_lengths(sm, sm.hist_ll, sm.len_ll, 286, 15, lane);
        build_lengths(sm, sm.hist_d, sm.len_d, 30, 15, lane);
        if (lane == 0) {
            nll = 286; while (nll > 257u && sm.len_ll[nll - 1u] == 0) --nll;
            nd = 30;   while (nd > 1u && sm.len_d[nd - 1u] == 0) --nd;
            // run-length code of the nll + nd code lengths (RFC 1951 3.2.7): 16 = repeat previous 3-6, 17 = 3-10 zeros, 18 = 11-138 zeros
            const uint32_t tot = nll + nd;
            uint32_t i = 0;
            while (i < tot) {
                const uint32_t v = i < nll ? sm.len_ll[i] : sm.len_d[i - nll];
                uint32_t run = 1;
                while (i + run < tot && (i + run < nll ? sm.len_ll[i + run] : sm.len_d[i + run - nll]) == v) ++run;
                i += run;
                if (v == 0) {
                    while (run >= 11u) { const uint32_t c = run < 138u ? run : 138u; sm.hb.rle[nrle++] = (uint16_t)(18u | ((c - 11u) << 5)); run -= c; }
                    if (run >= 3u) { sm.hb.rle[nrle++] = (uint16_t)(17u | ((run - 3u) << 5)); run = 0; }
                    while (run--) sm.hb.rle[nrle++] = 0;
                } else {
                    sm.hb.rle[nrle++] = (uint16_t)v; --run;
                    while (run >= 3u) { const uint32_t c = run < 6u ? run : 6u; sm.hb.rle[nrle++] = (uint16_t)(16u | ((c - 3u) << 5)); run -= c; }
                    while (run--) sm.hb.rle[nrle++] = (uint16_t)v;
                }
            }
            for (uint32_t k = 0; k < nrle; ++k) ++sm.hist_cl[sm.hb.rle[k] & 31u];
        }
        nll = __shfl_sync(FULL, nll, 0); nd = __shfl_sync(FULL, nd, 0); nrle = __shfl_sync(FULL, nrle, 0);
        __syncwarp();
        // the run-length symbols survive the next build: it uses hb.f / s / parent / depth only
        build_lengths(sm, sm.hist_cl, sm.len_cl, NCL, 7, lane);
        uint64_t bits = 0;
        for (uint32_t i = lane; i < 286u; i += 32u) bits += (uint64_t)sm.hist_ll[i] * sm.len_ll[i];
        if (lane < 30u) bits += (uint64_t)sm.hist_d[lane] * sm.len_d[lane];
        if (lane < (uint32_t)NCL) bits += (uint64_t)sm.hist_cl[lane] * (sm.len_cl[lane] + (lane == 16u ? 2u : lane == 17u ? 3u : lane == 18u ? 7u : 0u));
#pragma unroll
        for (int sft = 16; sft; sft >>= 1) bits += __shfl_xor_sync(FULL, bits, sft);
        __syncwarp();
        assign_codes(sm.len_ll, sm.code_ll, 286, lane);      // the histograms are gone from here on
        assign_codes(sm.len_d, sm.code_d, 30, lane);
        assign_codes(sm.len_cl, sm.code_cl, NCL, lane);
        if (lane == 0) {
            const uint8_t order[NCL] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
            ncl = NCL; while (ncl > 4u && sm.len_cl[order[ncl - 1u]] == 0) --ncl;
        }
        ncl = __shfl_sync(FULL, ncl, 0);
        total_bits = bits + extra_bits + 3u + 14u + 3u * ncl;
    }
    uint32_t E;          // bytes behind slot + 16: the two BSIZE bytes + the DEFLATE stream
    uint32_t *out32 = reinterpret_cast<uint32_t *>(slot + 16);
    if (!lz || (total_bits + 7u) / 8u >= (uint64_t)n + 5u) {
        // ---------------- stored block ----------------
        uint8_t *q = slot + 18;
        if (lane == 0) { q[0] = 1; q[1] = (uint8_t)n; q[2] = (uint8_t)(n >> 8); q[3] = (uint8_t)~n; q[4] = (uint8_t)(~n >> 8); }
        for (uint32_t i = lane; i < n; i += 32u) q[5 + i] = src[i];
        E = 2u + 5u + n;
    } else {
        // ---------------- dynamic block: header by lane 0 ----------------
        for (uint32_t i = lane; i < 128u; i += 32u) sm.win[i] = 0u;
        __syncwarp();
        uint32_t bitpos = 0, flushed = 0;
        if (lane == 0) {
            const uint8_t order[NCL] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
            BitWriter bw{out32, 0ull, 16u, 0u};
            bw.put(1u, 1); bw.put(2u, 2);
            bw.put(nll - 257u, 5); bw.put(nd - 1u, 5); bw.put(ncl - 4u, 4);
            for (uint32_t i = 0; i < ncl; ++i) bw.put(sm.len_cl[order[i]], 3);
            for (uint32_t k = 0; k < nrle; ++k) {
                const uint32_t e = sm.hb.rle[k], s = e & 31u, c = sm.code_cl[s];
                bw.put(c & 0xffffu, c >> 16);
                if (s == 16u) bw.put(e >> 5, 2); else if (s == 17u) bw.put(e >> 5, 3); else if (s == 18u) bw.put(e >> 5, 7);
            }
            sm.win[bw.widx & 127u] = (uint32_t)bw.acc;
            bitpos = bw.widx * 32u + bw.nacc;
            flushed = bw.widx;
        }
        bitpos = __shfl_sync(FULL, bitpos, 0);
        flushed = __shfl_sync(FULL, flushed, 0);
        __syncwarp();
        // ---------------- sweep 2: the symbols ----------------
        uint32_t mi = 0;
        for (uint32_t base = 0; base <= n; base += 32u) {
            const uint32_t p = base + lane;
            uint64_t val = 0;
            uint32_t nb = 0;
            uint32_t role = 3u;
            if (p < n) role = roles[p];
            const uint32_t mm = __ballot_sync(FULL, role == 2u);
            if (role == 0u) {
                const uint32_t c = sm.code_ll[src[p]];
                val = c & 0xffffu; nb = c >> 16;
            } else if (role == 2u) {
                const uint32_t m = matches[mi + __popc(mm & lt)];
                uint32_t ls, le, lx, ds, de, dx;
                len_code(m >> 16, ls, le, lx);
                dist_code(m & 0xffffu, ds, de, dx);
                const uint32_t c1 = sm.code_ll[ls], c2 = sm.code_d[ds];
                const uint32_t n1 = c1 >> 16, n2 = c2 >> 16;
                val = (uint64_t)(c1 & 0xffffu) | ((uint64_t)lx << n1) | ((uint64_t)(c2 & 0xffffu) << (n1 + le)) | ((uint64_t)dx << (n1 + le + n2));
                nb = n1 + le + n2 + de;
            } else if (p == n) {                      // end of block
                const uint32_t c = sm.code_ll[256];
                val = c & 0xffffu; nb = c >> 16;
            }
            mi += (uint32_t)__popc(mm);
            uint32_t inc = nb;
#pragma unroll
            for (int sft = 1; sft < 32; sft <<= 1) {
                const uint32_t t = __shfl_up_sync(FULL, inc, sft);
                if (lane >= (uint32_t)sft) inc += t;
            }
            if (nb) {
                const uint32_t q = bitpos + inc - nb, word = q >> 5, sh = q & 31u;
                const uint32_t w0 = (uint32_t)(val << sh);
                const uint64_t v1 = sh ? (val >> (32u - sh)) : (val >> 32);
                if (w0) atomicOr(&sm.win[word & 127u], w0);
                if ((uint32_t)v1) atomicOr(&sm.win[(word + 1u) & 127u], (uint32_t)v1);
                if ((uint32_t)(v1 >> 32)) atomicOr(&sm.win[(word + 2u) & 127u], (uint32_t)(v1 >> 32));
            }
            bitpos += __shfl_sync(FULL, inc, 31);
            __syncwarp();
            const uint32_t nw = bitpos >> 5;
            for (uint32_t i = flushed + lane; i < nw; i += 32u) { out32[i] = sm.win[i & 127u]; sm.win[i & 127u] = 0u; }
            flushed = nw;
            __syncwarp();
        }
        if ((bitpos & 31u) && lane == 0) out32[bitpos >> 5] = sm.win[(bitpos >> 5) & 127u];
        E = (bitpos + 7u) >> 3;
    }
    __syncwarp();
    // ---------------- BGZF header and footer (htslib/bgzf.c:640-676) ----------------
    const uint32_t bsize = 16u + E + 8u;
    if (lane < 18u) {
        const uint8_t hdr[16] = {31, 139, 8, 4, 0, 0, 0, 0, 0, 255, 6, 0, 'B', 'C', 2, 0};
        slot[lane] = lane < 16u ? hdr[lane] : (uint8_t)((bsize - 1u) >> (8u * (lane - 16u)));
    }
    if (lane < 8u) slot[16u + E + lane] = (uint8_t)((lane < 4u ? crc : n) >> (8u * (lane & 3u)));
    if (lane == 0) csize[b] = bsize;
}

// ---------------------------------------------------------------------------------------------------------------------
// pack the slots back to back: offsets = exclusive scan of the block sizes (one CTA), then one CTA per block copies
// ---------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) deflate_scan_kernel(const uint32_t *__restrict__ csize, uint32_t nblocks, uint64_t *__restrict__ offs) {
    __shared__ uint64_t part[1024];
    const uint32_t t = threadIdx.x, per = (nblocks + 1023u) / 1024u;
    const uint32_t lo = t * per, hi = lo + per < nblocks ? lo + per : nblocks;
    uint64_t sum = 0;
    for (uint32_t i = lo; i < hi; ++i) sum += csize[i];
    part[t] = sum;
    __syncthreads();
    for (uint32_t d = 1; d < 1024u; d <<= 1) {
        const uint64_t v = t >= d ? part[t - d] : 0;
        __syncthreads();
        part[t] += v;
        __syncthreads();
    }
    uint64_t run = t ? part[t - 1] : 0;
    for (uint32_t i = lo; i < hi; ++i) { offs[i] = run; run += csize[i]; }
    if (t == 1023u) offs[nblocks] = part[1023];
}

__global__ void __launch_bounds__(256) deflate_gather_kernel(const uint8_t *__restrict__ slots, const uint32_t *__restrict__ csize,
                                                             const uint64_t *__restrict__ offs, uint8_t *__restrict__ out) {
    const uint32_t b = blockIdx.x;
    const uint8_t *src = slots + (size_t)b * DEFLATE_SLOT_STRIDE;
    uint8_t *dst = out + offs[b];
    const uint32_t n = csize[b];
    // destination-aligned 4-byte stores; the head and tail bytes one by one
    const uint32_t head = (uint32_t)((4u - (reinterpret_cast<uintptr_t>(dst) & 3u)) & 3u);
    const uint32_t h = head < n ? head : n;
    if (threadIdx.x < h) dst[threadIdx.x] = src[threadIdx.x];
    const uint32_t words = (n - h) >> 2;
    uint32_t *d32 = reinterpret_cast<uint32_t *>(dst + h);
    for (uint32_t i = threadIdx.x; i < words; i += 256u) d32[i] = load32u(src + h + 4u * i);
    const uint32_t t0 = h + (words << 2);
    if (t0 + threadIdx.x < n && threadIdx.x < 4u) dst[t0 + threadIdx.x] = src[t0 + threadIdx.x];
}

}  // namespace dfl

size_t deflate_scratch_bytes(uint32_t nblocks) { return (size_t)nblocks * dfl::SCRATCH_STRIDE + 256u; }

static void build_tables(uint32_t *t) {      // as bgzf_inflate3.cu's build_crc3_tables
    uint32_t slice[8 * 256], x2n[32];
    for (int i = 0; i < 256; ++i) crc_table_column(slice, i);
    crc_x2n_init(x2n);
    for (int k = 0; k < 4; ++k)
        for (int i = 0; i < 256; ++i) t[k * 256 + i] = crc_shift(x2n, slice[i], 127u - (uint32_t)k);
    for (int i = 0; i < 256; ++i) t[4 * 256 + i] = slice[i];
    for (int l = 0; l < 32; ++l) t[5 * 256 + l] = crc_shift(x2n, 0x80000000u, 128u - 4u * (uint32_t)l);
}

cudaError_t launch_bgzf_deflate(const uint8_t *in, uint64_t in_bytes, uint32_t nblocks, int level, uint8_t *slots, uint8_t *scratch,
                                uint32_t *csize, uint64_t *offs, uint8_t *out, cudaStream_t stream) {
    using namespace dfl;
    static std::mutex mu;
    static bool ready[64] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    {
        std::lock_guard<std::mutex> lk(mu);
        if (!ready[dev]) {
            static uint32_t host_tabs[CRCW];
            build_tables(host_tabs);
            if ((e = cudaMemcpyToSymbol(g_crc_d, host_tabs, sizeof host_tabs)) != cudaSuccess) return e;
            if ((e = cudaFuncSetAttribute(bgzf_deflate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(sizeof(WarpSmem) * WARPS))) != cudaSuccess) return e;
            ready[dev] = true;
        }
    }
    if (nblocks == 0) return cudaMemsetAsync(offs, 0, sizeof(uint64_t), stream);
    bgzf_deflate_kernel<<<(nblocks + WARPS - 1) / WARPS, WARPS * 32, sizeof(WarpSmem) * WARPS, stream>>>(in, in_bytes, nblocks, slots, scratch, csize, level);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    deflate_scan_kernel<<<1, 1024, 0, stream>>>(csize, nblocks, offs);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    deflate_gather_kernel<<<nblocks, 256, 0, stream>>>(slots, csize, offs, out);
    return cudaGetLastError();
}

}  // namespace dh
