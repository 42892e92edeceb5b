// bgzf_resolve5.cuh — inflate phase 2, fifth generation: the per-warp body of bgzf_resolve5_kernel (bgzf_inflate3.cu).
//
// In a header of its own so that the SAME source is
//   * run by every warp of the kernel (one warp per BGZF block), and
//   * compiled for the host behind a SIMT emulation (tests/host/resolve5_host.cpp: 32 lanes as coroutines, every warp intrinsic a
//     rendezvous that also checks that all 32 lanes arrived at the same intrinsic), so the CPU box unit-tests the kernel's logic --
//     window slides, chunk and row bookkeeping, period folds, head / tail units, the CRC catch-up -- against zlib
//     (tests/test_resolve5_host.py).  The host build is test infrastructure: the product never runs it.
//
// =========================================================================================================
// phase 2, fifth generation: one warp per block, BYTE-parallel, a window of recent output in shared memory, 16-byte stores
// =========================================================================================================
// What the two earlier forms showed: the warp-per-block kernel spends 2.4 warp instructions per output byte on
// command-parallel byte loops (one lane per command: ragged lengths, dependency rounds, every long copy a serial whole-warp
// episode) and moves every back-reference byte through L2/DRAM as 1-byte LDG/STG (3.8x DRAM read amplification); the
// CTA-per-block kernel with the whole 64 KB window in shared memory fixes the traffic but keeps 3 blocks per SM in flight and
// loses 3x.  This form keeps the warp per block (32 blocks per SM in flight) and turns the work around: lanes own OUTPUT BYTES.
//   * 32 commands per step as before (one per lane; warp scans give each its destination); every command is two segments
//     (its literals, its match);
//   * the step's output span is done in chunks of <= 512 bytes: the segments that touch the chunk set a bit at their first byte
//     in a 512-bit start mask and write a descriptor {start, source position, source ADDRESS} at their ordinal (two ballots);
//   * rows of 32 bytes, one byte per lane: ordinal = population count of the start mask up to the byte, one 16-byte descriptor
//     load, source address = descriptor base + offset -- ONE generic load whatever the source is: the literal stream, the
//     warp's window of its last 1.5-3.5 KB of output in shared memory (linear, slid down when full), or the output buffer in global
//     memory for anything older; a byte whose source lies in the same row takes it from the owning lane's register by shuffle
//     (every earlier row is complete, so no other dependency exists); overlapping matches (distance < length) are folded to
//     their first period by a multiply-shift modulo;
//   * finished 16-byte units leave the window with st.global.v4 (north-star item (b)), aligned to the output buffer; only the
//     first and last partial unit of a block use byte stores;
//   * the CRC-32 (same lane-interleaved Horner form as resolve3) reads its 128-byte rows from the window as they complete: the
//     output is never read back from global memory.
// DRAM traffic = command + literal streams in, inflated bytes out, far sources.
#pragma once
#include <stdint.h>

#include "bgzf_inflate.cuh"
#include "crc32_core.cuh"
#include "inflate_core.cuh"

namespace dh {

// CRC tables of phase 2 (built on the host, uploaded once per device): [k*256 + i], k = 0..3: byte i at byte position k of a 32-bit
// register advanced by 128 bytes (multiplication by x^1024 mod P, split by byte); k = 4: the classic byte table;
// [1280 + l]: x^(8 * (128 - 4l)) for lane l.
constexpr int CRC3_WORDS = 5 * 256 + 32;
inline void build_crc3_tables(uint32_t *t) {
    uint32_t slice[8 * 256], x2n[32];
    for (int i = 0; i < 256; ++i) crc_table_column(slice, i);
    crc_x2n_init(x2n);
    for (int k = 0; k < 4; ++k)
        for (int i = 0; i < 256; ++i) t[k * 256 + i] = crc_shift(x2n, slice[i], 127u - (uint32_t)k);
    for (int i = 0; i < 256; ++i) t[4 * 256 + i] = slice[i];
    for (int l = 0; l < 32; ++l) t[5 * 256 + l] = crc_shift(x2n, 0x80000000u, 128u - 4u * (uint32_t)l);
}

#if defined(__CUDACC__)
#define DH_R5_DEV __device__ __forceinline__
DH_R5_DEV uint32_t r5_lanemask_le(uint32_t) {
    uint32_t m;
    asm volatile("mov.u32 %0, %%lanemask_le;" : "=r"(m));
    return m;
}
DH_R5_DEV void r5_prefetch(const uint8_t *p, int mode) {      // 1: into L2, else into L1
    if (mode == 1) asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
    else asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
}
#else
#define DH_R5_DEV static inline
DH_R5_DEV uint32_t r5_lanemask_le(uint32_t lane) { return 0xffffffffu >> (31u - lane); }
DH_R5_DEV void r5_prefetch(const uint8_t *, int) {}
#endif

constexpr uint32_t R5_CH = 512, R5_BUF = 3456, R5_KEEP = 1536, R5_LST = 512;
constexpr uint32_t R5_WARP_SMEM = R5_BUF + 64 * 16 + 64 * 4 + 64 + R5_LST;
constexpr uint32_t R5_PER = 1u << 31;
#ifndef DH_R5_ROWS
#define DH_R5_ROWS 2     // r2u/r2v: 2 rows 95.2 ms, 4 rows 94.9 ms (and spills): the loads of two rows in flight are enough
#endif
constexpr int R5_ROWS = DH_R5_ROWS;    // rows (of 32 bytes) per round of the row loop

// One warp resolves one block: d = the block, n = its command count, tabs_s = the CRC tables (CRC3 layout, bgzf_inflate3.cu) in shared
// memory, buf = R5_WARP_SMEM bytes of shared memory private to the warp (16-byte aligned), *status_b receives INF_ERR_ISIZE / INF_ERR_CRC.
DH_R5_DEV void resolve5_warp(const BlockDesc d, const uint32_t n, uint8_t *outbuf, const uint8_t *work, int *status_b, const uint32_t *tabs_s,
                             uint8_t *buf, const uint32_t lane, const int pf_mode) {
    constexpr unsigned FULL = 0xffffffffu;
    auto tab = [&](uint32_t i) -> uint32_t { return tabs_s[i]; };
    uint4 *desc = reinterpret_cast<uint4 *>(buf + R5_BUF);
    uint32_t *per = reinterpret_cast<uint32_t *>(buf + R5_BUF + 64 * 16);
    uint32_t *mask = per + 64;
    uint8_t *lstage = reinterpret_cast<uint8_t *>(mask + 16);     // the step's literals, staged by one coalesced 16-byte load per lane
    uint8_t *o = outbuf + d.out_off;
    const uint8_t *lit = work + d.cmd_off;
    const uint32_t area = inflate_work_cap(d.isize);
    const uint32_t *cmd_top = reinterpret_cast<const uint32_t *>(work + d.cmd_off + area) - 1;
    // q = p + A: position in the 128-byte aligned frame of the output buffer; the window holds q in [wlo, wlo + R5_BUF), wlo a
    // multiple of 128, so its 16-byte units and 128-byte rows are the buffer's
    const uint32_t A = (uint32_t)((uintptr_t)o & 127u);
    uint8_t *ob = o - A;
    const uint32_t endq = A + d.isize;
    const uint32_t fu0 = (A + 15u) >> 4;          // first unit that lies wholly inside the block
    uint32_t fu = fu0;                            // next unit to store
    bool head_pending = (A & 15u) != 0u;          // bytes in front of unit fu0 leave by byte stores
    const uint32_t head_end = (fu0 << 4) < endq ? (fu0 << 4) : endq;
    uint32_t a0q = (A + 127u) & ~127u;            // CRC: bytes [A, a0q) by lane 0, rows [a0q, rows_end), bytes [rows_end, endq) by lane 0
    if (a0q > endq) a0q = endq;
    const uint32_t rows_end = a0q + ((endq - a0q) & ~127u);
    uint32_t cq = A, r = 0xffffffffu, H = 0;
    bool first_row = true;
    uint32_t pos = 0, lpos = 0, wlo = 0;
    const uint32_t le_mask = r5_lanemask_le(lane);              // (read once: the compiler recomputed a shift form in every row)
    const uint32_t lt_mask = le_mask >> 1;
    uint32_t c_next = lane < n ? cmd_top[-(int)lane] : 0x80000000u;
    // the literals of a step are staged by one 16-byte load per lane, issued a whole step ahead (as soon as the scan has said where the
    // next step's literals begin) and stored to shared memory when that step starts
    auto stage_load = [&](uint32_t lp) -> uint4 {
        const uint32_t lso = (lp & ~15u) + 16u * lane;             // offset in the block's work area (16-byte aligned like the area)
        uint4 x = make_uint4(0u, 0u, 0u, 0u);
        if (lso + 16u <= area) x = *reinterpret_cast<const uint4 *>(lit + lso);
        return x;
    };
    uint4 lst = stage_load(0u);

    for (uint32_t k0 = 0; k0 < n; k0 += 32u) {
        const uint32_t c = c_next;
        {   // the next step's commands are on their way while the scan runs
            const uint32_t kn = k0 + 32u + lane;
            c_next = kn < n ? cmd_top[-(int)kn] : 0x80000000u;
        }
        const uint32_t lmis = lpos & 15u;
        const bool isrun = (c >> 31) != 0u;
        const uint32_t lr = isrun ? (c & 0x7fffffffu) : (c >> 23);
        const uint32_t len = isrun ? 0u : ((c >> 15) & 255u) + 3u;
        const uint32_t dist = (c & 0x7fffu) + 1u;
        uint32_t sc = (lr + len) | (lr << 16);       // one scan for both sums: <= 32 * 513 output bytes, <= 32 * 255 literals per step
#pragma unroll
        for (int sft = 1; sft < 32; sft <<= 1) {
            const uint32_t vs = __shfl_up_sync(FULL, sc, sft);
            if (lane >= (uint32_t)sft) sc += vs;
        }
        const uint32_t ae = sc & 0xffffu, le = sc >> 16;
        const uint32_t dl = pos + ae - lr - len;     // where this command's literals go
        const uint32_t dm = dl + lr;                 // where its match goes
        const uint32_t me = dm + len;
        const uint32_t src0 = dm - dist;             // first source byte of the match
        const uint8_t *lsrc = lit + (lpos + le - lr);   // its literals in the literal stream
        const uint32_t gend = pos + __shfl_sync(FULL, ae, 31);
        if (gend > d.isize) break;                   // cannot happen after a clean phase 1; the ISIZE check below reports it
        uint32_t permagic = 0u;
        if (dist < len) permagic = dist | ((131072u / dist + 1u) << 9);
        reinterpret_cast<uint4 *>(lstage)[lane] = lst;                // (the previous step's rows are behind a __syncwarp)
        lst = stage_load(lpos + __shfl_sync(FULL, le, 31));           // the next step's literals
        const uint32_t lrel = lpos + le - lr - (lpos - lmis);       // this command's literals inside the staged bytes
        const bool lstaged = lrel + lr <= R5_LST;

        for (uint32_t cb = pos; cb < gend; cb += R5_CH) {
            const uint32_t ce = cb + R5_CH < gend ? cb + R5_CH : gend;
            // ---- room in the window: slide the last R5_KEEP bytes (and whatever is not stored / summed yet) down ----
            if (cb + A + R5_CH > wlo + R5_BUF) {
                const uint32_t nlo = (cb + A - R5_KEEP) & ~127u;        // cb + A > R5_BUF - R5_CH here, so no underflow
                const uint32_t delta = nlo - wlo, units = (cb + A - nlo + 15u) >> 4;
                for (uint32_t u0 = 0; u0 < units; u0 += 32u) {
                    uint4 x = make_uint4(0u, 0u, 0u, 0u);
                    if (u0 + lane < units) x = *reinterpret_cast<const uint4 *>(buf + delta + ((u0 + lane) << 4));
                    __syncwarp();
                    if (u0 + lane < units) *reinterpret_cast<uint4 *>(buf + ((u0 + lane) << 4)) = x;
                    __syncwarp();
                }
                wlo = nlo;
            }
            // ---- the segments of this chunk: start bits, descriptors at their ordinal ----
            if (lane < 16u) mask[lane] = 0u;
            __syncwarp();
            const bool hasL = lr != 0u && dl < ce && dm > cb, hasM = len != 0u && dm < ce && me > cb;
            const uint32_t bl = __ballot_sync(FULL, hasL), bm = __ballot_sync(FULL, hasM);   // (also orders the mask clear)
            const uint32_t ord = __popc(bl & lt_mask) + __popc(bm & lt_mask);
            if (hasL) {
                const uint32_t st = (dl > cb ? dl : cb) - cb;
                atomicOr(&mask[st >> 5], 1u << (st & 31u));
                const uint64_t ga = (uint64_t)(uintptr_t)(lstaged ? lstage + lrel : lsrc);
                desc[ord] = make_uint4(dl, 0x80000000u, (uint32_t)ga, (uint32_t)(ga >> 32));
            }
            if (hasM) {
                const uint32_t st = (dm > cb ? dm : cb) - cb;
                atomicOr(&mask[st >> 5], 1u << (st & 31u));
                // sources inside the window come from there, older ones from the output buffer (all of those are stored: they lie
                // at least R5_KEEP - 128 bytes behind the chunk)
                const bool far = src0 + A < wlo;
                const uint8_t *sp = far ? o + src0 : buf + (src0 + A - wlo);
                if (far && pf_mode) {          // an old source: ask for its line(s) now, the rows that read it come a few hundred cycles later
                    r5_prefetch(sp, pf_mode);
                    r5_prefetch(sp + (len - 1u), pf_mode);
                }
                const uint64_t ga = (uint64_t)(uintptr_t)sp;
                const uint32_t oi = ord + (hasL ? 1u : 0u);
                desc[oi] = make_uint4(dm | (permagic ? R5_PER : 0u), src0, (uint32_t)ga, (uint32_t)(ga >> 32));
                if (permagic) per[oi] = permagic;
            }
            __syncwarp();
            // ---- rows of 32 bytes ----
            const uint32_t nrow = (ce - cb + 31u) >> 5;
            uint32_t acc = 0xffffffffu;                                     // (segments started in earlier rows) - 1
            uint8_t *wp = buf + (cb + A - wlo) + lane;
            // R5_ROWS rows per round: their loads are in flight together; a byte whose source lies inside the round's bytes takes it
            // from the owning lane's register -- rows in order: first from the rows in front (final by then), then from its own row
            for (uint32_t row = 0; row < nrow; row += R5_ROWS, wp += 32 * R5_ROWS) {
                const uint32_t rs = cb + (row << 5);
                uint32_t v[R5_ROWS], dd[R5_ROWS], off[R5_ROWS], ai[R5_ROWS];
                uint4 D[R5_ROWS];
                bool pend[R5_ROWS], valid[R5_ROWS];
                uint32_t anyper = 0u;
#pragma unroll
                for (int j = 0; j < R5_ROWS; ++j) {
                    const uint32_t m = mask[row + j];
                    ai[j] = acc + __popc(m & le_mask);
                    acc += __popc(m);
                    D[j] = desc[ai[j]];
                    off[j] = rs + 32u * j + lane - D[j].x;
                    anyper |= off[j];
                }
                if (__any_sync(FULL, (int)anyper < 0)) {           // R5_PER: overlapping matches in this round, fold them into their first period
#pragma unroll
                    for (int j = 0; j < R5_ROWS; ++j)
                        if ((int)off[j] < 0) {
                            const uint32_t pm_ = per[ai[j]];
                            off[j] &= 0x7fffffffu;
                            off[j] -= ((off[j] * (pm_ >> 9)) >> 17) * (pm_ & 511u);
                        }
                }
#pragma unroll
                for (int j = 0; j < R5_ROWS; ++j) {
                    const uint32_t sp = D[j].y + off[j];           // source position (literals: negative)
                    valid[j] = rs + 32u * j + lane < ce;
                    pend[j] = valid[j] && (int)sp >= (int)rs;
                    dd[j] = sp - rs;                               // relative to the round: 0 .. 32 j + lane - 1 = inside it
                    v[j] = 0u;
                    if (valid[j] && !pend[j]) v[j] = *reinterpret_cast<const uint8_t *>((uintptr_t)((((uint64_t)D[j].w << 32) | D[j].z) + off[j]));
                }
#pragma unroll
                for (int j = 0; j < R5_ROWS; ++j) {
                    uint32_t pm = __ballot_sync(FULL, pend[j]);
                    if (pm) {
#pragma unroll
                        for (int i = 0; i < j; ++i) {
                            const uint32_t vv = __shfl_sync(FULL, v[i], (int)(dd[j] & 31u));
                            if (pend[j] && (dd[j] >> 5) == (uint32_t)i) { v[j] = vv; pend[j] = false; }
                        }
                        pm = __ballot_sync(FULL, pend[j]);
                        for (int it = 0; pm && it < 32; ++it) {   // (every pass settles the lowest pending lane at least)
                            const uint32_t vv = __shfl_sync(FULL, v[j], (int)(dd[j] & 31u));
                            if (pend[j] && !((pm >> (dd[j] & 31u)) & 1u)) { v[j] = vv; pend[j] = false; }
                            pm = __ballot_sync(FULL, pend[j]);
                        }
                    }
                }
#pragma unroll
                for (int j = 0; j < R5_ROWS; ++j)
                    if (valid[j]) wp[32 * j] = (uint8_t)v[j];
                __syncwarp();
            }
            // ---- finished 16-byte units leave ----
            const uint32_t avail = ce + A;
            const uint32_t fend = avail >> 4;
            for (uint32_t u = fu + lane; u < fend; u += 32u)
                *reinterpret_cast<uint4 *>(ob + ((size_t)u << 4)) = *reinterpret_cast<const uint4 *>(buf + ((u << 4) - wlo));
            if (fend > fu) fu = fend;
            if (head_pending && avail >= head_end) {
                const uint32_t q = (A & ~15u) + lane;
                if (lane < 16u && q >= A && q < head_end) ob[q] = buf[q - wlo];
                head_pending = false;
            }
            // ---- CRC-32 catches up from the window ----
            if (cq < a0q) {
                const uint32_t lim = a0q < avail ? a0q : avail;
                if (lane == 0)
                    for (uint32_t q = cq; q < lim; ++q) r = tab(4 * 256 + ((r ^ buf[q - wlo]) & 0xffu)) ^ (r >> 8);
                cq = lim;
            }
            while (cq >= a0q && cq + 128u <= rows_end && cq + 128u <= avail) {
                uint32_t w = *reinterpret_cast<const uint32_t *>(buf + (cq - wlo) + 4u * lane);
                if (first_row && lane == 0) w ^= r;
                first_row = false;
                H = tab(H & 0xffu) ^ tab(256 + ((H >> 8) & 0xffu)) ^ tab(512 + ((H >> 16) & 0xffu)) ^ tab(768 + (H >> 24)) ^ w;
                cq += 128u;
            }
            __syncwarp();
        }
        pos = gend;
        lpos += __shfl_sync(FULL, le, 31);
    }

    // ---- the last partial unit, the end of the CRC, the footer checks ----
    if (pos == d.isize) {
        if (head_pending) {                            // a block that ends inside its first partial unit
            const uint32_t q = (A & ~15u) + lane;
            if (lane < 16u && q >= A && q < head_end) ob[q] = buf[q - wlo];
        }
        const uint32_t q = (fu << 4) + lane;
        if (lane < 16u && q >= head_end && q < endq) ob[q] = buf[q - wlo];
        if (rows_end > a0q) {
            uint32_t v = crc_mulmod(tab(5 * 256 + lane), H);
#pragma unroll
            for (int sft = 16; sft; sft >>= 1) v ^= __shfl_xor_sync(FULL, v, sft);
            r = v;
        }
        if (lane == 0)
            for (uint32_t q2 = rows_end; q2 < endq; ++q2) r = tab(4 * 256 + ((r ^ buf[q2 - wlo]) & 0xffu)) ^ (r >> 8);
    }
    if (lane == 0) {
        if (pos != d.isize) *status_b = INF_ERR_ISIZE;
        else if ((r ^ 0xffffffffu) != d.crc) *status_b = INF_ERR_CRC;
    }
}

}  // namespace dh
