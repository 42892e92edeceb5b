// bgzf_inflate.cu — hand-written sm_100a BGZF inflate kernel.
//
// Replaces bgzf_uncompress / inflate_block / bgzf_decode_func (htslib/bgzf.c:762-804, 808-824,
// 1373-1384): raw DEFLATE inflate of every BGZF block, CRC-32 of the output compared with the footer,
// inflated length compared with ISIZE — all in one persistent kernel.
//
// Execution model (B200: 148 SMs, 227 KB smem/CTA, warps of 32):
//   * One persistent CTA per SM, 256 threads = 32 LANE-GROUPS of G=8 lanes; a lane-group owns one BGZF
//     block at a time and pulls the next block index from a global ticket counter (file order, so
//     in-flight blocks are neighbours in HBM).  ~4.7k blocks are in flight chip-wide.
//   * DEFLATE decoding is bit-serial per block, so the parallelism is ACROSS blocks: the 4 groups of a
//     warp run the same decode instruction stream in SIMT lock-step on 4 different blocks (4x the issue
//     efficiency of one warp per block), and each group's 8 lanes decode REDUNDANTLY (identical
//     registers, broadcast shared-memory reads) so that they can split the byte work — literal runs and
//     LZ77 copies — without any shuffle traffic.
//   * Per group in shared memory: litlen table (9-bit root + subtables, 852 entries), distance table
//     (8-bit root, 402 entries), code-length scratch, and a 2x512 B input ring filled by TMA bulk copies
//     (cp.async.bulk global->shared, completion on an mbarrier per slot) that run ahead of the decoder.
//   * Output goes straight to the block's final place in the inflated buffer (no 64 KB staging in smem);
//     back-references re-read the group's own recent output through L1/L2.
//   * After end-of-stream the group computes CRC-32 over its output in 8 parallel segments
//     (slice-by-8 tables shared by the CTA) and merges them with GF(2) polynomial shifts.
#include <mutex>

#include "bgzf_inflate.cuh"
#include "crc32_core.cuh"
#include "inflate_core.cuh"

namespace dh {

constexpr int G = 8;                 // lanes per BGZF block
constexpr int CTA_THREADS = 256;     // 32 lane-groups
constexpr int GROUPS = CTA_THREADS / G;
constexpr uint32_t SLOT = 512;       // bytes per TMA bulk copy
constexpr uint32_t NSLOT = 2;        // ring slots per group
constexpr uint32_t RING = SLOT * NSLOT;
constexpr int NACC = 8;              // literal bytes buffered in registers before a store (<= G)

struct alignas(16) GroupSmem {
    InflateTables t;
    alignas(16) uint8_t ring[RING];
    alignas(8) uint64_t mbar[NSLOT];
};

struct alignas(16) CtaSmem {
    uint32_t crc_tab[8 * 256];
    uint32_t x2n[32];
    GroupSmem g[GROUPS];
};

// ---- PTX wrappers -----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
// TMA 1-D bulk copy global -> shared, completion counted in bytes on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void tma_bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

// ---- IO policy of the lane-group ------------------------------------------------------------------
// Every member function is called by all G lanes with identical arguments (group-uniform control flow).
struct GroupIO {
    // input ring
    GroupSmem *sm;
    uint32_t ring_s, mbar_s;      // shared-space addresses
    const uint8_t *payload;       // global: first byte of the DEFLATE payload
    const uint8_t *gbase;         // global, 16 B aligned: source of relative chunk 0
    int32_t pos0;                 // payload position that gbase corresponds to (<= current position)
    uint32_t in_len;
    uint32_t total;               // relative chunks that may ever be needed
    uint32_t chunk_base;          // absolute chunk id of relative chunk 0 (keeps slot/parity continuous)
    uint32_t issued, waited;      // relative chunk counters
    uint32_t fast_hi;             // relative offset below which load32 needs no ring service
    // output
    uint8_t *out;
    uint32_t outpos;
    uint64_t acc;
    uint32_t nacc;
    uint32_t lane, gmask;

    __device__ __forceinline__ bool leader() const { return lane == 0; }
    __device__ __forceinline__ void sync() const { __syncwarp(gmask); }
    __device__ __forceinline__ uint32_t produced() const { return outpos + nacc; }

    // Slow path of the ring, out of line (runs once per SLOT input bytes): recycle consumed slots with new
    // TMA bulk copies and wait for the chunks the next loads need.  Takes and returns plain values so that
    // none of the decoder's registers has to live in memory.  Returns issued | waited << 32.
    static __device__ __noinline__ uint64_t service(uint32_t a, uint32_t issued, uint32_t waited, uint32_t total,
                                                    uint32_t chunk_base, const uint8_t *gbase, uint32_t ring_s,
                                                    uint32_t mbar_s, uint32_t lane, uint32_t gmask) {
        uint32_t dead = a / SLOT;  // chunks below this one are fully consumed
        uint32_t lim = dead + NSLOT;
        if (lim > total) lim = total;
        if (issued < lim) {
            __syncwarp(gmask);  // every lane is done reading the slots being recycled
            if (lane == 0) {
                fence_proxy_async();
                for (uint32_t rel = issued; rel < lim; ++rel) {
                    uint32_t slot = (chunk_base + rel) % NSLOT;
                    uint32_t bar = mbar_s + 8u * slot;
                    mbar_arrive_expect_tx(bar, SLOT);
                    tma_bulk_g2s(ring_s + slot * SLOT, gbase + (size_t)rel * SLOT, SLOT, bar);
                }
            }
            issued = lim;
        }
        uint32_t need = (a + 7u) / SLOT;
        if (need >= total) need = total - 1;  // unreachable for positions the decoder may legally touch
        while (waited <= need) {
            uint32_t abs_c = chunk_base + waited;
            uint32_t bar = mbar_s + 8u * (abs_c % NSLOT);
            uint32_t parity = (abs_c / NSLOT) & 1u;
            while (!mbar_try_wait(bar, parity)) {}
            ++waited;
        }
        return (uint64_t)issued | ((uint64_t)waited << 32);
    }
    // (re)start the ring at payload position `pos`
    __device__ __forceinline__ void open_at(uint32_t pos) {
        const uint8_t *g = payload + pos;
        uint32_t mis = (uint32_t)((uintptr_t)g & 15u);
        gbase = g - mis;
        pos0 = (int32_t)pos - (int32_t)mis;
        total = (uint32_t)((int32_t)in_len + 24 - pos0 + (int32_t)SLOT - 1) / SLOT;
        issued = 0;
        waited = 0;
        fast_hi = 0;
    }
    // wait for everything in flight, so that the next stream can reuse the slots/barriers
    __device__ __forceinline__ void drain() {
        while (waited < issued) {
            uint32_t abs_c = chunk_base + waited;
            while (!mbar_try_wait(mbar_s + 8u * (abs_c % NSLOT), (abs_c / NSLOT) & 1u)) {}
            ++waited;
        }
        chunk_base += issued;
        issued = 0;
        waited = 0;
        fast_hi = 0;
    }

    __device__ __forceinline__ uint32_t load32(uint32_t bytepos) {
        uint32_t a = (uint32_t)((int32_t)bytepos - pos0);  // relative byte offset in the ring stream
        if (a >= fast_hi) {
            uint64_t iw = service(a, issued, waited, total, chunk_base, gbase, ring_s, mbar_s, lane, gmask);
            issued = (uint32_t)iw;
            waited = (uint32_t)(iw >> 32);
            // until `a` reaches fast_hi no slot can be recycled and every byte a load touches has landed
            uint32_t hi_dead = (a / SLOT + 1u) * SLOT;
            uint32_t hi_wait = waited * SLOT - 7u;
            fast_hi = hi_dead < hi_wait ? hi_dead : hi_wait;
        }
        uint32_t r = (chunk_base * SLOT + (a & ~3u)) & (RING - 1u);
        const uint32_t *ring32 = reinterpret_cast<const uint32_t *>(sm->ring);
        uint32_t w0 = ring32[r >> 2];
        uint32_t w1 = ring32[((r + 4u) & (RING - 1u)) >> 2];
        return __funnelshift_r(w0, w1, (a & 3u) * 8u);
    }

    __device__ __forceinline__ void flush() {
        if (nacc) {
            if (lane < nacc && outpos + lane < BGZF_MAX_OUT) out[outpos + lane] = (uint8_t)(acc >> (8u * lane));
            outpos += nacc;
            nacc = 0;
            acc = 0;
        }
    }
    __device__ __forceinline__ void literal(uint32_t b) {
        acc |= (uint64_t)b << (8u * nacc);
        if (++nacc == (uint32_t)NACC) flush();
    }
    __device__ __forceinline__ bool match(uint32_t len, uint32_t dist) {
        flush();
        if (dist > outpos || outpos + len > BGZF_MAX_OUT) return false;
        __syncwarp(gmask);  // earlier stores of all lanes are visible before anyone reads them back
        uint8_t *dst = out + outpos;
        const uint8_t *src = dst - dist;
        if (dist >= (uint32_t)G) {
            // pass p copies bytes [pG, pG+G); its sources lie below pG because dist >= G
            for (uint32_t base = 0; base < len; base += G) {
                uint32_t i = base + lane;
                if (i < len) dst[i] = src[i];
                __syncwarp(gmask);
            }
        } else {
            // short period: byte i repeats src[i mod dist], all of which predate this match
            uint32_t m = lane % dist;
            const uint32_t step = (uint32_t)G % dist;
            for (uint32_t base = 0; base < len; base += G) {
                uint32_t i = base + lane;
                if (i < len) dst[i] = src[m];
                m += step;
                if (m >= dist) m -= dist;
            }
        }
        outpos += len;
        return true;
    }
    __device__ __forceinline__ bool stored(uint32_t bytepos, uint32_t len) {
        flush();
        if (outpos + len > BGZF_MAX_OUT) return false;
        const uint8_t *s = payload + bytepos;
        uint8_t *dst = out + outpos;
        for (uint32_t i = lane; i < len; i += G) dst[i] = s[i];
        outpos += len;
        // the bit reader resumes behind the raw bytes: restart the ring there
        drain();
        open_at(bytepos + len);
        return true;
    }
};

// CRC-32 of out[0..n) by the lane-group; result valid in every lane
__device__ __forceinline__ uint32_t group_crc32(const CtaSmem *cs, const uint8_t *p, uint32_t n, uint32_t lane,
                                                uint32_t gmask) {
    uint32_t b, e;
    crc_segment(n, G, (int)lane, b, e);
    uint32_t r = crc_update(cs->crc_tab, lane == 0 ? 0xffffffffu : 0u, p + b, e - b);
    r = crc_shift(cs->x2n, r, n - e);
#pragma unroll
    for (int d = 1; d < G; d <<= 1) r ^= __shfl_xor_sync(gmask, r, d);
    return r ^ 0xffffffffu;
}

__global__ void __launch_bounds__(CTA_THREADS, 1)
bgzf_inflate_kernel(const uint8_t *__restrict__ comp, const BlockDesc *__restrict__ blocks, uint32_t nblocks,
                    uint8_t *__restrict__ outbuf, int *__restrict__ status, unsigned int *__restrict__ counter, int only_flagged) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    CtaSmem *cs = reinterpret_cast<CtaSmem *>(smem_raw);

    const uint32_t tid = threadIdx.x;
    const uint32_t grp = tid / G;
    const uint32_t lane = tid % G;
    const uint32_t gmask = ((1u << G) - 1u) << ((tid & 31u) / G * G);
    GroupSmem *gs = &cs->g[grp];

    // one-time CTA setup: CRC tables, x^(2^k) table, mbarriers
    crc_table_column(cs->crc_tab, (int)tid);  // CTA_THREADS == 256 columns
    if (tid == 0) crc_x2n_init(cs->x2n);
    if (lane == 0) {
        for (uint32_t s = 0; s < NSLOT; ++s) mbar_init(smem_u32(&gs->mbar[s]), 1);
        fence_mbar_init();
    }
    __syncthreads();

    GroupIO io;
    io.sm = gs;
    io.ring_s = smem_u32(gs->ring);
    io.mbar_s = smem_u32(gs->mbar);
    io.chunk_base = 0;
    io.issued = io.waited = 0;
    io.fast_hi = 0;
    io.lane = lane;
    io.gmask = gmask;

    for (;;) {
        uint32_t b = 0;
        if (lane == 0) b = atomicAdd(counter, 1u);
        b = __shfl_sync(gmask, b, 0, G);
        if (b >= nblocks) break;
        if (only_flagged && status[b] != INF_NEEDS_V1) continue;  // filter mode: redo what the lockstep kernel passed on
        const BlockDesc d = blocks[b];

        io.payload = comp + d.in_off;
        io.in_len = d.in_len;
        io.out = outbuf + d.out_off;
        io.outpos = 0;
        io.acc = 0;
        io.nacc = 0;
        io.open_at(0);

        Inflater<GroupIO> inf(io, gs->t, d.in_len);
        int rc = inf.run();
        io.flush();
        io.drain();
        __syncwarp(gmask);

        const uint32_t n = io.outpos;
        if (rc == INF_OK && n != d.isize) rc = INF_ERR_ISIZE;
        if (rc == INF_OK) {
            uint32_t crc = group_crc32(cs, io.out, n, lane, gmask);
            if (crc != d.crc) rc = INF_ERR_CRC;
        }
        if (lane == 0) status[b] = rc;
    }
}

cudaError_t launch_bgzf_inflate(const uint8_t *comp, const BlockDesc *blocks, uint32_t nblocks, uint8_t *out,
                                int *status, unsigned int *counter, int sm_count, cudaStream_t stream, bool only_flagged) {
    // the attribute is per device (one process may drive several GPUs, one dh_ctx each) and set once per device
    static std::mutex mu;
    static bool configured[64] = {};
    const size_t smem = sizeof(CtaSmem);
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    {
        std::lock_guard<std::mutex> lk(mu);
        if (!configured[dev]) {
            e = cudaFuncSetAttribute(bgzf_inflate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e != cudaSuccess) return e;
            configured[dev] = true;
        }
    }
    e = cudaMemsetAsync(counter, 0, sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    if (nblocks == 0) return cudaSuccess;
    // persistent: one CTA per SM; fewer if there is not enough work for every lane-group
    uint32_t ctas = (nblocks + GROUPS - 1) / GROUPS;
    if (ctas > (uint32_t)sm_count) ctas = (uint32_t)sm_count;
    bgzf_inflate_kernel<<<ctas, CTA_THREADS, smem, stream>>>(comp, blocks, nblocks, out, status, counter, only_flagged ? 1 : 0);
    return cudaGetLastError();
}

}  // namespace dh
