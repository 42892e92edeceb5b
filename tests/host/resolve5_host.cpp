// Host build of duckhts_b200/csrc/bgzf_resolve5.cuh -- the per-warp body of the inflate's phase-2 kernel -- behind a SIMT emulation.
// TEST INFRASTRUCTURE ONLY (tests/test_resolve5_host.py, -m "not gpu"); never linked into the product.
//
// The 32 lanes of one warp run as coroutines (ucontext) on one thread.  Every warp intrinsic the kernel uses (__shfl_sync,
// __shfl_up_sync, __shfl_xor_sync, __ballot_sync, __any_sync, __syncwarp) is a rendezvous: a lane parks there, and when every lane
// has parked the results are computed for all of them at once and the lanes go on.  The rendezvous also checks what the hardware
// would punish with a hang or garbage: all 32 lanes must arrive at the SAME kind of intrinsic, and no lane may leave the kernel
// while others wait (a *_sync inside a lane-divergent branch -- the bug that cost a GPU session in round 2 -- fails here in
// milliseconds).  Between two rendezvous the lanes run one after the other, so shared-memory traffic that the kernel orders with
// __syncwarp is ordered here too; what the emulation cannot show is a MISSING __syncwarp.
//
// Input of phase 2 = output of phase 1: a literal stream and a command stream per block.  They are produced here from a raw DEFLATE
// payload by the host build of the product's own decoder core (inflate_core.cuh) with an IO policy that records tokens in phase 1's
// format (bgzf_inflate3.cu: one word per match, lit_run << 23 | (len - 3) << 15 | (dist - 1); 0x80000000 | run for literal runs).
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <ucontext.h>

#include <vector>

#include <cuda_runtime.h>   // vector types (uint4, make_uint4) and the runtime types bgzf_inflate.cuh's declarations name; host headers only

// ---- SIMT emulation -----------------------------------------------------------------------------------------------------------
namespace simt {
constexpr int W = 32;
enum { OP_NONE = 0, OP_SHFL, OP_SHFL_UP, OP_SHFL_XOR, OP_BALLOT, OP_ANY, OP_SYNCWARP };
static const char *op_name[] = {"-", "__shfl_sync", "__shfl_up_sync", "__shfl_xor_sync", "__ballot_sync", "__any_sync", "__syncwarp"};
struct LaneCtx {
    ucontext_t ctx;
    std::vector<char> stack;
    bool done = false;
    int op = OP_NONE;
    uint32_t a = 0, b = 0, res = 0;
};
static LaneCtx L[W];
static ucontext_t sched;
static int cur = -1;
static char fail[256];
static uint64_t n_rendezvous;

static uint32_t rendezvous(int op, uint32_t a, uint32_t b) {
    LaneCtx &me = L[cur];
    me.op = op; me.a = a; me.b = b;
    swapcontext(&me.ctx, &sched);
    return me.res;
}
static void (*lane_fn)(uint32_t lane);
static void lane_main(int lane) {
    lane_fn((uint32_t)lane);
    L[lane].done = true;       // uc_link returns to the scheduler
}
// run fn on 32 lanes; false (and a message in fail) when the lanes diverge around an intrinsic
static bool run_warp(void (*fn)(uint32_t)) {
    lane_fn = fn;
    fail[0] = 0;
    for (int i = 0; i < W; ++i) {
        L[i].stack.resize(256 * 1024);
        L[i].done = false; L[i].op = OP_NONE;
        getcontext(&L[i].ctx);
        L[i].ctx.uc_stack.ss_sp = L[i].stack.data();
        L[i].ctx.uc_stack.ss_size = L[i].stack.size();
        L[i].ctx.uc_link = &sched;
        makecontext(&L[i].ctx, (void (*)())lane_main, 1, i);
    }
    for (;;) {
        int n_done = 0;
        for (int i = 0; i < W; ++i) {
            if (L[i].done) { ++n_done; continue; }
            cur = i;
            L[i].op = OP_NONE;
            swapcontext(&sched, &L[i].ctx);      // until the lane parks at an intrinsic or returns
            if (L[i].done) ++n_done;
        }
        if (n_done == W) return true;
        if (n_done) {
            int w = 0; while (L[w].done) ++w;
            snprintf(fail, sizeof fail, "%d lanes left the kernel while lane %d waits at %s", n_done, w, op_name[L[w].op]);
            return false;
        }
        for (int i = 1; i < W; ++i)
            if (L[i].op != L[0].op) {
                snprintf(fail, sizeof fail, "lanes diverge: lane 0 at %s, lane %d at %s (rendezvous %llu)", op_name[L[0].op], i, op_name[L[i].op],
                         (unsigned long long)n_rendezvous);
                return false;
            }
        ++n_rendezvous;
        uint32_t any = 0, ballot = 0, snap[W];
        for (int i = 0; i < W; ++i) { snap[i] = L[i].a; if (L[i].a) { any = 1; ballot |= 1u << i; } }
        for (int i = 0; i < W; ++i) {
            switch (L[0].op) {
            case OP_SHFL: L[i].res = snap[L[i].b & 31u]; break;
            case OP_SHFL_UP: L[i].res = (uint32_t)i >= L[i].b ? snap[i - (int)L[i].b] : snap[i]; break;
            case OP_SHFL_XOR: L[i].res = snap[(i ^ (int)L[i].b) & 31]; break;
            case OP_BALLOT: L[i].res = ballot; break;
            case OP_ANY: L[i].res = any; break;
            default: L[i].res = 0; break;
            }
        }
    }
}
}  // namespace simt

// the intrinsics, as the kernel source spells them
static inline uint32_t __shfl_sync(unsigned, uint32_t v, int src) { return simt::rendezvous(simt::OP_SHFL, v, (uint32_t)src); }
static inline uint32_t __shfl_up_sync(unsigned, uint32_t v, int d) { return simt::rendezvous(simt::OP_SHFL_UP, v, (uint32_t)d); }
static inline uint32_t __shfl_xor_sync(unsigned, uint32_t v, int m) { return simt::rendezvous(simt::OP_SHFL_XOR, v, (uint32_t)m); }
static inline uint32_t __ballot_sync(unsigned, bool p) { return simt::rendezvous(simt::OP_BALLOT, p ? 1u : 0u, 0u); }
static inline bool __any_sync(unsigned, bool p) { return simt::rendezvous(simt::OP_ANY, p ? 1u : 0u, 0u) != 0u; }
static inline void __syncwarp() { simt::rendezvous(simt::OP_SYNCWARP, 0u, 0u); }
static inline int __popc(uint32_t x) { return __builtin_popcount(x); }
static inline uint32_t atomicOr(uint32_t *p, uint32_t v) { uint32_t o = *p; *p = o | v; return o; }   // one thread runs all the lanes

#include "../../duckhts_b200/csrc/bgzf_resolve5.cuh"

// ---- phase 1's output format, from the product's decoder core -------------------------------------------------------------------
struct CmdIO {
    const uint8_t *in; uint32_t in_len;
    std::vector<uint8_t> lit;
    std::vector<uint32_t> cmd;
    uint32_t lit_run = 0, pos = 0;
    uint32_t load32(uint32_t p) const {
        uint32_t v = 0;
        for (int i = 0; i < 4; ++i) if (p + i < in_len) v |= (uint32_t)in[p + i] << (8 * i);
        return v;
    }
    void literal(uint32_t b) {
        lit.push_back((uint8_t)b); ++pos;
        if (++lit_run == 255u) { cmd.push_back(0x80000000u | 255u); lit_run = 0; }
    }
    bool match(uint32_t len, uint32_t dist) {
        if (dist > pos || pos + len > dh::BGZF_MAX_OUT) return false;
        cmd.push_back((lit_run << 23) | ((len - 3u) << 15) | (dist - 1u));
        lit_run = 0; pos += len;
        return true;
    }
    bool stored(uint32_t p, uint32_t len) {
        if (pos + len > dh::BGZF_MAX_OUT) return false;
        for (uint32_t i = 0; i < len; ++i) literal(in[p + i]);
        return true;
    }
    uint32_t produced() const { return pos; }
    bool leader() const { return true; }
    void sync() const {}
};

namespace {
struct Job {
    dh::BlockDesc d;
    uint32_t n;
    uint8_t *outbuf;
    const uint8_t *work;
    int status;
    int pf_mode;
    uint32_t tabs[dh::CRC3_WORDS];
    alignas(16) uint8_t smem[dh::R5_WARP_SMEM];
} job;
void lane_entry(uint32_t lane) {
    dh::resolve5_warp(job.d, job.n, job.outbuf, job.work, &job.status, job.tabs, job.smem, lane, job.pf_mode);
}
}  // namespace

// raw: one block's raw DEFLATE payload; isize / crc: what the BGZF footer would say (the kernel checks both); align: the block's output
// begins at a 128-byte aligned address + align (0..127).  out receives isize bytes.  Returns the status word the kernel leaves (0 or
// dh::INF_ERR_*), -1 when the lanes diverged (msg), -2 when the payload does not decode, -3 when bytes outside the block were written.
extern "C" int dh_host_resolve5(const uint8_t *raw, uint32_t raw_len, uint32_t isize, uint32_t crc, uint32_t align, uint8_t *out, int pf_mode,
                                uint64_t *n_rendezvous, char *msg, int msg_cap) {
    static dh::InflateTables t;
    CmdIO io{raw, raw_len};
    dh::Inflater<CmdIO> inf(io, t, raw_len);
    if (inf.run() != 0) return -2;
    if (io.lit_run) io.cmd.push_back(0x80000000u | io.lit_run);
    const uint32_t produced = io.pos;             // (the kernel compares with isize itself)
    const uint32_t area = dh::inflate_work_cap(produced > isize ? produced : isize);
    if (io.lit.size() + 4 * io.cmd.size() > area) return -2;
    uint8_t *work = (uint8_t *)aligned_alloc(256, area + 256);
    memset(work, 0xEE, area + 256);
    memcpy(work, io.lit.data(), io.lit.size());
    uint32_t *top = reinterpret_cast<uint32_t *>(work + area) - 1;
    for (size_t k = 0; k < io.cmd.size(); ++k) top[-(ptrdiff_t)k] = io.cmd[k];
    const size_t GUARD = 256;
    const size_t cap = GUARD + 128 + (size_t)(produced > isize ? produced : isize) + GUARD + 128;
    uint8_t *obuf = (uint8_t *)aligned_alloc(128, (cap + 127) & ~(size_t)127);
    memset(obuf, 0xA5, cap);
    memset(&job.d, 0, sizeof job.d);
    job.d.isize = isize; job.d.crc = crc; job.d.out_off = GUARD + (align & 127u); job.d.cmd_off = 0;
    job.n = (uint32_t)io.cmd.size();
    job.outbuf = obuf; job.work = work; job.status = 0; job.pf_mode = pf_mode;
    dh::build_crc3_tables(job.tabs);
    memset(job.smem, 0xCC, sizeof job.smem);
    simt::n_rendezvous = 0;
    const bool ok = simt::run_warp(lane_entry);
    if (n_rendezvous) *n_rendezvous = simt::n_rendezvous;
    int rc = job.status;
    if (!ok) { snprintf(msg, (size_t)msg_cap, "%s", simt::fail); rc = -1; }
    else {
        const uint8_t *o = obuf + job.d.out_off;
        const uint32_t wrote = produced < isize ? produced : isize;
        for (size_t i = 0; i < job.d.out_off && rc >= 0; ++i) if (obuf[i] != 0xA5) { snprintf(msg, (size_t)msg_cap, "byte %zu in front of the block was written", job.d.out_off - i); rc = -3; }
        if (rc == 0 || rc == dh::INF_ERR_CRC)
            for (size_t i = job.d.out_off + wrote; i < cap && rc >= 0; ++i) if (obuf[i] != 0xA5) { snprintf(msg, (size_t)msg_cap, "byte %zu behind the block was written", i - job.d.out_off - wrote); rc = -3; }
        memcpy(out, o, wrote);
    }
    free(work); free(obuf);
    return rc;
}

// the harness tested on itself: the two ways a warp goes wrong around an intrinsic must both be reported
namespace {
void bad_divergent(uint32_t lane) { if (lane & 1u) __syncwarp(); else (void)__ballot_sync(0xffffffffu, true); }
void bad_exit(uint32_t lane) { if (lane == 5u) return; (void)__shfl_sync(0xffffffffu, lane, 0); }
void good(uint32_t lane) {
    uint32_t v = __shfl_sync(0xffffffffu, lane * 3u, (int)((lane + 1u) & 31u));
    v += __shfl_up_sync(0xffffffffu, lane, 1) + __shfl_xor_sync(0xffffffffu, lane, 16);
    if (__ballot_sync(0xffffffffu, lane < 4u) != 0xfu || !__any_sync(0xffffffffu, lane == 31u) || __any_sync(0xffffffffu, false)) job.status = 1;
    const uint32_t up = lane ? lane - 1u : lane;
    if (v != ((lane + 1u) & 31u) * 3u + up + (lane ^ 16u)) job.status = 2;
}
}  // namespace
extern "C" int dh_host_simt_selftest(char *msg, int msg_cap) {
    int seen = 0;
    if (!simt::run_warp(bad_divergent) && strstr(simt::fail, "diverge")) seen |= 1;
    if (!simt::run_warp(bad_exit) && strstr(simt::fail, "left the kernel")) seen |= 2;
    job.status = 0;
    if (simt::run_warp(good) && job.status == 0) seen |= 4;
    snprintf(msg, (size_t)msg_cap, "%s", simt::fail);
    return seen;
}
