// dh_api.cu — C-ABI of libduckhts_cuda.so: contexts, BGZF framing (A1), inflate entry points (A2/A3).
// Declarations and the reference interfaces each one replaces: include/duckhts_cuda.h.
#include <errno.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <map>
#include <mutex>
#include <thread>
#include <vector>

#include "dh_internal.h"
#include "inflate_core.cuh"

namespace dh {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
}
int cuda_fail(cudaError_t e, const char *what) {
    set_error("CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
    return DH_ECUDA;
}

// ---- allocation cache -------------------------------------------------------------------------------------------------
// cudaMalloc / cudaHostAlloc of the pipeline's buffers (GBs of segment buffers, ~100 MB pinned batch arenas) cost more than a scan of a
// small file, and a session runs many scans: released buffers are kept (per device / pinned) and handed to the next reserve that
// fits.  DH_ALLOC_CACHE=0 turns it off; dh_trim() gives everything back.
namespace {
struct AllocCache {
    std::mutex mu;
    std::multimap<size_t, void *> dev[64];
    std::multimap<size_t, void *> pin;
    size_t dev_bytes[64] = {}, pin_bytes = 0;
    const size_t dev_cap = 48ull << 30, pin_cap = 8ull << 30;
    const bool on = !(getenv("DH_ALLOC_CACHE") && atoi(getenv("DH_ALLOC_CACHE")) == 0);
};
AllocCache &cache() { static AllocCache c; return c; }
void *cache_take(std::multimap<size_t, void *> &m, size_t &total, size_t n, size_t *got) {
    auto it = m.lower_bound(n);
    if (it == m.end() || it->first > n + n / 2 + (64u << 20)) return nullptr;     // do not burn a 3 GiB buffer on a 1 MiB request
    void *p = it->second;
    *got = it->first;
    total -= it->first;
    m.erase(it);
    return p;
}
}  // namespace

int DevBuf::reserve(size_t n) {
    if (n <= cap) return DH_OK;
    // growing means cudaFree + cudaMalloc (a device-wide synchronisation): leave room, sizes of successive batches creep up
    const bool regrow = cap != 0;
    if (regrow) cudaDeviceSynchronize();      // what cudaFree would have waited for: work in flight may still use the old buffer
    release();
    size_t want = std::max(regrow ? n + n / 4 : n, (size_t)1 << 20);
    int d = 0;
    cudaGetDevice(&d);
    dev = d;
    AllocCache &c = cache();
    if (c.on && d >= 0 && d < 64) {
        std::lock_guard<std::mutex> lk(c.mu);
        size_t got = 0;
        if (void *q = cache_take(c.dev[d], c.dev_bytes[d], want, &got)) { p = q; cap = got; return DH_OK; }
    }
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess && c.on) {           // out of memory with buffers parked in the cache: give them back and try once more
        cudaGetLastError();
        dh_trim();
        e = cudaMalloc(&p, want);
    }
    if (e != cudaSuccess) { p = nullptr; return cuda_fail(e, "cudaMalloc"); }
    cap = want;
    return DH_OK;
}
void DevBuf::release() {
    if (p) {
        AllocCache &c = cache();
        bool kept = false;
        if (c.on && dev >= 0 && dev < 64) {
            std::lock_guard<std::mutex> lk(c.mu);
            if (c.dev_bytes[dev] + cap <= c.dev_cap) { c.dev[dev].emplace(cap, p); c.dev_bytes[dev] += cap; kept = true; }
        }
        if (!kept) cudaFree(p);
    }
    p = nullptr;
    cap = 0;
}
int PinBuf::reserve(size_t n) {
    if (n <= cap) return DH_OK;
    const bool regrow = cap != 0;
    if (regrow) cudaDeviceSynchronize();
    release();
    size_t want = std::max(regrow ? n + n / 4 : n, (size_t)1 << 16);
    AllocCache &c = cache();
    if (c.on) {
        std::lock_guard<std::mutex> lk(c.mu);
        size_t got = 0;
        if (void *q = cache_take(c.pin, c.pin_bytes, want, &got)) { p = q; cap = got; return DH_OK; }
    }
    cudaError_t e = cudaHostAlloc(&p, want, cudaHostAllocPortable);
    if (e != cudaSuccess) { p = nullptr; return cuda_fail(e, "cudaHostAlloc"); }
    cap = want;
    return DH_OK;
}
void PinBuf::release() {
    if (p) {
        AllocCache &c = cache();
        bool kept = false;
        if (c.on) {
            std::lock_guard<std::mutex> lk(c.mu);
            if (c.pin_bytes + cap <= c.pin_cap) { c.pin.emplace(cap, p); c.pin_bytes += cap; kept = true; }
        }
        if (!kept) cudaFreeHost(p);
    }
    p = nullptr;
    cap = 0;
}

// ---- A1: framing -------------------------------------------------------------------------------------
// htslib/bgzf.c:896-903 check_header: ID1=31 ID2=139 CM=8, FLG&4, XLEN=6, SI1='B' SI2='C', SLEN=2.
// A plain-gzip header (check_header == -1) is not BGZF: the reference switches to a streaming gzip
// reader (bgzf.c:1168-1195), which is outside this path — we report DH_EFORMAT.
static inline bool bgzf_header_ok(const uint8_t *h) {
    return h[0] == 31 && h[1] == 139 && h[2] == 8 && (h[3] & 4) && (h[10] | (h[11] << 8)) == 6 && h[12] == 'B' &&
           h[13] == 'C' && (h[14] | (h[15] << 8)) == 2;
}
static inline uint32_t le32(const uint8_t *p) {
    return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
}

int walk_blocks(const uint8_t *data, uint64_t size, uint64_t start, BlockTable &bt, uint64_t *bad_off) {
    uint64_t off = start, u = bt.uoff.empty() ? 0 : bt.uoff.back();
    if (bt.uoff.empty()) bt.uoff.push_back(0);
    bt.eof_marker = false;
    while (off < size) {
        if (size - off < 18 || !bgzf_header_ok(data + off)) { if (bad_off) *bad_off = off; return DH_EFORMAT; }
        uint32_t bsize = (uint32_t)(data[off + 16] | (data[off + 17] << 8)) + 1;
        // bgzf.c:1198-1204: block_length < BLOCK_HEADER_LENGTH is an error; we also need room for the footer
        if (bsize < 18 + 8 || size - off < bsize) { if (bad_off) *bad_off = off; return DH_EFORMAT; }
        dh_bgzf_block b;
        b.coffset = off;
        b.csize = bsize;
        b.crc = le32(data + off + bsize - 8);
        b.isize = le32(data + off + bsize - 4);
        b.reserved = 0;
        if (b.isize > BGZF_MAX_OUT) { if (bad_off) *bad_off = off; return DH_EFORMAT; }
        bt.blocks.push_back(b);
        u += b.isize;
        bt.uoff.push_back(u);
        bt.eof_marker = (bsize == 28 && b.isize == 0);
        off += bsize;
    }
    bt.end_off = off;
    return DH_OK;
}

uint64_t fill_descs(const BlockTable &bt, size_t b0, size_t b1, uint64_t comp_base, size_t out_b0, uint64_t out_bias,
                    BlockDesc *d) {
    uint64_t area = 0;
    for (size_t i = b0; i < b1; ++i) {
        const dh_bgzf_block &b = bt.blocks[i];
        BlockDesc &o = d[i - b0];
        o.in_off = (uint32_t)(b.coffset + 18 - comp_base);
        o.in_len = b.csize - 26;
        o.isize = b.isize;
        o.crc = b.crc;
        o.out_off = out_bias + (bt.uoff[i] - bt.uoff[out_b0]);
        o.cmd_off = area;
        area += inflate_work_cap(b.isize);
    }
    return area;
}

int inflate_blocks(dh_ctx *c, const uint8_t *comp, const BlockDesc *descs, uint32_t n, uint64_t cmd_words, uint8_t *out,
                   int *status, DevBuf &work, unsigned int *counter, cudaStream_t stream, int sm_limit, cudaStream_t stream2, cudaEvent_t mid) {
    const int sms = sm_limit > 0 && sm_limit < c->sm_count ? sm_limit : c->sm_count;
    static const bool force_v1 = getenv("DH_INFLATE_V1") != nullptr;
    if (force_v1) {
        DH_CUDA(launch_bgzf_inflate(comp, descs, n, out, status, counter, c->sm_count, stream, false));
        c->launches += 1;
        return DH_OK;
    }
    int rc = work.reserve(inflate_work_bytes(n, cmd_words));
    if (rc != DH_OK) return rc;
    const bool tp = c->time_phases;
    if (tp) DH_CUDA(cudaEventRecord(c->ev_ph[0], stream));
    DH_CUDA(launch_bgzf_inflate_v3(comp, descs, n, out, status, (uint8_t *)work.p, counter, sms, stream, tp ? c->ev_ph[1] : nullptr, stream2, mid));
    cudaStream_t tail = stream2 && mid ? stream2 : stream;     // where phase 2 went
    if (tp) DH_CUDA(cudaEventRecord(c->ev_ph[2], tail));
    DH_CUDA(launch_bgzf_inflate(comp, descs, n, out, status, counter + 1, c->sm_count, tail, true));
    if (tp) DH_CUDA(cudaEventRecord(c->ev_ph[3], tail));
    c->launches += 3;   // phase 1 (main + retry CTAs in one grid), phase 2, the lane-group filter
    return DH_OK;
}

}  // namespace dh

using namespace dh;

extern "C" {

const char *dh_version(void) { return "duckhts_b200 0.1 (sm_100a)"; }
const char *dh_last_error(void) { return g_err; }

void dh_trim(void) {
    AllocCache &c = cache();
    std::lock_guard<std::mutex> lk(c.mu);
    int cur = 0;
    cudaGetDevice(&cur);
    for (int d = 0; d < 64; ++d) {
        if (c.dev[d].empty()) continue;
        cudaSetDevice(d);
        for (auto &e : c.dev[d]) cudaFree(e.second);
        c.dev[d].clear();
        c.dev_bytes[d] = 0;
    }
    cudaSetDevice(cur);
    for (auto &e : c.pin) cudaFreeHost(e.second);
    c.pin.clear();
    c.pin_bytes = 0;
}

int dh_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

dh_ctx *dh_ctx_create(int device) {
    int n = dh_device_count();
    if (n <= 0) { set_error("no CUDA device available: duckhts_b200 has no CPU fallback"); return nullptr; }
    if (device < 0 || device >= n) { set_error("device %d out of range (have %d)", device, n); return nullptr; }
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) { cuda_fail(e, "cudaSetDevice"); return nullptr; }
    dh_ctx *c = new dh_ctx();
    c->device = device;
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, device);
    if (e != cudaSuccess) { cuda_fail(e, "cudaGetDeviceProperties"); delete c; return nullptr; }
    if (prop.major < 10) {
        set_error("device %d is sm_%d%d; this library ships sm_100a code only", device, prop.major, prop.minor);
        delete c;
        return nullptr;
    }
    c->sm_count = prop.multiProcessorCount;
    bool ok = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithFlags(&c->inf2_stream, cudaStreamNonBlocking) == cudaSuccess;
    for (int i = 0; i < 3 && ok; ++i) {
        ok = cudaStreamCreateWithFlags(&c->copy_streams[i], cudaStreamNonBlocking) == cudaSuccess &&
             cudaEventCreate(&c->slots[i].ev_k0) == cudaSuccess && cudaEventCreate(&c->slots[i].ev_k1) == cudaSuccess &&
             cudaEventCreate(&c->slots[i].ev_done) == cudaSuccess;
    }
    ok = ok && cudaEventCreate(&c->ev_timer[0]) == cudaSuccess && cudaEventCreate(&c->ev_timer[1]) == cudaSuccess;
    ok = ok && c->counter.reserve(64 * sizeof(unsigned int)) == DH_OK;
    if (!ok) {
        if (!g_err[0]) set_error("failed to create CUDA streams/events");
        dh_ctx_destroy(c);
        return nullptr;
    }
    return c;
}

void dh_ctx_destroy(dh_ctx *c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    for (int i = 0; i < 3; ++i) {
        auto &s = c->slots[i];
        s.comp.release(); s.out.release(); s.desc.release(); s.status.release(); s.work.release();
        s.h_desc.release(); s.h_status.release(); s.h_comp.release(); s.h_out.release();
        if (s.ev_k0) cudaEventDestroy(s.ev_k0);
        if (s.ev_k1) cudaEventDestroy(s.ev_k1);
        if (s.ev_done) cudaEventDestroy(s.ev_done);
        if (c->copy_streams[i]) cudaStreamDestroy(c->copy_streams[i]);
    }
    c->l2_scratch.release();
    c->counter.release();
    for (int i = 0; i < 4; ++i) if (c->ev_ph[i]) cudaEventDestroy(c->ev_ph[i]);
    if (c->ev_timer[0]) cudaEventDestroy(c->ev_timer[0]);
    if (c->ev_timer[1]) cudaEventDestroy(c->ev_timer[1]);
    if (c->stream) cudaStreamDestroy(c->stream);
    if (c->inf2_stream) cudaStreamDestroy(c->inf2_stream);
    delete c;
}

int dh_ctx_sm_count(const dh_ctx *c) { return c ? c->sm_count : 0; }
uint64_t dh_ctx_launch_count(const dh_ctx *c) { return c ? c->launches : 0; }

void *dh_host_alloc(size_t bytes) {
    void *p = nullptr;
    // portable: pinned for every context of the process (one image feeds the pipelines of several GPUs)
    cudaError_t e = cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable);
    if (e != cudaSuccess) { cuda_fail(e, "cudaMallocHost"); return nullptr; }
    return p;
}
void dh_host_free(void *p) {
    if (p) cudaFreeHost(p);
}

int64_t dh_bgzf_walk(const uint8_t *data, uint64_t size, uint64_t start, dh_bgzf_block *out, int64_t cap,
                     uint64_t *end_off, int strict) {
    if (!data || start > size) { set_error("dh_bgzf_walk: bad arguments"); return DH_EINVAL; }
    BlockTable bt;
    uint64_t bad = 0;
    int rc = walk_blocks(data, size, start, bt, &bad);
    if (rc != DH_OK) {
        // blocks before the bad offset are still reported; a short tail is only an error when strict
        bool truncated_tail = bad + 18 > size || (bgzf_header_ok(data + bad) && size - bad < (uint64_t)(data[bad + 16] | (data[bad + 17] << 8)) + 1);
        if (strict || !truncated_tail) {
            set_error("invalid BGZF header at offset %llu", (unsigned long long)bad);
            return DH_EFORMAT;
        }
        bt.end_off = bad;
    }
    int64_t n = (int64_t)bt.blocks.size();
    if (out) {
        if (n > cap) n = cap;
        memcpy(out, bt.blocks.data(), (size_t)n * sizeof(dh_bgzf_block));
        if (n < (int64_t)bt.blocks.size()) bt.end_off = bt.blocks[n].coffset;
    }
    if (end_off) *end_off = bt.end_off;
    return n;
}

// ---- A2/A3 host -> host, pipelined over segments -------------------------------------------------------
static const uint64_t SEG_COMP = 384ull << 20;   // compressed bytes per segment
static const uint64_t SEG_OUT = 1536ull << 20;   // inflated bytes per segment
static const size_t SEG_BLOCKS = 1u << 17;

int dh_bgzf_inflate_host(dh_ctx *c, const uint8_t *comp, uint64_t comp_size, uint8_t *out, uint64_t out_cap,
                         uint64_t *out_size, dh_inflate_stats *st) {
    if (!c || (!comp && comp_size)) { set_error("dh_bgzf_inflate_host: bad arguments"); return DH_EINVAL; }
    DH_CUDA(cudaSetDevice(c->device));
    dh_inflate_stats local;
    if (!st) st = &local;
    memset(st, 0, sizeof *st);
    st->bad_block = -1;

    BlockTable bt;
    uint64_t bad_off = 0;
    int wrc = walk_blocks(comp, comp_size, 0, bt, &bad_off);
    if (wrc != DH_OK) {
        st->bad_block = (int64_t)bt.blocks.size();
        st->bad_code = INF_ERR_HEADER;
        set_error("invalid BGZF header at offset %llu", (unsigned long long)bad_off);
        return DH_EFORMAT;
    }
    const size_t nb = bt.blocks.size();
    st->n_blocks = nb;
    st->comp_bytes = bt.end_off;
    st->out_bytes = bt.uoff.back();
    st->eof_marker = bt.eof_marker;
    if (out_size) *out_size = bt.uoff.back();
    if (out && out_cap < bt.uoff.back()) { set_error("output buffer too small: need %llu", (unsigned long long)bt.uoff.back()); return DH_EINVAL; }

    cudaEvent_t ev_begin, ev_end;
    DH_CUDA(cudaEventCreate(&ev_begin));
    DH_CUDA(cudaEventCreate(&ev_end));
    DH_CUDA(cudaEventRecord(ev_begin, c->copy_streams[0]));

    struct Pending { size_t b0, b1; int slot; bool live; };
    Pending pend[3] = {};
    float kernel_ms = 0.f;
    int rc = DH_OK;
    auto retire = [&](Pending &p) -> int {
        if (!p.live) return DH_OK;
        auto &s = c->slots[p.slot];
        DH_CUDA(cudaEventSynchronize(s.ev_done));
        float ms = 0.f;
        DH_CUDA(cudaEventElapsedTime(&ms, s.ev_k0, s.ev_k1));
        kernel_ms += ms;
        const int *hs = (const int *)s.h_status.p;
        for (size_t i = p.b0; i < p.b1; ++i)
            if (hs[i - p.b0] != 0 && st->bad_block < 0) { st->bad_block = (int64_t)i; st->bad_code = hs[i - p.b0]; }
        p.live = false;
        return DH_OK;
    };

    size_t b = 0;
    int seg = 0;
    while (b < nb && rc == DH_OK) {
        // cut a segment on block boundaries
        size_t e = b;
        uint64_t cbase = bt.blocks[b].coffset & ~15ull;
        while (e < nb && e - b < SEG_BLOCKS && bt.blocks[e].coffset + bt.blocks[e].csize - cbase <= SEG_COMP &&
               bt.uoff[e + 1] - bt.uoff[b] <= SEG_OUT)
            ++e;
        if (e == b) e = b + 1;
        const int slot = seg % 3;
        auto &s = c->slots[slot];
        cudaStream_t strm = c->copy_streams[slot];
        if ((rc = retire(pend[slot])) != DH_OK) break;
        const uint64_t cend = bt.blocks[e - 1].coffset + bt.blocks[e - 1].csize;
        const uint64_t cbytes = cend - cbase, obytes = bt.uoff[e] - bt.uoff[b];
        if ((rc = s.comp.reserve(cbytes + DH_COMP_PAD)) != DH_OK) break;
        if ((rc = s.out.reserve(obytes + BGZF_MAX_OUT)) != DH_OK) break;
        if ((rc = s.desc.reserve((e - b) * sizeof(BlockDesc))) != DH_OK) break;
        if ((rc = s.status.reserve((e - b) * sizeof(int))) != DH_OK) break;
        if ((rc = s.h_desc.reserve((e - b) * sizeof(BlockDesc))) != DH_OK) break;
        if ((rc = s.h_status.reserve((e - b) * sizeof(int))) != DH_OK) break;
        const uint64_t cw = fill_descs(bt, b, e, cbase, b, 0, (BlockDesc *)s.h_desc.p);
        DH_CUDA(cudaMemcpyAsync(s.comp.p, comp + cbase, cbytes, cudaMemcpyHostToDevice, strm));
        DH_CUDA(cudaMemsetAsync((uint8_t *)s.comp.p + cbytes, 0, DH_COMP_PAD, strm));
        DH_CUDA(cudaMemcpyAsync(s.desc.p, s.h_desc.p, (e - b) * sizeof(BlockDesc), cudaMemcpyHostToDevice, strm));
        DH_CUDA(cudaEventRecord(s.ev_k0, strm));
        if ((rc = inflate_blocks(c, (const uint8_t *)s.comp.p, (const BlockDesc *)s.desc.p, (uint32_t)(e - b), cw, (uint8_t *)s.out.p,
                                 (int *)s.status.p, s.work, (unsigned int *)c->counter.p + 2 + 2 * slot, strm)) != DH_OK)
            break;
        DH_CUDA(cudaEventRecord(s.ev_k1, strm));
        if (out && obytes) DH_CUDA(cudaMemcpyAsync(out + bt.uoff[b], s.out.p, obytes, cudaMemcpyDeviceToHost, strm));
        DH_CUDA(cudaMemcpyAsync(s.h_status.p, s.status.p, (e - b) * sizeof(int), cudaMemcpyDeviceToHost, strm));
        DH_CUDA(cudaEventRecord(s.ev_done, strm));
        pend[slot] = Pending{b, e, slot, true};
        b = e;
        ++seg;
    }
    for (int i = 0; i < 3; ++i) {
        // retire in issue order so that bad_block is the FIRST failing block
        int slot = (seg + i) % 3;
        int r2 = retire(pend[slot]);
        if (rc == DH_OK) rc = r2;
    }
    // all slots done: stamp the end on every stream's tail by waiting for them explicitly
    for (int i = 0; i < 3; ++i) cudaStreamSynchronize(c->copy_streams[i]);
    cudaEventRecord(ev_end, c->copy_streams[0]);
    cudaEventSynchronize(ev_end);
    cudaEventElapsedTime(&st->total_ms, ev_begin, ev_end);
    cudaEventDestroy(ev_begin);
    cudaEventDestroy(ev_end);
    st->kernel_ms = kernel_ms;
    if (rc != DH_OK) return rc;
    if (st->bad_block >= 0) {
        set_error("BGZF block %lld failed inflate/CRC check (code %d)", (long long)st->bad_block, st->bad_code);
        return DH_EDATA;
    }
    return DH_OK;
}

// ---- A2/A3 file -> file, streaming (bgunzip) ---------------------------------------------------------------
// pread / pwrite of one big range by a few threads: one thread moves page-cache bytes at a fraction of what the GPU takes
bool dh_parallel_io(int fd, uint8_t *buf, uint64_t n, uint64_t off, bool write) {
    if (n == 0) return true;
    int nt = (int)std::min<uint64_t>(8, (n + (32ull << 20) - 1) / (32ull << 20));
    std::vector<int> ok((size_t)nt, 1);
    auto work = [&](int t) {
        uint64_t lo = n * (uint64_t)t / (uint64_t)nt, hi = n * (uint64_t)(t + 1) / (uint64_t)nt;
        while (lo < hi) {
            const size_t want = (size_t)std::min<uint64_t>(hi - lo, 64ull << 20);
            const ssize_t k = write ? pwrite(fd, buf + lo, want, (off_t)(off + lo)) : pread(fd, buf + lo, want, (off_t)(off + lo));
            if (k < 0 && errno == EINTR) continue;
            if (k <= 0) { ok[(size_t)t] = 0; return; }
            lo += (uint64_t)k;
        }
    };
    std::vector<std::thread> th;
    for (int t = 1; t < nt; ++t) th.emplace_back(work, t);
    work(0);
    for (auto &x : th) x.join();
    for (int v : ok) if (!v) return false;
    return true;
}

int dh_bgzf_inflate_fd(dh_ctx *c, int in_fd, int out_fd, dh_inflate_stats *st) {
    if (!c || in_fd < 0 || out_fd < 0) { set_error("dh_bgzf_inflate_fd: bad arguments"); return DH_EINVAL; }
    DH_CUDA(cudaSetDevice(c->device));
    dh_inflate_stats local;
    if (!st) st = &local;
    memset(st, 0, sizeof *st);
    st->bad_block = -1;
    struct stat sb;
    if (fstat(in_fd, &sb) != 0) { set_error("fstat: %s", strerror(errno)); return DH_EIO; }
    const uint64_t fsize = (uint64_t)sb.st_size;
    // per segment: compressed bytes read, inflated bytes written (DH_FD_SEG_MB shrinks both: tests drive many segments through small files)
    static const uint64_t seg_mb = getenv("DH_FD_SEG_MB") ? std::max(1, atoi(getenv("DH_FD_SEG_MB"))) : 192;
    const uint64_t F_COMP = seg_mb << 20, F_OUT = 4 * F_COMP;

    cudaEvent_t ev_begin, ev_end;
    DH_CUDA(cudaEventCreate(&ev_begin));
    DH_CUDA(cudaEventCreate(&ev_end));
    DH_CUDA(cudaEventRecord(ev_begin, c->copy_streams[0]));
    struct Pending { uint64_t blk0, nblk, out_off, obytes; bool live; };
    Pending pend[3] = {};
    float kernel_ms = 0.f;
    int rc = DH_OK;
    // wait for a segment, check its blocks, write its bytes
    auto retire = [&](int slot) -> int {
        Pending &p = pend[slot];
        if (!p.live) return DH_OK;
        auto &s = c->slots[slot];
        p.live = false;
        DH_CUDA(cudaEventSynchronize(s.ev_done));
        float ms = 0.f;
        DH_CUDA(cudaEventElapsedTime(&ms, s.ev_k0, s.ev_k1));
        kernel_ms += ms;
        const int *hs = (const int *)s.h_status.p;
        uint64_t good = p.obytes;
        for (uint64_t i = 0; i < p.nblk; ++i)
            if (hs[i] != 0) {
                if (st->bad_block < 0) { st->bad_block = (int64_t)(p.blk0 + i); st->bad_code = hs[i]; }
                good = ((const BlockDesc *)s.h_desc.p)[i].out_off;       // the bytes in front of the failing block are still written
                break;
            }
        if (st->bad_block >= 0 && (uint64_t)st->bad_block < p.blk0) return DH_OK;   // an earlier segment already failed
        if (!dh_parallel_io(out_fd, (uint8_t *)s.h_out.p, good, p.out_off, true)) { set_error("write: %s", strerror(errno)); return DH_EIO; }
        return DH_OK;
    };

    uint64_t foff = 0, out_off = 0, nblocks = 0;
    int seg = 0;
    bool format_error = false;
    while (foff < fsize && rc == DH_OK && st->bad_block < 0) {
        const int slot = seg % 3;
        auto &s = c->slots[slot];
        cudaStream_t strm = c->copy_streams[slot];
        if ((rc = retire(slot)) != DH_OK || st->bad_block >= 0) break;
        const uint64_t want = std::min<uint64_t>(F_COMP, fsize - foff);
        if ((rc = s.h_comp.reserve(F_COMP + 64)) != DH_OK || (rc = s.h_out.reserve(F_OUT + BGZF_MAX_OUT)) != DH_OK) break;
        if (!dh_parallel_io(in_fd, (uint8_t *)s.h_comp.p, want, foff, false)) { set_error("read: %s", strerror(errno)); rc = DH_EIO; break; }
        // frame the blocks that lie whole in the buffer (A1); the partial block at its end starts the next segment
        BlockTable bt;
        const uint8_t *buf = (const uint8_t *)s.h_comp.p;
        uint64_t off = 0, u = 0;
        bt.uoff.push_back(0);
        while (off < want) {
            if (want - off < 18) break;
            if (!bgzf_header_ok(buf + off)) { format_error = true; break; }
            const uint32_t bsize = (uint32_t)(buf[off + 16] | (buf[off + 17] << 8)) + 1;
            if (bsize < 18 + 8) { format_error = true; break; }
            if (want - off < bsize) break;
            dh_bgzf_block b;
            b.coffset = off;
            b.csize = bsize;
            b.crc = le32(buf + off + bsize - 8);
            b.isize = le32(buf + off + bsize - 4);
            b.reserved = 0;
            if (b.isize > BGZF_MAX_OUT) { format_error = true; break; }
            if (u + b.isize > F_OUT) break;
            bt.blocks.push_back(b);
            u += b.isize;
            bt.uoff.push_back(u);
            st->eof_marker = (bsize == 28 && b.isize == 0);
            off += bsize;
        }
        const size_t nb = bt.blocks.size();
        if (nb == 0) {
            // nothing whole in front of the end of the file (a cut block) or a bad header: the read error bgzf_read reports
            if (!format_error && foff + want < fsize) { set_error("internal: no BGZF block fits a segment"); rc = DH_EINVAL; break; }
            st->bad_block = (int64_t)nblocks;
            st->bad_code = INF_ERR_HEADER;
            break;
        }
        if ((rc = s.comp.reserve(off + DH_COMP_PAD)) != DH_OK || (rc = s.out.reserve(u + BGZF_MAX_OUT)) != DH_OK ||
            (rc = s.desc.reserve(nb * sizeof(BlockDesc))) != DH_OK || (rc = s.status.reserve(nb * sizeof(int))) != DH_OK ||
            (rc = s.h_desc.reserve(nb * sizeof(BlockDesc))) != DH_OK || (rc = s.h_status.reserve(nb * sizeof(int))) != DH_OK)
            break;
        const uint64_t cw = fill_descs(bt, 0, nb, 0, 0, 0, (BlockDesc *)s.h_desc.p);
        DH_CUDA(cudaMemcpyAsync(s.comp.p, buf, off, cudaMemcpyHostToDevice, strm));
        DH_CUDA(cudaMemsetAsync((uint8_t *)s.comp.p + off, 0, DH_COMP_PAD, strm));
        DH_CUDA(cudaMemcpyAsync(s.desc.p, s.h_desc.p, nb * sizeof(BlockDesc), cudaMemcpyHostToDevice, strm));
        DH_CUDA(cudaEventRecord(s.ev_k0, strm));
        if ((rc = inflate_blocks(c, (const uint8_t *)s.comp.p, (const BlockDesc *)s.desc.p, (uint32_t)nb, cw, (uint8_t *)s.out.p, (int *)s.status.p,
                                 s.work, (unsigned int *)c->counter.p + 2 + 2 * slot, strm)) != DH_OK)
            break;
        DH_CUDA(cudaEventRecord(s.ev_k1, strm));
        if (u) DH_CUDA(cudaMemcpyAsync(s.h_out.p, s.out.p, u, cudaMemcpyDeviceToHost, strm));
        DH_CUDA(cudaMemcpyAsync(s.h_status.p, s.status.p, nb * sizeof(int), cudaMemcpyDeviceToHost, strm));
        DH_CUDA(cudaEventRecord(s.ev_done, strm));
        pend[slot] = Pending{nblocks, nb, out_off, u, true};
        nblocks += nb;
        out_off += u;
        foff += off;
        st->comp_bytes += off;
        ++seg;
        if (format_error) { st->bad_block = (int64_t)nblocks; st->bad_code = INF_ERR_HEADER; }
    }
    // retire in issue order so that bad_block is the FIRST failing block and the bytes in front of it reach the file
    const int64_t framing_bad = st->bad_block;
    st->bad_block = -1;
    for (int i = 0; i < 3; ++i) {
        const int r2 = retire((seg + i) % 3);
        if (rc == DH_OK) rc = r2;
    }
    if (st->bad_block < 0) st->bad_block = framing_bad;
    for (int i = 0; i < 3; ++i) cudaStreamSynchronize(c->copy_streams[i]);
    cudaEventRecord(ev_end, c->copy_streams[0]);
    cudaEventSynchronize(ev_end);
    cudaEventElapsedTime(&st->total_ms, ev_begin, ev_end);
    cudaEventDestroy(ev_begin);
    cudaEventDestroy(ev_end);
    st->kernel_ms = kernel_ms;
    st->n_blocks = nblocks;
    st->out_bytes = out_off;
    if (rc != DH_OK) return rc;
    if (st->bad_block >= 0) {
        if (st->bad_code == INF_ERR_HEADER) { set_error("invalid or truncated BGZF block %lld", (long long)st->bad_block); return DH_EFORMAT; }
        set_error("BGZF block %lld failed inflate/CRC check (code %d)", (long long)st->bad_block, st->bad_code);
        return DH_EDATA;
    }
    return DH_OK;
}

// ---- device-resident variant -----------------------------------------------------------------------------
struct dh_bgzf_dev {
    dh_ctx *ctx;
    BlockTable bt;
    DevBuf comp, out, desc, status, work, work2;
    struct Seg { size_t b0, b1; uint64_t cbase, cmd_words; };
    std::vector<Seg> segs;
    cudaEvent_t ev_mid[2] = {}, ev_tail[2] = {};      // two-stream mode (DH_PHASE_STREAMS=2): phase 2 of launch i overlaps phase 1 of launch i + 1
};
static bool phase_streams_two() {
    static const bool two = getenv("DH_PHASE_STREAMS") && atoi(getenv("DH_PHASE_STREAMS")) == 2;
    return two;
}

dh_bgzf_dev *dh_bgzf_dev_load(dh_ctx *c, const uint8_t *comp, uint64_t comp_size) {
    if (!c || !comp) { set_error("dh_bgzf_dev_load: bad arguments"); return nullptr; }
    if (cudaSetDevice(c->device) != cudaSuccess) { set_error("cudaSetDevice failed"); return nullptr; }
    dh_bgzf_dev *d = new dh_bgzf_dev();
    d->ctx = c;
    uint64_t bad = 0;
    if (walk_blocks(comp, comp_size, 0, d->bt, &bad) != DH_OK) {
        set_error("invalid BGZF header at offset %llu", (unsigned long long)bad);
        delete d;
        return nullptr;
    }
    const size_t nb = d->bt.blocks.size();
    const uint64_t cbytes = d->bt.end_off;
    bool ok = d->comp.reserve(cbytes + DH_COMP_PAD) == DH_OK && d->out.reserve(d->bt.uoff.back() + BGZF_MAX_OUT) == DH_OK &&
              d->desc.reserve((nb + 1) * sizeof(BlockDesc)) == DH_OK && d->status.reserve((nb + 1) * sizeof(int)) == DH_OK;
    if (!ok) { dh_bgzf_dev_free(d); return nullptr; }
    // segments: in_off is 32-bit, so a launch covers at most 2 GiB of compressed bytes
    std::vector<BlockDesc> descs(nb);
    size_t b = 0;
    while (b < nb) {
        size_t e = b;
        uint64_t cbase = d->bt.blocks[b].coffset & ~15ull;
        // a launch covers at most 2 GiB of compressed bytes (32-bit in_off) and 2^17 blocks (bounded work buffer)
        const size_t quantum = bgzf_inflate_v3_round(c->sm_count);   // streams phase 1 decodes at a time: launch whole rounds
        const size_t max_blocks = phase_streams_two() ? 2u * quantum : (1u << 17);   // two streams: shorter launches interleave better
        while (e < nb && e - b < max_blocks && d->bt.blocks[e].coffset + d->bt.blocks[e].csize - cbase <= (2ull << 30)) ++e;
        if (e < nb && e - b > quantum) e = b + (e - b) / quantum * quantum;
        const uint64_t cw = fill_descs(d->bt, b, e, cbase, 0, 0, descs.data() + b);
        d->segs.push_back({b, e, cbase, cw});
        b = e;
    }
    cudaError_t e1 = cudaMemcpy(d->comp.p, comp, cbytes, cudaMemcpyHostToDevice);
    cudaError_t e2 = cudaMemset((uint8_t *)d->comp.p + cbytes, 0, DH_COMP_PAD);
    cudaError_t e3 = nb ? cudaMemcpy(d->desc.p, descs.data(), nb * sizeof(BlockDesc), cudaMemcpyHostToDevice) : cudaSuccess;
    if (e1 != cudaSuccess || e2 != cudaSuccess || e3 != cudaSuccess) {
        cuda_fail(e1 != cudaSuccess ? e1 : (e2 != cudaSuccess ? e2 : e3), "dh_bgzf_dev_load upload");
        dh_bgzf_dev_free(d);
        return nullptr;
    }
    return d;
}

void dh_bgzf_dev_free(dh_bgzf_dev *d) {
    if (!d) return;
    cudaSetDevice(d->ctx->device);
    cudaStreamSynchronize(d->ctx->stream);
    d->comp.release(); d->out.release(); d->desc.release(); d->status.release(); d->work.release(); d->work2.release();
    for (int i = 0; i < 2; ++i) { if (d->ev_mid[i]) cudaEventDestroy(d->ev_mid[i]); if (d->ev_tail[i]) cudaEventDestroy(d->ev_tail[i]); }
    delete d;
}
uint64_t dh_bgzf_dev_nblocks(const dh_bgzf_dev *d) { return d ? d->bt.blocks.size() : 0; }
uint64_t dh_bgzf_dev_comp_bytes(const dh_bgzf_dev *d) { return d ? d->bt.end_off : 0; }
uint64_t dh_bgzf_dev_out_bytes(const dh_bgzf_dev *d) { return d ? d->bt.uoff.back() : 0; }

int dh_bgzf_dev_inflate(dh_bgzf_dev *d) {
    if (!d) { set_error("dh_bgzf_dev_inflate: null handle"); return DH_EINVAL; }
    dh_ctx *c = d->ctx;
    DH_CUDA(cudaSetDevice(c->device));
    if (phase_streams_two() && d->segs.size() > 1) {
        // phase 1 of every launch on the context's stream, phase 2 (+ filter) on the second stream; two work areas in turn
        for (int i = 0; i < 2; ++i) {
            if (!d->ev_mid[i]) DH_CUDA(cudaEventCreateWithFlags(&d->ev_mid[i], cudaEventDisableTiming));
            if (!d->ev_tail[i]) DH_CUDA(cudaEventCreateWithFlags(&d->ev_tail[i], cudaEventDisableTiming));
        }
        for (size_t i = 0; i < d->segs.size(); ++i) {
            auto &s = d->segs[i];
            const int q = (int)(i & 1);
            if (i >= 2) DH_CUDA(cudaStreamWaitEvent(c->stream, d->ev_tail[q], 0));      // launch i - 2 has left this work area
            int rc = inflate_blocks(c, (const uint8_t *)d->comp.p + s.cbase, (const BlockDesc *)d->desc.p + s.b0, (uint32_t)(s.b1 - s.b0),
                                    s.cmd_words, (uint8_t *)d->out.p, (int *)d->status.p + s.b0, q ? d->work2 : d->work,
                                    (unsigned int *)c->counter.p + 4 * q, c->stream, 0, c->inf2_stream, d->ev_mid[q]);
            if (rc != DH_OK) return rc;
            DH_CUDA(cudaEventRecord(d->ev_tail[q], c->inf2_stream));
        }
        DH_CUDA(cudaStreamWaitEvent(c->stream, d->ev_tail[(d->segs.size() - 1) & 1], 0));   // the context's stream ends behind the last phase 2
        return DH_OK;
    }
    for (auto &s : d->segs) {
        int rc = inflate_blocks(c, (const uint8_t *)d->comp.p + s.cbase, (const BlockDesc *)d->desc.p + s.b0, (uint32_t)(s.b1 - s.b0),
                                s.cmd_words, (uint8_t *)d->out.p, (int *)d->status.p + s.b0, d->work, (unsigned int *)c->counter.p,
                                c->stream);
        if (rc != DH_OK) return rc;
    }
    return DH_OK;
}

int dh_bgzf_dev_inflate_timed(dh_bgzf_dev *d, float *ms) {
    if (!d || !ms) { set_error("dh_bgzf_dev_inflate_timed: bad arguments"); return DH_EINVAL; }
    dh_ctx *c = d->ctx;
    DH_CUDA(cudaSetDevice(c->device));
    for (int i = 0; i < 4; ++i)
        if (!c->ev_ph[i]) DH_CUDA(cudaEventCreate(&c->ev_ph[i]));
    ms[0] = ms[1] = ms[2] = 0.f;
    c->time_phases = true;
    int rc = DH_OK;
    for (auto &s : d->segs) {
        rc = inflate_blocks(c, (const uint8_t *)d->comp.p + s.cbase, (const BlockDesc *)d->desc.p + s.b0, (uint32_t)(s.b1 - s.b0),
                            s.cmd_words, (uint8_t *)d->out.p, (int *)d->status.p + s.b0, d->work, (unsigned int *)c->counter.p, c->stream);
        if (rc != DH_OK) break;
        if (cudaEventSynchronize(c->ev_ph[3]) != cudaSuccess) { rc = cuda_fail(cudaGetLastError(), "cudaEventSynchronize"); break; }
        for (int i = 0; i < 3; ++i) {
            float t = 0.f;
            cudaEventElapsedTime(&t, c->ev_ph[i], c->ev_ph[i + 1]);
            ms[i] += t;
        }
    }
    c->time_phases = false;
    return rc;
}

int dh_bgzf_dev_check(dh_bgzf_dev *d, int64_t *bad_block, int32_t *bad_code) {
    if (!d) { set_error("dh_bgzf_dev_check: null handle"); return DH_EINVAL; }
    dh_ctx *c = d->ctx;
    DH_CUDA(cudaSetDevice(c->device));
    const size_t nb = d->bt.blocks.size();
    std::vector<int> hs(nb);
    DH_CUDA(cudaStreamSynchronize(c->stream));
    if (nb) DH_CUDA(cudaMemcpy(hs.data(), d->status.p, nb * sizeof(int), cudaMemcpyDeviceToHost));
    if (bad_block) *bad_block = -1;
    if (bad_code) *bad_code = 0;
    for (size_t i = 0; i < nb; ++i)
        if (hs[i]) {
            if (bad_block) *bad_block = (int64_t)i;
            if (bad_code) *bad_code = hs[i];
            set_error("BGZF block %zu failed inflate/CRC check (code %d)", i, hs[i]);
            return DH_EDATA;
        }
    return DH_OK;
}

int dh_bgzf_dev_read(dh_bgzf_dev *d, uint64_t offset, uint8_t *dst, uint64_t n) {
    if (!d || !dst || offset + n > d->bt.uoff.back()) { set_error("dh_bgzf_dev_read: bad range"); return DH_EINVAL; }
    DH_CUDA(cudaSetDevice(d->ctx->device));
    DH_CUDA(cudaStreamSynchronize(d->ctx->stream));
    if (n) DH_CUDA(cudaMemcpy(dst, (const uint8_t *)d->out.p + offset, n, cudaMemcpyDeviceToHost));
    return DH_OK;
}

// ---- timing / measurement helpers -----------------------------------------------------------------------
int dh_ctx_sync(dh_ctx *c) {
    if (!c) return DH_EINVAL;
    DH_CUDA(cudaSetDevice(c->device));
    DH_CUDA(cudaStreamSynchronize(c->stream));
    for (int i = 0; i < 3; ++i) DH_CUDA(cudaStreamSynchronize(c->copy_streams[i]));
    DH_CUDA(cudaStreamSynchronize(c->inf2_stream));
    return DH_OK;
}
int dh_timer_start(dh_ctx *c) {
    if (!c) return DH_EINVAL;
    DH_CUDA(cudaSetDevice(c->device));
    DH_CUDA(cudaEventRecord(c->ev_timer[0], c->stream));
    return DH_OK;
}
int dh_timer_stop(dh_ctx *c, float *ms) {
    if (!c || !ms) return DH_EINVAL;
    DH_CUDA(cudaSetDevice(c->device));
    DH_CUDA(cudaEventRecord(c->ev_timer[1], c->stream));
    DH_CUDA(cudaEventSynchronize(c->ev_timer[1]));
    DH_CUDA(cudaEventElapsedTime(ms, c->ev_timer[0], c->ev_timer[1]));
    return DH_OK;
}
int dh_flush_l2(dh_ctx *c) {
    if (!c) return DH_EINVAL;
    DH_CUDA(cudaSetDevice(c->device));
    const size_t n = 256u << 20;  // 2x the 126 MB L2
    int rc = c->l2_scratch.reserve(n);
    if (rc != DH_OK) return rc;
    DH_CUDA(cudaMemsetAsync(c->l2_scratch.p, 0x5a, n, c->stream));
    return DH_OK;
}
int dh_measure_pcie(dh_ctx *c, uint64_t bytes, float *h2d_gbs, float *d2h_gbs) {
    if (!c || !bytes) return DH_EINVAL;
    DH_CUDA(cudaSetDevice(c->device));
    void *h = nullptr, *d = nullptr;
    DH_CUDA(cudaMallocHost(&h, bytes));
    cudaError_t e = cudaMalloc(&d, bytes);
    if (e != cudaSuccess) { cudaFreeHost(h); return cuda_fail(e, "cudaMalloc"); }
    memset(h, 1, bytes);
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    float best_h2d = 0.f, best_d2h = 0.f;
    for (int it = 0; it < 4; ++it) {
        float ms;
        cudaEventRecord(a, c->stream);
        cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, c->stream);
        cudaEventRecord(b, c->stream);
        cudaEventSynchronize(b);
        cudaEventElapsedTime(&ms, a, b);
        best_h2d = std::max(best_h2d, (float)(bytes / 1e6 / ms));
        cudaEventRecord(a, c->stream);
        cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, c->stream);
        cudaEventRecord(b, c->stream);
        cudaEventSynchronize(b);
        cudaEventElapsedTime(&ms, a, b);
        best_d2h = std::max(best_d2h, (float)(bytes / 1e6 / ms));
    }
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    cudaFree(d);
    cudaFreeHost(h);
    if (h2d_gbs) *h2d_gbs = best_h2d;
    if (d2h_gbs) *d2h_gbs = best_d2h;
    cudaError_t le = cudaGetLastError();
    if (le != cudaSuccess) return cuda_fail(le, "dh_measure_pcie");
    return DH_OK;
}

// both directions at once on two streams: what the read_bam pipeline sees (compressed bytes in, columns out)
int dh_measure_pcie_bidir(dh_ctx *c, uint64_t bytes, float *h2d_gbs, float *d2h_gbs) {
    if (!c || !bytes) return DH_EINVAL;
    DH_CUDA(cudaSetDevice(c->device));
    void *h1 = nullptr, *h2 = nullptr, *d1 = nullptr, *d2 = nullptr;
    DH_CUDA(cudaMallocHost(&h1, bytes));
    DH_CUDA(cudaMallocHost(&h2, bytes));
    if (cudaMalloc(&d1, bytes) != cudaSuccess || cudaMalloc(&d2, bytes) != cudaSuccess) {
        cudaFreeHost(h1); cudaFreeHost(h2); cudaFree(d1);
        return cuda_fail(cudaGetLastError(), "cudaMalloc");
    }
    memset(h1, 1, bytes);
    memset(h2, 2, bytes);
    cudaEvent_t a0, a1, b0, b1;
    cudaEventCreate(&a0); cudaEventCreate(&a1); cudaEventCreate(&b0); cudaEventCreate(&b1);
    float best_a = 0.f, best_b = 0.f;
    for (int it = 0; it < 3; ++it) {
        cudaEventRecord(a0, c->copy_streams[0]);
        cudaEventRecord(b0, c->copy_streams[1]);
        for (int r = 0; r < 4; ++r) {
            cudaMemcpyAsync(d1, h1, bytes, cudaMemcpyHostToDevice, c->copy_streams[0]);
            cudaMemcpyAsync(h2, d2, bytes, cudaMemcpyDeviceToHost, c->copy_streams[1]);
        }
        cudaEventRecord(a1, c->copy_streams[0]);
        cudaEventRecord(b1, c->copy_streams[1]);
        cudaEventSynchronize(a1);
        cudaEventSynchronize(b1);
        float ma, mb;
        cudaEventElapsedTime(&ma, a0, a1);
        cudaEventElapsedTime(&mb, b0, b1);
        best_a = std::max(best_a, (float)(4.0 * bytes / 1e6 / ma));
        best_b = std::max(best_b, (float)(4.0 * bytes / 1e6 / mb));
    }
    cudaEventDestroy(a0); cudaEventDestroy(a1); cudaEventDestroy(b0); cudaEventDestroy(b1);
    cudaFree(d1); cudaFree(d2); cudaFreeHost(h1); cudaFreeHost(h2);
    if (h2d_gbs) *h2d_gbs = best_a;
    if (d2h_gbs) *d2h_gbs = best_b;
    cudaError_t le = cudaGetLastError();
    if (le != cudaSuccess) return cuda_fail(le, "dh_measure_pcie_bidir");
    return DH_OK;
}

}  // extern "C"
