// dh_deflate.cu — C-ABI of the BGZF writer: dh_bgzf_deflate_host / dh_bgzf_deflate_fd.
//
// Mirrors what the reference's bgzip() does with htslib (src/bgzip.c:209-273: a read(2) / bgzf_write loop over 64 KB buffers,
// bgzf_close appends the EOF marker): the input travels in groups of DEFLATE_GROUP blocks of 0xff00 bytes through three
// staging slots — read -> H2D -> bgzf_deflate_kernel + pack -> D2H -> write — the host I/O of one group overlapping the GPU
// work of the others.  Nothing is held whole.  No CPU compressor anywhere: without a GPU the entry points fail.
#include <errno.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>

#include "bgzf_deflate.cuh"
#include "dh_internal.h"

using namespace dh;

namespace {

constexpr uint32_t DEFLATE_GROUP = 8192;       // blocks per launch: 534 MB of input
const uint8_t EOF_BLOCK[28] = {0x1f, 0x8b, 0x08, 0x04, 0, 0, 0, 0, 0, 0xff, 0x06, 0, 0x42, 0x43, 0x02, 0, 0x1b, 0, 0x03, 0, 0, 0, 0, 0, 0, 0, 0, 0};

struct Slot {
    DevBuf in, slots, scratch, csize, offs, out;
    PinBuf h_in, h_out, h_total;
    cudaEvent_t ev_k0 = nullptr, ev_k1 = nullptr, ev_o = nullptr;
    uint64_t in_bytes = 0, total = 0;
    uint32_t nblocks = 0;
    int stage = 0;            // 0 free, 1 kernels enqueued, 2 output copy enqueued
    void release() {
        in.release(); slots.release(); scratch.release(); csize.release(); offs.release(); out.release();
        h_in.release(); h_out.release(); h_total.release();
        for (cudaEvent_t e : {ev_k0, ev_k1, ev_o}) if (e) cudaEventDestroy(e);
    }
};

// where the bytes come from / go to
struct Source {
    const uint8_t *mem; int fd; uint64_t size;
    bool read(uint8_t *dst, uint64_t off, uint64_t n) const {
        if (fd < 0) { memcpy(dst, mem + off, n); return true; }
        return dh_parallel_io(fd, dst, n, off, false);
    }
};
struct Sink {
    uint8_t *mem; uint64_t cap; int fd; uint64_t pos;
    int write(const uint8_t *src, uint64_t n) {
        if (fd < 0) {
            if (pos + n > cap) { set_error("output buffer too small (%llu bytes needed so far, %llu given)", (unsigned long long)(pos + n), (unsigned long long)cap); return DH_ENOMEM; }
            memcpy(mem + pos, src, n);
        } else {
            uint64_t done = 0;
            while (done < n) {
                const ssize_t k = ::write(fd, src + done, (size_t)std::min<uint64_t>(n - done, 1ull << 30));
                if (k < 0 && errno == EINTR) continue;
                if (k <= 0) { set_error("write: %s", strerror(errno)); return DH_EIO; }
                done += (uint64_t)k;
            }
        }
        pos += n;
        return DH_OK;
    }
};

int deflate_stream(dh_ctx *c, const Source &src, Sink &sink, int level, dh_deflate_stats *st) {
    DH_CUDA(cudaSetDevice(c->device));
    dh_deflate_stats local;
    if (!st) st = &local;
    memset(st, 0, sizeof *st);
    if (level < -1 || level > 9) { set_error("compression level %d out of range (-1..9)", level); return DH_EINVAL; }
    if (level < 0) level = 6;
    static const uint32_t group_max = getenv("DH_DEFLATE_GROUP") ? (uint32_t)std::max(1, atoi(getenv("DH_DEFLATE_GROUP"))) : DEFLATE_GROUP;
    const uint64_t nb_total = (src.size + DEFLATE_BLOCK_IN - 1) / DEFLATE_BLOCK_IN;
    const uint32_t group = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(group_max, nb_total));     // small inputs take small buffers
    const uint64_t gbytes = (uint64_t)group * DEFLATE_BLOCK_IN;
    const uint64_t ngroups = (src.size + gbytes - 1) / gbytes;
    Slot sl[3];
    int rc = DH_OK;
    float kernel_ms = 0.f;
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr;
    auto cleanup = [&]() {
        for (int i = 0; i < 3; ++i) cudaStreamSynchronize(c->copy_streams[i]);
        for (auto &s : sl) s.release();
        if (ev_begin) cudaEventDestroy(ev_begin);
        if (ev_end) cudaEventDestroy(ev_end);
    };
    if (cudaEventCreate(&ev_begin) != cudaSuccess || cudaEventCreate(&ev_end) != cudaSuccess) { cleanup(); return cuda_fail(cudaGetLastError(), "cudaEventCreate"); }
    cudaEventRecord(ev_begin, c->copy_streams[0]);

    // stage A: read group g, upload, compress, pack; the packed size comes back in h_total
    auto start = [&](uint64_t g) -> int {
        Slot &s = sl[g % 3];
        cudaStream_t strm = c->copy_streams[g % 3];
        const uint64_t off = g * gbytes, nbytes = std::min<uint64_t>(gbytes, src.size - off);
        const uint32_t nb = (uint32_t)((nbytes + DEFLATE_BLOCK_IN - 1) / DEFLATE_BLOCK_IN);
        int r;
        if ((r = s.in.reserve(gbytes + DEFLATE_IN_PAD)) != DH_OK || (r = s.slots.reserve((size_t)group * DEFLATE_SLOT_STRIDE)) != DH_OK ||
            (r = s.scratch.reserve(deflate_scratch_bytes(group))) != DH_OK || (r = s.csize.reserve((size_t)group * 4)) != DH_OK ||
            (r = s.offs.reserve(((size_t)group + 1) * 8)) != DH_OK || (r = s.out.reserve((size_t)group * 65536u + 256)) != DH_OK ||
            (r = s.h_in.reserve(gbytes)) != DH_OK || (r = s.h_out.reserve((size_t)group * 65536u)) != DH_OK || (r = s.h_total.reserve(64)) != DH_OK)
            return r;
        if (!s.ev_k0) { DH_CUDA(cudaEventCreate(&s.ev_k0)); DH_CUDA(cudaEventCreate(&s.ev_k1)); DH_CUDA(cudaEventCreate(&s.ev_o)); }
        if (!src.read((uint8_t *)s.h_in.p, off, nbytes)) { set_error("read: %s", strerror(errno)); return DH_EIO; }
        s.in_bytes = nbytes;
        s.nblocks = nb;
        DH_CUDA(cudaMemcpyAsync(s.in.p, s.h_in.p, nbytes, cudaMemcpyHostToDevice, strm));
        DH_CUDA(cudaMemsetAsync((uint8_t *)s.in.p + nbytes, 0, DEFLATE_IN_PAD, strm));
        DH_CUDA(cudaEventRecord(s.ev_k0, strm));
        DH_CUDA(launch_bgzf_deflate((const uint8_t *)s.in.p, nbytes, nb, level, (uint8_t *)s.slots.p, (uint8_t *)s.scratch.p, (uint32_t *)s.csize.p,
                                    (uint64_t *)s.offs.p, (uint8_t *)s.out.p, strm));
        c->launches += 3;
        DH_CUDA(cudaEventRecord(s.ev_k1, strm));
        DH_CUDA(cudaMemcpyAsync(s.h_total.p, (const uint64_t *)s.offs.p + nb, 8, cudaMemcpyDeviceToHost, strm));
        DH_CUDA(cudaEventRecord(s.ev_o, strm));
        s.stage = 1;
        st->h2d_bytes += nbytes;
        st->n_blocks += nb;
        return DH_OK;
    };
    // stage B: the size is known, fetch the packed blocks
    auto fetch = [&](uint64_t g) -> int {
        Slot &s = sl[g % 3];
        cudaStream_t strm = c->copy_streams[g % 3];
        DH_CUDA(cudaEventSynchronize(s.ev_o));
        float ms = 0.f;
        DH_CUDA(cudaEventElapsedTime(&ms, s.ev_k0, s.ev_k1));
        kernel_ms += ms;
        s.total = *(const uint64_t *)s.h_total.p;
        if (s.total > (uint64_t)s.nblocks * 65536u) { set_error("internal: packed size %llu beyond the bound", (unsigned long long)s.total); return DH_EINVAL; }
        DH_CUDA(cudaMemcpyAsync(s.h_out.p, s.out.p, s.total, cudaMemcpyDeviceToHost, strm));
        DH_CUDA(cudaEventRecord(s.ev_o, strm));
        s.stage = 2;
        st->d2h_bytes += s.total + 8;
        return DH_OK;
    };
    // stage C: write them
    auto finish = [&](uint64_t g) -> int {
        Slot &s = sl[g % 3];
        DH_CUDA(cudaEventSynchronize(s.ev_o));
        s.stage = 0;
        return sink.write((const uint8_t *)s.h_out.p, s.total);
    };
    for (uint64_t g = 0; g < ngroups + 2 && rc == DH_OK; ++g) {
        if (g >= 2 && g - 2 < ngroups) rc = finish(g - 2);          // frees slot (g - 2) % 3 == (g + 1) % 3 ... and g % 3 was freed one round ago
        if (rc == DH_OK && g < ngroups) rc = start(g);
        if (rc == DH_OK && g >= 1 && g - 1 < ngroups) rc = fetch(g - 1);
    }
    if (rc == DH_OK) rc = sink.write(EOF_BLOCK, sizeof EOF_BLOCK);   // bgzf_close (htslib/bgzf.c:2138-2150)
    cudaEventRecord(ev_end, c->copy_streams[0]);
    for (int i = 0; i < 3; ++i) cudaStreamSynchronize(c->copy_streams[i]);
    cudaEventSynchronize(ev_end);
    cudaEventElapsedTime(&st->total_ms, ev_begin, ev_end);
    st->kernel_ms = kernel_ms;
    st->in_bytes = src.size;
    st->out_bytes = sink.pos;
    cleanup();
    return rc;
}

}  // namespace

extern "C" {

uint64_t dh_bgzf_deflate_bound(uint64_t in_size) { return (in_size + DEFLATE_BLOCK_IN - 1) / DEFLATE_BLOCK_IN * 65536ull + 28; }

int dh_bgzf_deflate_host(dh_ctx *c, const uint8_t *in, uint64_t in_size, int level, uint8_t *out, uint64_t out_cap, uint64_t *out_size,
                         dh_deflate_stats *stats) {
    if (!c || (!in && in_size) || !out || !out_size) { set_error("dh_bgzf_deflate_host: bad arguments"); return DH_EINVAL; }
    Source src{in, -1, in_size};
    Sink sink{out, out_cap, -1, 0};
    const int rc = deflate_stream(c, src, sink, level, stats);
    *out_size = sink.pos;
    return rc;
}

int dh_bgzf_deflate_fd(dh_ctx *c, int in_fd, int out_fd, int level, dh_deflate_stats *stats) {
    if (!c || in_fd < 0 || out_fd < 0) { set_error("dh_bgzf_deflate_fd: bad arguments"); return DH_EINVAL; }
    struct stat sb;
    if (fstat(in_fd, &sb) != 0) { set_error("fstat: %s", strerror(errno)); return DH_EIO; }
    Source src{nullptr, in_fd, (uint64_t)sb.st_size};
    Sink sink{nullptr, 0, out_fd, 0};
    return deflate_stream(c, src, sink, level, stats);
}

}  // extern "C"
