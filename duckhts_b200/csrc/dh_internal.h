// dh_internal.h — internals shared by the translation units of libduckhts_cuda.so (not part of the C-ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/duckhts_cuda.h"
#include "bgzf_inflate.cuh"

namespace dh {

void set_error(const char *fmt, ...);
int cuda_fail(cudaError_t e, const char *what);  // records the message, returns DH_ECUDA

#define DH_CUDA(call)                                             \
    do {                                                          \
        cudaError_t e__ = (call);                                 \
        if (e__ != cudaSuccess) return ::dh::cuda_fail(e__, #call); \
    } while (0)

// growable device / pinned-host buffers (never shrink; reused across calls)
struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    int dev = 0;     // device the buffer lives on
    int reserve(size_t n);
    void release();
};
struct PinBuf {
    void *p = nullptr;
    size_t cap = 0;
    int reserve(size_t n);
    void release();
};

// host-side block table for a BGZF image (A1)
struct BlockTable {
    std::vector<dh_bgzf_block> blocks;
    std::vector<uint64_t> uoff;  // inflated offset of each block; size nblocks+1
    uint64_t end_off = 0;
    bool eof_marker = false;
};
// walk data[start..size); returns DH_OK / DH_EFORMAT (bad header at *bad_off)
int walk_blocks(const uint8_t *data, uint64_t size, uint64_t start, BlockTable &bt, uint64_t *bad_off);

// Device block descriptors for blocks [b0,b1): compressed bytes were uploaded starting at file offset `comp_base`
// (16-byte aligned); block i's bytes go to out + out_bias + (uoff[i] - uoff[out_b0]).  Returns the work-area bytes the
// lockstep kernel's work buffer needs for these blocks.
uint64_t fill_descs(const BlockTable &bt, size_t b0, size_t b1, uint64_t comp_base, size_t out_b0, uint64_t out_bias,
                    BlockDesc *d);
// Inflate `n` described blocks on `stream`: lockstep kernels (phase 1 + 2), then the lane-group kernel in filter mode
// for blocks the compact tables could not hold.  `work` is grown as needed.  counter: >= 2 device words.
int inflate_blocks(dh_ctx *c, const uint8_t *comp, const BlockDesc *descs, uint32_t n, uint64_t cmd_words, uint8_t *out,
                   int *status, DevBuf &work, unsigned int *counter, cudaStream_t stream, int sm_limit = 0,
                   cudaStream_t stream2 = nullptr, cudaEvent_t mid = nullptr);   // stream2 + mid: phase 2 and the filter run on stream2

}  // namespace dh

// pread / pwrite of [off, off + n) by several threads (dh_api.cu)
extern "C" bool dh_parallel_io(int fd, uint8_t *buf, uint64_t n, uint64_t off, bool write);

struct dh_ctx {
    int device = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;        // the stream every kernel is launched on
    cudaStream_t inf2_stream = nullptr;   // phase 2 of the inflate when the scan pipeline splits the phases over two streams (DH_PHASE_STREAMS=2)
    cudaStream_t copy_streams[3] = {};    // pipeline slots for the *_host entry points
    cudaEvent_t ev_timer[2] = {};
    dh::DevBuf l2_scratch;
    dh::DevBuf counter;                   // inflate ticket counters (one word per pipeline slot + main)
    uint64_t launches = 0;
    // per-phase timing of the inflate (dh_bgzf_dev_inflate_timed): events before phase 1, between the phases, after
    // phase 2, after the fallback filter
    bool time_phases = false;
    cudaEvent_t ev_ph[4] = {};
    // pipeline slots of dh_bgzf_inflate_host
    struct Slot {
        dh::DevBuf comp, out, desc, status, work;
        dh::PinBuf h_desc, h_status;
        dh::PinBuf h_comp, h_out;            // staging of the streaming file-to-file inflate (dh_bgzf_inflate_fd)
        cudaEvent_t ev_k0 = nullptr, ev_k1 = nullptr, ev_done = nullptr;
    } slots[3];
};
