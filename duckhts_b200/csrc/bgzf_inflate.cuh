// bgzf_inflate.cuh — sm_100a BGZF inflate kernel: device-side types + launcher declaration.
//
// Hot-path row A2 of SURVEY.md §8: replaces bgzf_uncompress()/inflate_block()/bgzf_decode_func()
// (htslib/bgzf.c:762-804, 808-824, 1373-1384) — raw DEFLATE inflate of each BGZF block with the
// CRC-32 and ISIZE footer checks fused into the same kernel.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace dh {

// One BGZF block as the host framing pass (A1, bgzf_walk in dh_api.cu) describes it to the device.
struct BlockDesc {
    uint32_t in_off;   // offset of the raw DEFLATE payload (block start + 18) inside the compressed buffer
    uint32_t in_len;   // payload length = BSIZE - 18 - 8
    uint32_t isize;    // ISIZE footer: expected inflated length
    uint32_t crc;      // CRC-32 footer
    uint64_t out_off;  // where the block's bytes go in the inflated buffer (prefix sum of ISIZE)
    uint64_t cmd_off;  // lockstep kernels: byte offset of the block's literal/command area inside the work buffer
};

// Lockstep kernels, work area of one block of `isize` inflated bytes: the literal stream grows upwards from the start,
// the command stream (one 32-bit word per match, one per 255-literal run) downwards from the end.  With L literal
// bytes and M matches, L + 3M <= isize, so L + 4M + 4(L/255 + 1) <= 4/3 isize + 4; 64 bytes of slack, 16-byte aligned.
#if defined(__CUDACC__)
__host__ __device__
#endif
inline uint32_t inflate_work_cap(uint32_t isize) { return (isize + isize / 3u + 64u + 15u) & ~15u; }
// in front of the areas: the per-block command counts, then the retry queue of phase 1 (a 16-byte header + 32 bytes per block, in 64 + 32 n bytes)
inline size_t inflate_work_retry(uint32_t nblocks) { return (size_t)((nblocks + 63u) & ~63u) * 4u; }
inline size_t inflate_work_header(uint32_t nblocks) { return inflate_work_retry(nblocks) + 64u + (size_t)((nblocks + 63u) & ~63u) * 32u; }
// work-buffer bytes for a launch over `nblocks` blocks whose areas total `area_bytes`
inline size_t inflate_work_bytes(uint32_t nblocks, uint64_t area_bytes) { return inflate_work_header(nblocks) + (size_t)area_bytes + 256u; }

// Launch the persistent inflate kernel over `nblocks` blocks.
//   comp     device buffer holding the compressed bytes; must be 16-byte aligned and padded with
//            >= DH_COMP_PAD readable bytes after the last block (TMA prefetch runs ahead)
//   out      device buffer for the inflated bytes
//   status   one int per block: 0 = OK, else dh::INF_ERR_*
//   counter  device word used as the work-distribution ticket; zeroed by this call on `stream`
// Returns the cudaError of the launch.
cudaError_t launch_bgzf_inflate(const uint8_t *comp, const BlockDesc *blocks, uint32_t nblocks, uint8_t *out,
                                int *status, unsigned int *counter, int sm_count, cudaStream_t stream,
                                bool only_flagged = false);

// Third-generation ("lockstep") kernels (bgzf_inflate3.cu): phase 1 decodes one stream per lane into a literal stream +
// match commands, phase 2 (a warp per block) places literals, replays the matches and checks CRC-32/ISIZE.
// `work` = inflate_work_bytes(nblocks, sum of inflate_work_cap).  Blocks flagged INF_NEEDS_V1 must then be redone with
// launch_bgzf_inflate(..., only_flagged = true).
cudaError_t launch_bgzf_inflate_v3(const uint8_t *comp, const BlockDesc *blocks, uint32_t nblocks, uint8_t *out, int *status,
                                   uint8_t *work, unsigned int *counter, int sm_count, cudaStream_t stream,
                                   cudaEvent_t between_phases = nullptr, cudaStream_t stream2 = nullptr, cudaEvent_t mid = nullptr);

// blocks per launch that keep every SM busy to the end of phase 1: launch multiples of this (see bgzf_inflate3.cu)
uint32_t bgzf_inflate_v3_round(int sm_count);

// {index of the first block whose status is not 0, or -1; its status} -> out[0..1]; `out` may be pinned host memory
cudaError_t launch_status_first_bad(const int *status, uint32_t nblocks, int *out, cudaStream_t stream);

constexpr uint32_t DH_COMP_PAD = 8192;

}  // namespace dh
