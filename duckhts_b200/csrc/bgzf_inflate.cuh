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
    uint64_t cmd_off;  // lockstep kernel: first word of the block's match-command area inside the work buffer
};

// match-command words a block of `isize` inflated bytes can need: one per match (>= 3 bytes each) + one skip word per
// literal run longer than 255 bytes
inline uint64_t cmd_words_for(uint32_t isize) { return (uint64_t)isize / 3u + (uint64_t)isize / 256u + 4u; }
// work-buffer bytes for a launch over `nblocks` blocks whose command areas total `cmd_words`
inline size_t inflate_work_bytes(uint32_t nblocks, uint64_t cmd_words) { return ((size_t)((nblocks + 63u) & ~63u) + cmd_words) * 4u; }

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

// Second-generation ("lockstep") kernels: phase 1 decodes one stream per lane into literals + match commands,
// phase 2 replays the matches and checks CRC-32/ISIZE.  `work` = inflate_work_bytes(nblocks, sum of cmd areas).
// Blocks flagged INF_NEEDS_V1 must then be redone with launch_bgzf_inflate(..., only_flagged = true).
cudaError_t launch_bgzf_inflate_v2(const uint8_t *comp, const BlockDesc *blocks, uint32_t nblocks, uint8_t *out, int *status,
                                   uint32_t *work, unsigned int *counter, int sm_count, cudaStream_t stream);

constexpr uint32_t DH_COMP_PAD = 8192;

}  // namespace dh
