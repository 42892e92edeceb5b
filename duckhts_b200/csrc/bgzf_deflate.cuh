// bgzf_deflate.cuh — BGZF compression kernels: launcher + buffer geometry.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace dh {

constexpr uint32_t DEFLATE_BLOCK_IN = 0xff00u;            // BGZF_BLOCK_SIZE: input bytes per block (htslib/bgzf.c:95)
constexpr uint32_t DEFLATE_SLOT_STRIDE = 65536u + 256u;   // where a block is put together before the blocks are packed
constexpr uint32_t DEFLATE_IN_PAD = 64u;                  // readable bytes the kernels may touch behind the input

size_t deflate_scratch_bytes(uint32_t nblocks);

// Compress in[0, in_bytes) as ceil(in_bytes / 0xff00) BGZF blocks (the last one short; none for an empty input).
//   slots    nblocks * DEFLATE_SLOT_STRIDE bytes, 256-byte aligned
//   scratch  deflate_scratch_bytes(nblocks)
//   csize    nblocks words: the size of every finished block
//   offs     nblocks + 1 words: where each block starts in `out`; offs[nblocks] = the total
//   out      packed blocks (capacity >= nblocks * 65536)
// level 0 writes stored blocks, any other level runs the one LZ77 + dynamic-Huffman kernel.
cudaError_t launch_bgzf_deflate(const uint8_t *in, uint64_t in_bytes, uint32_t nblocks, int level, uint8_t *slots, uint8_t *scratch,
                                uint32_t *csize, uint64_t *offs, uint8_t *out, cudaStream_t stream);

}  // namespace dh
