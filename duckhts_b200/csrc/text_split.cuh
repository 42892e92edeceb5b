// text_split.cuh — line / field splitting kernels for the text path; see text_split.cu.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace dh {

struct FieldParams {
    const uint8_t *text;      // the segment's bytes (16-byte aligned, padded to a multiple of 16)
    const uint32_t *nl;       // offset of every '\n' (the last line of a file without one gets a '\n' written behind it)
    uint64_t n_lines;
    uint32_t n_cols;          // fields per line the schema has
    uint32_t n_proj;          // how many of them are returned
    const uint16_t *slot;     // [n_cols]: position of field k among the returned ones, 0xffff = not returned
    uint32_t *cells;          // [n_lines][n_proj][2]: begin, end offsets into text; begin = 0xffffffff: the line has no such field
    uint8_t *flags;           // [n_lines]: 1 = an empty line, 2 = a line that starts with the meta character
    int meta;                 // 0 = none
};

uint64_t text_chunks(uint64_t n);     // counts[] needs text_chunks(n) + 1 words
// counts[c] = number of '\n' in front of 4 KB chunk c, counts[text_chunks(n)] = the total
cudaError_t launch_text_count(const uint8_t *text, uint64_t n, uint32_t *counts, cudaStream_t stream);
cudaError_t launch_text_lines(const uint8_t *text, uint64_t n, const uint32_t *first, uint32_t *nl, cudaStream_t stream);
cudaError_t launch_text_fields(const FieldParams &p, cudaStream_t stream);

}  // namespace dh
