// text_split.cu — lines and tab-separated fields of an inflated text stream (the read_tabix / VCF.gz text path
// on the same inflate kernels).
//
// Replaces the per-line loop of the reference's tabix_scan (src/tabix_reader.c:852-1030): hts_getline (bgzf_getline,
// htslib/bgzf.c:2308-2370: bytes up to '\n', a trailing '\r' dropped) and get_field (src/tabix_reader.c:234-254: the idx-th
// '\t'-separated span of the line).  Three kernels over one segment of text:
//   text_count_kernel   one warp per 4 KB: number of '\n' bytes (16-byte loads, __vcmpeq4 + popc)
//   (an exclusive scan of the counts: text_scan_kernel, one CTA)
//   text_lines_kernel   the same walk again, now writing the offset of every '\n'
//   text_fields_kernel  one warp per line: tabs by ballot, every projected field's [begin, end) offsets, flags for empty
//                       lines (1) and lines that start with the meta character (2)
#include "text_split.cuh"

namespace dh {
namespace txt {

constexpr unsigned FULL = 0xffffffffu;
constexpr uint32_t CHUNK = 4096;

__device__ __forceinline__ uint32_t nl_mask16(const uint4 v) {       // bit i = byte i of the 16 is '\n'
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    uint32_t m = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint32_t e = __vcmpeq4(w[k], 0x0a0a0a0au);            // 0xff per equal byte
        m |= (((e >> 7) & 1u) | ((e >> 14) & 2u) | ((e >> 21) & 4u) | ((e >> 28) & 8u)) << (4 * k);
    }
    return m;
}

// text must be readable up to the next multiple of 16 behind n (the caller pads); bytes at and behind n are ignored
__global__ void __launch_bounds__(256) text_count_kernel(const uint8_t *__restrict__ text, uint64_t n, uint32_t *__restrict__ counts) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint64_t chunk = (uint64_t)blockIdx.x * 8u + (threadIdx.x >> 5);
    const uint64_t lo = chunk * CHUNK;
    if (lo >= n) return;
    uint32_t c = 0;
#pragma unroll
    for (uint32_t it = 0; it < CHUNK / 512u; ++it) {
        const uint64_t off = lo + it * 512u + lane * 16u;
        if (off < n) {
            uint32_t m = nl_mask16(*reinterpret_cast<const uint4 *>(text + off));
            if (n - off < 16u) m &= (1u << (uint32_t)(n - off)) - 1u;
            c += (uint32_t)__popc(m);
        }
    }
#pragma unroll
    for (int s = 16; s; s >>= 1) c += __shfl_xor_sync(FULL, c, s);
    if (lane == 0) counts[chunk] = c;
}

// exclusive scan of counts[0..n) in place + the total in counts[n]; one CTA
__global__ void __launch_bounds__(1024) text_scan_kernel(uint32_t *__restrict__ counts, uint32_t n) {
    __shared__ uint32_t part[1024];
    const uint32_t t = threadIdx.x, per = (n + 1023u) / 1024u;
    const uint32_t lo = t * per < n ? t * per : n, hi = lo + per < n ? lo + per : n;
    uint32_t sum = 0;
    for (uint32_t i = lo; i < hi; ++i) sum += counts[i];
    part[t] = sum;
    __syncthreads();
    for (uint32_t d = 1; d < 1024u; d <<= 1) {
        const uint32_t v = t >= d ? part[t - d] : 0u;
        __syncthreads();
        part[t] += v;
        __syncthreads();
    }
    uint32_t run = t ? part[t - 1] : 0u;
    for (uint32_t i = lo; i < hi; ++i) { const uint32_t c = counts[i]; counts[i] = run; run += c; }
    if (t == 1023u) counts[n] = part[1023];
}

__global__ void __launch_bounds__(256) text_lines_kernel(const uint8_t *__restrict__ text, uint64_t n, const uint32_t *__restrict__ first,
                                                         uint32_t *__restrict__ nl) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint64_t chunk = (uint64_t)blockIdx.x * 8u + (threadIdx.x >> 5);
    const uint64_t lo = chunk * CHUNK;
    if (lo >= n) return;
    uint32_t at = first[chunk];
    for (uint32_t it = 0; it < CHUNK / 512u; ++it) {
        const uint64_t off = lo + it * 512u + lane * 16u;
        uint32_t m = 0;
        if (off < n) {
            m = nl_mask16(*reinterpret_cast<const uint4 *>(text + off));
            if (n - off < 16u) m &= (1u << (uint32_t)(n - off)) - 1u;
        }
        const uint32_t c = (uint32_t)__popc(m);
        uint32_t inc = c;
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) {
            const uint32_t v = __shfl_up_sync(FULL, inc, s);
            if (lane >= (uint32_t)s) inc += v;
        }
        uint32_t k = at + inc - c;
        while (m) {
            const uint32_t b = (uint32_t)__ffs((int)m) - 1u;
            m &= m - 1u;
            nl[k++] = (uint32_t)(off + b);
        }
        at += __shfl_sync(FULL, inc, 31);
    }
}

__global__ void __launch_bounds__(256) text_fields_kernel(const FieldParams p) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint64_t line = (uint64_t)blockIdx.x * 8u + (threadIdx.x >> 5);
    if (line >= p.n_lines) return;
    const uint32_t start = line ? p.nl[line - 1] + 1u : 0u;
    uint32_t end = p.nl[line];
    if (end > start && p.text[end - 1] == '\r') --end;               // bgzf_getline drops the '\r' of "\r\n"
    uint32_t *cells = p.cells + line * p.n_proj * 2u;
    for (uint32_t i = lane; i < p.n_proj; i += 32u) cells[2u * i] = 0xffffffffu;      // absent until a tab or the line's end says otherwise
    __syncwarp();
    uint32_t ntab = 0;
    for (uint32_t base = start; base < end; base += 32u) {
        const uint32_t off = base + lane;
        const bool tab = off < end && p.text[off] == '\t';
        const uint32_t m = __ballot_sync(FULL, tab);
        if (tab) {
            const uint32_t k = ntab + (uint32_t)__popc(m & ((1u << lane) - 1u));       // this tab ends field k and field k + 1 begins behind it
            if (k < p.n_cols && p.slot[k] != 0xffffu) cells[2u * p.slot[k] + 1u] = off;
            if (k + 1u < p.n_cols && p.slot[k + 1u] != 0xffffu) cells[2u * p.slot[k + 1u]] = off + 1u;
        }
        ntab += (uint32_t)__popc(m);
    }
    if (lane == 0) {
        if (p.n_cols && p.slot[0] != 0xffffu) cells[2u * p.slot[0]] = start;
        if (ntab < p.n_cols && p.slot[ntab] != 0xffffu) cells[2u * p.slot[ntab] + 1u] = end;
        p.flags[line] = (uint8_t)((end == start ? 1u : 0u) | (end > start && p.meta && p.text[start] == (uint8_t)p.meta ? 2u : 0u));
    }
}

}  // namespace txt

cudaError_t launch_text_count(const uint8_t *text, uint64_t n, uint32_t *counts, cudaStream_t stream) {
    const uint64_t chunks = (n + txt::CHUNK - 1) / txt::CHUNK;
    if (chunks == 0) return cudaMemsetAsync(counts, 0, 4, stream);
    txt::text_count_kernel<<<(unsigned)((chunks + 7) / 8), 256, 0, stream>>>(text, n, counts);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    txt::text_scan_kernel<<<1, 1024, 0, stream>>>(counts, (uint32_t)chunks);
    return cudaGetLastError();
}
cudaError_t launch_text_lines(const uint8_t *text, uint64_t n, const uint32_t *first, uint32_t *nl, cudaStream_t stream) {
    const uint64_t chunks = (n + txt::CHUNK - 1) / txt::CHUNK;
    if (chunks == 0) return cudaSuccess;
    txt::text_lines_kernel<<<(unsigned)((chunks + 7) / 8), 256, 0, stream>>>(text, n, first, nl);
    return cudaGetLastError();
}
cudaError_t launch_text_fields(const FieldParams &p, cudaStream_t stream) {
    if (p.n_lines == 0) return cudaSuccess;
    txt::text_fields_kernel<<<(unsigned)((p.n_lines + 7) / 8), 256, 0, stream>>>(p);
    return cudaGetLastError();
}
uint64_t text_chunks(uint64_t n) { return (n + txt::CHUNK - 1) / txt::CHUNK; }

}  // namespace dh
