// dh_checksum.cu — per-column 64-bit checksums of a read_bam batch (dh_bam_batch_checksum, include/duckhts_cuda.h).
//
// Parity at BASELINE scale: a 100 M-record scan cannot be compared cell by cell with the reference, but the sum over
// rows of a per-cell value hash can be compared with the same sum over the inputs the scan is made of (bench.py: the
// image is `tiles` copies of a unit whose scan is checked cell by cell against the reference, so every column must sum
// to tiles x the unit's checksum modulo 2^64), between the two output layouts, between device-resident and host batches,
// and between one GPU and the ranges of a sharded scan.  One routine, compiled for the device (batches left in HBM) and
// for the host (batches that crossed PCIe), reads the cells exactly as a consumer would.
#include <thread>
#include <vector>

#include "bam_decode.cuh"

namespace dh {

namespace {

constexpr uint32_t LENF = 0x80000000u;
enum { K_SKIP = CK_SKIP, K_U16 = CK_U16, K_MAPQ = CK_MAPQ, K_POS = CK_POS, K_TLEN = CK_TLEN, K_STR = CK_STR, K_DICT = CK_DICT, K_SAMPLE = CK_SAMPLE,
       K_TAG_I = CK_TAG_I, K_TAG_F = CK_TAG_F, K_TAG_A = CK_TAG_A, K_TAG_B = CK_TAG_B };

__host__ __device__ inline uint32_t a4(uint32_t n) { return (n + 3u) & ~3u; }
__host__ __device__ inline bool bit(const unsigned long long *v, uint64_t r) { return !v || ((v[r >> 6] >> (r & 63)) & 1ull); }
__host__ __device__ inline uint32_t strnlen_u8(const uint8_t *p, uint32_t n) {
    uint32_t i = 0;
    while (i < n && p[i]) ++i;
    return i;
}

// heap bytes the compact-layout slot of column c takes in row r (0 for NULL cells)
__host__ __device__ inline uint32_t compact_slot(const ChecksumParams &p, int c, uint64_t r) {
    const uint8_t k = p.kind[c];
    if (k == K_STR) {
        if (!bit(p.dp.valid[c], r)) return 0u;
        return a4(reinterpret_cast<const uint32_t *>(p.dp.col[c])[r] & ~LENF);
    }
    if (k == K_TAG_B) {
        if (!bit(p.dp.valid[c], r)) return 0u;
        return 4u + 8u * (uint32_t)reinterpret_cast<const uint64_t *>(p.dp.col[c])[2 * r + 1];
    }
    return 0u;
}

// slot order of the compact layout: QNAME CIGAR SEQ QUAL, tag columns in request order, READ_GROUP_ID
__host__ __device__ inline int slot_rank_next(const ChecksumParams &p, int prev) {
    const int first_tag = BAM_CORE_COLS < (int)p.n_cols ? BAM_CORE_COLS : COL_RG;
    switch (prev) {
    case -1: return COL_QNAME;
    case COL_QNAME: return COL_CIGAR;
    case COL_CIGAR: return COL_SEQ;
    case COL_SEQ: return COL_QUAL;
    case COL_QUAL: return first_tag;
    case COL_RG: return -2;
    default: return prev + 1 < (int)p.n_cols ? prev + 1 : COL_RG;
    }
}

__host__ __device__ inline const uint8_t *compact_slot_ptr(const ChecksumParams &p, int c, uint64_t r) {
    uint64_t off = p.dp.rec_off[r];
    for (int q = slot_rank_next(p, -1); q >= 0 && q != c; q = slot_rank_next(p, q))
        if (p.dp.col[q]) off += compact_slot(p, q, r);
    return p.dp.heap + off;
}

// DuckDB-layout VARCHAR cell -> hash of its bytes.  `device` = the batch lives in HBM: heap pointers are translated, and
// dictionary strings longer than 12 bytes (host memory the device cannot read) go through the pointer -> hash table.
template <bool DEVICE>
__host__ __device__ inline uint64_t cell_hash(const ChecksumParams &p, const uint8_t *cell) {
    const uint32_t len = *reinterpret_cast<const uint32_t *>(cell);
    if (len <= 12u) return ck_bytes(cell + 4, len);
    const uint64_t ptr = *reinterpret_cast<const uint64_t *>(cell + 8);
    if (!DEVICE) return ck_bytes(reinterpret_cast<const uint8_t *>((uintptr_t)ptr), len);
    if (ptr >= p.dp.heap_host && ptr < p.dp.heap_host + p.heap_bytes) return ck_bytes(p.dp.heap + (ptr - p.dp.heap_host), len);
    uint32_t lo = 0, hi = p.n_dict;
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (p.dict_ptr[mid] < ptr) lo = mid + 1; else hi = mid;
    }
    return (lo < p.n_dict && p.dict_ptr[lo] == ptr) ? p.dict_hash[lo] : 0ull;
}

template <bool DEVICE>
__host__ __device__ inline uint64_t value_hash(const ChecksumParams &p, int c, uint64_t r) {
    const uint8_t k = p.kind[c];
    const uint8_t *col = p.dp.col[c];
    const bool compact = p.compact != 0u;
    switch (k) {
    case K_U16: return ck_mix(reinterpret_cast<const uint16_t *>(col)[r]);
    case K_MAPQ: return ck_mix(compact ? (uint64_t)col[r] : (uint64_t)(int64_t)reinterpret_cast<const int32_t *>(col)[r]);
    case K_POS: return ck_mix(compact ? (uint64_t)((int64_t)reinterpret_cast<const int32_t *>(col)[r] + 1) : reinterpret_cast<const uint64_t *>(col)[r]);
    case K_TLEN: return ck_mix(compact ? (uint64_t)(int64_t)reinterpret_cast<const int32_t *>(col)[r] : reinterpret_cast<const uint64_t *>(col)[r]);
    case K_STR: {
        if (!bit(p.dp.valid[c], r)) return CK_NULL;
        if (!compact) return cell_hash<DEVICE>(p, col + 16 * r);
        const uint32_t v = reinterpret_cast<const uint32_t *>(col)[r];
        const uint8_t *s = compact_slot_ptr(p, c, r);
        uint32_t n = v & ~LENF;
        if (c == COL_QNAME) n = (v & LENF) ? strnlen_u8(s, n) : n - 1u;
        else if (v & LENF) n = strnlen_u8(s, n);
        return ck_bytes(s, n);
    }
    case K_DICT: {
        if (!compact) return cell_hash<DEVICE>(p, col + 16 * r);
        const int32_t tid = reinterpret_cast<const int32_t *>(col)[r];
        return p.ref_hash[tid >= 0 && tid < p.n_ref ? tid : p.n_ref];
    }
    case K_SAMPLE: {
        if (!bit(p.dp.valid[c], r)) return CK_NULL;
        if (!compact) return cell_hash<DEVICE>(p, col + 16 * r);
        return p.sample_hash[reinterpret_cast<const int32_t *>(col)[r]];
    }
    case K_TAG_I: case K_TAG_F:
        return bit(p.dp.valid[c], r) ? ck_mix(reinterpret_cast<const uint64_t *>(col)[r]) : CK_NULL;
    case K_TAG_A: {
        if (!bit(p.dp.valid[c], r)) return CK_NULL;
        if (!compact) return cell_hash<DEVICE>(p, col + 16 * r);
        const uint8_t ch = col[r];
        return ck_bytes(&ch, ch ? 1u : 0u);
    }
    case K_TAG_B: {
        if (!bit(p.dp.valid[c], r)) return CK_NULL;
        const uint64_t off8 = reinterpret_cast<const uint64_t *>(col)[2 * r], n = reinterpret_cast<const uint64_t *>(col)[2 * r + 1];
        return ck_words(reinterpret_cast<const uint64_t *>(p.dp.heap) + off8, n);
    }
    default: return 0ull;
    }
}

__global__ void __launch_bounds__(256) bam_checksum_kernel(const __grid_constant__ ChecksumParams p) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x, t0 = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (int c = 0; c < (int)p.n_cols; ++c) {
        if (p.kind[c] == K_SKIP || !p.dp.col[c]) continue;
        uint64_t acc = 0;
        for (uint64_t r = t0; r < p.n_rows; r += stride) acc += value_hash<true>(p, c, r);
#pragma unroll
        for (int sft = 16; sft; sft >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, sft);
        if (lane == 0 && acc) atomicAdd(reinterpret_cast<unsigned long long *>(p.sums) + c, (unsigned long long)acc);
    }
}

}  // namespace

cudaError_t launch_bam_checksum(const ChecksumParams &p, cudaStream_t stream) {
    if (p.n_rows == 0) return cudaSuccess;
    uint64_t blocks = (p.n_rows + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    bam_checksum_kernel<<<(unsigned)blocks, 256, 0, stream>>>(p);
    return cudaGetLastError();
}

void bam_checksum_host(const ChecksumParams &p, int n_threads) {
    if (p.n_rows == 0) return;
    if (n_threads < 1) n_threads = 1;
    if ((uint64_t)n_threads > (p.n_rows + 4095) / 4096) n_threads = (int)((p.n_rows + 4095) / 4096);
    std::vector<std::vector<uint64_t>> part((size_t)n_threads, std::vector<uint64_t>(BAM_MAX_COLS, 0));
    auto work = [&](int ti) {
        const uint64_t lo = p.n_rows * (uint64_t)ti / (uint64_t)n_threads, hi = p.n_rows * (uint64_t)(ti + 1) / (uint64_t)n_threads;
        for (int c = 0; c < (int)p.n_cols; ++c) {
            if (p.kind[c] == K_SKIP || !p.dp.col[c]) continue;
            uint64_t acc = 0;
            for (uint64_t r = lo; r < hi; ++r) acc += value_hash<false>(p, c, r);
            part[(size_t)ti][(size_t)c] = acc;
        }
    };
    std::vector<std::thread> th;
    for (int i = 1; i < n_threads; ++i) th.emplace_back(work, i);
    work(0);
    for (auto &t : th) t.join();
    for (int i = 0; i < n_threads; ++i)
        for (int c = 0; c < (int)p.n_cols; ++c) p.sums[c] += part[(size_t)i][(size_t)c];
}

}  // namespace dh
