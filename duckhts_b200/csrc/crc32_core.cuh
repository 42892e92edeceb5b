// crc32_core.cuh — CRC-32 (IEEE 802.3, reflected, poly 0xEDB88320) pieces used by the fused
// CRC check of the inflate kernel.  Replaces zlib's crc32() call at htslib/bgzf.c:793
// (`crc32(crc32(0L, NULL, 0L), dst, *dlen)` compared with the BGZF footer).
//
// A lane-group splits the block's output into G contiguous segments; each lane runs a slice-by-8
// table CRC over its segment from raw state 0 (lane 0 from 0xFFFFFFFF), then multiplies its
// register by x^(8*bytes_after_segment) mod P in GF(2)[x]; the XOR over lanes, inverted, is the CRC.
#pragma once
#include <stdint.h>
#ifndef DH_HD
#if defined(__CUDACC__)
#define DH_HD __host__ __device__ __forceinline__
#else
#define DH_HD static inline
#endif
#endif

namespace dh {

constexpr uint32_t CRC_POLY = 0xEDB88320u;

// fill column i (0..255) of the 8 slice tables: tab[k*256+i]
DH_HD void crc_table_column(uint32_t *tab, int i) {
    uint32_t c = (uint32_t)i;
    for (int k = 0; k < 8; ++k) c = (c & 1u) ? (c >> 1) ^ CRC_POLY : (c >> 1);
    tab[i] = c;
    // tab[k][i] = CRC register after byte i followed by k zero bytes
    for (int k = 1; k < 8; ++k) {
        uint32_t z = c;
        for (int b = 0; b < 8; ++b) z = (z & 1u) ? (z >> 1) ^ CRC_POLY : (z >> 1);
        c = z;
        tab[k * 256 + i] = c;
    }
}

// a(x)*b(x) mod P(x); reflected representation: bit 31 is the coefficient of x^0
DH_HD uint32_t crc_mulmod(uint32_t a, uint32_t b) {
    uint32_t p = 0;
#pragma unroll 4
    for (int i = 0; i < 32; ++i) {
        p ^= (a & (0x80000000u >> i)) ? b : 0u;
        b = (b >> 1) ^ ((b & 1u) ? CRC_POLY : 0u);
    }
    return p;
}

// x2n[k] = x^(2^k) mod P
DH_HD void crc_x2n_init(uint32_t *x2n) {
    uint32_t p = 0x40000000u;  // x^1
    x2n[0] = p;
    for (int k = 1; k < 32; ++k) { p = crc_mulmod(p, p); x2n[k] = p; }
}

// register r advanced over nbytes zero bytes == r * x^(8*nbytes) mod P
DH_HD uint32_t crc_shift(const uint32_t *x2n, uint32_t r, uint32_t nbytes) {
    uint32_t e = nbytes;  // exponent in units of 8 bits: start at table index 3
    int k = 3;
    while (e) {
        if (e & 1u) r = crc_mulmod(x2n[k], r);
        e >>= 1;
        ++k;
    }
    return r;
}

// segment [b,e) of n bytes owned by `lane` of G; segment length is a multiple of 8
DH_HD void crc_segment(uint32_t n, int G, int lane, uint32_t &b, uint32_t &e) {
    uint32_t seg = (((n + G - 1) / G) + 7u) & ~7u;
    b = seg * (uint32_t)lane;
    e = b + seg;
    if (b > n) b = n;
    if (e > n) e = n;
}

// raw register update over p[0..n)
DH_HD uint32_t crc_update(const uint32_t *tab, uint32_t r, const uint8_t *p, uint32_t n) {
    // head: to 8-byte alignment
    while (n && ((uintptr_t)p & 7u)) { r = tab[(r ^ *p++) & 0xffu] ^ (r >> 8); --n; }
    while (n >= 8) {
        uint64_t w = *reinterpret_cast<const uint64_t *>(p);
        uint32_t lo = (uint32_t)w ^ r, hi = (uint32_t)(w >> 32);
        r = tab[7 * 256 + (lo & 0xffu)] ^ tab[6 * 256 + ((lo >> 8) & 0xffu)] ^ tab[5 * 256 + ((lo >> 16) & 0xffu)] ^
            tab[4 * 256 + (lo >> 24)] ^ tab[3 * 256 + (hi & 0xffu)] ^ tab[2 * 256 + ((hi >> 8) & 0xffu)] ^
            tab[1 * 256 + ((hi >> 16) & 0xffu)] ^ tab[hi >> 24];
        p += 8;
        n -= 8;
    }
    while (n) { r = tab[(r ^ *p++) & 0xffu] ^ (r >> 8); --n; }
    return r;
}

}  // namespace dh
