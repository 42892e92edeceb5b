// Host build of duckhts_b200/csrc/inflate_core.cuh with a trivial IO policy.
// TEST INFRASTRUCTURE ONLY: lets the CPU box unit-test the exact table-builder / decode-loop source the
// CUDA kernel runs (tests/test_inflate_core.py compares it with zlib).  Never linked into the product.
#include <string.h>
#include "../../duckhts_b200/csrc/inflate_core.cuh"
#include "../../duckhts_b200/csrc/crc32_core.cuh"

struct HostIO {
    const uint8_t *in; uint32_t in_len;
    uint8_t *out; uint32_t cap; uint32_t pos = 0;
    uint32_t load32(uint32_t p) const {
        uint32_t v = 0;
        for (int i = 0; i < 4; ++i) if (p + i < in_len) v |= (uint32_t)in[p + i] << (8 * i);
        return v;
    }
    void literal(uint32_t b) { if (pos < cap) out[pos] = (uint8_t)b; pos++; }
    bool match(uint32_t len, uint32_t dist) {
        if (dist > pos || pos + len > cap) return false;
        for (uint32_t i = 0; i < len; ++i) out[pos + i] = out[pos + i - dist];
        pos += len; return true;
    }
    bool stored(uint32_t p, uint32_t len) {
        if (pos + len > cap) return false;
        memcpy(out + pos, in + p, len); pos += len; return true;
    }
    uint32_t produced() const { return pos; }
    bool leader() const { return true; }
    void sync() const {}
};

extern "C" int dh_host_inflate(const uint8_t *in, uint32_t in_len, uint8_t *out, uint32_t cap, uint32_t *out_len) {
    static dh::InflateTables t;
    HostIO io{in, in_len, out, cap};
    dh::Inflater<HostIO> inf(io, t, in_len);
    int rc = inf.run();
    *out_len = io.pos;
    return rc;
}

// CRC-32 by G independent segments + GF(2) combine, exactly as the lane-group does it
extern "C" uint32_t dh_host_crc32_segments(const uint8_t *p, uint32_t n, int G) {
    static uint32_t tab[8 * 256]; static uint32_t x2n[32]; static int init = 0;
    if (!init) { for (int i = 0; i < 256; ++i) dh::crc_table_column(tab, i); dh::crc_x2n_init(x2n); init = 1; }
    uint32_t acc = 0;
    for (int lane = 0; lane < G; ++lane) {
        uint32_t b, e; dh::crc_segment(n, G, lane, b, e);
        uint32_t r = dh::crc_update(tab, lane == 0 ? 0xffffffffu : 0u, p + b, e - b);
        acc ^= dh::crc_shift(x2n, r, n - e);
    }
    return acc ^ 0xffffffffu;
}
