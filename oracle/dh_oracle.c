/* dh_oracle.c — CPU restatement of the reference's hot path.  TEST INFRASTRUCTURE ONLY (see dh_oracle.h).
 * Parity pinned by tests/test_oracle.py against reference-generated fixtures in tests/golden/. */
#include "dh_oracle.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <zlib.h>

static inline uint32_t le16(const uint8_t *p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8); }
static inline uint32_t le32(const uint8_t *p) {
    return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
}

/* htslib/bgzf.c:896-903 check_header */
static int check_header(const uint8_t *h) {
    if (h[0] != 31 || h[1] != 139 || h[2] != 8) return -2;
    return ((h[3] & 4) != 0 && le16(h + 10) == 6 && h[12] == 'B' && h[13] == 'C' && le16(h + 14) == 2) ? 0 : -1;
}

int64_t dho_bgzf_walk(const uint8_t *data, uint64_t size, dho_block *out, int64_t cap, uint64_t *end_off) {
    uint64_t off = 0;
    int64_t n = 0;
    while (off < size) {
        /* bgzf.c:1147-1163: a short or invalid 18-byte header is BGZF_ERR_HEADER */
        if (size - off < 18 || check_header(data + off) != 0) return -1;
        uint32_t bsize = le16(data + off + 16) + 1; /* bgzf.c:1198 */
        if (bsize < 18 + 8 || size - off < bsize) return -1; /* bgzf.c:1199-1212 */
        if (out && n < cap) {
            out[n].coffset = off;
            out[n].csize = bsize;
            out[n].crc = le32(data + off + bsize - 8);
            out[n].isize = le32(data + off + bsize - 4);
            out[n].reserved = 0;
        }
        ++n;
        off += bsize;
    }
    if (end_off) *end_off = off;
    return n;
}

/* htslib/bgzf.c:762-804 bgzf_uncompress (zlib variant), called as in inflate_block :808-824:
 * src = block + 18, slen = block_length - 18 (footer included), dlen = 65536. */
int dho_bgzf_inflate_block(const uint8_t *block, uint32_t csize, uint8_t *dst) {
    z_stream zs;
    memset(&zs, 0, sizeof zs);
    zs.next_in = (Bytef *)(block + 18);
    zs.avail_in = csize - 18;
    zs.next_out = dst;
    zs.avail_out = 65536;
    if (inflateInit2(&zs, -15) != Z_OK) return -2;
    if (inflate(&zs, Z_FINISH) != Z_STREAM_END) {
        inflateEnd(&zs);
        return -2;
    }
    if (inflateEnd(&zs) != Z_OK) return -2;
    uint32_t dlen = 65536 - zs.avail_out;
    uint32_t crc = (uint32_t)crc32(crc32(0L, NULL, 0L), dst, dlen);
    if (crc != le32(block + csize - 8)) return -3;
    return (int)dlen;
}

int dho_bgzf_inflate(const uint8_t *data, uint64_t size, uint8_t *out, uint64_t cap, uint64_t *out_len,
                     int64_t *bad_block) {
    uint64_t off = 0, o = 0;
    int64_t b = 0;
    uint8_t tmp[65536];
    if (bad_block) *bad_block = -1;
    while (off < size) {
        if (size - off < 18 || check_header(data + off) != 0) { if (bad_block) *bad_block = b; return -1; }
        uint32_t bsize = le16(data + off + 16) + 1;
        if (bsize < 26 || size - off < bsize) { if (bad_block) *bad_block = b; return -1; }
        int n = dho_bgzf_inflate_block(data + off, bsize, tmp);
        if (n < 0) { if (bad_block) *bad_block = b; if (out_len) *out_len = o; return n; }
        if (out) {
            if (o + (uint64_t)n > cap) { if (bad_block) *bad_block = b; return -4; }
            memcpy(out + o, tmp, (size_t)n);
        }
        o += (uint64_t)n;
        off += bsize;
        ++b;
    }
    if (out_len) *out_len = o;
    return 0;
}

#include "dh_oracle_bam.inc"
#include "dh_oracle_idx.inc"
