"""Parity of the sm_100a BGZF inflate kernel (through the C-ABI) against the oracle.  Bit-exact."""
import zlib

import numpy as np
import pytest

import duckhts_b200 as dh
from oracle import dh_oracle_py as orc
from tests import util

pytestmark = pytest.mark.gpu


def check(ctx, img):
    rc, ref, bad = orc.bgzf_inflate(img)
    assert rc == 0, (rc, bad)
    out, st = dh.bgzf_inflate(ctx, img)
    assert st.bad_block == -1
    assert st.out_bytes == ref.size
    assert np.array_equal(out, ref)
    return st


def test_synthetic_bam_matches_oracle(ctx):
    for n, seed, paired in ((1, 1, False), (500, 2, False), (20000, 3, True), (200000, 4, True)):
        st = check(ctx, util.synth_bam_bytes(n_records=n, seed=seed, paired=paired))
        assert st.eof_marker == 1


def test_unaligned_bgzf_stream_matches_oracle(ctx):
    check(ctx, np.fromfile(util.synth_path("bgzf", n_records=50000, seed=5, paired=True), dtype=np.uint8))


def test_block_types_and_edge_cases(ctx):
    rng = np.random.default_rng(0)
    acgt = bytes(rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), 400000))
    noise = rng.integers(0, 256, 200000, dtype=np.uint8).tobytes()
    cases = [
        util.make_bgzf(b""),                                     # EOF block only
        util.make_bgzf(b"", eof=False),                          # empty image
        util.make_bgzf(b"x"),
        util.make_bgzf(acgt, level=6),
        util.make_bgzf(acgt, level=1),
        util.make_bgzf(acgt, level=9),
        util.make_bgzf(noise, level=6),                          # stored-block fallback (bgzf.c:631-638)
        util.make_bgzf(noise, level=0, block=60000),             # level 0: stored blocks
        util.make_bgzf(acgt, strategy=zlib.Z_FIXED),             # BTYPE=01
        util.make_bgzf(acgt, strategy=zlib.Z_HUFFMAN_ONLY),
        util.make_bgzf(bytes(300000)),                           # dist=1 runs, 258-byte matches
        util.make_bgzf(acgt, block=1),                           # 400k one-byte blocks? too many: trimmed below
    ]
    cases[-1] = util.make_bgzf(acgt[:3000], block=1)
    for period in (1, 2, 3, 5, 7, 8, 9, 15, 16, 17, 31, 33, 258, 259, 32768):
        base = rng.integers(0, 256, period, dtype=np.uint8).tobytes()
        cases.append(util.make_bgzf((base * (200000 // period + 1))[:200000]))
    # ragged block sizes incl. 65536-byte blocks (BGZF_MAX_BLOCK_SIZE), mid-file empty blocks (bgzf.c:1145-1228)
    img = b"".join([util.bgzf_block(bytes(65536)), util.BGZF_EOF, util.bgzf_block(acgt[:12345]), util.bgzf_block(b""),
                    util.bgzf_block(acgt[:65536], level=9), util.BGZF_EOF])
    cases.append(img)
    for img in cases:
        check(ctx, img)


def test_corruption_is_detected_at_the_same_block(ctx):
    img = bytearray(util.synth_bam_bytes(n_records=20000, seed=3, paired=True))
    blocks, _ = dh.bgzf_walk(bytes(img))
    assert len(blocks) > 20
    # (a) flip a payload bit in block 7 -> zlib error or CRC mismatch in the reference; (b) CRC footer of block 11
    for blk, where in ((7, "payload"), (11, "crc"), (13, "isize")):
        bad = bytearray(img)
        off, cs, isz, crc = blocks[blk]
        if where == "payload":
            bad[off + 18 + (cs - 26) // 2] ^= 0x10
        elif where == "crc":
            bad[off + cs - 8] ^= 1
        else:
            bad[off + cs - 4] ^= 1
        rc, _, ref_bad = orc.bgzf_inflate(bytes(bad), want_output=False)
        with pytest.raises(dh.DuckhtsError) as ei:
            dh.bgzf_inflate(ctx, bytes(bad))
        st = ei.value.stats
        if where == "isize":
            # the reference ignores ISIZE (dlen comes from zlib, bgzf.c:791); we are stricter by design
            assert rc == 0 and st.bad_block == blk and st.bad_code == 14
        else:
            assert rc in (-2, -3) and st.bad_block == ref_bad == blk


def test_device_resident_path_and_repeatability(ctx):
    img = util.synth_bam_bytes(n_records=200000, seed=4, paired=True)
    rc, ref, _ = orc.bgzf_inflate(img)
    dev = dh.BgzfDevice(ctx, img)
    for _ in range(3):
        dev.inflate()
        ctx.sync()
        assert dev.check()[0] == 0
    assert np.array_equal(dev.read(), ref)
    dev.free()


def test_checksum_of_checksums_large(ctx):
    """Size-independent property at a larger size: every block's fused CRC32/ISIZE verdict is OK and the
    per-block CRCs recomputed on the host from the returned bytes equal the footers."""
    img = util.synth_bam_bytes(n_records=1000000, seed=6, paired=True)
    out, st = dh.bgzf_inflate(ctx, img)
    assert st.bad_block == -1
    blocks, _ = dh.bgzf_walk(img)
    pos = 0
    for k in range(0, len(blocks), max(1, len(blocks) // 200)):
        pos = sum(b[2] for b in blocks[:k])
        off, cs, isz, crc = blocks[k]
        assert zlib.crc32(out[pos:pos + isz].tobytes()) & 0xffffffff == crc
    assert st.out_bytes == sum(b[2] for b in blocks)


def test_table_pool_overflow_paths(ctx):
    """Streams whose Huffman tables outgrow the decoder's table pool: 1 chain fits the main pass (6 warps per SM, 589
    entries per stream), 2 chains are queued for the retry pass (4 warps, 892 entries), 3 chains need the lane-group
    kernel.  Mixed with ordinary blocks so that all three paths run inside one launch; the bytes must not differ."""
    rng = np.random.default_rng(5)
    acgt = bytes(rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), 300000))
    parts = []
    for rep in range(3):
        for k in (1, 2, 3, 2, 1, 3):
            payload, raw = util.chain_code_deflate(k, seed=10 * rep + k)
            parts.append(util.bgzf_block(payload, raw_deflate=raw))
            parts.append(util.make_bgzf(acgt[rep * 50000:(rep + 2) * 50000], eof=False))
    parts.append(util.BGZF_EOF)
    check(ctx, b"".join(parts))
    # a launch where EVERY block overflows the main pool (the retry queue holds all of them)
    blocks = []
    for i in range(70):
        payload, raw = util.chain_code_deflate(2, n_tokens=3000, seed=100 + i)
        blocks.append(util.bgzf_block(payload, raw_deflate=raw))
    check(ctx, b"".join(blocks) + util.BGZF_EOF)
