"""Host build of the EXACT decoder-core source the CUDA kernel runs (inflate_core.cuh, crc32_core.cuh) vs zlib."""
import ctypes as C
import os
import zlib

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "host", "libinflate_core_host.so")


@pytest.fixture(scope="module")
def core():
    if not os.path.exists(SO):
        import subprocess

        subprocess.check_call(["make", "-C", os.path.join(HERE, "host")])
    lib = C.CDLL(SO)
    lib.dh_host_inflate.restype = C.c_int
    lib.dh_host_inflate.argtypes = [C.c_char_p, C.c_uint32, C.c_void_p, C.c_uint32, C.POINTER(C.c_uint32)]
    lib.dh_host_crc32_segments.restype = C.c_uint32
    lib.dh_host_crc32_segments.argtypes = [C.c_char_p, C.c_uint32, C.c_int]
    return lib


def inflate(core, raw, cap=65536):
    out = np.zeros(cap + 8, dtype=np.uint8)
    n = C.c_uint32(0)
    rc = core.dh_host_inflate(raw, len(raw), out.ctypes.data, cap, C.byref(n))
    return rc, out[: n.value].tobytes()


def deflate(data, level=6, strategy=zlib.Z_DEFAULT_STRATEGY, mem=8):
    c = zlib.compressobj(level, zlib.DEFLATED, -15, mem, strategy)
    return c.compress(data) + c.flush()


def payloads():
    rng = np.random.default_rng(11)
    yield b""
    yield b"a"
    yield b"abc" * 20000
    yield bytes(65280)
    yield rng.integers(0, 256, 65280, dtype=np.uint8).tobytes()  # incompressible
    yield rng.integers(0, 4, 65280, dtype=np.uint8).tobytes()
    yield bytes(rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), 65280))
    text = (b"SYN:1:%d:abc\t" * 3000) % tuple(range(3000))
    yield text[:65280]
    # long-distance matches and every short period
    for period in (1, 2, 3, 4, 5, 7, 8, 9, 15, 16, 17, 31, 32, 33, 255, 258, 259, 4096, 32768):
        base = rng.integers(0, 256, period, dtype=np.uint8).tobytes()
        yield (base * (65280 // period + 1))[:65280]


@pytest.mark.parametrize("level,strategy", [(6, zlib.Z_DEFAULT_STRATEGY), (1, zlib.Z_DEFAULT_STRATEGY), (9, zlib.Z_DEFAULT_STRATEGY),
                                            (0, zlib.Z_DEFAULT_STRATEGY), (6, zlib.Z_FIXED), (6, zlib.Z_HUFFMAN_ONLY), (6, zlib.Z_RLE)])
def test_roundtrip_against_zlib(core, level, strategy):
    for p in payloads():
        raw = deflate(p, level, strategy)
        if len(raw) > 65536:
            continue
        rc, out = inflate(core, raw)
        assert rc == 0, (rc, len(p), level, strategy)
        assert out == p


def test_multi_block_streams_and_sync_flush(core):
    c = zlib.compressobj(6, zlib.DEFLATED, -15, 8)
    parts, raw = [], b""
    rng = np.random.default_rng(5)
    for i in range(6):
        part = rng.integers(0, 8, 5000, dtype=np.uint8).tobytes()
        parts.append(part)
        raw += c.compress(part) + c.flush(zlib.Z_SYNC_FLUSH if i % 2 else zlib.Z_FULL_FLUSH)  # stored empty blocks in between
    raw += c.flush()
    rc, out = inflate(core, raw)
    assert rc == 0 and out == b"".join(parts)


def test_corrupt_streams_are_rejected_like_zlib(core):
    p = bytes(np.random.default_rng(2).integers(0, 16, 30000, dtype=np.uint8))
    raw = bytearray(deflate(p))
    rng = np.random.default_rng(9)
    checked = 0
    for _ in range(300):
        bad = bytearray(raw)
        k = int(rng.integers(0, len(bad)))
        bad[k] ^= 1 << int(rng.integers(0, 8))
        try:
            d = zlib.decompressobj(-15)
            ref = d.decompress(bytes(bad), 65536)
            ok = d.eof and not d.unconsumed_tail
        except zlib.error:
            ok, ref = False, None
        rc, out = inflate(core, bytes(bad))
        if ok:
            assert rc == 0 and out == ref
        else:
            # zlib rejected (or ran out of input/over 64 KiB): we must not return success with different bytes
            assert rc != 0 or out != p
        checked += 1
    assert checked == 300
    # truncated input
    rc, _ = inflate(core, bytes(raw[: len(raw) // 2]))
    assert rc != 0
    # reserved block type
    rc, _ = inflate(core, b"\x07")
    assert rc == 1
    # stored block with bad NLEN
    rc, _ = inflate(core, b"\x01\x05\x00\x00\x00hello")
    assert rc == 2


def test_output_limit_is_64k(core):
    rc, out = inflate(core, deflate(bytes(65536)))
    assert rc == 0 and len(out) == 65536
    rc, _ = inflate(core, deflate(bytes(65537)))
    assert rc != 0


def test_crc32_segmented(core):
    rng = np.random.default_rng(4)
    for n in (0, 1, 7, 8, 9, 63, 64, 65, 1000, 65280, 65536):
        p = rng.integers(0, 256, n, dtype=np.uint8).tobytes()
        for g in (1, 2, 4, 8, 16, 32):
            assert core.dh_host_crc32_segments(p, n, g) == (zlib.crc32(p) & 0xffffffff), (n, g)
