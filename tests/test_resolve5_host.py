"""Host build of the EXACT per-warp source of the inflate's phase-2 kernel (duckhts_b200/csrc/bgzf_resolve5.cuh) behind a SIMT
emulation (tests/host/resolve5_host.cpp: 32 lanes as coroutines, every warp intrinsic a rendezvous that fails when the lanes
diverge) vs zlib.  Covers what the GPU suite covers only by whole files: window slides, chunk / row bookkeeping, period folds,
in-row and cross-row sources, the first and last partial 16-byte unit at every alignment, the CRC catch-up, the footer verdicts."""
import ctypes as C
import os
import subprocess
import zlib

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "host", "libresolve5_host.so")
INF_ERR_ISIZE, INF_ERR_CRC = 14, 15


@pytest.fixture(scope="module")
def r5():
    subprocess.check_call(["make", "-s", "-C", os.path.join(HERE, "host"), "libresolve5_host.so"])
    lib = C.CDLL(SO)
    lib.dh_host_resolve5.restype = C.c_int
    lib.dh_host_resolve5.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_int, C.POINTER(C.c_uint64), C.c_char_p, C.c_int]
    return lib


def deflate(data, level=6, strategy=zlib.Z_DEFAULT_STRATEGY, mem=8):
    c = zlib.compressobj(level, zlib.DEFLATED, -15, mem, strategy)
    return c.compress(data) + c.flush()


def resolve(r5, raw, isize, crc, align=0, pf_mode=1):
    out = np.zeros(max(isize, 1) + 16, dtype=np.uint8)
    msg = C.create_string_buffer(256)
    nr = C.c_uint64(0)
    rc = r5.dh_host_resolve5(raw, len(raw), isize, crc, align, out.ctypes.data, pf_mode, C.byref(nr), msg, 256)
    assert rc != -1, "SIMT emulation: " + msg.value.decode()
    assert rc != -3, "write outside the block: " + msg.value.decode()
    return rc, out[:isize].tobytes(), nr.value


def check(r5, data, align=0, **kw):
    rc, out, nr = resolve(r5, deflate(data, **kw), len(data), zlib.crc32(data), align)
    assert rc == 0, rc
    assert out == data
    return nr


def payloads():
    rng = np.random.default_rng(17)
    acgt = bytes(rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), 65280))
    yield "empty", b""
    yield "one byte", b"a"
    yield "15 bytes", b"0123456789abcde"
    yield "abc x 20000 (one overlapping match, distance 3)", b"abc" * 20000
    yield "zeros (distance 1)", bytes(65280)
    yield "noise (stored / literals only)", rng.integers(0, 256, 65280, dtype=np.uint8).tobytes()
    yield "4-letter alphabet", rng.integers(0, 4, 65280, dtype=np.uint8).tobytes()
    yield "ACGT", acgt
    yield "text lines", ((b"SYN:1:%d:abc\t%d\tchr%d\n" * 2500) % tuple(x for i in range(2500) for x in (i, i * 37, i % 22)))[:65280]
    yield "full 64 KiB block", acgt + acgt[:256]
    # far sources: a 3 KB motif repeated at distances beyond the shared-memory window (1.5 - 3.4 KB), up to the 32 KB limit
    motif = rng.integers(0, 256, 3000, dtype=np.uint8).tobytes()
    filler = rng.integers(0, 256, 40000, dtype=np.uint8).tobytes()
    yield "far matches", motif + filler[:5000] + motif + filler[5000:33000] + motif + motif[:1000]
    for period in (1, 2, 3, 5, 7, 8, 15, 16, 17, 31, 32, 33, 63, 64, 65, 127, 255, 256, 257, 258, 259, 511, 1537, 4000):
        base = rng.integers(0, 256, period, dtype=np.uint8).tobytes()
        yield f"period {period}", (base * (30000 // period + 1))[:30000]
    # short matches at short distances: sources inside the same 32-byte row and the row in front of it
    words = [rng.integers(97, 123, int(n), dtype=np.uint8).tobytes() for n in rng.integers(3, 9, 40)]
    yield "short words", b"".join(words[int(i)] for i in rng.integers(0, 40, 9000))


@pytest.mark.parametrize("name,data", list(payloads()), ids=[n for n, _ in payloads()])
def test_resolve5_matches_zlib(r5, name, data):
    nr = check(r5, data)
    assert nr > 0 or not data
    if len(data) > 300:
        check(r5, data, align=77)


def test_every_alignment_of_the_output(r5):
    """The first and the last partial 16-byte unit leave by byte stores, everything between as st.global.v4: every offset of the
    block in its 128-byte frame, sizes around the unit and row sizes, nothing written outside the block (guard bytes)."""
    rng = np.random.default_rng(5)
    text = bytes(rng.choice(np.frombuffer(b"ACGTN", dtype=np.uint8), 5000))
    for align in range(128):
        check(r5, text[: 700 + align], align=align)
    for n in (1, 2, 15, 16, 17, 31, 32, 33, 111, 112, 113, 127, 128, 129, 255, 256, 257, 511, 512, 513, 1023, 1024, 1025):
        for align in (0, 1, 15, 16, 17, 111, 127):
            check(r5, text[:n], align=align)


def test_compression_levels_and_strategies(r5):
    rng = np.random.default_rng(9)
    data = bytes(rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), 40000)) + bytes(2000) + rng.integers(0, 256, 3000, dtype=np.uint8).tobytes()
    for level in (1, 6, 9):
        check(r5, data, level=level, align=5)
    check(r5, data, strategy=zlib.Z_FIXED, align=9)
    check(r5, data, strategy=zlib.Z_HUFFMAN_ONLY)
    check(r5, data, strategy=zlib.Z_RLE, align=100)
    check(r5, data, level=0)                                   # stored blocks: 255-literal runs only


def test_footer_verdicts(r5):
    data = b"the quick brown fox jumps over the lazy dog " * 500
    raw = deflate(data)
    assert resolve(r5, raw, len(data), zlib.crc32(data))[0] == 0
    assert resolve(r5, raw, len(data), zlib.crc32(data) ^ 1)[0] == INF_ERR_CRC
    assert resolve(r5, raw, len(data) + 1, zlib.crc32(data))[0] == INF_ERR_ISIZE
    assert resolve(r5, raw, len(data) - 1, zlib.crc32(data))[0] == INF_ERR_ISIZE


def test_the_simt_emulation_reports_divergence(r5):
    """The harness on itself: lanes that arrive at different intrinsics, and a lane that leaves while the others wait at one (both hang
    or corrupt a real warp), are reported; a well-formed exchange gives the hardware's results."""
    r5.dh_host_simt_selftest.restype = C.c_int
    msg = C.create_string_buffer(256)
    assert r5.dh_host_simt_selftest(msg, 256) == 7, msg.value


def test_seeded_mixtures(r5):
    """60 seeded mixtures of runs, repeats at random distances (1 .. 32 KiB), alphabet noise and text, random sizes and alignments."""
    rng = np.random.default_rng(2026)
    for case in range(60):
        parts, total = [], 0
        target = int(rng.integers(1, 65281))
        while total < target:
            kind = int(rng.integers(0, 5))
            n = int(rng.integers(1, 3000))
            if kind == 0:
                p = bytes([int(rng.integers(0, 256))]) * n
            elif kind == 1 and total > 0:
                whole = b"".join(parts)
                dist = int(rng.integers(1, min(total, 32768) + 1))
                p = (whole[total - dist:] * (n // dist + 1))[:n]       # a copy from `dist` bytes back, overlapping when n > dist
            elif kind == 2:
                p = rng.integers(0, 256, n, dtype=np.uint8).tobytes()
            elif kind == 3:
                p = bytes(rng.choice(np.frombuffer(b"ACGTN\t\n0123", dtype=np.uint8), n))
            else:
                p = (b"chr%d\t%d\t.\tA\tG\t%d\tPASS\n" % (case % 22, total, n)) * (n // 20 + 1)
            parts.append(p[:n])
            total += len(parts[-1])
        data = b"".join(parts)[:target]
        check(r5, data, align=int(rng.integers(0, 128)), level=int(rng.choice([1, 6, 9])))
