"""dh_bgzf_inflate_noalloc (the inflate kernel on a fixed __device__ scratch array, no cudaMalloc) against the existing
dh_bgzf_inflate_host, through the C-ABI, byte for byte, on calls in a row.  Collected last on purpose (pytest -x)."""
import ctypes as C
import zlib

import numpy as np
import pytest

import duckhts_b200 as dh
from duckhts_b200 import _lib
from tests import util

pytestmark = pytest.mark.gpu

MAX_BLOCKS = 16384
PAD = 8192


def _walk(img):
    a = np.frombuffer(img, dtype=np.uint8) if not isinstance(img, np.ndarray) else img
    cap = a.size // 26 + 2
    blocks = (_lib.BgzfBlock * cap)()
    end = C.c_uint64(0)
    n = _lib.lib().dh_bgzf_walk(a.ctypes.data, a.size, 0, blocks, cap, C.byref(end), 1)
    assert n >= 0
    return a, blocks, int(n)


def _noalloc(img, stream=None):
    import torch

    a, blocks, n = _walk(img)
    total = sum(blocks[i].isize for i in range(n))
    d_comp = torch.zeros(a.size + PAD, dtype=torch.uint8, device="cuda:0")      # the caller's buffers: inputs and outputs
    d_comp[: a.size] = torch.from_numpy(a.copy())
    d_out = torch.full((max(total, 1),), 0xA5, dtype=torch.uint8, device="cuda:0")
    torch.cuda.synchronize()
    n_out, bad, code = C.c_uint64(0), C.c_int64(-1), C.c_int32(0)
    rc = _lib.lib().dh_bgzf_inflate_noalloc(0, d_comp.data_ptr(), d_comp.numel(), blocks, n, d_out.data_ptr(), d_out.numel(),
                                            C.byref(n_out), C.byref(bad), C.byref(code), stream)
    return rc, d_out[: n_out.value].cpu().numpy(), bad.value, code.value


def _images():
    rng = np.random.default_rng(11)
    acgt = bytes(rng.choice(np.frombuffer(b"ACGT", dtype=np.uint8), 900000))
    noise = rng.integers(0, 256, 150000, dtype=np.uint8).tobytes()
    text = b"".join(b"chr%d\t%d\tid%d\tA\tC\t%d\tPASS\tDP=%d\n" % (i % 22 + 1, i * 37, i, i % 60, i % 101) for i in range(40000))
    big = util.make_bgzf(acgt + noise + text + bytes(200000))                                   # dynamic, stored, long runs
    small = util.make_bgzf(text[:70000], strategy=zlib.Z_FIXED) + util.make_bgzf(b"x") + util.make_bgzf(b"")
    return big, small


def test_noalloc_matches_inflate_host_on_calls_in_a_row(ctx):
    big, small = _images()
    for img in (big, small, big, small):           # the scratch array keeps the last call's contents: shrink, grow, repeat
        ref, st = dh.bgzf_inflate(ctx, img)
        rc, out, bad, code = _noalloc(img)
        assert rc == 0, _lib.lib().dh_last_error()
        assert bad == -1 and code == 0
        assert out.size == st.out_bytes and np.array_equal(out, ref)


def test_noalloc_on_a_caller_stream(ctx):
    import torch

    big, _ = _images()
    ref, _ = dh.bgzf_inflate(ctx, big)
    s = torch.cuda.Stream(device="cuda:0")
    for _ in range(2):
        rc, out, bad, _ = _noalloc(big, stream=s.cuda_stream)
        assert rc == 0 and bad == -1 and np.array_equal(out, ref)


def test_noalloc_reports_the_same_bad_block(ctx):
    big, _ = _images()
    a, blocks, n = _walk(big)
    img = a.copy()
    k = n // 2
    img[blocks[k].coffset + 18 + blocks[k].csize // 2] ^= 0x5A         # damage one DEFLATE payload
    with pytest.raises(dh.DuckhtsError) as e:
        dh.bgzf_inflate(ctx, img)
    rc, _, bad, code = _noalloc(img)
    assert rc == -4 and bad == e.value.stats.bad_block == k and code != 0
    rc, out, bad, _ = _noalloc(big)                                      # and the next call is clean again
    ref, _ = dh.bgzf_inflate(ctx, big)
    assert rc == 0 and bad == -1 and np.array_equal(out, ref)


def test_noalloc_rejects_sizes_beyond_the_scratch(ctx):
    import torch

    lib = _lib.lib()
    blocks = (_lib.BgzfBlock * (MAX_BLOCKS + 1))()
    buf = torch.zeros(1 << 16, dtype=torch.uint8, device="cuda:0")
    n_out, bad, code = C.c_uint64(0), C.c_int64(-1), C.c_int32(0)
    rc = lib.dh_bgzf_inflate_noalloc(0, buf.data_ptr(), buf.numel(), blocks, MAX_BLOCKS + 1, buf.data_ptr(), buf.numel(),
                                     C.byref(n_out), C.byref(bad), C.byref(code), None)
    assert rc == -1 and b"DH_NOALLOC_MAX_BLOCKS" in lib.dh_last_error()
    # a block that leaves no readable pad behind it is refused too, before anything is launched
    a, blocks, n = _walk(util.make_bgzf(b"hello"))
    rc = lib.dh_bgzf_inflate_noalloc(0, buf.data_ptr(), a.size, blocks, n, buf.data_ptr(), buf.numel(),
                                     C.byref(n_out), C.byref(bad), C.byref(code), None)
    assert rc == -1 and b"readable bytes" in lib.dh_last_error()
