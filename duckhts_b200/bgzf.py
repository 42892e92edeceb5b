"""Host-side mirror of the BGZF part of the C-ABI (rows A1-A3 of SURVEY.md §8)."""
import ctypes as C

import numpy as np

from ._lib import BgzfBlock, DeflateStats, DuckhtsError, InflateStats, check, lib


class PinnedBuffer:
    """Page-locked host memory from dh_host_alloc, exposed as a numpy uint8 array."""

    def __init__(self, nbytes):
        self.nbytes = int(nbytes)
        self.ptr = lib().dh_host_alloc(max(self.nbytes, 1))
        if not self.ptr:
            raise DuckhtsError("dh_host_alloc failed: " + lib().dh_last_error().decode())
        self.array = np.ctypeslib.as_array(C.cast(self.ptr, C.POINTER(C.c_uint8)), shape=(max(self.nbytes, 1),))[: self.nbytes]

    def free(self):
        if self.ptr:
            self.array = None
            lib().dh_host_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def _as_u8(data):
    if isinstance(data, PinnedBuffer):
        return data.array
    if isinstance(data, np.ndarray):
        assert data.dtype == np.uint8 and data.flags["C_CONTIGUOUS"]
        return data
    return np.frombuffer(data, dtype=np.uint8)


def bgzf_walk(data, start=0, strict=True):
    """A1: list of (coffset, csize, isize, crc) for every BGZF block of `data` (htslib/bgzf.c:896-903,1147-1216)."""
    a = _as_u8(data)
    cap = a.size // 26 + 1
    arr = (BgzfBlock * cap)()
    end = C.c_uint64(0)
    n = lib().dh_bgzf_walk(a.ctypes.data, a.size, start, arr, cap, C.byref(end), 1 if strict else 0)
    check(n, "dh_bgzf_walk")
    return [(arr[i].coffset, arr[i].csize, arr[i].isize, arr[i].crc) for i in range(n)], end.value


class Context:
    """One per GPU (dh_ctx)."""

    def __init__(self, device=0):
        self.h = lib().dh_ctx_create(device)
        if not self.h:
            raise DuckhtsError("dh_ctx_create failed: " + lib().dh_last_error().decode())

    def close(self):
        if self.h:
            lib().dh_ctx_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def sm_count(self):
        return lib().dh_ctx_sm_count(self.h)

    @property
    def launches(self):
        return lib().dh_ctx_launch_count(self.h)

    def sync(self):
        check(lib().dh_ctx_sync(self.h), "dh_ctx_sync")

    def timer_start(self):
        check(lib().dh_timer_start(self.h), "dh_timer_start")

    def timer_stop(self):
        ms = C.c_float(0)
        check(lib().dh_timer_stop(self.h, C.byref(ms)), "dh_timer_stop")
        return ms.value

    def flush_l2(self):
        check(lib().dh_flush_l2(self.h), "dh_flush_l2")

    def measure_pcie(self, nbytes=1 << 30):
        a, b = C.c_float(0), C.c_float(0)
        check(lib().dh_measure_pcie(self.h, nbytes, C.byref(a), C.byref(b)), "dh_measure_pcie")
        return a.value, b.value

    def measure_pcie_bidir(self, nbytes=1 << 30):
        a, b = C.c_float(0), C.c_float(0)
        check(lib().dh_measure_pcie_bidir(self.h, nbytes, C.byref(a), C.byref(b)), "dh_measure_pcie_bidir")
        return a.value, b.value


def bgzf_inflate(ctx, data, out=None, want_output=True):
    """A2/A3: inflate a whole BGZF image host->host (src/bgzip.c:154-200 equivalent).  Returns (bytes-like, stats)."""
    a = _as_u8(data)
    st = InflateStats()
    n = C.c_uint64(0)
    # first pass for the size only when the caller gave no buffer: framing is cheap
    if want_output and out is None:
        blocks, _ = bgzf_walk(a)
        total = sum(b[2] for b in blocks)
        out = np.empty(total, dtype=np.uint8)
    o = _as_u8(out) if out is not None else None
    rc = lib().dh_bgzf_inflate_host(ctx.h, a.ctypes.data, a.size, o.ctypes.data if o is not None and o.size else None,
                                    o.size if o is not None else 0, C.byref(n), C.byref(st))
    if rc < 0:
        err = DuckhtsError(f"dh_bgzf_inflate_host rc={rc}: {lib().dh_last_error().decode()}")
        err.rc, err.stats = rc, st
        raise err
    return (o[: n.value] if o is not None else None), st


def bgzf_deflate(ctx, data, level=-1, out=None):
    """N1: compress bytes into a BGZF stream host->host (bgzip(), src/bgzip.c:209-273).  Returns (uint8 array, stats)."""
    a = _as_u8(data)
    cap = lib().dh_bgzf_deflate_bound(a.size)
    o = _as_u8(out) if out is not None else np.empty(cap, dtype=np.uint8)
    st = DeflateStats()
    n = C.c_uint64(0)
    rc = lib().dh_bgzf_deflate_host(ctx.h, a.ctypes.data if a.size else None, a.size, int(level), o.ctypes.data, o.size, C.byref(n), C.byref(st))
    if rc < 0:
        raise DuckhtsError(f"dh_bgzf_deflate_host rc={rc}: {lib().dh_last_error().decode()}")
    return o[: n.value], st


def bgzf_deflate_fd(ctx, fd_in, fd_out, level=-1):
    st = DeflateStats()
    rc = lib().dh_bgzf_deflate_fd(ctx.h, int(fd_in), int(fd_out), int(level), C.byref(st))
    if rc < 0:
        raise DuckhtsError(f"dh_bgzf_deflate_fd rc={rc}: {lib().dh_last_error().decode()}")
    return st


class BgzfDevice:
    """Device-resident BGZF image (dh_bgzf_dev_*): upload once, inflate repeatedly; used for kernel timing."""

    def __init__(self, ctx, data):
        a = _as_u8(data)
        self.ctx = ctx
        self.h = lib().dh_bgzf_dev_load(ctx.h, a.ctypes.data, a.size)
        if not self.h:
            raise DuckhtsError("dh_bgzf_dev_load failed: " + lib().dh_last_error().decode())

    nblocks = property(lambda s: lib().dh_bgzf_dev_nblocks(s.h))
    comp_bytes = property(lambda s: lib().dh_bgzf_dev_comp_bytes(s.h))
    out_bytes = property(lambda s: lib().dh_bgzf_dev_out_bytes(s.h))

    def inflate(self):
        check(lib().dh_bgzf_dev_inflate(self.h), "dh_bgzf_dev_inflate")

    def inflate_timed(self):
        """One pass with CUDA events between the kernels -> (phase1_ms, phase2_ms, fallback_ms)."""
        ms = (C.c_float * 3)()
        check(lib().dh_bgzf_dev_inflate_timed(self.h, ms), "dh_bgzf_dev_inflate_timed")
        return ms[0], ms[1], ms[2]

    def check(self):
        bb, bc = C.c_int64(-1), C.c_int32(0)
        rc = lib().dh_bgzf_dev_check(self.h, C.byref(bb), C.byref(bc))
        return rc, bb.value, bc.value

    def read(self, offset=0, n=None):
        n = self.out_bytes - offset if n is None else n
        out = np.empty(n, dtype=np.uint8)
        check(lib().dh_bgzf_dev_read(self.h, offset, out.ctypes.data, n), "dh_bgzf_dev_read")
        return out

    def free(self):
        if self.h:
            lib().dh_bgzf_dev_free(self.h)
            self.h = None
