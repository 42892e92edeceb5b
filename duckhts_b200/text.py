"""ctypes mirror of the text path (dh_text_*): a bgzip'ed tab-separated file as batches of lines whose fields
the GPU has located.  Used by the tests; the product's consumer is read_tabix in host/duckhts_ext.c."""
import ctypes as C

import numpy as np

from ._lib import DuckhtsError, TextBatch, check, lib


def _lib():
    return lib()


class TextFile:
    def __init__(self, ctx, path):
        self.ctx = ctx
        self.h = _lib().dh_text_open_file(ctx.h, str(path).encode())
        if not self.h:
            raise DuckhtsError("dh_text_open_file failed: " + lib().dh_last_error().decode())

    def close(self):
        if self.h:
            _lib().dh_text_close(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def head(self, cap=1 << 20):
        buf = np.empty(cap, dtype=np.uint8)
        n = C.c_uint64(0)
        check(_lib().dh_text_head(self.h, buf.ctypes.data, cap, C.byref(n)), "dh_text_head")
        return buf[: n.value].tobytes()

    def scan(self, n_cols, proj=None, meta=ord("#")):
        """-> (lines, status): lines = [(flags, [field bytes or None for each projected column])]"""
        proj = list(range(n_cols)) if proj is None else list(proj)
        arr = (C.c_uint32 * max(1, len(proj)))(*proj)
        check(_lib().dh_text_configure(self.h, n_cols, arr, len(proj), meta), "dh_text_configure")
        check(_lib().dh_text_scan_begin(self.h), "dh_text_scan_begin")
        out, batches = [], 0
        b = TextBatch()
        while True:
            rc = _lib().dh_text_next_batch(self.h, C.byref(b))
            if rc < 0:
                raise DuckhtsError(f"dh_text_next_batch rc={rc}: {lib().dh_last_error().decode()}")
            if rc == 0:
                self.batches = batches
                return out, b.status
            batches += 1
            assert b.line0 == len(out) and b.n_proj == len(proj)
            n = b.n_lines
            text = np.ctypeslib.as_array(C.cast(b.text, C.POINTER(C.c_uint8)), shape=(b.text_bytes,)).tobytes()
            flags = np.ctypeslib.as_array(C.cast(b.flags, C.POINTER(C.c_uint8)), shape=(n,)).copy()
            cells = np.ctypeslib.as_array(C.cast(b.cells, C.POINTER(C.c_uint32)), shape=(n, len(proj), 2)).copy() if proj else np.zeros((n, 0, 2), np.uint32)
            for i in range(n):
                row = []
                for k in range(len(proj)):
                    lo, hi = int(cells[i, k, 0]), int(cells[i, k, 1])
                    row.append(None if lo == 0xFFFFFFFF else text[lo:hi])
                out.append((int(flags[i]), row))
