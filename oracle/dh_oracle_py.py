"""ctypes binding of oracle/libdh_oracle.so — TEST INFRASTRUCTURE ONLY.

May be imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs;
never by duckhts_b200/.  See oracle/dh_oracle.h for what each function restates.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATH = os.path.join(_HERE, "libdh_oracle.so")
_lib = None


class Block(C.Structure):
    _fields_ = [("coffset", C.c_uint64), ("csize", C.c_uint32), ("isize", C.c_uint32), ("crc", C.c_uint32),
                ("reserved", C.c_uint32)]


class StrCol(C.Structure):
    _fields_ = [("off", C.POINTER(C.c_uint64)), ("heap", C.POINTER(C.c_char)), ("valid", C.POINTER(C.c_uint8)),
                ("heap_len", C.c_uint64), ("heap_cap", C.c_uint64)]


class TagReq(C.Structure):
    _fields_ = [("tag", C.c_char * 2), ("type", C.c_char)]


class TagCol(C.Structure):
    _fields_ = [("i", C.POINTER(C.c_int64)), ("f", C.POINTER(C.c_double)), ("s", StrCol), ("valid", C.POINTER(C.c_uint8))]


class BamColumns(C.Structure):
    _fields_ = [("n", C.c_uint64), ("flag", C.POINTER(C.c_uint16)), ("pos", C.POINTER(C.c_int64)),
                ("mapq", C.POINTER(C.c_int32)), ("pnext", C.POINTER(C.c_int64)), ("tlen", C.POINTER(C.c_int64)),
                ("tid", C.POINTER(C.c_int32)), ("mtid", C.POINTER(C.c_int32)),
                ("qname", StrCol), ("rname", StrCol), ("cigar", StrCol), ("rnext", StrCol), ("seq", StrCol),
                ("qual", StrCol), ("rg", StrCol), ("sample", StrCol),
                ("n_tags", C.c_int), ("tags", C.POINTER(TagCol)),
                ("n_ref", C.c_int32), ("ref_name", C.POINTER(C.c_char_p)), ("ref_len", C.POINTER(C.c_uint32)),
                ("text", C.POINTER(C.c_char)), ("l_text", C.c_uint32), ("first_record", C.c_uint64),
                ("rec_off", C.POINTER(C.c_uint64)), ("status", C.c_int)]


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_PATH):
            subprocess.check_call(["make", "-C", _HERE, "port"], stdout=subprocess.DEVNULL)
        _lib = C.CDLL(_PATH)
        _lib.dho_bgzf_walk.restype = C.c_int64
        _lib.dho_bgzf_walk.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(Block), C.c_int64, C.POINTER(C.c_uint64)]
        _lib.dho_bgzf_inflate.restype = C.c_int
        _lib.dho_bgzf_inflate.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64),
                                          C.POINTER(C.c_int64)]
        _lib.dho_bgzf_inflate_block.restype = C.c_int
        _lib.dho_bgzf_inflate_block.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
        _lib.dho_bam_decode.restype = C.c_int
        _lib.dho_bam_decode.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(TagReq), C.c_int, C.POINTER(BamColumns)]
        _lib.dho_bam_free.restype = None
        _lib.dho_bam_free.argtypes = [C.POINTER(BamColumns)]
    return _lib


def _u8(data):
    if isinstance(data, np.ndarray):
        return np.ascontiguousarray(data, dtype=np.uint8)
    return np.frombuffer(bytes(data), dtype=np.uint8)


def bgzf_walk(data):
    a = _u8(data)
    cap = a.size // 26 + 1
    arr = (Block * cap)()
    end = C.c_uint64(0)
    n = lib().dho_bgzf_walk(a.ctypes.data, a.size, arr, cap, C.byref(end))
    if n < 0:
        return None, 0
    return [(arr[i].coffset, arr[i].csize, arr[i].isize, arr[i].crc) for i in range(n)], end.value


def bgzf_inflate(data, want_output=True):
    """-> (rc, np.uint8 array | None, bad_block).  rc: 0 ok, -1 header, -2 zlib, -3 crc."""
    a = _u8(data)
    n = C.c_uint64(0)
    bad = C.c_int64(-1)
    rc = lib().dho_bgzf_inflate(a.ctypes.data, a.size, None, 0, C.byref(n), C.byref(bad))
    if rc != 0 or not want_output:
        return rc, None, bad.value
    out = np.empty(n.value, dtype=np.uint8)
    rc = lib().dho_bgzf_inflate(a.ctypes.data, a.size, out.ctypes.data, out.size, C.byref(n), C.byref(bad))
    return rc, out, bad.value


def _strcol(sc, n):
    off = np.ctypeslib.as_array(sc.off, shape=(n + 1,)).copy() if n + 1 else np.zeros(1, np.uint64)
    heap = np.ctypeslib.as_array(C.cast(sc.heap, C.POINTER(C.c_uint8)), shape=(max(int(sc.heap_len), 1),))[: int(sc.heap_len)].copy()
    valid = np.ctypeslib.as_array(sc.valid, shape=(max(n, 1),))[:n].copy().astype(bool)
    return {"off": off, "heap": heap, "valid": valid}


def strcol_list(col):
    """[bytes | None] for a string column dict."""
    off, heap, valid = col["off"], col["heap"].tobytes(), col["valid"]
    return [heap[int(off[i]):int(off[i + 1])] if valid[i] else None for i in range(len(valid))]


def bam_decode(inflated, tags=()):
    """Reference-order columns of a whole inflated BAM stream.  tags: iterable of (b'NM', 'i') ..."""
    a = _u8(inflated)
    reqs = (TagReq * max(len(tags), 1))()
    for k, (tg, ty) in enumerate(tags):
        reqs[k].tag = tg if isinstance(tg, bytes) else tg.encode()
        reqs[k].type = ty.encode()
    cols = BamColumns()
    rc = lib().dho_bam_decode(a.ctypes.data, a.size, reqs, len(tags), C.byref(cols))
    if rc != 0:
        lib().dho_bam_free(C.byref(cols))
        raise ValueError("dho_bam_decode: invalid BAM header")
    n = int(cols.n)

    def arr(p, dt):
        return np.ctypeslib.as_array(p, shape=(max(n, 1),))[:n].copy().astype(dt, copy=False)

    out = {"n": n, "status": cols.status, "first_record": int(cols.first_record),
           "flag": arr(cols.flag, np.uint16), "pos": arr(cols.pos, np.int64), "mapq": arr(cols.mapq, np.int32),
           "pnext": arr(cols.pnext, np.int64), "tlen": arr(cols.tlen, np.int64), "tid": arr(cols.tid, np.int32),
           "mtid": arr(cols.mtid, np.int32), "rec_off": arr(cols.rec_off, np.uint64),
           "ref_names": [cols.ref_name[i] for i in range(cols.n_ref)],
           "ref_lens": [cols.ref_len[i] for i in range(cols.n_ref)],
           "text": C.string_at(cols.text, cols.l_text)}
    for name in ("qname", "rname", "cigar", "rnext", "seq", "qual", "rg", "sample"):
        out[name] = _strcol(getattr(cols, name), n)
    out["tags"] = []
    for k, (tg, ty) in enumerate(tags):
        tc = cols.tags[k]
        d = {"valid": np.ctypeslib.as_array(tc.valid, shape=(max(n, 1),))[:n].copy().astype(bool)}
        if ty == "i":
            d["i"] = np.ctypeslib.as_array(tc.i, shape=(max(n, 1),))[:n].copy()
        elif ty == "f":
            d["f"] = np.ctypeslib.as_array(tc.f, shape=(max(n, 1),))[:n].copy()
        else:
            d["s"] = _strcol(tc.s, n)
        out["tags"].append(d)
    lib().dho_bam_free(C.byref(cols))
    return out
