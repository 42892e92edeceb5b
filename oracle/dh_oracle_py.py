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
                ("rec_off", C.POINTER(C.c_uint64)), ("status", C.c_int), ("endpos", C.POINTER(C.c_int64))]


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
        _lib.dho_index_parse.restype = C.c_void_p
        _lib.dho_index_parse.argtypes = [C.c_void_p, C.c_uint64]
        _lib.dho_index_free.restype = None
        _lib.dho_index_free.argtypes = [C.c_void_p]
        _lib.dho_index_n_ref.restype = C.c_int32
        _lib.dho_index_n_ref.argtypes = [C.c_void_p]
        _lib.dho_itr_query_rows.restype = C.c_int64
        _lib.dho_itr_query_rows.argtypes = [C.c_void_p, C.c_uint64] + [C.c_void_p] * 5 + [C.c_int, C.c_int64, C.c_int64, C.c_void_p]
        _lib.dho_itr_regions_rows.restype = C.c_int64
        _lib.dho_itr_regions_rows.argtypes = [C.c_void_p, C.c_uint64] + [C.c_void_p] * 5 + [C.c_int] + [C.c_void_p] * 4
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


# the SAMtags columns of standard_tags := true (src/bam_reader.c:54-70), the tags the AUXILIARY_TAGS map then leaves out
STD_TAGS = ("AM AS BC BQ BZ CB CC CG CM CO CP CQ CR CS CT CY E2 FI FS FZ H0 H1 H2 HI IH LB MC MD MI ML MM MN MQ NH NM OA OC OP OQ OX PG "
            "PQ PT PU Q2 QT QX R2 RG RX SA SM TC TS U2 UQ").split()


def set_aux_exclude(tags=()):
    ex = b"".join(t if isinstance(t, bytes) else t.encode() for t in tags)
    lib().dho_set_aux_exclude(ex, len(ex) // 2)


def bam_decode(inflated, tags=()):
    """Reference-order columns of a whole inflated BAM stream.  tags: iterable of (b'NM', 'i') ..."""
    a = _u8(inflated)
    reqs = (TagReq * max(len(tags), 1))()
    for k, (tg, ty) in enumerate(tags):
        reqs[k].tag = tg if isinstance(tg, bytes) else tg.encode()
        reqs[k].type = ty[:1].encode()
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
           "mtid": arr(cols.mtid, np.int32), "rec_off": arr(cols.rec_off, np.uint64), "endpos": arr(cols.endpos, np.int64),
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
        elif ty == "M":
            # AUXILIARY_TAGS: key NUL value NUL ... -> [(key, value), ...]
            rows = []
            for x in strcol_list(_strcol(tc.s, n)):
                if x is None:
                    rows.append(None)
                else:
                    f = x.split(b"\0")[:-1]
                    rows.append(list(zip(f[0::2], f[1::2])))
            d["m"] = rows
        elif ty[0] == "B":
            d["b"] = [None if x is None else np.frombuffer(x, dtype=np.int64) for x in strcol_list(_strcol(tc.s, n))]
        else:
            d["s"] = _strcol(tc.s, n)
        out["tags"].append(d)
    lib().dho_bam_free(C.byref(cols))
    return out


# ---- A8: indexed scans --------------------------------------------------------------------------------------
HTS_POS_MAX = (1 << 63) - 1


def _parse_decimal(s):
    """hts_parse_decimal with HTS_PARSE_THOUSANDS_SEP (htslib/hts.c) -> (value, rest)."""
    i, n = 0, len(s)
    while i < n and s[i] in " \t\n\v\f\r":
        i += 1
    sign, v, digits, decimals, e = 1, 0, 0, 0, 0
    if i < n and s[i] in "+-":
        sign = -1 if s[i] == "-" else 1
        i += 1
    while i < n and (s[i].isdigit() or s[i] == ","):
        if s[i] != ",":
            v = v * 10 + int(s[i]); digits += 1
        i += 1
    if i < n and s[i] == ".":
        i += 1
        while i < n and s[i].isdigit():
            v = v * 10 + int(s[i]); digits += 1; decimals += 1; i += 1
    if i < n and s[i] in "eE":
        i += 1
        es = 1
        if i < n and s[i] in "+-":
            es = -1 if s[i] == "-" else 1
            i += 1
        while i < n and s[i].isdigit():
            e = e * 10 + int(s[i]); i += 1
        e *= es
    elif i < n and s[i] in "kK":
        e += 3; i += 1
    elif i < n and s[i] in "mM":
        e += 6; i += 1
    elif i < n and s[i] in "gG":
        e += 9; i += 1
    e -= decimals
    if e > 0:
        v *= 10 ** e
    elif e < 0:
        v //= 10 ** (-e)
    return sign * v, (s[i:] if digits else s)


def parse_region(s, names):
    """hts_parse_region (htslib/hts.c) without HTS_PARSE_LIST -> (tid, beg, end) or None."""
    def tid_of(x):
        return names.index(x) if x in names else -1
    if s.startswith("{"):
        close = s.find("}")
        if close < 0:
            return None
        name, rest = s[1:close], s[close + 1:]
        if not rest.startswith(":"):
            t = tid_of(s[1:len(s) - 1])      # no colon after the brace: the remainder minus one trailing character
            return (t, 0, HTS_POS_MAX) if t >= 0 else None
        coords = rest[1:]
    else:
        c = s.rfind(":")
        if c < 0:
            t = tid_of(s)
            return (t, 0, HTS_POS_MAX) if t >= 0 else None
        t = tid_of(s)
        if t >= 0:
            return None if tid_of(s[:c]) >= 0 else (t, 0, HTS_POS_MAX)
        name, coords = s[:c], s[c + 1:]
    t = tid_of(name)
    if t < 0:
        return None
    v, rest = _parse_decimal(coords)
    beg = v - 1
    if beg < 0:
        if beg != -1 and rest[:1] == "-" and coords != "":
            return None
        if rest[:1].isdigit() or rest == "" or rest[:1] == ",":
            return (t, 0, HTS_POS_MAX if beg == -1 else -(beg + 1))
        elif beg < -1:
            return None
    if rest == "":
        end = HTS_POS_MAX
    elif rest[0] == "-":
        end, rest2 = _parse_decimal(rest[1:])
        if rest2 != "" and rest2[0] != ",":
            return None
    else:
        return None
    if end == 0:
        end = HTS_POS_MAX
    return (t, beg, end) if beg < end else None


def compact_regions(regions, names):
    """hts_reglist_create + reg_compact (htslib/region.c:84-122,177-260) -> sorted [(tid, beg, end)] or None."""
    per = {}
    for r in regions:
        pr = parse_region(r, names)
        if pr is None:
            continue
        per.setdefault(pr[0], []).append([pr[1], pr[2]])
    out = []
    for tid in sorted(per):
        iv = sorted(per[tid])
        m = [iv[0]]
        for b, e in iv[1:]:
            if m[-1][1] < b:
                m.append([b, e])
            elif m[-1][1] < e:
                m[-1][1] = e
        out += [(tid, b, e) for b, e in m]
    return out or None


class Index:
    def __init__(self, index_bytes):
        a = _u8(index_bytes)
        if a.size >= 2 and a[0] == 0x1f and a[1] == 0x8b:      # .csi is BGZF-compressed
            rc, a, _ = bgzf_inflate(a)
            assert rc == 0
        self._keep = a
        self.h = lib().dho_index_parse(a.ctypes.data, a.size)
        if not self.h:
            raise ValueError("dho_index_parse failed")
        self.n_ref = lib().dho_index_n_ref(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            lib().dho_index_free(self.h)
            self.h = None


def record_voffsets(bam_bytes, rec_off):
    """Virtual offset (htslib/bgzf.h:261) of every record, in the normalised form (block, in-block offset < isize)."""
    blocks, _ = bgzf_walk(bam_bytes)
    co = np.array([b[0] for b in blocks], dtype=np.uint64)
    isz = np.array([b[2] for b in blocks], dtype=np.uint64)
    uoff = np.concatenate([[0], np.cumsum(isz)]).astype(np.uint64)
    nz = np.nonzero(isz)[0]
    rec_off = np.asarray(rec_off, dtype=np.uint64)
    k = np.clip(np.searchsorted(uoff[nz], rec_off, side="right") - 1, 0, max(len(nz) - 1, 0))
    blk = nz[k]
    return (co[blk] << np.uint64(16)) | (rec_off - uoff[blk]), blocks


def _iter_arrays(bam_bytes, dec):
    voff, blocks = record_voffsets(bam_bytes, dec["rec_off"])
    n = dec["n"]
    end_of_data = np.uint64((blocks[-1][0] + blocks[-1][1]) << 16)
    nxt = np.concatenate([voff[1:], [end_of_data]]).astype(np.uint64) if n else np.zeros(0, np.uint64)
    return (np.ascontiguousarray(voff), np.ascontiguousarray(nxt), np.ascontiguousarray(dec["tid"], dtype=np.int32),
            np.ascontiguousarray(dec["pos"] - 1, dtype=np.int64), np.ascontiguousarray(dec["endpos"], dtype=np.int64))


def contig_scan_rows(bam_bytes, index_bytes, dec, tid_lo=0, tid_hi=None):
    """Row indices (into the sequential decode `dec`) of the reference's index-driven scan without a region:
    one sam_itr_queryi(tid, 0, HTS_POS_MAX) per contig in order (src/bam_reader.c:689-716).  Empty contigs are
    skipped (the reference's early end there is SURVEY Q2, a documented divergence)."""
    ix = Index(index_bytes)
    voff, nxt, tid, beg, end = _iter_arrays(bam_bytes, dec)
    n = dec["n"]
    out = np.zeros(max(n, 1), np.uint64)
    rows = []
    hi = len(dec["ref_names"]) if tid_hi is None else tid_hi
    for t in range(tid_lo, hi):
        k = lib().dho_itr_query_rows(ix.h, n, voff.ctypes.data, nxt.ctypes.data, tid.ctypes.data, beg.ctypes.data, end.ctypes.data,
                                     t, 0, HTS_POS_MAX, out.ctypes.data)
        rows += [int(x) for x in out[:k]]
    return rows


def region_scan_rows(bam_bytes, index_bytes, dec, regions):
    """Row indices of read_bam(region := ...) = sam_itr_regarray + sam_itr_next (src/bam_reader.c:660-671,749-752).
    None = the reference fails at init ("No reads found for region(s)")."""
    names = [x.decode() for x in dec["ref_names"]]
    iv = compact_regions(regions, names)
    if iv is None:
        return None
    ix = Index(index_bytes)
    voff, nxt, tid, beg, end = _iter_arrays(bam_bytes, dec)
    n = dec["n"]
    out = np.zeros(max(n, 1), np.uint64)
    it = np.array([v[0] for v in iv], np.int32)
    ib = np.array([v[1] for v in iv], np.int64)
    ie = np.array([min(v[2], HTS_POS_MAX) for v in iv], np.int64)
    k = lib().dho_itr_regions_rows(ix.h, n, voff.ctypes.data, nxt.ctypes.data, tid.ctypes.data, beg.ctypes.data, end.ctypes.data,
                                   len(iv), it.ctypes.data, ib.ctypes.data, ie.ctypes.data, out.ctypes.data)
    return [int(x) for x in out[:k]]
