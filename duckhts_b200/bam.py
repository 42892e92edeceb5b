"""Host-side mirror of the read_bam part of the C-ABI (rows A4-A15 of SURVEY.md §8; reference src/bam_reader.c).

Column names, order and NULL rules are the reference's (bam_reader.c:189-204, 509-549)."""
import ctypes as C

import numpy as np

from ._lib import DH_BAM_CORE_COLS, BamBatch, BamScanStats, DuckhtsError, TagReq, check, lib
from .bgzf import _as_u8

VOFF_MAX = 0xFFFFFFFFFFFFFFFF
CORE_NAMES = ("qname", "flag", "rname", "pos", "mapq", "cigar", "rnext", "pnext", "tlen", "seq", "qual", "rg", "sample")
CORE_DTYPES = {"flag": np.uint16, "pos": np.int64, "mapq": np.int32, "pnext": np.int64, "tlen": np.int64}
# the reference's standard_tags table (bam_reader.c:54-70): tag -> column type
STD_TAGS = {"AM": "i", "AS": "i", "BC": "Z", "BQ": "Z", "BZ": "Z", "CB": "Z", "CC": "Z", "CM": "i", "CO": "Z", "CP": "i",
            "CQ": "Z", "CR": "Z", "CS": "Z", "CT": "Z", "CY": "Z", "E2": "Z", "FI": "i", "FS": "Z", "H0": "i", "H1": "i",
            "H2": "i", "HI": "i", "IH": "i", "LB": "Z", "MC": "Z", "MD": "Z", "MI": "Z", "MM": "Z", "MN": "i", "MQ": "i",
            "NH": "i", "NM": "i", "OA": "Z", "OC": "Z", "OP": "i", "OQ": "Z", "OX": "Z", "PG": "Z", "PQ": "i", "PT": "Z",
            "PU": "Z", "Q2": "Z", "QT": "Z", "QX": "Z", "R2": "Z", "RG": "Z", "RX": "Z", "SA": "Z", "SM": "i", "TC": "i",
            "TS": "A", "U2": "Z", "UQ": "i"}


def cells_to_list(ptr, n, validity=None):
    """duckdb_string_t[n] at host address ptr -> [bytes | None]; follows the pointers exactly as DuckDB would."""
    if n == 0:
        return []
    raw = C.string_at(ptr, 16 * n)
    a = np.frombuffer(raw, dtype=np.uint8).reshape(n, 16)
    lens = a[:, :4].copy().view(np.uint32).ravel()
    ptrs = a[:, 8:].copy().view(np.uint64).ravel()
    out = [None] * n
    for i in range(n):
        if validity is not None and not validity[i]:
            continue
        ln = int(lens[i])
        out[i] = raw[16 * i + 4:16 * i + 4 + ln] if ln <= 12 else C.string_at(int(ptrs[i]), ln)
    return out


def validity_to_bool(ptr, n):
    if not ptr:
        return None
    words = np.frombuffer(C.string_at(ptr, ((n + 63) // 64) * 8), dtype=np.uint64)
    return np.unpackbits(words.view(np.uint8), bitorder="little")[:n].astype(bool)


DH_ENOREGION = -7


class Index:
    """A BAI or CSI index image (dh_index; reference hts_idx_load3).  CSI is BGZF: the GPU inflates it."""

    def __init__(self, ctx, data):
        """ctx may be None for a BAI (plain bytes, parsed on the host); a CSI needs a context to be inflated."""
        self.a = _as_u8(data)
        self.h = lib().dh_index_load(ctx.h if ctx is not None else None, self.a.ctypes.data, self.a.size)
        if not self.h:
            raise DuckhtsError("dh_index_load failed: " + lib().dh_last_error().decode())

    @property
    def n_ref(self):
        return lib().dh_index_n_ref(self.h)

    @property
    def is_csi(self):
        return bool(lib().dh_index_is_csi(self.h))

    def split_contigs(self, n_parts):
        """dh_index_split_contigs: n_parts + 1 contig bounds of about equal compressed size (host code, no GPU)."""
        v = (C.c_int32 * (n_parts + 1))()
        check(lib().dh_index_split_contigs(self.h, n_parts, v), "dh_index_split_contigs")
        return list(v)

    def close(self):
        if self.h:
            lib().dh_index_free(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


class BamFile:
    """One BAM image bound to a Context (dh_bam)."""

    def __init__(self, ctx, data):
        self.a = _as_u8(data)  # keep alive
        self.ctx = ctx
        self.h = lib().dh_bam_open(ctx.h, self.a.ctypes.data, self.a.size)
        if not self.h:
            raise DuckhtsError("dh_bam_open failed: " + lib().dh_last_error().decode())
        self.tags = []

    @classmethod
    def open_file(cls, ctx, path):
        """dh_bam_open_file: the file is streamed through pinned staging buffers instead of being held as one image."""
        o = cls.__new__(cls)
        o.a, o.ctx, o.tags = None, ctx, []
        o.h = lib().dh_bam_open_file(ctx.h, path.encode() if isinstance(path, str) else path)
        if not o.h:
            raise DuckhtsError("dh_bam_open_file failed: " + lib().dh_last_error().decode())
        return o

    def close(self):
        if self.h:
            lib().dh_bam_close(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def ref_names(self):
        return [lib().dh_bam_ref_name(self.h, i) for i in range(lib().dh_bam_n_ref(self.h))]

    @property
    def ref_lens(self):
        return [lib().dh_bam_ref_len(self.h, i) for i in range(lib().dh_bam_n_ref(self.h))]

    @property
    def header_text(self):
        n = C.c_uint32(0)
        p = lib().dh_bam_header_text(self.h, C.byref(n))
        return C.string_at(p, n.value)

    @property
    def first_voffset(self):
        return lib().dh_bam_first_voffset(self.h)

    def set_projection(self, col_mask=None, tags=()):
        self.tags = [(t if isinstance(t, bytes) else t.encode(), ty) for t, ty in tags]
        reqs = (TagReq * max(len(self.tags), 1))()
        for k, (tg, ty) in enumerate(self.tags):
            reqs[k].tag = tg
            reqs[k].type = ty[:1].encode()          # "Bi" / "Bf": LIST column with BIGINT / DOUBLE children
            reqs[k].pad = ty[1:2].encode() if len(ty) > 1 else b"\0"
        if col_mask is None:
            col_mask = (1 << (DH_BAM_CORE_COLS + len(self.tags))) - 1
        self.col_mask = col_mask
        check(lib().dh_bam_set_projection(self.h, col_mask & 0xFFFFFFFFFFFFFFFF, reqs, len(self.tags)), "dh_bam_set_projection")

    def set_aux_map(self, enable=True, exclude=()):
        """AUXILIARY_TAGS MAP column; exclude = two-byte tags left out (the typed standard_tags columns)."""
        ex = b"".join(t if isinstance(t, bytes) else t.encode() for t in exclude)
        self.aux_map = bool(enable)
        check(lib().dh_bam_set_aux_map(self.h, 1 if enable else 0, ex, len(ex) // 2), "dh_bam_set_aux_map")

    def configure(self, seg_comp=0, batch_rows=0, carry=0):
        check(lib().dh_bam_configure(self.h, seg_comp, batch_rows, carry), "dh_bam_configure")

    def set_mode(self, device_input=False, device_output=False, compact=False):
        """compact: batches in the compact wire layout (DH_BAM_COMPACT_OUTPUT) instead of DuckDB vector layout."""
        check(lib().dh_bam_set_mode(self.h, (1 if device_input else 0) | (2 if device_output else 0) | (4 if compact else 0)), "dh_bam_set_mode")

    def set_filter(self, flag_require=0, flag_exclude=0, min_mapq=0):
        check(lib().dh_bam_set_filter(self.h, flag_require, flag_exclude, min_mapq), "dh_bam_set_filter")

    def set_slots(self, n):
        check(lib().dh_bam_set_slots(self.h, n), "dh_bam_set_slots")

    def prefetch(self):
        check(lib().dh_bam_prefetch(self.h), "dh_bam_prefetch")

    def clone(self, ctx):
        """A second handle on the same image for another context (another GPU): dh_bam_clone."""
        o = BamFile.__new__(BamFile)
        o.a, o.ctx, o.tags = self.a, ctx, list(self.tags)
        o.h = lib().dh_bam_clone(self.h, ctx.h)
        if not o.h:
            raise DuckhtsError("dh_bam_clone failed: " + lib().dh_last_error().decode())
        if hasattr(self, "col_mask"):
            o.col_mask = self.col_mask
        return o

    @property
    def rg(self):
        """[(id bytes, sample bytes | None)] in header order: the dictionary SAMPLE_ID codes of the compact layout index."""
        out = []
        for i in range(lib().dh_bam_n_rg(self.h)):
            n = C.c_uint32(0)
            p = lib().dh_bam_rg_id(self.h, i, C.byref(n))
            out.append((C.string_at(p, n.value), lib().dh_bam_rg_sample(self.h, i)))
        return out

    def index_build(self):
        """dh_bam_index_build: the BAI image of this (coordinate-sorted) file, built from one GPU scan -> bytes."""
        p, n = C.c_void_p(), C.c_uint64(0)
        check(lib().dh_bam_index_build(self.h, C.byref(p), C.byref(n)), "dh_bam_index_build")
        try:
            return C.string_at(p.value, n.value)
        finally:
            lib().dh_free(p)

    def acquire(self):
        """dh_bam_acquire_batch: the batch stays valid until release(batch.slot)."""
        b = BamBatch()
        rc = check(lib().dh_bam_acquire_batch(self.h, C.byref(b)), "dh_bam_acquire_batch")
        return b if rc == 1 else None

    def release(self, slot):
        check(lib().dh_bam_release_batch(self.h, slot), "dh_bam_release_batch")

    def checksum(self, batch, sums=None):
        """Adds the per-column checksums of a held batch to `sums` (numpy uint64[DH_BAM_MAX_COLS])."""
        if sums is None:
            sums = np.zeros(DH_BAM_CORE_COLS + 63, dtype=np.uint64)
        check(lib().dh_bam_batch_checksum(self.h, C.byref(batch), sums.ctypes.data_as(C.POINTER(C.c_uint64))), "dh_bam_batch_checksum")
        return sums

    def drain_checksums(self):
        """Run the scan to the end; returns (rows, per-column checksums)."""
        sums = np.zeros(DH_BAM_CORE_COLS + 63, dtype=np.uint64)
        rows = 0
        while True:
            b = self.next_batch()
            if b is None:
                return rows, sums
            rows += b.n_rows
            self.checksum(b, sums)

    def split(self, n_parts):
        v = (C.c_uint64 * (n_parts + 1))()
        check(lib().dh_bam_split(self.h, n_parts, v), "dh_bam_split")
        return list(v)

    def scan_begin(self, voff_beg=0, voff_end=VOFF_MAX):
        check(lib().dh_bam_scan_begin(self.h, voff_beg, voff_end), "dh_bam_scan_begin")

    def scan_regions(self, index, regions):
        """read_bam(region := ...): `regions` = the parameter split at commas (src/bam_reader.c:318-348).
        Returns False when no region names a known contig (the reference's "No reads found for region(s)")."""
        arr = (C.c_char_p * max(len(regions), 1))(*[r.encode() if isinstance(r, str) else r for r in regions])
        rc = lib().dh_bam_scan_regions(self.h, index.h, arr, len(regions))
        if rc == DH_ENOREGION:
            return False
        check(rc, "dh_bam_scan_regions")
        return True

    def scan_contigs(self, index, tid_lo=0, tid_hi=1 << 30):
        check(lib().dh_bam_scan_contigs(self.h, index.h, tid_lo, min(tid_hi, 1 << 30)), "dh_bam_scan_contigs")

    def split_contigs(self, index, n_parts):
        v = (C.c_int32 * (n_parts + 1))()
        check(lib().dh_bam_split_contigs(self.h, index.h, n_parts, v), "dh_bam_split_contigs")
        return list(v)

    def next_batch(self):
        b = BamBatch()
        rc = check(lib().dh_bam_next_batch(self.h, C.byref(b)), "dh_bam_next_batch")
        return b if rc == 1 else None

    def stats(self):
        st = BamScanStats()
        check(lib().dh_bam_scan_stats_get(self.h, C.byref(st)), "dh_bam_scan_stats_get")
        return st

    def drain(self):
        """Run the scan to the end without looking at the rows (timing)."""
        rows = 0
        while True:
            b = self.next_batch()
            if b is None:
                return rows
            rows += b.n_rows


COMPACT = 4
_COMPACT_FIX = {"flag": np.uint16, "pos": np.int32, "mapq": np.uint8, "pnext": np.int32, "tlen": np.int32}


def _compact_batch_to_columns(bf, b):
    """A batch in the compact wire layout (include/duckhts_cuda.h, DH_BAM_COMPACT_OUTPUT) -> the same dict as the DuckDB
    layout gives: the expansion a consumer does (widening, dictionary lookup, one string per length word)."""
    n = int(b.n_rows)
    out = {"n": n}
    heap = C.string_at(b.heap, int(b.heap_bytes)) if b.heap_bytes else b""
    rec_off = np.frombuffer(C.string_at(b.rec_off, 4 * n), dtype=np.uint32) if n else np.zeros(0, np.uint32)
    names, rgs = bf.ref_names, bf.rg

    def arr(c, dt):
        dt = np.dtype(dt)
        return np.frombuffer(C.string_at(b.data[c], n * dt.itemsize), dtype=dt).copy() if n else np.zeros(0, dt)

    on = lambda c: c >= 64 or (bf.col_mask >> c) & 1  # noqa: E731
    for c, name in enumerate(CORE_NAMES):
        if not on(c):
            continue
        if name in _COMPACT_FIX:
            v = arr(c, _COMPACT_FIX[name])
            out[name] = (v.astype(np.int64) + 1) if name in ("pos", "pnext") else v.astype(CORE_DTYPES[name])
        elif name in ("rname", "rnext"):
            out[name] = [names[t] if 0 <= t < len(names) else b"*" for t in arr(c, np.int32)]
        elif name == "sample":
            valid = validity_to_bool(b.validity[c], n)
            out[name] = [rgs[k][1] if valid[i] else None for i, k in enumerate(arr(c, np.int32))]
    # string slots of every row, in slot order
    slot_cols = [(c, CORE_NAMES[c], "q" if c == 0 else "u" if c == 10 else "s") for c in (0, 5, 9, 10) if on(c)]
    for k, (tg, ty) in enumerate(bf.tags):
        c = DH_BAM_CORE_COLS + k
        if on(c) and (ty == "Z" or ty[0] == "B"):
            slot_cols.append((c, k, "b" if ty[0] == "B" else "s"))
    if on(11):
        slot_cols.append((11, "rg", "s"))
    lens = {c: (arr(c, np.uint32) if kind != "b" else
                (np.frombuffer(C.string_at(b.data[c], 16 * n), dtype=np.uint64).reshape(n, 2) if n else np.zeros((0, 2), np.uint64)))
            for c, _, kind in slot_cols}
    valids = {c: validity_to_bool(b.validity[c], n) for c, _, _ in slot_cols}
    vals = {c: [None] * n for c, _, _ in slot_cols}
    for r in range(n):
        off = int(rec_off[r])
        for c, _, kind in slot_cols:
            if valids[c] is not None and not valids[c][r]:
                continue
            if kind == "b":
                o8, ln = int(lens[c][r, 0]), int(lens[c][r, 1])
                vals[c][r] = np.frombuffer(heap[8 * o8:8 * o8 + 8 * ln], dtype=np.int64).copy()
                off += 4 + 8 * ln
                continue
            v = int(lens[c][r])
            cap, flagged = v & 0x7fffffff, v >> 31
            raw = heap[off:off + cap]
            if kind == "q":
                raw = raw.split(b"\0")[0] if flagged else raw[:cap - 1]
            elif flagged:
                raw = raw.split(b"\0")[0]
            vals[c][r] = raw
            off += (cap + 3) & ~3
    for c, key, kind in slot_cols:
        if c < DH_BAM_CORE_COLS:
            out[CORE_NAMES[c]] = vals[c]
    tags = []
    for k, (tg, ty) in enumerate(bf.tags):
        c = DH_BAM_CORE_COLS + k
        if not on(c):
            tags.append(None)
            continue
        valid = validity_to_bool(b.validity[c], n)
        d = {"valid": valid if valid is not None else np.zeros(0, bool)}
        if ty in ("i", "c"):        # "c": a CIGAR metric as a derived BIGINT column (dh_bam_set_projection, N3)
            d["i"] = arr(c, np.int64)
        elif ty == "f":
            d["f"] = arr(c, np.float64)
        elif ty[0] == "B":
            d["b"] = vals[c]
        elif ty == "A":
            ch = arr(c, np.uint8)
            d["s"] = [(bytes([x]) if x else b"") if valid[i] else None for i, x in enumerate(ch)]
        else:
            d["s"] = vals[c]
        tags.append(d)
    out["tags"] = tags
    return out


def batch_to_columns(bf, b):
    """One dh_bam_batch -> dict of numpy arrays / lists of bytes (None = NULL)."""
    if b.layout == COMPACT:
        return _compact_batch_to_columns(bf, b)
    n = int(b.n_rows)
    out = {"n": n}
    for c, name in enumerate(CORE_NAMES):
        if not (bf.col_mask >> c) & 1:
            continue
        if name in CORE_DTYPES:
            dt = np.dtype(CORE_DTYPES[name])
            out[name] = np.frombuffer(C.string_at(b.data[c], n * dt.itemsize), dtype=dt).copy() if n else np.zeros(0, dt)
        else:
            out[name] = cells_to_list(b.data[c], n, validity_to_bool(b.validity[c], n))
    tags = []
    for k, (tg, ty) in enumerate(bf.tags):
        c = DH_BAM_CORE_COLS + k
        if c < 64 and not (bf.col_mask >> c) & 1:
            tags.append(None)
            continue
        valid = validity_to_bool(b.validity[c], n)
        d = {"valid": valid if valid is not None else np.zeros(0, bool)}
        if ty in ("i", "c"):        # "c": a CIGAR metric as a derived BIGINT column (dh_bam_set_projection, N3)
            d["i"] = np.frombuffer(C.string_at(b.data[c], n * 8), dtype=np.int64).copy() if n else np.zeros(0, np.int64)
        elif ty == "f":
            d["f"] = np.frombuffer(C.string_at(b.data[c], n * 8), dtype=np.float64).copy() if n else np.zeros(0, np.float64)
        elif ty[0] == "B":
            # duckdb_list_entry cells; children = 8-byte words at heap + 8 * offset (raw bits: int64, or double bits)
            ent = np.frombuffer(C.string_at(b.data[c], n * 16), dtype=np.uint64).reshape(n, 2) if n else np.zeros((0, 2), np.uint64)
            heap = np.frombuffer(C.string_at(b.heap, int(b.heap_bytes)), dtype=np.uint8) if b.heap_bytes else np.zeros(0, np.uint8)
            lists = []
            for r in range(n):
                if valid is not None and not valid[r]:
                    lists.append(None)
                    continue
                o, ln = int(ent[r, 0]) * 8, int(ent[r, 1])
                lists.append(heap[o:o + 8 * ln].view(np.int64).copy())
            d["b"] = lists
        else:
            d["s"] = cells_to_list(b.data[c], n, valid)
        tags.append(d)
    out["tags"] = tags
    if b.aux_map:
        # MAP(VARCHAR, VARCHAR): per row None or [(key, value), ...]
        na = int(b.n_aux)
        keys, vals = cells_to_list(b.aux_keys, na), cells_to_list(b.aux_values, na)
        ent = np.frombuffer(C.string_at(b.aux_map, n * 16), dtype=np.uint64).reshape(n, 2) if n else np.zeros((0, 2), np.uint64)
        valid = validity_to_bool(b.aux_map_validity, n)
        out["aux_map"] = [None if not valid[r] else [(keys[i], vals[i]) for i in range(int(ent[r, 0]), int(ent[r, 0] + ent[r, 1]))]
                          for r in range(n)]
    return out


def bam_scan(ctx, data, tags=(), col_mask=None, voff=(0, VOFF_MAX), seg_comp=0, batch_rows=0, carry=0, index=None,
             regions=None, contigs=None, aux_map=None, compact=False, filt=None):
    """Whole scan -> one dict of columns in row order (test/inspection helper; the product API is batch-wise).
    index = index file bytes; with it, regions = list of region strings, or contigs = (tid_lo, tid_hi) / True."""
    with (BamFile.open_file(ctx, data) if isinstance(data, str) else BamFile(ctx, data)) as bf:
        bf.set_projection(col_mask, tags)
        if aux_map is not None:          # aux_map = iterable of excluded tags (may be empty)
            bf.set_aux_map(True, aux_map)
        bf.configure(seg_comp, batch_rows, carry)
        if compact:
            bf.set_mode(compact=True)
        if filt:
            bf.set_filter(*filt)
        if index is not None:
            with Index(ctx, index) as ix:
                if regions is not None:
                    if not bf.scan_regions(ix, regions):
                        return None
                else:
                    bf.scan_contigs(ix, *(contigs if isinstance(contigs, tuple) else ()))
        else:
            bf.scan_begin(*voff)
        parts = []
        status = 0
        while True:
            b = bf.next_batch()
            if b is None:
                break
            parts.append(batch_to_columns(bf, b))
            status = b.status
        st = bf.stats()
        res = {"n": sum(p["n"] for p in parts), "status": st.status, "ref_names": bf.ref_names, "ref_lens": bf.ref_lens,
               "text": bf.header_text, "n_batches": len(parts), "stats": st, "last_batch_status": status}
        names = [nm for c, nm in enumerate(CORE_NAMES) if (bf.col_mask >> c) & 1]
        for nm in names:
            if nm in CORE_DTYPES:
                res[nm] = np.concatenate([p[nm] for p in parts]) if parts else np.zeros(0, CORE_DTYPES[nm])
            else:
                res[nm] = [x for p in parts for x in p[nm]]
        if aux_map is not None:
            res["aux_map"] = [x for p in parts for x in p["aux_map"]]
        res["tags"] = []
        for k, (tg, ty) in enumerate(bf.tags):
            if DH_BAM_CORE_COLS + k < 64 and not (bf.col_mask >> (DH_BAM_CORE_COLS + k)) & 1:
                res["tags"].append(None)
                continue
            d = {"valid": np.concatenate([p["tags"][k]["valid"] for p in parts]) if parts else np.zeros(0, bool)}
            for key in ("i", "f"):
                if parts and key in parts[0]["tags"][k]:
                    d[key] = np.concatenate([p["tags"][k][key] for p in parts])
                elif not parts and (ty == key or (ty == "c" and key == "i")):
                    d[key] = np.zeros(0, np.int64 if key == "i" else np.float64)
            if ty[0] == "B":
                d["b"] = [x for p in parts for x in p["tags"][k]["b"]]
            elif ty not in ("i", "f", "c"):
                d["s"] = [x for p in parts for x in p["tags"][k]["s"]]
            res["tags"].append(d)
        return res
