"""Host-side mirror of the read_bcf part of the C-ABI (rows B1-B10 of SURVEY.md §8; reference src/bcf_reader.c).

Column names, order, types and NULL rules are the reference's (bcf_reader.c:552-758).  `rows_text` renders a scan in
the text form harness/fake_duckdb prints for the reference extension, so the two can be compared cell for cell."""
import ctypes as C

import numpy as np

from ._lib import BamScanStats, BcfBatch, BcfField, DuckhtsError, check, lib

DH_ENOREGION = -7
from .bam import cells_to_list, validity_to_bool
from .bgzf import _as_u8

CORE_NAMES = ("CHROM", "POS", "ID", "REF", "ALT", "QUAL", "FILTER")
TYPE_NAMES = ("BOOLEAN", "INTEGER", "FLOAT", "VARCHAR")


class BcfFile:
    """One BCF image bound to a Context (dh_bcf)."""

    def __init__(self, ctx, data):
        self.a = _as_u8(data)
        self.ctx = ctx
        self.h = lib().dh_bcf_open(ctx.h, self.a.ctypes.data, self.a.size)
        if not self.h:
            raise DuckhtsError("dh_bcf_open failed: " + lib().dh_last_error().decode())
        L = lib()
        self.samples = [L.dh_bcf_sample_name(self.h, i) for i in range(L.dh_bcf_n_samples(self.h))]
        self.contigs = [L.dh_bcf_contig_name(self.h, i) for i in range(L.dh_bcf_n_contigs(self.h))]
        self.info, self.format = [], []
        for n, get, dst in ((L.dh_bcf_n_info(self.h), L.dh_bcf_info_field, self.info), (L.dh_bcf_n_format(self.h), L.dh_bcf_format_field, self.format)):
            for i in range(n):
                f = BcfField()
                check(get(self.h, i, C.byref(f)), "dh_bcf_*_field")
                dst.append((f.name.decode(), int(f.type), bool(f.is_list)))
        self.core_mask = 0x7f
        self.info_want = [True] * len(self.info)
        self.format_want = [True] * len(self.format)

    def close(self):
        if self.h:
            lib().dh_bcf_close(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def has_vep(self):
        return bool(lib().dh_bcf_has_vep(self.h))

    def set_projection(self, core_mask=0x7f, info_want=None, format_want=None):
        self.core_mask = core_mask
        self.info_want = [True] * len(self.info) if info_want is None else [bool(x) for x in info_want]
        self.format_want = [True] * len(self.format) if format_want is None else [bool(x) for x in format_want]
        iw = np.array(self.info_want + [False], dtype=np.uint8)
        fw = np.array(self.format_want + [False], dtype=np.uint8)
        check(lib().dh_bcf_set_projection(self.h, core_mask, iw.ctypes.data, fw.ctypes.data), "dh_bcf_set_projection")

    def configure(self, seg_comp=0, batch_records=0, string_bytes=0):
        check(lib().dh_bcf_configure(self.h, seg_comp, batch_records, string_bytes), "dh_bcf_configure")

    def set_mode(self, device_input=False, device_output=False, compact=False):
        check(lib().dh_bcf_set_mode(self.h, (1 if device_input else 0) | (2 if device_output else 0) | (4 if compact else 0)), "dh_bcf_set_mode")

    def scan_begin(self):
        check(lib().dh_bcf_scan_begin(self.h), "dh_bcf_scan_begin")

    def scan_region(self, index, region):
        """One region through the CSI index (bcf_itr_querys); False when the region names no contig of the file."""
        rc = lib().dh_bcf_scan_region(self.h, index.h, region if isinstance(region, bytes) else region.encode())
        if rc == DH_ENOREGION:
            return False
        check(rc, "dh_bcf_scan_region")
        return True

    def next_batch(self):
        b = BcfBatch()
        rc = check(lib().dh_bcf_next_batch(self.h, C.byref(b)), "dh_bcf_next_batch")
        return b if rc == 1 else None

    def stats(self):
        st = BamScanStats()
        check(lib().dh_bcf_scan_stats_get(self.h, C.byref(st)), "dh_bcf_scan_stats_get")
        return st

    def drain(self):
        n = 0
        while True:
            b = self.next_batch()
            if b is None:
                return n
            n += b.n_records


class _NoIndex:
    def __enter__(self):
        return None

    def __exit__(self, *a):
        return False


def _esc(b):
    out = []
    for c in b:
        if c == 9:
            out.append("\\t")
        elif c == 10:
            out.append("\\n")
        elif c == 92:
            out.append("\\\\")
        elif c < 32 or c == 127:
            out.append("\\x%02x" % c)
        else:
            out.append(chr(c))
    return "".join(out)


def _column_text(ptr, vptr, n, typ, is_list, heap):
    """n cells of one column -> list of strings in fake_duckdb's text form."""
    valid = validity_to_bool(vptr, n) if vptr else None
    if n == 0:
        return []
    if is_list:
        ent = np.frombuffer(C.string_at(ptr, 16 * n), dtype=np.uint64).reshape(n, 2)
        out = []
        for i in range(n):
            if valid is not None and not valid[i]:
                out.append("\\N")
                continue
            off, ln = int(ent[i, 0]), int(ent[i, 1])
            if typ == 3:
                kids = [_esc(x) for x in cells_to_list(heap + 16 * off, ln)]
            elif typ == 1:
                kids = [str(int(v)) for v in np.frombuffer(C.string_at(heap + 4 * off, 4 * ln), dtype=np.int32)]
            else:
                kids = ["%.9g" % float(v) for v in np.frombuffer(C.string_at(heap + 4 * off, 4 * ln), dtype=np.float32)]
            out.append("[" + ",".join(kids) + "]")
        return out
    if typ == 3:
        return ["\\N" if x is None else _esc(x) for x in cells_to_list(ptr, n, valid)]
    if typ == 0:
        vals = ["true" if v else "false" for v in np.frombuffer(C.string_at(ptr, n), dtype=np.uint8)]
    elif typ == 1:
        vals = [str(int(v)) for v in np.frombuffer(C.string_at(ptr, 4 * n), dtype=np.int32)]
    elif typ == 2:
        vals = ["%.9g" % float(v) for v in np.frombuffer(C.string_at(ptr, 4 * n), dtype=np.float32)]
    elif typ == "i64":
        vals = [str(int(v)) for v in np.frombuffer(C.string_at(ptr, 8 * n), dtype=np.int64)]
    else:
        vals = ["%.17g" % float(v) for v in np.frombuffer(C.string_at(ptr, 8 * n), dtype=np.float64)]
    if valid is not None:
        vals = [v if ok else "\\N" for v, ok in zip(vals, valid)]
    return vals


def _compact_format_text(ptr, vptr, n, f, heap, rec_ptr, n_rec, ns):
    """n FORMAT cells of the compact wire layout -> text (include/duckhts_cuda.h, dh_bcf_set_mode): GT cells are packed in 4 bytes,
    numeric LIST cells are an element count with the children at a per-record base; anything else has the DuckDB layout."""
    name, typ, is_list = f
    valid = validity_to_bool(vptr, n) if vptr else None
    if typ == 3 and not is_list and name == "GT":
        cells = np.frombuffer(C.string_at(ptr, 4 * n), dtype=np.uint32)
        out = []
        for i in range(n):
            if valid is not None and not valid[i]:
                out.append("\\N")
                continue
            v = int(cells[i])
            if not v >> 31:
                out.append(_esc(bytes([v & 0xff, (v >> 8) & 0xff, (v >> 16) & 0xff])[:(v >> 24) & 3]))
            else:
                off = v & 0x7fffffff
                ln = int(np.frombuffer(C.string_at(heap + off, 4), dtype=np.uint32)[0])
                out.append(_esc(C.string_at(heap + off + 4, ln)))
        return out
    if is_list and typ in (1, 2):
        cnt = np.frombuffer(C.string_at(ptr, 2 * n), dtype=np.uint16)
        rec = np.frombuffer(C.string_at(rec_ptr, 8 * n_rec), dtype=np.uint32).reshape(n_rec, 2)
        out = []
        for i in range(n):
            if valid is not None and not valid[i]:
                out.append("\\N")
                continue
            r, smp = divmod(i, ns)
            first = int(rec[r, 0]) + smp * int(rec[r, 1])
            raw = C.string_at(heap + 4 * first, 4 * int(cnt[i]))
            kids = [str(int(v)) for v in np.frombuffer(raw, dtype=np.int32)] if typ == 1 else ["%.9g" % float(v) for v in np.frombuffer(raw, dtype=np.float32)]
            out.append("[" + ",".join(kids) + "]")
        return out
    return _column_text(ptr, vptr, n, typ, is_list, heap)


def schema_names(bf, tidy):
    names = list(CORE_NAMES) + ["INFO_" + f[0] for f in bf.info]
    if bf.samples:
        if tidy:
            names += ["SAMPLE_ID"] + ["FORMAT_" + f[0] for f in bf.format]
        else:
            names += ["FORMAT_%s_%s" % (f[0], s.decode()) for s in bf.samples for f in bf.format]
    return names


def rows_text(ctx, data, tidy=True, batch_records=0, seg_comp=0, index=None, regions=None, compact=False):
    """Whole scan as text lines (header line first), like `fake_duckdb EXT read_bcf FILE [tidy_format=true] --dump`.
    index = CSI bytes + regions = list of region strings: the reference's chained single-region scans (bcf_reader.c:917-951)."""
    from .bam import Index
    with BcfFile(ctx, data) as bf, (Index(ctx, index) if index is not None else _NoIndex()) as ix:
        bf.configure(seg_comp, batch_records, 0)
        if compact:
            bf.set_mode(compact=True)
        lines = ["\t".join(schema_names(bf, tidy))]
        ns = len(bf.samples)
        status = 0

        def batches():
            if regions is None:
                bf.scan_begin()
                while (bt := bf.next_batch()) is not None:
                    yield bt
                return
            for rg in regions:                      # chained single-region scans, in the order given
                if not bf.scan_region(ix, rg):
                    continue                        # "Region query returned no iterator; skipping region"
                while (bt := bf.next_batch()) is not None:
                    yield bt

        for b in batches():
            status = b.status
            n = int(b.n_records)
            heap = b.heap or 0
            core_types = (3, "i64", 3, 3, 3, "f64", 3)
            site = [_column_text(b.core[c], b.core_validity[c], n, core_types[c], c in (4, 6), heap) for c in range(7)]
            site += [_column_text(b.info[i], b.info_validity[i], n, f[1], f[2], heap) for i, f in enumerate(bf.info)]
            if b.layout == 4:
                fmt = [_compact_format_text(b.format[i], b.format_validity[i], n * ns, f, heap, b.format_rec[i], n, ns) if b.format[i] else
                       _column_text(b.format[i], b.format_validity[i], n * ns, f[1], f[2], heap) for i, f in enumerate(bf.format)]
            else:
                fmt = [_column_text(b.format[i], b.format_validity[i], n * ns, f[1], f[2], heap) for i, f in enumerate(bf.format)]
            for r in range(n):
                srow = [col[r] for col in site]
                if not ns:
                    lines.append("\t".join(srow))
                elif tidy:
                    for s in range(ns):
                        lines.append("\t".join(srow + [_esc(bf.samples[s])] + [col[r * ns + s] for col in fmt]))
                else:
                    lines.append("\t".join(srow + [col[r * ns + s] for s in range(ns) for col in fmt]))
        st = bf.stats()
        return lines, st.status, status
