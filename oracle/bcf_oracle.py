"""bcf_oracle.py — CPU restatement of the reference's read_bcf path.  TEST INFRASTRUCTURE ONLY:
imported by tests/ (and smoke()); never by duckhts_b200/.

Parity pinned: tests/test_oracle_bcf.py checks it against dumps the REFERENCE EXTENSION itself produced under the fake
DuckDB host (tests/golden/*.bcf.{tidy,wide}.txt; tests/golden/make_golden.sh) and, where the prebuilt reference is
present, against live runs on seeded synthetic files.

Pure-Python loops: meant for files of a few thousand (record, sample) cells.  Follows, in order:
  bcf_hdr_read / bcf_hdr_parse / bcf_hdr_register_hrec     htslib/vcf.c:1710+, 1410+, 831-1010   (dictionaries)
  bcf_read_bind                                           src/bcf_reader.c:552-758              (schema)
  vcf_validate_*_field / vcf_is_list_type                 src/include/vcf_types.h:25-205
  bcf_read1_core, bcf_record_check                        htslib/vcf.c:1874-1911, 2040-2212
  bcf_unpack, bcf_fmt_array                               vcf.c:4178-4302
  bcf_get_info_values / bcf_get_format_values / bcf_get_format_string   vcf.c:6056-6248
  column writers incl. the GT string builder              src/bcf_reader.c:1381-1462, 1542-1733, 1738-1998
Output: text rows exactly as harness/fake_duckdb.c prints chunks (format_cell)."""
import re
import struct

HT_FLAG, HT_INT, HT_REAL, HT_STR = 0, 1, 2, 3
SHIFT = [0, 0, 1, 2, 3, 2, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0]          # bcf_type_shift, vcf.c:91
I32_MISSING, I32_VEND, F32_MISSING, F32_VEND = 0x80000000, 0x80000001, 0x7F800001, 0x7F800002
INFO_LIST = {"AD", "ADF", "ADR", "AC", "AF", "CIGAR"}
INFO_SCALAR = {"AA", "AN", "BQ", "DB", "DP", "END", "H2", "H3", "MQ", "MQ0", "NS", "SB", "SOMATIC", "VALIDATED", "1000G"}
FMT_LIST = {"AD", "ADF", "ADR", "EC", "GL", "GP", "PL", "PP"}
FMT_SCALAR = {"DP", "LEN", "FT", "GQ", "GT", "HQ", "MQ", "PQ", "PS"}


class Header:
    def __init__(self, text):
        self.ids = []            # dictionary: dict(name, has={cls}, type={cls: ht}, fixed={cls: bool}) or None for holes
        self.ctg = []
        self.samples = []
        self.version = (4, 2)
        self._add("FILTER", {"ID": "PASS"})                       # bcf_hdr_parse registers PASS first
        for line in text.split("\n"):
            line = line.rstrip("\r")
            if line.startswith("##"):
                key, eq, val = line[2:].partition("=")
                if not eq:
                    continue
                if key == "fileformat":
                    m = re.match(r"VCFv(\d+)\.(\d+)", val)
                    if m:
                        self.version = (int(m.group(1)), int(m.group(2)))
                elif val.startswith("<"):
                    kv = self._struct(val)
                    if key == "contig":
                        self._add_ctg(kv)
                    elif key in ("FILTER", "INFO", "FORMAT"):
                        self._add(key, kv)
            elif line.startswith("#"):
                self.samples = [s for s in line.split("\t")[9:] if s]
                break

    @staticmethod
    def _struct(v):
        kv, p = {}, 1
        while p < len(v) and v[p] != ">":
            while p < len(v) and v[p] in " ,":
                p += 1
            ks = p
            while p < len(v) and v[p] not in "=,>":
                p += 1
            if p >= len(v) or v[p] != "=":
                if p < len(v) and v[p] == ">":
                    break
                p += 1
                continue
            key = v[ks:p].rstrip(" ")
            p += 1
            val = ""
            if p < len(v) and v[p] == '"':
                p += 1
                while p < len(v) and v[p] != '"':
                    if v[p] == "\\" and p + 1 < len(v):
                        val += v[p:p + 2]
                        p += 2
                        continue
                    val += v[p]
                    p += 1
                p += 1
            else:
                depth = 0
                while p < len(v) and not (depth == 0 and v[p] in ",>"):
                    depth += (v[p] == "<") - (v[p] == ">")
                    val += v[p]
                    p += 1
            kv.setdefault(key, val)
        return kv

    def _add_ctg(self, kv):
        if "ID" not in kv or kv["ID"] in [c for c in self.ctg if c is not None]:
            return
        idx = int(kv["IDX"]) if kv.get("IDX", "").isdigit() else len(self.ctg)
        while len(self.ctg) <= idx:
            self.ctg.append(None)
        if self.ctg[idx] is None:
            self.ctg[idx] = kv["ID"]

    def _add(self, cls, kv):
        if "ID" not in kv:
            return
        ht, fixed = HT_STR, False
        if cls != "FILTER":
            ht = {"Integer": HT_INT, "Float": HT_REAL, "Flag": HT_FLAG}.get(kv.get("Type"), HT_STR)
            num = kv.get("Number", ".")
            fixed = bool(re.match(r"[+-]?\d+", num)) and not (cls == "FORMAT" and num in ("P", "LA", "LR", "LG", "M"))
            if num in ("A", "R", "G", "."):
                fixed = False
            if ht == HT_FLAG:
                fixed = True
        at = next((i for i, d in enumerate(self.ids) if d and d["name"] == kv["ID"]), None)
        if at is None:
            at = int(kv["IDX"]) if kv.get("IDX", "").isdigit() else len(self.ids)
            while len(self.ids) <= at:
                self.ids.append(None)
            if self.ids[at] is not None:
                return
            self.ids[at] = {"name": kv["ID"], "has": set(), "type": {}, "fixed": {}}
        d = self.ids[at]
        if cls in d["has"]:
            return                                                  # first definition wins
        d["has"].add(cls)
        d["type"][cls] = ht
        d["fixed"][cls] = fixed

    def fields(self, cls):
        out = []
        for i, d in enumerate(self.ids):
            if d and cls in d["has"]:
                name = d["name"]
                scalar, lst = (INFO_SCALAR, INFO_LIST) if cls == "INFO" else (FMT_SCALAR, FMT_LIST)
                is_list = False if name in scalar else True if name in lst else not d["fixed"][cls]
                out.append((name, i, d["type"][cls], is_list))
        return out


def _esc(b):
    out = []
    for c in b:
        out.append("\\t" if c == 9 else "\\n" if c == 10 else "\\\\" if c == 92 else "\\x%02x" % c if (c < 32 or c == 127) else chr(c))
    return "".join(out)


def _typed_int1(u, p, end):
    if end - p < 2:
        raise ValueError
    t = u[p] & 15
    p += 1
    if t == 1:
        return struct.unpack_from("<b", u, p)[0], p + 1
    if end - p < (1 << SHIFT[t]):
        raise ValueError
    if t == 2:
        return struct.unpack_from("<h", u, p)[0], p + 2
    if t == 3:
        return struct.unpack_from("<i", u, p)[0], p + 4
    raise ValueError


def _dec_size(u, p, end):
    if p >= end:
        raise ValueError
    b = u[p]
    if b >> 4 != 15:
        return b >> 4, b & 15, p + 1
    num, q = _typed_int1(u, p + 1, end)
    if num < 0:
        raise ValueError
    return num, b & 15, q


def _word(u, base, j, t):
    """value j of a typed vector as the 32-bit word bcf_get_*_values store (vcf.c:6098-6131, 6219-6244)"""
    if t == 1:
        v = u[base + j]
        return I32_MISSING if v == 0x80 else I32_VEND if v == 0x81 else struct.unpack_from("<b", u, base + j)[0] & 0xffffffff
    if t == 2:
        v = struct.unpack_from("<H", u, base + 2 * j)[0]
        return I32_MISSING if v == 0x8000 else I32_VEND if v == 0x8001 else struct.unpack_from("<h", u, base + 2 * j)[0] & 0xffffffff
    return struct.unpack_from("<I", u, base + 4 * j)[0]


def _cstr(u, p, n):
    s = bytes(u[p:p + n])
    k = s.find(b"\0")
    return s if k < 0 else s[:k]


def _num_text(w, ht):
    if ht == HT_INT:
        return str(struct.unpack("<i", struct.pack("<I", w))[0])
    return "%.9g" % struct.unpack("<f", struct.pack("<I", w))[0]


def scan(u, tidy, regions=None):
    """u: inflated bytes of a BCF file.  Returns (lines, status).
    regions: list of region strings = the reference's chained single-region scans (src/bcf_reader.c:917-951, 1318-1333):
    one pass per region that names a contig, in the order given, keeping the records bcf_itr_next returns -- those with
    rid == tid, pos < end and pos + rlen > beg (bcf_readrec, htslib/vcf.c:2267-2276; hts_itr_next, hts.c:4304-4313).  On a
    coordinate-sorted file (the only kind that can be indexed) the index only decides which blocks are read, not which
    records qualify."""
    if regions is not None:
        from oracle.dh_oracle_py import parse_region
        head = _scan(u, tidy, lambda rid, pos, rlen: False)[0][:1]
        names = _contig_names(u)
        lines, status = list(head), 0
        for rg in regions:
            iv = parse_region(rg, names)
            if iv is None or iv[0] < 0:
                continue
            tid, beg, end = iv
            part, status = _scan(u, tidy, lambda rid, pos, rlen: rid == tid and pos < end and pos + rlen > beg)
            lines += part[1:]
        return lines, status
    return _scan(u, tidy, None)


def _contig_names(u):
    u = bytes(u)
    l_text = struct.unpack_from("<I", u, 5)[0]
    h = Header(u[9:9 + l_text].split(b"\0")[0].decode("utf-8", "replace"))
    return [c if c is not None else "" for c in h.ctg]


def _scan(u, tidy, keep):
    u = bytes(u)
    if u[:3] != b"BCF" or u[3] != 2:
        raise ValueError("not BCF")
    l_text = struct.unpack_from("<I", u, 5)[0]
    text = u[9:9 + l_text].split(b"\0")[0].decode("utf-8", "replace")
    h = Header(text)
    info, fmt = h.fields("INFO"), h.fields("FORMAT") if h.samples else []
    ns = len(h.samples)
    default_gt = bool(ns) and not fmt
    names = ["CHROM", "POS", "ID", "REF", "ALT", "QUAL", "FILTER"] + ["INFO_" + f[0] for f in info]
    fmt_names = ["GT"] if default_gt else [f[0] for f in fmt]
    if ns:
        names += (["SAMPLE_ID"] + ["FORMAT_" + f for f in fmt_names]) if tidy else ["FORMAT_%s_%s" % (f, s) for s in h.samples for f in fmt_names]
    lines = ["\t".join(names)]
    gt_id = next((i for i, d in enumerate(h.ids) if d and d["name"] == "GT"), -1) if h.version < (4, 4) else -1
    p, n, status = 9 + l_text, len(u), 0
    id_ok = lambda k: 0 <= k < len(h.ids) and h.ids[k] is not None
    while p < n:
        try:
            if n - p < 32:
                raise ValueError
            l_shared, l_indiv, rid, pos, rlen, qual_bits, n_info, n_allele, w = struct.unpack_from("<IIiIiIHHI", u, p)
            if l_shared < 24:
                raise ValueError
            end = p + 8 + l_shared + l_indiv
            if end > n:
                raise ValueError
            n_sample, n_fmt = w & 0xffffff, w >> 24
            if (not l_indiv or not n_sample) and n_fmt:
                n_fmt = 0
            send = p + 8 + l_shared
            err = not (0 <= rid < len(h.ctg) and h.ctg[rid] is not None)
            q = p + 32
            num, t, q = _dec_size(u, q, send)                      # ID
            err |= t != 7
            if send - q < num << SHIFT[t]:
                raise ValueError
            ident = b"." if num == 0 else _cstr(u, q, num)
            q += num << SHIFT[t]
            err |= n_allele < 1
            alleles = []
            for _ in range(n_allele):
                num, t, q = _dec_size(u, q, send)
                err |= t != 7
                if send - q < num << SHIFT[t]:
                    raise ValueError
                alleles.append(b"." if num == 0 else _cstr(u, q, num))
                q += num << SHIFT[t]
            num, t, q = _dec_size(u, q, send)                      # FILTER
            flt = []
            if num > 0:
                if send - q < num << SHIFT[t]:
                    raise ValueError
                if t not in (1, 2, 3):
                    err = True
                else:
                    for j in range(num):
                        k = struct.unpack_from(("<b", "<h", "<i")[t - 1], u, q + (j << SHIFT[t]))[0]
                        err |= not id_ok(k)
                        flt.append(k)
                q += num << SHIFT[t]
            infos = []
            for _ in range(n_info):
                key, q = _typed_int1(u, q, send)
                err |= not id_ok(key)
                num, t, q = _dec_size(u, q, send)
                err |= t not in (0, 1, 2, 3, 5, 7) or (t == 0 and num > 0)
                if send - q < num << SHIFT[t]:
                    raise ValueError
                infos.append((key, num, t, q))
                q += num << SHIFT[t]
            fmts, q = [], send
            for _ in range(n_fmt):
                key, q = _typed_int1(u, q, end)
                err |= not id_ok(key)
                num, t, q = _dec_size(u, q, end)
                err |= t not in (0, 1, 2, 3, 5, 7) or (t == 0 and num > 0)
                nbytes = (num << SHIFT[t]) * n_sample
                if end - q < nbytes:
                    raise ValueError
                fmts.append((key, num, t, q))
                q += nbytes
            if n_fmt and n_sample != ns:
                err = True
            if err:
                raise ValueError
        except (ValueError, struct.error):
            status = -2                                            # bcf_read < -1: the scan ends silently (bcf_reader.c:1311-1338)
            break
        if keep is not None and not keep(rid, -1 if pos == 0xffffffff else pos, rlen):
            p = end
            continue
        # ---- site-level cells ----
        site = [_esc(h.ctg[rid].encode()), str((-1 if pos == 0xffffffff else pos) + 1), "\\N" if ident == b"." else _esc(ident), _esc(alleles[0]),
                "[" + ",".join(_esc(a) for a in alleles[1:]) + "]",
                "\\N" if qual_bits == F32_MISSING else "%.17g" % struct.unpack("<f", struct.pack("<I", qual_bits))[0],
                "[" + (",".join(_esc(h.ids[k]["name"].encode()) for k in flt) if flt else "PASS") + "]"]
        for name, fid, ht, is_list in info:
            ent = next((e for e in infos if e[0] == fid), None)     # first entry with this key (vcf.c:6063-6065)
            if ht == HT_FLAG:
                site.append("true" if ent else "false")
                continue
            cell = "\\N"
            if ent:
                _, num, t, vp = ent
                if ht == HT_STR:
                    if num > 0:
                        s = _cstr(u, vp, num)
                        if s != b".":
                            if not is_list:
                                cell = _esc(s)
                            else:                                    # process_comma_separated_list (bcf_reader.c:996-1061)
                                toks = s.split(b",")
                                if toks and toks[-1] == b"":
                                    toks.pop()
                                cell = "[" + ",".join(_esc(x) for x in toks) + "]"
                elif t in (1, 2, 3, 5):
                    vend_rec = F32_VEND if t == 5 else I32_VEND
                    words = []
                    for j in range(num):
                        wj = _word(u, vp, j, t)
                        if wj == vend_rec:
                            break
                        words.append(wj)
                    miss, vend = (I32_MISSING, I32_VEND) if ht == HT_INT else (F32_MISSING, F32_VEND)
                    if words:
                        if is_list:
                            cell = "[" + ",".join(_num_text(x, ht) for x in words if x not in (miss, vend)) + "]"
                        elif words[0] != miss:
                            cell = _num_text(words[0], ht)
            site.append(cell)
        if not ns:
            lines.append("\t".join(site))
            p = end
            continue
        # ---- FORMAT cells per sample ----
        per_sample = [[] for _ in range(ns)]
        for name, fid, ht, is_list in ([("GT", -1, HT_STR, False)] if default_gt else fmt):
            ent = next((e for e in fmts if e[0] == fid), None)
            for s in range(ns):
                cell = "\\N"
                if ent:
                    _, num, t, base = ent
                    vp = base + s * (num << SHIFT[t])
                    if ht == HT_STR and name != "GT":
                        if not is_list:                              # bcf_get_format_string: n bytes per sample, C string
                            cell = _esc(_cstr(u, base + s * num, num))
                    elif num > 0 and t in (1, 2, 3, 5):
                        if name == "GT" and ht == HT_STR:           # GT string builder (bcf_reader.c:1904-1957)
                            out = ""
                            for a in range(num if t != 5 else 0):
                                wv = _word(u, vp, a, t)
                                if wv == I32_VEND:
                                    break
                                v = struct.unpack("<i", struct.pack("<I", wv))[0]
                                if a:
                                    out += "|" if v & 1 else "/"
                                out += "." if (v >> 1) == 0 else str((v >> 1) - 1)
                            if out:
                                cell = out
                        elif ht in (HT_INT, HT_REAL):
                            miss, vend = (I32_MISSING, I32_VEND) if ht == HT_INT else (F32_MISSING, F32_VEND)
                            vend_rec = F32_VEND if t == 5 else I32_VEND
                            if not is_list:
                                w0 = _word(u, vp, 0, t)
                                if w0 != miss:
                                    cell = _num_text(w0, ht)
                            else:
                                vals = []
                                for j in range(num):
                                    wj = _word(u, vp, j, t)
                                    if wj == vend_rec:
                                        break
                                    if wj not in (miss, vend):
                                        vals.append(_num_text(wj, ht))
                                cell = "[" + ",".join(vals) + "]"
                per_sample[s].append(cell)
        if tidy:
            for s in range(ns):
                lines.append("\t".join(site + [_esc(h.samples[s].encode())] + per_sample[s]))
        else:
            lines.append("\t".join(site + [c for s in range(ns) for c in per_sample[s]]))
        p = end
    return lines, status
