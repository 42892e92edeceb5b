"""Shared helpers for tests, smoke() and bench.py: synthetic inputs and column comparison."""
import hashlib
import os
import struct
import subprocess
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GEN = os.path.join(ROOT, "harness", "_bin", "gen_synth")
CACHE = os.environ.get("DUCKHTS_CACHE", "/tmp/duckhts_b200_cache")
GOLDEN = os.path.join(ROOT, "tests", "golden")

BGZF_EOF = bytes.fromhex("1f8b08040000000000ff0600424302001b0003000000000000000000")


def bgzf_block(payload, level=6, raw_deflate=None):
    """One BGZF block around `payload` (<= 65536 bytes).  raw_deflate overrides the compressor output."""
    if raw_deflate is None:
        c = zlib.compressobj(level, zlib.DEFLATED, -15, 8)
        raw_deflate = c.compress(payload) + c.flush()
    bsize = 18 + len(raw_deflate) + 8
    assert bsize <= 65536, bsize
    hdr = b"\x1f\x8b\x08\x04\0\0\0\0\0\xff\x06\0BC\x02\0" + struct.pack("<H", bsize - 1)
    return hdr + raw_deflate + struct.pack("<II", zlib.crc32(payload) & 0xffffffff, len(payload))


def make_bgzf(data, level=6, block=0xff00, eof=True, strategy=zlib.Z_DEFAULT_STRATEGY):
    out = []
    for i in range(0, len(data), block):
        chunk = bytes(data[i:i + block])
        c = zlib.compressobj(level, zlib.DEFLATED, -15, 8, strategy)
        raw = c.compress(chunk) + c.flush()
        if 18 + len(raw) + 8 > 65536:  # htslib/bgzf.c:631-638 falls back to a stored block
            c = zlib.compressobj(0, zlib.DEFLATED, -15, 8)
            raw = c.compress(chunk) + c.flush()
        out.append(bgzf_block(chunk, raw_deflate=raw))
    if eof:
        out.append(BGZF_EOF)
    return b"".join(out)


def chain_code_deflate(n_chains, n_tokens=20000, seed=0):
    """(payload, raw deflate) of one dynamic-Huffman block whose literal/length code is built to need large decode
    tables: 128 - n_chains codes of 7 bits plus n_chains "chains" of lengths 8, 9, ..., 14, 15, 15 (a chain fills one
    7-bit prefix and forces a 256-entry sub-table behind a 7-bit root).  Exercises the table-pool overflow paths of the
    inflate kernel (main pass / retry pass with the larger pool / lane-group kernel).  RFC 1951 3.2.7 encoder, no RLE
    of the code lengths (code-length code: symbols 0..15, 4 bits each)."""
    rng = np.random.default_rng(seed)
    lit_len = [0] * 286
    short = list(range(0, 128 - n_chains - 6)) + [256, 257, 258, 259, 260, 261]      # 7-bit codes; 257..261 = lengths 3..7
    for sym in short:
        lit_len[sym] = 7
    chain_syms = []
    nxt = 128 - n_chains - 6
    for _ in range(n_chains):
        for l in (8, 9, 10, 11, 12, 13, 14, 15, 15):
            lit_len[nxt] = l
            chain_syms.append(nxt)
            nxt += 1
    assert nxt <= 256 and abs(sum(2.0 ** -l for l in lit_len if l) - 1.0) < 1e-12
    dist_len = [1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 10]                                  # distance symbols 0..10 (1..48)
    assert abs(sum(2.0 ** -l for l in dist_len) - 1.0) < 1e-12

    def canon(lens):
        code, nc = 0, {}
        for l in range(1, 16):
            nc[l] = code
            code = (code + sum(1 for x in lens if x == l)) << 1
        out = {}
        for sym, l in enumerate(lens):
            if l:
                out[sym] = (nc[l], l)
                nc[l] += 1
        return out

    bits = []

    def put(v, n):                       # LSB-first field
        for i in range(n):
            bits.append((v >> i) & 1)

    def put_code(c):                     # Huffman code, MSB-first
        code, l = c
        for i in range(l - 1, -1, -1):
            bits.append((code >> i) & 1)

    order = [16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15]
    pre_len = [4] * 16 + [0, 0, 0]
    pre = canon(pre_len)
    put(1, 1); put(2, 2)
    put(len(lit_len) - 257, 5); put(len(dist_len) - 1, 5); put(19 - 4, 4)
    for sym in order:
        put(pre_len[sym], 3)
    for l in lit_len + dist_len:
        put_code(pre[l])
    lc, dc = canon(lit_len), canon(dist_len)
    lits = [s_ for s_ in short if s_ < 256] + chain_syms
    dist_base = [1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33]
    dist_extra = [0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4]
    out = bytearray()
    for _ in range(n_tokens):
        if len(out) > 64 and rng.random() < 0.3:
            ln = int(rng.integers(3, 8))
            ds = int(rng.integers(0, len(dist_len)))
            ex = int(rng.integers(0, 1 << dist_extra[ds]))
            d = dist_base[ds] + ex
            put_code(lc[257 + ln - 3]); put_code(dc[ds]); put(ex, dist_extra[ds])
            for _k in range(ln):
                out.append(out[-d])
        else:
            b = lits[int(rng.integers(0, len(lits)))] if rng.random() < 0.5 else chain_syms[int(rng.integers(0, len(chain_syms)))]
            put_code(lc[b])
            out.append(b)
        if len(out) > 65000:
            break
    put_code(lc[256])
    while len(bits) % 8:
        bits.append(0)
    raw = np.packbits(np.array(bits, dtype=np.uint8), bitorder="little").tobytes()
    return bytes(out), raw


def have_gen():
    return os.path.exists(GEN)


def synth_path(kind="bam", n_records=1000, seed=1, paired=False, contigs=22, extra=(), index=False):
    """Path of a cached synthetic file written by harness/_bin/gen_synth (reference htslib, BGZF level 6)."""
    if not have_gen():
        if os.path.exists("/root/reference/src/bam_reader.c"):
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "-j8", "ref"], stdout=subprocess.DEVNULL)
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "harness"), "_bin/gen_synth"], stdout=subprocess.DEVNULL)
        else:
            raise FileNotFoundError(f"{GEN} missing: run __graft_entry__.build() where /root/reference exists")
    os.makedirs(CACHE, exist_ok=True)
    key = f"{kind}_{n_records}_{seed}_{int(paired)}_{contigs}_{'_'.join(extra)}".replace("/", "-").replace("--", "")
    path = os.path.join(CACHE, key + (".bam" if kind == "bam" else ".bgzf"))
    if not os.path.exists(path) or (index and not os.path.exists(path + ".bai")):
        cmd = [GEN, kind, path + ".tmp", str(n_records), hex(seed), "--contigs", str(contigs)] + (["--paired"] if paired else []) + list(extra)
        if index:
            cmd.append("--index")
        subprocess.check_call(cmd)
        os.replace(path + ".tmp", path)
        if index:
            os.replace(path + ".tmp.bai", path + ".bai")
    return path


def synth_bam_bytes(**kw):
    return np.fromfile(synth_path("bam", **kw), dtype=np.uint8)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


STR_COLS = ("qname", "rname", "cigar", "rnext", "seq", "qual", "rg", "sample")
FIX_COLS = ("flag", "pos", "mapq", "pnext", "tlen")


def assert_bam_equal(got, exp, tags=True):
    """got: dict from duckhts_b200.bam_scan; exp: dict from oracle.dh_oracle_py.bam_decode.  Bit-exact."""
    from oracle import dh_oracle_py as orc

    assert got["n"] == exp["n"], (got["n"], exp["n"])
    assert got["status"] == exp["status"], (got["status"], exp["status"])
    for c in FIX_COLS:
        assert got[c].dtype == exp[c].dtype, (c, got[c].dtype, exp[c].dtype)
        if not np.array_equal(got[c], exp[c]):
            i = int(np.nonzero(got[c] != exp[c])[0][0])
            raise AssertionError(f"column {c} differs first at row {i}: {got[c][i]} != {exp[c][i]}")
    for c in STR_COLS:
        g, e = got[c], orc.strcol_list(exp[c])
        assert len(g) == len(e), c
        if g != e:
            i = next(k for k in range(len(e)) if g[k] != e[k])
            raise AssertionError(f"column {c} differs first at row {i}: {g[i]!r} != {e[i]!r}")
    if tags:
        assert len(got["tags"]) == len(exp["tags"])
        for k, (gt, et) in enumerate(zip(got["tags"], exp["tags"])):
            assert np.array_equal(gt["valid"], et["valid"]), f"tag {k} validity"
            if "i" in et:
                assert np.array_equal(gt["i"][et["valid"]], et["i"][et["valid"]]), f"tag {k} int values"
            elif "f" in et:
                assert np.array_equal(gt["f"][et["valid"]].view(np.uint64), et["f"][et["valid"]].view(np.uint64)), f"tag {k} f"
            elif "b" in et:
                assert len(gt["b"]) == len(et["b"])
                for r, (g, e) in enumerate(zip(gt["b"], et["b"])):
                    assert (g is None) == (e is None) and (e is None or np.array_equal(g, e)), f"tag {k} list, row {r}: {g} != {e}"
            else:
                assert gt["s"] == orc.strcol_list(et["s"]), f"tag {k} strings"


# ---- hand-made BAM streams for edge cases ---------------------------------------------------------------
def bam_header(refs, text=b"@HD\tVN:1.6\n"):
    out = [b"BAM\x01", struct.pack("<I", len(text)), text, struct.pack("<I", len(refs))]
    for name, ln in refs:
        nb = name + b"\0"
        out += [struct.pack("<I", len(nb)), nb, struct.pack("<I", ln)]
    return b"".join(out)


CIG_OPS = "MIDNSHP=XB"
NT16 = "=ACMGRSVTWYHKDBN"


def bam_record(qname=b"r", flag=0, tid=0, pos=0, mapq=0, cigar=(), mtid=-1, mpos=-1, tlen=0, seq="", qual=None, aux=b"",
               raw_qname=None, bin_=0):
    """cigar: iterable of (len, opchar).  qual: bytes of phred values, None -> 0xff * len.  raw_qname overrides the
    NUL-terminated name bytes."""
    qn = raw_qname if raw_qname is not None else qname + b"\0"
    cig = b"".join(struct.pack("<I", (l << 4) | CIG_OPS.index(o)) for l, o in cigar)
    l_seq = len(seq)
    packed = bytearray((l_seq + 1) // 2)
    for i, ch in enumerate(seq):
        packed[i >> 1] |= NT16.index(ch) << (4 if i % 2 == 0 else 0)
    q = bytes([0xff] * l_seq) if qual is None else bytes(qual)
    assert len(q) == l_seq
    body = struct.pack("<iiBBHHHiiii", tid, pos, len(qn), mapq, bin_, len(cigar), flag, l_seq, mtid, mpos, tlen) + qn + cig + bytes(packed) + q + aux
    return struct.pack("<I", len(body)) + body


def aux_tag(tag, ty, val):
    t = tag if isinstance(tag, bytes) else tag.encode()
    fmt = {"c": "<b", "C": "<B", "s": "<h", "S": "<H", "i": "<i", "I": "<I", "f": "<f", "d": "<d"}
    if ty in fmt:
        return t + ty.encode() + struct.pack(fmt[ty], val)
    if ty == "A":
        return t + b"A" + (val if isinstance(val, bytes) else val.encode())
    if ty in ("Z", "H"):
        return t + ty.encode() + (val if isinstance(val, bytes) else val.encode()) + b"\0"
    if ty == "B":
        sub, vals = val
        return t + b"B" + sub.encode() + struct.pack("<I", len(vals)) + b"".join(struct.pack(fmt[sub], v) for v in vals)
    raise ValueError(ty)


# ---- canonical text form shared with harness/ref_dump.c (the reference-generated golden vectors) -------------
GOLDEN_TAGS = [(b"NM", "i"), (b"MD", "Z"), (b"AS", "i"), (b"XS", "i"), (b"XA", "A"), (b"XZ", "Z"), (b"Xf", "f"), (b"F0", "f"),
               (b"I1", "i"), (b"ZA", "A"), (b"H0", "Z")]


def _esc(b):
    return "".join(chr(c) if 0x21 <= c <= 0x7e and c != 0x5c else "\\x%02x" % c for c in b)


def dump_lines(res, tags, from_oracle):
    """Rows of a scan result (oracle dict or duckhts_b200.bam_scan dict) in ref_dump's text form."""
    from oracle import dh_oracle_py as orc

    def strs(name):
        return orc.strcol_list(res[name]) if from_oracle else res[name]

    S = {n: strs(n) for n in STR_COLS}
    tag_cols = []
    for k, (tg, ty) in enumerate(tags):
        t = res["tags"][k]
        if ty == "i":
            tag_cols.append([str(int(v)) if ok else "\\N" for v, ok in zip(t["i"], t["valid"])])
        elif ty == "f":
            tag_cols.append(["%016x" % int(np.float64(v).view(np.uint64)) if ok else "\\N" for v, ok in zip(t["f"], t["valid"])])
        else:
            s = orc.strcol_list(t["s"]) if from_oracle else t["s"]
            tag_cols.append([("\\N" if x is None else (_esc(x) if x else "\\e")) for x in s])
    lines = ["#n_ref %d" % len(res["ref_names"])]
    for nm, ln in zip(res["ref_names"], res["ref_lens"]):
        lines.append("#ref %s %d" % (_esc(nm), ln))
    for r in range(res["n"]):
        f = [_esc(S["qname"][r]), str(int(res["flag"][r])), _esc(S["rname"][r]), str(int(res["pos"][r])), str(int(res["mapq"][r])),
             _esc(S["cigar"][r]), _esc(S["rnext"][r]), str(int(res["pnext"][r])), str(int(res["tlen"][r])), _esc(S["seq"][r]),
             _esc(S["qual"][r]), "\\N" if S["rg"][r] is None else _esc(S["rg"][r]), "\\N" if S["sample"][r] is None else _esc(S["sample"][r])]
        f += [tc[r] for tc in tag_cols]
        lines.append("\t".join(f))
    lines.append("#status %d %d" % (res["status"], res["n"]))
    return lines


def golden_names():
    return sorted(f[:-4] for f in os.listdir(GOLDEN) if f.endswith(".bam"))


# ---- per-column checksums (dh_bam_batch_checksum, duckhts_b200/csrc/dh_checksum.cu), restated with numpy over the oracle's columns ----
CK_NULL = np.uint64(0x9e3779b97f4a7c15)
CK_K = np.uint64(0x100000001b3)


def ck_mix(x):
    x = np.asarray(x, dtype=np.uint64).copy()
    with np.errstate(over="ignore"):
        x ^= x >> np.uint64(30); x *= np.uint64(0xbf58476d1ce4e5b9)
        x ^= x >> np.uint64(27); x *= np.uint64(0x94d049bb133111eb)
        x ^= x >> np.uint64(31)
    return x


def _ck_segments(values, off):
    """Per row r: mix(H ^ n * CK_NULL), H = Horner polynomial of values[off[r]:off[r+1]] (+1 each) in CK_K, n = its length."""
    off = np.asarray(off, dtype=np.int64)
    n = len(off) - 1
    lens = off[1:] - off[:-1]
    total = int(off[-1] - off[0])
    with np.errstate(over="ignore"):
        if total:
            maxlen = int(lens.max())
            kpow = np.ones(maxlen + 1, dtype=np.uint64)
            for i in range(1, maxlen + 1):
                kpow[i] = kpow[i - 1] * CK_K
            row = np.repeat(np.arange(n), lens)
            pos = np.arange(total) - np.repeat(off[:-1] - off[0], lens)
            terms = (np.asarray(values[off[0]:off[-1]], dtype=np.uint64) + np.uint64(1)) * kpow[lens[row] - 1 - pos]
            cs = np.concatenate([[np.uint64(0)], np.cumsum(terms, dtype=np.uint64)])
            h = cs[off[1:] - off[0]] - cs[off[:-1] - off[0]]
        else:
            h = np.zeros(n, dtype=np.uint64)
        return ck_mix(h ^ (lens.astype(np.uint64) * CK_NULL))


def ck_strcol(col, rows=None):
    """Checksum of an oracle string column {off, heap, valid} (NULL rows count CK_NULL)."""
    cell = _ck_segments(col["heap"], col["off"])
    cell = np.where(col["valid"], cell, CK_NULL)
    if rows is not None:
        cell = cell[rows]
    with np.errstate(over="ignore"):
        return int(cell.sum(dtype=np.uint64))


def oracle_checksums(exp, tags, rows=None):
    """Per-column checksums (13 core columns, then the tag columns) of an oracle bam_decode result; rows = boolean mask / index."""
    def tot(cell):
        cell = np.asarray(cell, dtype=np.uint64)
        if rows is not None:
            cell = cell[rows]
        with np.errstate(over="ignore"):
            return int(cell.sum(dtype=np.uint64))
    out = []
    for name in ("qname", "flag", "rname", "pos", "mapq", "cigar", "rnext", "pnext", "tlen", "seq", "qual", "rg", "sample"):
        if name in FIX_COLS:
            out.append(tot(ck_mix(exp[name].astype(np.int64).view(np.uint64))))
        else:
            out.append(ck_strcol(exp[name], rows))
    for k, (tg, ty) in enumerate(tags):
        t = exp["tags"][k]
        if ty == "i":
            out.append(tot(np.where(t["valid"], ck_mix(t["i"].astype(np.int64).view(np.uint64)), CK_NULL)))
        elif ty == "f":
            out.append(tot(np.where(t["valid"], ck_mix(t["f"].view(np.uint64)), CK_NULL)))
        elif ty[0] == "B":
            lens = np.array([0 if x is None else len(x) for x in t["b"]], dtype=np.int64)
            vals = np.concatenate([np.zeros(0, np.int64)] + [x for x in t["b"] if x is not None]).view(np.uint64)
            cell = _ck_segments_words(vals, np.concatenate([[0], np.cumsum(lens)]))
            out.append(tot(np.where(t["valid"], cell, CK_NULL)))
        else:
            out.append(ck_strcol(t["s"], rows))
    return out


def _ck_segments_words(values, off):
    """Like _ck_segments for 8-byte children: the addends are the words themselves (+1, modulo 2^64)."""
    off = np.asarray(off, dtype=np.int64)
    n = len(off) - 1
    lens = off[1:] - off[:-1]
    total = int(off[-1])
    with np.errstate(over="ignore"):
        if total:
            maxlen = int(lens.max())
            kpow = np.ones(maxlen + 1, dtype=np.uint64)
            for i in range(1, maxlen + 1):
                kpow[i] = kpow[i - 1] * CK_K
            row = np.repeat(np.arange(n), lens)
            pos = np.arange(total) - np.repeat(off[:-1], lens)
            terms = (np.asarray(values, dtype=np.uint64) + np.uint64(1)) * kpow[lens[row] - 1 - pos]
            cs = np.concatenate([[np.uint64(0)], np.cumsum(terms, dtype=np.uint64)])
            h = cs[off[1:]] - cs[off[:-1]]
        else:
            h = np.zeros(n, dtype=np.uint64)
        return ck_mix(h ^ (lens.astype(np.uint64) * CK_NULL))
