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
            else:
                assert gt["s"] == orc.strcol_list(et["s"]), f"tag {k} strings"
