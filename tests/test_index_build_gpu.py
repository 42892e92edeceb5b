"""N2: bam_index on the GPU (dh_bam_index_build; reference src/hts_index_builder.c:110-160 -> sam_index_build3 -> hts_idx_push /
hts_idx_finish / hts_idx_save, htslib/hts.c:2400-2860).  The BAI must hold what htslib's holds: the same bins with the same chunk
lists, the same linear index, the same per-contig mapped / unmapped counts and n_no_coor (htslib writes the bins in its hash table's
order, we write them ascending: compared as structures), and the REFERENCE must answer region queries from our index as from its own."""
import os
import struct
import subprocess

import numpy as np
import pytest

import duckhts_b200 as dh
from tests import util

pytestmark = pytest.mark.gpu

ROOT = util.ROOT
HOST = os.path.join(ROOT, "harness", "_bin", "fake_duckdb")
REF_EXT = os.path.join(ROOT, "oracle", "_ref", "libduckhts_ref.so")
OUR_EXT = os.path.join(ROOT, "duckhts_b200", "lib", "duckhts.duckdb_extension.so")
REF_DUMP = os.path.join(ROOT, "harness", "_bin", "ref_dump")


def parse_bai(b):
    """BAI bytes -> ([{"bins": {bin: [(beg, end), ...]}, "linear": [ioffset, ...]}, ...], n_no_coor)"""
    assert b[:4] == b"BAI\x01"
    n_ref, = struct.unpack_from("<i", b, 4)
    p = 8
    refs = []
    for _ in range(n_ref):
        n_bin, = struct.unpack_from("<i", b, p)
        p += 4
        bins = {}
        for _ in range(n_bin):
            bin_, n_chunk = struct.unpack_from("<Ii", b, p)
            p += 8
            chunks = [struct.unpack_from("<QQ", b, p + 16 * k) for k in range(n_chunk)]
            p += 16 * n_chunk
            assert bin_ not in bins
            bins[bin_] = chunks
        n_intv, = struct.unpack_from("<i", b, p)
        p += 4
        lin = list(struct.unpack_from("<%dQ" % n_intv, b, p))
        p += 8 * n_intv
        refs.append({"bins": bins, "linear": lin})
    n_no_coor = struct.unpack_from("<Q", b, p)[0] if p + 8 <= len(b) else None
    return refs, n_no_coor


def assert_same_index(ours, theirs):
    (ro, no), (rt, nt) = parse_bai(ours), parse_bai(theirs)
    assert len(ro) == len(rt) and no == nt
    for t, (a, b) in enumerate(zip(ro, rt)):
        assert sorted(a["bins"]) == sorted(b["bins"]), (t, sorted(set(a["bins"]) ^ set(b["bins"]))[:10])
        for k in b["bins"]:
            assert a["bins"][k] == b["bins"][k], (t, k, a["bins"][k][:3], b["bins"][k][:3])
        assert a["linear"] == b["linear"], t


@pytest.mark.parametrize("kw", [dict(n_records=30000, seed=31, paired=True, extra=("--unplaced", "25")), dict(n_records=20000, seed=11),
                                dict(n_records=120000, seed=21, paired=True), dict(n_records=6000, seed=41, paired=True, contigs=3)])
def test_bai_holds_what_htslibs_holds(ctx, kw):
    path = util.synth_path("bam", index=True, **kw)          # PATH.bai written by the reference's htslib (sam_index_build)
    theirs = open(path + ".bai", "rb").read()
    img = np.fromfile(path, dtype=np.uint8)
    with dh.BamFile(ctx, img) as bf:                          # memory image, small segments: runs continue across blocks and segments
        bf.configure(1 << 18, 0, 0)
        assert_same_index(bf.index_build(), theirs)
    with dh.BamFile.open_file(ctx, path) as bf:               # streamed file
        ours = bf.index_build()
        assert_same_index(ours, theirs)
        # the handle is still good for scans afterwards, and the index we built drives them
        bf.set_projection(1 | (1 << 3), ())
        with dh.Index(ctx, np.frombuffer(ours, dtype=np.uint8)) as ix:
            assert bf.scan_regions(ix, ["chr2:1-5000000"])
            n = bf.drain()
    with dh.BamFile(ctx, img) as bf, dh.Index(ctx, np.frombuffer(theirs, dtype=np.uint8)) as ix:
        bf.set_projection(1 | (1 << 3), ())
        assert bf.scan_regions(ix, ["chr2:1-5000000"]) and bf.drain() == n


def test_reference_fixture_range_bam(ctx):
    """test/data/range.bam and its .bai as shipped by the reference."""
    path = os.path.join(util.GOLDEN, "range.bam")
    with dh.BamFile.open_file(ctx, path) as bf:
        assert_same_index(bf.index_build(), open(path + ".bai", "rb").read())


def test_unsorted_and_truncated_files_are_refused(ctx, tmp_path):
    from tests.util import bam_header, bam_record
    recs = [bam_record(b"r%d" % i, 0, 0, 1000 - i, 1, [(4, "M")], seq="ACGT", qual=[1] * 4) for i in range(50)]
    img = np.frombuffer(util.make_bgzf(bam_header([(b"c1", 100000)]) + b"".join(recs)), dtype=np.uint8)
    with dh.BamFile(ctx, img) as bf, pytest.raises(dh.DuckhtsError, match="unsorted"):
        bf.index_build()
    recs = [bam_record(b"r%d" % i, 0, i % 2, 10 + i, 1, [(4, "M")], seq="ACGT", qual=[1] * 4) for i in range(50)]      # contigs alternate
    img = np.frombuffer(util.make_bgzf(bam_header([(b"c1", 100000), (b"c2", 100000)]) + b"".join(recs)), dtype=np.uint8)
    with dh.BamFile(ctx, img) as bf, pytest.raises(dh.DuckhtsError, match="continuous"):
        bf.index_build()
    src = util.synth_bam_bytes(n_records=20000, seed=11)
    with dh.BamFile(ctx, src[: src.size // 2]) as bf, pytest.raises(dh.DuckhtsError):
        bf.index_build()


@pytest.mark.skipif(not (os.path.exists(HOST) and os.path.exists(REF_EXT)), reason="prebuilt fake host / reference extension absent")
def test_bam_index_table_function_and_the_reference_reading_our_index(tmp_path):
    import shutil
    src = util.synth_path("bam", n_records=30000, seed=31, paired=True, extra=("--unplaced", "25"))
    path = str(tmp_path / "x.bam")
    shutil.copyfile(src, path)

    def run(ext, func, *args):
        p = subprocess.run([HOST, ext, func, *args], stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=600)
        return p.returncode, p.stdout.decode()

    ours, theirs = str(tmp_path / "ours.bai"), str(tmp_path / "theirs.bai")
    rc_r, out_r = run(REF_EXT, "bam_index", path, "index_path=" + theirs)
    rc_o, out_o = run(OUR_EXT, "bam_index", path, "index_path=" + ours)
    assert rc_r == 0 and rc_o == 0
    assert out_o.replace("ours.bai", "X") == out_r.replace("theirs.bai", "X")         # success, index_path, index_format
    assert_same_index(open(ours, "rb").read(), open(theirs, "rb").read())
    # default index path: PATH.bai
    rc_o, out_o = run(OUR_EXT, "bam_index", path)
    assert rc_o == 0 and os.path.exists(path + ".bai") and out_o.splitlines()[1].split("\t")[1] == path + ".bai"
    # the reference answers region queries (and its contig-parallel scan) from our index exactly as from its own
    for args in (("region=chr3:10,000,000-20,000,000",), ("region=chr2:1-5000000,chr2:4000000-9000000,chr7:49000000-",), ()):
        a = run(REF_EXT, "read_bam", path, "index_path=" + ours, *args, "--hash")
        b = run(REF_EXT, "read_bam", path, "index_path=" + theirs, *args, "--hash")
        assert a == b and a[0] == 0 and int(a[1].splitlines()[0].split("\t")[1]) > 100
    # the reference's own index reader on our file: per-contig mapped / unmapped counts, n_no_coor
    os.replace(ours, path + ".bai")
    a = run(REF_EXT, "read_hts_index", path)
    os.replace(theirs, path + ".bai")
    b = run(REF_EXT, "read_hts_index", path)
    assert a == b and a[0] == 0 and len(a[1].splitlines()) == 23
    # errors read like the reference's
    bad = str(tmp_path / "missing.bam")
    assert run(OUR_EXT, "bam_index", bad) == run(REF_EXT, "bam_index", bad)
    assert run(OUR_EXT, "bam_index", "") == run(REF_EXT, "bam_index", "")
