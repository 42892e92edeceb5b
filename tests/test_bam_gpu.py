"""Parity of the read_bam scan (BGZF inflate -> boundary walk -> columnar decode, through the C-ABI) against the
oracle's restatement of bam_read1 + bam_read_function.  Bit-exact: values, row order, NULLs, stop-on-error."""
import os
import struct

import numpy as np
import pytest

import duckhts_b200 as dh
from oracle import dh_oracle_py as orc
from tests import util
from tests.util import aux_tag, bam_header, bam_record

pytestmark = pytest.mark.gpu

TAGS = [(b"NM", "i"), (b"MD", "Z"), (b"AS", "i")]


def both(ctx, img, tags=TAGS, **kw):
    rc, infl, bad = orc.bgzf_inflate(img)
    assert rc == 0
    exp = orc.bam_decode(infl, tags)
    got = dh.bam_scan(ctx, img, tags=tags, **kw)
    util.assert_bam_equal(got, exp)
    assert got["ref_names"] == exp["ref_names"] and got["ref_lens"] == exp["ref_lens"] and got["text"] == exp["text"]
    return got, exp


def test_synthetic_single_end(ctx):
    got, exp = both(ctx, util.synth_bam_bytes(n_records=20000, seed=11))
    assert got["n"] == 20000 and got["status"] == 0


def test_synthetic_paired_small_segments_and_batches(ctx):
    img = util.synth_bam_bytes(n_records=60000, seed=12, paired=True)
    got, _ = both(ctx, img, seg_comp=1 << 16, batch_rows=777)
    assert got["n_batches"] > 50
    both(ctx, img, seg_comp=1 << 20, batch_rows=5000)
    both(ctx, img)


def test_records_straddling_blocks_and_segments(ctx):
    """gen_synth bgzf: the BAM byte stream cut into BGZF blocks at arbitrary byte positions (foreign-writer case,
    htslib/test/bgzf_boundaries): exercises speculative entries, the exit/entry chain check and the carry."""
    img = np.fromfile(util.synth_path("bgzf", n_records=30000, seed=13, paired=True), dtype=np.uint8)
    got, _ = both(ctx, img, seg_comp=1 << 16, batch_rows=4096)
    both(ctx, img)


def test_b_array_list_tags(ctx):
    """A15: CG / FZ / ML LIST columns (bam_reader.c:106-138, 954-962): every B sub-type, float arrays under a BIGINT list,
    empty arrays; children cross batch and segment boundaries; mixed with long-string tag slots in the same heap."""
    img = util.synth_bam_bytes(n_records=30000, seed=79, paired=True, extra=("--btags",))
    tags = [(b"ML", "Bi"), (b"MD", "Z"), (b"CG", "Bi"), (b"XZ", "Z"), (b"FZ", "Bi"), (b"NM", "i"), (b"XD", "Bf"), (b"XB", "Bi")]
    got, exp = both(ctx, img, tags=tags)
    assert sum(x is not None for x in got["tags"][0]["b"]) > 10000
    both(ctx, img, tags=tags, seg_comp=1 << 16, batch_rows=333)
    # projection: only list columns / a list column whose tag never occurs
    sub = dh.bam_scan(ctx, img, tags=[(b"ML", "Bi"), (b"ZZ", "Bf")], col_mask=(1 << 13) | (1 << 14) | 1)
    assert sub["n"] == exp["n"] and sub["qname"] == orc.strcol_list(exp["qname"]) and all(x is None for x in sub["tags"][1]["b"])
    assert all((g is None) == (e is None) and (e is None or np.array_equal(g, e)) for g, e in zip(sub["tags"][0]["b"], exp["tags"][0]["b"]))
    # hand-made: sub-type 'd' is no B sub-type (skip_aux accepts it, the getter returns 0), first of two equal tags wins
    recs = [util.bam_record(qname=b"q1", aux=b"MLBd" + struct.pack("<I", 2) + struct.pack("<dd", 1.5, -2.0) + b"MLBC" + struct.pack("<I", 1) + b"\x07"),
            util.bam_record(qname=b"q2", aux=b"XXZ" + b"y" * 30 + b"\0" + b"MLBs" + struct.pack("<Ihh", 2, -5, 300)),
            util.bam_record(qname=b"q3", aux=b"MLBI" + struct.pack("<I", 0))]
    img2 = np.frombuffer(util.make_bgzf(util.bam_header([(b"c1", 1000)]) + b"".join(recs)), dtype=np.uint8)
    got, exp = both(ctx, img2, tags=[(b"ML", "Bi"), (b"XX", "Z")])
    assert [None if x is None else x.tolist() for x in got["tags"][0]["b"]] == [[0, 0], [-5, 300], []]


def map_both(ctx, img, exclude, **kw):
    rc, infl, _ = orc.bgzf_inflate(img)
    assert rc == 0
    orc.set_aux_exclude(exclude)
    try:
        exp = orc.bam_decode(infl, [(b"\0\0", "M")])
    finally:
        orc.set_aux_exclude(())
    got = dh.bam_scan(ctx, img, tags=(), col_mask=1, aux_map=exclude, **kw)
    assert got["n"] == exp["n"] and got["status"] == exp["status"]
    e = exp["tags"][0]["m"]
    for r, (g, x) in enumerate(zip(got["aux_map"], e)):
        assert g == x, (r, g, x)
    return got, e


def test_auxiliary_tags_map(ctx):
    """A16: MAP(VARCHAR, VARCHAR) of every aux entry (bam_reader.c:967-1031): all aux types, B arrays, the standard_tags
    exclusion, NULL rows, long values in the heap next to the other columns' slots."""
    img = util.synth_bam_bytes(n_records=30000, seed=80, paired=True, extra=("--btags",))
    got, e = map_both(ctx, img, orc.STD_TAGS)
    assert any(x is None for x in e) and any(x and any(len(v) > 12 for _, v in x) for x in e)
    map_both(ctx, img, (), seg_comp=1 << 16, batch_rows=501)
    # with every other column projected too (shared heap)
    tags = [(b"ML", "Bi"), (b"MD", "Z"), (b"XZ", "Z")]
    full = dh.bam_scan(ctx, img, tags=tags, aux_map=orc.STD_TAGS, batch_rows=4000)
    assert full["aux_map"] == got["aux_map"]
    rc, infl, _ = orc.bgzf_inflate(img)
    util.assert_bam_equal(full, orc.bam_decode(infl, tags))
    # edge records: CG consumed by bam_tag2cigar, unknown aux type, value cut by the record end, unterminated Z
    stream = bam_header(REFS, TEXT) + b"".join(edge_records())
    for block in (0xff00, 333):
        map_both(ctx, util.make_bgzf(stream, block=block), (), batch_rows=5)
        map_both(ctx, util.make_bgzf(stream, block=block), orc.STD_TAGS, batch_rows=5)


def test_reference_sqllogictest_standard_and_auxiliary_tags(ctx):
    """test/sql/duckhts.test:180-185: RG, NM, map_extract(AUXILIARY_TAGS, 'XZ') of aux_tags, standard_tags + auxiliary_tags -> x1, 2, [foo]."""
    img = np.fromfile(os.path.join(util.GOLDEN, "aux_tags.bam"), dtype=np.uint8)
    got = dh.bam_scan(ctx, img, tags=[(b"RG", "Z"), (b"NM", "i")], aux_map=orc.STD_TAGS)
    assert got["tags"][0]["s"][0] == b"x1" and got["tags"][1]["i"][0] == 2
    assert [v for k, v in got["aux_map"][0] if k == b"XZ"] == [b"foo"]
    assert all(k.decode() not in orc.STD_TAGS for k, _ in got["aux_map"][0])


def test_auxiliary_tags_float_text_is_printf_g(ctx):
    """\"%g\" on the GPU (exact big-integer digit generation) against the C library's, bit pattern by bit pattern: random doubles
    and floats, powers of ten, round-half cases, subnormals, infinities and NaNs."""
    rng = np.random.default_rng(5)
    dbl = [0.0, -0.0, 1.0, 0.1, 0.5, 1e-5, 1e-4, 9.9999949999e-5, 0.000123456, 999999.5, 999999.4999, 1000005.0, 1000015.0, 100000.5,
           100001.5, 2.5e-5, 1e22, 1e23, 1.7976931348623157e308, 5e-324, 2.2250738585072014e-308, 123456.5, 123455.5, 0.3, 2.0 / 3.0,
           float("inf"), float("-inf"), float("nan"), 1e100, 1e-100, 299792458.0, 6.02214076e23, 1.5, 2.5, 1234565.0, 1234575.0]
    dbl += [10.0 ** k for k in range(-30, 31)] + [-(10.0 ** k) * 1.0000005 for k in range(-10, 11)]
    dbl += list(rng.integers(0, 1 << 64, 3000, dtype=np.uint64).view(np.float64))
    dbl += list((rng.random(2000) * 10.0 ** rng.integers(-8, 9, 2000)).astype(np.float64))
    flt = list(rng.integers(0, 1 << 32, 3000, dtype=np.uint64).astype(np.uint32).view(np.float32)) + [np.float32(x) for x in (3.14159274, 1e-45, 3.4028235e38, 1e-5, 1e-4, 16777216.0, 0.1)]
    recs = []
    for i in range(0, len(dbl), 25):
        aux = b"".join(b"D" + bytes([65 + k]) + b"d" + struct.pack("<d", v) for k, v in enumerate(dbl[i:i + 25]))
        recs.append(bam_record(b"d%d" % i, aux=aux))
    for i in range(0, len(flt), 25):
        aux = b"".join(b"F" + bytes([65 + k]) + b"f" + struct.pack("<f", v) for k, v in enumerate(flt[i:i + 25]))
        aux += b"FBBf" + struct.pack("<I", len(flt[i:i + 25])) + b"".join(struct.pack("<f", v) for v in flt[i:i + 25])
        recs.append(bam_record(b"f%d" % i, aux=aux))
    got, e = map_both(ctx, util.make_bgzf(bam_header(REFS, TEXT) + b"".join(recs)), ())
    assert (b"DA", b"0") in e[0] and (b"DB", b"-0") in e[0]


def test_unplaced_reads_are_returned_in_sequential_mode(ctx):
    img = util.synth_bam_bytes(n_records=3000, seed=14, paired=True, extra=("--unplaced", "50"))
    got, exp = both(ctx, img)
    assert got["rname"][-1] == b"*" and got["n"] == 3050


def test_projection_subsets(ctx):
    img = util.synth_bam_bytes(n_records=5000, seed=15, paired=True)
    rc, infl, _ = orc.bgzf_inflate(img)
    exp = orc.bam_decode(infl, TAGS)
    for mask in (0, 1 << 1, (1 << 3) | (1 << 9), (1 << 12), (1 << 11) | (1 << 14), 0xffff & ~(1 << 10)):
        got = dh.bam_scan(ctx, img, tags=TAGS, col_mask=mask)
        assert got["n"] == exp["n"]
        for c, name in enumerate(dh.bam.CORE_NAMES):
            if (mask >> c) & 1:
                if name in util.FIX_COLS:
                    assert np.array_equal(got[name], exp[name])
                else:
                    assert got[name] == orc.strcol_list(exp[name]), name
        for k in range(3):
            if (mask >> (13 + k)) & 1:
                assert np.array_equal(got["tags"][k]["valid"], exp["tags"][k]["valid"])


REFS = [(b"chr1", 1000000), (b"a_contig_name_longer_than_12", 5000), (b"c", 10)]
TEXT = (b"@HD\tVN:1.6\tSO:unsorted\n@SQ\tSN:chr1\tLN:1000000\n@SQ\tSN:a_contig_name_longer_than_12\tLN:5000\n@SQ\tSN:c\tLN:10\n"
        b"@RG\tID:g1\tSM:s1\n@RG\tID:nosm\tPL:x\n@RG\tID:a_long_read_group_identifier\tSM:a_sample_name_longer_than_12_bytes\n"
        b"@RG\tID:g1\tSM:shadowed\n@CO\tx\n")


def edge_records():
    r = []
    seq = "ACGTNACGTMRSVWYHKDB=" * 3
    r.append(bam_record(b"plain", 0, 0, 99, 60, [(60, "M")], seq=seq, qual=range(60), aux=aux_tag("RG", "Z", "g1") + aux_tag("NM", "C", 3)))
    r.append(bam_record(b"unmapped", 4, -1, -1, 0, [], seq="ACGT", qual=[1, 2, 3, 4]))
    r.append(bam_record(b"noseq", 0, 0, 5, 1, [], seq=""))
    r.append(bam_record(b"noqual", 0, 1, 5, 1, [(13, "M")], seq="ACGTACGTACGTA", aux=aux_tag("RG", "Z", "nosm")))
    r.append(bam_record(b"short12chars", 16, 2, 0, 7, [(12, "M")], mtid=1, mpos=41, tlen=-77, seq="ACGTACGTACGT", qual=[40] * 12,
                        aux=aux_tag("RG", "Z", "a_long_read_group_identifier")))
    r.append(bam_record(b"q" * 13, 0, 0, 1, 1, [(5, "S"), (1000, "M"), (3, "I"), (20, "D"), (992, "M")], seq="A" * 2000, qual=[30] * 2000,
                        aux=aux_tag("MD", "Z", "1000^" + "A" * 20 + "992") + aux_tag("AS", "s", -5) + aux_tag("RG", "Z", "unknown_rg")))
    r.append(bam_record(b"n" * 254, 0, 0, 1, 1, [(1, "M")], seq="C", qual=[0]))
    ops = [(1 + (i % 3), "MID"[i % 3]) for i in range(100)]
    qlen = sum(l for l, o in ops if o in "MI")
    r.append(bam_record(b"manyops", 0, 0, 1, 1, ops, seq="G" * qlen, qual=[2] * qlen, aux=aux_tag("NM", "i", -7) + aux_tag("NM", "i", 9)))
    r.append(bam_record(b"alltypes", 0, 0, 1, 1, [(4, "M")], seq="ACGT", qual=[9] * 4,
                        aux=aux_tag("XA", "A", "Q") + aux_tag("Xc", "c", -128) + aux_tag("XC", "C", 255) + aux_tag("Xs", "s", -32768) +
                        aux_tag("XS", "S", 65535) + aux_tag("Xi", "i", -2147483648) + aux_tag("XI", "I", 4294967295) + aux_tag("Xf", "f", 3.25) +
                        aux_tag("Xd", "d", -1e300) + aux_tag("XZ", "Z", "") + aux_tag("XH", "H", "DEADBEEF") + aux_tag("XB", "B", ("S", [1, 2, 3])) +
                        aux_tag("Xb", "B", ("f", [1.5])) + aux_tag("TS", "A", "+") + aux_tag("AS", "f", 2.5) + aux_tag("NM", "Z", "notint") +
                        aux_tag("MD", "i", 5) + aux_tag("RG", "Z", "g1")))
    r.append(bam_record(b"qual223", 0, 0, 1, 1, [(20, "M")], seq="A" * 20, qual=[30] * 15 + [223] + [30] * 4))
    r.append(bam_record(b"qual223early", 0, 0, 1, 1, [(20, "M")], seq="A" * 20, qual=[30] * 3 + [223] + [30] * 16))
    r.append(bam_record(b"qual223first", 0, 0, 1, 1, [(20, "M")], seq="A" * 20, qual=[223] + [30] * 19))
    r.append(bam_record(b"nul\0inside", 0, 0, 1, 1, [(1, "M")], seq="T", qual=[5]))
    r.append(bam_record(raw_qname=b"nonul", flag=0, tid=0, pos=1, mapq=1, cigar=[(1, "M")], seq="T", qual=[5]))
    # CG:B,I long-CIGAR convention (bam_tag2cigar, sam.c:675-730): placeholder kS mN + real CIGAR in the tag
    real = [(10, "M"), (2, "I"), (8, "M")]
    cg = aux_tag("CG", "B", ("I", [(l << 4) | util.CIG_OPS.index(o) for l, o in real]))
    r.append(bam_record(b"cgtag", 0, 0, 100, 1, [(20, "S"), (18, "N")], seq="A" * 20, qual=[1] * 20, aux=aux_tag("NM", "C", 1) + cg + aux_tag("RG", "Z", "g1")))
    r.append(bam_record(b"hugeread", 0, 0, 1, 1, [(150000, "M")], seq="ACGT" * 37500, qual=[20] * 150000, aux=aux_tag("MD", "Z", "150000")))
    r.append(bam_record(b"after_huge", 0, 0, 2, 2, [(3, "M")], seq="AAA", qual=[1, 2, 3], aux=aux_tag("RG", "H", "AB")))
    r.append(bam_record(b"badaux", 0, 0, 2, 2, [(3, "M")], seq="AAA", qual=[1, 2, 3], aux=aux_tag("NM", "C", 1) + b"ZZ?" + aux_tag("AS", "C", 9)))
    r.append(bam_record(b"unterminatedZ", 0, 0, 2, 2, [(3, "M")], seq="AAA", qual=[1, 2, 3], aux=aux_tag("NM", "C", 1) + b"MDZabc"))
    return r


EDGE_TAGS = [(b"NM", "i"), (b"MD", "Z"), (b"AS", "i"), (b"AS", "f"), (b"TS", "A"), (b"XA", "A"), (b"Xf", "f"), (b"Xd", "f"), (b"XZ", "Z"),
             (b"XH", "Z"), (b"XI", "i"), (b"Xi", "i"), (b"Xs", "i"), (b"XS", "i"), (b"Xc", "i"), (b"XC", "i"), (b"XB", "Z"), (b"XB", "i"),
             (b"XA", "i"), (b"ZZ", "Z"), (b"Xf", "i")]


def test_edge_case_records(ctx):
    stream = bam_header(REFS, TEXT) + b"".join(edge_records())
    for block in (0xff00, 4096, 333):
        img = util.make_bgzf(stream, block=block)
        got, exp = both(ctx, img, tags=EDGE_TAGS, seg_comp=1 << 16, batch_rows=5)
        assert got["n"] == len(edge_records()) and got["status"] == 0
    assert got["rname"][4] == b"c" and got["rnext"][4] == b"a_contig_name_longer_than_12"
    assert got["sample"][0] == b"s1" and got["sample"][3] is None and got["sample"][4] == b"a_sample_name_longer_than_12_bytes"
    assert got["cigar"][14] == b"10M2I8M" and got["qual"][9] == b"?" * 15 and got["qual"][11] == b""


@pytest.mark.parametrize("case", ["block_len_small", "cigar_qlen", "tid_range", "mtid_range", "l_qname0", "sizes", "trunc_core",
                                  "trunc_body", "trunc_len", "neg_lseq"])
def test_invalid_record_ends_scan_silently_with_same_rows_and_code(ctx, case):
    good = [bam_record(b"g%d" % i, 0, 0, i, 1, [(4, "M")], seq="ACGT", qual=[1] * 4) for i in range(300)]
    bad = bam_record(b"bad", 0, 0, 5, 1, [(4, "M")], seq="ACGT", qual=[1] * 4)
    tail = good[:7]
    if case == "block_len_small":
        bad = struct.pack("<I", 31) + bad[4:]
    elif case == "cigar_qlen":
        bad = bam_record(b"bad", 0, 0, 5, 1, [(5, "M")], seq="ACGT", qual=[1] * 4)
    elif case == "tid_range":
        bad = bam_record(b"bad", 0, 3, 5, 1, [(4, "M")], seq="ACGT", qual=[1] * 4)
    elif case == "mtid_range":
        bad = bam_record(b"bad", 0, 0, 5, 1, [(4, "M")], mtid=-2, seq="ACGT", qual=[1] * 4)
    elif case == "l_qname0":
        bad = bytearray(bad); bad[12] = 0; bad = bytes(bad)
    elif case == "sizes":
        bad = bytearray(bad); bad[20:24] = struct.pack("<i", 1000); bad = bytes(bad)
    elif case == "neg_lseq":
        bad = bytearray(bad); bad[20:24] = struct.pack("<i", -1); bad = bytes(bad)
    elif case == "trunc_core":
        bad, tail = bad[:20], []
    elif case == "trunc_body":
        bad, tail = bad[:-3], []
    elif case == "trunc_len":
        bad, tail = bad[:2], []
    stream = bam_header(REFS, TEXT) + b"".join(good) + bad + b"".join(tail)
    for block in (0xff00, 700):
        got, exp = both(ctx, util.make_bgzf(stream, block=block), seg_comp=1 << 16, batch_rows=64)
        assert got["n"] == 300 and got["status"] == exp["status"] < 0


def test_header_only_and_tiny_files(ctx):
    got, _ = both(ctx, util.make_bgzf(bam_header(REFS, TEXT)))
    assert got["n"] == 0 and got["status"] == 0
    got, _ = both(ctx, util.make_bgzf(bam_header([], b"")))
    assert got["n"] == 0
    got, _ = both(ctx, util.make_bgzf(bam_header(REFS, TEXT) + edge_records()[0]))
    assert got["n"] == 1
    # a header larger than several BGZF blocks
    refs = [(b"contig_%05d" % i, 1000 + i) for i in range(20000)]
    got, _ = both(ctx, util.make_bgzf(bam_header(refs, b"@CO\t" + b"x" * 200000 + b"\n") + bam_record(b"r", 0, 19999, 1, 1, [(1, "M")], seq="A", qual=[1])))
    assert got["n"] == 1 and got["rname"][0] == b"contig_19999"
    with pytest.raises(dh.DuckhtsError):
        dh.bam_scan(ctx, util.make_bgzf(b"NOTBAM" * 10))


def test_range_scans_concatenate_to_the_sequential_scan(ctx):
    """Offset-range sharding (the multi-GPU unit): ranges from dh_bam_split scanned independently and merged in
    order give exactly the sequential scan; each range's end_voffset is the next range's start."""
    img = util.synth_bam_bytes(n_records=80000, seed=16, paired=True)
    rc, infl, _ = orc.bgzf_inflate(img)
    exp = orc.bam_decode(infl, TAGS)
    for parts in (2, 3, 8):
        with dh.BamFile(ctx, img) as bf:
            voffs = bf.split(parts)
        assert voffs[0] == dh.BamFile(ctx, img).first_voffset and voffs[-1] == dh.VOFF_MAX
        qn, pos, total = [], [], 0
        for p in range(parts):
            if voffs[p] == dh.VOFF_MAX:
                continue
            got = dh.bam_scan(ctx, img, tags=TAGS, voff=(voffs[p], voffs[p + 1]), seg_comp=1 << 18)
            assert got["status"] == 0
            if voffs[p + 1] != dh.VOFF_MAX:
                assert got["stats"].end_voffset == voffs[p + 1]
            qn += got["qname"]
            pos.append(got["pos"])
            total += got["n"]
        assert total == exp["n"] and qn == orc.strcol_list(exp["qname"]) and np.array_equal(np.concatenate(pos), exp["pos"])


def test_large_scan_properties(ctx):
    """1M records: row count, per-column checksums vs the oracle, monotone row0, heap pointers inside the heap."""
    img = util.synth_bam_bytes(n_records=1000000, seed=6, paired=True)
    rc, infl, _ = orc.bgzf_inflate(img)
    exp = orc.bam_decode(infl, TAGS)
    with dh.BamFile(ctx, img) as bf:
        bf.set_projection(None, TAGS)
        bf.scan_begin()
        rows, sums = 0, {k: 0 for k in util.FIX_COLS}
        seq_heap_bytes = 0
        while True:
            b = bf.next_batch()
            if b is None:
                break
            assert b.row0 == rows
            cols = dh.batch_to_columns(bf, b) if rows == 0 else None
            n = int(b.n_rows)
            import ctypes as C
            for c, name in enumerate(dh.bam.CORE_NAMES):
                if name in util.FIX_COLS:
                    dt = np.dtype(dh.bam.CORE_DTYPES[name])
                    sums[name] += int(np.frombuffer(C.string_at(b.data[c], n * dt.itemsize), dtype=dt).astype(np.int64).sum())
            cells = np.frombuffer(C.string_at(b.data[9], 16 * n), dtype=np.uint8).reshape(n, 16)
            ptr = cells[:, 8:].copy().view(np.uint64).ravel()
            ln = cells[:, :4].copy().view(np.uint32).ravel()
            assert (ln == 150).all() and (ptr >= b.heap).all() and (ptr + 150 <= b.heap + b.heap_bytes).all()
            if cols is not None:
                k = min(n, 2000)
                assert cols["seq"][:k] == orc.strcol_list(exp["seq"])[:k]
            rows += n
        assert rows == exp["n"] == 1000000
        for k in util.FIX_COLS:
            assert sums[k] == int(exp[k].astype(np.int64).sum()), k
        assert bf.stats().status == 0


@pytest.mark.parametrize("name", util.golden_names())
def test_reference_golden_vectors(ctx, name):
    """The reference's own fixtures (test/data/range.bam, htslib/test/bgzf_boundaries*, auxf#values, ce#unmap*, c1#noseq,
    ce#large_seq ...): GPU scan vs the cells the reference's htslib produced (tests/golden/*.tsv)."""
    import hashlib
    import os
    img = np.fromfile(os.path.join(util.GOLDEN, name + ".bam"), dtype=np.uint8)
    out, st = dh.bgzf_inflate(ctx, img)
    assert hashlib.sha256(out.tobytes()).hexdigest() == open(os.path.join(util.GOLDEN, name + ".sha256")).read().strip()
    got = dh.bam_scan(ctx, img, tags=util.GOLDEN_TAGS, seg_comp=1 << 16, batch_rows=50)
    lines = util.dump_lines(got, util.GOLDEN_TAGS, False)
    exp = open(os.path.join(util.GOLDEN, name + ".tsv")).read().split("\n")[:-1]
    assert len(lines) == len(exp)
    for i, (a, b) in enumerate(zip(lines, exp)):
        assert a == b, f"{name} line {i}"


def test_cigar_metric_columns(ctx):
    """N3: the reference's CIGAR UDFs (src/kmer_udf.c:197-293, :1231-1238) as derived columns of the decode kernel (tag requests of
    type 'c'), against the restatement of those UDFs applied to the CIGAR column -- both layouts, clipped / padded / placeholder /
    zero-length / 'B' operations, CIGARs of thousands of operations, mixed with real tag columns."""
    from oracle import cigar_udf as cu

    metrics = [(b"S", "c"), (b"H", "c"), (b"l", "c"), (b"r", "c"), (b"q", "c"), (b"a", "c"), (b"R", "c"), (b"oS", "c"), (b"oi", "c"), (b"oN", "c")]
    tags = [(b"NM", "i")] + metrics + [(b"MD", "Z")]

    def check(img, **kw):
        rc, infl, _ = orc.bgzf_inflate(img)
        assert rc == 0
        exp = orc.bam_decode(infl, [(b"NM", "i"), (b"MD", "Z")])
        cig = orc.strcol_list(exp["cigar"])
        for compact in (False, True):
            got = dh.bam_scan(ctx, img, tags=tags, compact=compact, **kw)
            assert got["n"] == exp["n"] and got["cigar"] == cig
            assert np.array_equal(got["tags"][0]["i"][exp["tags"][0]["valid"]], exp["tags"][0]["i"][exp["tags"][0]["valid"]])
            assert got["tags"][-1]["s"] == orc.strcol_list(exp["tags"][1]["s"])
            for k, (tg, _) in enumerate(metrics):
                col = got["tags"][1 + k]
                want = [cu.cigar_has_op(c, chr(tg[1])) if tg[:1] == b"o" else cu.cigar_metric(c, tg.decode()) for c in cig]
                assert [bool(v) for v in col["valid"]] == [w is not None for w in want], (tg, compact)
                assert [int(x) for x, w in zip(col["i"], want) if w is not None] == [w for w in want if w is not None], (tg, compact)
        return exp

    check(util.synth_bam_bytes(n_records=20000, seed=31, paired=True), batch_rows=3000)
    for name in util.golden_names():
        check(np.fromfile(os.path.join(util.GOLDEN, name + ".bam"), dtype=np.uint8))
    ops = [[(5, "S"), (90, "M"), (5, "S")], [(5, "H"), (95, "M")], [(10, "M"), (2, "I"), (3, "D"), (7, "N"), (4, "P"), (8, "="), (9, "X")],
           [(3, "S")], [(0, "M"), (5, "M")], [(5, "M"), (0, "I"), (3, "S")], [(4, "M"), (3, "B"), (2, "M")], [],
           [(1, "M"), (1, "I")] * 1500, [(268435455, "N"), (1, "M")], [(2, "H"), (3, "S"), (10, "M"), (4, "S"), (1, "H")]]
    recs = []
    for k, cg in enumerate(ops):
        qlen = sum(l for l, o in cg if o in "MIS=X")
        recs.append(bam_record(qname=b"q%d" % k, cigar=cg, seq="A" * min(qlen, 4000) if cg else "ACGT", pos=100 + k))
    img = util.make_bgzf(bam_header([(b"chr1", 1 << 29)]) + b"".join(recs))
    exp = check(img)
    assert exp["n"] >= 8        # (records whose CIGAR and SEQ lengths disagree end the scan like bam_read1: kept out of the list above)
    with pytest.raises(dh.DuckhtsError):
        dh.bam_scan(ctx, img, tags=[(b"oQ", "c")])
    with pytest.raises(dh.DuckhtsError):
        dh.bam_scan(ctx, img, tags=[(b"z", "c")])
