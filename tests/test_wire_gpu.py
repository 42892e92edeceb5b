"""Round-2 C-ABI features against the oracle (through libduckhts_cuda.so on the GPU): the compact wire layout, per-column
checksums (device kernel = host threads = numpy restatement over the oracle's columns), explicit batch lifetimes driven by
several threads, predicate pushdown, handles cloned onto a second context (the multi-GPU unit, one GPU used twice), two
scans running concurrently in two contexts, and truncated files."""
import threading

import numpy as np
import pytest

import duckhts_b200 as dh
from oracle import dh_oracle_py as orc
from tests import util
from tests.test_bam_gpu import EDGE_TAGS, REFS, TAGS, TEXT, edge_records
from tests.util import bam_header, bam_record

pytestmark = pytest.mark.gpu


def oracle_of(img, tags):
    rc, infl, _ = orc.bgzf_inflate(img)
    assert rc == 0
    return orc.bam_decode(infl, tags)


# ---- compact wire layout ---------------------------------------------------------------------------------------------
def test_compact_layout_synthetic(ctx):
    img = util.synth_bam_bytes(n_records=60000, seed=12, paired=True)
    exp = oracle_of(img, TAGS)
    for kw in ({}, {"seg_comp": 1 << 16, "batch_rows": 777}, {"batch_rows": 5000}):
        got = dh.bam_scan(ctx, img, tags=TAGS, compact=True, **kw)
        util.assert_bam_equal(got, exp)
    # fewer bytes cross PCIe than in the DuckDB layout
    a = dh.bam_scan(ctx, img, tags=TAGS)["stats"].d2h_bytes
    b = dh.bam_scan(ctx, img, tags=TAGS, compact=True)["stats"].d2h_bytes
    assert b < 0.85 * a, (a, b)


def test_compact_layout_edge_records(ctx):
    stream = bam_header(REFS, TEXT) + b"".join(edge_records())
    tags = [t for t in EDGE_TAGS]
    for block in (0xff00, 4096, 333):
        img = util.make_bgzf(stream, block=block)
        exp = oracle_of(img, tags)
        got = dh.bam_scan(ctx, img, tags=tags, compact=True, seg_comp=1 << 16, batch_rows=5)
        util.assert_bam_equal(got, exp)
        assert got["n"] == len(edge_records())
    assert got["qual"][9] == b"?" * 15 and got["qual"][11] == b"" and got["qname"][12] == b"nul" and got["qname"][13] == b"nonul"


def test_compact_layout_foreign_blocks_btags_and_projections(ctx):
    img = np.fromfile(util.synth_path("bgzf", n_records=30000, seed=13, paired=True), dtype=np.uint8)
    util.assert_bam_equal(dh.bam_scan(ctx, img, tags=TAGS, compact=True, seg_comp=1 << 16, batch_rows=4096), oracle_of(img, TAGS))
    img = util.synth_bam_bytes(n_records=30000, seed=79, paired=True, extra=("--btags",))
    tags = [(b"ML", "Bi"), (b"MD", "Z"), (b"CG", "Bi"), (b"XZ", "Z"), (b"FZ", "Bi"), (b"NM", "i"), (b"XD", "Bf"), (b"XB", "Bi"), (b"RG", "Z"), (b"TS", "A")]
    exp = oracle_of(img, tags)
    util.assert_bam_equal(dh.bam_scan(ctx, img, tags=tags, compact=True), exp)
    util.assert_bam_equal(dh.bam_scan(ctx, img, tags=tags, compact=True, seg_comp=1 << 16, batch_rows=333), exp)
    img = util.synth_bam_bytes(n_records=5000, seed=15, paired=True)
    exp = oracle_of(img, TAGS)
    for mask in (0, 1 << 1, (1 << 3) | (1 << 9), (1 << 12), (1 << 11) | (1 << 14), 0xffff & ~(1 << 10), 1 | (1 << 10), (1 << 5) | (1 << 2) | (1 << 6)):
        got = dh.bam_scan(ctx, img, tags=TAGS, col_mask=mask, compact=True)
        assert got["n"] == exp["n"]
        for c, name in enumerate(dh.bam.CORE_NAMES):
            if (mask >> c) & 1:
                if name in util.FIX_COLS:
                    assert np.array_equal(got[name], exp[name]), name
                else:
                    assert got[name] == orc.strcol_list(exp[name]), name
        for k in range(3):
            if (mask >> (13 + k)) & 1:
                assert np.array_equal(got["tags"][k]["valid"], exp["tags"][k]["valid"])


@pytest.mark.parametrize("name", util.golden_names())
def test_compact_layout_reference_golden_vectors(ctx, name):
    import os
    img = np.fromfile(os.path.join(util.GOLDEN, name + ".bam"), dtype=np.uint8)
    got = dh.bam_scan(ctx, img, tags=util.GOLDEN_TAGS, seg_comp=1 << 16, batch_rows=50, compact=True)
    lines = util.dump_lines(got, util.GOLDEN_TAGS, False)
    exp = open(os.path.join(util.GOLDEN, name + ".tsv")).read().split("\n")[:-1]
    assert lines == exp


def test_compact_layout_refuses_the_map_column(ctx):
    img = util.synth_bam_bytes(n_records=3000, seed=7)
    with dh.BamFile(ctx, img) as bf:
        bf.set_projection(1, ())
        bf.set_aux_map(True, ())
        bf.set_mode(compact=True)
        with pytest.raises(dh.DuckhtsError):
            bf.scan_begin()


# ---- checksums -----------------------------------------------------------------------------------------------------------
def scan_sums(bf, **mode):
    bf.set_mode(**mode)
    bf.scan_begin()
    rows, sums = bf.drain_checksums()
    return rows, [int(x) for x in sums]


def test_checksums_agree_across_layouts_sides_and_the_oracle(ctx):
    tags = [(b"NM", "i"), (b"MD", "Z"), (b"AS", "i"), (b"RG", "Z"), (b"XA", "A"), (b"Xf", "f")]
    img = util.synth_bam_bytes(n_records=120000, seed=21, paired=True)
    exp = oracle_of(img, tags)
    want = util.oracle_checksums(exp, tags)
    with dh.BamFile(ctx, img) as bf:
        bf.set_projection(None, tags)
        bf.configure(0, 30000, 0)
        for mode in ({}, {"compact": True}, {"device_output": True}, {"device_output": True, "compact": True}, {"device_input": True, "device_output": True, "compact": True}):
            rows, sums = scan_sums(bf, **mode)
            assert rows == exp["n"]
            assert sums[:13 + len(tags)] == want, mode
    # edge records: long contig / sample names (dictionary strings the device cannot follow), B lists, NULLs, truncated strings
    stream = bam_header(REFS, TEXT) + b"".join(edge_records())
    etags = list(EDGE_TAGS) + [(b"XB", "Bi"), (b"Xb", "Bf")]
    img = util.make_bgzf(stream, block=4096)
    exp = oracle_of(img, etags)
    want = util.oracle_checksums(exp, etags)
    with dh.BamFile(ctx, img) as bf:
        bf.set_projection(None, etags)
        bf.configure(1 << 16, 7, 0)
        for mode in ({}, {"compact": True}, {"device_output": True}, {"device_output": True, "compact": True}):
            rows, sums = scan_sums(bf, **mode)
            assert rows == exp["n"] and sums[:13 + len(etags)] == want, mode


def test_checksums_of_a_tiled_image_are_tiles_times_the_unit(ctx):
    """The size-independent property bench.py asserts at 100 M records: record blocks repeated k times => every column's
    checksum is k x the unit's (mod 2^64), whatever the batch and segment cuts."""
    img = util.synth_bam_bytes(n_records=40000, seed=23, paired=True)
    blocks, end = dh.bgzf_walk(img)
    with dh.BamFile(ctx, img) as bf:
        first = bf.first_voffset >> 16
        bf.set_projection(None, TAGS)
        rows1, s1 = scan_sums(bf, device_input=True, device_output=True)
    hb = next(i for i, b in enumerate(blocks) if b[0] == first)
    body = img[blocks[hb][0]:blocks[-1][0]]
    k = 5
    tiled = np.concatenate([img[:blocks[hb][0]]] + [body] * k + [np.frombuffer(util.BGZF_EOF, dtype=np.uint8)])
    with dh.BamFile(ctx, tiled) as bf:
        bf.set_projection(None, TAGS)
        bf.configure(1 << 20, 50000, 0)
        for mode in ({"device_input": True, "device_output": True}, {"compact": True}):
            rows, s = scan_sums(bf, **mode)
            assert rows == k * rows1
            assert s == [(x * k) & 0xFFFFFFFFFFFFFFFF for x in s1], mode


# ---- explicit batch lifetimes, several consumer threads ---------------------------------------------------------------------
def test_acquire_release_from_several_threads(ctx):
    img = util.synth_bam_bytes(n_records=120000, seed=21, paired=True)
    exp = oracle_of(img, TAGS)
    want = util.oracle_checksums(exp, TAGS)
    for compact in (False, True):
        with dh.BamFile(ctx, img) as bf:
            bf.set_projection(None, TAGS)
            bf.configure(1 << 20, 4096, 0)
            bf.set_slots(8)
            bf.set_mode(compact=compact)
            bf.scan_begin()
            bf.prefetch()
            lock = threading.Lock()
            sums = np.zeros(13 + 63, dtype=np.uint64)
            rows, seen, errors = [0], [], []

            def worker():
                try:
                    held = []
                    while True:
                        b = bf.acquire()
                        if b is None:
                            break
                        held.append(b)
                        if len(held) == 2:          # keep two batches alive at once, give the older one back
                            old = held.pop(0)
                            mine = bf.checksum(old)
                            with lock:
                                sums[:] += mine
                                rows[0] += int(old.n_rows)
                                seen.append(int(old.row0))
                            bf.release(old.slot)
                    for old in held:
                        mine = bf.checksum(old)
                        with lock:
                            sums[:] += mine
                            rows[0] += int(old.n_rows)
                            seen.append(int(old.row0))
                        bf.release(old.slot)
                except Exception as e:  # noqa: BLE001
                    errors.append(e)

            th = [threading.Thread(target=worker) for _ in range(3)]
            for t in th:
                t.start()
            for t in th:
                t.join()
            assert not errors, errors
            assert rows[0] == exp["n"] and len(set(seen)) == len(seen)
            assert [int(x) for x in sums[:16]] == want


def test_all_slots_held_is_reported_not_deadlocked(ctx):
    img = util.synth_bam_bytes(n_records=60000, seed=12, paired=True)
    with dh.BamFile(ctx, img) as bf:
        bf.set_projection(1, ())
        bf.configure(1 << 20, 2000, 0)
        bf.set_slots(3)
        bf.scan_begin()
        held = [bf.acquire() for _ in range(3)]
        assert all(h is not None for h in held)
        rc = dh.lib().dh_bam_acquire_batch(bf.h, __import__("ctypes").byref(dh.bam.BamBatch()))
        assert rc == -8                                   # DH_EBUSY
        bf.release(held[0].slot)
        assert bf.acquire() is not None
        with pytest.raises(dh.DuckhtsError):
            bf.release(17)


# ---- N3: predicate pushdown -----------------------------------------------------------------------------------------------
@pytest.mark.parametrize("filt", [(0, 0x4, 30), (0x40, 0x10, 0), (0x1, 0, 59), (0, 0, 61)])
def test_filter_pushdown_returns_the_rows_a_filter_over_the_full_scan_keeps(ctx, filt):
    img = util.synth_bam_bytes(n_records=60000, seed=12, paired=True)
    exp = oracle_of(img, TAGS)
    req, exc, mq = filt
    keep = ((exp["flag"] & req) == req) & ((exp["flag"] & exc) == 0) & (exp["mapq"] >= mq)
    for compact in (False, True):
        got = dh.bam_scan(ctx, img, tags=TAGS, filt=filt, compact=compact, batch_rows=9000)
        assert got["n"] == int(keep.sum())
        assert np.array_equal(got["pos"], exp["pos"][keep]) and np.array_equal(got["flag"], exp["flag"][keep])
        qn = orc.strcol_list(exp["qname"])
        assert got["qname"] == [q for q, k in zip(qn, keep) if k]
        assert np.array_equal(got["tags"][0]["i"], exp["tags"][0]["i"][keep])
    # fewer rows, fewer bytes over PCIe
    if 0 < keep.sum() < exp["n"]:
        full = dh.bam_scan(ctx, img, tags=TAGS)["stats"].d2h_bytes
        assert got["stats"].d2h_bytes < full


# ---- one image, several pipelines: clone + ranges / contig groups --------------------------------------------------------------
def test_cloned_handles_scan_ranges_that_merge_to_the_sequential_scan(ctx):
    """The multi-GPU unit with one GPU used twice: two contexts, two handles on one pinned image, each scanning its part
    of dh_bam_split / dh_bam_split_contigs concurrently; merged in range order = the sequential scan; checksums add up."""
    path = util.synth_path("bam", n_records=80000, seed=16, paired=True, index=True)
    img = np.fromfile(path, dtype=np.uint8)
    bai = np.fromfile(path + ".bai", dtype=np.uint8)
    exp = oracle_of(img, TAGS)
    want = util.oracle_checksums(exp, TAGS)
    ctx2 = dh.Context(0)
    try:
        with dh.BamFile(ctx, img) as a:
            a.set_projection(None, TAGS)
            b = a.clone(ctx2)
            b.set_projection(None, TAGS)
            assert b.ref_names == a.ref_names and b.first_voffset == a.first_voffset and b.rg == a.rg
            voffs = a.split(2)
            res = [None, None]

            def run(i, bf):
                bf.set_mode(compact=bool(i))
                bf.configure(1 << 20, 20000, 0)
                bf.scan_begin(voffs[i], voffs[i + 1])
                parts, sums = [], np.zeros(13 + 63, dtype=np.uint64)
                while True:
                    bt = bf.next_batch()
                    if bt is None:
                        break
                    parts.append(dh.batch_to_columns(bf, bt))
                    bf.checksum(bt, sums)
                res[i] = (parts, sums)

            th = [threading.Thread(target=run, args=(i, bf)) for i, bf in enumerate((a, b))]
            for t in th:
                t.start()
            for t in th:
                t.join()
            qn = [q for parts, _ in res for p in parts for q in p["qname"]]
            pos = np.concatenate([p["pos"] for parts, _ in res for p in parts])
            assert qn == orc.strcol_list(exp["qname"]) and np.array_equal(pos, exp["pos"])
            total = (res[0][1] + res[1][1])[:16]
            assert [int(x) for x in total] == want
            # contig groups from the index (BASELINE configs[3]): the product's partition, both parts, merged
            with dh.Index(ctx, bai) as ix:
                bounds = a.split_contigs(ix, 2)
                v = (__import__("ctypes").c_int32 * 3)()
                assert dh.lib().dh_index_split_contigs(ix.h, 2, v) == 0 and list(v) == bounds
                rows = 0
                sums = np.zeros(13 + 63, dtype=np.uint64)
                for i, bf in enumerate((a, b)):
                    bf.scan_contigs(ix, bounds[i], bounds[i + 1])
                    r, s = bf.drain_checksums()
                    rows += r
                    sums += s
                a.scan_contigs(ix)
                r_all, s_all = a.drain_checksums()
                assert rows == r_all == exp["n"] and np.array_equal(sums, s_all)
            b.close()
    finally:
        ctx2.close()


def test_two_scans_in_two_contexts_run_concurrently(ctx):
    """scans that may run at the same time need a context each (streams, ticket counters, staging slots);
    two threads, two contexts on one device, many small segments each: both results exact."""
    imgs = [util.synth_bam_bytes(n_records=60000, seed=12, paired=True), util.synth_bam_bytes(n_records=80000, seed=16, paired=True)]
    exps = [oracle_of(i, TAGS) for i in imgs]
    ctx2 = dh.Context(0)
    out, errors = [None, None], []

    def run(i, c):
        try:
            for _ in range(3):
                out[i] = dh.bam_scan(c, imgs[i], tags=TAGS, seg_comp=1 << 18, batch_rows=3000, compact=bool(i))
        except Exception as e:  # noqa: BLE001
            errors.append(e)

    try:
        th = [threading.Thread(target=run, args=(i, c)) for i, c in enumerate((ctx, ctx2))]
        for t in th:
            t.start()
        for t in th:
            t.join()
        assert not errors, errors
        for got, exp in zip(out, exps):
            util.assert_bam_equal(got, exp)
    finally:
        ctx2.close()


# ---- truncated files: what lies in front of the broken block, then the read error -------------------------------------------
def test_truncated_bam_returns_the_rows_in_front_of_the_cut(ctx):
    img = util.synth_bam_bytes(n_records=20000, seed=11)
    blocks, _ = dh.bgzf_walk(img)
    cut = blocks[len(blocks) // 2]
    rc, infl, _ = orc.bgzf_inflate(img[:cut[0]])
    exp = orc.bam_decode(infl, TAGS)
    for tail in (img[cut[0]:cut[0] + cut[1] // 2], np.frombuffer(b"garbage, not a BGZF block at all......", dtype=np.uint8)):
        bad = np.concatenate([img[:cut[0]], tail])
        got = dh.bam_scan(ctx, bad, tags=TAGS)
        assert got["n"] == exp["n"] and got["status"] < 0
        assert got["qname"] == orc.strcol_list(exp["qname"]) and np.array_equal(got["pos"], exp["pos"])
    # cut exactly at a block boundary, no EOF marker: a clean end to htslib too (it only warns)
    got = dh.bam_scan(ctx, img[:cut[0]], tags=TAGS)
    assert got["n"] == exp["n"] and got["status"] == 0
    with pytest.raises(dh.DuckhtsError):
        dh.bam_scan(ctx, np.frombuffer(b"not a BGZF file", dtype=np.uint8))


def test_long_record_at_a_range_end_is_finished_from_the_look_ahead(ctx):
    """a range scan whose last record is longer than two BGZF blocks (a 150 kb read) used to fail."""
    recs = [bam_record(b"r%d" % i, 0, 0, i, 1, [(4, "M")], seq="ACGT", qual=[1] * 4) for i in range(3000)]
    huge = bam_record(b"hugeread", 0, 0, 5000, 1, [(300000, "M")], seq="ACGT" * 75000, qual=[20] * 300000)
    stream = bam_header(REFS, TEXT) + b"".join(recs) + huge + b"".join(recs[:50])
    img = np.frombuffer(util.make_bgzf(stream, block=0xff00), dtype=np.uint8)
    exp = oracle_of(img, TAGS)
    blocks, _ = dh.bgzf_walk(img)
    # end the first range in the block where the huge record starts + 1: the record straddles the range end by several blocks
    with dh.BamFile(ctx, img) as bf:
        first = bf.first_voffset
    for cut_blk in range(2, len(blocks) - 1):
        lo = dh.bam_scan(ctx, img, tags=TAGS, voff=(first, blocks[cut_blk][0] << 16))
        hi = dh.bam_scan(ctx, img, tags=TAGS, voff=(lo["stats"].end_voffset, dh.VOFF_MAX))
        assert lo["status"] == 0 and hi["status"] == 0
        assert lo["n"] + hi["n"] == exp["n"], cut_blk
        assert lo["qname"] + hi["qname"] == orc.strcol_list(exp["qname"])


# ---- streamed files (dh_bam_open_file): same rows as the memory image, whatever the segment size ---------------------------------
def test_streamed_file_scans_match_the_memory_image(ctx, tmp_path):
    path = util.synth_path("bam", n_records=120000, seed=21, paired=True, index=True)
    img = np.fromfile(path, dtype=np.uint8)
    exp = oracle_of(img, TAGS)
    for kw in ({}, {"seg_comp": 1 << 16, "batch_rows": 777}, {"seg_comp": 1 << 20, "batch_rows": 50000, "compact": True}, {"seg_comp": 3 << 20}):
        got = dh.bam_scan(ctx, path, tags=TAGS, **kw)
        util.assert_bam_equal(got, exp)
        assert got["ref_names"] == exp["ref_names"] and got["text"] == exp["text"]
    # ranges of dh_bam_split on the file handle (block boundaries found by header search + record check), merged in order
    with dh.BamFile.open_file(ctx, path) as bf, dh.BamFile(ctx, img) as mem:
        assert bf.first_voffset == mem.first_voffset
        for parts in (2, 5):
            voffs = bf.split(parts)
            assert voffs[0] == mem.first_voffset and voffs[-1] == dh.VOFF_MAX
            assert all(a <= b for a, b in zip(voffs, voffs[1:]))
            qn, total = [], 0
            for p in range(parts):
                if voffs[p] == dh.VOFF_MAX:
                    continue
                got = dh.bam_scan(ctx, path, tags=TAGS, voff=(voffs[p], voffs[p + 1]), seg_comp=1 << 18, compact=bool(p & 1))
                assert got["status"] == 0
                if voffs[p + 1] != dh.VOFF_MAX:
                    assert got["stats"].end_voffset == voffs[p + 1]
                qn += got["qname"]
                total += got["n"]
            assert total == exp["n"] and qn == orc.strcol_list(exp["qname"])
    # index-driven scans on the file handle: regions and the contig mode
    bai = np.fromfile(path + ".bai", dtype=np.uint8)
    for regions in (["chr3:10,000,000-20,000,000"], ["chr2:1-5000000", "chr7:49000000-"]):
        a = dh.bam_scan(ctx, path, tags=TAGS, index=bai, regions=regions, seg_comp=1 << 18)
        m = dh.bam_scan(ctx, img, tags=TAGS, index=bai, regions=regions)
        assert a["n"] == m["n"] > 0 and a["qname"] == m["qname"] and np.array_equal(a["pos"], m["pos"])
    a = dh.bam_scan(ctx, path, tags=TAGS, index=bai, contigs=True, seg_comp=1 << 19)
    m = dh.bam_scan(ctx, img, tags=TAGS, index=bai, contigs=True)
    assert a["n"] == m["n"] == exp["n"] and a["qname"] == m["qname"]
    # edge records, a header of several blocks, tiny files, records straddling blocks: written to disk, streamed back
    stream = bam_header(REFS, TEXT) + b"".join(edge_records())
    refs = [(b"contig_%05d" % i, 1000 + i) for i in range(20000)]
    big_header = bam_header(refs, b"@CO\t" + b"x" * 200000 + b"\n") + bam_record(b"r", 0, 19999, 1, 1, [(1, "M")], seq="A", qual=[1])
    for i, data in enumerate((util.make_bgzf(stream, block=4096), util.make_bgzf(stream, block=333), util.make_bgzf(big_header), util.make_bgzf(bam_header(REFS, TEXT)),
                              util.make_bgzf(bam_header([], b"")))):
        p = str(tmp_path / f"f{i}.bam")
        open(p, "wb").write(data)
        e = oracle_of(np.frombuffer(data, dtype=np.uint8), EDGE_TAGS)
        util.assert_bam_equal(dh.bam_scan(ctx, p, tags=EDGE_TAGS, seg_comp=1 << 16, batch_rows=5), e)
    # truncated on disk
    blocks, _ = dh.bgzf_walk(img)
    cut = blocks[len(blocks) // 2]
    p = str(tmp_path / "cut.bam")
    open(p, "wb").write(img[:cut[0] + cut[1] // 2].tobytes())
    e = oracle_of(img[:cut[0]], TAGS)
    got = dh.bam_scan(ctx, p, tags=TAGS, seg_comp=1 << 20)
    assert got["n"] == e["n"] and got["status"] < 0 and got["qname"] == orc.strcol_list(e["qname"])
    with pytest.raises(dh.DuckhtsError):
        open(str(tmp_path / "x.bam"), "wb").write(b"not a BGZF file, not at all")
        dh.bam_scan(ctx, str(tmp_path / "x.bam"))
    with pytest.raises(dh.DuckhtsError):
        dh.bam_scan(ctx, str(tmp_path / "missing.bam"))
