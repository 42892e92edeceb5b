"""The DuckDB extension (duckhts_b200/host/duckhts_ext.c) against the REFERENCE extension, both driven by the same
minimal DuckDB host (harness/fake_duckdb.c): same table function, same arguments, output compared cell for cell.
The reference extension is the unchanged reference source compiled by oracle/Makefile (oracle/_ref/libduckhts_ref.so)."""
import os
import subprocess

import pytest

from tests import util

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "harness", "_bin", "fake_duckdb")
REF_EXT = os.path.join(ROOT, "oracle", "_ref", "libduckhts_ref.so")
OUR_EXT = os.path.join(ROOT, "duckhts_b200", "lib", "duckhts.duckdb_extension.so")
GOLD = os.path.join(ROOT, "tests", "golden")

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not (os.path.exists(HOST) and os.path.exists(REF_EXT)), reason="prebuilt fake host / reference extension absent")]

# SAMtags LIST columns (CG, FZ, ML; B-array tags, A15)
B_COLS = (13 + 7, 13 + 19, 13 + 29)


def run(ext, func, path, *args, env=None):
    e = dict(os.environ)
    e.update(env or {})
    p = subprocess.run([HOST, ext, func, path, *args], stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=900, env=e)
    return p.returncode, p.stdout.decode("utf-8", "replace"), p.stderr.decode("utf-8", "replace")


def multiset(out):
    """rows + the order-insensitive per-column hashes of a --hash run (third field of every column line)."""
    lines = out.splitlines()
    return [lines[0]] + [ln.split("\t")[0] + " " + ln.split("\t")[2] for ln in lines[1:]]


def both(func, path, *args):
    r = run(REF_EXT, func, path, *args)
    o = run(OUR_EXT, func, path, *args)
    return r, o


@pytest.mark.parametrize("name", util.golden_names())
def test_read_bam_matches_reference_extension_on_reference_fixtures(name):
    path = os.path.join(GOLD, name + ".bam")
    (rc_r, out_r, _), (rc_o, out_o, err_o) = both("read_bam", path)
    assert rc_r == 0 and rc_o == 0, err_o
    assert out_o == out_r


@pytest.mark.parametrize("name", ["aux_tags", "auxf_values", "range", "rg", "ce_unmap2"])
def test_standard_tags_columns(name):
    path = os.path.join(GOLD, name + ".bam")
    cols = ",".join(str(c) for c in range(13 + 56))
    (rc_r, out_r, _), (rc_o, out_o, err_o) = both("read_bam", path, "standard_tags=true", "--cols", cols)
    assert rc_r == 0 and rc_o == 0, err_o
    assert out_o == out_r
    # schema (header line) with every column, LIST columns included
    (_, full_r, _), (_, full_o, _) = both("read_bam", path, "standard_tags=true")
    assert full_o.splitlines()[0] == full_r.splitlines()[0]


def test_b_array_list_columns(tmp_path):
    # B arrays of every sub-type under CG / FZ / ML, float arrays under the BIGINT lists, empty arrays; small batches
    # so that list children cross batch and vector boundaries
    path = util.synth_path("bam", n_records=20000, seed=77, paired=True, extra=("--btags",))
    allc = ",".join(str(c) for c in range(13 + 56))
    for cols in (",".join(str(c) for c in B_COLS), "42,0,20", allc):
        (rc_r, out_r, _), (rc_o, out_o, err_o) = both("read_bam", path, "standard_tags=true", "--cols", cols)
        assert rc_r == 0 and rc_o == 0, err_o
        assert out_o == out_r, cols
    assert "[" in out_r


@pytest.mark.parametrize("std", [True, False])
def test_auxiliary_tags_map_column(std):
    # A16: MAP(VARCHAR, VARCHAR) column, alone and next to every other column; schema line included
    extra = ("standard_tags=true",) if std else ()
    paths = [util.synth_path("bam", n_records=20000, seed=77, paired=True, extra=("--btags",))] + [os.path.join(GOLD, n + ".bam") for n in ("auxf_values", "aux_tags", "range", "ce_unmap2")]
    for path in paths:
        for cols in ((str(69 if std else 13),), ()):
            args = ("read_bam", path, "auxiliary_tags=true") + extra + (("--cols", cols[0]) if cols else ())
            (rc_r, out_r, _), (rc_o, out_o, err_o) = both(*args)
            assert rc_r == 0 and rc_o == 0, err_o
            assert out_o == out_r, (path, cols)


@pytest.mark.parametrize("cols", ["0", "1,3,4", "9,10", "12,11,2", "5,0,8,6", "3"])
def test_projection_pushdown(cols):
    path = os.path.join(GOLD, "range.bam")
    (rc_r, out_r, _), (rc_o, out_o, err_o) = both("read_bam", path, "--cols", cols)
    assert rc_r == 0 and rc_o == 0, err_o
    assert out_o == out_r


def test_synthetic_paired_file_hashes(tmp_path):
    path = util.synth_path("bam", n_records=60000, seed=31, paired=True)
    cols = ",".join(str(c) for c in range(13 + 56))
    (rc_r, out_r, _), (rc_o, out_o, err_o) = both("read_bam", path, "standard_tags=true", "--cols", cols, "--hash")
    assert rc_r == 0 and rc_o == 0, err_o
    assert out_r.splitlines()[0].startswith("rows\t")
    assert out_o == out_r


def test_bind_errors_read_like_the_reference(tmp_path):
    for args in (("read_bam", str(tmp_path / "missing.bam")), ("read_bam", "")):
        (rc_r, out_r, _), (rc_o, out_o, _) = both(*args)
        assert rc_r == 1 and rc_o == 1
        assert out_o == out_r
    (rc_r, out_r, _), (rc_o, out_o, _) = both("read_bam", os.path.join(GOLD, "range.bam"), "no_such_parameter=1")
    assert out_o == out_r


def test_bgunzip_writes_identical_bytes(tmp_path):
    src = os.path.join(GOLD, "range.bam")
    outs = []
    for tag, ext in (("ref", REF_EXT), ("our", OUR_EXT)):
        out = str(tmp_path / f"{tag}.bin")
        rc, text, err = run(ext, "bgunzip", src, f"output_path={out}")
        assert rc == 0, err
        rows = text.splitlines()
        assert rows[0] == "success\toutput_path\tbytes_in\tbytes_out"
        outs.append((open(out, "rb").read(), rows[1].split("\t")))
    assert outs[0][0] == outs[1][0]
    assert outs[0][1][0] == outs[1][1][0] == "true"
    assert outs[0][1][2:] == outs[1][1][2:]          # bytes_in, bytes_out
    # refuses to overwrite, with the reference's message
    msgs = []
    for tag, ext in (("ref", REF_EXT), ("our", OUR_EXT)):
        rc, text, _ = run(ext, "bgunzip", src, f"output_path={tmp_path / (tag + '.bin')}")
        assert rc == 1
        msgs.append(text.replace(tag + ".bin", "X.bin"))
    assert msgs[0] == msgs[1]


# ---- A8: region queries and the index-driven contig scan, cell for cell against the reference extension -------------
@pytest.mark.parametrize("region", ["CHROMOSOME_I", "CHROMOSOME_I:1-1000", "CHROMOSOME_I:1-1000,CHROMOSOME_I:1-1000",
                                    "CHROMOSOME_II:2000-2500,CHROMOSOME_IV:1000-1100,CHROMOSOME_II:1-1200", "CHROMOSOME_V",
                                    "nosuchcontig,CHROMOSOME_IV:-1500", "CHROMOSOME_II:1136-1136"])
def test_region_queries_match_reference_extension(region):
    path = os.path.join(GOLD, "range.bam")       # range.bam.bai sits next to it (reference test/data)
    (rc_r, out_r, _), (rc_o, out_o, err_o) = both("read_bam", path, "region=" + region)
    assert rc_r == 0 and rc_o == 0, err_o
    assert out_o == out_r


def test_region_errors_read_like_the_reference():
    # no region names a known contig -> init error; region on a file without index -> init error
    for path, region in ((os.path.join(GOLD, "range.bam"), "nosuchcontig"), (os.path.join(GOLD, "rg.bam"), "xx:1-10")):
        (rc_r, out_r, _), (rc_o, out_o, _) = both("read_bam", path, "region=" + region)
        assert rc_r == 1 and rc_o == 1
        assert out_o == out_r


def test_indexed_synthetic_file_regions_and_contig_mode():
    path = util.synth_path("bam", n_records=30000, seed=31, paired=True, extra=("--unplaced", "25"), index=True)
    cols = ",".join(str(c) for c in range(13 + 56))
    for args in ((), ("region=chr3:10,000,000-20,000,000",), ("region=chr2:1-5000000,chr2:4000000-9000000,chr7:49000000-",),
                 ("region=chr22:-1000000,chr5:25m-26m",)):
        (rc_r, out_r, _), (rc_o, out_o, err_o) = both("read_bam", path, "standard_tags=true", "--cols", cols, "--hash", *args)
        assert rc_r == 0 and rc_o == 0, err_o
        assert out_o == out_r, args
    # explicit index_path (a copy under another name)
    import shutil
    alt = path + ".alt_index"
    shutil.copyfile(path + ".bai", alt)
    (rc_r, out_r, _), (rc_o, out_o, err_o) = both("read_bam", path, "index_path=" + alt, "region=chr9", "--hash")
    assert rc_r == 0 and rc_o == 0, err_o
    assert out_o == out_r


# ---- B1-B10: read_bcf, the two extensions under the same host ------------------------------------------------------
GEN_BCF = os.path.join(ROOT, "harness", "_bin", "gen_bcf")


@pytest.mark.parametrize("name", ["vcf_file", "edge_small", "plain_small"])
@pytest.mark.parametrize("args", [(), ("tidy_format=true",)])
def test_read_bcf_matches_reference_extension(name, args):
    path = os.path.join(GOLD, name + ".bcf")
    (rc_r, out_r, _), (rc_o, out_o, err_o) = both("read_bcf", path, *args)
    assert rc_r == 0 and rc_o == 0, err_o
    assert out_o == out_r


@pytest.mark.parametrize("cols", ["0,1", "16,17", "4,6,2", "18,20,22", "7,9,11,13", "19"])
def test_read_bcf_projection_pushdown(cols):
    path = os.path.join(GOLD, "edge_small.bcf")
    (rc_r, out_r, _), (rc_o, out_o, err_o) = both("read_bcf", path, "tidy_format=true", "--cols", cols)
    assert rc_r == 0 and rc_o == 0, err_o
    assert out_o == out_r


@pytest.mark.skipif(not os.path.exists(GEN_BCF), reason="generator absent")
def test_read_bcf_synthetic_2504_samples_hashes():
    os.makedirs(util.CACHE, exist_ok=True)
    path = os.path.join(util.CACHE, "ext_bcf_500x2504.bcf")
    if not os.path.exists(path):
        subprocess.check_call([GEN_BCF, path + ".tmp", "500", "2504", "0x71"])
        os.replace(path + ".tmp", path)
    (rc_r, out_r, _), (rc_o, out_o, err_o) = both("read_bcf", path, "tidy_format=true", "--hash")
    assert rc_r == 0 and rc_o == 0, err_o
    assert out_r.splitlines()[0].startswith("rows\t1252000")
    assert out_o == out_r


def test_read_bcf_region_queries(tmp_path):
    # chained single-region iterators through the CSI index (src/bcf_reader.c:917-951); explicit index_path; the errors
    from tests.test_oracle_bcf import BCF_REGIONS, indexed_bcf
    path = indexed_bcf(tmp_path, n_sites=5000, n_samples=9, seed=13, contigs=6)
    for region in BCF_REGIONS:
        for extra in ((), ("tidy_format=true",)):
            (rc_r, out_r, _), (rc_o, out_o, err_o) = both("read_bcf", path, "region=" + region, *extra)
            assert rc_r == 0 and rc_o == 0, err_o
            assert out_o == out_r, (region, extra)
    import shutil
    alt = str(tmp_path / "elsewhere.csi")
    shutil.move(path + ".csi", alt)
    (rc_r, out_r, _), (rc_o, out_o, err_o) = both("read_bcf", path, "region=chr2:1-9000000", "index_path=" + alt)
    assert rc_r == 0 and rc_o == 0 and out_o == out_r and len(out_r.splitlines()) > 10, err_o
    (rc_r, out_r, _), (rc_o, out_o, _) = both("read_bcf", path, "region=chr2")      # no index next to the file any more
    assert rc_r == 1 and rc_o == 1 and out_o == out_r


def test_read_bcf_bind_errors(tmp_path):
    for args in (("read_bcf", str(tmp_path / "missing.bcf")), ("read_bcf", "")):
        (rc_r, out_r, _), (rc_o, out_o, _) = both(*args)
        assert rc_r == 1 and rc_o == 1
        assert out_o == out_r


# ---- round 2: worker threads, layouts, several pipelines per query, BASELINE-scale parity -------------------------------------------
ALL69 = ",".join(str(c) for c in range(13 + 56))
BENCH_COLS = "0,1,2,3,4,5,6,7,8,9,10,11,12,47,40,14"      # 13 core columns + NM, MD, AS (bench.py's projection)


@pytest.mark.parametrize("layout", ["compact", "duckdb"])
def test_both_wire_layouts_fill_the_same_cells(layout):
    env = {"DUCKHTS_LAYOUT": layout, "DUCKHTS_BATCH_ROWS": "5000"}
    for path, args in ((os.path.join(GOLD, "range.bam"), ()), (os.path.join(GOLD, "auxf_values.bam"), ("standard_tags=true",)),
                       (util.synth_path("bam", n_records=20000, seed=77, paired=True, extra=("--btags",)), ("standard_tags=true", "--cols", ALL69)),
                       (util.synth_path("bam", n_records=20000, seed=77, paired=True, extra=("--btags",)), ("--cols", "10,0,0,3"))):
        rc_r, out_r, _ = run(REF_EXT, "read_bam", path, *args)
        rc_o, out_o, err_o = run(OUR_EXT, "read_bam", path, *args, env=env)
        assert rc_r == 0 and rc_o == 0, err_o
        assert out_o == out_r, (path, args)


@pytest.mark.parametrize("threads", [2, 4, 16])
def test_worker_threads_return_the_reference_rows(threads):
    # several scan threads claim chunks of shared batches; row order across threads is not defined: multiset hashes
    path = util.synth_path("bam", n_records=60000, seed=31, paired=True)
    rc_r, out_r, _ = run(REF_EXT, "read_bam", path, "standard_tags=true", "--cols", ALL69, "--hash")
    for env in ({}, {"DUCKHTS_BATCH_ROWS": "3000", "DUCKHTS_SLOTS": "4"}, {"DUCKHTS_LAYOUT": "duckdb", "DUCKHTS_BATCH_ROWS": "7000"}):
        rc_o, out_o, err_o = run(OUR_EXT, "read_bam", path, "standard_tags=true", "--cols", ALL69, "--hash", "--threads", str(threads), env=env)
        assert rc_r == 0 and rc_o == 0, err_o
        assert multiset(out_o) == multiset(out_r), env
    # the reference's own contig-parallel mode (index next to the file) against ours, both with several threads
    path = util.synth_path("bam", n_records=30000, seed=31, paired=True, extra=("--unplaced", "25"), index=True)
    rc_r, out_r, _ = run(REF_EXT, "read_bam", path, "--hash", "--threads", str(threads))
    rc_o, out_o, err_o = run(OUR_EXT, "read_bam", path, "--hash", "--threads", str(threads))
    assert rc_r == 0 and rc_o == 0, err_o
    assert multiset(out_o) == multiset(out_r)


def test_one_query_over_several_pipelines_keeps_file_order():
    """One source per listed device (here: the same GPU three times), ranges from dh_bam_split / contig groups from
    dh_bam_split_contigs; one scan thread walks the sources in order = the reference's row order, cell for cell."""
    # DUCKHTS_TEST_DEVICES=0,1,... runs this on real GPUs (a multi-GPU box); the default uses GPU 0 three times
    env = {"DUCKHTS_DEVICES": os.environ.get("DUCKHTS_TEST_DEVICES", "0,0,0"), "DUCKHTS_MIN_SHARD_BYTES": "1", "DUCKHTS_BATCH_ROWS": "4096"}
    seq = util.synth_path("bam", n_records=60000, seed=31, paired=True)
    idx = util.synth_path("bam", n_records=30000, seed=31, paired=True, extra=("--unplaced", "25"), index=True)
    for path in (seq, idx, os.path.join(GOLD, "range.bam")):
        rc_r, out_r, _ = run(REF_EXT, "read_bam", path, "standard_tags=true", "--cols", BENCH_COLS)
        rc_o, out_o, err_o = run(OUR_EXT, "read_bam", path, "standard_tags=true", "--cols", BENCH_COLS, env=env)
        assert rc_r == 0 and rc_o == 0, err_o
        assert out_o == out_r, path
        rc_o, out_o, err_o = run(OUR_EXT, "read_bam", path, "standard_tags=true", "--cols", BENCH_COLS, "--hash", "--threads", "6", env=env)
        rc_r, out_r, _ = run(REF_EXT, "read_bam", path, "standard_tags=true", "--cols", BENCH_COLS, "--hash")
        assert rc_o == 0 and multiset(out_o) == multiset(out_r), err_o
    # a region query stays on one pipeline
    (rc_r, out_r, _) = run(REF_EXT, "read_bam", idx, "region=chr3:10,000,000-20,000,000")
    (rc_o, out_o, err_o) = run(OUR_EXT, "read_bam", idx, "region=chr3:10,000,000-20,000,000", env=env)
    assert rc_r == 0 and rc_o == 0 and out_o == out_r, err_o


def test_invalid_record_in_an_early_range_ends_the_whole_scan(tmp_path):
    """Q8: the reference stops at the first invalid record; with the file cut into ranges, rows of later ranges must not appear."""
    from tests.util import bam_header, bam_record
    good = [bam_record(b"g%d" % i, 0, 0, i, 1, [(4, "M")], seq="ACGT", qual=[1] * 4) for i in range(4000)]
    bad = bam_record(b"bad", 0, 0, 5, 1, [(5, "M")], seq="ACGT", qual=[1] * 4)            # CIGAR / l_seq mismatch: bam_read1 -> -4
    # blocks hold whole records, like htslib writes them (the cut points of dh_bam_split must be record boundaries)
    recs = good[:700] + [bad] + good[700:]
    blocks = [util.bgzf_block(bam_header([(b"c1", 100000)]))] + [util.bgzf_block(b"".join(recs[i:i + 50])) for i in range(0, len(recs), 50)]
    path = str(tmp_path / "bad.bam")
    open(path, "wb").write(b"".join(blocks) + util.BGZF_EOF)
    env = {"DUCKHTS_DEVICES": os.environ.get("DUCKHTS_TEST_DEVICES", "0,0,0,0"), "DUCKHTS_MIN_SHARD_BYTES": "1", "DUCKHTS_BATCH_ROWS": "256"}
    rc_r, out_r, _ = run(REF_EXT, "read_bam", path)
    rc_o, out_o, err_o = run(OUR_EXT, "read_bam", path, env=env)
    assert rc_r == 0 and rc_o == 0, err_o
    assert out_o == out_r and len(out_r.splitlines()) == 701


def test_truncated_bam_like_the_reference(tmp_path):
    """A BAM cut in the middle of a block: htslib returns every record in front of the broken block, then its read fails (-2) and the
    scan ends silently.  Pinned against htslib itself (harness/ref_dump: sam_read1 without a thread pool).  The reference EXTENSION reads
    with 2 htslib threads per scan and loses the blocks its reader had in flight when the error arrived (fewer rows, timing-dependent):
    its rows must be a prefix of ours."""
    ref_dump = os.path.join(ROOT, "harness", "_bin", "ref_dump")
    src = util.synth_path("bam", n_records=20000, seed=11)
    data = open(src, "rb").read()
    for cut in (len(data) // 2, len(data) - 40):
        path = str(tmp_path / f"cut{cut}.bam")
        open(path, "wb").write(data[:cut])
        rc_r, out_r, _ = run(REF_EXT, "read_bam", path, "--cols", "0,1,3,9")
        rc_o, out_o, err_o = run(OUR_EXT, "read_bam", path, "--cols", "0,1,3,9")
        assert rc_r == 0 and rc_o == 0, err_o
        rows_r, rows_o = out_r.splitlines(), out_o.splitlines()
        assert 1000 < len(rows_r) <= len(rows_o) and rows_o[:len(rows_r)] == rows_r
        if os.path.exists(ref_dump):
            p = subprocess.run([ref_dump, "dump", path], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL)
            lines = p.stdout.decode().split("\n")[:-1]
            assert lines[-1].startswith("#status -2 ")
            body = [ln for ln in lines if not ln.startswith("#")]
            assert len(body) == len(rows_o) - 1 == int(lines[-1].split()[2])
            assert [ln.split("\t")[0] for ln in body] == [ln.split("\t")[0] for ln in rows_o[1:]]


def test_predicate_pushdown_parameters():
    path = util.synth_path("bam", n_records=60000, seed=31, paired=True)
    rc_r, out_r, _ = run(REF_EXT, "read_bam", path, "--cols", "0,1,3,4")
    rows = [ln.split("\t") for ln in out_r.splitlines()[1:]]
    keep = [r for r in rows if (int(r[1]) & 0x4) == 0 and (int(r[1]) & 0x40) == 0x40 and int(r[3]) >= 30]
    rc_o, out_o, err_o = run(OUR_EXT, "read_bam", path, "flag_exclude=4", "flag_require=64", "min_mapq=30", "--cols", "0,1,3,4")
    assert rc_o == 0, err_o
    assert [ln.split("\t") for ln in out_o.splitlines()[1:]] == keep and 0 < len(keep) < len(rows)


def _unit(n, paired):
    import bench
    return bench.gen_unit("bam", n, 0x9E3779B97F4A7C15, paired, os.cpu_count() or 1)


def test_config1_one_million_single_end_records_cell_for_cell():
    """BASELINE configs[0]: full-column scan of a 1 M-record single-end BAM; chained per-column hashes = every cell, in order."""
    path = _unit(1000000, False)
    (rc_r, out_r, _), (rc_o, out_o, err_o) = both("read_bam", path, "--hash")
    assert rc_r == 0 and rc_o == 0, err_o
    assert out_o == out_r and out_r.splitlines()[0] == "rows\t1000000"


def test_bench_unit_all_columns_against_the_reference_extension():
    """The 4 M-record unit bench.py tiles to 100 M records, all 69 columns (standard_tags), through both extensions; then
    with 8 scan threads on both sides (the reference in its sequential mode has one)."""
    path = _unit(4000000, True)
    (rc_r, out_r, _), (rc_o, out_o, err_o) = both("read_bam", path, "standard_tags=true", "--cols", ALL69, "--hash")
    assert rc_r == 0 and rc_o == 0, err_o
    assert out_o == out_r and int(out_r.splitlines()[0].split("\t")[1]) > 3900000
    rc_o, out_o, err_o = run(OUR_EXT, "read_bam", path, "standard_tags=true", "--cols", ALL69, "--hash", "--threads", "8")
    assert rc_o == 0 and multiset(out_o) == multiset(out_r), err_o


@pytest.mark.parametrize("threads", [1, 4])
def test_read_bcf_worker_threads_and_wide_mode(threads):
    path = os.path.join(util.CACHE, "ext_bcf_500x2504.bcf")
    if not os.path.exists(path):
        os.makedirs(util.CACHE, exist_ok=True)
        subprocess.check_call([GEN_BCF, path + ".tmp", "500", "2504", "0x71"])
        os.replace(path + ".tmp", path)
    for args in (("tidy_format=true",), ("tidy_format=true", "--cols", "0,1,4,6,8,12,13,14,16"), ("--cols", "0,1,2,3,4,5,6,7,8,9,10,11,12,13,14,15,16,17,18,19,20,5000,5001,9000")):
        rc_r, out_r, _ = run(REF_EXT, "read_bcf", path, *args, "--hash")
        rc_o, out_o, err_o = run(OUR_EXT, "read_bcf", path, *args, "--hash", "--threads", str(threads))
        assert rc_r == 0 and rc_o == 0, err_o
        assert multiset(out_o) == multiset(out_r), args
        if threads == 1:
            assert out_o == out_r, args
    edge = os.path.join(GOLD, "edge_small.bcf")
    for args in ((), ("tidy_format=true",)):
        rc_r, out_r, _ = run(REF_EXT, "read_bcf", edge, *args, "--hash")
        for env in ({}, {"DUCKHTS_LAYOUT": "duckdb"}):      # compact FORMAT cells (default) and ready-made DuckDB vectors
            rc_o, out_o, err_o = run(OUR_EXT, "read_bcf", edge, *args, "--hash", "--threads", str(threads), env=env)
            assert rc_r == 0 and rc_o == 0 and multiset(out_o) == multiset(out_r), err_o


def test_bgunzip_streams_in_segments(tmp_path):
    # several staging segments (DH_FD_SEG_MB shrinks them), output compared with the reference's byte for byte
    src = util.synth_path("bam", n_records=60000, seed=31, paired=True)
    outs = []
    for tag, ext, env in (("ref", REF_EXT, {}), ("our", OUR_EXT, {"DH_FD_SEG_MB": "1"})):
        out = str(tmp_path / f"{tag}.bin")
        rc, text, err = run(ext, "bgunzip", src, f"output_path={out}", env=env)
        assert rc == 0, err
        outs.append((open(out, "rb").read(), text.splitlines()[1].split("\t")[2:]))
    assert outs[0] == outs[1] and len(outs[0][0]) > 10_000_000
    # a corrupt block in a later segment: read error like the reference's, same message
    data = bytearray(open(src, "rb").read())
    data[len(data) // 2] ^= 0x55
    bad = str(tmp_path / "bad.bam")
    open(bad, "wb").write(bytes(data))
    msgs = []
    for tag, ext, env in (("ref", REF_EXT, {}), ("our", OUR_EXT, {"DH_FD_SEG_MB": "1"})):
        rc, text, _ = run(ext, "bgunzip", bad, f"output_path={tmp_path / (tag + '2.bin')}", env=env)
        msgs.append((rc, text))
    assert msgs[0] == msgs[1] and msgs[0][0] == 1


def test_files_whose_blocks_do_not_start_records_are_not_cut(tmp_path):
    """Foreign writers cut BGZF blocks anywhere (htslib/test/bgzf_boundaries): dh_bam_split finds no verifiable record boundary
    and the query stays on one pipeline; blocks behind a read longer than a block are skipped as cut points."""
    env = {"DUCKHTS_DEVICES": "0,0,0", "DUCKHTS_MIN_SHARD_BYTES": "1", "DUCKHTS_BATCH_ROWS": "4096"}
    for path in (util.synth_path("bgzf", n_records=30000, seed=13, paired=True), os.path.join(GOLD, "ce_large_seq.bam"), os.path.join(GOLD, "bgzf_boundaries2.bam")):
        rc_r, out_r, _ = run(REF_EXT, "read_bam", path, "--hash")
        rc_o, out_o, err_o = run(OUR_EXT, "read_bam", path, "--hash", env=env)
        assert rc_r == 0 and rc_o == 0, err_o
        assert out_o == out_r, path


def test_corrupt_aux_behind_a_placeholder_cigar_ends_the_scan_like_htslib(tmp_path):
    """bam_tag2cigar's fast path (first CIGAR op = l_seq S) looks for a CG tag; aux data that is broken in front of it (or a
    CG value cut short) is "bad aux data": bam_tag2cigar returns -1 (sam.c:690-694), bam_read1 -4, the reference scan ends there."""
    import struct
    from tests.util import aux_tag, bam_header, bam_record
    good = [bam_record(b"g%d" % i, 0, 0, i, 1, [(4, "M")], seq="ACGT", qual=[1] * 4) for i in range(40)]
    cases = {"unknown_type_in_front": aux_tag("NM", "C", 1) + b"ZZ?" + aux_tag("AS", "C", 9),
             "cg_value_cut_short": aux_tag("NM", "C", 1) + b"CGBI" + struct.pack("<I", 5) + struct.pack("<II", 1, 2),
             "no_cg_at_all": aux_tag("NM", "C", 1) + aux_tag("AS", "C", 9),
             "corrupt_behind_cg": b"CGBI" + struct.pack("<I", 2) + struct.pack("<II", (3 << 4) | 0, (2 << 4) | 1) + b"ZZ?x"}
    for name, aux in cases.items():
        placeholder = bam_record(b"ph", 0, 0, 7, 1, [(5, "S"), (9, "N")], seq="ACGTA", qual=[2] * 5, aux=aux)
        path = str(tmp_path / f"{name}.bam")
        open(path, "wb").write(util.make_bgzf(bam_header([(b"c1", 100000)]) + b"".join(good[:25]) + placeholder + b"".join(good[25:])))
        (rc_r, out_r, _), (rc_o, out_o, err_o) = both("read_bam", path)
        assert rc_r == 0 and rc_o == 0, err_o
        assert out_o == out_r, name
        n = len(out_r.splitlines()) - 1
        assert n == (41 if name in ("no_cg_at_all", "corrupt_behind_cg") else 25), (name, n)
