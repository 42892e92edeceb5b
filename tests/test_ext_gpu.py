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


def run(ext, func, path, *args):
    p = subprocess.run([HOST, ext, func, path, *args], stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=600)
    return p.returncode, p.stdout.decode("utf-8", "replace"), p.stderr.decode("utf-8", "replace")


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
