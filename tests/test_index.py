"""A8 — index-driven read_bam scans (BAI/CSI, region queries, per-contig scans, contig-group sharding).

CPU part pins the oracle's restatement of htslib's iterators (oracle/dh_oracle_idx.inc) against golden vectors the
reference produced (tests/golden/range.region*.tsv, range.contigs.tsv: sam_itr_regarray / sam_itr_queryi through
harness/_bin/ref_dump) and against live ref_dump runs on seeded multi-block files.  GPU part (-m gpu) compares the
product (range scans + overlap filter inside the boundary-walk kernel) with the oracle, row for row."""
import glob
import os
import subprocess

import numpy as np
import pytest

from oracle import dh_oracle_py as orc
from tests import util

REF_DUMP = os.path.join(util.ROOT, "harness", "_bin", "ref_dump")
TAGS3 = [(b"NM", "i"), (b"MD", "Z"), (b"AS", "i")]
SYNTH_REGIONS = ["chr1", "chr3:10,000,000-20,000,000", "chr2:1-5000000,chr2:4000000-9000000,chr7:49000000-", "chr22:-1000000,chr5:25m-26m",
                 "chr9:12345678-12345678", "nosuch,chr4:1-100", "chr11:30000000-30200000,chr10:1-2000000,chr11:1-100000"]


def load(path):
    img = np.fromfile(path, dtype=np.uint8)
    rc, infl, bad = orc.bgzf_inflate(img)
    assert rc == 0
    return img, infl


def subset_lines(lines, rows):
    head = [l for l in lines if l.startswith("#n_ref") or l.startswith("#ref")]
    data = [l for l in lines if not l.startswith("#")]
    return head + [data[i] for i in rows] + ["#status 0 %d" % len(rows)]


def golden_region_cases():
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(util.GOLDEN, "range.region*.txt")))


@pytest.mark.parametrize("case", golden_region_cases())
def test_oracle_region_rows_match_reference_golden(case):
    img, infl = load(os.path.join(util.GOLDEN, "range.bam"))
    bai = np.fromfile(os.path.join(util.GOLDEN, "range.bam.bai"), dtype=np.uint8)
    dec = orc.bam_decode(infl, util.GOLDEN_TAGS)
    seq_lines = util.dump_lines(dec, util.GOLDEN_TAGS, True)
    regions = open(os.path.join(util.GOLDEN, case + ".txt")).read().strip().split(",")
    exp = open(os.path.join(util.GOLDEN, case + ".tsv")).read().split("\n")[:-1]
    rows = orc.region_scan_rows(img, bai, dec, regions)
    if exp[-1] == "#no-iterator":
        assert rows is None
        return
    assert subset_lines(seq_lines, rows) == exp


def test_oracle_contig_rows_match_reference_golden():
    img, infl = load(os.path.join(util.GOLDEN, "range.bam"))
    bai = np.fromfile(os.path.join(util.GOLDEN, "range.bam.bai"), dtype=np.uint8)
    dec = orc.bam_decode(infl, util.GOLDEN_TAGS)
    rows = orc.contig_scan_rows(img, bai, dec)
    exp = open(os.path.join(util.GOLDEN, "range.contigs.tsv")).read().split("\n")[:-1]
    assert subset_lines(util.dump_lines(dec, util.GOLDEN_TAGS, True), rows) == exp


def test_known_answers_of_duckhts_sqllogictest_regions():
    """test/sql/duckhts.test:139-160: region := 'CHROMOSOME_I' -> 18 rows, 'CHROMOSOME_I:1-1000' -> 2, duplicates dedup."""
    img, infl = load(os.path.join(util.GOLDEN, "range.bam"))
    bai = np.fromfile(os.path.join(util.GOLDEN, "range.bam.bai"), dtype=np.uint8)
    dec = orc.bam_decode(infl, ())
    assert len(orc.region_scan_rows(img, bai, dec, ["CHROMOSOME_I"])) == 18
    assert len(orc.region_scan_rows(img, bai, dec, ["CHROMOSOME_I:1-1000"])) == 2
    assert len(orc.region_scan_rows(img, bai, dec, ["CHROMOSOME_I:1-1000", "CHROMOSOME_I:1-1000"])) == 2


def synth_indexed(n=30000, seed=31, unplaced=25, csi=False):
    path = util.synth_path("bam", n_records=n, seed=seed, paired=True, extra=("--unplaced", str(unplaced)), index=True)
    if csi and not os.path.exists(path + ".csi"):
        subprocess.check_call([REF_DUMP, "index", path, "14"])
    return path


@pytest.mark.skipif(not os.path.exists(REF_DUMP), reason="ref_dump binary absent")
def test_oracle_matches_live_reference_on_indexed_synthetic():
    path = synth_indexed()
    img, infl = load(path)
    bai = np.fromfile(path + ".bai", dtype=np.uint8)
    dec = orc.bam_decode(infl, TAGS3)
    seq_lines = util.dump_lines(dec, TAGS3, True)
    assert len(orc.bgzf_walk(img)[0]) > 50          # many BGZF blocks: the chunk logic matters
    for reg in SYNTH_REGIONS:
        ref = subprocess.run([REF_DUMP, "dumpregion", path, reg, "NM:i", "MD:Z", "AS:i"], stdout=subprocess.PIPE,
                             stderr=subprocess.DEVNULL).stdout.decode().split("\n")[:-1]
        rows = orc.region_scan_rows(img, bai, dec, reg.split(","))
        assert subset_lines(seq_lines, rows) == ref, reg
    ref = subprocess.check_output([REF_DUMP, "dumpcontigs", path, "NM:i", "MD:Z", "AS:i"]).decode().split("\n")[:-1]
    rows = orc.contig_scan_rows(img, bai, dec)
    assert len(rows) == dec["n"] - 25               # reads without a contig are never returned in this mode
    assert subset_lines(seq_lines, rows) == ref


# ---------------------------------------------------------------------------------------------------- GPU
def take_rows(dec, rows, tags):
    """Oracle decode restricted to `rows`, in dump_lines text form."""
    return subset_lines(util.dump_lines(dec, tags, True), rows)


@pytest.fixture(scope="module")
def ctx():
    import duckhts_b200 as dh
    with dh.Context(0) as c:
        yield c


@pytest.mark.gpu
@pytest.mark.parametrize("case", golden_region_cases())
def test_gpu_region_scan_matches_reference_golden(ctx, case):
    import duckhts_b200 as dh
    img = np.fromfile(os.path.join(util.GOLDEN, "range.bam"), dtype=np.uint8)
    bai = np.fromfile(os.path.join(util.GOLDEN, "range.bam.bai"), dtype=np.uint8)
    regions = open(os.path.join(util.GOLDEN, case + ".txt")).read().strip().split(",")
    exp = open(os.path.join(util.GOLDEN, case + ".tsv")).read().split("\n")[:-1]
    got = dh.bam_scan(ctx, img, tags=util.GOLDEN_TAGS, index=bai, regions=regions)
    if exp[-1] == "#no-iterator":
        assert got is None
        return
    assert util.dump_lines(got, util.GOLDEN_TAGS, False) == exp


@pytest.mark.gpu
def test_gpu_contig_scan_matches_reference_golden(ctx):
    import duckhts_b200 as dh
    img = np.fromfile(os.path.join(util.GOLDEN, "range.bam"), dtype=np.uint8)
    bai = np.fromfile(os.path.join(util.GOLDEN, "range.bam.bai"), dtype=np.uint8)
    got = dh.bam_scan(ctx, img, tags=util.GOLDEN_TAGS, index=bai, contigs=True)
    assert util.dump_lines(got, util.GOLDEN_TAGS, False) == open(os.path.join(util.GOLDEN, "range.contigs.tsv")).read().split("\n")[:-1]


@pytest.mark.gpu
@pytest.mark.parametrize("csi", [False, True])
def test_gpu_region_scans_match_oracle_on_indexed_synthetic(ctx, csi):
    import duckhts_b200 as dh
    if csi and not os.path.exists(REF_DUMP):
        pytest.skip("ref_dump binary absent")
    path = synth_indexed(csi=csi)
    img, infl = load(path)
    idx = np.fromfile(path + (".csi" if csi else ".bai"), dtype=np.uint8)
    dec = orc.bam_decode(infl, TAGS3)
    for reg in SYNTH_REGIONS:
        rows = orc.region_scan_rows(img, idx, dec, reg.split(","))
        # small segments and batches: ranges that span several segments, batches that end inside a range
        got = dh.bam_scan(ctx, img, tags=TAGS3, index=idx, regions=reg.split(","), seg_comp=1 << 18, batch_rows=700)
        assert util.dump_lines(got, TAGS3, False) == take_rows(dec, rows, TAGS3), reg
    rows = orc.contig_scan_rows(img, idx, dec)
    got = dh.bam_scan(ctx, img, tags=TAGS3, index=idx, contigs=True)
    assert got["n"] == dec["n"] - 25
    assert util.dump_lines(got, TAGS3, False) == take_rows(dec, rows, TAGS3)


@pytest.mark.gpu
@pytest.mark.parametrize("n_parts", [2, 8])
def test_gpu_contig_group_shards_concatenate_to_the_contig_scan(ctx, n_parts):
    """BASELINE configs[3]: region-sharded scan, contig groups balanced by compressed bytes, merged in order on the host."""
    import duckhts_b200 as dh
    path = synth_indexed()
    img, infl = load(path)
    bai = np.fromfile(path + ".bai", dtype=np.uint8)
    dec = orc.bam_decode(infl, TAGS3)
    with dh.BamFile(ctx, img) as bf, dh.Index(ctx, bai) as ix:
        bounds = bf.split_contigs(ix, n_parts)
    assert bounds[0] == 0 and bounds[-1] == 22 and all(a <= b for a, b in zip(bounds, bounds[1:]))
    assert max(b - a for a, b in zip(bounds, bounds[1:])) <= 22 // n_parts + 2      # balanced: contigs are equally loaded here
    lines = []
    for a, b in zip(bounds, bounds[1:]):
        got = dh.bam_scan(ctx, img, tags=TAGS3, index=bai, contigs=(a, b))
        exp_rows = orc.contig_scan_rows(img, bai, dec, a, b)
        part = util.dump_lines(got, TAGS3, False)
        assert part == take_rows(dec, exp_rows, TAGS3), (a, b)
        lines += [l for l in part if not l.startswith("#")]
    whole = [l for l in take_rows(dec, orc.contig_scan_rows(img, bai, dec), TAGS3) if not l.startswith("#")]
    assert lines == whole
