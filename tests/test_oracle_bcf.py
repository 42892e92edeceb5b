"""Pins oracle/bcf_oracle.py (CPU restatement of rows B1-B10) against the REFERENCE: dumps the reference extension
produced under the fake DuckDB host (tests/golden/*.bcf.*.txt) and live runs on seeded synthetic files."""
import os
import subprocess

import numpy as np
import pytest

from oracle import bcf_oracle
from oracle import dh_oracle_py as orc
from tests import util

HOST = os.path.join(util.ROOT, "harness", "_bin", "fake_duckdb")
REF_EXT = os.path.join(util.ROOT, "oracle", "_ref", "libduckhts_ref.so")
GEN = os.path.join(util.ROOT, "harness", "_bin", "gen_bcf")


def inflated(path):
    rc, infl, bad = orc.bgzf_inflate(np.fromfile(path, dtype=np.uint8))
    assert rc == 0, (rc, bad)
    return infl.tobytes()


@pytest.mark.parametrize("name", ["vcf_file", "edge_small", "plain_small"])
@pytest.mark.parametrize("tidy", [True, False])
def test_bcf_oracle_matches_reference_golden(name, tidy):
    exp = open(os.path.join(util.GOLDEN, "%s.bcf.%s.txt" % (name, "tidy" if tidy else "wide"))).read().split("\n")[:-1]
    got, status = bcf_oracle.scan(inflated(os.path.join(util.GOLDEN, name + ".bcf")), tidy)
    assert status == 0
    assert len(got) == len(exp)
    for i, (a, b) in enumerate(zip(got, exp)):
        assert a == b, "line %d" % i


def test_known_answers_of_duckhts_sqllogictest_bcf():
    """test/sql/duckhts.test:15-40: vcf_file.bcf has 15 records; the first is 1:3000150 C>T."""
    got, _ = bcf_oracle.scan(inflated(os.path.join(util.GOLDEN, "vcf_file.bcf")), False)
    assert len(got) == 16
    assert got[1].split("\t")[:5] == ["1", "3000150", "\\N", "C", "[T]"]


def test_known_answers_of_duckhts_sqllogictest_bcf_regions():
    """test/sql/duckhts.test:86-105: region := '1:3000150-3000151' has 2 records; a comma-separated list is the sum of its parts."""
    u = inflated(os.path.join(util.GOLDEN, "vcf_file.bcf"))
    n = lambda regs: len(bcf_oracle.scan(u, False, regions=regs)[0]) - 1
    assert n(["1:3000150-3000151"]) == 2
    assert n(["1:3000150-3000151", "1:3062915-3062915"]) == n(["1:3000150-3000151"]) + n(["1:3062915-3062915"]) > 2


@pytest.mark.skipif(not (os.path.exists(HOST) and os.path.exists(REF_EXT) and os.path.exists(GEN)), reason="prebuilt reference extension / generator absent")
@pytest.mark.parametrize("n_sites,n_samples,seed", [(400, 12, 0x81), (150, 0, 0x82), (60, 120, 0x83)])
def test_bcf_oracle_matches_live_reference_on_synthetic(n_sites, n_samples, seed, tmp_path):
    path = str(tmp_path / "s.bcf")
    subprocess.check_call([GEN, path, str(n_sites), str(n_samples), hex(seed), "--edge", "--contigs", "5"])
    u = inflated(path)
    for tidy in (True, False):
        args = ["tidy_format=true"] if tidy else []
        ref = subprocess.run([HOST, REF_EXT, "read_bcf", path, *args, "--dump"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL).stdout.decode("utf-8", "replace").split("\n")[:-1]
        got, status = bcf_oracle.scan(u, tidy)
        assert status == 0 and got == ref, tidy


def test_bcf_oracle_stops_at_a_bad_record():
    u = bytearray(inflated(os.path.join(util.GOLDEN, "plain_small.bcf")))
    full, _ = bcf_oracle.scan(bytes(u), False)
    import struct
    l_text = struct.unpack_from("<I", u, 5)[0]
    p = 9 + l_text
    for _ in range(7):                                              # corrupt the 8th record's contig id
        l_shared, l_indiv = struct.unpack_from("<II", u, p)
        p += 8 + l_shared + l_indiv
    struct.pack_into("<i", u, p + 8, 9999)
    got, status = bcf_oracle.scan(bytes(u), False)
    assert status == -2 and got == full[:8]


BCF_REGIONS = ["chr1", "chr2:1000000-8000000", "chr3:5,000,000-9000000,chr1:1-3000000,chr3:8000000-12000000", "nosuch,chr4:100-",
               " chr2:7,000,000-7,900,000 ,, chr2:7,500,000-", "chr9", "chr4:-2000000"]


def indexed_bcf(tmp_path, n_sites=3000, n_samples=5, seed=11, contigs=4):
    path = str(tmp_path / "r.bcf")
    subprocess.check_call([GEN, path, str(n_sites), str(n_samples), str(seed), "--contigs", str(contigs), "--edge", "--index"], stderr=subprocess.DEVNULL)
    return path


def split_regions(region):      # parse_regions_duckdb (src/bcf_reader.c:286-329)
    return [t.strip(" \t") for t in region.split(",") if t.strip(" \t")]


@pytest.mark.skipif(not (os.path.exists(HOST) and os.path.exists(REF_EXT) and os.path.exists(GEN)), reason="prebuilt reference extension / generator absent")
def test_bcf_oracle_region_queries_match_live_reference(tmp_path):
    """read_bcf(region := ...) through the CSI index: chained single-region iterators, duplicates kept, unknown contigs skipped."""
    path = indexed_bcf(tmp_path)
    u = inflated(path)
    full = subprocess.run([HOST, REF_EXT, "read_bcf", path, "--cols", "0,1,3"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL).stdout.decode().split("\n")[1:-1]
    # a record with a multi-base REF: queried by its last base only (pos + rlen > beg), and just past it
    chrom, pos, ref = next(r.split("\t") for r in full if len(r.split("\t")[2]) > 2)
    last = int(pos) + len(ref) - 1
    extra = ["%s:%d-%d" % (chrom, last, last), "%s:%d-%d" % (chrom, last + 1, last + 1)]
    for region in BCF_REGIONS + extra:
        for tidy in (False, True):
            args = ["region=" + region] + (["tidy_format=true"] if tidy else [])
            ref_out = subprocess.run([HOST, REF_EXT, "read_bcf", path, *args, "--dump"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL).stdout.decode("utf-8", "replace").split("\n")[:-1]
            got, status = bcf_oracle.scan(u, tidy, regions=split_regions(region))
            assert status == 0 and got == ref_out, (region, tidy)
    assert len(bcf_oracle.scan(u, False, regions=[extra[0]])[0]) >= 2
