"""B1-B10 — read_bcf on the GPU against the REFERENCE: golden dumps the reference extension produced under the fake
DuckDB host (tests/golden/*.bcf.{tidy,wide}.txt, make_golden.sh), and — where the prebuilt reference extension is
present — live runs of it on seeded synthetic files (harness/_bin/gen_bcf), cell for cell."""
import os
import subprocess

import numpy as np
import pytest

from tests import util

ROOT = util.ROOT
HOST = os.path.join(ROOT, "harness", "_bin", "fake_duckdb")
REF_EXT = os.path.join(ROOT, "oracle", "_ref", "libduckhts_ref.so")
GEN = os.path.join(ROOT, "harness", "_bin", "gen_bcf")
GOLD = util.GOLDEN

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    import duckhts_b200 as dh
    with dh.Context(0) as c:
        yield c


def first_diff(a, b):
    for i, (x, y) in enumerate(zip(a, b)):
        if x != y:
            return "line %d:\n  got %s\n  exp %s" % (i, x[:600], y[:600])
    return "length %d vs %d" % (len(a), len(b))


@pytest.mark.parametrize("name", ["vcf_file", "edge_small", "plain_small"])
@pytest.mark.parametrize("tidy", [True, False])
def test_read_bcf_matches_reference_golden(ctx, name, tidy):
    from duckhts_b200 import bcf
    img = np.fromfile(os.path.join(GOLD, name + ".bcf"), dtype=np.uint8)
    exp = open(os.path.join(GOLD, "%s.bcf.%s.txt" % (name, "tidy" if tidy else "wide"))).read().split("\n")[:-1]
    got, status, _ = bcf.rows_text(ctx, img, tidy=tidy)
    assert status == 0
    assert got == exp, first_diff(got, exp)
    got, status, _ = bcf.rows_text(ctx, img, tidy=tidy, compact=True, batch_records=3)      # compact FORMAT cells (GT 4 bytes, numeric lists as counts)
    assert status == 0 and got == exp, first_diff(got, exp)


def ref_dump(path, *args):
    p = subprocess.run([HOST, REF_EXT, "read_bcf", path, *args, "--dump"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, timeout=900)
    assert p.returncode == 0
    return p.stdout.decode("utf-8", "replace").split("\n")[:-1]


def synth_bcf(n_sites, n_samples, seed, edge, contigs=22):
    os.makedirs(util.CACHE, exist_ok=True)
    path = os.path.join(util.CACHE, "bcf_%d_%d_%d_%d_%d.bcf" % (n_sites, n_samples, seed, int(edge), contigs))
    if not os.path.exists(path):
        subprocess.check_call([GEN, path + ".tmp", str(n_sites), str(n_samples), hex(seed), "--contigs", str(contigs)] + (["--edge"] if edge else []))
        os.replace(path + ".tmp", path)
    return path


@pytest.mark.skipif(not (os.path.exists(HOST) and os.path.exists(REF_EXT) and os.path.exists(GEN)), reason="prebuilt reference extension / generator absent")
@pytest.mark.parametrize("n_sites,n_samples,seed,edge,kw", [
    (1500, 40, 0x61, True, {}),                                   # every record shape, records straddling BGZF blocks
    (400, 300, 0x62, True, dict(batch_records=37, seg_comp=1 << 17)),   # many small segments and batches, carries between segments
    (300, 2504, 0x63, False, dict(batch_records=64)),             # the config-5 shape: 2,504 samples
    (200, 0, 0x64, True, {}),                                     # sites only
])
def test_read_bcf_matches_live_reference_on_synthetic(ctx, n_sites, n_samples, seed, edge, kw):
    from duckhts_b200 import bcf
    path = synth_bcf(n_sites, n_samples, seed, edge)
    img = np.fromfile(path, dtype=np.uint8)
    exp = ref_dump(path, "tidy_format=true")
    got, status, _ = bcf.rows_text(ctx, img, tidy=True, **kw)
    assert status == 0
    assert len(got) == len(exp) == 1 + n_sites * max(n_samples, 1)
    assert got == exp, first_diff(got, exp)
    if n_samples <= 300:
        got, status, _ = bcf.rows_text(ctx, img, tidy=True, compact=True, **kw)
        assert status == 0 and got == exp, first_diff(got, exp)


# ---- against the oracle (oracle/bcf_oracle.py), incl. inputs the reference stops on ---------------------------------
def oracle_lines(img, tidy):
    from oracle import bcf_oracle
    from oracle import dh_oracle_py as orc
    rc, infl, _ = orc.bgzf_inflate(img)
    assert rc == 0
    return bcf_oracle.scan(infl.tobytes(), tidy), infl


@pytest.mark.parametrize("name", ["vcf_file", "edge_small", "plain_small"])
def test_read_bcf_matches_oracle(ctx, name):
    from duckhts_b200 import bcf
    img = np.fromfile(os.path.join(GOLD, name + ".bcf"), dtype=np.uint8)
    (exp, est), _ = oracle_lines(img, True)
    got, status, _ = bcf.rows_text(ctx, img, tidy=True, batch_records=7)
    assert status == est == 0 and got == exp, first_diff(got, exp)


@pytest.mark.parametrize("what", ["rid", "fmt_type", "truncated", "l_shared", "info_key"])
def test_invalid_records_end_the_scan_like_the_reference(ctx, what):
    """bcf_read returns -2 (bcf_read1_core / bcf_record_check); the rows before the bad record are returned."""
    import struct
    from oracle import bcf_oracle
    from duckhts_b200 import bcf
    img = np.fromfile(os.path.join(GOLD, "edge_small.bcf"), dtype=np.uint8)
    (_, _), infl = oracle_lines(img, True)
    u = bytearray(infl.tobytes())
    l_text = struct.unpack_from("<I", u, 5)[0]
    p = 9 + l_text
    for _ in range(23):
        l_shared, l_indiv = struct.unpack_from("<II", u, p)
        p += 8 + l_shared + l_indiv
    l_shared, l_indiv = struct.unpack_from("<II", u, p)
    if what == "rid":
        struct.pack_into("<i", u, p + 8, 77)
    elif what == "fmt_type":
        u[p + 8 + l_shared + 2] = (u[p + 8 + l_shared + 2] & 0xf0) | 6      # an undefined value type in the first FORMAT field
    elif what == "truncated":
        u = u[:p + 8 + l_shared + 5]
    elif what == "l_shared":
        struct.pack_into("<I", u, p, 20)
    elif what == "info_key":
        u[p + 32] = 0x13                                                       # the ID descriptor becomes an int vector
    bad = np.frombuffer(util.make_bgzf(bytes(u), block=30000), dtype=np.uint8)
    exp, est = bcf_oracle.scan(bytes(u), True)
    got, status, last = bcf.rows_text(ctx, bad, tidy=True, batch_records=5)
    assert est == -2 and status == -2 and last == -2
    assert len(exp) == 1 + 23 * 5
    assert got == exp, first_diff(got, exp)


def test_projection_returns_only_requested_columns(ctx):
    import duckhts_b200 as dh
    img = np.fromfile(os.path.join(GOLD, "edge_small.bcf"), dtype=np.uint8)
    with dh.BcfFile(ctx, img) as bf:
        bf.set_projection(core_mask=0b0000010, info_want=[False] * len(bf.info), format_want=[f[0] == "GT" for f in bf.format])
        bf.scan_begin()
        b = bf.next_batch()
        assert b.n_records == 60 and b.core[1] and not b.core[0] and not b.core[4]
        gt = [i for i, f in enumerate(bf.format) if f[0] == "GT"][0]
        assert all((b.format[i] is not None) == (i == gt) for i in range(len(bf.format)))
        assert all(b.info[i] is None for i in range(len(bf.info)))
        assert bf.next_batch() is None


# ---- region queries through the CSI index (bcf_itr_querys / bcf_itr_next; src/bcf_reader.c:917-951, 1318-1333) ------------
@pytest.mark.skipif(not os.path.exists(GEN), reason="generator absent")
def test_read_bcf_region_queries_match_oracle(ctx, tmp_path):
    from duckhts_b200 import bcf
    from tests.test_oracle_bcf import BCF_REGIONS, indexed_bcf, split_regions
    path = indexed_bcf(tmp_path, n_sites=6000, n_samples=7, seed=12, contigs=5)
    img = np.fromfile(path, dtype=np.uint8)
    csi = np.fromfile(path + ".csi", dtype=np.uint8)
    (_, _), infl = oracle_lines(img, False)
    from oracle import bcf_oracle
    for region in BCF_REGIONS:
        regs = split_regions(region)
        for tidy, kw in ((False, {}), (True, dict(batch_records=64, seg_comp=1 << 16))):
            exp, est = bcf_oracle.scan(infl.tobytes(), tidy, regions=regs)
            got, status, _ = bcf.rows_text(ctx, img, tidy=tidy, index=csi, regions=regs, **kw)
            assert status == est == 0 and got == exp, (region, tidy, first_diff(got, exp))


def test_reference_sqllogictest_region_answers(ctx):
    """test/sql/duckhts.test:86-105 on the reference's own vcf_file.bcf + .csi (test/data, bcftools index)."""
    from duckhts_b200 import bcf
    img = np.fromfile(os.path.join(GOLD, "vcf_file.bcf"), dtype=np.uint8)
    csi = np.fromfile(os.path.join(GOLD, "vcf_file.bcf.csi"), dtype=np.uint8)
    n = lambda regs: len(bcf.rows_text(ctx, img, tidy=False, index=csi, regions=regs)[0]) - 1
    assert n(["1:3000150-3000151"]) == 2
    assert n(["1:3000150-3000151", "1:3062915-3062915"]) == n(["1:3000150-3000151"]) + n(["1:3062915-3062915"]) > 2
    from oracle import bcf_oracle
    (_, _), infl = oracle_lines(img, False)
    for regs in (["1"], ["2", "1:3062915-3062915"], ["4:3258448-"], ["X"], ["20"]):
        exp, _ = bcf_oracle.scan(infl.tobytes(), True, regions=regs)
        got, status, _ = bcf.rows_text(ctx, img, tidy=True, index=csi, regions=regs)
        assert status == 0 and got == exp, (regs, first_diff(got, exp))
