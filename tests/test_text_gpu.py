"""N4: the text path (dh_text_*; read_tabix in the extension) against (a) a plain Python restatement of hts_getline + get_field
(bgzf_getline, htslib/bgzf.c:2308-2370; src/tabix_reader.c:234-254), (b) the reference's own fixtures and known answers
(test/sql/duckhts.test:422-499) as dumped from the unchanged reference extension (tests/golden/tabix.*.tsv, make_golden.sh), and
(c) the reference extension run live on seeded synthetic files where it is present."""
import gzip
import os
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
G = util.GOLDEN


def oracle_lines(data, n_cols, proj, meta=b"#"):
    """hts_getline per line ('\\n' ends a line, a '\\r' in front of it is dropped, a last line without '\\n' counts), get_field per cell."""
    parts = data.split(b"\n")
    if parts and parts[-1] == b"":
        parts.pop()
    out = []
    for ln in parts:
        if ln.endswith(b"\r"):
            ln = ln[:-1]
        fields = ln.split(b"\t")
        flags = 1 if len(ln) == 0 else (2 if meta and ln[:1] == meta else 0)
        out.append((flags, [fields[c] if c < len(fields) and c < n_cols else None for c in proj]))
    return out


def write_bgzf(path, data, block):
    open(path, "wb").write(util.make_bgzf(data, block=block))


def synth_table(n, seed, crlf=False, last_newline=True):
    rng = np.random.default_rng(seed)
    lines = [b"#meta line one", b"#another\twith\ttabs", b"", b"chr0\t1\t2\tA\t0.5"]     # the first data line sets the number of columns
    for i in range(n):
        k = int(rng.integers(0, 12))
        if k == 0:
            lines.append(b"")                                              # empty line
        elif k == 1:
            lines.append(b"#comment %d" % i)
        elif k == 2:
            lines.append(b"chr%d\t%d" % (i % 5, i))                        # fewer fields than the schema
        elif k == 3:
            lines.append(b"\t".join([b"x%d" % j for j in range(9)]))       # more fields than the schema
        elif k == 4:
            lines.append(b"chr1\t.\t\t%d\t" % i)                           # ".", empty field, trailing tab
        else:
            lines.append(b"chr%d\t%d\t%d\t%s\t%.3f" % (i % 23, 100 * i, 100 * i + int(rng.integers(1, 500)), b"ACGT"[i % 4: i % 4 + 1] * int(rng.integers(1, 60)), rng.random()))
    nl = b"\r\n" if crlf else b"\n"
    data = nl.join(lines)
    return data + (nl if last_newline else b"")


@pytest.mark.parametrize("block,seg_kb,crlf,last_nl", [(0xff00, 0, False, True), (1000, 0, False, False), (333, 64, True, True), (4096, 64, False, True), (0xff00, 64, True, False)])
def test_lines_and_fields_against_the_restatement(ctx, tmp_path, monkeypatch, block, seg_kb, crlf, last_nl):
    data = synth_table(40000, 7 + block, crlf, last_nl)
    path = str(tmp_path / "t.tsv.gz")
    write_bgzf(path, data, block)
    if seg_kb:
        monkeypatch.setenv("DH_TEXT_SEG_KB", str(seg_kb))                  # many segments: lines carried from one to the next
    for n_cols, proj in ((5, None), (5, [4, 0]), (3, [1]), (7, [6, 2, 0]), (5, [])):
        with dh.TextFile(ctx, path) as tf:
            got, status = tf.scan(n_cols, proj)
            if seg_kb:
                assert tf.batches > 2
        exp = oracle_lines(data, n_cols, list(range(n_cols)) if proj is None else proj)
        assert status == 0 and len(got) == len(exp)
        assert got == exp


def test_long_lines_and_degenerate_files(ctx, tmp_path, monkeypatch):
    monkeypatch.setenv("DH_TEXT_SEG_KB", "64")
    rng = np.random.default_rng(3)
    wide = b"\t".join(b"%d|%d:%d" % (i % 3, i % 2, int(rng.integers(0, 99))) for i in range(120000))      # one ~1 MB line: many blocks, several segments
    cases = {
        "one_long_line": wide + b"\n",
        "long_then_short": wide + b"\nshort\tline\n" + wide[:300000] + b"\nlast",
        "only_newlines": b"\n\n\n\n",
        "only_meta": b"#a\n#b\n",
        "single_no_newline": b"a\tb\tc",
        "tabs_only": b"\t\t\t\n\t\n",
    }
    for name, data in cases.items():
        path = str(tmp_path / (name + ".gz"))
        write_bgzf(path, data, 0xff00)
        with dh.TextFile(ctx, path) as tf:
            got, status = tf.scan(4, [0, 3, 1])
        assert status == 0 and got == oracle_lines(data, 4, [0, 3, 1]), name
    # an empty file (just the EOF marker) has no line; something that is not BGZF is refused when it is opened
    path = str(tmp_path / "empty.gz")
    open(path, "wb").write(util.BGZF_EOF)
    with dh.TextFile(ctx, path) as tf:
        assert tf.scan(1) == ([], 0)
    path = str(tmp_path / "plain.gz")
    open(path, "wb").write(gzip.compress(b"a\tb\n"))
    with pytest.raises(dh.DuckhtsError, match="not BGZF"):
        dh.TextFile(ctx, path)


def test_a_cut_file_returns_the_lines_in_front(ctx, tmp_path):
    data = synth_table(30000, 11)
    img = util.make_bgzf(data, block=8000)
    path = str(tmp_path / "cut.gz")
    open(path, "wb").write(img[: len(img) * 2 // 3])
    with dh.TextFile(ctx, path) as tf:
        got, status = tf.scan(5)
    assert status == -2 and 0 < len(got) < data.count(b"\n")
    exp = oracle_lines(data, 5, list(range(5)))
    assert got[:-1] == exp[: len(got) - 1]                                   # the last one may be the cut line's front


# ---- the table function ----------------------------------------------------------------------------------------------------------
def run(ext, func, *args):
    p = subprocess.run([HOST, ext, func, *args], stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=600)
    return p.returncode, p.stdout.decode()


GOLDEN_CASES = [("tabix.meta.tsv", "meta_tabix.tsv.gz", ()), ("tabix.meta_names.tsv", "meta_tabix.tsv.gz", ("header_names=chrom,pos,value",)),
                ("tabix.meta_auto.tsv", "meta_tabix.tsv.gz", ("auto_detect=true",)), ("tabix.header.tsv", "header_tabix.tsv.gz", ("header=true",)),
                ("tabix.header_types.tsv", "header_tabix.tsv.gz", ("header=true", "column_types=VARCHAR,BIGINT,VARCHAR")),
                ("tabix.gff.tsv", "gff_file.gff.gz", ()), ("tabix.gff_proj.tsv", "gff_file.gff.gz", ("--cols", "0,2")),
                ("tabix.vcf.tsv", "formatcols.vcf.gz", ()), ("tabix.no_contig_auto.tsv", "no_contig.vcf.gz", ("auto_detect=true",))]


@pytest.mark.skipif(not os.path.exists(HOST), reason="prebuilt fake host absent")
@pytest.mark.parametrize("dump,fixture,args", GOLDEN_CASES, ids=[c[0] for c in GOLDEN_CASES])
def test_read_tabix_reference_fixtures(dump, fixture, args):
    """The reference's fixtures (test/data) and its answers for them, incl. the known answers of test/sql/duckhts.test:426-499."""
    rc, out = run(OUR_EXT, "read_tabix", os.path.join(G, fixture), *args, "--dump", "--types")
    assert rc == 0
    assert out == open(os.path.join(G, dump)).read()


@pytest.mark.skipif(not (os.path.exists(HOST) and os.path.exists(REF_EXT)), reason="prebuilt fake host / reference extension absent")
def test_read_tabix_matches_the_live_reference(tmp_path):
    data = synth_table(60000, 21)
    path = str(tmp_path / "big.tsv.gz")
    write_bgzf(path, data, 0xff00)
    hdr = b"chrom\tstart\tend\tseq\tscore\n" + b"\n".join(b"chr%d\t%d\t%d\t%s\t%s" % (i % 7, i, i + 9, b"ACGT" * (i % 5), [b"1.5", b"2", b".", b"x", b"1e3", b" 7", b"0x10", b"nan"][i % 8]) for i in range(5000)) + b"\n"
    hpath = str(tmp_path / "hdr.tsv.gz")
    write_bgzf(hpath, hdr, 5000)
    for p_, args in ((path, ("--hash",)), (path, ("--cols", "4,1", "--hash")), (path, ("auto_detect=true", "--dump", "--types")),
                     (hpath, ("header=true", "--dump", "--types")), (hpath, ("header=true", "auto_detect=true", "--dump", "--types")),
                     (hpath, ("header=true", "column_types=VARCHAR,BIGINT,BIGINT,VARCHAR,DOUBLE", "--dump", "--types")),
                     (hpath, ("header_names=a,b,c,d,e", "--dump")), (hpath, ("header=true", "header_names=a,b", "--dump")),
                     (hpath, ("column_types=VARCHAR,BIGINT", "--dump"))):
        a, b = run(OUR_EXT, "read_tabix", p_, *args), run(REF_EXT, "read_tabix", p_, *args)
        assert a == b, (args, a[1][:300], b[1][:300])
        assert a[0] == 0 or "column_types" in args[0]
    assert run(OUR_EXT, "read_tabix", "")[1] == run(REF_EXT, "read_tabix", "")[1]
    assert run(OUR_EXT, "read_tabix", str(tmp_path / "missing.gz")) == run(REF_EXT, "read_tabix", str(tmp_path / "missing.gz"))
    rc, out = run(OUR_EXT, "read_tabix", path, "region=chr1:1-100")
    assert rc != 0 and "region queries are not served" in out
