"""N1: BGZF compression on the GPU (dh_bgzf_deflate_host / _fd, the extension's bgzip(); reference src/bgzip.c:209-273 ->
bgzf_write / bgzf_close, htslib/bgzf.c:624-683, 2138-2150).

No two DEFLATE compressors write the same bits, so the oracle here is the inverse: every stream we write must inflate — by zlib
(python's gzip, which checks each member's CRC-32 and ISIZE), by this library's own GPU inflate and by the REFERENCE's readers —
to exactly the input, in blocks cut where htslib's bgzf_write cuts them (0xff00 input bytes each, the 28-byte EOF marker last)."""
import gzip
import os
import subprocess
import sys

import numpy as np
import pytest

import duckhts_b200 as dh
from tests import util

pytestmark = pytest.mark.gpu

ROOT = util.ROOT
HOST = os.path.join(ROOT, "harness", "_bin", "fake_duckdb")
REF_EXT = os.path.join(ROOT, "oracle", "_ref", "libduckhts_ref.so")
OUR_EXT = os.path.join(ROOT, "duckhts_b200", "lib", "duckhts.duckdb_extension.so")
BLOCK = 0xFF00


def check_stream(ctx, data, out, stored=False):
    """out = our BGZF stream of `data`: framing as htslib writes it, and three inflaters agree on the bytes."""
    data = np.ascontiguousarray(data, dtype=np.uint8)
    out = np.ascontiguousarray(out, dtype=np.uint8)
    blocks, end = dh.bgzf_walk(out)
    assert end == out.size
    nb = (data.size + BLOCK - 1) // BLOCK
    assert len(blocks) == nb + 1
    assert bytes(out[-28:]) == util.BGZF_EOF and blocks[-1][1] == 28 and blocks[-1][2] == 0
    for i, b in enumerate(blocks[:-1]):
        assert b[2] == (BLOCK if i < nb - 1 else data.size - BLOCK * (nb - 1)), (i, b)
        assert 26 <= b[1] <= 65536
        if stored:
            assert b[1] == b[2] + 18 + 5 + 8
    assert gzip.decompress(out.tobytes()) == data.tobytes()                    # zlib: every member's CRC-32 and ISIZE checked
    back, st = dh.bgzf_inflate(ctx, out)                                        # our own inflate kernels
    assert back.size == data.size and np.array_equal(back, data) and st.bad_block == -1 and st.eof_marker == 1


def corpus():
    rng = np.random.default_rng(5)
    text = ("\n".join("chr%d\t%d\t%d\tgene_%d\t%s\t+" % (i % 23, 1000 * i, 1000 * i + 500, i, "ACGT"[i % 4] * (i % 37)) for i in range(60000))).encode()
    yield "empty", np.zeros(0, dtype=np.uint8)
    yield "one_byte", np.frombuffer(b"x", dtype=np.uint8)
    yield "three_bytes", np.frombuffer(b"abc", dtype=np.uint8)
    yield "short_run", np.frombuffer(b"aaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaaab", dtype=np.uint8)
    yield "exactly_one_block", rng.integers(65, 70, BLOCK, dtype=np.uint8)
    yield "one_block_plus_one", rng.integers(65, 70, BLOCK + 1, dtype=np.uint8)
    yield "zeros", np.zeros(300000, dtype=np.uint8)                             # length-258 matches at distance 1
    yield "random", rng.integers(0, 256, 1 << 20, dtype=np.uint8)              # incompressible: stored blocks
    yield "text", np.frombuffer(text, dtype=np.uint8)
    yield "period_40000", np.tile(rng.integers(0, 256, 40000, dtype=np.uint8), 8)      # repeats at distance 40000: beyond DEFLATE's window, must not be used
    yield "period_20000", np.tile(rng.integers(0, 256, 20000, dtype=np.uint8), 16)     # repeats at distance 20000: far matches (found only while the hash table remembers)
    yield "period_1500", np.tile(rng.integers(0, 256, 1500, dtype=np.uint8), 200)      # repeats at distance 1500: long matches
    yield "two_symbols", rng.integers(0, 2, 200000, dtype=np.uint8)            # a literal alphabet of two codes
    yield "one_symbol_no_match", np.frombuffer(b"zzz", dtype=np.uint8)
    skew = rng.choice(np.arange(256, dtype=np.uint8), size=400000, p=np.r_[[0.9], np.full(255, 0.1 / 255)])
    yield "skewed", skew                                                         # a Huffman tree deeper than 15 before the repair


@pytest.mark.parametrize("name,data", list(corpus()), ids=[n for n, _ in corpus()])
def test_round_trip(ctx, name, data):
    out, st = dh.bgzf_deflate(ctx, data)
    check_stream(ctx, data, out)
    assert st.in_bytes == data.size and st.out_bytes == out.size and st.n_blocks == (data.size + BLOCK - 1) // BLOCK
    if name in ("zeros", "text", "two_symbols", "skewed", "period_1500"):
        assert out.size < 0.7 * data.size, (name, out.size, data.size)
    if name == "random":
        assert out.size <= data.size + 31 * st.n_blocks + 28                   # never worse than stored blocks
    out0, _ = dh.bgzf_deflate(ctx, data, level=0)
    check_stream(ctx, data, out0, stored=True)


def test_bam_and_bcf_payloads_and_the_ratio_against_zlib(ctx):
    """The bytes the scan path cares about: an inflated BAM and an inflated BCF, recompressed.  zlib level 6 (what htslib's bgzf_write
    runs by default) is the yardstick for the size."""
    import zlib
    from tests.test_bcf_gpu import synth_bcf
    for kind, img in (("bam", util.synth_bam_bytes(n_records=60000, seed=12, paired=True)),
                      ("bcf", np.fromfile(synth_bcf(300, 2504, 0x63, False), dtype=np.uint8))):
        raw, _ = dh.bgzf_inflate(ctx, img)
        out, st = dh.bgzf_deflate(ctx, raw)
        check_stream(ctx, raw, out)
        z = sum(len(zlib.compress(raw[o:o + BLOCK].tobytes(), 6)) + 20 for o in range(0, raw.size, BLOCK))
        print(f"{kind}: {raw.size} -> ours {out.size} ({out.size / raw.size:.3f}), zlib -6 {z} ({z / raw.size:.3f})", file=sys.stderr)
        assert out.size < 1.35 * z, (kind, out.size, z)


def test_a_recompressed_bam_scans_like_the_original(ctx, tmp_path):
    """inflate(BAM) -> our deflate -> a BAM whose every record the scan (ours, and the reference's) returns unchanged."""
    from oracle import dh_oracle_py as orc
    from tests.test_bam_gpu import TAGS
    img = util.synth_bam_bytes(n_records=30000, seed=13, paired=True)
    raw, _ = dh.bgzf_inflate(ctx, img)
    out, _ = dh.bgzf_deflate(ctx, raw)
    rc, infl, _ = orc.bgzf_inflate(img)
    exp = orc.bam_decode(infl, TAGS)
    util.assert_bam_equal(dh.bam_scan(ctx, out, tags=TAGS), exp)
    if os.path.exists(HOST) and os.path.exists(REF_EXT):
        a, b = str(tmp_path / "orig.bam"), str(tmp_path / "ours.bam")
        img.tofile(a)
        out.tofile(b)
        ha = subprocess.run([HOST, REF_EXT, "read_bam", a, "--hash"], stdout=subprocess.PIPE, timeout=600).stdout
        hb = subprocess.run([HOST, REF_EXT, "read_bam", b, "--hash"], stdout=subprocess.PIPE, timeout=600).stdout
        assert ha == hb and ha.startswith(b"rows\t30000")


def test_several_launch_groups_and_the_fd_form(tmp_path):
    """Groups of 3 blocks (DH_DEFLATE_GROUP is read once per process, hence the child): 40 blocks = 14 launches through the three slots."""
    rng = np.random.default_rng(9)
    data = np.concatenate([rng.integers(65, 91, 17 * BLOCK, dtype=np.uint8), np.zeros(5 * BLOCK, dtype=np.uint8), rng.integers(0, 256, 18 * BLOCK - 777, dtype=np.uint8)])
    src, dst = str(tmp_path / "in.bin"), str(tmp_path / "out.gz")
    data.tofile(src)
    code = ("import os, duckhts_b200 as dh\n"
            "ctx = dh.Context(0)\n"
            f"fi, fo = os.open({src!r}, os.O_RDONLY), os.open({dst!r}, os.O_WRONLY | os.O_CREAT | os.O_TRUNC, 0o644)\n"
            "st = dh.bgzf_deflate_fd(ctx, fi, fo)\n"
            "os.close(fi); os.close(fo)\n"
            "print(st.in_bytes, st.out_bytes, st.n_blocks)\n")
    env = dict(os.environ, DH_DEFLATE_GROUP="3", PYTHONPATH=ROOT)
    p = subprocess.run([sys.executable, "-c", code], stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=env, timeout=600)
    assert p.returncode == 0, p.stderr.decode()[-2000:]
    out = np.fromfile(dst, dtype=np.uint8)
    a, b, c = (int(x) for x in p.stdout.split())
    assert (a, b, c) == (data.size, out.size, 40)
    assert gzip.decompress(out.tobytes()) == data.tobytes()
    blocks, end = dh.bgzf_walk(out)
    assert end == out.size and [x[2] for x in blocks[:-1]] == [BLOCK] * 39 + [BLOCK - 777]


@pytest.mark.skipif(not (os.path.exists(HOST) and os.path.exists(REF_EXT)), reason="prebuilt fake host / reference extension absent")
def test_bgzip_table_function_against_the_reference(tmp_path):
    def run(ext, func, *args):
        p = subprocess.run([HOST, ext, func, *args, "--dump"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=600)
        return p.returncode, p.stdout.decode()

    rng = np.random.default_rng(3)
    text = ("\n".join("%d\t%s" % (i, "".join("ACGT"[int(x)] for x in rng.integers(0, 4, 40))) for i in range(30000)) + "\n").encode()
    src = str(tmp_path / "table.tsv")
    open(src, "wb").write(text)
    ours, theirs = str(tmp_path / "ours.gz"), str(tmp_path / "theirs.gz")
    rc_o, out_o = run(OUR_EXT, "bgzip", src, "output_path=" + ours)
    rc_r, out_r = run(REF_EXT, "bgzip", src, "output_path=" + theirs)
    assert rc_o == 0 and rc_r == 0
    row_o, row_r = out_o.splitlines()[1].split("\t"), out_r.splitlines()[1].split("\t")
    assert out_o.splitlines()[0] == out_r.splitlines()[0]                                      # success, output_path, bytes_in, bytes_out
    assert row_o[0] == row_r[0] == "true" and row_o[1] == ours and row_o[2] == row_r[2] == str(len(text))
    assert int(row_o[3]) == os.path.getsize(ours)
    # same blocks in inflated space as htslib wrote
    bo, _ = dh.bgzf_walk(np.fromfile(ours, dtype=np.uint8))
    bt, _ = dh.bgzf_walk(np.fromfile(theirs, dtype=np.uint8))
    assert [b[2] for b in bo] == [b[2] for b in bt]
    # the reference's bgunzip reads ours back to the input; ours reads theirs
    back = str(tmp_path / "back.tsv")
    assert run(REF_EXT, "bgunzip", ours, "output_path=" + back)[0] == 0 and open(back, "rb").read() == text
    back2 = str(tmp_path / "back2.tsv")
    assert run(OUR_EXT, "bgunzip", theirs, "output_path=" + back2)[0] == 0 and open(back2, "rb").read() == text
    # default output path, keep := false, level := 0, errors
    assert run(OUR_EXT, "bgzip", src)[1].splitlines()[1].split("\t")[1] == src + ".gz" and os.path.exists(src)
    assert run(OUR_EXT, "bgzip", src) == run(REF_EXT, "bgzip", src)                              # output exists: same bind error
    assert run(OUR_EXT, "bgzip", str(tmp_path / "nope")) == run(REF_EXT, "bgzip", str(tmp_path / "nope"))
    assert run(OUR_EXT, "bgzip", "") == run(REF_EXT, "bgzip", "")
    rc, out = run(OUR_EXT, "bgzip", src, "overwrite=true", "level=0", "keep=false")
    assert rc == 0 and int(out.splitlines()[1].split("\t")[3]) > len(text) and not os.path.exists(src)
    assert gzip.decompress(open(src + ".gz", "rb").read()) == text
    empty = str(tmp_path / "empty")
    open(empty, "wb").close()
    assert run(OUR_EXT, "bgzip", empty)[1] == run(REF_EXT, "bgzip", empty, "overwrite=true")[1]
    assert open(empty + ".gz", "rb").read() == util.BGZF_EOF
