#!/usr/bin/env python
"""bench.py — the driver's measurement contract for duckhts_b200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One JSON line on stdout (rank 0).  A "step" is one pass of the hot path over one rank's whole input:
  * `value`     device-timed throughput with the compressed input already resident in HBM (batches left in HBM),
  * `e2e`       the same scan through the C-ABI with HOST buffers (pinned): H2D + kernels + D2H inside the timed region,
  * `ext_e2e`   the same scan through the DuckDB extension entry point (duckhts_init_c_api) under harness/fake_duckdb:
                file on disk in, filled DuckDB vectors out; the reference extension under the same host beside it,
  * `roofline`  for the dominant kernel pair (BGZF inflate), `cpu_baseline`, `clocks`, `gpu_launches`,
  * `parity`    per-column checksums of the 100 M-record scan == tiles x the unit's, device == host, both layouts.
Inputs are synthetic: a seeded unit written by the reference's htslib at BGZF level 6 (harness/_bin/gen_synth), its
record blocks tiled up to the configured size (BGZF members concatenate).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
P2_KERNEL = "bgzf_resolve3_kernel" if os.environ.get("DH_RESOLVE5") == "0" else "bgzf_resolve5_kernel"   # inflate phase 2 (bgzf_inflate3.cu)
sys.path.insert(0, ROOT)

from tests import util  # noqa: E402  (input generation helpers; no oracle import at module level)

HARNESS = os.path.join(ROOT, "harness", "_bin")
CACHE = util.CACHE
FAKE_HOST = os.path.join(HARNESS, "fake_duckdb")
REF_EXT = os.path.join(ROOT, "oracle", "_ref", "libduckhts_ref.so")
OUR_EXT = os.path.join(ROOT, "duckhts_b200", "lib", "duckhts.duckdb_extension.so")
M64 = (1 << 64) - 1


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# ---------------------------------------------------------------------------------------------------------
# synthetic input
# ---------------------------------------------------------------------------------------------------------
def gen_unit(kind, n_records, seed, paired, procs):
    """Seeded unit written through the reference htslib at level 6, by `procs` parallel writers (by contig
    groups; content depends only on (seed, contig)).  Returns the path of the concatenated file."""
    os.makedirs(CACHE, exist_ok=True)
    path = os.path.join(CACHE, f"unit_{kind}_{n_records}_{seed}_{int(paired)}.{'bam' if kind == 'bam' else 'bgzf'}")
    if os.path.exists(path):
        return path
    gen = os.path.join(HARNESS, "gen_synth")
    if not os.path.exists(gen):
        raise FileNotFoundError(f"{gen} missing — run __graft_entry__.build() where /root/reference exists")
    K = max(1, min(procs, 22))
    t0 = time.time()
    ps, parts = [], []
    for p in range(K):
        part = f"{path}.part{p}"
        cmd = [gen, kind, part, str(n_records), hex(seed), "--part", f"{p}/{K}"]
        if paired:
            cmd.append("--paired")
        if p > 0:
            cmd.append("--no-header")
        if p < K - 1:
            cmd.append("--no-eof")
        ps.append(subprocess.Popen(cmd))
        parts.append(part)
    for p in ps:
        if p.wait() != 0:
            raise RuntimeError("gen_synth failed")
    with open(path + ".tmp", "wb") as out:
        for part in parts:
            with open(part, "rb") as f:
                while True:
                    b = f.read(1 << 24)
                    if not b:
                        break
                    out.write(b)
            os.unlink(part)
    os.replace(path + ".tmp", path)
    log(f"[bench] generated {path}: {os.path.getsize(path) / 1e6:.1f} MB in {time.time() - t0:.1f}s with {K} writers")
    return path


def ensure_bai(path):
    """BAI next to the unit, built by the reference's htslib (sam_index_build through harness/ref_dump)."""
    exe = os.path.join(HARNESS, "ref_dump")
    if not os.path.exists(path + ".bai") and os.path.exists(exe):
        subprocess.check_call([exe, "index", path, "0"])
    return os.path.exists(path + ".bai")


def tile_image(dh, unit, target_bytes, header_blocks=0):
    """Pinned host image = unit's header blocks + its record blocks repeated to >= target_bytes + EOF block."""
    blocks, end = dh.bgzf_walk(unit)
    assert end == unit.size
    has_eof = blocks[-1][1] == 28 and blocks[-1][2] == 0
    body_lo = blocks[header_blocks][0] if header_blocks < len(blocks) else end
    body_hi = blocks[-1][0] if has_eof else end
    body = unit[body_lo:body_hi]
    reps = max(1, int(round(target_bytes / max(body.size, 1))))
    total = body_lo + reps * body.size + 28
    from duckhts_b200.bgzf import PinnedBuffer

    pin = PinnedBuffer(total)
    a = pin.array
    a[:body_lo] = unit[:body_lo]
    for r in range(reps):
        a[body_lo + r * body.size: body_lo + (r + 1) * body.size] = body
    a[total - 28:] = np.frombuffer(util.BGZF_EOF, dtype=np.uint8)
    return pin, reps


# ---------------------------------------------------------------------------------------------------------
# clocks
# ---------------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device, self.rows, self.p = device, [], None

    def __enter__(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                       "-i", str(self.device)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.p = None
        return self

    def _read(self):
        for line in self.p.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def __exit__(self, *a):
        if self.p:
            self.p.terminate()
            try:
                self.p.wait(timeout=2)
            except Exception:
                self.p.kill()

    def summary(self):
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 9:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ---------------------------------------------------------------------------------------------------------
# the table functions under the fake DuckDB host (both extensions, same binary, same arguments)
# ---------------------------------------------------------------------------------------------------------
WORKLOAD = ("read_bam full-column scan (13 core columns + typed tags NM, MD, AS) of a synthetic 100M-record 2x150 bp paired "
            "WGS-like BAM (BASELINE configs[2]); BGZF inflate GB/s of the same bytes reported alongside (configs[1])")
TAGS = [(b"NM", "i"), (b"MD", "Z"), (b"AS", "i")]
STD = ("AM AS BC BQ BZ CB CC CG CM CO CP CQ CR CS CT CY E2 FI FS FZ H0 H1 H2 HI IH LB MC MD MI ML MM MN MQ NH NM OA OC OP OQ OX PG "
       "PQ PT PU Q2 QT QX R2 RG RX SA SM TC TS U2 UQ").split()
# schema columns of read_bam(standard_tags := true) that bench.py's projection names: 13 core + NM, MD, AS
EXT_COLS = ",".join(str(c) for c in list(range(13)) + [13 + STD.index("NM"), 13 + STD.index("MD"), 13 + STD.index("AS")])


def host_run(ext, func, path, *args, threads=1, env=None, timeout=900, repeat=1):
    """fake_duckdb EXT FUNC PATH ... --time -> {"rows", "seconds" (first run of the process: cold), "best_seconds" (fastest of `repeat` runs in the
    same process: a session that has run the function before), "bind_seconds", ...}"""
    e = dict(os.environ)
    e.update(env or {})
    out = subprocess.check_output([FAKE_HOST, ext, func, path, *args, "--threads", str(threads), "--time", "--repeat", str(repeat)], stderr=subprocess.DEVNULL, text=True,
                                  env=e, timeout=timeout)
    return json.loads(out.strip().splitlines()[-1])


def ref_read_bam(path, threads, steps, warmup):
    """The unchanged reference extension (oracle/_ref/libduckhts_ref.so = /root/reference/src/*.c + htslib 1.23, zlib) scanning
    `path` under the fake host: with the .bai next to it the reference runs its contig-parallel mode (src/bam_reader.c:577-585,
    689-716: up to 16 scan threads x 2 htslib threads); without, one scan thread."""
    for _ in range(warmup):
        host_run(REF_EXT, "read_bam", path, "standard_tags=true", "--cols", EXT_COLS, threads=threads)
    res = [host_run(REF_EXT, "read_bam", path, "standard_tags=true", "--cols", EXT_COLS, threads=threads, repeat=2) for _ in range(max(steps, 1))]
    secs = float(np.mean([r["seconds"] for r in res]))
    warm = float(np.mean([r["best_seconds"] for r in res]))
    return {"value": res[0]["rows"] / 1e6 / secs, "unit": "Mrecords/s", "rows": res[0]["rows"], "seconds": secs, "threads": res[0]["threads"],
            "max_threads": res[0]["max_threads"], "warm_value": res[0]["rows"] / 1e6 / warm, "warm_seconds": warm}


def run_ref_bench(mode, path, threads, limit):
    exe = os.path.join(HARNESS, "ref_bench")
    out = subprocess.check_output([exe, mode, path, str(threads), str(limit)], text=True)
    return json.loads(out.strip().splitlines()[-1])


def common_config(args):
    """The same dict in both arms: what is scanned and how it is asked for."""
    return {"workload": WORKLOAD,
            "input": f"seeded unit ({args.unit_records} paired 2x150 records, 22 contigs, tags NM MD AS RG, written by the reference htslib at BGZF level 6); "
                     "our arm tiles its record blocks to the target size, the reference arm scans the unit itself (bounded sample)",
            "projection": "13 core columns + NM, MD, AS (read_bam(path, standard_tags := true), schema columns " + EXT_COLS + ")",
            "records_per_gpu_target": int(args.records), "sharding": "contiguous virtual-offset range per GPU, no collective"}


def bench_bgzip(dh, ctx, unit_path, tiles, ncpu):
    """N1: bgzip of the inflated unit BAM stream tiled `tiles` times (a file in the page cache -> /dev/null), dh_bgzf_deflate_fd
    against the reference's bgzip() (htslib bgzf_write + bgzf_mt with all cores) on the same file."""
    import ctypes as C
    from duckhts_b200._lib import InflateStats, lib as L
    raw_path = os.path.join(CACHE, "bench_bgzip_in.bin")
    out = {"what": "bgzip of the inflated record stream of the unit BAM (file in the page cache -> compressed stream to /dev/null): pread, H2D, "
                   "bgzf_deflate_kernel + pack, D2H, write(2)"}
    try:
        fd_in, fd_out = os.open(unit_path, os.O_RDONLY), os.open(raw_path + ".unit", os.O_WRONLY | os.O_CREAT | os.O_TRUNC, 0o644)
        sti = InflateStats()
        rc = L().dh_bgzf_inflate_fd(ctx.h, fd_in, fd_out, C.byref(sti))
        os.close(fd_in)
        os.close(fd_out)
        assert rc == 0, rc
        with open(raw_path, "wb") as f:
            unit = open(raw_path + ".unit", "rb").read()
            for _ in range(tiles):
                f.write(unit)
        os.unlink(raw_path + ".unit")
        best, st = None, None
        for _ in range(3):
            fi, fo = os.open(raw_path, os.O_RDONLY), os.open("/dev/null", os.O_WRONLY)
            t0 = time.perf_counter()
            st = dh.bgzf_deflate_fd(ctx, fi, fo)
            dt = time.perf_counter() - t0
            os.close(fi)
            os.close(fo)
            best = dt if best is None else min(best, dt)
        out.update({"metric": "bgzip_input_GBps_host_to_host", "value": st.in_bytes / 1e9 / best, "unit": "GB/s", "seconds": best, "input_gb": st.in_bytes / 1e9,
                    "output_gb": st.out_bytes / 1e9, "ratio": st.out_bytes / st.in_bytes, "blocks": int(st.n_blocks), "kernel_ms": st.kernel_ms,
                    "kernel_input_GBps": st.in_bytes / 1e9 / (st.kernel_ms / 1e3)})
        if os.path.exists(REF_EXT) and os.path.exists(FAKE_HOST):
            ref_out = raw_path + ".ref.gz"
            r = host_run(REF_EXT, "bgzip", raw_path, "output_path=" + ref_out, f"threads={ncpu}", "overwrite=true", threads=1)
            out["reference"] = {"value": st.in_bytes / 1e9 / r["seconds"], "unit": "GB/s", "seconds": r["seconds"], "htslib_threads": ncpu,
                                "output_gb": os.path.getsize(ref_out) / 1e9, "ratio": os.path.getsize(ref_out) / st.in_bytes,
                                "what": "the unchanged reference extension's bgzip() (bgzf_write, zlib default level, bgzf_mt) on the same file, output to a file in /tmp"}
            os.unlink(ref_out)
    except Exception as e:  # noqa: BLE001
        out["error"] = str(e)[:300]
    if os.path.exists(raw_path):
        os.unlink(raw_path)
    return out


def bench_tabix(dh, ctx):
    """N4: read_tabix over a bgzip'ed 5-column table, every column, through both extensions under the fake host."""
    out = {"what": "read_tabix(path), all 5 VARCHAR columns, 10,000,000 lines (a seeded 200,000-line unit tiled x50, compressed by dh_bgzf_deflate); "
                   "file -> filled DuckDB vectors under harness/fake_duckdb, one scan thread (the reference's read_tabix is single-threaded)"}
    path = os.path.join(CACHE, "bench_tabix.tsv.gz")
    try:
        rng = np.random.default_rng(0x7ab1)
        pos = np.cumsum(rng.integers(1, 400, 200000))
        unit = "".join("chr%d\t%d\t%d\tfeature_%d\t%.3f\n" % (1 + i * 22 // 200000, p, p + 150, i, s) for i, (p, s) in enumerate(zip(pos.tolist(), rng.random(200000).tolist()))).encode()
        text = np.frombuffer(unit * 50, dtype=np.uint8)
        comp, st = dh.bgzf_deflate(ctx, text)
        comp.tofile(path)
        n_lines = 200000 * 50
        r = host_run(OUR_EXT, "read_tabix", path, threads=1, repeat=3)
        assert r["rows"] == n_lines, r
        out.update({"metric": "read_tabix_Mlines_per_s", "value": n_lines / 1e6 / r["best_seconds"], "unit": "Mlines/s", "seconds": r["best_seconds"],
                    "cold_seconds": r["seconds"], "lines": n_lines, "text_gb": text.size / 1e9, "compressed_gb": comp.size / 1e9,
                    "text_GBps": text.size / 1e9 / r["best_seconds"]})
        if os.path.exists(REF_EXT):
            q = host_run(REF_EXT, "read_tabix", path, threads=1)
            out["reference"] = {"value": q["rows"] / 1e6 / q["seconds"], "unit": "Mlines/s", "seconds": q["seconds"], "rows": q["rows"]}
    except Exception as e:  # noqa: BLE001
        out["error"] = str(e)[:300]
    if os.path.exists(path):
        os.unlink(path)
    return out


def gen_bcf(n_sites, n_samples, seed):
    os.makedirs(CACHE, exist_ok=True)
    path = os.path.join(CACHE, f"bench_bcf_{n_sites}_{n_samples}_{seed}.bcf")
    if not os.path.exists(path):
        gen = os.path.join(HARNESS, "gen_bcf")
        t0 = time.time()
        subprocess.check_call([gen, path + ".tmp", str(n_sites), str(n_samples), hex(seed)])
        os.replace(path + ".tmp", path)
        log(f"[bench] generated {path}: {os.path.getsize(path) / 1e6:.1f} MB in {time.time() - t0:.1f}s")
    return path


def bench_bcf(dh, ctx, args, rank, barrier, rmax, rsum):
    """BASELINE configs[4]: tidy FORMAT/GT + INFO decode of 1,000,000 sites x 2,504 samples, every column: sites/s and tidy rows/s."""
    ns = 2504
    if rank == 0:
        gen_bcf(args.bcf_sites, ns, 0x5eed)
    barrier()
    path = gen_bcf(args.bcf_sites, ns, 0x5eed)

    raw = np.fromfile(path, dtype=np.uint8)
    # bcf_hdr_write flushes the BGZF stream after the header (htslib/vcf.c:1802): the record blocks can be tiled like the BAM unit's
    blocks, _ = dh.bgzf_walk(raw)
    import zlib
    first = zlib.decompressobj(-15).decompress(raw[blocks[0][0] + 18: blocks[0][0] + blocks[0][1] - 8].tobytes())   # input preparation only
    hdr_bytes = 9 + int.from_bytes(first[5:9], "little")
    hb, acc = 0, 0
    while acc < hdr_bytes:
        acc += blocks[hb][2]
        hb += 1
    pin, reps = tile_image(dh, raw, raw.size * max(1, args.bcf_tiles), header_blocks=hb)
    n_sites = args.bcf_sites * reps
    bf = dh.BcfFile(ctx, pin)
    out = {"workload": f"read_bcf tidy_format scan, all columns (7 core + {len(bf.info)} INFO + SAMPLE_ID + {len(bf.format)} FORMAT), "
                       f"{n_sites} biallelic sites x {ns} samples (BASELINE configs[4]; a seeded {args.bcf_sites}-site unit written by the "
                       f"reference htslib at BGZF level 6, its record blocks tiled x{reps})",
           "sites": n_sites, "compressed_mb": pin.array.size / 1e6}

    def once():
        bf.scan_begin()
        n = bf.drain()
        return n, bf.stats()

    bf.set_mode(device_input=True, device_output=True, compact=True)
    for _ in range(min(args.warmup, 2)):
        n, st = once()
    assert st.status == 0 and n == n_sites, (st.status, n, n_sites)
    barrier()
    ctx.timer_start()
    k = [0.0, 0.0, 0.0]
    for _ in range(args.steps):
        n, st = once()
        k = [k[0] + st.inflate_ms, k[1] + st.walk_ms, k[2] + st.decode_ms]
    ms = rmax(ctx.timer_stop()) / args.steps
    sites = rsum(float(n))
    out.update({"value": sites / (ms / 1e3), "unit": "sites/s", "rows_per_s": sites * ns / (ms / 1e3), "ms_per_step": ms,
                "inflated_mb": st.inflated_bytes / 1e6,
                "kernel_ms_per_step": {"bgzf_inflate": k[0] / args.steps, "bcf_walk": k[1] / args.steps, "bcf_decode": k[2] / args.steps}})
    def e2e(compact, steps):
        bf.set_mode(device_input=False, device_output=False, compact=compact)
        walls = []
        for it in range(1 + steps):
            barrier()
            t0 = time.perf_counter()
            n, s = once()
            if it:
                walls.append(rmax((time.perf_counter() - t0) * 1e3))
        w = float(np.mean(walls))
        return {"value": sites / (w / 1e3), "unit": "sites/s", "rows_per_s": sites * ns / (w / 1e3), "ms_per_step": w,
                "h2d_bytes_per_step": int(s.h2d_bytes), "d2h_bytes_per_step": int(s.d2h_bytes), "d2h_bytes_per_site": s.d2h_bytes / max(n, 1),
                "d2h_gbs": s.d2h_bytes / 1e9 / (w / 1e3)}

    out["e2e"] = e2e(True, min(args.steps, 2))
    out["e2e"]["layout"] = "compact FORMAT cells (GT 4 bytes, numeric lists as 2-byte counts + children)"
    out["e2e"]["duckdb_layout"] = e2e(False, 1)
    bf.close()
    del pin
    if rank == 0 and os.path.exists(FAKE_HOST):
        # the table function itself (duckhts_init_c_api under the fake host) on the unit; the unchanged reference extension on a bounded sample
        try:
            ncpu = os.cpu_count() or 1
            r = host_run(OUR_EXT, "read_bcf", path, "tidy_format=true", threads=min(ncpu, 16), repeat=3)
            out["ext_e2e"] = {"value": r["rows"] / ns / r["best_seconds"], "unit": "sites/s", "rows_per_s": r["rows"] / r["best_seconds"], "threads": r["threads"],
                              "sites": r["rows"] // ns, "seconds": r["best_seconds"], "bind_seconds": r["best_bind_seconds"],
                              "cold_value": r["rows"] / ns / r["seconds"], "cold_seconds": r["seconds"]}
        except Exception as e:  # noqa: BLE001
            out["ext_e2e"] = {"error": str(e)[:300]}
        if os.path.exists(REF_EXT):
            sample = gen_bcf(min(args.bcf_sites, 3000), ns, 0x5eed)
            r = host_run(REF_EXT, "read_bcf", sample, "tidy_format=true", threads=1)
            out["cpu_baseline"] = {"value": r["rows"] / ns / r["seconds"], "unit": "sites/s", "rows_per_s": r["rows"] / r["seconds"], "cores": 1, "kind": "reference",
                                   "sample": f"{r['rows'] // ns} sites x {ns} samples; the unchanged reference extension (read_bcf, tidy_format) under harness/fake_duckdb, "
                                             "1 scan thread (no index => the reference scans sequentially)"}
    return out


def bench_contig_shards(dh, ctx, args, unit_path, rank, world, barrier, rmax, rsum):
    """BASELINE configs[3]: BAI-driven scan, contig groups balanced by compressed bytes, one group per GPU (no collective)."""
    if rank == 0:
        ensure_bai(unit_path)
    barrier()
    if not os.path.exists(unit_path + ".bai"):
        return None
    from duckhts_b200.bgzf import PinnedBuffer

    raw = np.fromfile(unit_path, dtype=np.uint8)
    pin = PinnedBuffer(raw.size)
    pin.array[:] = raw
    bai = np.fromfile(unit_path + ".bai", dtype=np.uint8)
    bf = dh.BamFile(ctx, pin)
    bf.set_projection(None, TAGS)
    bf.set_mode(compact=True)
    ix = dh.Index(ctx, bai)
    bounds = bf.split_contigs(ix, world)
    walls, rows = [], 0
    for it in range(1 + args.steps):
        barrier()
        t0 = time.perf_counter()
        bf.scan_contigs(ix, bounds[rank], bounds[rank + 1])
        rows = bf.drain()
        if it:
            walls.append(rmax((time.perf_counter() - t0) * 1e3))
    st = bf.stats()
    total = rsum(float(rows))
    w = float(np.mean(walls))
    bf.close()
    ix.close()
    return {"workload": "read_bam over contig groups from the BAI index (chr1-chr22), host buffers in and out; the file is split over the GPUs "
                        "(strong scaling on the 4M-record unit: the tiled 100M image is not coordinate-sorted, so it has no BAI)", "value": total / 1e6 / (w / 1e3),
            "unit": "Mrecords/s", "ms_per_step": w, "records": int(total), "contig_bounds": bounds, "status": int(st.status)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--records", type=float, default=100e6, help="records per GPU (BASELINE configs[2]: 100M)")
    ap.add_argument("--unit-records", type=int, default=4000000)
    ap.add_argument("--seed", type=lambda s: int(s, 0), default=0x9E3779B97F4A7C15)
    ap.add_argument("--bcf-sites", type=int, default=20000, help="sites of the seeded 2,504-sample BCF unit (0 = skip the sub-benchmark)")
    ap.add_argument("--bcf-tiles", type=int, default=50, help="how many times the unit's record blocks are repeated in the scanned image (50 x 20,000 = 1M sites)")
    ap.add_argument("--no-ext", action="store_true", help="skip the extension-level runs (fake DuckDB host)")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    ncpu = os.cpu_count() or 1
    config = common_config(args)

    if args.impl == "reference":
        # The reference's own CPU implementation of the path, stock code: the unchanged extension under the fake DuckDB host, the .bai
        # next to the file, all the threads its global init allows (min(n_contigs, 16) scan threads, 2 htslib threads each).
        if rank != 0:
            return 0
        unit_path = gen_unit("bam", args.unit_records, args.seed, True, ncpu)
        have_bai = ensure_bai(unit_path)
        thr = min(16, ncpu)
        par = ref_read_bam(unit_path, thr, args.steps, min(args.warmup, 1))
        one = ref_read_bam(unit_path, 1, 1, 0)
        extra = {}
        try:
            r = run_ref_bench("sam_read1", unit_path, ncpu, 2000000)
            extra["htslib_loop"] = {"value": r["records"] / 1e6 / r["seconds"], "unit": "Mrecords/s", "cores": ncpu,
                                    "what": "harness/ref_bench: sam_read1 + the conversions of src/bam_reader.c:779-966, one consumer, hts_set_threads(all cores), first 2M records"}
        except Exception as e:  # noqa: BLE001
            extra["htslib_loop"] = {"error": str(e)[:200]}
        v = par["value"]
        cb = {"value": v, "unit": "Mrecords/s", "cores": ncpu, "kind": "reference",
              "sample": f"the whole {par['rows']}-record unit BAM per step ({'with' if have_bai else 'WITHOUT'} its .bai => contig-parallel mode), unchanged reference "
                        f"extension under harness/fake_duckdb, {par['threads']} scan threads (max_threads {par['max_threads']}) x 2 htslib threads, zlib",
              "threads_1": one, "threads_16": par}
        cb.update(extra)
        line = {"impl": "reference", "metric": "read_bam_Mrecords_per_s", "value": v, "unit": "Mrecords/s", "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * par["seconds"], "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": config, "cpu_baseline": cb,
                "e2e": {"value": v, "unit": "Mrecords/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0, "host_cores": ncpu}
        print(json.dumps(line), flush=True)
        return 0

    import torch
    import torch.distributed as dist

    import duckhts_b200 as dh
    from duckhts_b200.bgzf import PinnedBuffer  # noqa: F401

    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(local_rank)

    def reduce(x, op):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=f"cuda:{local_rank}")
        dist.all_reduce(t, op=op)
        return float(t.item())

    rmax = lambda x: reduce(x, dist.ReduceOp.MAX) if world > 1 else x  # noqa: E731
    rsum = lambda x: reduce(x, dist.ReduceOp.SUM) if world > 1 else x  # noqa: E731

    def sum_u64(vals):
        """Sum of 64-bit checksums over the ranks, modulo 2^64 (32-bit halves ride in int64 words)."""
        vals = [int(v) & M64 for v in vals]
        if world == 1:
            return vals
        t = torch.tensor([[v & 0xffffffff, v >> 32] for v in vals], dtype=torch.int64, device=f"cuda:{local_rank}")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return [(int(lo) + (int(hi) << 32)) & M64 for lo, hi in t.tolist()]

    torch.cuda.set_device(local_rank)
    if rank == 0:
        gen_unit("bam", args.unit_records, args.seed, True, ncpu)
    barrier()
    unit_path = gen_unit("bam", args.unit_records, args.seed, True, ncpu)
    unit = np.fromfile(unit_path, dtype=np.uint8)

    ctx = dh.Context(local_rank)
    # ---- the unit: where its alignment blocks start, how many records it holds, its per-column checksums ------------------------
    with dh.BamFile(ctx, unit) as ub:
        first_voff = ub.first_voffset
        ub.set_projection(None, TAGS)
        ub.set_mode(device_input=True, device_output=True, compact=True)
        ub.scan_begin()
        unit_rows, unit_sums = ub.drain_checksums()
        unit_sums = [int(x) for x in unit_sums]
    blocks, _ = dh.bgzf_walk(unit)
    header_blocks = next(i for i, b in enumerate(blocks) if b[0] == first_voff >> 16)
    assert first_voff & 0xffff == 0, "htslib flushes after the header: first record starts a block"
    body_bytes = blocks[-1][0] - blocks[header_blocks][0]
    reps = max(1, int(round(args.records / unit_rows)))
    pin, reps = tile_image(dh, unit, reps * body_bytes, header_blocks=header_blocks)
    want_sums = [(s * reps) & M64 for s in unit_sums]
    h2d_gbs, d2h_gbs = ctx.measure_pcie(1 << 30)
    bf = dh.BamFile(ctx, pin)
    bf.set_projection(None, TAGS)
    parity = {"what": "per-column 64-bit checksums (dh_bam_batch_checksum: sum over rows of a hash of every cell's value) of the tiled scan == tiles x the "
                      "unit's, modulo 2^64; the unit itself is compared cell by cell with the reference extension in tests/test_ext_gpu.py",
              "tiles": reps, "columns": 13 + len(TAGS), "checks": []}

    def check_sums(name, rows, sums, want_rows, want):
        ok = rows == want_rows and [int(x) & M64 for x in sums] == want
        parity["checks"].append({"scan": name, "rows": int(rows), "ok": bool(ok)})
        if not ok:
            bad = [c for c, (a, b) in enumerate(zip(sums, want)) if (int(a) & M64) != b]
            raise AssertionError(f"PARITY FAILURE in {name}: rows {rows} (want {want_rows}), columns that differ: {bad}")

    def scan_once():
        bf.scan_begin()
        rows = bf.drain()
        return rows, bf.stats()

    def timed_value(compact):
        """compressed image resident in HBM, batches left in HBM; CUDA events on the kernels' stream."""
        bf.set_mode(device_input=True, device_output=True, compact=compact)
        # warm-up scan 1 carries the parity check: the checksum kernel runs on every batch while it is held
        bf.scan_begin()
        rows, sums = bf.drain_checksums()
        check_sums(f"device-resident, {'compact' if compact else 'DuckDB'} layout", rows, sums, reps * unit_rows, want_sums)
        for _ in range(max(args.warmup - 1, 0)):
            rows, st = scan_once()
        assert st.status == 0 and rows == reps * unit_rows, (st.status, rows, reps * unit_rows)
        barrier()
        with ClockSampler(local_rank) as clk:
            l0 = ctx.launches
            ctx.timer_start()
            kms = [0.0, 0.0, 0.0]
            for _ in range(args.steps):
                rows, st = scan_once()
                kms = [kms[0] + st.inflate_ms, kms[1] + st.walk_ms, kms[2] + st.decode_ms]
            ms = ctx.timer_stop()
            launches = ctx.launches - l0
            barrier()
        return rmax(ms) / args.steps, rows, st, kms, launches, clk

    # ---- value: the layout the product ships (compact), then the DuckDB layout for continuity with round 1 ----------------------
    ms_per_step, rows, st, kms, launches, clk = timed_value(True)
    total_rows = rsum(float(rows))
    value = total_rows / 1e6 / (ms_per_step / 1e3)
    nblocks = st.n_blocks
    comp_bytes = pin.nbytes
    ms_duck, _, _, kms_duck, _, _ = timed_value(False)

    # ---- e2e: pinned host image in, column batches in pinned host memory out (C-ABI) ---------------------------------------------
    def timed_e2e(compact):
        bf.set_mode(device_input=False, device_output=False, compact=compact)
        walls, d2h, h2d = [], 0, 0
        for it in range(1 + args.steps):
            barrier()
            t0 = time.perf_counter()
            if it == 0:      # untimed pass: host-side checksums of what crossed PCIe
                bf.scan_begin()
                r, sums = bf.drain_checksums()
                check_sums(f"host-to-host, {'compact' if compact else 'DuckDB'} layout", r, sums, reps * unit_rows, want_sums)
                continue
            r, s = scan_once()
            walls.append(rmax((time.perf_counter() - t0) * 1e3))
            d2h, h2d = s.d2h_bytes, s.h2d_bytes
        return float(np.mean(walls)), int(d2h), int(h2d)

    e2e_step, d2h_bytes, h2d_bytes = timed_e2e(True)
    e2e_val = total_rows / 1e6 / (e2e_step / 1e3)
    e2e_duck_step, d2h_duck, _ = timed_e2e(False)

    # ---- strong scaling: ONE image, one contiguous virtual-offset range per GPU, merged rows and checksums == the whole scan ------
    strong = None
    if world > 1:
        voffs = bf.split(world)
        bf.set_mode(device_input=False, device_output=False, compact=True)
        walls = []
        for it in range(1 + args.steps):
            barrier()
            t0 = time.perf_counter()
            if voffs[rank] != dh.VOFF_MAX:
                bf.scan_begin(voffs[rank], voffs[rank + 1])
                if it == 0:
                    r, sums = bf.drain_checksums()
                else:
                    r = bf.drain()
            else:
                r, sums = 0, np.zeros(76, dtype=np.uint64)
            w = rmax((time.perf_counter() - t0) * 1e3)
            if it == 0:
                merged_rows = int(rsum(float(r)))
                merged = sum_u64(list(sums)[:13 + len(TAGS)])
                check_sums(f"strong scaling, {world} ranges merged", merged_rows, merged + [0] * (76 - len(merged)), reps * unit_rows, want_sums)
            else:
                walls.append(w)
        sw = float(np.mean(walls))
        strong = {"workload": f"ONE {reps * unit_rows}-record image cut into {world} contiguous virtual-offset ranges (dh_bam_split), one per GPU, host buffers in and out, "
                              "compact layout; merged row count and per-column checksums equal the single-GPU scan", "value": reps * unit_rows / 1e6 / (sw / 1e3),
                  "unit": "Mrecords/s", "ms_per_step": sw, "scaling": "strong", "records": reps * unit_rows}
    bf.close()

    # ---- the dominant kernel pair alone (BGZF inflate of the same bytes, device-resident): roofline + inflate GB/s ----------------
    dev = dh.BgzfDevice(ctx, pin)
    ib_comp, ib_out = dev.comp_bytes, dev.out_bytes
    for _ in range(args.warmup):
        dev.inflate()
    ctx.sync()
    assert dev.check()[0] == 0
    barrier()
    l0 = ctx.launches
    ctx.timer_start()
    for _ in range(args.steps):
        dev.inflate()
    ims = rmax(ctx.timer_stop())
    k_inf = ctx.launches - l0
    ph1, ph2, ph3 = dev.inflate_timed()      # one extra pass with events between the kernels (outside the timed region)
    dev.free()
    inflate_gbs = rsum(float(ib_out)) / 1e9 / (ims / args.steps / 1e3)
    peak, peak_src = measured_peaks()
    # roofline of the dominant kernel pair: one inflate launch group = phase 1 (bgzf_decode3_kernel) + phase 2
    # (bgzf_resolve5_kernel; DH_RESOLVE5=0: bgzf_resolve3_kernel) over <= 2^17 BGZF blocks.  Algorithmic HBM bytes of a BGZF inflate = compressed bytes read +
    # inflated bytes written (c + u per record), whatever the kernels exchange in between.
    groups = max(k_inf // 3, 1)
    alg_bytes = (ib_comp + ib_out) * args.steps / groups
    kernel_ms = ims / groups
    achieved = alg_bytes / 1e9 / (kernel_ms / 1e3)
    bi_h2d, bi_d2h = ctx.measure_pcie_bidir(1 << 30)
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "inflate_dram_traffic.json")   # dram bytes per inflated byte from the committed ncu capture
    if os.path.exists(tpath):
        traffic = json.load(open(tpath))["dram_bytes_per_inflated_byte"] * ib_out * args.steps / groups

    # ---- configs[1] end to end + the table function itself: the tiled image as a FILE ---------------------------------------------
    ext, bgunzip = None, None
    big_path = os.path.join(CACHE, f"bench_tiled_{rank}.bam")
    if rank == 0 and world == 1 and not args.no_ext and os.path.exists(FAKE_HOST):
        try:
            t0 = time.time()
            with open(big_path, "wb") as f:
                a = pin.array
                for o in range(0, a.size, 1 << 28):
                    f.write(memoryview(a[o:o + (1 << 28)]))
            log(f"[bench] wrote {big_path}: {os.path.getsize(big_path) / 1e9:.2f} GB in {time.time() - t0:.1f}s")
            # bgunzip-equivalent (configs[1]): file (page cache) -> pinned staging -> GPU inflate -> pinned staging -> write(2) to /dev/null
            import ctypes as C
            from duckhts_b200._lib import InflateStats, lib as L
            fd_in, fd_out = os.open(big_path, os.O_RDONLY), os.open("/dev/null", os.O_WRONLY)
            sti = InflateStats()
            best = None
            for _ in range(2):
                t0 = time.perf_counter()
                rc = L().dh_bgzf_inflate_fd(ctx.h, fd_in, fd_out, C.byref(sti))
                dt = time.perf_counter() - t0
                assert rc == 0, rc
                best = dt if best is None else min(best, dt)
            os.close(fd_in)
            os.close(fd_out)
            bgunzip = {"metric": "bgunzip_decompressed_GBps_host_to_host", "value": sti.out_bytes / 1e9 / best, "unit": "GB/s", "seconds": best,
                       "compressed_gb": sti.comp_bytes / 1e9, "inflated_gb": sti.out_bytes / 1e9, "blocks": int(sti.n_blocks),
                       "what": "dh_bgzf_inflate_fd (what the bgunzip table function runs): pread from the page cache, H2D, inflate + CRC, D2H, write(2) to /dev/null; "
                               "3 pinned staging slots, nothing held whole",
                       "d2h_bound_GBps": d2h_gbs}
        except Exception as e:  # noqa: BLE001
            bgunzip = {"error": str(e)[:300]}
    del pin
    if rank == 0 and world == 1 and not args.no_ext and os.path.exists(FAKE_HOST):
        ext = {"what": "read_bam(path, standard_tags := true) projected on the bench columns, through duckhts_init_c_api under harness/fake_duckdb "
                       "(bind: file -> pinned image + header; scan: GPU pipeline -> filled DuckDB vectors); seconds include bind",
               "host_threads": min(16, ncpu)}
        try:
            ensure_bai(unit_path)
            thr = min(16, ncpu)
            for name, path in (("unit_4M", unit_path), ("tiled_100M", big_path)):
                if not os.path.exists(path):
                    continue
                o = {}
                for t in (1, thr):
                    # cold = the first query of a fresh process (CUDA context, buffers and staging created inside the timed region);
                    # warm = the fastest of three runs in one process (a DuckDB session that has called the function before)
                    r = host_run(OUR_EXT, "read_bam", path, "standard_tags=true", "--cols", EXT_COLS, threads=t, repeat=3 if (t == thr or name == "unit") else 1)
                    o[f"threads_{t}"] = {"value": r["rows"] / 1e6 / r["best_seconds"], "unit": "Mrecords/s", "rows": r["rows"], "seconds": r["best_seconds"],
                                         "bind_seconds": r["best_bind_seconds"], "cold_value": r["rows"] / 1e6 / r["seconds"], "cold_seconds": r["seconds"],
                                         "cold_bind_seconds": r["bind_seconds"]}
                ext[name] = o
            if os.path.exists(REF_EXT):
                ext["reference_unit_4M"] = {"threads_1": ref_read_bam(unit_path, 1, 1, 0), f"threads_{thr}": ref_read_bam(unit_path, thr, 2, 1)}
            best = ext.get("tiled_100M", ext.get("unit_4M", {})).get(f"threads_{thr}")
            if best:
                ext["value"] = best["value"]
                ext["cold_value"] = best["cold_value"]
                ext["unit"] = "Mrecords/s"
        except Exception as e:  # noqa: BLE001
            ext["error"] = str(e)[:300]
    if os.path.exists(big_path):
        os.unlink(big_path)
    bgzip = bench_bgzip(dh, ctx, unit_path, 4, ncpu) if rank == 0 and world == 1 and not args.no_ext else None
    tabix = bench_tabix(dh, ctx) if rank == 0 and world == 1 and not args.no_ext and os.path.exists(FAKE_HOST) else None

    bcf_res = bench_bcf(dh, ctx, args, rank, barrier, rmax, rsum) if args.bcf_sites > 0 and os.path.exists(os.path.join(HARNESS, "gen_bcf")) else None
    shard_res = bench_contig_shards(dh, ctx, args, unit_path, rank, world, barrier, rmax, rsum)

    if rank == 0:
        cpu = None
        if world == 1 and os.path.exists(REF_EXT) and os.path.exists(FAKE_HOST):
            have_bai = ensure_bai(unit_path)
            thr = min(16, ncpu)
            par = (ext or {}).get("reference_unit_4M", {}).get(f"threads_{thr}") or ref_read_bam(unit_path, thr, 1, 0)
            cpu = {"value": par["value"], "unit": "Mrecords/s", "cores": ncpu, "kind": "reference",
                   "sample": f"the whole {par['rows']}-record unit BAM ({'with' if have_bai else 'WITHOUT'} its .bai => contig-parallel mode), unchanged reference extension under "
                             f"harness/fake_duckdb, {par['threads']} scan threads x 2 htslib threads, zlib"}
        rec_c = comp_bytes / rows
        run = {"records_per_gpu": int(rows), "compressed_gb_per_gpu": comp_bytes / 1e9, "inflated_gb_per_gpu": ib_out / 1e9,
               "bgzf_blocks_per_gpu": int(nblocks), "tiles": reps, "compressed_bytes_per_record": rec_c, "layout": "compact wire layout (DH_BAM_COMPACT_OUTPUT)",
               "l2": "inputs (>= 10 GB per GPU) far exceed the 126 MB L2; no flush needed"}
        line = {"metric": "read_bam_Mrecords_per_s", "value": value, "unit": "Mrecords/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "u8", "data": "synthetic", "config": config, "run": run, "parity": parity,
                "kernel_ms_per_step": {"bgzf_inflate": kms[0] / args.steps, "bam_walk": kms[1] / args.steps, "bam_decode": kms[2] / args.steps},
                "value_duckdb_layout": {"value": total_rows / 1e6 / (ms_duck / 1e3), "unit": "Mrecords/s", "ms_per_step": ms_duck,
                                        "bam_decode_ms": kms_duck[2] / args.steps},
                "inflate": {"metric": "bgzf_inflate_decompressed_GBps", "value": inflate_gbs, "unit": "GB/s",
                            "ms_per_step": ims / args.steps, "compressed_gb": ib_comp / 1e9, "inflated_gb": ib_out / 1e9,
                            "phase_ms": {"bgzf_decode3_kernel": ph1, P2_KERNEL: ph2, "fallback_filter": ph3},
                            "pcie_ingest_bound_GBps": h2d_gbs * ib_out / ib_comp},
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": traffic, "kernel": f"bgzf_decode3_kernel + {P2_KERNEL} (one inflate launch group)",
                             "algorithmic_bytes_per_launch": alg_bytes,
                             "peak_source": peak_src, "kernel_ms": kernel_ms},
                "e2e": {"value": e2e_val, "unit": "Mrecords/s", "h2d_bytes_per_step": int(h2d_bytes), "d2h_bytes_per_step": int(d2h_bytes),
                        "ms_per_step": e2e_step, "d2h_bytes_per_record": d2h_bytes / max(rows, 1), "pcie_h2d_gbs": h2d_gbs, "pcie_d2h_gbs": d2h_gbs,
                        "pcie_ingest_roofline_Mrec_s": h2d_gbs * 1e3 / rec_c,
                        "pcie_ingest_roofline_frac": e2e_val / (h2d_gbs * 1e3 / rec_c),
                        "pcie_return_roofline_frac": (d2h_bytes / 1e9 / (e2e_step / 1e3)) / d2h_gbs,
                        "pcie_bidir_h2d_gbs": bi_h2d, "pcie_bidir_d2h_gbs": bi_d2h,
                        "pcie_return_roofline_frac_bidir": (d2h_bytes / 1e9 / (e2e_step / 1e3)) / bi_d2h,
                        "duckdb_layout": {"value": total_rows / 1e6 / (e2e_duck_step / 1e3), "unit": "Mrecords/s", "ms_per_step": e2e_duck_step,
                                          "d2h_bytes_per_step": int(d2h_duck), "d2h_bytes_per_record": d2h_duck / max(rows, 1)}},
                "ext_e2e": ext, "bgunzip_e2e": bgunzip, "bgzip_e2e": bgzip, "read_tabix_e2e": tabix, "strong_scaling": strong,
                "read_bcf": bcf_res, "contig_sharded": shard_res,
                "cpu_baseline": cpu, "clocks": clk.summary(), "gpu_launches": int(launches), "host_cores": ncpu}
        print(json.dumps(line), flush=True)
    barrier()
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
