#!/usr/bin/env python
"""bench.py — the driver's measurement contract for duckhts_b200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload read_bam|inflate]

One JSON line on stdout (rank 0).  A "step" is one pass of the hot path over one rank's whole input:
  * `value`   device-timed throughput with the compressed input already resident in HBM,
  * `e2e`     the same metric through the C-ABI with HOST buffers (pinned), H2D + kernels + D2H all inside
              the timed region,
  * `roofline` for the dominant kernel (BGZF inflate), `cpu_baseline` (reference htslib on the host cores,
              bounded sample), `clocks`, `gpu_launches`.
Inputs are synthetic: a seeded unit written by the reference's htslib at BGZF level 6
(harness/_bin/gen_synth), tiled up to the configured size (BGZF members concatenate).
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
sys.path.insert(0, ROOT)

from tests import util  # noqa: E402  (input generation helpers; no oracle import at module level)

HARNESS = os.path.join(ROOT, "harness", "_bin")
CACHE = util.CACHE


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
# reference arm / cpu baseline
# ---------------------------------------------------------------------------------------------------------
def run_ref_bench(mode, path, threads, limit):
    exe = os.path.join(HARNESS, "ref_bench")
    out = subprocess.check_output([exe, mode, path, str(threads), str(limit)], text=True)
    return json.loads(out.strip().splitlines()[-1])


def cpu_baseline_inflate(unit_path, threads, out_limit):
    """Reference htslib bgzf_read loop (src/bgzip.c:154-200) with bgzf_mt(threads) on a bounded sample."""
    if os.path.exists(os.path.join(HARNESS, "ref_bench")):
        r = run_ref_bench("bgunzip", unit_path, threads, out_limit)
        return {"value": r["bytes_out"] / 1e9 / r["seconds"], "unit": "GB/s decompressed", "cores": threads, "kind": "reference",
                "sample": f"first {r['bytes_out'] / 1e9:.2f} GB (inflated) of the seeded unit, htslib 1.23 bgzf_mt({threads}) + zlib"}
    from oracle import dh_oracle_py as orc  # cpu_baseline leg: the one place bench.py may execute oracle/

    data = np.fromfile(unit_path, dtype=np.uint8)
    blocks, _ = orc.bgzf_walk(data)
    cut, acc = 0, 0
    for b in blocks:
        acc += b[2]
        cut = b[0] + b[1]
        if acc >= out_limit:
            break
    t0 = time.time()
    rc, _, _ = orc.bgzf_inflate(data[:cut], want_output=False)
    dt = time.time() - t0
    return {"value": acc / 1e9 / dt, "unit": "GB/s decompressed", "cores": 1, "kind": "port",
            "sample": f"first {acc / 1e9:.2f} GB (inflated) of the seeded unit, oracle port (zlib), 1 thread"}


# ---------------------------------------------------------------------------------------------------------
WORKLOAD = ("read_bam full-column scan (13 core columns + typed tags NM, MD, AS) of a synthetic 100M-record 2x150 bp paired "
            "WGS-like BAM (BASELINE configs[2]); BGZF inflate GB/s of the same bytes reported alongside (configs[1])")
TAGS = [(b"NM", "i"), (b"MD", "Z"), (b"AS", "i")]


def cpu_baseline_read_bam(unit_path, threads, max_records):
    """The reference's htslib path (sam_read1 + the column conversions of src/bam_reader.c:779-966 through the same
    htslib calls) with hts_set_threads(threads), on a bounded sample."""
    r = run_ref_bench("sam_read1", unit_path, threads, max_records)
    return {"value": r["records"] / 1e6 / r["seconds"], "unit": "Mrecords/s", "cores": threads, "kind": "reference",
            "sample": f"first {r['records']} records of the seeded unit BAM; htslib 1.23 (zlib) sam_read1 + bam_reader.c conversions, "
                      f"hts_set_threads({threads})"}


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
    """BASELINE configs[4] shape (tidy FORMAT/GT + INFO decode, 2,504 samples), every column: sites/s and tidy rows/s."""
    ns = 2504
    if rank == 0:
        gen_bcf(args.bcf_sites, ns, 0x5eed)
    barrier()
    path = gen_bcf(args.bcf_sites, ns, 0x5eed)
    from duckhts_b200.bgzf import PinnedBuffer

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
                       f"{n_sites} biallelic sites x {ns} samples (BASELINE configs[4] shape; a seeded {args.bcf_sites}-site unit written by the "
                       f"reference htslib at BGZF level 6, its record blocks tiled x{reps})",
           "compressed_mb": pin.array.size / 1e6}

    def once():
        bf.scan_begin()
        n = bf.drain()
        return n, bf.stats()

    bf.set_mode(device_input=True, device_output=True)
    for _ in range(args.warmup):
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
    bf.set_mode(device_input=False, device_output=False)
    walls = []
    for it in range(1 + args.steps):
        barrier()
        t0 = time.perf_counter()
        n, st = once()
        if it:
            walls.append(rmax((time.perf_counter() - t0) * 1e3))
    w = float(np.mean(walls))
    out["e2e"] = {"value": sites / (w / 1e3), "unit": "sites/s", "rows_per_s": sites * ns / (w / 1e3), "ms_per_step": w,
                  "h2d_bytes_per_step": int(st.h2d_bytes), "d2h_bytes_per_step": int(st.d2h_bytes),
                  "d2h_gbs": st.d2h_bytes / 1e9 / (w / 1e3)}
    bf.close()
    host, ref = os.path.join(HARNESS, "fake_duckdb"), os.path.join(ROOT, "oracle", "_ref", "libduckhts_ref.so")
    if rank == 0 and os.path.exists(host) and os.path.exists(ref):
        sample = gen_bcf(min(args.bcf_sites, 3000), ns, 0x5eed)
        r = json.loads(subprocess.check_output([host, ref, "read_bcf", sample, "tidy_format=true", "--time"], stderr=subprocess.DEVNULL, text=True).strip().splitlines()[-1])
        out["cpu_baseline"] = {"value": r["rows"] / ns / r["seconds"], "unit": "sites/s", "rows_per_s": r["rows"] / r["seconds"], "cores": 1, "kind": "reference",
                               "sample": f"{r['rows'] // ns} sites x {ns} samples; the unchanged reference extension (read_bcf, tidy_format) under harness/fake_duckdb, "
                                         "1 scan thread (no index => the reference scans sequentially)"}
    return out


def bench_contig_shards(dh, ctx, args, unit_path, rank, world, barrier, rmax, rsum):
    """BASELINE configs[3]: BAI-driven scan, contig groups balanced by compressed bytes, one group per GPU (no collective)."""
    exe = os.path.join(HARNESS, "ref_dump")
    if rank == 0 and not os.path.exists(unit_path + ".bai") and os.path.exists(exe):
        subprocess.check_call([exe, "index", unit_path, "0"])
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
                        "(strong scaling on one 4M-record file)", "value": total / 1e6 / (w / 1e3), "unit": "Mrecords/s", "ms_per_step": w,
            "records": int(total), "contig_bounds": bounds, "status": int(st.status)}


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
    ap.add_argument("--bcf-tiles", type=int, default=5, help="how many times the unit's record blocks are repeated in the scanned image")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    ncpu = os.cpu_count() or 1
    config = {"workload": WORKLOAD, "records_per_gpu_target": int(args.records),
              "input": f"seeded unit ({args.unit_records} paired 2x150 records, 22 contigs, written by the reference htslib at BGZF level 6) "
                       "with its record blocks tiled to the target size", "sharding": "contiguous virtual-offset range per GPU, no collective"}

    if args.impl == "reference":
        if rank != 0:
            return 0
        unit_path = gen_unit("bam", args.unit_records, args.seed, True, ncpu)
        limit = 2000000
        for _ in range(min(args.warmup, 1)):
            run_ref_bench("sam_read1", unit_path, ncpu, 200000)
        res = [run_ref_bench("sam_read1", unit_path, ncpu, limit) for _ in range(args.steps)]
        secs = sum(r["seconds"] for r in res)
        recs = sum(r["records"] for r in res)
        v = recs / 1e6 / secs
        cb = {"value": v, "unit": "Mrecords/s", "cores": ncpu, "kind": "reference",
              "sample": f"first {res[0]['records']} records of the seeded unit BAM per step; htslib 1.23 (zlib) sam_read1 + "
                        f"src/bam_reader.c column conversions, hts_set_threads({ncpu})"}
        line = {"impl": "reference", "metric": "read_bam_Mrecords_per_s", "value": v, "unit": "Mrecords/s", "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs / max(args.steps, 1), "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": config, "cpu_baseline": cb,
                "e2e": {"value": v, "unit": "Mrecords/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
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

    torch.cuda.set_device(local_rank)
    if rank == 0:
        gen_unit("bam", args.unit_records, args.seed, True, ncpu)
    barrier()
    unit_path = gen_unit("bam", args.unit_records, args.seed, True, ncpu)
    unit = np.fromfile(unit_path, dtype=np.uint8)

    ctx = dh.Context(local_rank)
    with dh.BamFile(ctx, unit) as ub:   # where do the alignment blocks start, how many records does the unit hold
        first_voff = ub.first_voffset
        ub.set_projection(0, ())
        ub.scan_begin()
        unit_rows = ub.drain()
    blocks, _ = dh.bgzf_walk(unit)
    header_blocks = next(i for i, b in enumerate(blocks) if b[0] == first_voff >> 16)
    assert first_voff & 0xffff == 0, "htslib flushes after the header: first record starts a block"
    body_bytes = blocks[-1][0] - blocks[header_blocks][0]
    reps = max(1, int(round(args.records / unit_rows)))
    pin, reps = tile_image(dh, unit, reps * body_bytes, header_blocks=header_blocks)
    h2d_gbs, d2h_gbs = ctx.measure_pcie(1 << 30)
    bf = dh.BamFile(ctx, pin)
    bf.set_projection(None, TAGS)

    def scan_once():
        bf.scan_begin()
        rows = bf.drain()
        return rows, bf.stats()

    # ---- value: compressed image resident in HBM, columns left in HBM; CUDA events on the kernels' stream --------
    bf.set_mode(device_input=True, device_output=True)
    for _ in range(args.warmup):
        rows, st = scan_once()
    assert st.status == 0 and rows == reps * unit_rows, (st.status, rows, reps * unit_rows)
    barrier()
    with ClockSampler(local_rank) as clk:
        launches0 = ctx.launches
        ctx.timer_start()
        kms = [0.0, 0.0, 0.0]
        for _ in range(args.steps):
            rows, st = scan_once()
            kms = [kms[0] + st.inflate_ms, kms[1] + st.walk_ms, kms[2] + st.decode_ms]
        ms = ctx.timer_stop()
        launches = ctx.launches - launches0
        barrier()
    ms = rmax(ms)
    ms_per_step = ms / args.steps
    total_rows = rsum(float(rows))
    value = total_rows / 1e6 / (ms_per_step / 1e3)
    comp_bytes, out_bytes, nblocks = st.h2d_bytes, st.inflated_bytes, st.n_blocks
    comp_bytes = pin.nbytes

    # ---- e2e: pinned host image in, DuckDB-layout column batches in pinned host memory out -------------------------
    bf.set_mode(device_input=False, device_output=False)
    e2e_ms, d2h_bytes, h2d_bytes = [], 0, 0
    for it in range(1 + args.steps):
        barrier()
        t0 = time.perf_counter()
        rows, st = scan_once()
        wall = (time.perf_counter() - t0) * 1e3
        if it >= 1:
            e2e_ms.append(rmax(wall))
            d2h_bytes, h2d_bytes = st.d2h_bytes, st.h2d_bytes
    e2e_step = float(np.mean(e2e_ms))
    e2e_val = total_rows / 1e6 / (e2e_step / 1e3)
    bf.close()

    # ---- the dominant kernel alone (BGZF inflate of the same bytes, device-resident): roofline + inflate GB/s -------
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
    # (bgzf_resolve3_kernel) over <= 2^17 BGZF blocks.  Algorithmic HBM bytes of a BGZF inflate = compressed bytes read +
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

    bcf_res = bench_bcf(dh, ctx, args, rank, barrier, rmax, rsum) if args.bcf_sites > 0 and os.path.exists(os.path.join(HARNESS, "gen_bcf")) else None
    shard_res = bench_contig_shards(dh, ctx, args, unit_path, rank, world, barrier, rmax, rsum)

    if rank == 0:
        cpu = cpu_baseline_read_bam(unit_path, ncpu, 2000000) if world == 1 else None
        rec_c = comp_bytes / rows
        config.update({"records_per_gpu": int(rows), "compressed_gb_per_gpu": comp_bytes / 1e9, "inflated_gb_per_gpu": ib_out / 1e9,
                       "bgzf_blocks_per_gpu": int(nblocks), "tiles": reps, "compressed_bytes_per_record": rec_c,
                       "l2": "inputs (>= 10 GB per GPU) far exceed the 126 MB L2; no flush needed"})
        line = {"metric": "read_bam_Mrecords_per_s", "value": value, "unit": "Mrecords/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "u8", "data": "synthetic", "config": config,
                "kernel_ms_per_step": {"bgzf_inflate": kms[0] / args.steps, "bam_walk": kms[1] / args.steps, "bam_decode": kms[2] / args.steps},
                "inflate": {"metric": "bgzf_inflate_decompressed_GBps", "value": inflate_gbs, "unit": "GB/s",
                            "ms_per_step": ims / args.steps, "compressed_gb": ib_comp / 1e9, "inflated_gb": ib_out / 1e9,
                            "phase_ms": {"bgzf_decode3_kernel": ph1, "bgzf_resolve3_kernel": ph2, "fallback_filter": ph3},
                            "pcie_ingest_bound_GBps": h2d_gbs * ib_out / ib_comp},
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": traffic, "kernel": "bgzf_decode3_kernel + bgzf_resolve3_kernel (one inflate launch group)",
                             "algorithmic_bytes_per_launch": alg_bytes,
                             "peak_source": peak_src, "kernel_ms": kernel_ms},
                "e2e": {"value": e2e_val, "unit": "Mrecords/s", "h2d_bytes_per_step": int(h2d_bytes), "d2h_bytes_per_step": int(d2h_bytes),
                        "ms_per_step": e2e_step, "pcie_h2d_gbs": h2d_gbs, "pcie_d2h_gbs": d2h_gbs,
                        "pcie_ingest_roofline_Mrec_s": h2d_gbs * 1e3 / rec_c,
                        "pcie_ingest_roofline_frac": e2e_val / (h2d_gbs * 1e3 / rec_c),
                        "pcie_return_roofline_frac": (d2h_bytes / 1e9 / (e2e_step / 1e3)) / d2h_gbs,
                        "pcie_bidir_h2d_gbs": bi_h2d, "pcie_bidir_d2h_gbs": bi_d2h,
                        "pcie_return_roofline_frac_bidir": (d2h_bytes / 1e9 / (e2e_step / 1e3)) / bi_d2h},
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
