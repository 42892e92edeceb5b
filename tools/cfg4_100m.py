#!/usr/bin/env python
"""BASELINE configs[3] at its stated size, and the 100 M-record scan checked cell for cell against the reference.

A coordinate-sorted 100,000,000-record paired BAM (22 contigs, seeded, written by the reference's htslib at BGZF level 6 — every
record distinct, nothing tiled) goes through:
  1. bam_index: our GPU build (dh_bam_index_build through the extension's bam_index) and htslib's sam_index_build (the reference's
     bam_index), timed; the two .bai compared as structures (tests/test_index_build_gpu.py::assert_same_index),
  2. read_bam (13 core columns + NM, MD, AS) through the UNCHANGED reference extension at 16 scan threads with its own .bai
     (contig-parallel mode) -> rows + order-insensitive hash of every cell,
  3. the same query through our extension, driven by OUR .bai: one source per listed GPU, each scanning a contiguous group of contigs
     through the index's virtual offsets, batches merged in contig order on the host -> rows + hash must equal (2),
  4. with --ordered: both at one scan thread -> the ORDERED hash must be equal too (the reference takes about a minute).
usage: python tools/cfg4_100m.py [--records 100e6] [--devices 0,0] [--ordered] [--out gpurun_out/r2h_cfg4.json]
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from tests.test_index_build_gpu import assert_same_index  # noqa: E402


def host(ext, func, path, *args, threads=16, env=None, hashed=False, repeat=1, timeout=3600):
    e = dict(os.environ)
    e.update(env or {})
    cmd = [bench.FAKE_HOST, ext, func, path, *args, "--threads", str(threads), "--time", "--repeat", str(repeat)] + (["--hash"] if hashed else [])
    t0 = time.time()
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, env=e, timeout=timeout)
    if p.returncode != 0:
        raise RuntimeError(f"{' '.join(cmd)} -> {p.returncode}\n{p.stdout[-2000:]}\n{p.stderr[-2000:]}")
    lines = p.stdout.strip().splitlines()
    res = json.loads(lines[-1])
    res["wall"] = time.time() - t0
    if hashed:
        res["hash_lines"] = lines[:-1]          # rows, then per column: name, chained hash (one thread only), multiset hash
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--records", type=float, default=100e6)
    ap.add_argument("--devices", default="0")
    ap.add_argument("--ordered", action="store_true")
    ap.add_argument("--skip-index", action="store_true", help="reuse the two .bai files of an earlier run")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "r2h_cfg4.json"))
    a = ap.parse_args()
    n = int(a.records)
    rep = {"records": n, "devices": a.devices}
    t0 = time.time()
    path = bench.gen_unit("bam", n, 0x9e3779b97f4a7c15, True, os.cpu_count() or 16)
    rep["file_bytes"] = os.path.getsize(path)
    rep["generate_seconds"] = time.time() - t0
    ours_bai, ref_bai = path + ".ours.bai", path + ".bai"
    env = {"DUCKHTS_DEVICES": a.devices}
    # 1. index build, both sides (the reference's writes PATH.bai, which its read_bam then finds by itself)
    if not (a.skip_index and os.path.exists(ours_bai) and os.path.exists(ref_bai)):
        for p_ in (ours_bai, ref_bai):
            if os.path.exists(p_):
                os.unlink(p_)
        r = host(bench.OUR_EXT, "bam_index", path, "index_path=" + ours_bai, threads=1, env=env, repeat=2)
        rep["bam_index_ours"] = {"seconds_cold": r["seconds"], "seconds_warm": r["best_seconds"], "Mrecords_per_s_warm": n / 1e6 / r["best_seconds"]}
        r = host(bench.REF_EXT, "bam_index", path, "index_path=" + ref_bai, "threads=16", threads=1)
        rep["bam_index_reference"] = {"seconds": r["seconds"], "Mrecords_per_s": n / 1e6 / r["seconds"], "htslib_threads": 16}
        assert_same_index(open(ours_bai, "rb").read(), open(ref_bai, "rb").read())
        rep["bai_equal_as_structures"] = True
        rep["bai_bytes"] = os.path.getsize(ours_bai)
        bench.log("[cfg4] index:", json.dumps({k: rep[k] for k in ("bam_index_ours", "bam_index_reference", "bai_bytes")}))
    # 2/3. the scan, every cell hashed
    q = ("standard_tags=true", "--cols", bench.EXT_COLS)
    ref = host(bench.REF_EXT, "read_bam", path, *q, threads=16, hashed=True)
    ours = host(bench.OUR_EXT, "read_bam", path, "index_path=" + ours_bai, *q, threads=16, env=env, hashed=True)
    rep["multiset_hash_equal"] = ref["hash_lines"] == ours["hash_lines"] and ref["rows"] == ours["rows"] and ours["rows"] > 0.99 * n
    rep["hash_lines"] = {"reference": ref["hash_lines"], "ours": ours["hash_lines"]}
    # the same two queries timed without the hashing consumer
    ref_t = host(bench.REF_EXT, "read_bam", path, *q, threads=16)
    ours_t = host(bench.OUR_EXT, "read_bam", path, "index_path=" + ours_bai, *q, threads=16, env=env, repeat=3)
    rep["reference_16_threads"] = {"rows": ref_t["rows"], "seconds": ref_t["seconds"], "Mrecords_per_s": ref_t["rows"] / 1e6 / ref_t["seconds"]}
    rep["ours_16_threads"] = {"rows": ours_t["rows"], "seconds_cold": ours_t["seconds"], "seconds_warm": ours_t["best_seconds"],
                              "Mrecords_per_s_cold": ours_t["rows"] / 1e6 / ours_t["seconds"], "Mrecords_per_s_warm": ours_t["rows"] / 1e6 / ours_t["best_seconds"]}
    bench.log("[cfg4] scan:", json.dumps({k: rep[k] for k in ("reference_16_threads", "ours_16_threads", "multiset_hash_equal")}))
    if a.ordered:
        ref1 = host(bench.REF_EXT, "read_bam", path, *q, threads=1, hashed=True)
        ours1 = host(bench.OUR_EXT, "read_bam", path, "index_path=" + ours_bai, *q, threads=1, env=env, hashed=True)
        rep["reference_1_thread"] = {"rows": ref1["rows"], "seconds": ref1["seconds"], "hash": ref1["hash_lines"]}
        rep["ours_1_thread"] = {"rows": ours1["rows"], "seconds": ours1["seconds"], "hash": ours1["hash_lines"]}
        rep["ordered_hash_equal"] = ref1["hash_lines"] == ours1["hash_lines"]
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    json.dump(rep, open(a.out, "w"), indent=1)
    print(json.dumps(rep))
    ok = rep["multiset_hash_equal"] and rep.get("ordered_hash_equal", True)
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
