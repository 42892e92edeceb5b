#!/usr/bin/env python
"""Compress ~180 MB of BAM payload (the inflated stream of a seeded 60,000-record BAM, tiled) once or a few times: the program the
ncu captures of bgzf_deflate_kernel run.  usage: python tools/prof_deflate.py [repeats]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import duckhts_b200 as dh  # noqa: E402
from tests import util  # noqa: E402

ctx = dh.Context(0)
img = util.synth_bam_bytes(n_records=60000, seed=12, paired=True)
raw, _ = dh.bgzf_inflate(ctx, img)
data = np.tile(raw, 10)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 1
for r in range(reps):
    t0 = time.perf_counter()
    out, st = dh.bgzf_deflate(ctx, data)
    dt = time.perf_counter() - t0
    print(f"run {r}: {data.size / 1e6:.1f} MB -> {out.size / 1e6:.1f} MB ({out.size / data.size:.3f}), kernel {st.kernel_ms:.2f} ms = {data.size / 1e6 / st.kernel_ms:.2f} GB/s, wall {dt * 1e3:.1f} ms")
