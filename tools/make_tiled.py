#!/usr/bin/env python
"""Writes bench.py's tiled BAM (the seeded unit's record blocks repeated to ~N records) to a file: tools/make_tiled.py OUT [records]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import duckhts_b200 as dh  # noqa: E402

out, records = sys.argv[1], float(sys.argv[2]) if len(sys.argv) > 2 else 100e6
unit_path = bench.gen_unit("bam", 4000000, 0x9E3779B97F4A7C15, True, os.cpu_count() or 1)
unit = np.fromfile(unit_path, dtype=np.uint8)
blocks, end = dh.bgzf_walk(unit)
with dh.Context(0) as ctx, dh.BamFile(ctx, unit) as ub:
    first = ub.first_voffset >> 16
hb = next(i for i, b in enumerate(blocks) if b[0] == first)
body = unit[blocks[hb][0]:blocks[-1][0]]
reps = max(1, int(round(records / 4e6)))
with open(out, "wb") as f:
    f.write(unit[:blocks[hb][0]].tobytes())
    for _ in range(reps):
        f.write(memoryview(body))
    f.write(bench.util.BGZF_EOF)
print(out, os.path.getsize(out), reps)
