import os, sys, time
import numpy as np
sys.path.insert(0, "/root/repo")
import bench
import duckhts_b200 as dh
ctx = dh.Context(0)
unit_path = bench.gen_unit("bam", 4000000, 0x9E3779B97F4A7C15, True, os.cpu_count())
unit = np.fromfile(unit_path, dtype=np.uint8)
blocks, _ = dh.bgzf_walk(unit)
with dh.BamFile(ctx, unit) as ub:
    first_voff = ub.first_voffset
hb = next(i for i, b in enumerate(blocks) if b[0] == first_voff >> 16)
body = blocks[-1][0] - blocks[hb][0]
pin, reps = bench.tile_image(dh, unit, 12 * body, header_blocks=hb)
bf = dh.BamFile(ctx, pin)
bf.set_projection(None, bench.TAGS)
bf.set_mode(device_input=True, device_output=True)
for seg in (0, 512 << 20, 768 << 20, 1536 << 20, 2047 << 20):
    bf.configure(seg, 0, 0)
    ts = []
    for it in range(3):
        t0 = time.perf_counter(); bf.scan_begin(); rows = bf.drain(); ts.append((time.perf_counter() - t0) * 1e3)
    st = bf.stats()
    print(f"seg_comp={seg>>20} MB: ms {[round(t,1) for t in ts]} inf {st.inflate_ms:.0f} walk {st.walk_ms:.0f} dec {st.decode_ms:.0f}", flush=True)
