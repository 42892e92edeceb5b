"""Scratch: where does the host-to-host read_bam pipeline lose time?  Times the scan with device/host input and output."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import duckhts_b200 as dh

recs = float(sys.argv[1]) if len(sys.argv) > 1 else 50e6
ctx = dh.Context(0)
print("pcie one direction at a time (h2d, d2h):", ctx.measure_pcie(1 << 30))
print("pcie both directions at once  (h2d, d2h):", ctx.measure_pcie_bidir(1 << 30))
unit_path = bench.gen_unit("bam", 4000000, 0x9E3779B97F4A7C15, True, os.cpu_count())
unit = np.fromfile(unit_path, dtype=np.uint8)
with dh.BamFile(ctx, unit) as ub:
    first_voff = ub.first_voffset
    ub.set_projection(0, ())
    ub.scan_begin()
    unit_rows = ub.drain()
blocks, _ = dh.bgzf_walk(unit)
hb = next(i for i, b in enumerate(blocks) if b[0] == first_voff >> 16)
body = blocks[-1][0] - blocks[hb][0]
reps = max(1, int(round(recs / unit_rows)))
pin, reps = bench.tile_image(dh, unit, reps * body, header_blocks=hb)
bf = dh.BamFile(ctx, pin)
bf.set_projection(None, bench.TAGS)
for di, do in ((True, True), (True, False), (False, True), (False, False)):
    bf.set_mode(device_input=di, device_output=do)
    ts = []
    for it in range(3):
        t0 = time.perf_counter()
        bf.scan_begin()
        rows = bf.drain()
        ts.append((time.perf_counter() - t0) * 1e3)
    st = bf.stats()
    print(f"device_input={di} device_output={do}: rows {rows} ms {[round(t, 1) for t in ts]}  h2d {st.h2d_bytes / 1e9:.2f} GB d2h {st.d2h_bytes / 1e9:.2f} GB "
          f"kernels inf {st.inflate_ms:.0f} walk {st.walk_ms:.0f} dec {st.decode_ms:.0f}")
bf.close()
