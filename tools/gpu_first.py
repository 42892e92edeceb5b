"""First GPU bring-up script (scratch tool): inflate synthetic BAMs, compare with zlib, time."""
import sys, time, zlib, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import duckhts_b200 as dh

def ref_inflate(data):
    blocks, _ = dh.bgzf_walk(data)
    out = []
    for off, cs, isz, crc in blocks:
        raw = zlib.decompress(bytes(data[off + 18: off + cs - 8]), -15)
        assert len(raw) == isz and zlib.crc32(raw) == crc
        out.append(raw)
    return b"".join(out), blocks

ctx = dh.Context(0)
print("sm_count", ctx.sm_count)
print("pcie GB/s (h2d, d2h):", ctx.measure_pcie(1 << 30))
for path in sys.argv[1:]:
    data = np.fromfile(path, dtype=np.uint8)
    t0 = time.time(); ref, blocks = ref_inflate(data); t1 = time.time()
    print(path, "bytes", data.size, "blocks", len(blocks), "out", len(ref), "zlib s", round(t1 - t0, 2))
    out, st = dh.bgzf_inflate(ctx, data)
    ok = bytes(out) == ref
    print("  host->host ok", ok, "kernel_ms", st.kernel_ms, "total_ms", st.total_ms, "GB/s out(kernel)", st.out_bytes / 1e6 / max(st.kernel_ms, 1e-9))
    if not ok:
        o = np.frombuffer(ref, dtype=np.uint8); d = np.nonzero(o != out)[0]
        print("  first diff at", d[:5], "ndiff", d.size)
    dev = dh.BgzfDevice(ctx, data)
    for it in range(3):
        ctx.flush_l2(); ctx.sync()
        ctx.timer_start(); dev.inflate(); ms = ctx.timer_stop()
        print("  dev inflate ms", ms, "GB/s out", dev.out_bytes / 1e6 / ms, "GB/s in", dev.comp_bytes / 1e6 / ms)
    print("  dev check", dev.check(), "equal", bytes(dev.read()) == ref)
    dev.free()
