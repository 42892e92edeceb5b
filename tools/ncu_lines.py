#!/usr/bin/env python
"""Per-CUDA-source-line instruction / stall-sample shares from `ncu --page source --csv --print-source cuda,sass`.
usage: tools/ncu_lines.py report.ncu-rep kernel_regex [min_pct]"""
import csv, subprocess, sys, io
rep, kre = sys.argv[1], sys.argv[2]
minp = float(sys.argv[3]) if len(sys.argv) > 3 else 0.5
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", f"regex:{kre}"],
                     stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
fname, hdr, lines = None, None, []
for r in rows:
    if not r: continue
    if r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if r[0] == "Function Name": continue
    if r[0] == "Line No": hdr = r; ii = hdr.index("Instructions Executed"); si = hdr.index("# Samples"); ti = hdr.index("Thread Instructions Executed"); continue
    if hdr and r[0] not in ("", "0") and r[0].isdigit():
        try: lines.append((fname, int(r[0]), r[1], float(r[ii]), float(r[si]), float(r[ti])))
        except ValueError: pass
ti_ = sum(l[3] for l in lines); ts = sum(l[4] for l in lines)
print(f"total warp-instr {ti_:.0f}  samples {ts:.0f}")
for f, n, src, i, s, t in lines:
    if i >= minp / 100 * ti_ or s >= minp / 100 * ts:
        print(f"{f}:{n:<4d} instr {i / ti_ * 100:5.1f}%  samples {s / ts * 100:5.1f}%  thr/instr {t / max(i, 1):4.1f} | {src.strip()[:100]}")
