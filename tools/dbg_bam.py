import sys, numpy as np
sys.path.insert(0, '/root/repo')
import duckhts_b200 as dh
from oracle import dh_oracle_py as orc
from tests import util
ctx = dh.Context(0)
img = np.fromfile(util.synth_path("bgzf", n_records=30000, seed=13, paired=True), dtype=np.uint8)
rc, infl, _ = orc.bgzf_inflate(img)
TAGS = [(b"NM", "i"), (b"MD", "Z"), (b"AS", "i")]
exp = orc.bam_decode(infl, TAGS)
blocks, _ = dh.bgzf_walk(img)
uoff = np.cumsum([0] + [b[2] for b in blocks])
for seg in (1 << 16, 1 << 18, 1 << 26):
    got = dh.bam_scan(ctx, img, tags=TAGS, seg_comp=seg, batch_rows=4096)
    e = orc.strcol_list(exp["rg"])
    bad = [i for i in range(exp["n"]) if got["rg"][i] != e[i]]
    print("seg", seg, "n", got["n"], "bad rg rows", len(bad), bad[:10], "rewalks", got["stats"].rewalk_rounds, "batches", got["n_batches"])
    for i in bad[:5]:
        o = int(exp["rec_off"][i]); blk = int(np.searchsorted(uoff, o, side="right") - 1)
        nxt = int(exp["rec_off"][i + 1]) if i + 1 < exp["n"] else -1
        print("  row", i, "off", o, "block", blk, "in-block", o - uoff[blk], "block end", uoff[blk + 1], "rec end", nxt, "got", got["rg"][i], got["sample"][i],
              [ (t["valid"][i]) for t in got["tags"]], got["qname"][i])
    for c in ("sample",):
        e2 = orc.strcol_list(exp[c]); print("  sample bad", sum(1 for i in range(exp["n"]) if got[c][i] != e2[i]))
    for k in range(3):
        print("  tag", k, "valid mismatches", int((got["tags"][k]["valid"] != exp["tags"][k]["valid"]).sum()))
