"""Multi-process host logic of the sharded scans, world_size 2 over gloo on CPU: every rank takes its
share of the file (a contig group from the BAI index — BASELINE configs[3] — or a virtual-offset range), scans it (here:
the oracle stands in for the GPU), and rank 0 merges in rank order.  The merged rows must be the unsharded scan.
No collective on the data path: torch.distributed only carries the barrier, the max-over-ranks time and the test's gather."""
import os
import sys

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

from tests import util

REF_DUMP = os.path.join(util.ROOT, "harness", "_bin", "ref_dump")


def contig_bounds(weights, n_parts):
    """The balanced contiguous partition dh_bam_split_contigs computes (duckhts_b200/csrc/dh_bam.cu), restated."""
    pre = np.concatenate([[0], np.cumsum(weights)])
    out = [0]
    for p in range(1, n_parts):
        target = int(pre[-1]) // n_parts * p
        t = int(np.searchsorted(pre, target, side="left"))
        if t > 0 and target - pre[t - 1] < pre[min(t, len(weights))] - target:
            t -= 1
        out.append(min(max(t, out[-1]), len(weights)))
    return out + [len(weights)]


def worker(rank, world, port, path, q):
    sys.path.insert(0, util.ROOT)
    from oracle import dh_oracle_py as orc
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    img = np.fromfile(path, dtype=np.uint8)
    bai = np.fromfile(path + ".bai", dtype=np.uint8)
    rc, infl, _ = orc.bgzf_inflate(img)
    dec = orc.bam_decode(infl, [])
    n_ref = len(dec["ref_names"])
    weights = [int(np.sum(dec["tid"] == t)) for t in range(n_ref)]      # stands in for the index's compressed spans
    bounds = contig_bounds(weights, world)
    dist.barrier()
    rows = orc.contig_scan_rows(img, bai, dec, bounds[rank], bounds[rank + 1])
    gathered = [None] * world
    dist.all_gather_object(gathered, rows)
    import torch
    t = torch.tensor([float(len(rows))], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        merged = [r for part in gathered for r in part]
        whole = orc.contig_scan_rows(img, bai, dec)
        q.put((merged == whole, len(merged), bounds, float(t.item())))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(not util.have_gen(), reason="generator binary absent")
def test_contig_group_shards_merge_in_rank_order_world2():
    path = util.synth_path("bam", n_records=6000, seed=41, paired=True, extra=("--unplaced", "10"), index=True)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    ps = [ctx.Process(target=worker, args=(r, 2, port, path, q)) for r in range(2)]
    for p in ps:
        p.start()
    ok, n, bounds, mx = q.get(timeout=240)
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok and n == 6000 and bounds[0] == 0 and bounds[-1] == 22 and 8 <= bounds[1] <= 14
    assert mx >= n / 2


def test_contig_bounds_partition_is_contiguous_and_balanced():
    w = [10, 10, 10, 10, 40, 10, 10]
    b = contig_bounds(w, 4)
    assert b[0] == 0 and b[-1] == 7 and all(x <= y for x, y in zip(b, b[1:]))
    assert contig_bounds([5] * 22, 8)[-1] == 22 and len(contig_bounds([5] * 22, 8)) == 9
