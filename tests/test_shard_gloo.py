"""Multi-process host logic of the sharded scans, world_size 2 over gloo on CPU: every rank takes its
share of the file (a contig group from the BAI index — BASELINE configs[3] — computed by the PRODUCT's partition function
dh_index_split_contigs, which is host code), scans it (here: the oracle stands in for the GPU; the -m gpu version of this test,
tests/test_wire_gpu.py::test_cloned_handles_scan_ranges_that_merge_to_the_sequential_scan, merges real dh_bam_next_batch output),
and rank 0 merges in rank order.  The merged rows must be the unsharded scan.
No collective on the data path: torch.distributed only carries the barrier, the max-over-ranks time and the test's gather."""
import os
import sys

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

from tests import util

REF_DUMP = os.path.join(util.ROOT, "harness", "_bin", "ref_dump")


def product_bounds(bai, n_parts):
    """The partition the product computes: dh_index_split_contigs (duckhts_b200/csrc/dh_bam.cu) on the BAI image -- host code of
    libduckhts_cuda.so, no GPU involved (a BAI is parsed on the host)."""
    import duckhts_b200 as dh
    with dh.Index(None, bai) as ix:
        return ix.split_contigs(n_parts)


def worker(rank, world, port, path, q):
    sys.path.insert(0, util.ROOT)
    from oracle import dh_oracle_py as orc
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    img = np.fromfile(path, dtype=np.uint8)
    bai = np.fromfile(path + ".bai", dtype=np.uint8)
    rc, infl, _ = orc.bgzf_inflate(img)
    dec = orc.bam_decode(infl, [])
    bounds = product_bounds(bai, world)          # every rank computes the same partition from the index, no communication
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


@pytest.mark.skipif(not util.have_gen(), reason="generator binary absent")
def test_product_partition_is_contiguous_balanced_and_covers_every_contig():
    path = util.synth_path("bam", n_records=6000, seed=41, paired=True, extra=("--unplaced", "10"), index=True)
    bai = np.fromfile(path + ".bai", dtype=np.uint8)
    for parts in (1, 2, 3, 8, 22, 40):
        b = product_bounds(bai, parts)
        assert len(b) == parts + 1 and b[0] == 0 and b[-1] == 22 and all(x <= y for x, y in zip(b, b[1:]))
    b8 = product_bounds(bai, 8)
    sizes = [y - x for x, y in zip(b8, b8[1:])]
    assert max(sizes) <= 5 and min(sizes) >= 1          # 22 equal contigs over 8 parts: 2 or 3 (4 at a rounding edge) each
