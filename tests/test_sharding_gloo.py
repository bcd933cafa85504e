"""N>1 host logic on CPU: world size 2 over gloo.  Each rank takes its contiguous slice of one batch,
runs the kernels' per-thread DP in host emulation (tests/emu), rank 0 merges by index and compares with
the single-process oracle; the timing/work reductions bench.py uses are checked as well."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    for p in (os.path.join(ROOT, "oracle"), os.path.join(ROOT, "bwa-mem2-arm_b200"), os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import common, emu, ora, sharding
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        d = ora.gen_bsw(1500, seed=0xA5)
        P, r, qer = d["left"]
        mine, lo, hi = sharding.slice_workload(P, rank, world)
        res, cells, _ = emu.bsw(mine, r, qer)
        merged = sharding.gather_results(dist, torch, res, len(P), rank, world, ora.SEQPAIR)
        ms, sums = sharding.reduce_step_stats(dist, torch, "cpu", [10.0 + rank, 3.0 - rank], [float(cells.sum()), float(len(mine))])
        if rank == 0:
            o = P.copy()
            tot = ora.bsw_batch(o, r, qer, ora.make_params(), 100, 2)
            ok = not common.diff_fields(o, merged, common.BSW_FIELDS).any() and (merged["id"] == P["id"]).all()
            q.put((ok, ms, sums, tot, len(P)))
    finally:
        dist.destroy_process_group()


def test_two_rank_shard_and_merge():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    ok, ms, sums, tot, n = q.get(timeout=180)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok
    assert ms == [11.0, 3.0]                    # max over ranks
    assert sums == [float(tot), float(n)]       # cells and pairs summed over ranks


def test_shard_range_tiles():
    sys.path.insert(0, os.path.join(ROOT, "bwa-mem2-arm_b200"))
    import sharding
    for n in (0, 1, 7, 64, 1000003):
        for world in (1, 2, 3, 8):
            cuts = [sharding.shard_range(n, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(cuts[i][1] == cuts[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in cuts]
            assert max(sizes) - min(sizes) <= 1
