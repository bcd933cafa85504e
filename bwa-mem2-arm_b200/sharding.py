"""Host-side plumbing of the multi-GPU path: pairs are independent, so a batch is cut
into contiguous index slices, one per rank / device, and results are merged by index.  There is no
collective on the data path; torch.distributed is only used to gather results (when one process wants
them all) and to reduce the timing (max over ranks) and the work counters (sum over ranks)."""
import numpy as np


def shard_range(n, rank, world):
    """Contiguous slice [lo, hi) of n items owned by `rank`; slices tile [0, n) with sizes differing by <= 1
    (same arithmetic as the C library's slicing of a multi-device handle)."""
    lo = n * rank // world
    hi = n * (rank + 1) // world
    return lo, hi


def slice_workload(pairs, rank, world):
    """This rank's SeqPairs (a copy, offsets untouched: every rank indexes the same sequence buffers)."""
    lo, hi = shard_range(len(pairs), rank, world)
    return np.ascontiguousarray(pairs[lo:hi]), lo, hi


def reduce_step_stats(dist, torch, device, ms_values, sum_values):
    """max over ranks of the time-like values, sum over ranks of the work counters (bench.py)."""
    t = torch.tensor(list(ms_values), dtype=torch.float64, device=device)
    s = torch.tensor(list(sum_values), dtype=torch.float64, device=device)
    if dist is not None and dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(s, op=dist.ReduceOp.SUM)
    return [float(x) for x in t.tolist()], [float(x) for x in s.tolist()]


def gather_results(dist, torch, local_pairs, n_total, rank, world, dtype):
    """Merge per-rank result slices into one array on rank 0 (host-side merge by index)."""
    buf = torch.from_numpy(local_pairs.view(np.uint8).reshape(-1).copy())
    sizes = [(shard_range(n_total, r, world)[1] - shard_range(n_total, r, world)[0]) * dtype.itemsize for r in range(world)]
    if world == 1:
        return local_pairs.copy()
    outs = [torch.empty(sz, dtype=torch.uint8) for sz in sizes] if rank == 0 else None
    dist.gather(buf, outs, dst=0)
    if rank != 0:
        return None
    return np.concatenate([o.numpy() for o in outs]).view(dtype)
