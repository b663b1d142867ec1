"""Block partition of the cell range across ranks and the tiny global reductions.

The source path has no exchange step: cells are independent (MOC.C:106-122, QMOM.C:226-264), so
each rank owns the contiguous range [r*ncells/R, (r+1)*ncells/R) of every class / moment field
and no data-path collective exists.  The only cross-rank traffic is what the reference gets from
OpenFOAM's implicit MPI reductions: sums (global moments, printAvgMaxMin
numerators), mins (MOC.C:181-183) and maxes (MOC.C:238,276-279; QMOM.C:171) of a few doubles.
`torch.distributed` (NCCL on the GPU box, gloo in the CPU tests) carries them.
"""
from __future__ import annotations

from typing import Tuple

import torch
import torch.distributed as dist


def shard_range(ncells: int, rank: int, world: int) -> Tuple[int, int]:
    """[begin, end) of rank's cells: contiguous, balanced to within one cell, covers [0, ncells)."""
    assert 0 <= rank < world
    base, rem = divmod(int(ncells), int(world))
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def reduce_diagnostics(sums: torch.Tensor | None = None, mins: torch.Tensor | None = None,
                       maxs: torch.Tensor | None = None, group=None, async_op: bool = False):
    """All-reduce the three kinds of diagnostics in place; returns the work handles when async.

    Messages are <= 2.4 KB (N doubles at most): latency-bound, so bench.py issues them on a side
    stream and never waits for them inside the timed kernel stream."""
    works = []
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return works
    for t, op in ((sums, dist.ReduceOp.SUM), (mins, dist.ReduceOp.MIN), (maxs, dist.ReduceOp.MAX)):
        if t is not None and t.numel():
            w = dist.all_reduce(t, op=op, group=group, async_op=async_op)
            if async_op:
                works.append(w)
    return works


def avg_max_min(partials: torch.Tensor, group=None):
    """Finish Utility.C:4-10 printAvgMaxMin from pbe_field_avg_max_min's 4 partials:
    (sum f V, sum V, max f, min f) -> (volume average, global max, global min)."""
    s = partials[:2].clone()
    mx = partials[2:3].clone()
    mn = partials[3:4].clone()
    reduce_diagnostics(s, mn, mx, group)
    return (s[0] / s[1]).item(), mx.item(), mn.item()
