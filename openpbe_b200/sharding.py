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


# ---- transport: decomposed meshes and the halo exchange ------------------------
# The only data-path exchange of the engine.  Cells stay block-partitioned exactly as on the source
# path; a face whose two cells live on different ranks becomes a processor-patch face on both
# (appended after the physical boundary faces, grouped by neighbouring rank, in ascending global
# face order on both sides -- OpenFOAM's decomposePar convention), and before every transport step
# each rank sends the field values of its cells along those faces to the rank across.
import numpy as np  # noqa: E402


def rank_of_cell(cells, ncells: int, world: int):
    begins = np.array([shard_range(ncells, r, world)[0] for r in range(world)] + [ncells])
    return np.searchsorted(begins, cells, side="right") - 1


def partition_mesh(mesh: dict, rank: int, world: int) -> dict:
    """Sub-domain `rank` of a mesh dict (owner, neighbour, boundary_owner, vcell, phi as
    synthetic.block_mesh returns them).  Adds: coupled = (first boundary face, count),
    neighbours = [(rank, first coupled face, count)], send_cells (local, one per coupled face),
    halo_cells_global (the cell across each coupled face), cell_range, phys_faces."""
    own, nei, bown = (np.asarray(mesh[k], dtype=np.int64) for k in ("owner", "neighbour", "boundary_owner"))
    V, phi = np.asarray(mesh["vcell"]), np.asarray(mesh["phi"])
    nc, ni = V.size, own.size
    lo, hi = shard_range(nc, rank, world)
    o_in, n_in = (own >= lo) & (own < hi), (nei >= lo) & (nei < hi)
    internal, cut = o_in & n_in, o_in ^ n_in
    cut_faces = np.nonzero(cut)[0]
    mine = np.where(o_in, own, nei)[cut]
    across = np.where(o_in, nei, own)[cut]
    sign = np.where(o_in[cut], 1.0, -1.0)  # phi is positive owner -> neighbour; patch flux is outward
    other = rank_of_cell(across, nc, world)
    order = np.lexsort((cut_faces, other))
    phys = np.nonzero((bown >= lo) & (bown < hi))[0]
    groups, start = [], 0
    for q in np.unique(other):
        cnt = int(np.sum(other == q))
        groups.append((int(q), start, cnt))
        start += cnt
    return {
        "owner": (own[internal] - lo).astype(np.int32), "neighbour": (nei[internal] - lo).astype(np.int32),
        "boundary_owner": np.concatenate([bown[phys] - lo, mine[order] - lo]).astype(np.int32),
        "vcell": V[lo:hi].copy(),
        "phi": np.concatenate([phi[:ni][internal], phi[ni:][phys], (sign * phi[:ni][cut])[order]]),
        "coupled": (int(phys.size), int(cut_faces.size)), "neighbours": groups,
        "send_cells": (mine[order] - lo).astype(np.int32), "halo_cells_global": across[order],
        "cell_range": (lo, hi), "phys_faces": phys,
    }


class HaloExchange:
    """Per-step exchange of the values along the processor patches: pack (device gather kernel, or
    plain indexing for CPU tensors in the gloo tests), one send + one receive per neighbouring rank
    (`torch.distributed` point-to-point: NCCL over NVLink between GPUs), unpack into the
    [fields][coupled faces] layout pbe_transport_explicit_halo reads."""

    def __init__(self, part: dict, device, mesh=None, group=None):
        self.groups = part["neighbours"]
        self.n = part["coupled"][1]
        self.mesh = mesh  # engine.Mesh (device packing) or None (CPU tensors)
        self.group = group
        self.send_cells = torch.from_numpy(part["send_cells"]).to(device)

    def pack(self, psi: torch.Tensor, q_start: int, q_count: int) -> torch.Tensor:
        idx = self.send_cells[q_start:q_start + q_count]
        if self.mesh is not None and psi.is_cuda:
            return self.mesh.gather_cells(psi, idx)
        return psi[:, idx.long()].contiguous()

    def exchange(self, psi: torch.Tensor) -> torch.Tensor:
        return self.finish(self.start(psi))

    def start(self, psi: torch.Tensor):
        """Pack and post the sends / receives; returns a handle for finish().  Work enqueued on the
        current stream between start() and finish() (the regular cells' update, which reads no
        halo value) overlaps the transfers."""
        halo = torch.empty((psi.shape[0], self.n), dtype=psi.dtype, device=psi.device)
        if not self.groups:
            return halo, [], []
        ops, recvs = [], []
        for q, start, count in self.groups:
            s = self.pack(psi, start, count)
            r = torch.empty_like(s)
            recvs.append((r, start, count))
            ops.append(dist.P2POp(dist.isend, s, q, self.group))
            ops.append(dist.P2POp(dist.irecv, r, q, self.group))
        return halo, dist.batch_isend_irecv(ops), recvs

    def finish(self, handle) -> torch.Tensor:
        halo, works, recvs = handle
        for w in works:
            w.wait()
        for r, start, count in recvs:
            halo[:, start:start + count] = r
        return halo


class PeerHaloExchange:
    """The halo exchange as stores into peer memory: every rank's halo buffer lives in symmetric
    memory (torch.distributed._symmetric_memory: cudaMalloc'd, peer-mapped over NVLink / NVSwitch),
    and the pack kernel (pbe_gather_cells) of the sending rank writes the values straight into the
    receiving rank's buffer -- pack and transfer are one kernel, there is no send / receive pair and
    no unpack copy.  One device-side barrier per step (symmetric-memory signal pads) orders the
    stores before the reads.  Two buffers alternate, so a rank that runs ahead writes step t+1's
    values while its neighbour still reads step t's.  Same start() / finish() protocol as
    HaloExchange; needs a CUDA device per rank (no CPU form)."""

    def __init__(self, part: dict, device, mesh, nfields: int, group=None):
        import torch.distributed._symmetric_memory as symm_mem

        self.mesh, self.K = mesh, int(nfields)
        self.groups = part["neighbours"]
        self.n = part["coupled"][1]
        group = group if group is not None else dist.group.WORLD
        rank = dist.get_rank(group)
        everyone = [None] * dist.get_world_size(group)
        dist.all_gather_object(everyone, (self.n, self.groups), group=group)
        self.nmax = max(1, max(n for n, _ in everyone))
        # where my values land in each neighbour's buffer: that neighbour's group for me
        self.targets = []
        for q, start, count in self.groups:
            back = [g for g in everyone[q][1] if g[0] == rank]
            assert len(back) == 1 and back[0][2] == count, "processor patches of the two sides differ"
            self.targets.append((q, start, count, back[0][1]))
        self.buf = symm_mem.empty((2, self.K, self.nmax), dtype=torch.float64, device=device)
        self.hdl = symm_mem.rendezvous(self.buf, group)
        self.ptrs = list(self.hdl.buffer_ptrs)
        self.send_cells = torch.from_numpy(part["send_cells"]).to(device)
        self.parity = 0
        self.hdl.barrier(channel=0)

    def start(self, psi: torch.Tensor):
        from . import capi
        from .engine import _stream_ptr

        p = self.parity
        self.parity ^= 1
        K, ncells = psi.shape
        assert K == self.K and psi.is_contiguous()
        for q, start, count, rstart in self.targets:
            out = self.ptrs[q] + ((p * K) * self.nmax + rstart) * 8
            capi.check(capi.lib().pbe_gather_cells(
                self.mesh.ctx.handle, K, psi.stride(0), psi.data_ptr(), count,
                self.send_cells.data_ptr() + 4 * start, out, self.nmax, _stream_ptr(psi.device)))
        return p

    def finish(self, p: int) -> torch.Tensor:
        self.hdl.barrier(channel=0)  # every rank's packs of this step have landed
        return self.buf[p][:, :self.n]  # [K][n coupled faces], row stride nmax

    def exchange(self, psi: torch.Tensor) -> torch.Tensor:
        return self.finish(self.start(psi))
