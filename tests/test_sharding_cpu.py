"""World-size-2 gloo tests of the multi-rank host logic (no GPU): the block partition, the
rank-independent synthetic shards and the sum / min / max reductions.  The per-rank source
evaluation is done by the oracle here -- this file tests the plumbing around the kernels."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from openpbe_b200 import sharding, synthetic  # noqa: E402


def test_shard_range_partitions():
    for ncells in (0, 1, 7, 128, 1000, 16_000_000, 128_000_001):
        for world in (1, 2, 3, 4, 8):
            prev = 0
            sizes = []
            for r in range(world):
                b, e = sharding.shard_range(ncells, r, world)
                assert b == prev and e >= b
                sizes.append(e - b)
                prev = e
            assert prev == ncells and max(sizes) - min(sizes) <= 1


def _worker(rank, world, port, ncells, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import cpu_oracle as orc

    w = synthetic.workload("moc30_ct", ncells)
    b, e = sharding.shard_range(ncells, rank, world)
    f = synthetic.moc_fields(w, offset=b, ncells=e - b)
    spec = orc.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    S, _ = orc.moc_source(spec, w.N, w.xi1, w.relax, f["n"], f["eps"], f["rho"], f["nu"], w.dt)
    n1 = f["n"].copy()
    orc.moc_update(n1, S, w.dt)
    xi = w.xi1 * np.arange(1, w.N + 1)
    mom = torch.tensor([np.sum(f["vcell"] * np.sum(xi[:, None] ** k * n1, axis=0)) for k in range(4)])
    mn = torch.from_numpy(n1.min(axis=1).copy())
    mx = torch.from_numpy(n1.max(axis=1).copy())
    sharding.reduce_diagnostics(mom, mn, mx)
    part = torch.tensor([np.sum(n1[0] * f["vcell"]), np.sum(f["vcell"]), n1[0].max(), n1[0].min()])
    avg = sharding.avg_max_min(part)
    if rank == 0:
        np.savez(out, mom=mom.numpy(), mn=mn.numpy(), mx=mx.numpy(), avg=np.array(avg))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_reductions_match_single_process(tmp_path, oracle):
    ncells, world = 1001, 2
    out = str(tmp_path / "r.npz")
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(world, port, ncells, out), nprocs=world, join=True)
    z = np.load(out)
    # single-process reference over the same shards
    w = synthetic.workload("moc30_ct", ncells)
    spec = oracle.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    xi = w.xi1 * np.arange(1, w.N + 1)
    mom = np.zeros(4)
    parts = []
    for r in range(world):
        b, e = sharding.shard_range(ncells, r, world)
        f = synthetic.moc_fields(w, offset=b, ncells=e - b)
        S, _ = oracle.moc_source(spec, w.N, w.xi1, w.relax, f["n"], f["eps"], f["rho"], f["nu"], w.dt)
        n1 = f["n"].copy()
        oracle.moc_update(n1, S, w.dt)
        mom += [np.sum(f["vcell"] * np.sum(xi[:, None] ** k * n1, axis=0)) for k in range(4)]
        parts.append((n1, f["vcell"]))
    alln = np.concatenate([p[0] for p in parts], axis=1)
    allv = np.concatenate([p[1] for p in parts])
    np.testing.assert_allclose(z["mom"], mom, rtol=1e-14)
    np.testing.assert_array_equal(z["mn"], alln.min(axis=1))
    np.testing.assert_array_equal(z["mx"], alln.max(axis=1))
    np.testing.assert_allclose(z["avg"], [np.sum(alln[0] * allv) / allv.sum(), alln[0].max(), alln[0].min()],
                               rtol=1e-14)


def test_reduce_is_noop_without_process_group():
    t = torch.tensor([1.0, 2.0])
    assert sharding.reduce_diagnostics(t, t.clone(), t.clone()) == []
    assert sharding.avg_max_min(torch.tensor([6.0, 3.0, 5.0, 1.0])) == (2.0, 5.0, 1.0)
