"""Seeded synthetic per-cell fields and the named workloads of BASELINE.json.

Shared by tests/ and bench.py so that the checker, the CPU baseline and the CUDA path see
identical inputs.  Pure NumPy: this module never touches the GPU and never computes a source.

Constants come from the reference tree (cited; the files themselves are not needed at run time):
  * benchmark operands: benchmarks/coalescence.cpp:28-33, benchmarks/breakup.cpp:29-33
  * c01 validation case: tests/validation/simmons_and_azzopardi/c01/constant/phaseProperties:24-50,
    .../thermophysicalProperties.acps:38,47, .../system/controlDict:28, .../0/epsilon.acps:20
  * log-normal drop-size initial condition: tests/validation/simmons_and_azzopardi/setupCases.py:22-40
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, Tuple

import numpy as np

SEED = 20240611


@dataclass
class Workload:
    name: str
    method: str                 # "MOC" | "QMOM"
    N: int                      # classes (MOC) or quadrature order (QMOM)
    ncells: int
    xi1: float = 0.0            # MOC only
    dt: float = 2.5e-4
    coalescence: str = "noCoalescence"
    coal_coeffs: Tuple[float, ...] = ()
    breakup: str = "noBreakup"
    br_coeffs: Tuple[float, ...] = ()
    rho: float = 1000.0
    nu: float = 1e-6
    relax: float = 1.0
    extra: Dict = field(default_factory=dict)


# CT coefficient order everywhere: (c1, c2, gamma, sigma)
_BENCH_CT = (0.0152, 0.0678, 0.050000000000000003, 0.04282)
_C01_COAL = (4.5e6, 1.891e12, 0.0, 0.01)
_C01_BR = (0.1, 0.06354, 0.0, 0.01)
_C01_RHO = 1166.0
_C01_NU = 1.599752e-3 / 1166.0


def workload(name: str, ncells: int | None = None) -> Workload:
    """The BASELINE.json configs by short name (ncells can be overridden for tests)."""
    if name == "moc30_ct":      # configs[0]: MOC 30 classes, 1e5 cells, benchmark constants
        w = Workload(name, "MOC", 30, 100_000, xi1=3e-10 / 30, coalescence="CoulaloglouTavlaridesC",
                     coal_coeffs=_BENCH_CT, breakup="CoulaloglouTavlarides", br_coeffs=_BENCH_CT,
                     rho=972.0, nu=1e-3)
    elif name == "qmom3_ct":    # configs[1]: QMOM 6 moments, 1e7 cells
        w = Workload(name, "QMOM", 3, 10_000_000, coalescence="CoulaloglouTavlaridesC",
                     coal_coeffs=_C01_COAL, breakup="CoulaloglouTavlarides", br_coeffs=_C01_BR,
                     rho=_C01_RHO, nu=_C01_NU)
    elif name == "moc100_ct":   # configs[2]: MOC 100 classes, 1.6e7 cells, c01 constants
        w = Workload(name, "MOC", 100, 16_000_000, xi1=3e-10 / 100, coalescence="CoulaloglouTavlaridesC",
                     coal_coeffs=_C01_COAL, breakup="CoulaloglouTavlarides", br_coeffs=_C01_BR,
                     rho=_C01_RHO, nu=_C01_NU)
    elif name == "moc100_shard":  # configs[3]: 1.28e8 cells over all ranks
        w = workload("moc100_ct")
        w.name, w.ncells = name, 128_000_000
    elif name == "replay_moc":  # configs[4]: correct() replay, MOC side (c01: 50 classes)
        w = Workload(name, "MOC", 50, 4_000_000, xi1=3.5918876006043295e-12,
                     coalescence="CoulaloglouTavlaridesC", coal_coeffs=_C01_COAL,
                     breakup="CoulaloglouTavlarides", br_coeffs=_C01_BR, rho=_C01_RHO, nu=_C01_NU)
    elif name == "replay_qmom":
        w = workload("qmom3_ct")
        w.name, w.ncells = name, 4_000_000
    elif name == "moc100_const":  # cell-independent variant (the kernels of the 0-D analytical solutions)
        w = Workload(name, "MOC", 100, 16_000_000, xi1=0.1, dt=1e-3, coalescence="constantCoalescence",
                     coal_coeffs=(0.1,), breakup="binaryBreakup")
    elif name == "qmom3_const":
        w = Workload(name, "QMOM", 3, 10_000_000, dt=1e-3, coalescence="constantCoalescence",
                     coal_coeffs=(0.1,), breakup="binaryBreakup")
    else:
        raise KeyError(name)
    if ncells is not None:
        w.ncells = int(ncells)
    return w


def _lognormal_volume_pdf(v: np.ndarray, dmean=5.0e-4, dstd=5.0e-5) -> np.ndarray:
    """Number pdf in volume space of a log-normal DIAMETER distribution (setupCases.py:22-40)."""
    dvar = dstd ** 2
    mu = np.log(dmean / np.sqrt(1 + dvar / dmean ** 2))
    sigma = np.sqrt(np.log(1 + dvar / dmean ** 2))
    d = np.cbrt(6.0 * v / np.pi)
    pdf_d = np.exp(-(np.log(d) - mu) ** 2 / (2 * sigma ** 2)) / (d * sigma * np.sqrt(2 * np.pi))
    dd_dv = (2.0 / (np.pi * d ** 2))  # d(d)/dv for v = pi d^3 / 6
    return pdf_d * dd_dv


def cell_fields(ncells: int, rho: float, nu: float, seed: int = SEED, offset: int = 0):
    """eps log-uniform [1e-3, 1], alpha uniform [0.01, 0.10], V = 1e-9, uniform rho / nu fields.

    `offset` = index of the first cell in the global numbering: every rank of a sharded run
    draws its own block from a generator seeded with (seed, offset), so results do not depend
    on how many ranks there are as long as shard boundaries coincide."""
    rng = np.random.default_rng([seed, offset])
    eps = 10.0 ** rng.uniform(-3.0, 0.0, ncells)
    alpha = rng.uniform(0.01, 0.10, ncells)
    vcell = np.full(ncells, 1e-9)
    return {"eps": eps, "alpha": alpha, "vcell": vcell,
            "rho": np.full(ncells, float(rho)), "nu": np.full(ncells, float(nu))}


def moc_fields(w: Workload, seed: int = SEED, offset: int = 0, ncells: int | None = None):
    """n_i[c] = alpha V p_i / sum_j p_j xi_j, times a per-(class, cell) jitter in [0.5, 2]."""
    nc = w.ncells if ncells is None else ncells
    f = cell_fields(nc, w.rho, w.nu, seed, offset)
    xi = w.xi1 * np.arange(1, w.N + 1)
    if w.coalescence == "constantCoalescence":
        # the 0-D analytical test's initial condition shape (tests/integration/MOC/setupCases.py:41-54)
        N0, v0 = 2.0, 0.5
        p = (N0 / v0) * (xi / v0) * np.exp(-xi / v0) * w.xi1
        base = np.repeat(p[:, None], nc, axis=1)
    else:
        p = _lognormal_volume_pdf(xi) + 1e-300
        base = (f["alpha"] * f["vcell"])[None, :] * (p / np.sum(p * xi))[:, None]
    rng = np.random.default_rng([seed + 1, offset])
    jitter = 2.0 ** rng.uniform(-1.0, 1.0, (w.N, nc))
    f["n"] = np.ascontiguousarray(base * jitter)
    return f


def qmom_fields(w: Workload, seed: int = SEED, offset: int = 0, ncells: int | None = None,
                degenerate_fraction: float = 0.01):
    """Moments of a random `order`-node quadrature per cell (realisable by construction);
    a fraction of the cells gets two nodes within 1e-6 relative (near the realisability
    boundary) to exercise the degenerate branch of the inversion."""
    nc = w.ncells if ncells is None else ncells
    order = w.N
    f = cell_fields(nc, w.rho, w.nu, seed, offset)
    rng = np.random.default_rng([seed + 2, offset])
    if w.coalescence == "constantCoalescence":
        nodes = np.sort(rng.uniform(0.05, 3.0, (order, nc)), axis=0)
        weights = rng.uniform(0.1, 2.0, (order, nc))
    else:
        nodes = np.sort(10.0 ** rng.uniform(-12.0, np.log10(3e-10), (order, nc)), axis=0)
        weights = rng.uniform(1e5, 1e7, (order, nc))
    if order >= 2 and degenerate_fraction > 0:
        deg = rng.random(nc) < degenerate_fraction
        nodes[1, deg] = nodes[0, deg] * (1.0 + 1e-6)
        nodes = np.sort(nodes, axis=0)
    m = np.empty((2 * order, nc))
    pw = np.ones_like(nodes)
    for k in range(2 * order):
        m[k] = np.sum(weights * pw, axis=0)
        pw = pw * nodes
    f["m"] = m
    f["nodes"] = nodes
    f["weights"] = weights
    return f


def block_mesh(nx: int, ny: int, nz: int, lengths=(1.0, 1.0, 1.0), velocity=(1.0, 0.0, 0.0)):
    """A blockMesh-style hex block in OpenFOAM face addressing: internal faces ordered by owner,
    then neighbour (upper-triangular order), boundary faces patch by patch (x-, x+, y-, y+, z-,
    z+).  Returns owner, neighbour, boundary_owner, vcell, phi (uniform velocity dotted with the
    face area vectors; boundary values outward-positive), and the patch start offsets."""
    dx, dy, dz = lengths[0] / nx, lengths[1] / ny, lengths[2] / nz
    idx = np.arange(nx * ny * nz, dtype=np.int64).reshape(nz, ny, nx)
    u, v, w = velocity
    own, nei, ph, sf = [], [], [], []
    for sl_o, sl_n, flux, area in (
            ((slice(None), slice(None), slice(0, nx - 1)), (slice(None), slice(None), slice(1, nx)), u * dy * dz, (dy * dz, 0.0, 0.0)),
            ((slice(None), slice(0, ny - 1), slice(None)), (slice(None), slice(1, ny), slice(None)), v * dx * dz, (0.0, dx * dz, 0.0)),
            ((slice(0, nz - 1), slice(None), slice(None)), (slice(1, nz), slice(None), slice(None)), w * dx * dy, (0.0, 0.0, dx * dy))):
        o, n = idx[sl_o].ravel(), idx[sl_n].ravel()
        own.append(o); nei.append(n); ph.append(np.full(o.size, flux)); sf.append(np.tile(np.array(area), (o.size, 1)))
    own, nei, ph, sf = np.concatenate(own), np.concatenate(nei), np.concatenate(ph), np.concatenate(sf)
    order = np.lexsort((nei, own))
    own, nei, ph, sf = own[order], nei[order], ph[order], sf[order]
    bown, bph, bsf, starts = [], [], [], [0]
    for cells, flux, area in ((idx[:, :, 0], -u * dy * dz, (-dy * dz, 0.0, 0.0)), (idx[:, :, -1], u * dy * dz, (dy * dz, 0.0, 0.0)),
                              (idx[:, 0, :], -v * dx * dz, (0.0, -dx * dz, 0.0)), (idx[:, -1, :], v * dx * dz, (0.0, dx * dz, 0.0)),
                              (idx[0, :, :], -w * dx * dy, (0.0, 0.0, -dx * dy)), (idx[-1, :, :], w * dx * dy, (0.0, 0.0, dx * dy))):
        bown.append(cells.ravel()); bph.append(np.full(cells.size, flux)); starts.append(starts[-1] + cells.size)
        bsf.append(np.tile(np.array(area), (cells.size, 1)))
    bown, bph, bsf = np.concatenate(bown), np.concatenate(bph), np.concatenate(bsf)
    # geometry for the limited schemes (fvMesh::Sf(), C(), weights()): outward-from-owner face area
    # vectors, cell centres, linear interpolation weights (1/2 on a uniform block)
    kz, ky, kx = np.meshgrid(np.arange(nz), np.arange(ny), np.arange(nx), indexing="ij")
    centres = np.stack([(kx.ravel() + 0.5) * dx, (ky.ravel() + 0.5) * dy, (kz.ravel() + 0.5) * dz], axis=1)
    return {"owner": own.astype(np.int32), "neighbour": nei.astype(np.int32),
            "boundary_owner": bown.astype(np.int32), "vcell": np.full(nx * ny * nz, dx * dy * dz),
            "phi": np.concatenate([ph, bph]), "patch_start": np.array(starts), "shape": (nx, ny, nz),
            "Sf": np.ascontiguousarray(np.concatenate([sf, bsf])), "C": np.ascontiguousarray(centres),
            "weights": np.full(own.size, 0.5)}


def block_slab_part(n: int, rank: int, world: int, velocity=(1.0, 0.5, 0.25)):
    """Sub-domain `rank` of an n x n x (n world) hex block cut into z-slabs of n^3 cells, built
    directly (what sharding.partition_mesh gives for the global block, without generating it):
    the z- / z+ patches of the local block are the processor patches towards rank-1 / rank+1."""
    m = block_mesh(n, n, n, lengths=(1.0, 1.0, 1.0), velocity=velocity)
    ps = m["patch_start"]
    lo_c, hi_c = rank > 0, rank < world - 1
    first = ps[4] if lo_c else ps[5]
    count = (ps[5] - ps[4] if lo_c else 0) + (ps[6] - ps[5] if hi_c else 0)
    groups, start = [], 0
    if lo_c:
        groups.append((rank - 1, 0, int(ps[5] - ps[4]))); start = int(ps[5] - ps[4])
    if hi_c:
        groups.append((rank + 1, start, int(ps[6] - ps[5])))
    send = m["boundary_owner"][first:first + count].copy()
    m.update({"coupled": (int(first) if count else int(ps[6]), int(count)), "neighbours": groups,
              "send_cells": send.astype(np.int32), "cell_range": (rank * n ** 3, (rank + 1) * n ** 3)})
    return m
