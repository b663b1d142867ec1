"""Row f2 (first slice): explicit upwind transport + source update on OpenFOAM face addressing.

CPU part: properties of the oracle restatement (parity unpinned: OpenFOAM is not on this machine,
so the restated MULES::explicitSolve / upwind / surfaceIntegrate loops are checked against what the
scheme must do analytically).  GPU part: the device kernels against the oracle, bit for bit.
"""
import numpy as np
import pytest

from oracle import cpu_oracle as orc
from openpbe_b200 import capi, synthetic


def chain_mesh(n, u=1.0, dx=1.0):
    m = synthetic.block_mesh(n, 1, 1, lengths=(n * dx, 1.0, 1.0), velocity=(u, 0.0, 0.0))
    return m


def random_mesh(rng, ncells, nint, nbnd):
    """Any owner < neighbour pair list is a valid face addressing for the update (topology only)."""
    a = rng.integers(0, ncells, size=nint)
    b = (a + 1 + rng.integers(0, ncells - 1, size=nint)) % ncells
    own, nei = np.minimum(a, b), np.maximum(a, b)
    order = np.lexsort((nei, own))
    return {"owner": own[order].astype(np.int32), "neighbour": nei[order].astype(np.int32),
            "boundary_owner": rng.integers(0, ncells, size=nbnd).astype(np.int32),
            "vcell": rng.uniform(0.5, 2.0, ncells), "phi": rng.normal(size=nint + nbnd)}


def test_block_mesh_addressing():
    m = synthetic.block_mesh(4, 3, 2)
    assert m["owner"].size == 3 * 3 * 2 + 4 * 2 * 2 + 4 * 3 * 1
    assert np.all(m["owner"] < m["neighbour"])
    assert np.all(np.diff(m["owner"]) >= 0)  # upper-triangular order
    assert m["boundary_owner"].size == 2 * (3 * 2 + 4 * 2 + 4 * 3)
    # a uniform velocity is divergence-free: per-cell flux sums vanish
    div = np.zeros(24)
    ni = m["owner"].size
    np.add.at(div, m["owner"], m["phi"][:ni]); np.add.at(div, m["neighbour"], -m["phi"][:ni])
    np.add.at(div, m["boundary_owner"], m["phi"][ni:])
    assert np.abs(div).max() < 1e-15


def test_oracle_cfl_one_shift_is_exact():
    # 1-D chain, u dt / dx = 1: upwind moves every value exactly one cell downstream
    n = 16
    m = chain_mesh(n)
    psi = np.arange(1.0, n + 1)[None, :].repeat(2, 0) * np.array([[1.0], [3.0]])
    psi_b = np.zeros((2, m["boundary_owner"].size)); psi_b[:, 0] = [7.0, 9.0]  # inflow values at x-
    psi_b[:, 1] = psi[:, -1]  # zero-gradient at the outlet; side patches carry no flux
    out = orc.transport_upwind_explicit(m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"], psi,
                                        m["phi"], 1.0, psi_b=psi_b)
    np.testing.assert_array_equal(out[:, 1:], psi[:, :-1])
    np.testing.assert_array_equal(out[:, 0], [7.0, 9.0])


def test_oracle_conservation_and_source():
    rng = np.random.default_rng(5)
    m = synthetic.block_mesh(7, 5, 3, velocity=(0.3, -0.2, 0.1))
    ni = m["owner"].size
    phi = m["phi"].copy(); phi[:ni] = rng.normal(size=ni) * 1e-3; phi[ni:] = 0.0  # closed box
    psi = rng.uniform(0.0, 1.0, (3, 105)); S = rng.normal(size=(3, 105))
    dt = 0.01
    out = orc.transport_upwind_explicit(m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"], psi, phi, dt, S=S)
    V = m["vcell"]
    np.testing.assert_allclose((out * V).sum(1), (psi * V).sum(1) + dt * (S * V).sum(1), rtol=1e-13)
    # zero flux: the MULES form of n + dt S (MOC.C:259-272 as the 0-D integration tests run it)
    out0 = orc.transport_upwind_explicit(m["owner"], m["neighbour"], m["boundary_owner"], V, psi, 0 * phi, dt, S=S)
    np.testing.assert_allclose(out0, psi + dt * S, rtol=1e-14)


def test_oracle_upwind_is_bounded():
    # divergence-free flux, CFL < 1, zero-gradient outflow / fixed inflow: no new extrema
    rng = np.random.default_rng(6)
    m = synthetic.block_mesh(9, 8, 4, velocity=(1.0, 0.5, -0.25))
    psi = rng.uniform(1.0, 2.0, (2, 288))
    psi_b = rng.uniform(1.0, 2.0, (2, m["boundary_owner"].size))
    outflow = m["phi"][m["owner"].size:] >= 0
    psi_b[:, outflow] = psi[:, m["boundary_owner"][outflow]]  # zeroGradient where the flow leaves
    dt = 0.4 * (1 / 9) / 1.75
    out = orc.transport_upwind_explicit(m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"], psi, m["phi"], dt, psi_b=psi_b)
    assert out.min() >= 1.0 - 1e-12 and out.max() <= 2.0 + 1e-12


def test_div_scheme_names():
    out = capi.C.c_int(-1)
    assert capi.lib().pbe_div_scheme_type(b"Gauss upwind", capi.C.byref(out)) == capi.PBE_OK and out.value == capi.DIV_UPWIND
    # the scheme of the reference's validation case (simmons_and_azzopardi/c01/system/fvSchemes:34)
    assert capi.lib().pbe_div_scheme_type(b"Gauss vanLeer", capi.C.byref(out)) == capi.PBE_OK and out.value == capi.DIV_VANLEER
    # the scheme the reference's MOC cases put on the implicit term, div(U,nk) (tests/integration/MOC/template/system/fvSchemes:44)
    assert capi.lib().pbe_div_scheme_type(b"Gauss limitedLinear 1", capi.C.byref(out)) == capi.PBE_OK and out.value == capi.DIV_LIMITEDLINEAR
    assert capi.lib().pbe_div_scheme_type(b"Gauss limitedLinear 0.5", capi.C.byref(out)) == capi.PBE_E_UNKNOWN_TYPE  # only k = 1
    rc = capi.lib().pbe_div_scheme_type(b"Gauss QUICK", capi.C.byref(out))
    assert rc == capi.PBE_E_UNKNOWN_TYPE
    msg = capi.last_error()
    assert "QUICK" in msg and "Valid schemes are" in msg and "upwind" in msg and "vanLeer" in msg and "limitedLinear 1" in msg


# ------------------------------------------------------------------------------------------------
gpu = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    from openpbe_b200 import engine
    return engine.Context(0)


def _run_device(ctx, m, psi, phi, dt, psi_b=None, S=None):
    import torch
    from openpbe_b200 import engine
    dev = torch.device("cuda:0")
    T = lambda a: None if a is None else torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    mesh = engine.Mesh(ctx, m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"])
    out = mesh.transport_explicit(T(psi), T(phi), dt, psi_b=T(psi_b), S=T(S))
    return out.cpu().numpy(), mesh.info()


@gpu
@pytest.mark.parametrize("shape", [(12, 9, 7), (33, 1, 1), (17, 13, 1), (300, 40, 9)])
@pytest.mark.parametrize("K", [1, 6, 30])
def test_device_matches_oracle_block(ctx, shape, K):
    rng = np.random.default_rng(shape[0] * 100 + K)
    m = synthetic.block_mesh(*shape, velocity=(1.0, -0.7, 0.3))
    nc = m["vcell"].size
    phi = m["phi"] * rng.uniform(0.5, 1.5, m["phi"].size) * rng.choice([1.0, 1.0, -1.0, 0.0], m["phi"].size)
    psi = rng.uniform(0.0, 1e8, (K, nc)) * rng.choice([1.0, 1e-20], (K, nc))
    S = rng.normal(size=(K, nc)) * 1e9
    psi_b = rng.uniform(0.0, 1e8, (K, m["boundary_owner"].size))
    dt = 2.5e-4
    for kw in ({}, {"psi_b": psi_b}, {"S": S}, {"psi_b": psi_b, "S": S}):
        want = orc.transport_upwind_explicit(m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"], psi, phi, dt, **kw)
        got, info = _run_device(ctx, m, psi, phi, dt, **kw)
        assert info["max_faces_per_cell"] == 6
        # every hexahedron has six faces, boundary ones included: all cells take the ELL kernel
        assert info["regular_cells"] == nc and info["faces_per_regular_cell"] == 6
        np.testing.assert_array_equal(got, want)


@gpu
@pytest.mark.parametrize("nint,nbnd", [(0, 0), (40, 3), (900, 120), (5000, 0)])
def test_device_matches_oracle_random_topology(ctx, nint, nbnd):
    # polyhedral-style connectivity: many faces per cell (generic kernel), cells without faces
    rng = np.random.default_rng(nint + nbnd)
    m = random_mesh(rng, 301, nint, nbnd)
    psi = rng.normal(size=(5, 301)); S = rng.normal(size=(5, 301))
    psi_b = rng.normal(size=(5, nbnd)) if nbnd else None
    want = orc.transport_upwind_explicit(m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"], psi, m["phi"], 0.01, psi_b=psi_b, S=S)
    got, _ = _run_device(ctx, m, psi, m["phi"], 0.01, psi_b=psi_b, S=S)
    np.testing.assert_array_equal(got, want)


@gpu
def test_device_nonfinite_pattern_matches(ctx):
    rng = np.random.default_rng(9)
    m = synthetic.block_mesh(10, 6, 4, velocity=(1.0, 1.0, 1.0))
    psi = rng.uniform(size=(3, 240)); psi[0, 17] = np.nan; psi[1, 100] = np.inf; psi[2, 5] = -np.inf
    want = orc.transport_upwind_explicit(m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"], psi, m["phi"], 0.01)
    got, _ = _run_device(ctx, m, psi, m["phi"], 0.01)
    np.testing.assert_array_equal(got, want)  # NaN == NaN positionally in assert_array_equal


@gpu
def test_device_argument_errors(ctx):
    import torch
    from openpbe_b200 import engine
    m = synthetic.block_mesh(4, 4, 4)
    mesh = engine.Mesh(ctx, m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"])
    psi = torch.zeros(2, 64, dtype=torch.float64, device="cuda:0")
    phi = torch.zeros(m["phi"].size, dtype=torch.float64, device="cuda:0")
    with pytest.raises(capi.PbeError, match="alias"):
        mesh.transport_explicit(psi, phi, 0.1, out=psi)
    with pytest.raises(capi.PbeError, match="pbe_transport_mules"):  # limited schemes take the MULES path
        mesh.transport_explicit(psi, phi, 0.1, scheme="Gauss vanLeer")
    with pytest.raises(capi.PbeError, match="owner"):
        engine.Mesh(ctx, [0, 70], [1, 2], [], np.ones(64))
    assert mesh.transport_explicit(psi[:0], phi, 0.1).shape == (0, 64)


# ---- decomposed meshes (row f2 x row e): partition, halo exchange, coupled faces ----------------
def _decomposed_oracle(m, world, psi, dt, S=None, psi_b=None):
    """Every sub-domain through the coupled oracle, halos taken straight from the global field."""
    from openpbe_b200 import sharding
    out = np.empty_like(psi)
    parts = []
    for r in range(world):
        p = sharding.partition_mesh(m, r, world)
        lo, hi = p["cell_range"]
        nphys, ncpl = p["coupled"]
        pb = None
        if psi_b is not None:
            pb = np.zeros((psi.shape[0], nphys + ncpl)); pb[:, :nphys] = psi_b[:, p["phys_faces"]]
        out[:, lo:hi] = orc.transport_upwind_explicit(
            p["owner"], p["neighbour"], p["boundary_owner"], p["vcell"], psi[:, lo:hi], p["phi"], dt, psi_b=pb,
            S=None if S is None else np.ascontiguousarray(S[:, lo:hi]), coupled=p["coupled"],
            psi_halo=psi[:, p["halo_cells_global"]])
        parts.append(p)
    return out, parts


@pytest.mark.parametrize("world", [2, 3])
def test_partition_matches_serial_to_rounding(world):
    rng = np.random.default_rng(world)
    m = synthetic.block_mesh(6, 5, 8, velocity=(0.3, -0.2, 0.4))
    m["phi"] = m["phi"] * rng.uniform(0.5, 1.5, m["phi"].size) * rng.choice([1.0, -1.0], m["phi"].size)
    psi = rng.uniform(1.0, 2.0, (3, 240)); S = rng.normal(size=(3, 240))
    psi_b = rng.uniform(1.0, 2.0, (3, m["boundary_owner"].size))
    serial = orc.transport_upwind_explicit(m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"], psi, m["phi"], 0.01, psi_b=psi_b, S=S)
    dec, parts = _decomposed_oracle(m, world, psi, 0.01, S=S, psi_b=psi_b)
    np.testing.assert_allclose(dec, serial, rtol=1e-13)
    # only cells on a processor patch may differ (coupled interpolation rounds differently)
    touched = np.zeros(240, bool)
    for r, p in enumerate(parts):
        touched[p["cell_range"][0] + p["send_cells"]] = True
        # both sides list the shared faces in the same order
        for q, start, cnt in p["neighbours"]:
            back = [g for g in parts[q]["neighbours"] if g[0] == r][0]
            np.testing.assert_array_equal(p["halo_cells_global"][start:start + cnt],
                                          parts[q]["cell_range"][0] + parts[q]["send_cells"][back[1]:back[1] + back[2]])
    np.testing.assert_array_equal(dec[:, ~touched], serial[:, ~touched])
    assert touched.sum() < 240


def test_block_slab_part_is_the_partition_of_the_global_block():
    from openpbe_b200 import sharding
    n, world = 4, 3
    g = synthetic.block_mesh(n, n, n * world, lengths=(1.0, 1.0, float(world)), velocity=(1.0, 0.5, 0.25))
    for r in range(world):
        a = sharding.partition_mesh(g, r, world)
        b = synthetic.block_slab_part(n, r, world)
        for k in ("owner", "neighbour", "send_cells"):
            np.testing.assert_array_equal(a[k], b[k])
        assert a["coupled"][1] == b["coupled"][1] and a["neighbours"] == b["neighbours"]
        # same faces; the global block lists its z patches after the side patches just like the slab
        nphys = a["coupled"][0]
        assert sorted(a["boundary_owner"][:nphys]) == sorted(np.delete(b["boundary_owner"], np.arange(b["coupled"][0], b["coupled"][0] + b["coupled"][1])))
        np.testing.assert_array_equal(a["boundary_owner"][nphys:], b["boundary_owner"][b["coupled"][0]:b["coupled"][0] + b["coupled"][1]])
        np.testing.assert_allclose(a["phi"][nphys + a["owner"].size:], b["phi"][b["owner"].size + b["coupled"][0]:][:b["coupled"][1]], rtol=1e-15)


def _halo_worker(rank, world, port, out):
    import os
    import torch
    import torch.distributed as dist
    from openpbe_b200 import sharding
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(11)
    m = synthetic.block_mesh(5, 4, 6, velocity=(0.3, 0.2, 0.4))
    psi = rng.uniform(1.0, 2.0, (4, 120))
    p = sharding.partition_mesh(m, rank, world)
    lo, hi = p["cell_range"]
    hx = sharding.HaloExchange(p, "cpu")
    halo = hx.exchange(torch.from_numpy(np.ascontiguousarray(psi[:, lo:hi]))).numpy()
    np.testing.assert_array_equal(halo, psi[:, p["halo_cells_global"]])
    new = orc.transport_upwind_explicit(p["owner"], p["neighbour"], p["boundary_owner"], p["vcell"], psi[:, lo:hi],
                                        p["phi"], 0.02, coupled=p["coupled"], psi_halo=halo)
    parts = [None] * world
    dist.all_gather_object(parts, (lo, hi, new))
    if rank == 0:
        full = np.empty_like(psi)
        for a, b, v in parts:
            full[:, a:b] = v
        np.save(out, full)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_halo_exchange_gloo(tmp_path):
    import os
    import torch.multiprocessing as mp
    out = str(tmp_path / "t.npy")
    port = 31500 + (os.getpid() % 2000)
    mp.spawn(_halo_worker, args=(2, port, out), nprocs=2, join=True)
    rng = np.random.default_rng(11)
    m = synthetic.block_mesh(5, 4, 6, velocity=(0.3, 0.2, 0.4))
    psi = rng.uniform(1.0, 2.0, (4, 120))
    want, _ = _decomposed_oracle(m, 2, psi, 0.02)
    np.testing.assert_array_equal(np.load(out), want)


@gpu
@pytest.mark.parametrize("world", [2, 3])
def test_device_decomposed_matches_oracle(ctx, world):
    import torch
    from openpbe_b200 import engine, sharding
    rng = np.random.default_rng(40 + world)
    m = synthetic.block_mesh(11, 9, 12, velocity=(0.3, -0.2, 0.4))
    m["phi"] = m["phi"] * rng.uniform(0.5, 1.5, m["phi"].size) * rng.choice([1.0, -1.0, 0.0], m["phi"].size)
    nc = 11 * 9 * 12
    psi = rng.uniform(0.0, 2.0, (7, nc)); S = rng.normal(size=(7, nc))
    psi_b = rng.uniform(1.0, 2.0, (7, m["boundary_owner"].size))
    want, parts = _decomposed_oracle(m, world, psi, 0.01, S=S, psi_b=psi_b)
    dev = torch.device("cuda:0")
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    for p in parts:
        lo, hi = p["cell_range"]
        nphys, ncpl = p["coupled"]
        mesh = engine.Mesh(ctx, p["owner"], p["neighbour"], p["boundary_owner"], p["vcell"])
        mesh.set_coupled(nphys, ncpl)
        pb = np.zeros((7, nphys + ncpl)); pb[:, :nphys] = psi_b[:, p["phys_faces"]]
        loc = T(psi[:, lo:hi])
        # what this rank would send is what its neighbours' halos hold
        sent = mesh.gather_cells(loc, T(p["send_cells"])).cpu().numpy()
        np.testing.assert_array_equal(sent, psi[:, lo + p["send_cells"]])
        with pytest.raises(capi.PbeError, match="coupled"):
            mesh.transport_explicit(loc, T(p["phi"]), 0.01)
        got = mesh.transport_explicit(loc, T(p["phi"]), 0.01, psi_b=T(pb), S=T(S[:, lo:hi]),
                                      psi_halo=T(psi[:, p["halo_cells_global"]]))
        np.testing.assert_array_equal(got.cpu().numpy(), want[:, lo:hi])
        # the same in two parts (interior first, then boundary / processor-patch cells): the overlap form
        two = torch.full_like(got, float("nan"))
        mesh.transport_explicit(loc, T(p["phi"]), 0.01, psi_b=T(pb), S=T(S[:, lo:hi]), out=two, cells=capi.CELLS_REGULAR)
        mesh.transport_explicit(loc, T(p["phi"]), 0.01, psi_b=T(pb), S=T(S[:, lo:hi]), out=two,
                                psi_halo=T(psi[:, p["halo_cells_global"]]), cells=capi.CELLS_IRREGULAR)
        np.testing.assert_array_equal(two.cpu().numpy(), want[:, lo:hi])


@gpu
def test_qmom_correct_with_transport_sequence(ctx):
    """QMOM::correct() with mesh transport (QMOM.C:85-129, 131-197): sources -> NaN scrub ->
    explicit upwind transport of the 2*order moments -> clamp to SMALL -> d, each device step
    against its restatement."""
    import torch
    from openpbe_b200 import engine
    wq = synthetic.workload("qmom3_ct", 12 * 10 * 8)
    fq = synthetic.qmom_fields(wq)
    tm = synthetic.block_mesh(12, 10, 8, velocity=(1.0, -0.5, 0.25))
    dev = torch.device("cuda:0")
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    q = engine.QMOM(ctx, wq.N, engine.make_spec(wq.coalescence, wq.coal_coeffs, wq.breakup, wq.br_coeffs))
    m = T(fq["m"])
    S = q.source(m, T(fq["eps"]), T(fq["rho"]), T(fq["nu"]))
    S_raw = S.cpu().numpy().copy()
    S_raw[0, 3] = np.nan  # at least one NaN to scrub, whatever the synthetic cells give
    S = T(S_raw)
    ctx.scrub_nan(S)
    S_scrubbed = np.where(np.isnan(S_raw), 0.0, S_raw)
    np.testing.assert_array_equal(S.cpu().numpy(), S_scrubbed)
    mesh = engine.Mesh(ctx, tm["owner"], tm["neighbour"], tm["boundary_owner"], tm["vcell"])
    dt = 1e-3
    m_new = mesh.transport_explicit(m, T(tm["phi"]), dt, S=S)
    want = orc.transport_upwind_explicit(tm["owner"], tm["neighbour"], tm["boundary_owner"], tm["vcell"], fq["m"],
                                         tm["phi"], dt, S=S_scrubbed)
    np.testing.assert_array_equal(m_new.cpu().numpy(), want)
    m_new[2, 7] = -1.0  # something to clamp
    m_new[3, 9] = float("nan")
    d = ctx.qmom_clamp_d(m_new)
    want[2, 7] = -1.0; want[3, 9] = np.nan
    clamped = np.where(want > 1e-15, want, 1e-15)
    np.testing.assert_array_equal(m_new.cpu().numpy(), clamped)
    np.testing.assert_allclose(d.cpu().numpy(), np.cbrt(6.0 / np.pi * clamped[1] / clamped[0]), rtol=4e-16)


@gpu
@pytest.mark.parametrize("with_patch_values", [False, True])
def test_moc_correct_host_equals_resident_path(ctx, with_patch_values):
    """pbe_moc_correct_host (the adapter's entry: host buffers, chunked uploads, sources and
    transport on the device) gives exactly what the resident calls give."""
    import torch
    from openpbe_b200 import engine
    tm = synthetic.block_mesh(20, 16, 12, velocity=(1.0, -0.5, 0.25))
    nc = 20 * 16 * 12
    w = synthetic.workload("moc30_ct", nc)
    f = synthetic.moc_fields(w)
    moc = engine.MOC(ctx, w.N, w.xi1, engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs), w.relax)
    mesh = engine.Mesh(ctx, tm["owner"], tm["neighbour"], tm["boundary_owner"], tm["vcell"])
    dev = torch.device("cuda:0")
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    psi_b = None
    if with_patch_values:
        psi_b = np.ascontiguousarray(f["n"][:, tm["boundary_owner"]] * 0.5)
    n = T(f["n"])
    S = moc.source(n, T(f["eps"]), T(f["rho"]), T(f["nu"]), w.dt)
    want = mesh.transport_explicit(n, T(tm["phi"]), w.dt, S=S, psi_b=None if psi_b is None else T(psi_b))
    want_d = moc.d(want).cpu().numpy()
    n_host = torch.from_numpy(f["n"].copy())
    d = mesh.moc_correct_host(moc, n_host, f["eps"], f["rho"], f["nu"], tm["phi"], w.dt, psi_b=psi_b)
    np.testing.assert_array_equal(n_host.numpy(), want.cpu().numpy())
    np.testing.assert_array_equal(d.numpy(), want_d)
    # and against the oracle's face loop on the device sources
    ref = orc.transport_upwind_explicit(tm["owner"], tm["neighbour"], tm["boundary_owner"], tm["vcell"], f["n"],
                                        tm["phi"], w.dt, S=S.cpu().numpy(), psi_b=psi_b)
    np.testing.assert_array_equal(n_host.numpy(), ref)


def circulant_mesh(rng, n, degree):
    """Every cell has exactly `degree` internal faces and no boundary face (offsets 1, 2, (3 |
    antipode)): the three register-resident kernels (F = 4, 5, 6) each get a mesh of their own."""
    offs = {4: (1, 2), 5: (1, 2), 6: (1, 2, 3)}[degree]
    pairs = {(min(c, (c + o) % n), max(c, (c + o) % n)) for c in range(n) for o in offs}
    if degree == 5:
        pairs |= {(c, c + n // 2) for c in range(n // 2)}
    own, nei = map(np.array, zip(*sorted(pairs)))
    return {"owner": own.astype(np.int32), "neighbour": nei.astype(np.int32),
            "boundary_owner": np.zeros(0, np.int32), "vcell": rng.uniform(0.5, 2.0, n),
            "phi": rng.normal(size=own.size)}


@gpu
@pytest.mark.parametrize("degree", [4, 5, 6])
@pytest.mark.parametrize("kernel", ["0", "2", "3"])
def test_device_regular_kernels_every_face_count(ctx, degree, kernel, monkeypatch):
    monkeypatch.setenv("PBE_TRANSPORT_KERNEL", kernel)  # register-staged / cp.async 2 and 3 stages
    rng = np.random.default_rng(degree)
    m = circulant_mesh(rng, 1000, degree)
    psi = rng.normal(size=(9, 1000)); S = rng.normal(size=(9, 1000))
    want = orc.transport_upwind_explicit(m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"], psi, m["phi"], 0.05, S=S)
    got, info = _run_device(ctx, m, psi, m["phi"], 0.05, S=S)
    assert info["faces_per_regular_cell"] == degree and info["regular_cells"] == 1000
    np.testing.assert_array_equal(got, want)


def test_c_oracle_equals_numpy_restatement():
    from oracle import numpy_restatement as npo
    rng = np.random.default_rng(77)
    m = synthetic.block_mesh(9, 7, 5, velocity=(0.3, -0.2, 0.4))
    m["phi"] = m["phi"] * rng.uniform(0.5, 1.5, m["phi"].size) * rng.choice([1.0, -1.0, 0.0], m["phi"].size)
    psi = rng.uniform(0.0, 2.0, (4, 315)); S = rng.normal(size=(4, 315))
    psi_b = rng.uniform(0.0, 2.0, (4, m["boundary_owner"].size))
    for kw in ({}, {"S": S}, {"psi_b": psi_b, "S": S}):
        a = orc.transport_upwind_explicit(m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"], psi, m["phi"], 0.01, **kw)
        b = npo.transport_upwind_explicit(m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"], psi, m["phi"], 0.01, **kw)
        np.testing.assert_array_equal(a, b)
    # a sub-domain with processor-patch faces
    from openpbe_b200 import sharding
    p = sharding.partition_mesh(m, 1, 3)
    lo, hi = p["cell_range"]
    halo = psi[:, p["halo_cells_global"]]
    a = orc.transport_upwind_explicit(p["owner"], p["neighbour"], p["boundary_owner"], p["vcell"], psi[:, lo:hi], p["phi"],
                                      0.01, coupled=p["coupled"], psi_halo=halo)
    b = npo.transport_upwind_explicit(p["owner"], p["neighbour"], p["boundary_owner"], p["vcell"], psi[:, lo:hi], p["phi"],
                                      0.01, coupled=p["coupled"], psi_halo=halo)
    np.testing.assert_array_equal(a, b)


def test_oracle_first_order_convergence_1d():
    # smooth profile advected over a fixed time: the upwind error halves when dx and dt halve
    errs = []
    for n in (50, 100, 200):
        m = chain_mesh(n, u=1.0, dx=1.0 / n)
        x = (np.arange(n) + 0.5) / n
        psi = np.exp(-200.0 * (x - 0.3) ** 2)[None, :]
        dt = 0.5 / n
        for _ in range(int(round(0.3 / dt))):
            psi = orc.transport_upwind_explicit(m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"], psi, m["phi"], dt)
        exact = np.exp(-200.0 * (x - 0.6) ** 2)
        errs.append(np.abs(psi[0] - exact).sum() / n)
    assert errs[1] < 0.75 * errs[0] and errs[2] < 0.7 * errs[1]


@gpu
def test_device_padded_strides_through_the_c_abi(ctx):
    """stride > ncells and stride_b > boundary faces (fields that are rows of bigger allocations)."""
    import torch
    from openpbe_b200 import engine
    rng = np.random.default_rng(123)
    m = synthetic.block_mesh(13, 9, 6, velocity=(1.0, -0.7, 0.3))
    nc, nb, K, pad, padb = 13 * 9 * 6, m["boundary_owner"].size, 5, 37, 11
    psi = rng.normal(size=(K, nc)); S = rng.normal(size=(K, nc)); psi_b = rng.normal(size=(K, nb))
    want = orc.transport_upwind_explicit(m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"], psi, m["phi"], 0.02,
                                         psi_b=psi_b, S=S)
    dev = torch.device("cuda:0")
    big = lambda a, p: torch.from_numpy(np.pad(a, ((0, 0), (0, p)), constant_values=np.nan)).to(dev)
    Psi, Ss, Pb = big(psi, pad), big(S, pad), big(psi_b, padb)
    out = torch.full((K, nc + pad), -7.0, dtype=torch.float64, device=dev)
    phi = torch.from_numpy(m["phi"]).to(dev)
    mesh = engine.Mesh(ctx, m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"])
    capi.check(capi.lib().pbe_transport_explicit(mesh._h, capi.DIV_UPWIND, K, nc + pad, Psi.data_ptr(), phi.data_ptr(),
                                                 Pb.data_ptr(), nb + padb, Ss.data_ptr(), 0.02, out.data_ptr(), None))
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    np.testing.assert_array_equal(got[:, :nc], want)
    assert np.all(got[:, nc:] == -7.0)  # the padding is never written
    # halo packing with padded rows on both sides
    cells = torch.from_numpy(rng.integers(0, nc, 50).astype(np.int32)).to(dev)
    packed = torch.full((K, 64), -1.0, dtype=torch.float64, device=dev)
    capi.check(capi.lib().pbe_gather_cells(ctx.handle, K, nc + pad, Psi.data_ptr(), 50, cells.data_ptr(),
                                           packed.data_ptr(), 64, None))
    torch.cuda.synchronize()
    np.testing.assert_array_equal(packed.cpu().numpy()[:, :50], psi[:, cells.cpu().numpy()])
    assert np.all(packed.cpu().numpy()[:, 50:] == -1.0)
