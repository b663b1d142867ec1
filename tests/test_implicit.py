"""Row f2, third slice: the implicit transport MOC::solveWithFVM (MOC.C:186-198) / QMOM::solveWithFVM
(QMOM.C:201-209) -- the reference's default branch (usingMULES false).

PARITY UNPINNED: fvMatrix and the linear solvers are OpenFOAM's.  What is checked:
  * CPU: oracle/fvm_implicit.py (OpenFOAM's assembly restated, solved directly) has the properties the scheme must
    have -- zero flux reduces to the explicit update, the update conserves, stays positive at any Courant number,
    relax(1) leaves a dominant matrix alone;
  * CPU: the reference's OWN MOC.C / QMOM.C with usingMULES false, compiled unmodified (oracle/_ref) and driven through
    the shimmed fvm:: calls, land on the same solution as the direct solve of the restated matrix;
  * GPU: pbe_transport_implicit (all fields at once, BiCGStab on the device) reaches the direct solution to the
    tolerance asked, reports OpenFOAM's normalised residuals, and stops by OpenFOAM's criterion.
"""
import numpy as np
import pytest

from oracle import fvm_implicit as fvm
from oracle import ref_oracle as ref
from openpbe_b200 import synthetic

needs_ref = pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built (no /root/reference at build time)")


def random_mesh(rng, ncells, nint, nbnd):
    a = rng.integers(0, ncells, size=nint)
    b = (a + 1 + rng.integers(0, ncells - 1, size=nint)) % ncells
    own, nei = np.minimum(a, b), np.maximum(a, b)
    order = np.lexsort((nei, own))
    return {"owner": own[order].astype(np.int32), "neighbour": nei[order].astype(np.int32),
            "boundary_owner": rng.integers(0, ncells, size=nbnd).astype(np.int32),
            "vcell": rng.uniform(0.5, 2.0, ncells), "phi": rng.normal(size=nint + nbnd) * 0.05}


def fields(rng, K, nc):
    return rng.uniform(0.1, 2.0, size=(K, nc)) * (1.0 + np.sin(np.arange(nc) * 0.3)[None, :] ** 2)


# ---- CPU: properties of the restatement -------------------------------------------------------------------
def test_zero_flux_is_the_explicit_update():
    rng = np.random.default_rng(1)
    m = synthetic.block_mesh(5, 4, 3, velocity=(0.0, 0.0, 0.0))
    psi = fields(rng, 3, 60)
    S = rng.normal(size=psi.shape)
    out = fvm.solve(m, m["phi"], 0.01, psi, S=S)
    assert np.allclose(out, psi + 0.01 * S, rtol=1e-14, atol=0)


@pytest.mark.parametrize("courant", [0.3, 5.0, 80.0])
def test_implicit_upwind_conserves_and_stays_positive(courant):
    rng = np.random.default_rng(2)
    m = synthetic.block_mesh(12, 9, 1, velocity=(1.0, -0.5, 0.0))
    nc = 12 * 9
    dx = 1.0 / 12
    dt = courant * dx / 1.0
    psi = fields(rng, 4, nc)
    psi[:, rng.integers(0, nc, 20)] = 0.0
    out = fvm.solve(m, m["phi"], dt, psi)  # zero-gradient patches
    assert out.min() >= -1e-15  # M-matrix: positivity at any Courant number
    # conservation: d/dt sum(psi V) = - boundary outflow of the NEW field
    ni = m["owner"].size
    bflux = np.zeros(4)
    for k in range(4):
        bflux[k] = np.sum(m["phi"][ni:] * out[k][m["boundary_owner"]])
    lhs = ((out - psi) * m["vcell"]).sum(axis=1) / dt
    assert np.allclose(lhs, -bflux, rtol=1e-10, atol=1e-12)


def test_relaxation_one_keeps_a_dominant_matrix_and_damps_otherwise():
    rng = np.random.default_rng(3)
    m = synthetic.block_mesh(8, 6, 1, velocity=(1.0, 0.3, 0.0))
    psi = fields(rng, 2, 48)
    dt = 0.01
    pb = rng.uniform(0.0, 1.0, size=(2, m["boundary_owner"].size))
    a = fvm.solve(m, m["phi"], dt, psi, psi_b=pb)
    b = fvm.solve(m, m["phi"], dt, psi, psi_b=pb, relaxation=1.0)
    assert np.allclose(a, b, rtol=1e-13)  # uniform velocity, fixed-value patches: rows dominant, nothing to do
    # zero-gradient patches with inflow: relax() adds |phi_b| where the patch coefficient is negative (OpenFOAM's
    # "maximum magnitude diagonal contribution to ensure stability"): the matrix changes even at factor 1
    a = fvm.solve(m, m["phi"], dt, psi)
    assert not np.allclose(a, fvm.solve(m, m["phi"], dt, psi, relaxation=1.0), rtol=1e-6)
    c = fvm.solve(m, m["phi"], dt, psi, relaxation=0.5)
    # under-relaxed: the new field moves about half as far from the old one
    assert np.all(np.abs(c - psi).sum(axis=1) < 0.62 * np.abs(a - psi).sum(axis=1))
    assert np.all(np.abs(c - psi).sum(axis=1) > 0.38 * np.abs(a - psi).sum(axis=1))


def test_residual_normalisation_is_openfoams():
    rng = np.random.default_rng(4)
    m = synthetic.block_mesh(6, 5, 4, velocity=(0.4, 0.2, -0.7))
    psi = fields(rng, 1, 120)[0]
    A, b = fvm.assemble(m, m["phi"], 0.05, psi)
    x = np.asarray(np.linalg.solve(A.toarray(), b))
    assert fvm.residual(A, b, x, psi) < 1e-14
    r0 = fvm.residual(A, b, psi)
    assert 1e-4 < r0 < 1.0  # the old field is no solution, and the factor makes it O(1) at most


def test_limited_weights_are_the_mules_restatements(oracle):
    """The NumPy weights of the implicit term against the C restatement the MULES path is checked with: the vanLeer
    face flux phi (w (P - N) + N) must be what orc_div_flux gives; limitedLinear 1 stays between upwind and linear."""
    rng = np.random.default_rng(5)
    m = synthetic.block_mesh(9, 7, 5, velocity=(1.0, -0.5, 0.25))
    ni = m["owner"].size
    psi = fields(rng, 1, 9 * 7 * 5)[0]
    w = fvm.limited_weights(m, "Gauss vanLeer", m["phi"], psi)
    P, N = psi[m["owner"]], psi[m["neighbour"]]
    flux = oracle.div_flux(m, "Gauss vanLeer", m["phi"], psi)
    assert np.allclose(m["phi"][:ni] * (w * (P - N) + N), flux[:ni], rtol=1e-13, atol=1e-15)
    wl = fvm.limited_weights(m, "Gauss limitedLinear 1", m["phi"], psi)
    up = (m["phi"][:ni] >= 0).astype(float)
    lo, hi = np.minimum(up, m["weights"]), np.maximum(up, m["weights"])
    assert np.all(wl >= lo - 1e-15) and np.all(wl <= hi + 1e-15)
    assert np.any(wl == up) and np.any(np.abs(wl - m["weights"]) < 1e-15)  # both branches of the limiter occur


@pytest.mark.parametrize("scheme", ["Gauss vanLeer", "Gauss limitedLinear 1"])
def test_limited_implicit_is_less_diffusive_and_conserves(scheme):
    rng = np.random.default_rng(6)
    m = synthetic.block_mesh(40, 1, 1, velocity=(1.0, 0.0, 0.0))
    x = (np.arange(40) + 0.5) / 40
    psi = np.exp(-((x - 0.3) / 0.08) ** 2)[None, :]
    dt = 0.2 / 40
    a, b = psi, psi
    for _ in range(20):
        a = fvm.solve(m, m["phi"], dt, a)
        b = fvm.solve(m, m["phi"], dt, b, scheme=scheme)
    assert b.max() > a.max() * 1.05  # the peak survives better
    assert np.allclose((a * m["vcell"]).sum(), (b * m["vcell"]).sum(), rtol=1e-6)  # nothing has reached the outlet yet


@needs_ref
@pytest.mark.parametrize("method", ["moc", "qmom"])
def test_reference_solve_with_fvm_lands_on_the_direct_solution(oracle, method):
    """The reference's own sources with usingMULES false (its default): MOC.C:177-180 -> solveWithFVM -> the shimmed
    fvm::ddt + fvm::div == S, relax(), solve()."""
    tm = synthetic.block_mesh(9, 7, 5, velocity=(1.0, -0.5, 0.25))
    nc = 9 * 7 * 5
    dt = 0.02
    if method == "moc":
        w = synthetic.workload("moc30_ct", nc)
        w.N, w.xi1 = 6, 3e-10 / 6
        f = synthetic.moc_fields(w)
        f0 = f["n"] * (1.0 + 5.0 * np.sin(np.arange(nc) * 0.7)[None, :] ** 2)
        txt = ref.pbe_dict("methodOfClasses", w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs, N=w.N, xi1=w.xi1,
                           relax=w.relax, usingMULES=False)
        stem = "n"
    else:
        w = synthetic.workload("qmom3_const", nc)
        f = synthetic.qmom_fields(w, degenerate_fraction=0.0)
        f0 = f["m"]
        txt = ref.pbe_dict("quadratureMethodOfMoments", w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs, order=w.N,
                           usingMULES=False)
        stem = "m"
    try:
        ref.set_mesh(nc, tm["owner"], tm["neighbour"], tm["boundary_owner"], Sf=tm["Sf"], C_=tm["C"], weights=tm["weights"],
                     div_scheme="Gauss upwind")
        f_ref, S_ref, _, _ = ref.method_steps(txt, stem, f0, f["eps"], f["rho"], f["nu"], dt, alpha=f.get("alpha"),
                                              vcell=tm["vcell"], phi=tm["phi"])
    finally:
        ref.clear_mesh()
    mine = fvm.solve(tm, tm["phi"], dt, f0, S=S_ref)
    if method == "qmom":
        mine = np.maximum(mine, 1e-15)  # QMOM.C:117-123
    scale = np.abs(f0).max(axis=1, keepdims=True)
    assert np.all(np.abs(f_ref - mine) <= 1e-11 * scale)


def _decomposed_cpu_rank(rank, world, port, out):
    import os
    import torch
    import torch.distributed as dist
    from openpbe_b200 import sharding
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(31)
    g = synthetic.block_mesh(7, 5, 8, velocity=(0.4, -0.3, 1.0))
    nc = 7 * 5 * 8
    psi = fields(rng, 3, nc)
    S = rng.normal(size=psi.shape) * 0.1
    p = sharding.partition_mesh(g, rank, world)
    lo, hi = p["cell_range"]
    hx = sharding.HaloExchange(p, "cpu")

    def exchange(f):
        return hx.exchange(torch.from_numpy(np.ascontiguousarray(f))).numpy()

    def allreduce(a):
        t = torch.from_numpy(np.array(a, dtype=np.float64, copy=True))
        dist.all_reduce(t)
        return t.numpy()

    x, it, (r0, r1) = fvm.jacobi_decomposed(p, 0.08, psi[:, lo:hi], S[:, lo:hi], exchange, allreduce, nc)
    parts = [None] * world
    dist.all_gather_object(parts, (lo, hi, x, it, r0, r1))
    if rank == 0:
        full = np.empty_like(psi)
        for a, b, v, *_ in parts:
            full[:, a:b] = v
        np.savez(out, x=full, it=[q[3] for q in parts], r0=np.stack([q[4] for q in parts]), r1=np.stack([q[5] for q in parts]))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_decomposed_jacobi_gloo(tmp_path):
    """The decomposed implicit solve's logic on the CPU, two ranks over gloo: processor patches as coupled rows, a halo
    exchange of the correction per sweep, all-reduced residual sums -- the gathered result is the direct solution of
    the undecomposed mesh, both ranks see the same residuals and stop at the same sweep, and the initial residual is
    the global one."""
    import os
    import torch.multiprocessing as mp
    out = str(tmp_path / "dec.npz")
    port = 33500 + (os.getpid() % 2000)
    mp.spawn(_decomposed_cpu_rank, args=(2, port, out), nprocs=2, join=True)
    rng = np.random.default_rng(31)
    g = synthetic.block_mesh(7, 5, 8, velocity=(0.4, -0.3, 1.0))
    psi = fields(rng, 3, 280)
    S = rng.normal(size=psi.shape) * 0.1
    want = fvm.solve(g, g["phi"], 0.08, psi, S=S)
    r = np.load(out)
    assert np.all(np.abs(r["x"] - want).sum(axis=1) <= 1e-11 * np.abs(want).sum(axis=1))
    assert r["it"][0] == r["it"][1] and np.array_equal(r["r0"][0], r["r0"][1]) and np.array_equal(r["r1"][0], r["r1"][1])
    for k in range(3):
        A, b = fvm.assemble(g, g["phi"], 0.08, psi[k], S=S[k])
        assert r["r0"][0][k] == pytest.approx(fvm.residual(A, b, psi[k]), rel=1e-12)


# ---- GPU ---------------------------------------------------------------------------------------------------
torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def ctx():
    from openpbe_b200 import engine
    assert torch.cuda.is_available(), "-m gpu tests need a CUDA device"
    c = engine.Context(0)
    yield c
    c.close()


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).cuda()


def device_solve(ctx, m, psi, phi, dt, S=None, psi_b=None, **kw):
    from openpbe_b200 import engine
    mesh = engine.Mesh(ctx, m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"])
    out, perf = mesh.transport_implicit(T(psi), T(phi), dt, S=None if S is None else T(S),
                                        psi_b=None if psi_b is None else T(psi_b), **kw)
    out = out.cpu().numpy()
    mesh.close()
    return out, perf


@pytest.mark.gpu
@pytest.mark.parametrize("shape,vel", [((12, 9, 7), (1.0, -0.5, 0.25)), ((40, 1, 1), (1.0, 0.0, 0.0)),
                                       ((17, 13, 1), (-0.3, 0.8, 0.0)), ((32, 24, 16), (0.7, 0.4, -0.9))])
@pytest.mark.parametrize("courant", [0.4, 6.0])
@pytest.mark.parametrize("solver", ["jacobi", "bicgstab"])
def test_device_implicit_reaches_the_direct_solution_block(ctx, shape, vel, courant, solver):
    rng = np.random.default_rng(10)
    m = synthetic.block_mesh(*shape, velocity=vel)
    nc = shape[0] * shape[1] * shape[2]
    K = 5
    psi = fields(rng, K, nc)
    S = rng.normal(size=psi.shape) * 0.1
    dt = courant / (max(abs(v) for v in vel) * max(shape))
    want = fvm.solve(m, m["phi"], dt, psi, S=S)
    got, perf = device_solve(ctx, m, psi, m["phi"], dt, S=S, tolerance=1e-13, solver=solver)
    scale = np.abs(want).max(axis=1, keepdims=True)
    # a normalised L1 residual of 1e-13 bounds the error of one cell by ~1e-13 x ncells (all of it in one place)
    assert np.all(np.abs(got - want) <= 1e-13 * max(nc, 100) * scale)
    assert np.all(np.abs(got - want).sum(axis=1) <= 1e-11 * np.abs(want).sum(axis=1))
    for k, (r0, r1, it, state) in enumerate(perf):
        A, b = fvm.assemble(m, m["phi"], dt, psi[k], S=S[k])
        assert state == 1 and it >= 1 and r1 < 1e-13
        assert r0 == pytest.approx(fvm.residual(A, b, psi[k]), rel=1e-10)
        assert fvm.residual(A, b, got[k], psi[k]) < 1e-12


@pytest.mark.gpu
@pytest.mark.parametrize("nint,nbnd,fixed", [(0, 4, False), (60, 5, True), (900, 120, False), (6000, 300, True)])
def test_device_implicit_random_topology_and_patches(ctx, nint, nbnd, fixed):
    """any face count per cell (the slot tables' wide form), fixed-value patches in the source, zero-gradient patches
    on the diagonal"""
    rng = np.random.default_rng(11 + nint)
    nc = 200
    m = random_mesh(rng, nc, nint, nbnd)
    m["vcell"] *= 1.0  # V / dt dominates the random (non-solenoidal) fluxes
    K = 3
    psi = fields(rng, K, nc)
    psi_b = rng.uniform(0.0, 1.0, size=(K, nbnd)) if fixed else None
    dt = 0.5
    want = fvm.solve(m, m["phi"], dt, psi, psi_b=psi_b)
    got, perf = device_solve(ctx, m, psi, m["phi"], dt, psi_b=psi_b, tolerance=1e-13)
    assert np.all(np.abs(got - want) <= 1e-10 * np.abs(want).max())
    assert all(p[3] == 1 for p in perf)


@pytest.mark.gpu
@pytest.mark.parametrize("scheme", ["Gauss vanLeer", "Gauss limitedLinear 1"])
@pytest.mark.parametrize("shape,vel", [((12, 9, 7), (1.0, -0.5, 0.25)), ((40, 1, 1), (1.0, 0.0, 0.0)), ((30, 24, 12), (0.7, 0.4, -0.9))])
def test_device_implicit_limited_schemes(ctx, scheme, shape, vel):
    """one matrix per field: the weights of the implicit term come from the field itself (limitedScheme::weights)"""
    from openpbe_b200 import engine
    rng = np.random.default_rng(15)
    m = synthetic.block_mesh(*shape, velocity=vel)
    nc = shape[0] * shape[1] * shape[2]
    K = 4
    psi = fields(rng, K, nc)
    S = rng.normal(size=psi.shape) * 0.1
    dt = 0.3 / sum(abs(v) * n for v, n in zip(vel, shape))
    for pb in (None, rng.uniform(0.1, 2.0, size=(K, m["boundary_owner"].size))):
        want = fvm.solve(m, m["phi"], dt, psi, S=S, psi_b=pb, scheme=scheme)
        assert np.abs(want - fvm.solve(m, m["phi"], dt, psi, S=S, psi_b=pb)).max() > 1e-4  # not the upwind answer
        mesh = engine.Mesh(ctx, m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"])
        with pytest.raises(engine.capi.PbeError, match="geometry"):
            mesh.transport_implicit(T(psi), T(m["phi"]), dt, scheme=scheme)
        mesh.set_geometry(m["Sf"], m["C"], m["weights"])
        got, perf = mesh.transport_implicit(T(psi), T(m["phi"]), dt, S=T(S), psi_b=None if pb is None else T(pb),
                                            scheme=scheme, tolerance=1e-13)
        got = got.cpu().numpy()
        assert all(p[3] == 1 for p in perf)
        assert np.all(np.abs(got - want).sum(axis=1) <= 1e-10 * np.abs(want).sum(axis=1))
        for k in range(K):
            A, b = fvm.assemble(m, m["phi"], dt, psi[k], S=S[k], psi_b=None if pb is None else pb[k], scheme=scheme)
            assert perf[k][0] == pytest.approx(fvm.residual(A, b, psi[k]), rel=1e-9)
        gb, pfb = mesh.transport_implicit(T(psi), T(m["phi"]), dt, S=T(S), psi_b=None if pb is None else T(pb),
                                          scheme=scheme, tolerance=1e-13, solver="bicgstab")
        assert all(p[3] == 1 for p in pfb)
        assert np.all(np.abs(gb.cpu().numpy() - want).sum(axis=1) <= 1e-10 * np.abs(want).sum(axis=1))
        mesh.close()


@pytest.mark.gpu
@pytest.mark.parametrize("courant", [1.3, 8.0])
def test_device_implicit_limited_scheme_beyond_diagonal_dominance(ctx, courant):
    """Downwind weights put positive coefficients off the diagonal: beyond a cell Courant number of ~1 the fields'
    matrices are not diagonally dominant and neither iteration is guaranteed to converge (OpenFOAM's DILU-preconditioned
    solvers are sturdier there).  AUTO measures the dominance and switches to BiCGStab; whatever happens, a field is
    either solved to the tolerance or the call fails saying which field diverged -- never numbers without a warning."""
    from openpbe_b200 import engine
    rng = np.random.default_rng(16)
    shape, vel = (24, 20, 10), (1.0, -0.5, 0.25)
    m = synthetic.block_mesh(*shape, velocity=vel)
    x = m["C"][:, 0]
    psi = 1.0 + 0.5 * np.sin(2 * np.pi * x)[None, :] * np.array([[1.0], [0.5], [0.25]])  # smooth: the limiter stays near linear
    dt = courant / sum(abs(v) * n for v, n in zip(vel, shape))
    want = fvm.solve(m, m["phi"], dt, psi, scheme="Gauss vanLeer")
    mesh = engine.Mesh(ctx, m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"])
    mesh.set_geometry(m["Sf"], m["C"], m["weights"])
    for solver in ("auto", "jacobi", "bicgstab"):
        try:
            got, perf = mesh.transport_implicit(T(psi), T(m["phi"]), dt, scheme="Gauss vanLeer", tolerance=1e-12, solver=solver)
        except engine.capi.PbeError as e:
            assert "diverged" in str(e) and courant > 1.0
            continue
        for k, p in enumerate(perf):
            if p[3] == 1:
                assert np.abs(got[k].cpu().numpy() - want[k]).sum() <= 1e-8 * np.abs(want[k]).sum()
            else:
                assert p[1] <= p[0]  # stagnated short of the tolerance, not worse than it started
    mesh.close()


@pytest.mark.gpu
@pytest.mark.parametrize("relaxation", [1.0, 0.7])
def test_device_implicit_relaxation(ctx, relaxation):
    rng = np.random.default_rng(12)
    m = random_mesh(rng, 300, 1500, 40)  # random fluxes: rows are NOT dominant, relax() changes the matrix
    psi = fields(rng, 4, 300)
    dt = 4.0
    want = fvm.solve(m, m["phi"], dt, psi, relaxation=relaxation)
    plain = fvm.solve(m, m["phi"], dt, psi)
    assert np.abs(want - plain).max() > 1e-3
    got, _ = device_solve(ctx, m, psi, m["phi"], dt, relaxation=relaxation, tolerance=1e-13)
    assert np.all(np.abs(got - want) <= 1e-10 * np.abs(want).max())


@pytest.mark.gpu
def test_device_implicit_stops_by_openfoams_criterion(ctx):
    """tolerance 1e-5, relTol 0, minIter 1 (the reference's fvSolution): every field stops at the first iterate below
    the tolerance -- fields already below it still take one iteration -- and the answer is within what the tolerance
    allows of the exact one"""
    rng = np.random.default_rng(13)
    m = synthetic.block_mesh(20, 16, 12, velocity=(1.0, -0.5, 0.25))
    nc = 20 * 16 * 12
    psi = fields(rng, 6, nc)
    psi[5] = 1.0  # uniform field, zero-gradient patches, solenoidal flux: already the solution
    dt = 0.5 / 20
    got, perf = device_solve(ctx, m, psi, m["phi"], dt)
    want = fvm.solve(m, m["phi"], dt, psi)
    for k, (r0, r1, it, state) in enumerate(perf[:5]):
        assert state == 1 and r1 < 1e-5 and 1 <= it <= 20
        A, b = fvm.assemble(m, m["phi"], dt, psi[k])
        assert fvm.residual(A, b, got[k], psi[k]) == pytest.approx(r1, rel=1e-6, abs=1e-16)
    # the uniform field: OpenFOAM's normalisation is 0 / 0 there (the factor measures the deviation from the mean), so
    # the residual says nothing; the field stays what it is
    assert np.all(np.abs(got[5] - 1.0) <= 1e-13)
    assert np.all(np.abs(got - want) <= 1e-4 * np.abs(want).max())
    # min_iter 0 and a tolerance every field already meets: nothing is touched
    got0, perf0 = device_solve(ctx, m, psi[:5], m["phi"], dt, min_iter=0, tolerance=1.0)
    assert all(p[2] == 0 and p[3] == 1 for p in perf0) and np.array_equal(got0, psi[:5])
    # the two solvers stop at different iterates, both within the tolerance of the exact solution
    gj, pj = device_solve(ctx, m, psi[:5], m["phi"], dt, solver="jacobi")
    gb, pb = device_solve(ctx, m, psi[:5], m["phi"], dt, solver="bicgstab")
    assert all(p[3] == 1 for p in pj + pb)
    assert np.all(np.abs(gj - gb) <= 1e-4 * np.abs(want).max())
    # rel_tol: stop at 1e-2 of the initial residual
    _, perf_r = device_solve(ctx, m, psi, m["phi"], dt, tolerance=0.0, rel_tol=1e-2)
    for r0, r1, it, state in perf_r[:5]:
        assert state == 1 and r1 < 1e-2 * r0
    # max_iter
    _, perf_m = device_solve(ctx, m, psi, m["phi"], dt, tolerance=1e-300, max_iter=2)
    assert all(p[3] == 3 and p[2] == 2 for p in perf_m[:5])


@pytest.mark.gpu
def test_device_implicit_zero_flux_is_the_explicit_update(ctx):
    rng = np.random.default_rng(14)
    m = synthetic.block_mesh(6, 5, 4, velocity=(0.0, 0.0, 0.0))
    psi = fields(rng, 3, 120)
    S = rng.normal(size=psi.shape)
    got, perf = device_solve(ctx, m, psi, m["phi"], 0.01, S=S)
    assert np.allclose(got, psi + 0.01 * S, rtol=1e-14, atol=0)
    assert all(p[2] == 1 for p in perf)


@pytest.mark.gpu
def test_device_implicit_argument_errors(ctx):
    from openpbe_b200 import capi, engine
    m = synthetic.block_mesh(4, 3, 2)
    mesh = engine.Mesh(ctx, m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"])
    psi = T(np.ones((2, 24)))
    with pytest.raises(capi.PbeError, match="geometry"):
        mesh.transport_implicit(psi, T(m["phi"]), 0.1, scheme="vanLeer")
    with pytest.raises(capi.PbeError, match="alias"):
        mesh.transport_implicit(psi, T(m["phi"]), 0.1, out=psi)
    with pytest.raises(capi.PbeError, match="dt"):
        mesh.transport_implicit(psi, T(m["phi"]), 0.0)
    with pytest.raises(capi.PbeError, match="controls"):
        mesh.transport_implicit(psi, T(m["phi"]), 0.1, max_iter=0)
    with pytest.raises(capi.PbeError, match="solver"):
        mesh.transport_implicit(psi, T(m["phi"]), 0.1, solver=7)
    mesh.close()


@pytest.mark.gpu
def test_resident_moc_correct_with_the_implicit_solve(ctx, oracle):
    """MOC::correct() with usingMULES false (the reference's default) through the resident C-ABI call: the device's own
    sources, then the implicit solve of every class; two steps, compared with the direct solution."""
    from openpbe_b200 import engine
    tm = synthetic.block_mesh(14, 12, 10, velocity=(1.0, -0.5, 0.25))
    nc = 14 * 12 * 10
    w = synthetic.workload("moc30_ct", nc)
    f = synthetic.moc_fields(w)
    f["n"] *= 1.0 + 5.0 * np.sin(np.arange(nc) * 0.7)[None, :] ** 2
    dt = 0.03
    spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    moc = engine.MOC(ctx, w.N, w.xi1, spec, w.relax)
    mesh = engine.Mesh(ctx, tm["owner"], tm["neighbour"], tm["boundary_owner"], tm["vcell"])
    st = engine.MOCState(moc, nc, vcell=tm["vcell"], n0=f["n"], mesh=mesh, scheme="Gauss upwind")
    with pytest.raises(engine.capi.PbeError, match="no implicit step"):
        st.solver_performance()
    st.set_implicit(tolerance=1e-13)
    n = f["n"]
    for _ in range(2):
        S_dev = moc.source(T(n), T(f["eps"]), T(f["rho"]), T(f["nu"]), dt).cpu().numpy()
        want = fvm.solve(tm, tm["phi"], dt, n, S=S_dev)
        d, diag = st.correct(f["eps"], f["rho"], f["nu"], f["alpha"], dt, phi=tm["phi"], minmax=True)
        n_g = st.download().numpy()
        scale = np.abs(want).max(axis=1, keepdims=True)
        assert np.all(np.abs(n_g - want) <= 1e-10 * scale)
        assert np.allclose(d.numpy(), oracle.moc_d(n_g, w.xi1), rtol=1e-12)
        assert np.array_equal(diag["class_min"], n_g.min(axis=1))
        perf = st.solver_performance()
        assert len(perf) == w.N and all(p[3] == 1 and p[1] < 1e-13 for p in perf)
        n = n_g
    # back to the explicit update
    st.set_implicit(False)
    st.correct(f["eps"], f["rho"], f["nu"], f["alpha"], dt, phi=tm["phi"])
    S_dev = moc.source(T(n), T(f["eps"]), T(f["rho"]), T(f["nu"]), dt)
    want = mesh.transport_explicit(T(n), T(tm["phi"]), dt, S=S_dev).cpu().numpy()
    assert np.array_equal(st.download().numpy(), want)
    st.close(); mesh.close(); moc.close()


@pytest.mark.gpu
def test_resident_qmom_correct_with_the_implicit_solve(ctx, oracle):
    from openpbe_b200 import engine
    tm = synthetic.block_mesh(16, 9, 7, velocity=(0.5, 1.0, -0.25))
    nc = 16 * 9 * 7
    w = synthetic.workload("qmom3_const", nc)
    f = synthetic.qmom_fields(w, degenerate_fraction=0.0)
    dt = 0.02
    spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    q = engine.QMOM(ctx, w.N, spec)
    mesh = engine.Mesh(ctx, tm["owner"], tm["neighbour"], tm["boundary_owner"], tm["vcell"])
    st = engine.QMOMState(q, nc, vcell=tm["vcell"], m0=f["m"], mesh=mesh, scheme="Gauss upwind")
    st.set_implicit(tolerance=1e-13, min_iter=1)
    d, _ = st.correct(f["eps"], f["rho"], f["nu"], dt, phi=tm["phi"])
    m_g = st.download().numpy()
    S_dev = q.source(T(f["m"]), T(f["eps"]), T(f["rho"]), T(f["nu"])).cpu().numpy()
    S_dev = np.where(np.isnan(S_dev), 0.0, S_dev)
    want = np.maximum(fvm.solve(tm, tm["phi"], dt, f["m"], S=S_dev), 1e-15)  # QMOM.C:117-123
    assert np.all(np.abs(m_g - want) <= 1e-10 * np.abs(want).max(axis=1, keepdims=True))
    assert np.allclose(d.numpy(), np.cbrt(6.0 / np.pi * m_g[1] / m_g[0]), rtol=1e-13)
    assert all(p[3] == 1 for p in st.solver_performance())
    # the same step with a limited scheme in the implicit term
    mesh.set_geometry(tm["Sf"], tm["C"], tm["weights"])
    st2 = engine.QMOMState(q, nc, vcell=tm["vcell"], m0=f["m"], mesh=mesh, scheme="Gauss vanLeer")
    st2.set_implicit(tolerance=1e-13)
    st2.correct(f["eps"], f["rho"], f["nu"], dt, phi=tm["phi"])
    want = np.maximum(fvm.solve(tm, tm["phi"], dt, f["m"], S=S_dev, scheme="Gauss vanLeer"), 1e-15)
    assert np.all(np.abs(st2.download().numpy() - want) <= 1e-10 * np.abs(want).max(axis=1, keepdims=True))
    st2.close(); st.close(); mesh.close(); q.close()


# ---- decomposed meshes: the ranks solve ONE system (halo exchange of z per sweep, all-reduced sums) ---------------------
def _periodic_pair(nx, ny, nz, vel):
    """A block whose z- / z+ patches are processor patches towards ITSELF (send_cells = the cell across the periodic
    link), and the same domain as one global mesh with the periodic links as internal faces."""
    m = synthetic.block_mesh(nx, ny, nz, velocity=vel)
    ps = m["patch_start"]
    ni = m["owner"].size
    zlo, zhi = np.arange(ps[4], ps[5]), np.arange(ps[5], ps[6])
    bot, top = m["boundary_owner"][zlo], m["boundary_owner"][zhi]
    assert np.array_equal(bot + nx * ny * (nz - 1), top)
    local = dict(m)
    local["coupled"] = (int(ps[4]), int(ps[6] - ps[4]))
    local["send_cells"] = np.concatenate([top, bot]).astype(np.int32)  # what lies ACROSS each coupled face
    keep = np.arange(0, ps[4])
    glob = {"owner": np.concatenate([m["owner"], bot]), "neighbour": np.concatenate([m["neighbour"], top]),
            "boundary_owner": m["boundary_owner"][keep], "vcell": m["vcell"],
            "phi": np.concatenate([m["phi"][:ni], m["phi"][ni + zlo], m["phi"][ni + keep]])}
    return local, glob


def _one_rank_comm(c):
    import ctypes as C
    from openpbe_b200 import capi
    ident = C.create_string_buffer(128)
    capi.check(capi.lib().pbe_comm_unique_id(ident))
    capi.check(capi.lib().pbe_comm_init(c.handle, 1, 0, ident))


@pytest.mark.gpu
@pytest.mark.parametrize("psib,relaxation", [(False, 0.0), (True, 0.0), (False, 0.8)])
def test_device_implicit_decomposed_path_on_one_rank(psib, relaxation):
    """The whole decomposed machinery on ONE device: a one-rank NCCL communicator, a block that is its own neighbour
    across z (periodic), so that every sweep packs the processor-patch values, exchanges them (ncclSend / ncclRecv to
    itself), reads them as extra columns of the rows, and all-reduces the sums -- against the direct solve of the
    periodic domain as one mesh."""
    from openpbe_b200 import capi, engine
    rng = np.random.default_rng(20)
    shape, vel = (9, 7, 6), (0.4, -0.3, 1.0)
    local, glob = _periodic_pair(*shape, vel)
    nc = 9 * 7 * 6
    K = 4
    psi = fields(rng, K, nc)
    S = rng.normal(size=psi.shape) * 0.1
    dt = 1.5 / sum(abs(v) * n for v, n in zip(vel, shape))
    nb = local["boundary_owner"].size
    pb_local = rng.uniform(0.1, 2.0, size=(K, nb)) if psib else None
    pb_glob = None if pb_local is None else pb_local[:, :local["coupled"][0]]
    want = fvm.solve(glob, glob["phi"], dt, psi, S=S, psi_b=pb_glob, relaxation=relaxation)  # relax(): coupled-patch rule
    assert np.abs(want - fvm.solve(local, local["phi"], dt, psi, S=S, psi_b=pb_local, relaxation=relaxation)).max() > 1e-3
    c = engine.Context(0)
    _one_rank_comm(c)
    mesh = engine.Mesh(c, local["owner"], local["neighbour"], local["boundary_owner"], local["vcell"])
    mesh.set_coupled(*local["coupled"])
    with pytest.raises(capi.PbeError, match="halo plan"):
        mesh.transport_implicit(T(psi), T(local["phi"]), dt)
    mesh.set_halo_plan([(0, 0, local["coupled"][1])], local["send_cells"], nc)
    got, perf = mesh.transport_implicit(T(psi), T(local["phi"]), dt, S=T(S), psi_b=None if pb_local is None else T(pb_local),
                                        tolerance=1e-13, relaxation=relaxation)
    got = got.cpu().numpy()
    assert all(p[3] == 1 for p in perf)
    assert np.all(np.abs(got - want).sum(axis=1) <= 1e-10 * np.abs(want).sum(axis=1))
    for k in range(K):
        A, b = fvm.assemble(glob, glob["phi"], dt, psi[k], S=S[k], psi_b=None if pb_glob is None else pb_glob[k],
                            relaxation=relaxation)
        assert perf[k][0] == pytest.approx(fvm.residual(A, b, psi[k]), rel=1e-9)
    with pytest.raises(capi.PbeError, match="Jacobi iteration only"):
        mesh.transport_implicit(T(psi), T(local["phi"]), dt, solver="bicgstab")
    mesh.close()
    capi.check(capi.lib().pbe_comm_destroy(c.handle))
    c.close()


def _implicit_rank(rank, world, port, out):
    import os
    import ctypes as C
    import torch
    import torch.distributed as dist
    from openpbe_b200 import capi, engine, sharding
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)  # only to hand the NCCL id around and to gather
    ident = C.create_string_buffer(128)
    if rank == 0:
        capi.check(capi.lib().pbe_comm_unique_id(ident))
    box = [bytes(ident.raw)]
    dist.broadcast_object_list(box, src=0)
    ident = C.create_string_buffer(box[0], 128)
    c = engine.Context(rank)
    capi.check(capi.lib().pbe_comm_init(c.handle, world, rank, ident))
    rng = np.random.default_rng(21)
    g = synthetic.block_mesh(10, 8, 12, velocity=(0.4, -0.3, 1.0))
    nc = 10 * 8 * 12
    psi = fields(rng, 5, nc)
    S = rng.normal(size=psi.shape) * 0.1
    dt = 1.2 / (0.4 * 10 + 0.3 * 8 + 1.0 * 12)
    p = sharding.partition_mesh(g, rank, world)
    lo, hi = p["cell_range"]
    dev = torch.device(f"cuda:{rank}")
    D = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).to(dev)
    mesh = engine.Mesh(c, p["owner"], p["neighbour"], p["boundary_owner"], p["vcell"])
    mesh.set_coupled(*p["coupled"])
    mesh.set_halo_plan(p["neighbours"], p["send_cells"], nc)
    got, perf = mesh.transport_implicit(D(psi[:, lo:hi]), D(p["phi"]), dt, S=D(S[:, lo:hi]), tolerance=1e-13)
    parts = [None] * world
    dist.all_gather_object(parts, (lo, hi, got.cpu().numpy(), perf))
    if rank == 0:
        full = np.empty_like(psi)
        for a, b, v, _ in parts:
            full[:, a:b] = v
        np.save(out, full)
        assert all(pp == parts[0][3] for _, _, _, pp in parts), "the ranks disagree about residuals / iterations"
        assert all(q[3] == 1 for q in parts[0][3])
    dist.barrier()
    mesh.close()
    capi.check(capi.lib().pbe_comm_destroy(c.handle))
    c.close()
    dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.timeout(600)
def test_device_implicit_two_ranks_solve_one_system(tmp_path):
    """two GPUs, two sub-domains of one block (sharding.partition_mesh), NCCL through the C ABI: the gathered result
    is the direct solution of the undecomposed mesh"""
    import os
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2); the one-rank test above covers the same code on one")
    out = str(tmp_path / "imp2.npy")
    port = 32500 + (os.getpid() % 2000)
    mp.spawn(_implicit_rank, args=(2, port, out), nprocs=2, join=True)
    rng = np.random.default_rng(21)
    g = synthetic.block_mesh(10, 8, 12, velocity=(0.4, -0.3, 1.0))
    psi = fields(rng, 5, 960)
    S = rng.normal(size=psi.shape) * 0.1
    dt = 1.2 / (0.4 * 10 + 0.3 * 8 + 1.0 * 12)
    want = fvm.solve(g, g["phi"], dt, psi, S=S)
    got = np.load(out)
    assert np.all(np.abs(got - want).sum(axis=1) <= 1e-10 * np.abs(want).sum(axis=1))
