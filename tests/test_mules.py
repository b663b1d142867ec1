"""Row f2 (second slice): the MULES-limited transport of MOC::solveWithMULES (MOC.C:198-280) and
QMOM::solveWithMULES (QMOM.C:131-197) for the scheme the reference's cases select, `Gauss vanLeer`
(tests/validation/simmons_and_azzopardi/c01/system/fvSchemes:34).

PARITY UNPINNED: OpenFOAM's MULES / limitedScheme / gaussGrad are third-party code absent from the
reference tree and from this image; oracle/pbe_oracle_mules.c restates OpenFOAM 3.0.x.  What is
checked:
  * CPU: the restatement does what the algorithm must do (upwind scheme = no correction, boundedness
    with the limiter, conservation, second-order accuracy on a smooth profile, limitSum makes the
    corrections of all fields sum to zero on every face);
  * CPU: the reference's OWN MOC.C / QMOM.C (compiled unmodified, oracle/_ref), driving the shimmed
    OpenFOAM calls, gives bit for bit what the port of its call sequence (orc_transport_mules) gives;
  * GPU: pbe_transport_mules and the resident correct() are bit-identical to the port.
"""
import numpy as np
import pytest

from oracle import cpu_oracle as orc
from oracle import ref_oracle as ref
from openpbe_b200 import synthetic

CT = (0.0152, 0.0678, 0.05, 0.04282)


def bump_fields(m, K, rng, lo=0.0, hi=1.0):
    """smooth bumps + a step: gradients of both signs, local extrema, flat regions"""
    C = m["C"]
    cmin, cmax = C.min(axis=0), C.max(axis=0)
    ext = np.where(cmax > cmin, cmax - cmin, 1.0)
    X = (C - cmin) / ext  # unit box in every direction that has more than one cell layer
    psi = np.empty((K, C.shape[0]))
    for k in range(K):
        c0 = rng.uniform(0.2, 0.8, 3) * (cmax > cmin)
        r2 = ((X - c0) ** 2).sum(axis=1)
        psi[k] = lo + (hi - lo) * (np.exp(-r2 / rng.uniform(0.01, 0.05)) + (X[:, 0] > rng.uniform(0.3, 0.7)) * rng.uniform(0.2, 0.6))
    return psi


def random_geometry_mesh(rng, ncells, nint, nbnd):
    """random face addressing with fabricated geometry: parity needs numbers, not a valid mesh"""
    a = rng.integers(0, ncells, size=nint)
    b = (a + 1 + rng.integers(0, ncells - 1, size=nint)) % ncells
    own, nei = np.minimum(a, b), np.maximum(a, b)
    order = np.lexsort((nei, own))
    return {"owner": own[order].astype(np.int32), "neighbour": nei[order].astype(np.int32),
            "boundary_owner": rng.integers(0, ncells, size=nbnd).astype(np.int32),
            "vcell": rng.uniform(0.5, 2.0, ncells), "phi": rng.normal(size=nint + nbnd) * 0.05,
            "Sf": rng.normal(size=(nint + nbnd, 3)), "C": rng.uniform(size=(ncells, 3)),
            "weights": rng.uniform(0.3, 0.7, nint)}


# ---- CPU: properties of the restatement -------------------------------------------------------------------
def test_upwind_scheme_through_mules_is_the_plain_upwind_update():
    rng = np.random.default_rng(1)
    m = synthetic.block_mesh(7, 6, 5, velocity=(0.4, -0.3, 0.2))
    psi = bump_fields(m, 4, rng)
    S = rng.normal(size=psi.shape)
    a = orc.transport_mules(m, "Gauss upwind", psi, m["phi"], 0.05, np.full(4, 10.0), S=S)
    b = orc.transport_upwind_explicit(m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"], psi, m["phi"], 0.05, S=S)
    assert np.array_equal(a, b)


def test_vanleer_mules_is_bounded_and_conservative():
    rng = np.random.default_rng(2)
    n = 24
    m = synthetic.block_mesh(n, n, 1, lengths=(1.0, 1.0, 1.0 / n), velocity=(1.0, 0.5, 0.0))
    psi = bump_fields(m, 3, rng)
    ni = m["owner"].size
    phi = m["phi"].copy()
    phi[ni:] = 0.0  # closed box: nothing crosses the boundary
    # a discretely divergence-free vortex tangent to the walls: face fluxes are differences of the stream
    # function sin(pi x) sin(pi y) between the face's end points (MULES bounds presuppose div(phi) = 0)
    C = m["C"]
    own, nei = m["owner"], m["neighbour"]
    fc = 0.5 * (C[own] + C[nei])
    h = 1.0 / n
    sf = lambda x, y: np.sin(np.pi * x) * np.sin(np.pi * y)  # noqa: E731
    xn = np.abs(m["Sf"][:ni, 0]) > 0
    phi[:ni] = np.where(xn, (sf(fc[:, 0], fc[:, 1] + h / 2) - sf(fc[:, 0], fc[:, 1] - h / 2)) * h,
                        -(sf(fc[:, 0] + h / 2, fc[:, 1]) - sf(fc[:, 0] - h / 2, fc[:, 1])) * h)
    dt = 0.3 / (np.pi * n)  # CFL ~ 0.3 (|u| <= pi)
    tot0 = (psi * m["vcell"]).sum(axis=1)
    cur = psi.copy()
    for _ in range(20):
        cur = orc.transport_mules(m, "Gauss vanLeer", cur, phi, dt, np.full(3, psi.max()), psi_min=0.0)
        assert cur.min() >= -1e-14 and cur.max() <= psi.max() * (1 + 1e-12)   # MULES bounds [psiMin, psiMax]
    assert np.allclose((cur * m["vcell"]).sum(axis=1), tot0, rtol=1e-12)       # fluxes telescope
    up = psi.copy()
    for _ in range(20):
        up = orc.transport_upwind_explicit(m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"], up, phi, dt)
    # the limited scheme is less diffusive than upwind: it keeps more of the fields' second moment
    assert np.all((cur ** 2 * m["vcell"]).sum(axis=1) > 1.01 * (up ** 2 * m["vcell"]).sum(axis=1))


def test_vanleer_is_second_order_on_a_smooth_profile():
    errs = {}
    for scheme in ("Gauss upwind", "Gauss vanLeer"):
        e = []
        for n in (32, 64, 128):
            m = synthetic.block_mesh(n, 1, 1, lengths=(1.0, 1.0 / n, 1.0 / n), velocity=(1.0, 0.0, 0.0))
            x = m["C"][:, 0]
            psi = (1.0 + 0.5 * np.sin(2 * np.pi * x))[None, :]
            nb = m["boundary_owner"].size
            f = fvc = orc.div_flux(m, scheme, m["phi"], psi[0], psi_b=np.concatenate([[psi[0, -1]], [psi[0, 0]], np.zeros(nb - 2)]))
            # face flux error against the exact face value on interior faces
            ni = m["owner"].size
            xf = 0.5 * (x[m["owner"]] + x[m["neighbour"]])
            exact = 1.0 + 0.5 * np.sin(2 * np.pi * xf)
            e.append(np.abs(f[:ni] / m["phi"][:ni] - exact)[2:-2].mean())  # error of the face VALUE
        errs[scheme] = e
    up, vl = errs["Gauss upwind"], errs["Gauss vanLeer"]
    assert 1.7 < up[0] / up[1] < 2.3 and 1.7 < up[1] / up[2] < 2.3        # first order: error / 2 per refinement
    assert vl[0] / vl[1] > 3.0 and vl[1] / vl[2] > 3.0                    # ~second order away from the extrema
    assert vl[2] < 0.1 * up[2]


def test_limit_sum_cancels_the_corrections_of_all_fields():
    rng = np.random.default_rng(3)
    corr = rng.normal(size=(5, 200))
    corr[:, :10] = np.abs(corr[:, :10])  # all positive on some faces: nothing balances them, all removed
    out = corr.copy()
    fn = orc.lib().orc_mules_limit_sum
    import ctypes as C
    fn.argtypes = [C.c_int, C.c_int64, C.c_int64, C.POINTER(C.c_double)]
    fn(5, 200, 200, out.ctypes.data_as(C.POINTER(C.c_double)))
    assert np.all(out[:, :10] == 0.0)
    assert np.abs(out.sum(axis=0)).max() < 1e-14
    assert np.all(np.abs(out) <= np.abs(corr) + 1e-300) and np.all(out[:, 10:] * corr[:, 10:] >= 0.0)


# ---- CPU: the reference's own call sequence == the port of it -----------------------------------------------
needs_ref = pytest.mark.skipif(not ref.build(), reason="oracle/_ref/libpbe_ref.so not built (needs /root/reference)")


@needs_ref
@pytest.mark.parametrize("scheme", ["Gauss upwind", "Gauss vanLeer"])
def test_reference_moc_solve_with_mules_equals_port(oracle, scheme):
    tm = synthetic.block_mesh(9, 7, 5, velocity=(1.0, -0.5, 0.25))
    nc = 9 * 7 * 5
    w = synthetic.workload("moc30_ct", nc)
    w.N, w.xi1 = 6, 3e-10 / 6
    f = synthetic.moc_fields(w)
    f["n"] *= 1.0 + 5.0 * np.sin(np.arange(nc) * 0.7)[None, :] ** 2
    dt = 0.02
    try:
        ref.set_mesh(nc, tm["owner"], tm["neighbour"], tm["boundary_owner"], Sf=tm["Sf"], C_=tm["C"], weights=tm["weights"],
                     div_scheme=scheme)
        txt = ref.pbe_dict("methodOfClasses", w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs, N=w.N, xi1=w.xi1,
                           relax=w.relax)
        n_ref, S_ref, d_ref, _ = ref.method_steps(txt, "n", f["n"], f["eps"], f["rho"], f["nu"], dt, alpha=f["alpha"],
                                                  vcell=tm["vcell"], phi=tm["phi"], nsteps=2)
    finally:
        ref.clear_mesh()
    sp = oracle.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    xi = w.xi1 * np.arange(1, w.N + 1)
    pmax = np.array([np.max(f["alpha"] * tm["vcell"] / x) for x in xi])  # max(phase_ * mesh.V() / xi_k), MOC.C:238
    n = f["n"]
    for _ in range(2):
        S, _ = oracle.moc_source(sp, w.N, w.xi1, w.relax, n, f["eps"], f["rho"], f["nu"], dt)
        n = oracle.transport_mules(tm, scheme, n, tm["phi"], dt, pmax, S=S)
    assert np.array_equal(S_ref, S) and np.array_equal(n_ref, n)
    assert np.allclose(d_ref, oracle.moc_d(n, w.xi1), rtol=1e-14)


@needs_ref
def test_reference_qmom_solve_with_mules_equals_port(oracle):
    tm = synthetic.block_mesh(8, 6, 4, velocity=(0.5, 1.0, -0.25))
    nc = 8 * 6 * 4
    w = synthetic.workload("qmom3_const", nc)
    f = synthetic.qmom_fields(w, degenerate_fraction=0.0)
    dt = 0.01
    try:
        ref.set_mesh(nc, tm["owner"], tm["neighbour"], tm["boundary_owner"], Sf=tm["Sf"], C_=tm["C"], weights=tm["weights"],
                     div_scheme="Gauss vanLeer")
        txt = ref.pbe_dict("quadratureMethodOfMoments", w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs, order=w.N)
        m_ref, S_ref, _, _ = ref.method_steps(txt, "m", f["m"], f["eps"], f["rho"], f["nu"], dt, vcell=tm["vcell"], phi=tm["phi"])
    finally:
        ref.clear_mesh()
    # the port's sequence on the reference's own (scrubbed) sources: 1.2 max(m_k) bound (QMOM.C:171), limit,
    # limitSum, explicitSolve, clamp to SMALL (QMOM.C:117-123)
    pmax = 1.2 * f["m"].max(axis=1)
    m_port = np.maximum(oracle.transport_mules(tm, "Gauss vanLeer", f["m"], tm["phi"], dt, pmax, S=S_ref), 1e-15)
    assert np.array_equal(m_ref, m_port)


# ---- GPU: the device path against the port, bit for bit -----------------------------------------------------
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


def device_mules(ctx, m, scheme, psi, phi, dt, pmax, S=None, psi_b=None, n_iter=3):
    from openpbe_b200 import engine
    mesh = engine.Mesh(ctx, m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"])
    mesh.set_geometry(m["Sf"], m["C"], m["weights"])
    out = mesh.transport_mules(T(psi), T(phi), dt, T(pmax), scheme=scheme, S=None if S is None else T(S),
                               psi_b=None if psi_b is None else T(psi_b), n_limiter_iter=n_iter)
    out = out.cpu().numpy()
    mesh.close()
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(12, 9, 7), (40, 1, 1), (17, 13, 1), (64, 48, 20)])
@pytest.mark.parametrize("scheme", ["Gauss vanLeer", "Gauss upwind"])
def test_device_mules_matches_port_block(ctx, shape, scheme, monkeypatch):
    rng = np.random.default_rng(sum(shape))
    m = synthetic.block_mesh(*shape, velocity=(1.0, -0.5, 0.25))
    K = 7
    psi = bump_fields(m, K, rng, 0.0, 3.0)
    S = rng.normal(size=psi.shape) * 0.3
    pmax = psi.max(axis=1) * rng.uniform(0.7, 1.3, K)  # some bounds bind
    dt = 0.3 / max(shape)
    want = orc.transport_mules(m, scheme, psi, m["phi"], dt, pmax, S=S)
    monkeypatch.setenv("PBE_MULES_CHUNK", "3")  # several field chunks
    got = device_mules(ctx, m, scheme, psi, m["phi"], dt, pmax, S=S)
    assert np.array_equal(got, want)
    nb = m["boundary_owner"].size
    psi_b = rng.uniform(0.0, 3.0, (K, nb))
    want = orc.transport_mules(m, scheme, psi, m["phi"], dt, pmax, psi_b=psi_b, n_limiter_iter=2)
    got = device_mules(ctx, m, scheme, psi, m["phi"], dt, pmax, psi_b=psi_b, n_iter=2)
    assert np.array_equal(got, want)


@pytest.mark.gpu
@pytest.mark.parametrize("nint,nbnd", [(0, 4), (60, 5), (900, 120), (6000, 0)])
def test_device_mules_matches_port_random_topology(ctx, nint, nbnd):
    rng = np.random.default_rng(nint + nbnd)
    m = random_geometry_mesh(rng, 257, nint, nbnd)
    K = 5
    psi = rng.uniform(0.0, 2.0, (K, 257))
    S = rng.normal(size=psi.shape)
    pmax = np.full(K, 1.8)
    want = orc.transport_mules(m, "Gauss vanLeer", psi, m["phi"], 0.1, pmax, S=S)
    got = device_mules(ctx, m, "Gauss vanLeer", psi, m["phi"], 0.1, pmax, S=S)
    assert np.array_equal(got, want)


@pytest.mark.gpu
@pytest.mark.parametrize("n_iter", [1, 3, 5])
def test_device_mules_fused_path_equals_per_field_path(ctx, n_iter, monkeypatch):
    """The fused kernels (field chunks per thread, lambda rebuilt from the lm / lp rows of the earlier sweeps,
    limitSum applied in the solve) against the per-field kernels they replaced and against the port."""
    rng = np.random.default_rng(n_iter)
    m = synthetic.block_mesh(21, 14, 9, velocity=(0.8, -0.6, 0.3))
    K = 11
    psi = bump_fields(m, K, rng, 0.0, 3.0)
    S = rng.normal(size=psi.shape) * 0.3
    pmax = psi.max(axis=1) * rng.uniform(0.6, 1.2, K)
    psi_b = rng.uniform(0.0, 3.0, (K, m["boundary_owner"].size))
    want = orc.transport_mules(m, "Gauss vanLeer", psi, m["phi"], 0.02, pmax, S=S, psi_b=psi_b, n_limiter_iter=n_iter)
    monkeypatch.setenv("PBE_MULES_CHUNK", "4")
    fused = device_mules(ctx, m, "Gauss vanLeer", psi, m["phi"], 0.02, pmax, S=S, psi_b=psi_b, n_iter=n_iter)
    monkeypatch.setenv("PBE_MULES_PATH", "0")
    plain = device_mules(ctx, m, "Gauss vanLeer", psi, m["phi"], 0.02, pmax, S=S, psi_b=psi_b, n_iter=n_iter)
    assert np.array_equal(fused, want) and np.array_equal(plain, want)
    # polyhedral cells with more faces than the fused kernels keep in registers
    m = random_geometry_mesh(rng, 40, 400, 30)
    psi = rng.uniform(0.0, 2.0, (K, 40))
    pmax = np.full(K, 1.7)
    want = orc.transport_mules(m, "Gauss vanLeer", psi, m["phi"], 0.1, pmax, n_limiter_iter=n_iter)
    monkeypatch.delenv("PBE_MULES_PATH")
    assert np.array_equal(device_mules(ctx, m, "Gauss vanLeer", psi, m["phi"], 0.1, pmax, n_iter=n_iter), want)


@pytest.mark.gpu
def test_device_psi_max_helpers(ctx):
    from openpbe_b200 import engine
    rng = np.random.default_rng(9)
    nc, N = 5000, 13
    alpha, V = rng.uniform(0.01, 0.1, nc), rng.uniform(1e-10, 1e-8, nc)
    moc = engine.MOC(ctx, N, 3e-12, engine.make_spec("noCoalescence", (), "noBreakup", ()))
    got = moc.psi_max(T(alpha), T(V)).cpu().numpy()
    xi = 3e-12 * np.arange(1, N + 1)
    assert np.array_equal(got, np.array([np.max(alpha * V / x) for x in xi]))  # MOC.C:238
    moc.close()


@pytest.mark.gpu
def test_resident_moc_correct_with_vanleer_matches_port(ctx, oracle):
    """MOC::correct() with solveWithMULES and Gauss vanLeer through the resident C-ABI call: sources to
    1e-12 sum|terms|; given the same sources the transport is bit-identical, so compare after feeding the
    device's own sources to the port."""
    from openpbe_b200 import engine
    tm = synthetic.block_mesh(14, 12, 10, velocity=(1.0, -0.5, 0.25))
    nc = 14 * 12 * 10
    w = synthetic.workload("moc30_ct", nc)
    f = synthetic.moc_fields(w)
    f["n"] *= 1.0 + 5.0 * np.sin(np.arange(nc) * 0.7)[None, :] ** 2
    dt = 0.01
    spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    moc = engine.MOC(ctx, w.N, w.xi1, spec, w.relax)
    mesh = engine.Mesh(ctx, tm["owner"], tm["neighbour"], tm["boundary_owner"], tm["vcell"])
    mesh.set_geometry(tm["Sf"], tm["C"], tm["weights"])
    st = engine.MOCState(moc, nc, vcell=tm["vcell"], n0=f["n"], mesh=mesh, scheme="Gauss vanLeer")
    d, diag = st.correct(f["eps"], f["rho"], f["nu"], f["alpha"], dt, phi=tm["phi"], minmax=True)
    n_g = st.download().numpy()
    S_dev = moc.source(T(f["n"]), T(f["eps"]), T(f["rho"]), T(f["nu"]), dt).cpu().numpy()
    xi = w.xi1 * np.arange(1, w.N + 1)
    pmax = np.array([np.max(f["alpha"] * tm["vcell"] / x) for x in xi])
    want = oracle.transport_mules(tm, "Gauss vanLeer", f["n"], tm["phi"], dt, pmax, S=S_dev)
    assert np.array_equal(n_g, want)
    sp = oracle.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    So, scale = oracle.moc_source(sp, w.N, w.xi1, w.relax, f["n"], f["eps"], f["rho"], f["nu"], dt)
    assert np.all(np.abs(S_dev - So) <= 1e-12 * scale + 1e-300)
    assert np.allclose(d.numpy(), oracle.moc_d(n_g, w.xi1), rtol=1e-12)
    assert np.array_equal(diag["class_min"], n_g.min(axis=1))
    # without alpha the limiter's bound cannot be formed
    with pytest.raises(engine.capi.PbeError, match="needs alpha"):
        st.correct(f["eps"], f["rho"], f["nu"], None, dt, phi=tm["phi"])
    st.close(); mesh.close(); moc.close()


@pytest.mark.gpu
def test_resident_qmom_correct_with_vanleer_matches_port(ctx, oracle):
    from openpbe_b200 import engine
    tm = synthetic.block_mesh(16, 9, 7, velocity=(0.5, 1.0, -0.25))
    nc = 16 * 9 * 7
    w = synthetic.workload("qmom3_const", nc)
    f = synthetic.qmom_fields(w, degenerate_fraction=0.0)
    dt = 0.01
    spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    q = engine.QMOM(ctx, w.N, spec)
    mesh = engine.Mesh(ctx, tm["owner"], tm["neighbour"], tm["boundary_owner"], tm["vcell"])
    mesh.set_geometry(tm["Sf"], tm["C"], tm["weights"])
    st = engine.QMOMState(q, nc, vcell=tm["vcell"], m0=f["m"], mesh=mesh, scheme="Gauss vanLeer")
    d, _ = st.correct(f["eps"], f["rho"], f["nu"], dt, phi=tm["phi"])
    m_g = st.download().numpy()
    S_dev = q.source(T(f["m"]), T(f["eps"]), T(f["rho"]), T(f["nu"])).cpu().numpy()
    S_dev = np.where(np.isnan(S_dev), 0.0, S_dev)
    want = np.maximum(oracle.transport_mules(tm, "Gauss vanLeer", f["m"], tm["phi"], dt, 1.2 * f["m"].max(axis=1), S=S_dev), 1e-15)
    assert np.array_equal(m_g, want)
    st.close(); mesh.close(); q.close()


@pytest.mark.gpu
def test_device_mules_argument_errors(ctx):
    from openpbe_b200 import engine
    m = synthetic.block_mesh(4, 3, 2)
    mesh = engine.Mesh(ctx, m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"])
    psi = T(np.ones((2, 24)))
    with pytest.raises(engine.capi.PbeError, match="needs the mesh geometry"):
        mesh.transport_mules(psi, T(m["phi"]), 0.1, T(np.ones(2)))
    mesh.set_geometry(m["Sf"], m["C"], m["weights"])
    with pytest.raises(engine.capi.PbeError, match="must not alias"):
        mesh.transport_mules(psi, T(m["phi"]), 0.1, T(np.ones(2)), out=psi)
    with pytest.raises(engine.capi.PbeError, match="Unknown or not yet supported divergence scheme"):
        engine.div_scheme("Gauss QUICK")
    mesh.set_coupled(0, 2)
    with pytest.raises(engine.capi.PbeError, match="processor patches"):
        mesh.transport_mules(psi, T(m["phi"]), 0.1, T(np.ones(2)))
    mesh.close()
