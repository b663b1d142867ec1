"""fixed-pivot method of classes on a geometric class grid (extension, PARITY UNPINNED:
the reference only has the uniform grid, MOC.C:56,79; Kumar & Ramkrishna 1996 is named at MOC.C:147-148).

What pins it instead: (1) on the uniform volumes the coalescence terms are the reference's own
(oracle.fixed_pivot == numpy_restatement.moc_source); (2) conservation properties of the technique;
(3) the device kernel == the CPU restatement to 1e-12 of the summed term magnitudes."""
import numpy as np
import pytest

from oracle import fixed_pivot as fp
from oracle import numpy_restatement as npo

gpu = pytest.mark.gpu

CT = {"coalescence": "CoulaloglouTavlaridesC", "coal_c": (0.00487, 8.0e7, 0.1, 0.047),
      "breakup": "CoulaloglouTavlarides", "br_c": (0.0118, 0.055, 0.1, 0.047)}
CONST = {"coalescence": "constantCoalescence", "coal_c": (0.7,), "breakup": "binaryBreakup"}
COAL_ONLY = {"coalescence": "constantCoalescence", "coal_c": (0.7,), "breakup": "noBreakup"}
BR_ONLY = {"coalescence": "noCoalescence", "breakup": "binaryBreakup"}


def fields(rng, N, nc, x):
    n = rng.uniform(0.0, 1.0, (N, nc)) * rng.choice([1.0, 1e-6, 0.0], (N, nc), p=[0.8, 0.1, 0.1]) / x[:, None]
    return n, 10.0 ** rng.uniform(-3, 0, nc), rng.uniform(800.0, 1200.0, nc), rng.uniform(1e-6, 1e-5, nc)


def test_uniform_pivots_reduce_to_the_reference_coalescence():
    rng = np.random.default_rng(3)
    N, nc, xi1 = 14, 9, 0.3
    x = xi1 * (np.arange(N) + 1)
    n, eps, rho, nu = fields(rng, N, nc, x)
    for spec in (dict(CT, breakup="noBreakup"), COAL_ONLY):
        got, _ = fp.moc_source(spec, x, 1.0, n, eps, rho, nu, 1e-3, limiter=False)
        want = npo.moc_source(spec, N, xi1, 1e30, n, eps, rho, nu, 1e-3)  # relax 1e30: limiter off
        scale = np.abs(want).max(axis=0) + 1e-300
        assert np.all(np.abs(got - want) <= 1e-12 * scale)
    # every pair lands on a pivot: eta == 1 (to rounding) and t == i + j + 1 (MOC.C:115-121), none is lost
    for xi1_ in (0.25, 0.3, 1e-11):
        pl = fp.pair_list(xi1_ * (np.arange(N) + 1))
        assert len(pl) == sum(1 for i in range(N) for j in range(i, N) if i + j + 1 <= N - 1)
        assert all(t == i + j + 1 and abs(eta - 1.0) < 1e-14 for i, j, t, eta in pl)


@pytest.mark.parametrize("ratio", [1.15, 1.5, 2.0, 2.7])
def test_conservation(ratio):
    rng = np.random.default_rng(int(ratio * 100))
    N, nc = 24, 7
    x = fp.geometric_pivots(N, 1e-3, ratio)
    n, eps, rho, nu = fields(rng, N, nc, x)
    # coalescence: volume exactly conserved, number lost at the pair rate
    S, scale = fp.moc_source(COAL_ONLY, x, 1.0, n, eps, rho, nu, 1e-3, limiter=False)
    assert np.all(np.abs(x @ S) <= 1e-13 * (x @ scale))
    rate = sum((0.5 if i == j else 1.0) * 0.7 * n[i] * n[j] for i, j, _, _ in fp.pair_list(x))
    assert np.allclose(S.sum(axis=0), -rate, rtol=1e-12, atol=0)
    # every birth lands between the pivots that bracket it
    for i, j, t, eta in fp.pair_list(x):
        v = x[i] + x[j]
        assert x[t] <= v * (1 + fp.SNAP) and (t == N - 1 or v < x[t + 1]) and 0.0 < eta <= 1.0
        if t + 1 < N:
            assert eta * x[t] + (1 - eta) * x[t + 1] == pytest.approx(v, rel=1e-14)
    # breakup: volume exactly conserved; each breakup of class k > 0 adds 1 - x_0 / x_k drops
    S, scale = fp.moc_source(BR_ONLY, x, 1.0, n, eps, rho, nu, 1e-3, limiter=False)
    assert np.all(np.abs(x @ S) <= 1e-13 * (x @ scale))
    g = x * x
    want = ((1.0 - x[0] / x)[:, None] * g[:, None] * n).sum(axis=0)
    assert np.allclose(S.sum(axis=0), want, rtol=1e-12, atol=0)
    # the limiter keeps n + dt S >= (1 - relax) n
    S, _ = fp.moc_source(CONST, x, 0.6, n * 1e3, eps, rho, nu, 10.0)
    assert np.all(n * 1e3 + 10.0 * S >= 0.4 * n * 1e3 * (1 - 1e-12))


def _dev(ctx, spec, x, relax=1.0):
    from openpbe_b200 import engine
    ks = engine.make_spec(spec["coalescence"], spec.get("coal_c", ()), spec["breakup"], spec.get("br_c", ()))
    return engine.MOC(ctx, x.size, 0.0, ks, relax, pivots=x)


@pytest.fixture(scope="module")
def ctx():
    from openpbe_b200 import engine
    return engine.Context(0)


@gpu
@pytest.mark.parametrize("spec", [CT, CONST, COAL_ONLY, BR_ONLY], ids=["ct", "const_binary", "coal", "breakup"])
@pytest.mark.parametrize("N,ratio", [(30, 1.5), (100, 1.12), (17, 2.0)])
def test_device_source_vs_restatement(ctx, spec, N, ratio):
    import torch
    rng = np.random.default_rng(N)
    nc = 300  # not a multiple of the 128-cell batch
    x = fp.geometric_pivots(N, 1e-11, ratio)
    n, eps, rho, nu = fields(rng, N, nc, x)
    n *= 1e-3 if spec is CT else 1.0
    dt = 2.5e-4
    want, scale = fp.moc_source(spec, x, 1.0, n, eps, rho, nu, dt)
    moc = _dev(ctx, spec, x)
    assert moc.kernel()["kernel"] == "pivot" and not moc.grid()[1] and np.array_equal(moc.grid()[0], x)
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    got = moc.source(T(n), T(eps), T(rho), T(nu), dt).cpu().numpy()
    assert np.all(np.abs(got - want) <= 1e-12 * scale + 1e-300), np.max(np.abs(got - want) / (scale + 1e-300))
    # fused step, d and moments
    nn = T(n.copy())
    d = torch.empty(nc, dtype=torch.float64, device="cuda")
    mom = torch.zeros(3, dtype=torch.float64, device="cuda")
    moc.step(nn, T(eps), T(rho), T(nu), dt, d=d, moments=mom)
    n1 = n + dt * want
    assert np.all(np.abs(nn.cpu().numpy() - n1) <= dt * 1e-12 * scale + 1e-15 * np.abs(n1))
    dv = (6.0 / np.pi * x) ** (1.0 / 3.0)
    np.testing.assert_allclose(d.cpu().numpy(), (dv @ n1) / n1.sum(axis=0), rtol=1e-12)
    np.testing.assert_allclose(mom.cpu().numpy(), [n1.sum(), (x @ n1).sum(), ((x * x) @ n1).sum()], rtol=1e-11)
    moc.close()


@gpu
def test_device_uniform_pivots_equal_uniform_kernel(ctx):
    import torch
    from openpbe_b200 import engine
    rng = np.random.default_rng(8)
    N, nc, xi1 = 40, 500, 1e-11
    x = xi1 * (np.arange(N) + 1)
    n, eps, rho, nu = fields(rng, N, nc, x)
    n *= 1e-3
    spec = dict(CT, breakup="noBreakup")  # the breakup quadratures differ by design (module docstring of oracle.fixed_pivot)
    ks = engine.make_spec(spec["coalescence"], spec["coal_c"], spec["breakup"], ())
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    a = engine.MOC(ctx, N, xi1, ks).source(T(n), T(eps), T(rho), T(nu), 1e30)  # huge dt: limiter off
    b = _dev(ctx, spec, x).source(T(n), T(eps), T(rho), T(nu), 1e30)
    _, scale = fp.moc_source(spec, x, 1.0, n, eps, rho, nu, 1e30)
    assert np.all(np.abs((a - b).cpu().numpy()) <= 1e-12 * scale + 1e-300)


@gpu
def test_device_errors(ctx):
    from openpbe_b200 import capi, engine
    ks = engine.make_spec("constantCoalescence", (1.0,), "noBreakup", ())
    with pytest.raises(capi.PbeError, match="must increase"):
        engine.MOC(ctx, 3, 0.0, ks, pivots=np.array([1.0, 3.0, 2.0]))
    with pytest.raises(capi.PbeError, match="positive"):
        engine.MOC(ctx, 2, 0.0, ks, pivots=np.array([0.0, 1.0]))
    moc = engine.MOC(ctx, 4, 0.0, ks, pivots=np.array([1.0, 2.0, 4.0, 8.0]))
    with pytest.raises(capi.PbeError, match="one kernel"):
        moc.set_kernel("smem")
    st = engine.MOCState(moc, 10)  # the zero-flux resident state takes any grid
    st.close()


@gpu
@pytest.mark.parametrize("scheme", ["Gauss upwind", "Gauss vanLeer"])
def test_resident_correct_with_transport_on_a_geometric_grid(ctx, scheme):
    """solveWithMULES on fixed pivots: the continuity check and the MULES bound maxNk = max(alpha V) / x_k read the
    handle's class volumes.  Given the device's own sources the transport is bit-identical to the C restatement."""
    import torch
    from openpbe_b200 import engine, synthetic
    from oracle import cpu_oracle as orc
    tm = synthetic.block_mesh(9, 8, 6, velocity=(1.0, -0.5, 0.25))
    nc, N, dt = 9 * 8 * 6, 16, 0.01
    rng = np.random.default_rng(5)
    x = fp.geometric_pivots(N, 1e-3, 1.35)
    n, eps, rho, nu = fields(rng, N, nc, x)
    n *= 1.0 + 3.0 * np.sin(np.arange(nc) * 0.7)[None, :] ** 2
    alpha = (x @ n) / tm["vcell"] * rng.uniform(0.9, 1.1, nc)
    moc = _dev(ctx, CONST, x)
    mesh = engine.Mesh(ctx, tm["owner"], tm["neighbour"], tm["boundary_owner"], tm["vcell"])
    mesh.set_geometry(tm["Sf"], tm["C"], tm["weights"])
    st = engine.MOCState(moc, nc, vcell=tm["vcell"], n0=n, mesh=mesh, scheme=scheme)
    d, diag = st.correct(eps, rho, nu, alpha, dt, phi=tm["phi"], minmax=True, n_moments=2)
    n_g = st.download().numpy()
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    S_dev = moc.source(T(n), T(eps), T(rho), T(nu), dt).cpu().numpy()
    pmax = np.array([np.max(alpha * tm["vcell"] / xk) for xk in x])
    want = orc.transport_mules(tm, scheme, n, tm["phi"], dt, pmax, S=S_dev)
    assert np.array_equal(n_g, want)
    want_S, scale = fp.moc_source(CONST, x, 1.0, n, eps, rho, nu, dt)
    assert np.all(np.abs(S_dev - want_S) <= 1e-12 * scale + 1e-300)
    assert diag["cont_err_max"] == pytest.approx(np.abs((x @ n_g) / tm["vcell"] - alpha).max(), rel=1e-9)
    np.testing.assert_allclose(diag["moments"], [(n_g.sum(axis=0) * tm["vcell"]).sum(), ((x @ n_g) * tm["vcell"]).sum()], rtol=1e-11)
    st.close(); mesh.close(); moc.close()


@gpu
def test_resident_correct_on_a_geometric_grid(ctx):
    """MOC::correct() through the resident state (class fields owned on the device) on fixed pivots: the same
    numbers as the device-pointer step, diagnostics included."""
    import torch
    rng = np.random.default_rng(21)
    N, nc, dt = 40, 3000, 2.5e-4
    x = fp.geometric_pivots(N, 1e-11, 1.2)
    n, eps, rho, nu = fields(rng, N, nc, x)
    n *= 1e-3
    vcell = rng.uniform(0.5, 2.0, nc)
    alpha = (x @ n) / vcell
    moc = _dev(ctx, CT, x)
    from openpbe_b200 import engine
    st = engine.MOCState(moc, nc, vcell=vcell, n0=n)
    d = torch.empty(nc, dtype=torch.float64).pin_memory()
    _, diag = st.correct(eps, rho, nu, alpha, dt, d_out=d, minmax=True, n_moments=3)
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    nn = T(n.copy())
    dd = torch.empty(nc, dtype=torch.float64, device="cuda")
    moc.step(nn, T(eps), T(rho), T(nu), dt, d=dd)
    assert torch.equal(st.download(), nn.cpu()) and torch.equal(d, dd.cpu())
    want, scale = fp.moc_source(CT, x, 1.0, n, eps, rho, nu, dt)
    n1 = n + dt * want
    assert np.all(np.abs(st.download().numpy() - n1) <= dt * 1e-12 * scale + 1e-15 * np.abs(n1))
    np.testing.assert_allclose(np.asarray(diag["class_min"]), n1.min(axis=1), rtol=1e-9, atol=1e-300)
    assert diag["cont_err_max"] == pytest.approx(np.abs((x @ n1) / vcell - alpha).max(), rel=1e-6, abs=1e-18)
    st.close()
