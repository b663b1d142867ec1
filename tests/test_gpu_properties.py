"""GPU tests that need no oracle run at full size: the reference's 0-D analytical acceptance
tests executed on the device, and size-independent properties at BASELINE.json's full sizes.

 * O1 / O2 (tests/integration/MOC/testConvergence.py:15-23, 34-47, 54-60): pure coalescence vs
   Scott, pure breakup vs Ziff-McGrady, N in {10,20,40,80}; L2 error monotone in N and ~0 at one
   decimal -- the reference's own acceptance criterion, with pbe_moc_step as the time stepper.
 * coalescence conserves volume exactly on the uniform grid (xi_i + xi_j = xi_{i+j+1}, death
   truncated at the grid edge, MOC.C:107-121): sum_i xi_i C_i = 0 up to rounding;
 * scaling n by 2 scales coalescence by 4 and breakup by 2 EXACTLY (powers of two commute with
   every rounding): a bitwise check that needs no reference values;
 * Wheeler round trip: the moments of the inverted quadrature reproduce the input moments.
"""
import numpy as np
import pytest
from scipy.special import gamma as sp_gamma

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from openpbe_b200 import engine, synthetic  # noqa: E402


@pytest.fixture(scope="module")
def ctx():
    assert torch.cuda.is_available()
    c = engine.Context(0)
    yield c
    c.close()


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).cuda()


def scott(xi, t, C, N0, xi0):
    Tt = C * N0 * t
    x = xi / xi0
    phi3 = sum((x * 2) ** (2 * (k + 1)) / sp_gamma(2 * (k + 1)) * (Tt / (Tt + 2)) ** k
               for k in range(100)) * 4.0 * np.exp(-2 * x) / (x * (Tt + 2) ** 2)
    return N0 / xi0 * phi3


def ziff_mcgrady(x, t, l):
    return np.piecewise(x, [x < l, x == l, x > l],
                        [lambda x: 2.0 * t * l * np.exp(-t * x ** 2), lambda x: np.exp(-t * x ** 2),
                         lambda x: 0.0])


def l2rel(x, f, g):
    return np.sqrt(np.trapezoid((f - g) ** 2, x=x)) / np.sqrt(np.trapezoid(f ** 2, x=x))


def run_0d_device(ctx, spec, N, dv, n0, dt, t_end, ncells=3):
    moc = engine.MOC(ctx, N, dv, spec)
    n = T(np.repeat(n0[:, None], ncells, axis=1))
    one = torch.ones(ncells, dtype=torch.float64, device="cuda")
    for _ in range(int(round(t_end / dt))):
        moc.step(n, one, one, one, dt)
    out = n.cpu().numpy()
    assert np.array_equal(out[:, 0], out[:, -1])   # cell-independent kernels: identical columns
    return out[:, 0]


def test_o1_pure_coalescence_vs_scott_on_device(ctx):
    N0, v0, Cc, vmax = 2.0, 0.5, 0.1, 10.0
    spec = engine.make_spec("constantCoalescence", (Cc,), "noBreakup")
    errs = []
    for N in (10, 20, 40, 80):
        dv = vmax / N
        v = dv + dv * np.arange(N)
        n0 = (N0 / v0) * (v / v0) * np.exp(-v / v0) * dv
        nT = run_0d_device(ctx, spec, N, dv, n0, 1e-3, 1.0)
        xi_n = np.linspace(dv, N * dv, N)
        errs.append(l2rel(xi_n, nT / dv, scott(xi_n, 1.0, Cc, N0, 2.0 * v0)))
        if N == 80:  # the coarse grids do not even hold the initial number; the finest one does
            assert np.sum(nT) == pytest.approx(2 * N0 / (Cc * N0 * 1.0 + 2), rel=0.05)  # total number, :10-12
    assert all(errs[k] < errs[k - 1] for k in range(1, 4))
    np.testing.assert_almost_equal(errs[-1], 0.0, decimal=1)


def test_o2_pure_breakup_vs_ziff_mcgrady_on_device(ctx):
    spec = engine.make_spec("noCoalescence", (), "binaryBreakup")
    errs = []
    for N in (10, 20, 40, 80):
        dv = 1.0 / N
        v = dv + dv * np.arange(N)
        n0 = np.where(v < 1.0, 0.0, 1.0)
        nT = run_0d_device(ctx, spec, N, dv, n0, 5e-3, 10.0)
        xi_n = np.linspace(dv, N * dv, N)
        errs.append(l2rel(xi_n, nT / dv, ziff_mcgrady(xi_n, 10.0, 1.0)))
    assert all(errs[k] < errs[k - 1] for k in range(1, 4))
    np.testing.assert_almost_equal(errs[-1], 0.0, decimal=1)


def test_qmom_0d_constant_coalescence_moments(ctx):
    """tests/integration/QMOM: constant kernel, dm0/dt = -C m0^2/2, m1 conserved."""
    Cc, dt, steps = 0.1, 1e-3, 1000
    q = engine.QMOM(ctx, 2, engine.make_spec("constantCoalescence", (Cc,), "noBreakup"))
    N0, v0 = 2.0, 0.5
    from math import factorial
    m0 = np.array([N0 * factorial(k + 1) * (v0 / 2) ** k for k in range(4)])  # setupCases.py:44-48
    m = T(np.repeat(m0[:, None], 5, axis=1))
    one = torch.ones(5, dtype=torch.float64, device="cuda")
    for _ in range(steps):
        q.step(m, one, one, one, dt)
    out = m.cpu().numpy()[:, 0]
    assert out[0] == pytest.approx(2 * N0 / (Cc * N0 * 1.0 + 2), rel=2e-3)   # 2 N0 / (T + 2)
    assert out[1] == pytest.approx(m0[1], rel=1e-10)                           # volume conserved


# ---- full-size, oracle-free properties (BASELINE configs[2]: 100 classes, 1.6e7 cells) ---------
FULL_MOC = 16_000_000
FULL_QMOM = 10_000_000


def _device_moc_fields(w, ncells, seed=1):
    g = torch.Generator(device="cuda").manual_seed(seed)
    xi = w.xi1 * torch.arange(1, w.N + 1, dtype=torch.float64, device="cuda")
    p = torch.from_numpy(synthetic._lognormal_volume_pdf(xi.cpu().numpy()) + 1e-300).cuda()
    alpha = 0.01 + 0.09 * torch.rand(ncells, generator=g, dtype=torch.float64, device="cuda")
    n = (alpha * 1e-9)[None, :] * (p / torch.sum(p * xi))[:, None]
    n *= torch.exp2(2 * torch.rand((w.N, 1), generator=g, dtype=torch.float64, device="cuda") - 1)
    eps = 10 ** (-3 * torch.rand(ncells, generator=g, dtype=torch.float64, device="cuda"))
    rho = torch.full((ncells,), w.rho, dtype=torch.float64, device="cuda")
    nu = torch.full((ncells,), w.nu, dtype=torch.float64, device="cuda")
    return n.contiguous(), eps, rho, nu, xi


def test_full_size_volume_conservation_and_scaling(ctx):
    w = synthetic.workload("moc100_ct")
    n, eps, rho, nu, xi = _device_moc_fields(w, FULL_MOC)
    coal_only = engine.MOC(ctx, w.N, w.xi1, engine.make_spec(w.coalescence, w.coal_coeffs), 1e300)
    S = coal_only.source(n, eps, rho, nu, w.dt)
    # volume conservation of the coalescence operator, relative to the gross volume exchanged
    net = torch.einsum("i,ic->c", xi, S)
    gross = torch.einsum("i,ic->c", xi, S.abs())
    assert float((net.abs() / (gross + 1e-300)).max()) < 1e-11
    assert bool(torch.isfinite(S).all())
    # births only in the top class, deaths only in class 0 (MOC.C:107-121 truncation quirks)
    assert bool((S[-1] >= 0).all()) and bool((S[0] <= 0).all())
    chk = torch.einsum("ic->i", S)            # checksum of checksums, reused below
    # exact quadratic scaling: C(2n) == 4 C(n) bit for bit
    S2 = coal_only.source(n * 2.0, eps, rho, nu, w.dt)
    assert bool((S2 == 4.0 * S).all())
    del S2
    # breakup is linear in n: B(2n) == 2 B(n) bit for bit
    br_only = engine.MOC(ctx, w.N, w.xi1, engine.make_spec("noCoalescence", (), w.breakup, w.br_coeffs), 1e300)
    B = br_only.source(n, eps, rho, nu, w.dt)
    B2 = br_only.source(n * 2.0, eps, rho, nu, w.dt)
    assert bool((B2 == 2.0 * B).all())
    del B2
    # the full operator is the sum of its parts when the limiter is off
    full = engine.MOC(ctx, w.N, w.xi1, engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs), 1e300)
    F = full.source(n, eps, rho, nu, w.dt)
    tol = 1e-13 * (S.abs() + B.abs())
    assert bool(((F - (S + B)).abs() <= tol + 1e-300).all())
    # checksum of checksums; F - B cancels where breakup dominates, so the bound is on the gross sums
    tol_sum = 1e-13 * torch.einsum("ic->i", S.abs() + B.abs())
    assert bool(((torch.einsum("ic->i", F - B) - chk).abs() <= tol_sum).all())


def test_full_size_step_realisable_and_idempotent_layout(ctx):
    w = synthetic.workload("moc100_ct")
    n, eps, rho, nu, xi = _device_moc_fields(w, FULL_MOC, seed=2)
    moc = engine.MOC(ctx, w.N, w.xi1, engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs))
    n0 = n.clone()
    S = torch.empty_like(n)
    mom = torch.zeros(2, dtype=torch.float64, device="cuda")
    moc.step(n, eps, rho, nu, 10.0, S=S, moments=mom)   # huge dt: limiter active in many cells
    # MOC realisability (MOC.C:87-91): a fully limited class ends at n + dt RN(-n/dt), which is 0
    # up to one rounding of n -- in the reference exactly as here
    assert bool((n >= -4e-16 * n0).all()), "MOC realisability: n >= 0 after the limited step"
    assert bool((n == n0 + 10.0 * S).all())
    # fused global moments agree with a separate reduction of the updated field
    np.testing.assert_allclose(mom.cpu().numpy(), [float(n.sum()), float(torch.einsum("i,ic->", xi, n))], rtol=1e-10)
    # a sub-range launched on its own gives bit-identical results (cells are independent: sharding)
    lo, hi = 5_000_037, 5_000_037 + 262_144
    Ss = moc.source(n0[:, lo:hi].contiguous(), eps[lo:hi], rho[lo:hi], nu[lo:hi], 10.0)
    assert bool((Ss == S[:, lo:hi]).all())


def test_full_size_wheeler_round_trip(ctx):
    g = torch.Generator(device="cuda").manual_seed(3)
    nc = FULL_QMOM
    nodes, _ = torch.sort(10 ** (-12 + 2.477 * torch.rand((3, nc), generator=g, dtype=torch.float64, device="cuda")), dim=0)
    # keep the nodes apart (well-conditioned inversion)
    nodes[1] = torch.maximum(nodes[1], nodes[0] * 1.5)
    nodes[2] = torch.maximum(nodes[2], nodes[1] * 1.5)
    wts = 1e5 + 9.9e6 * torch.rand((3, nc), generator=g, dtype=torch.float64, device="cuda")
    m = torch.stack([(wts * nodes ** k).sum(0) for k in range(6)]).contiguous()
    x, w = ctx.wheeler_batch(m)
    assert bool(torch.isfinite(x).all()) and bool((w > 0).all()) and bool((x[1:] >= x[:-1]).all())
    back = torch.stack([(w * x ** k).sum(0) for k in range(6)])
    assert float(((back - m).abs() / m).max()) < 1e-8
    assert float(((x - nodes).abs() / nodes).max()) < 1e-5


def test_large_fixed_pivot_conserves_volume_and_scales(ctx):
    """Fixed-pivot kernel (row f4) at N = 100 on a geometric grid, 3e5 cells: the two-pivot redistribution
    conserves the volume of every coalescence event (sum_i x_i C_i = 0 against the gross volume exchanged), the
    operator is exactly quadratic (powers of two commute with every rounding), breakup conserves volume."""
    w = synthetic.workload("moc100_ct")
    nc = 300_000
    n, eps, rho, nu, _ = _device_moc_fields(w, nc, seed=4)
    x = w.xi1 * 1.08 ** np.arange(w.N)
    xd = T(x)
    coal = engine.MOC(ctx, w.N, 0.0, engine.make_spec(w.coalescence, w.coal_coeffs), 1e300, pivots=x)
    assert coal.kernel()["kernel"] == "pivot"
    S = coal.source(n, eps, rho, nu, w.dt)
    net = torch.einsum("i,ic->c", xd, S)
    gross = torch.einsum("i,ic->c", xd, S.abs())
    assert float((net.abs() / (gross + 1e-300)).max()) < 1e-11 and bool(torch.isfinite(S).all())
    assert bool((S[0] <= 0).all())  # the smallest class can only lose drops to coalescence
    S2 = coal.source(n * 2.0, eps, rho, nu, w.dt)
    assert bool((S2 == 4.0 * S).all())
    br = engine.MOC(ctx, w.N, 0.0, engine.make_spec("noCoalescence", (), w.breakup, w.br_coeffs), 1e300, pivots=x)
    B = br.source(n, eps, rho, nu, w.dt)
    netb = torch.einsum("i,ic->c", xd, B)
    grossb = torch.einsum("i,ic->c", xd, B.abs())
    assert float((netb.abs() / (grossb + 1e-300)).max()) < 1e-11
    assert bool((br.source(n * 2.0, eps, rho, nu, w.dt) == 2.0 * B).all())
    coal.close()
    br.close()


def test_large_mules_update_is_bounded_and_form_independent(ctx, monkeypatch):
    """MULES-limited transport (Gauss vanLeer, 3 limiter sweeps, limitSum) on a 96^3 block x 8 fields: the update
    stays inside [psiMin, psiMax] (what the limiter is for), and the fused kernels give bit for bit what the
    per-field kernels give."""
    nx = 96
    m = synthetic.block_mesh(nx, nx, nx, velocity=(1.0, 0.5, 0.25))
    mesh = engine.Mesh(ctx, m["owner"], m["neighbour"], m["boundary_owner"], m["vcell"])
    mesh.set_geometry(m["Sf"], m["C"], m["weights"])
    nc, K = nx ** 3, 8
    g = torch.Generator(device="cuda").manual_seed(11)
    psi = torch.rand(K, nc, dtype=torch.float64, device="cuda", generator=g)
    psi *= (torch.rand(K, nc, dtype=torch.float64, device="cuda", generator=g) > 0.3)  # sharp fronts
    pmax = torch.full((K,), 1.0, dtype=torch.float64, device="cuda")
    phi = T(m["phi"])
    dt = 0.3 / (nx * 1.75)
    out = mesh.transport_mules(psi, phi, dt, pmax)
    assert bool(torch.isfinite(out).all())
    assert float(out.min()) >= -1e-12 and float(out.max()) <= 1.0 + 1e-12
    monkeypatch.setenv("PBE_MULES_PATH", "0")
    ref = mesh.transport_mules(psi, phi, dt, pmax)
    assert bool((out == ref).all())
    mesh.close()
