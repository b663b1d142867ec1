"""The oracle port (oracle/pbe_oracle.c) against the reference ITSELF.

oracle/_ref/libpbe_ref.so holds the reference's own sources -- PBEMethod.C, MOC.C, QMOM.C,
MomentInversion.C, Utility.C, the coalescence / breakup kernels and uniformBinaryBreakup.C -- compiled
unmodified from /root/reference against the OpenFOAM / Eigen API shim of oracle/ref_shim
(`make -C oracle ref`).  These tests pin the restatement on it:
  * the reference's unit-test goldens come out of its own compiled functors
    (tests/unit_tests/{breakup,coalescence,moment_inversion}.cpp);
  * MOC::classSourceTerm (MOC.C:87-157) of the port is BIT-IDENTICAL to the reference's;
  * QMOM::momentSourceTerm (QMOM.C:215-267) agrees to 1e-12 sum|terms| (the symmetric eigen-solve is
    third-party in both: Eigen there -- restated by Jacobi rotations in the shim -- implicit QL in the port)
    with the identical NaN pattern;
  * one correct() (sources, zero-flux update, clamp, d) agrees to the update's rounding.
CPU only; skipped where the library was not built (no /root/reference and no prebuilt file).
"""
import json
import os

import numpy as np
import pytest

from openpbe_b200 import synthetic
from oracle import ref_oracle as ref

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_unit_tests.json")
CT = (0.0152, 0.0678, 0.05, 0.04282)

ref.build()
pytestmark = pytest.mark.skipif(not ref.available(), reason="oracle/_ref/libpbe_ref.so not built (needs /root/reference)")


@pytest.fixture(scope="module")
def L():
    return ref.lib()


def test_reference_goldens_from_its_own_functors(L):
    # tests/unit_tests/breakup.cpp:6-59
    assert L.ref_binary_breakup_S(6.0) == 36.0
    assert L.ref_no_breakup_S(6.0) == 0.0
    g3 = ref.ct_breakup_S((0.0152, 0.0678, 0.05, 0.04282), 9.738310380055584507e-11, 972.0, 0.12935361098784559)
    assert abs(g3 - 1.122262582852271567e-02) <= 4 * np.spacing(1.122262582852271567e-02)  # ASSERT_DOUBLE_EQ
    # tests/unit_tests/coalescence.cpp:6-78 (the CT literal there is stale: it includes "/ Vcell")
    assert L.ref_constant_coalescence_S(5.0) == 5.0
    assert L.ref_no_coalescence_S() == 0.0
    g6 = ref.ct_coalescence_S((1.06e-12, 5.13e13, 0.05, 0.04282), 5.071423503403169858e-10, 2.885425297794245861e-11,
                              0.12935361098784559, 1000.0, 1e-6, 0.012)
    assert g6 == pytest.approx(3.839617077589699e-23, rel=1e-14)
    assert g6 / 0.012 == pytest.approx(3.199680897991421460e-21, rel=1e-6)  # ASSERT_FLOAT_EQ on the stale literal
    # daughterParticleDistributions/uniformBinaryBreakup.C:55-69
    assert L.ref_uniform_binary_beta(1.0, 2.0) == 1.0 and L.ref_uniform_binary_beta(2.0, 2.0) == 1.0  # pos(0) = 1
    assert L.ref_uniform_binary_beta(3.0, 2.0) == 0.0
    assert L.ref_uniform_binary_moment(2.0, 3) == 2 * 8.0 / 4
    # tests/unit_tests/moment_inversion.cpp:15-49 (Marchisio & Fox examples 3.1, 3.2), ||delta|| < 1e-4
    w1, w2 = 1 / (4 * np.sqrt(6) + 12), 1 / (12 - 4 * np.sqrt(6))
    x, w = ref.wheeler([1, 5, 26, 140, 778, 4450, 26140, 157400])
    assert np.linalg.norm(x - [2.6656, 4.2580, 5.7420, 7.3344]) < 1e-4 and np.linalg.norm(w - [w1, w2, w2, w1]) < 1e-4
    x, w = ref.wheeler([1, 0, 1, 0, 3, 0, 15, 0])
    r1, r2 = np.sqrt(3 + np.sqrt(6)), np.sqrt(3 - np.sqrt(6))
    assert np.linalg.norm(x - [-r1, -r2, r2, r1]) < 1e-4 and np.linalg.norm(w - [w1, w2, w2, w1]) < 1e-4


def test_port_functors_equal_reference_functors(oracle):
    rng = np.random.default_rng(11)
    for _ in range(200):
        xi1, xi2 = 10.0 ** rng.uniform(-13, -9, 2)
        eps, rho, nu = 10.0 ** rng.uniform(-3, 0), rng.uniform(700, 1200), 10.0 ** rng.uniform(-6, -3)
        assert oracle.lib().orc_coal_ct(*CT, xi1, xi2, eps, rho, nu, 1.0, 0) == ref.ct_coalescence_S(CT, xi1, xi2, eps, rho, nu)
        assert oracle.lib().orc_break_ct(*CT, xi1, rho, eps) == ref.ct_breakup_S(CT, xi1, rho, eps)


@pytest.mark.parametrize("name,ncells", [("moc30_ct", 257), ("moc100_ct", 40), ("replay_moc", 64), ("moc100_const", 33)])
def test_port_moc_sources_bit_identical_to_reference(oracle, name, ncells):
    w = synthetic.workload(name, ncells)
    f = synthetic.moc_fields(w)
    S = ref.moc_source(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs, w.N, w.xi1, w.relax, f["n"], f["eps"],
                       f["rho"], f["nu"], w.dt)
    sp = oracle.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    So, _ = oracle.moc_source(sp, w.N, w.xi1, w.relax, f["n"], f["eps"], f["rho"], f["nu"], w.dt)
    assert np.array_equal(S, So)


@pytest.mark.parametrize("N", [1, 2, 3, 7, 16])
@pytest.mark.parametrize("kernels", [("constantCoalescence", (0.1,), "binaryBreakup", ()),
                                     ("noCoalescence", (), "CoulaloglouTavlarides", CT),
                                     ("CoulaloglouTavlaridesC", CT, "noBreakup", ())])
def test_port_moc_small_grids_and_limiter(oracle, N, kernels):
    coal, cc, br, bc = kernels
    rng = np.random.default_rng(N)
    nc = 50
    cell_indep = coal != "CoulaloglouTavlaridesC" and br != "CoulaloglouTavlarides"
    n = rng.uniform(0.0, 1.0, (N, nc)) if cell_indep else rng.uniform(1e5, 1e7, (N, nc))
    n[:, ::7] *= 1e-6  # nearly empty cells
    xi1 = 0.25 if cell_indep else 1e-11
    eps = 10.0 ** rng.uniform(-3, 0, nc)
    rho, nu = np.full(nc, 972.0), np.full(nc, 1e-3)
    dt = 50.0 if cell_indep else 1e3  # huge steps: the limiter max(C + B, -relax n / dt) binds (MOC.C:87-91)
    for relax in (1.0, 0.5):
        S = ref.moc_source(coal, cc, br, bc, N, xi1, relax, n, eps, rho, nu, dt)
        So, _ = oracle.moc_source(oracle.make_spec(coal, cc, br, bc), N, xi1, relax, n, eps, rho, nu, dt)
        assert np.array_equal(S, So)
        if N > 2 and br != "noBreakup":
            assert np.any(S == -relax * n / dt)  # the limiter did bind somewhere


def test_reference_correct_step_and_d(oracle):
    """MOC::correct() + d() (MOC.C:159-184, 282-290), zero flux: n <- (n rDeltaT + S) / rDeltaT in
    MULES::explicitSolve's form vs the port's n + dt S: equal to the update's rounding; both
    transport branches (usingMULES true / false) see the same sources."""
    w = synthetic.workload("moc30_ct", 101)
    f = synthetic.moc_fields(w)
    txt = ref.pbe_dict("methodOfClasses", w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs, N=w.N, xi1=w.xi1,
                       relax=w.relax)
    n1, S1, d1, _ = ref.method_steps(txt, "n", f["n"], f["eps"], f["rho"], f["nu"], w.dt, alpha=f["alpha"], vcell=f["vcell"])
    sp = oracle.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    So, _ = oracle.moc_source(sp, w.N, w.xi1, w.relax, f["n"], f["eps"], f["rho"], f["nu"], w.dt)
    n_o = f["n"].copy()
    oracle.moc_update(n_o, So, w.dt)
    assert np.array_equal(S1, So)
    assert np.all(np.abs(n1 - n_o) <= 4.5e-16 * np.abs(n_o))
    assert np.allclose(d1, oracle.moc_d(n1, w.xi1), rtol=1e-14)
    txt_fvm = ref.pbe_dict("methodOfClasses", w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs, N=w.N, xi1=w.xi1,
                           relax=w.relax, usingMULES=False)
    n2, S2, _, _ = ref.method_steps(txt_fvm, "n", f["n"], f["eps"], f["rho"], f["nu"], w.dt, vcell=f["vcell"])
    assert np.array_equal(S2, So)
    assert np.all(np.abs(n2 - n_o) <= 9e-16 * np.abs(n_o))
    # three steps in one call = three calls
    n3, _, _, _ = ref.method_steps(txt, "n", f["n"], f["eps"], f["rho"], f["nu"], w.dt, nsteps=3, want_sources=False)
    n_it = f["n"]
    for _ in range(3):
        n_it, _, _, _ = ref.method_steps(txt, "n", n_it, f["eps"], f["rho"], f["nu"], w.dt, want_sources=False)
    assert np.array_equal(n3, n_it)


@pytest.mark.parametrize("name,ncells", [("qmom3_ct", 2000), ("qmom3_const", 500)])
def test_port_qmom_sources_match_reference(oracle, name, ncells):
    w = synthetic.workload(name, ncells)
    f = synthetic.qmom_fields(w)
    S = ref.qmom_source(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs, w.N, f["m"], f["eps"], f["rho"], f["nu"], w.dt)
    sp = oracle.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    So, scale = oracle.qmom_source(sp, w.N, f["m"], f["eps"], f["rho"], f["nu"], reference_shaped=True)
    # the reference scrubs NaN sources to 0 (QMOM.C:101-107) before the probe sees them
    nan_port = np.isnan(So)
    assert np.array_equal(nan_port, (S == 0.0) & nan_port)
    So = np.where(nan_port, 0.0, So)
    sep = np.ones(ncells, dtype=bool)
    for k in range(w.N - 1):
        sep &= f["nodes"][k + 1] / f["nodes"][k] > 1.001
    assert np.all(np.abs(S - So)[:, sep] <= 1e-12 * scale[:, sep] + 1e-300)


@pytest.mark.parametrize("order", [1, 2, 3, 4])
def test_port_qmom_orders_and_correct(oracle, order):
    w = synthetic.workload("qmom3_const", 300)
    w.N = order
    f = synthetic.qmom_fields(w, degenerate_fraction=0.0)
    txt = ref.pbe_dict("quadratureMethodOfMoments", w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs, order=order)
    m1, S1, d1, _ = ref.method_steps(txt, "m", f["m"], f["eps"], f["rho"], f["nu"], w.dt)
    sp = oracle.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    So, scale = oracle.qmom_source(sp, order, f["m"], f["eps"], f["rho"], f["nu"], reference_shaped=True)
    assert np.all(np.abs(S1 - So) <= 1e-10 * scale + 1e-300)
    m_o = f["m"].copy()
    d_o = oracle.qmom_update(m_o, So.copy(), w.dt)
    assert np.all(np.abs(m1 - m_o) <= 1e-10 * scale * w.dt + 9e-16 * np.abs(m_o))
    assert np.all(m1 >= 1e-15)  # clamp to SMALL, QMOM.C:117-123
    if order >= 1:
        assert np.allclose(d1, d_o, rtol=1e-9)


def test_port_wheeler_matches_reference(oracle):
    rng = np.random.default_rng(3)
    for order in (1, 2, 3, 4):
        for _ in range(100):
            nodes = np.cumprod(rng.uniform(1.5, 4.0, order)) * 0.05  # well separated: the inversion is well conditioned
            wts = rng.uniform(0.1, 2.0, order)
            m = np.array([np.sum(wts * nodes ** k) for k in range(2 * order)])
            xr, wr = ref.wheeler(m)
            xo, wo = oracle.wheeler(m)
            assert np.allclose(xr, xo, rtol=1e-9) and np.allclose(wr, wo, rtol=1e-7, atol=1e-12)
            assert np.allclose(xr, nodes, rtol=1e-5)


def test_reference_rejects_unknown_types():
    txt = ref.pbe_dict("methodOfClasses", "noCoalescence", (), "noBreakup", (), N=2, xi1=1.0)
    z = np.zeros(3)
    with pytest.raises(RuntimeError, match="Unknown pbeModelType type"):  # PBEMethod.C:97-105
        ref.method_steps(txt.replace("methodOfClasses;", "methodOfClassez;"), "n", np.ones((2, 3)), z, z, z, 1.0)
    with pytest.raises(RuntimeError, match="Unknown coalescenceKernelType type"):  # coalescenceKernel.C:57-65
        ref.method_steps(txt.replace("coalescenceKernel noCoalescence", "coalescenceKernel Madsen"), "n", np.ones((2, 3)),
                         z, z, z, 1.0)
    with pytest.raises(RuntimeError, match="noCoalescenceCoeffs is undefined"):  # subDict is fatal, coalescenceKernel.C:66-70
        ref.method_steps(txt.replace("noCoalescenceCoeffs", "otherCoeffs"), "n", np.ones((2, 3)), z, z, z, 1.0)


def test_parallel_workers_equal_serial():
    w = synthetic.workload("moc30_ct", 400)
    f = synthetic.moc_fields(w)
    a = (w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs, w.N, w.xi1, w.relax, f["n"], f["eps"], f["rho"], f["nu"], w.dt)
    S1 = ref.moc_source(*a)
    S3, secs = ref.moc_source(*a, nthreads=3, return_seconds=True)
    assert np.array_equal(S1, S3) and secs > 0.0
