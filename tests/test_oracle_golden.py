"""Pin the oracle on the reference's own known-answer tests.

Every expected value below is quoted from /root/reference/tests/unit_tests/*.cpp (file:line in
each test); nothing here reads the reference tree at run time.
"""
import math

import numpy as np
import pytest

from oracle import numpy_restatement as npo


def ulp_diff(a, b):
    ia = np.array([a], dtype=np.float64).view(np.int64)[0]
    ib = np.array([b], dtype=np.float64).view(np.int64)[0]
    return abs(int(ia) - int(ib))


# G1: tests/unit_tests/breakup.cpp:6-17
def test_g1_binary_breakup(oracle):
    assert oracle.lib().orc_break_binary(6.0) == 36.0


# G2: tests/unit_tests/breakup.cpp:19-32
def test_g2_no_breakup(oracle):
    assert oracle.lib().orc_break_none(6.0) == 0.0


# G3: tests/unit_tests/breakup.cpp:34-59 (ASSERT_DOUBLE_EQ = 4 ulp)
G3 = dict(c1=0.0152, c2=0.0678, sigma=0.04282, gamma=0.050000000000000003,
          eps=0.12935361098784559, rho_d=972.0, xi=9.738310380055584507e-11,
          expected=1.122262582852271567e-02)


def test_g3_ct_breakup(oracle):
    g = G3
    r = oracle.lib().orc_break_ct(g["c1"], g["c2"], g["gamma"], g["sigma"], g["xi"], g["rho_d"], g["eps"])
    assert ulp_diff(r, g["expected"]) <= 4
    r2 = float(npo.break_ct(g["c1"], g["c2"], g["gamma"], g["sigma"], g["xi"], g["rho_d"], g["eps"]))
    assert r2 == pytest.approx(g["expected"], rel=2e-15)  # numpy SIMD pow: a few ulp


# G4: tests/unit_tests/coalescence.cpp:6-22 ; G5: :24-38
def test_g4_g5_constant_and_none(oracle):
    assert oracle.lib().orc_coal_constant(5.0) == 5.0
    assert oracle.lib().orc_coal_none() == 0.0


# G6: tests/unit_tests/coalescence.cpp:40-78.  The file's expected value still contains the
# "/ Vcell" factor that coalescenceKernels/CoulaloglouTavlarides.C:78-83 has since commented out
# (ASSERT_FLOAT_EQ = 4 float ulp); HEAD therefore returns expected * Vcell.
G6 = dict(c1=1.06e-12, c2=5.13e13, sigma=0.04282, gamma=0.050000000000000003,
          eps=0.12935361098784559, rho=1000.0, nu=1e-6, xi1=5.071423503403169858e-10,
          xi2=2.885425297794245861e-11, V=0.012, file_expected=3.199680897991421460e-21,
          head_expected=3.839617077589699e-23)


def test_g6_ct_coalescence_head_and_stale(oracle):
    g = G6
    L = oracle.lib()
    head = L.orc_coal_ct(g["c1"], g["c2"], g["gamma"], g["sigma"], g["xi1"], g["xi2"], g["eps"],
                         g["rho"], g["nu"], g["V"], 0)
    stale = L.orc_coal_ct(g["c1"], g["c2"], g["gamma"], g["sigma"], g["xi1"], g["xi2"], g["eps"],
                          g["rho"], g["nu"], g["V"], 1)
    # the reference's literal golden, at its own float tolerance, with the historical /Vcell
    assert np.float32(stale) == pytest.approx(np.float32(g["file_expected"]), rel=4 * 1.2e-7)
    assert stale == pytest.approx(g["file_expected"], rel=1e-13)
    # HEAD behaviour
    assert head == pytest.approx(g["head_expected"], rel=1e-13)
    assert head == pytest.approx(g["file_expected"] * g["V"], rel=1e-13)
    np_head = float(npo.coal_ct(g["c1"], g["c2"], g["gamma"], g["sigma"], g["xi1"], g["xi2"],
                                g["eps"], g["rho"], g["nu"]))
    assert np_head == pytest.approx(head, rel=1e-14)


# G7: tests/unit_tests/moment_inversion.cpp:15-30 (Marchisio & Fox example 3.1)
def test_g7_wheeler_example31(oracle):
    m = np.array([1, 5, 26, 140, 778, 4450, 26140, 157400], dtype=np.float64)
    w1 = 1 / (4 * math.sqrt(6) + 12)
    w2 = 1 / (12 - 4 * math.sqrt(6))
    x, w = oracle.wheeler(m)
    assert np.linalg.norm(x - np.array([2.6656, 4.2580, 5.7420, 7.3344])) < 1e-4
    assert np.linalg.norm(w - np.array([w1, w2, w2, w1])) < 1e-4
    xn, wn = npo.wheeler(m)
    np.testing.assert_allclose(x, xn, rtol=1e-10)
    np.testing.assert_allclose(w, wn, rtol=1e-8)


# G8: tests/unit_tests/moment_inversion.cpp:32-49 (example 3.2)
def test_g8_wheeler_example32(oracle):
    m = np.array([1, 0, 1, 0, 3, 0, 15, 0], dtype=np.float64)
    xi1 = math.sqrt(math.sqrt(6) + 3)
    xi2 = math.sqrt(3 - math.sqrt(6))
    w1 = 1 / (4 * math.sqrt(6) + 12)
    w2 = 1 / (12 - 4 * math.sqrt(6))
    x, w = oracle.wheeler(m)
    assert np.linalg.norm(x - np.array([-xi1, -xi2, xi2, xi1])) < 1e-4
    assert np.linalg.norm(w - np.array([w1, w2, w2, w1])) < 1e-4
    # analytic values are exact for this example: hold the oracle to 1e-12
    np.testing.assert_allclose(x, [-xi1, -xi2, xi2, xi1], rtol=1e-12, atol=1e-14)
    np.testing.assert_allclose(w, [w1, w2, w2, w1], rtol=1e-12)


# G9: tests/unit_tests/integration.cpp:6-13 (ASSERT_DOUBLE_EQ = 4 ulp)
def test_g9_integrate_exp(oracle):
    r = oracle.lib().orc_integrate_exp_neg(0.0)
    assert ulp_diff(r, 1.0) <= 4
    assert oracle.lib().orc_integrate_exp_neg(2.0) == pytest.approx(math.exp(-2.0), abs=1e-7)


# daughter distribution: uniformBinaryBreakup.C:55-69
def test_daughter_distribution(oracle):
    L = oracle.lib()
    assert L.orc_daughter_beta(1.0, 4.0) == 0.5
    assert L.orc_daughter_beta(4.0, 4.0) == 0.5      # pos(0) = 1
    assert L.orc_daughter_beta(5.0, 4.0) == 0.0
    assert L.orc_daughter_moment(3.0, 2) == 6.0
    assert L.orc_daughter_moment(3.0, 0) == 2.0


def test_wheeler_orders_and_random_vs_eigh(oracle):
    rng = np.random.default_rng(7)
    for order in (1, 2, 3, 4, 5, 6):
        for _ in range(50):
            x0 = np.cumsum(rng.uniform(0.4, 1.0, order))  # well separated nodes
            w0 = rng.uniform(0.1, 2.0, order)
            m = np.array([np.sum(w0 * x0 ** k) for k in range(2 * order)])
            x, w = oracle.wheeler(m)
            xn, wn = npo.wheeler(m)
            tol = 1e-9 if order <= 4 else 1e-5
            np.testing.assert_allclose(x, xn, rtol=tol)
            np.testing.assert_allclose(w, wn, rtol=tol * 100, atol=1e-12)
            if order <= 3:
                np.testing.assert_allclose(x, x0, rtol=1e-7)
                np.testing.assert_allclose(w, w0, rtol=1e-6)


def test_wheeler_degenerate_gives_nan(oracle):
    # two-node measure asked for three nodes: sigma[3,2] = 0 -> 0/0 in a[2] (MomentInversion.C:26-30)
    x0 = np.array([1.0, 2.0])
    w0 = np.array([1.0, 1.0])
    m = np.array([np.sum(w0 * x0 ** k) for k in range(6)])
    x, w = oracle.wheeler(m)
    assert np.all(np.isnan(x)) and np.all(np.isnan(w))


def test_tridiag_matches_eigh(oracle):
    rng = np.random.default_rng(3)
    for n in (2, 3, 4, 8, 12):
        d = rng.normal(size=n)
        e = rng.normal(size=n - 1)
        rc, lam, z = oracle.tridiag_eig_firstrow(d, e)
        assert rc == 0
        J = np.diag(d) + np.diag(e, 1) + np.diag(e, -1)
        l2, V = np.linalg.eigh(J)
        np.testing.assert_allclose(lam, l2, rtol=1e-12, atol=1e-13)
        np.testing.assert_allclose(z ** 2, V[0] ** 2, rtol=1e-9, atol=1e-13)
