"""Host instantiation of openpbe_b200/csrc/pbe_eig.cuh (the arithmetic the CUDA QMOM kernels run:
Wheeler sigma recurrence + the fixed-work order-2/3 Jacobi eigen-solves + the generic QL) against
the oracle's restatement of MomentInversion.C:9-51, on the CPU.  The harness tests/host/
eig_harness.cpp is compiled with g++ -ffp-contract=off; no GPU and no CUDA runtime involved.

Tolerances as in tests/test_gpu_qmom.py: nodes rel 1e-10, weights rel 1e-9 on
well-conditioned cells; NaN pattern identical on every cell.
"""
import os
import subprocess

import numpy as np
import pytest

from openpbe_b200 import synthetic

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def harness(tmp_path_factory):
    d = tmp_path_factory.mktemp("eig")
    exe = d / "eig_harness"
    subprocess.check_call(["/usr/bin/g++", "-O2", "-std=c++17", "-ffp-contract=off", "-I",
                           os.path.join(ROOT, "openpbe_b200", "csrc"),
                           os.path.join(ROOT, "tests", "host", "eig_harness.cpp"), "-o", str(exe)])

    def run(m):
        order, n = m.shape[0] // 2, m.shape[1]
        fin, fout = d / "in.bin", d / "out.bin"
        np.ascontiguousarray(m, dtype=np.float64).tofile(fin)
        out = subprocess.check_output([str(exe), str(order), str(n), str(fin), str(fout)], text=True)
        r = np.fromfile(fout).reshape(2, order, n)
        return r[0], r[1], int(out.split()[-1])
    return run


def _oracle(oracle, m):
    order, n = m.shape[0] // 2, m.shape[1]
    xo, wo = np.empty((order, n)), np.empty((order, n))
    for c in range(n):
        xo[:, c], wo[:, c] = oracle.wheeler(m[:, c])
    return xo, wo


@pytest.mark.parametrize("order", [1, 2, 3, 4])
def test_well_conditioned_matches_oracle(harness, oracle, order):
    rng = np.random.default_rng(40 + order)
    nc = 4000
    x0 = np.cumsum(rng.uniform(0.4, 1.0, (order, nc)), axis=0)
    w0 = rng.uniform(0.1, 2.0, (order, nc))
    m = np.stack([np.sum(w0 * x0 ** k, axis=0) for k in range(2 * order)])
    x, w, fallbacks = harness(m)
    xo, wo = _oracle(oracle, m)
    np.testing.assert_allclose(x, xo, rtol=1e-10)
    np.testing.assert_allclose(w, wo, rtol=1e-9, atol=1e-13)
    assert np.all(w >= 0) and np.all(np.diff(x, axis=0) >= 0)
    assert fallbacks == 0                      # the fixed-work path handles every such cell itself


def test_bench_distribution_order3(harness, oracle):
    """configs[1] fields: nodes over 2.5 decades, 1 % near-degenerate cells."""
    w = synthetic.workload("qmom3_ct", 20000)
    f = synthetic.qmom_fields(w)
    x, wt, fallbacks = harness(f["m"])
    xo, wo = _oracle(oracle, f["m"])
    assert np.array_equal(np.isnan(x), np.isnan(xo)) and np.array_equal(np.isnan(wt), np.isnan(wo))
    well = (f["nodes"][1] / f["nodes"][0] > 1.001) & np.isfinite(xo).all(axis=0)
    np.testing.assert_allclose(x[:, well], xo[:, well], rtol=1e-10)
    np.testing.assert_allclose(wt[:, well], wo[:, well], rtol=1e-9)
    # both quadratures reproduce the moments they were built from
    fin = np.isfinite(xo).all(axis=0) & (xo > 0).all(axis=0)
    rec = np.stack([np.sum(wt * x ** k, axis=0) for k in range(6)])
    assert np.max(np.abs(rec - f["m"])[:, fin] / f["m"][:, fin]) < 1e-11
    # the generic QL re-run is the exception (nearly decoupled Jacobi matrices), not the rule
    assert fallbacks < 0.01 * w.ncells


def test_degenerate_nan_pattern(harness, oracle):
    cases = [
        [2.0, 3.0, 5.0, 9.0, 17.0, 33.0], [0.0] * 6, [1.0] * 6, [1.0, 2.0, 4.0, 8.0, 16.0, 32.0],
        [1.0, 0.5, 1.0, 2.0, 7.0, 1.0], [np.nan, 1.0, 1.0, 1.0, 1.0, 1.0], [np.inf, 1.0, 1.0, 1.0, 1.0, 1.0],
        [3.0, 6.0, 14.0, 36.0, 98.0, 276.0],
        [2.0, 2.0, 2.0, 2.0, 2.0, 2.0 + 1e-12],     # almost a single node
        [1.0, 1.0, 2.0, 4.0, 8.0, 16.0],
    ]
    m = np.array(cases, dtype=np.float64).T.copy()
    x, w, _ = harness(m)
    xo, wo = _oracle(oracle, m)
    assert np.array_equal(np.isnan(x), np.isnan(xo)) and np.array_equal(np.isnan(w), np.isnan(wo))
    np.testing.assert_allclose(x[:, 7], [1, 2, 3], rtol=1e-12)
    np.testing.assert_allclose(w[:, 7], [1, 1, 1], rtol=1e-11)
