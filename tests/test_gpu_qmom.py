"""GPU parity tests for the QMOM path: batched Wheeler inversion + quadrature moment sources.

Oracle = oracle/pbe_oracle.c (MomentInversion.C:9-51 with an implicit-QL eigen-solve in place of
Eigen; QMOM.C:215-267, 98-128).  Tolerances: nodes rel 1e-10, weights rel 1e-9 on
well-conditioned inputs; sources |dS| <= 1e-9 * scale; 1e-4 (l2) against the reference's own
Marchisio & Fox goldens; the NaN pattern of degenerate inputs must match exactly.
"""
import json
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from openpbe_b200 import capi, engine, synthetic  # noqa: E402

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CT = (0.0152, 0.0678, 0.05, 0.04282)


@pytest.fixture(scope="module")
def ctx():
    assert torch.cuda.is_available()
    c = engine.Context(0)
    yield c
    c.close()


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).cuda()


def test_wheeler_reference_goldens(ctx):
    g = json.load(open(os.path.join(GOLD, "reference_unit_tests.json")))
    for key in ("G7_wheeler_example_3_1", "G8_wheeler_example_3_2"):
        m = np.array(g[key]["moments"], dtype=np.float64)[:, None]
        x, w = ctx.wheeler_batch(T(m))
        x, w = x.cpu().numpy()[:, 0], w.cpu().numpy()[:, 0]
        assert np.linalg.norm(x - np.array(g[key]["nodes"])) < 1e-4      # the reference's tolerance
        assert np.linalg.norm(w - np.array(g[key]["weights"])) < 1e-4
    # example 3.2 is exact analytically: hold the device to 1e-12
    x, w = ctx.wheeler_batch(T(np.array(g["G8_wheeler_example_3_2"]["moments"], dtype=np.float64)[:, None]))
    np.testing.assert_allclose(x.cpu().numpy()[:, 0], g["G8_wheeler_example_3_2"]["nodes"], rtol=1e-12, atol=1e-14)
    np.testing.assert_allclose(w.cpu().numpy()[:, 0], g["G8_wheeler_example_3_2"]["weights"], rtol=1e-12)


@pytest.mark.parametrize("order", [1, 2, 3, 4])
def test_wheeler_vs_oracle_random(ctx, oracle, order):
    rng = np.random.default_rng(40 + order)
    nc = 3001
    x0 = np.cumsum(rng.uniform(0.4, 1.0, (order, nc)), axis=0)
    w0 = rng.uniform(0.1, 2.0, (order, nc))
    m = np.stack([np.sum(w0 * x0 ** k, axis=0) for k in range(2 * order)])
    x, w = ctx.wheeler_batch(T(m))
    x, w = x.cpu().numpy(), w.cpu().numpy()
    xo = np.empty_like(x)
    wo = np.empty_like(w)
    for c in range(nc):
        xo[:, c], wo[:, c] = oracle.wheeler(m[:, c])
    np.testing.assert_allclose(x, xo, rtol=1e-10)
    np.testing.assert_allclose(w, wo, rtol=1e-9, atol=1e-13)
    # realisability: weights >= 0, nodes ascending and inside the support
    assert np.all(w >= 0) and np.all(np.diff(x, axis=0) >= 0)
    assert np.all(x >= x0.min(axis=0) * (1 - 1e-6)) and np.all(x <= x0.max(axis=0) * (1 + 1e-6))
    if order <= 3:
        np.testing.assert_allclose(x, x0, rtol=1e-6)
        np.testing.assert_allclose(w, w0, rtol=1e-5)


def test_wheeler_degenerate_nan_pattern(ctx, oracle):
    cases = [
        [2.0, 3.0, 5.0, 9.0, 17.0, 33.0],        # two-node measure, three nodes asked: 0/0
        [0.0, 0.0, 0.0, 0.0, 0.0, 0.0],          # empty cell
        [1.0, 1.0, 1.0, 1.0, 1.0, 1.0],          # single node
        [1.0, 2.0, 4.0, 8.0, 16.0, 32.0],
        [1.0, 0.5, 1.0, 2.0, 7.0, 1.0],          # unrealisable but finite
        [np.nan, 1.0, 1.0, 1.0, 1.0, 1.0],
        [np.inf, 1.0, 1.0, 1.0, 1.0, 1.0],
        [3.0, 6.0, 14.0, 36.0, 98.0, 276.0],     # regular: nodes 1,2,3 weights 1,1,1
    ]
    m = np.array(cases, dtype=np.float64).T.copy()
    x, w = ctx.wheeler_batch(T(m))
    x, w = x.cpu().numpy(), w.cpu().numpy()
    for c in range(m.shape[1]):
        xo, wo = oracle.wheeler(m[:, c])
        assert np.array_equal(np.isnan(x[:, c]), np.isnan(xo)), f"case {c}: {x[:, c]} vs {xo}"
        assert np.array_equal(np.isnan(w[:, c]), np.isnan(wo)), f"case {c}: {w[:, c]} vs {wo}"
    np.testing.assert_allclose(x[:, -1], [1, 2, 3], rtol=1e-9)
    np.testing.assert_allclose(w[:, -1], [1, 1, 1], rtol=1e-8)


@pytest.mark.parametrize("order", [1, 2, 3, 4])
@pytest.mark.parametrize("kern", ["ct", "const", "mixed", "none"])
def test_source_parity(ctx, oracle, order, kern):
    w = synthetic.workload("qmom3_ct" if kern in ("ct", "mixed") else "qmom3_const", 5003)
    w.N = order
    f = synthetic.qmom_fields(w, seed=order * 7 + 1)
    sel = {"ct": ("CoulaloglouTavlaridesC", w.coal_coeffs, "CoulaloglouTavlarides", w.br_coeffs),
           "const": ("constantCoalescence", (0.1,), "binaryBreakup", ()),
           "mixed": ("constantCoalescence", (1e-9,), "CoulaloglouTavlarides", w.br_coeffs),
           "none": ("noCoalescence", (), "noBreakup", ())}[kern]
    q = engine.QMOM(ctx, order, engine.make_spec(*sel))
    S = q.source(T(f["m"]), T(f["eps"]), T(f["rho"]), T(f["nu"])).cpu().numpy()
    So, scale = oracle.qmom_source(oracle.make_spec(*sel), order, f["m"], f["eps"], f["rho"], f["nu"], nthreads=8)
    assert np.array_equal(np.isnan(S), np.isnan(So)), "NaN pattern (cells later scrubbed, QMOM.C:101-107)"
    well = np.ones(w.ncells, bool) if order < 2 else (f["nodes"][1] / f["nodes"][0] > 1.001)
    ok = np.isfinite(So) & well[None, :]
    assert np.all(np.abs(S - So)[ok] <= 1e-9 * scale[ok] + 1e-300), \
        f"max {np.max(np.abs(S - So)[ok] / (scale[ok] + 1e-300)):.3e}"
    if kern == "none":
        assert np.all(S[np.isfinite(S)] == 0.0)


def test_reference_shaped_oracle_gives_same_sources(oracle):
    """The reference re-inverts for every moment (QMOM.C:231); inverting once is the same result."""
    w = synthetic.workload("qmom3_ct", 500)
    f = synthetic.qmom_fields(w)
    sp = oracle.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    a, _ = oracle.qmom_source(sp, 3, f["m"], f["eps"], f["rho"], f["nu"], reference_shaped=True)
    b, _ = oracle.qmom_source(sp, 3, f["m"], f["eps"], f["rho"], f["nu"], reference_shaped=False)
    np.testing.assert_array_equal(a, b)


def test_step_scrub_clamp_diameter(ctx, oracle):
    w = synthetic.workload("qmom3_ct", 4001)
    f = synthetic.qmom_fields(w, degenerate_fraction=0.05)
    m_h = f["m"].copy()
    m_h[:, 7] = 0.0                     # empty cell: NaN sources -> scrub -> clamp to SMALL
    m_h[:, 11] = [1, 1, 1, 1, 1, 1]     # single-node cell
    spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    ospec = oracle.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    q = engine.QMOM(ctx, 3, spec)
    m = T(m_h)
    S = torch.empty_like(m)
    d = torch.empty(w.ncells, dtype=torch.float64, device="cuda")
    dt = 1e-3
    S_raw = q.source(m, T(f["eps"]), T(f["rho"]), T(f["nu"])).cpu().numpy()   # before the scrub
    q.step(m, T(f["eps"]), T(f["rho"]), T(f["nu"]), dt, S=S, d=d)
    So, scale = oracle.qmom_source(ospec, 3, m_h, f["eps"], f["rho"], f["nu"], nthreads=8)
    nan_o = np.isnan(So)
    m_o = m_h.copy()
    d_o = oracle.qmom_update(m_o, So, dt)   # scrubs So in place
    S, m, d = S.cpu().numpy(), m.cpu().numpy(), d.cpu().numpy()
    assert not np.isnan(S).any() and not np.isnan(m).any()
    assert np.array_equal(np.isnan(S_raw), nan_o) and nan_o.any()   # same entries are NaN ...
    assert np.all(S[nan_o] == 0.0) and np.all(So[nan_o] == 0.0)     # ... and scrubbed to 0 (QMOM.C:101-107)
    assert np.all(m >= 1e-15)                                  # QMOM realisability floor (SMALL)
    assert np.all(m[:, 7] == 1e-15)
    well = (f["nodes"][1] / f["nodes"][0] > 1.001)
    well[[7, 11]] = False
    assert np.all(np.abs(S - So)[:, well] <= 1e-9 * scale[:, well] + 1e-300)
    np.testing.assert_allclose(m[:, well], m_o[:, well], rtol=1e-9)
    np.testing.assert_allclose(d[well], d_o[well], rtol=1e-9)
    assert d[7] == pytest.approx((6 / np.pi) ** (1 / 3), rel=1e-14)


@pytest.mark.parametrize("name", ["qmom3_ct_1000", "qmom3_const_500"])
def test_golden_fixture(ctx, name):
    z = np.load(os.path.join(GOLD, name + ".npz"))
    w = synthetic.workload(str(z["workload"]))
    q = engine.QMOM(ctx, int(z["order"]), engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs))
    m = T(z["m"])
    e, r, v = T(z["eps"]), T(z["rho"]), T(z["nu"])
    S = q.source(m, e, r, v).cpu().numpy()
    assert np.array_equal(np.isnan(S), np.isnan(z["S"]))
    ok = np.isfinite(z["S"]) & z["well"][None, :]
    assert np.all(np.abs(S - z["S"])[ok] <= 1e-9 * z["scale"][ok] + 1e-300)
    x, wts = ctx.wheeler_batch(m)
    wl = z["well"]
    np.testing.assert_allclose(x.cpu().numpy()[:, wl], z["x"][:, wl], rtol=1e-10)
    np.testing.assert_allclose(wts.cpu().numpy()[:, wl], z["w"][:, wl], rtol=1e-9)
    d = torch.empty(m.shape[1], dtype=torch.float64, device="cuda")
    q.step(m, e, r, v, float(z["dt"]), d=d)
    np.testing.assert_allclose(m.cpu().numpy()[:, wl], z["m_after"][:, wl], rtol=1e-9)
    np.testing.assert_allclose(d.cpu().numpy()[wl], z["d_after"][wl], rtol=1e-9)


def test_source_host_and_errors(ctx):
    w = synthetic.workload("qmom3_ct", 70_001)
    f = synthetic.qmom_fields(w)
    spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    q = engine.QMOM(ctx, 3, spec)
    a = q.source_host(f["m"], f["eps"], f["rho"], f["nu"]).numpy()
    b = q.source(T(f["m"]), T(f["eps"]), T(f["rho"]), T(f["nu"])).cpu().numpy()
    np.testing.assert_array_equal(a, b)
    with pytest.raises(capi.PbeError) as e:
        engine.QMOM(ctx, 5, spec)
    assert e.value.code == capi.PBE_E_UNSUPPORTED
    with pytest.raises(capi.PbeError):
        engine.QMOM(ctx, 3, engine.make_spec("table", (), "noBreakup", (), coal_table=np.ones((3, 3))))
    m0 = torch.empty((6, 0), dtype=torch.float64, device="cuda")
    e0 = torch.empty(0, dtype=torch.float64, device="cuda")
    assert q.source(m0, e0, e0, e0).shape == (6, 0)
