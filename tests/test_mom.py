"""methodOfMoments (first slice): MOM::correct()'s per-cell path, MOM.C:148-236,
259-362.  PARITY UNPINNED in the reference (no MOM test, GSL / Boost.Math absent): the checkers are
oracle/mom_oracle.py -- nested adaptive QUADPACK integrals in the reference's shape (fixture
tests/golden/mom_adaptive.json), the same fixed-depth rule as the device kernel in NumPy, and the
closed forms for constantCoalescence + binaryBreakup.

Tolerances: device vs the NumPy fixed rule |dS| <= 1e-10 * scale (scale = largest |S_k| of the cell
and 1; exp/log/lgamma differ by ulps, S_1 is a cancelling sum); fixed rule vs adaptive 1e-6 * scale
(the reference asks QAGIU for epsabs 1e-7 per integral, nested).
"""
import json
import os

import numpy as np
import pytest

from oracle import mom_oracle as mo

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "mom_adaptive.json")
CT = (1.0, 1.0, 0.0, 1.0)
SPECS = {
    "ct_ct": {"coalescence": "CoulaloglouTavlaridesC", "coal_c": CT,
              "breakup": "CoulaloglouTavlarides", "br_c": CT},
    "const_binary": {"coalescence": "constantCoalescence", "coal_c": (0.1,),
                     "breakup": "binaryBreakup", "br_c": ()},
    "ct_none": {"coalescence": "CoulaloglouTavlaridesC", "coal_c": CT, "breakup": "noBreakup", "br_c": ()},
    "none_ct": {"coalescence": "noCoalescence", "coal_c": (), "breakup": "CoulaloglouTavlarides", "br_c": CT},
}


def random_cells(rng, nc):
    m0 = rng.uniform(0.5, 3.0, nc)
    k = rng.uniform(1.5, 4.0, nc)
    theta = rng.uniform(0.15, 0.6, nc)
    m = np.stack([m0, m0 * k * theta, m0 * k * (k + 1) * theta ** 2])
    return m, rng.uniform(0.5, 2.0, nc), rng.uniform(0.8, 1.2, nc), rng.uniform(0.5, 1.5, nc)


def scale(S):
    return np.maximum(np.max(np.abs(S), axis=0), 1.0)


# ---------------------------------------------------------------------------- CPU: the oracle
def test_gamma_parameters_and_pdf():
    c0, k, theta = mo.gamma_params(2.0, 1.0, 0.75)  # MOM.C:151-161
    assert c0 == 2.0 and abs(k - 2.0) < 1e-12 and theta == 0.25
    x = np.array([0.1, 0.5, 2.0])
    np.testing.assert_allclose(mo.gamma_pdf(2.0, 0.25, x), x * np.exp(-x / 0.25) / 0.25 ** 2, rtol=1e-13)
    assert np.isnan(mo.gamma_pdf(-1.0, 0.25, 0.3))  # Boost raises a domain error


def test_adaptive_oracle_reproduces_closed_form():
    m = (2.0, 1.0, 0.75)
    S = mo.mom_source_cell(SPECS["const_binary"], m, 1.0, 1.0, 1.0)
    coal, br = mo.closed_form_constant_binary(m, 0.1)
    np.testing.assert_allclose(S, coal + br, atol=1e-6)  # nested epsabs 1e-7


def test_fixed_rule_converges_to_closed_form():
    m = (1.3, 0.9, 0.85)
    coal, br = mo.closed_form_constant_binary(m, 0.1)
    err = [np.max(np.abs(mo.mom_source_cell_fixed(SPECS["const_binary"], m, 1.0, 1.0, 1.0, p) - coal - br))
           for p in (2, 4, 16)]
    assert err[0] > err[1] > err[2] or err[1] < 1e-11
    assert err[2] < 1e-11


def test_fixed_rule_matches_adaptive_fixture():
    g = json.load(open(GOLD))
    assert len(g["cases"]) >= 9
    for case in g["cases"]:
        S = mo.mom_source_cell_fixed(SPECS[case["spec"]], case["m"], case["eps"], case["rho"], case["nu"], 12)
        ref = np.array(case["S"])
        assert np.max(np.abs(S - ref)) <= 1e-6 * max(1.0, np.max(np.abs(ref))), (case, S)


def test_step_restatement():
    S, m, d = mo.mom_step_cell(np.array([np.nan, -1e9, 0.5]), [2.0, 1.0, 0.75], 0.1)
    assert S[0] == 0.0 and m[0] == 2.0 and m[1] == mo.SMALL and m[2] == 0.8
    assert d == (6 / np.pi * mo.SMALL / 2.0) ** (1 / 3)


# ---------------------------------------------------------------------------- GPU: the kernel
torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def ctx():
    from openpbe_b200 import engine
    assert torch.cuda.is_available()
    c = engine.Context(0)
    yield c
    c.close()


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).cuda()


def dev_spec(name):
    from openpbe_b200 import engine
    s = SPECS[name]
    return engine.make_spec(s["coalescence"], s["coal_c"], s["breakup"], s["br_c"])


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(SPECS))
@pytest.mark.parametrize("panels", [0, 5])
def test_device_matches_fixed_rule(ctx, name, panels):
    from openpbe_b200 import engine
    rng = np.random.default_rng(7 + panels)
    nc = 37  # ragged: not a multiple of the 4 cells a CTA takes
    m, eps, rho, nu = random_cells(rng, nc)
    mom = engine.MOM(ctx, dev_spec(name), panels)
    gam = torch.empty(3, nc, dtype=torch.float64, device="cuda")
    S = mom.source(T(m), T(eps), T(rho), T(nu), gamma=gam).cpu().numpy()
    want = np.stack([mo.mom_source_cell_fixed(SPECS[name], m[:, c], eps[c], rho[c], nu[c], panels or 12)
                     for c in range(nc)], axis=1)
    assert np.all(np.abs(S - want) <= 1e-10 * scale(want)), np.max(np.abs(S - want) / scale(want))
    gp = np.array([mo.gamma_params(*m[:, c]) for c in range(nc)]).T
    np.testing.assert_allclose(gam.cpu().numpy(), gp, rtol=1e-14)
    mom.close()


@pytest.mark.gpu
def test_device_matches_adaptive_fixture(ctx):
    from openpbe_b200 import engine
    g = json.load(open(GOLD))
    for name in sorted({c["spec"] for c in g["cases"]}):
        cases = [c for c in g["cases"] if c["spec"] == name]
        m = np.array([c["m"] for c in cases]).T
        mom = engine.MOM(ctx, dev_spec(name))
        S = mom.source(T(m), T([c["eps"] for c in cases]), T([c["rho"] for c in cases]),
                       T([c["nu"] for c in cases])).cpu().numpy()
        ref = np.array([c["S"] for c in cases]).T
        assert np.all(np.abs(S - ref) <= 1e-6 * scale(ref)), (name, S, ref)
        mom.close()


@pytest.mark.gpu
def test_device_closed_form_many_cells(ctx):
    from openpbe_b200 import engine
    rng = np.random.default_rng(11)
    nc = 5003
    m, eps, rho, nu = random_cells(rng, nc)
    mom = engine.MOM(ctx, dev_spec("const_binary"))
    S = mom.source(T(m), T(eps), T(rho), T(nu)).cpu().numpy()
    want = np.empty_like(S)
    for c in range(nc):
        coal, br = mo.closed_form_constant_binary(m[:, c], 0.1)
        want[:, c] = coal + br
    assert np.all(np.abs(S - want) <= 1e-10 * scale(want))
    mom.close()


@pytest.mark.gpu
def test_device_step_scrub_clamp_diameter(ctx):
    from openpbe_b200 import engine
    rng = np.random.default_rng(13)
    nc = 64
    m, eps, rho, nu = random_cells(rng, nc)
    m[2, 5] = 0.5 * m[1, 5] ** 2 / m[0, 5]  # unrealisable: k < 0 -> NaN sources (Boost domain error)
    stride = 80                             # padded rows through the C ABI
    buf = torch.full((3, stride), -7.0, dtype=torch.float64, device="cuda")
    buf[:, :nc] = T(m)
    mv = buf[:, :nc]
    mom = engine.MOM(ctx, dev_spec("const_binary"))
    S0 = mom.source(T(m), T(eps), T(rho), T(nu)).cpu().numpy()
    assert np.all(np.isnan(S0[:, 5])) and not np.any(np.isnan(np.delete(S0, 5, axis=1)))
    S = torch.empty(3, stride, dtype=torch.float64, device="cuda")
    d = torch.empty(nc, dtype=torch.float64, device="cuda")
    dt = 2.0  # large enough to push some moments below SMALL
    from openpbe_b200 import capi
    eps_d, rho_d, nu_d = T(eps), T(rho), T(nu)
    capi.check(capi.lib().pbe_mom_step(mom._h, nc, stride, mv.data_ptr(), eps_d.data_ptr(), rho_d.data_ptr(),
                                       nu_d.data_ptr(), dt, S.data_ptr(), d.data_ptr(), None, None))
    torch.cuda.synchronize()
    got, Sd, dd = buf.cpu().numpy(), S.cpu().numpy()[:, :nc], d.cpu().numpy()
    assert np.all(got[:, nc:] == -7.0)
    clamped = 0
    for c in range(nc):
        Sc, mc, _ = mo.mom_step_cell(S0[:, c], m[:, c], dt)
        np.testing.assert_array_equal(Sd[:, c], Sc)
        assert np.all(np.abs(got[:, c] - mc) <= 4e-16 * (np.abs(m[:, c]) + np.abs(dt * Sc)))
        clamped += int(np.sum(got[:, c] == mo.SMALL))
        np.testing.assert_allclose(dd[c], (6 / np.pi * got[1, c] / got[0, c]) ** (1 / 3), rtol=1e-14)
    assert clamped > 0
    assert np.all(got[:, :nc] >= mo.SMALL)
    np.testing.assert_array_equal(got[:, 5], m[:, 5])  # scrubbed sources leave the cell alone
    mom.close()


@pytest.mark.gpu
def test_device_empty_input_and_errors(ctx):
    from openpbe_b200 import capi, engine
    mom = engine.MOM(ctx, dev_spec("ct_ct"))
    e = torch.empty(3, 0, dtype=torch.float64, device="cuda")
    z = torch.empty(0, dtype=torch.float64, device="cuda")
    assert mom.source(e, z, z, z).shape == (3, 0)
    mom.step(e, z, z, z, 1e-3)
    with pytest.raises(capi.PbeError, match="panels"):
        engine.MOM(ctx, dev_spec("ct_ct"), 99)
    with pytest.raises(capi.PbeError, match="not available for methodOfMoments"):
        engine.MOM(ctx, engine.make_spec("table", (), "noBreakup", (), coal_table=np.zeros((4, 4))))
    m = torch.ones(3, 8, dtype=torch.float64, device="cuda")
    with pytest.raises(capi.PbeError, match="NULL field pointer"):
        capi.check(capi.lib().pbe_mom_source(mom._h, 8, 8, m.data_ptr(), None, None, None, m.data_ptr(), None, None))
    mom.close()


@pytest.mark.gpu
def test_device_coverage_diagnostic_and_wider_range(ctx):
    """the rule is fixed where QAGIU adapts, so every call checks that it gives back the
    density's own m0 and m1.  (a) Shapes k in (0.3, 20) -- below 1 the integrable singularity at
    x = 0 the origin panel exists for -- and mean volumes from 1e-5 to 30 are covered and meet the closed
    forms (which involve m4); (b) densities at SI drop volumes (1e-10) or far beyond the last panel are
    COUNTED, not silently integrated wrong."""
    from openpbe_b200 import engine
    rng = np.random.default_rng(17)
    nc = 600
    m0 = rng.uniform(0.5, 3.0, nc)
    k = np.concatenate([rng.uniform(0.3, 1.0, nc // 2), rng.uniform(1.0, 20.0, nc - nc // 2)])
    mean = 10.0 ** rng.uniform(-5.0, np.log10(30.0), nc)  # k theta
    theta = mean / k
    m = np.stack([m0, m0 * k * theta, m0 * k * (k + 1) * theta ** 2])
    one = np.ones(nc)
    mom = engine.MOM(ctx, dev_spec("const_binary"))
    mom.coverage(reset=True)
    S = mom.source(T(m), T(one), T(one), T(one)).cpu().numpy()
    bad, worst = mom.coverage()
    assert bad == 0, (bad, worst)
    want = np.stack([np.sum(mo.closed_form_constant_binary(m[:, c], 0.1), axis=0) for c in range(nc)], axis=1)
    # m4 of the k = 0.3, mean 30 corner is the least resolved (4e-7): the documented 1e-6
    assert np.all(np.abs(S - want) <= 1e-6 * scale(want)), np.max(np.abs(S - want) / scale(want))
    # (b) out of range: SI volumes and a mean volume of 1e5
    for mean_bad in (1e-10, 1e6):
        kb = np.full(50, 2.5)
        mb = np.stack([np.ones(50), kb * (mean_bad / kb), kb * (kb + 1) * (mean_bad / kb) ** 2])
        mom.source(T(mb), T(np.ones(50)), T(np.ones(50)), T(np.ones(50)))
        bad, worst = mom.coverage()
        assert bad == 50 and worst > 1e-3, (mean_bad, bad, worst)
    assert mom.coverage() == (0, 0.0)  # reset
    mom.close()


def test_fixed_rule_coverage_map():
    """The validated range include/pbe_b200.h states for the default rule: m0 and m1 of a gamma density
    come back to 1e-6 for mean volumes in [1e-5, 30] and shapes in [0.3, 20]; outside (SI volumes) the
    deviation is large, which is what pbe_mom_coverage counts."""
    x, w = mo.fixed_rule(12)
    assert x.size == 15 * (12 + mo.GRADED_LEVELS + mo.TAIL_LEVELS) and np.all(np.diff(np.sort(x)) > 0)

    def dev(k, mean):
        f = w * mo.gamma_pdf(k, mean / k, x)
        return max(abs(f.sum() - 1.0), abs((f * x).sum() / mean - 1.0))

    for k in (0.3, 0.5, 0.9, 1.0, 1.5, 3.0, 6.0, 20.0):
        for mean in (1e-5, 1e-4, 1e-2, 1.0, 10.0, 30.0):
            assert dev(k, mean) < 1e-6, (k, mean, dev(k, mean))
    assert dev(2.5, 1e-10) > 1e-2 and dev(2.5, 1e6) > 1e-2
