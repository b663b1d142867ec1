"""GPU parity tests for the method-of-classes path (through the C ABI of libpbe_b200.so).

Oracle = oracle/pbe_oracle.c (the CPU restatement of MOC.C:87-184, pinned on the reference's
goldens).  Tolerance, FP64:
    |S_gpu - S_oracle| <= 1e-12 * scale,   scale = sum of |every product that entered S_i|
(the device evaluates each unordered class pair once, factorises the CT kernel into tables and
uses a suffix sum for breakup birth, so it re-associates; it does not change the arithmetic).
"""
import json
import math
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from openpbe_b200 import capi, engine, synthetic  # noqa: E402

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TOL = 1e-12
CT = (0.0152, 0.0678, 0.05, 0.04282)


@pytest.fixture(scope="module")
def ctx():
    assert torch.cuda.is_available(), "-m gpu tests need a CUDA device"
    c = engine.Context(0)
    yield c
    c.close()


@pytest.fixture(autouse=True, params=["smem", "tmem"])
def moc_kernel(request, monkeypatch):
    """Every test runs against BOTH device kernels: the shared-memory one and the tensor-memory
    one (pbe_moc_create honours PBE_MOC_KERNEL; where the TMEM kernel cannot serve an N the
    library keeps the shared-memory kernel, which test_kernel_selection pins down)."""
    monkeypatch.setenv("PBE_MOC_KERNEL", request.param)
    return request.param


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).cuda()


def run_pair(ctx, oracle, N, xi1, n, eps, rho, nu, dt, coal, cc, br, bc, relax=1.0, coal_table=None,
             br_table=None):
    spec = engine.make_spec(coal, cc, br, bc, coal_table=coal_table, br_table=br_table)
    ospec = oracle.make_spec(coal, cc, br, bc, coal_table=coal_table, br_table=br_table)
    moc = engine.MOC(ctx, N, xi1, spec, relax)
    S = moc.source(T(n), T(eps), T(rho), T(nu), dt).cpu().numpy()
    So, scale = oracle.moc_source(ospec, N, xi1, relax, n, eps, rho, nu, dt, nthreads=8)
    moc.close()
    return S, So, scale


def assert_parity(S, So, scale, tol=TOL):
    assert S.shape == So.shape
    assert np.array_equal(np.isnan(S), np.isnan(So))
    err = np.abs(S - So)
    bad = err > tol * scale + 1e-300
    assert not bad.any(), f"max |dS|/scale = {np.nanmax(err / (scale + 1e-300)):.3e} at {np.argwhere(bad)[:3]}"


def rand_fields(rng, N, ncells, lo=1e5, hi=1e7):
    eps = np.exp(rng.uniform(np.log(1e-3), np.log(1.0), ncells))
    return (rng.uniform(lo, hi, (N, ncells)), eps, np.full(ncells, 972.0), np.full(ncells, 1e-3))


# ---- class counts: every tiling edge (N odd/even, < tile, 128- and 64-cell CTAs) ---------------
@pytest.mark.parametrize("N", [1, 2, 3, 4, 5, 7, 8, 9, 10, 16, 20, 30, 33, 40, 50, 64, 80, 100, 101, 104, 105,
                               128, 160])
def test_source_parity_ct_all_class_counts(ctx, oracle, N):
    rng = np.random.default_rng(100 + N)
    ncells = 133 if N <= 50 else 67
    n, eps, rho, nu = rand_fields(rng, N, ncells)
    S, So, scale = run_pair(ctx, oracle, N, 3e-10 / N, n, eps, rho, nu, 2.5e-4,
                            "CoulaloglouTavlaridesC", CT, "CoulaloglouTavlarides", CT)
    assert_parity(S, So, scale)


@pytest.mark.parametrize("coal,cc,br,bc", [
    ("noCoalescence", (), "noBreakup", ()),
    ("constantCoalescence", (0.1,), "noBreakup", ()),
    ("noCoalescence", (), "binaryBreakup", ()),
    ("constantCoalescence", (0.1,), "binaryBreakup", ()),
    ("CoulaloglouTavlaridesC", CT, "noBreakup", ()),
    ("noCoalescence", (), "CoulaloglouTavlarides", CT),
    ("constantCoalescence", (5.0,), "CoulaloglouTavlarides", CT),
    ("CoulaloglouTavlaridesC", CT, "binaryBreakup", ()),
])
@pytest.mark.parametrize("N", [12, 31])
def test_source_parity_kernel_combinations(ctx, oracle, N, coal, cc, br, bc):
    rng = np.random.default_rng(7)
    cell_indep = coal != "CoulaloglouTavlaridesC" and br != "CoulaloglouTavlarides"
    n, eps, rho, nu = rand_fields(rng, N, 200, *( (0.0, 1.0) if cell_indep else (1e5, 1e7)))
    xi1 = 0.25 if cell_indep else 1e-11
    S, So, scale = run_pair(ctx, oracle, N, xi1, n, eps, rho, nu, 1e-3, coal, cc, br, bc)
    assert_parity(S, So, scale)


# ---- cells far outside the synthetic range: efficiencies from 1 down to below 2^-1022, the
# exponent clamps of the tensor-memory kernel (negK_floor, negKb_floor) and its 2^k floor ----------
@pytest.mark.parametrize("N", [30, 100])
def test_source_parity_extreme_cells(ctx, oracle, N, moc_kernel):
    rng = np.random.default_rng(N)
    ncells = 160
    n = rng.uniform(1e5, 1e7, (N, ncells))
    eps = 10.0 ** rng.uniform(-12.0, 12.0, ncells)        # exp(-K H): K H from ~1e-20 to ~1e40
    rho = 10.0 ** rng.uniform(-2.0, 5.0, ncells)
    nu = 10.0 ** rng.uniform(-9.0, 0.0, ncells)
    eps[:4] = [1e-300, 1e300, 1e-30, 1e30]
    S, So, scale = run_pair(ctx, oracle, N, 3e-10 / N, n, eps, rho, nu, 2.5e-4,
                            "CoulaloglouTavlaridesC", CT, "CoulaloglouTavlarides", CT)
    assert np.isfinite(So).all()
    assert_parity(S, So, scale)


# ---- ragged / empty cell counts, padded stride -------------------------------------------------
@pytest.mark.parametrize("ncells", [1, 2, 31, 32, 33, 127, 128, 129, 255, 256, 257, 1000, 148 * 128 + 5, 148 * 256 + 5])
def test_source_parity_ragged_cells(ctx, oracle, ncells):
    rng = np.random.default_rng(ncells)
    N = 30
    n, eps, rho, nu = rand_fields(rng, N, ncells)
    S, So, scale = run_pair(ctx, oracle, N, 1e-11, n, eps, rho, nu, 2.5e-4,
                            "CoulaloglouTavlaridesC", CT, "CoulaloglouTavlarides", CT)
    assert_parity(S, So, scale)


def test_empty_input_is_a_noop(ctx):
    spec = engine.make_spec("constantCoalescence", (0.1,), "binaryBreakup")
    moc = engine.MOC(ctx, 10, 0.1, spec)
    n = torch.empty((10, 0), dtype=torch.float64, device="cuda")
    e = torch.empty(0, dtype=torch.float64, device="cuda")
    before = ctx.launch_count()
    S = moc.source(n, e, e, e, 1e-3)
    assert S.shape == (10, 0) and ctx.launch_count() == before


def test_padded_stride_leaves_padding_untouched(ctx, oracle):
    rng = np.random.default_rng(3)
    N, ncells, stride = 20, 77, 96
    n, eps, rho, nu = rand_fields(rng, N, ncells)
    spec = engine.make_spec("CoulaloglouTavlaridesC", CT, "CoulaloglouTavlarides", CT)
    moc = engine.MOC(ctx, N, 1e-11, spec)
    nbuf = torch.full((N, stride), float("nan"), dtype=torch.float64, device="cuda")
    nbuf[:, :ncells] = T(n)
    Sbuf = torch.full((N, stride), -7.0, dtype=torch.float64, device="cuda")
    nv, Sv = nbuf[:, :ncells], Sbuf[:, :ncells]
    te, tr, tv = T(eps), T(rho), T(nu)
    torch.cuda.synchronize()
    capi.check(capi.lib().pbe_moc_source(moc._h, ncells, stride, nv.data_ptr(), te.data_ptr(),
                                         tr.data_ptr(), tv.data_ptr(), 2.5e-4, Sv.data_ptr(), None))
    torch.cuda.synchronize()
    assert torch.all(Sbuf[:, ncells:] == -7.0)
    ospec = oracle.make_spec("CoulaloglouTavlaridesC", CT, "CoulaloglouTavlarides", CT)
    So, scale = oracle.moc_source(ospec, N, 1e-11, 1.0, n, eps, rho, nu, 2.5e-4)
    assert_parity(Sbuf[:, :ncells].cpu().numpy(), So, scale)


def test_kernel_selection(ctx, moc_kernel, monkeypatch):
    spec = engine.make_spec("CoulaloglouTavlaridesC", CT, "CoulaloglouTavlarides", CT)
    m = engine.MOC(ctx, 100, 3e-12, spec)
    assert m.kernel()["kernel"] == moc_kernel           # both kernels serve N = 100
    assert m.kernel()["ctas_per_sm"] == (2 if moc_kernel == "tmem" else 1)   # 8 vs 4 warps per SM
    big = engine.MOC(ctx, 160, 3e-12, spec)              # accumulators exceed 256 TMEM columns
    assert big.kernel()["kernel"] == "smem"
    with pytest.raises(capi.PbeError) as ei:
        big.set_kernel("tmem")
    assert ei.value.code == capi.PBE_E_UNSUPPORTED
    monkeypatch.delenv("PBE_MOC_KERNEL")
    assert engine.MOC(ctx, 100, 3e-12, spec).kernel()["kernel"] == "tmem"   # auto: smem-capped N
    k30 = engine.MOC(ctx, 30, 1e-11, spec).kernel()          # auto: whichever keeps more warps per SM
    assert k30["kernel"] in ("smem", "tmem") and k30["ctas_per_sm"] >= 2
    m.set_kernel("smem")
    assert m.kernel() == {"kernel": "smem", "cells_per_cta": 128, "ctas_per_sm": 1}


# ---- the limiter (MOC realisability, MOC.C:87-91) and the fused step ---------------------------
def test_limiter_and_nonnegativity(ctx, oracle):
    rng = np.random.default_rng(13)
    N, ncells, xi1, dt = 8, 300, 0.125, 50.0  # absurd dt: the limiter must engage
    n, eps, rho, nu = rand_fields(rng, N, ncells, 0.0, 1.0)
    S, So, scale = run_pair(ctx, oracle, N, xi1, n, eps, rho, nu, dt, "constantCoalescence", (0.1,),
                            "binaryBreakup", ())
    assert_parity(S, So, scale)
    lim = -n / dt
    assert np.any(So == lim)
    assert np.array_equal(S == lim, So == lim)   # same cells limited
    # n + dt*(-n/dt) is 0 up to one rounding of n (the reference's own update has the same ulp)
    assert np.all(n + dt * S >= -4e-16 * np.abs(n))


@pytest.mark.parametrize("relax", [1.0, 0.5])
def test_step_fused_outputs(ctx, oracle, relax):
    w = synthetic.workload("moc30_ct", 777)
    f = synthetic.moc_fields(w)
    spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    ospec = oracle.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    moc = engine.MOC(ctx, w.N, w.xi1, spec, relax)
    dt = 0.05  # large enough that some classes hit the limiter
    n = T(f["n"])
    S = torch.empty_like(n)
    d = torch.empty(777, dtype=torch.float64, device="cuda")
    mom = torch.zeros(4, dtype=torch.float64, device="cuda")
    cerr = torch.zeros(1, dtype=torch.float64, device="cuda")
    moc.step(n, T(f["eps"]), T(f["rho"]), T(f["nu"]), dt, S=S, d=d, alpha=T(f["alpha"]), vcell=T(f["vcell"]),
             moments=mom, cont_err_max=cerr)
    So, scale = oracle.moc_source(ospec, w.N, w.xi1, relax, f["n"], f["eps"], f["rho"], f["nu"], dt)
    assert_parity(S.cpu().numpy(), So, scale)
    n1 = f["n"].copy()
    oracle.moc_update(n1, So, dt)
    got = n.cpu().numpy()
    assert np.all(np.abs(got - n1) <= dt * TOL * scale + 1e-15 * np.abs(n1))
    assert np.all(got >= (1 - relax) * f["n"] * (1 - 1e-12) - 1e-300)      # realisability
    np.testing.assert_allclose(d.cpu().numpy(), oracle.moc_d(n1, w.xi1), rtol=1e-12)
    xi = w.xi1 * np.arange(1, w.N + 1)
    ref_m = [np.sum(f["vcell"] * np.sum(xi[:, None] ** k * n1, axis=0)) for k in range(4)]
    np.testing.assert_allclose(mom.cpu().numpy(), ref_m, rtol=1e-11)
    mx, _ = oracle.moc_continuity(n1, w.xi1, f["alpha"], f["vcell"])
    assert float(cerr) == pytest.approx(mx, rel=1e-10)
    # stand-alone d and per-class min/max entry points
    np.testing.assert_allclose(moc.d(n).cpu().numpy(), oracle.moc_d(n1, w.xi1), rtol=1e-12)
    mn, mx2 = moc.class_minmax(n)
    np.testing.assert_array_equal(mn.cpu().numpy(), got.min(axis=1))
    np.testing.assert_array_equal(mx2.cpu().numpy(), got.max(axis=1))


# ---- user kernels through the table plug-in path -------------------------------------------------
def test_table_kernels_symmetric_and_asymmetric(ctx, oracle):
    rng = np.random.default_rng(21)
    N, ncells = 23, 150
    n, eps, rho, nu = rand_fields(rng, N, ncells, 0.0, 2.0)
    xi = 0.2 * np.arange(1, N + 1)
    sym = np.add.outer(np.cbrt(xi), np.cbrt(xi)) ** 2       # a sum-type kernel, symmetric
    asym = sym * (1.0 + 0.3 * np.subtract.outer(xi, xi) / xi.max())
    g = 0.7 * xi ** 1.5
    for tab in (sym, asym):
        S, So, scale = run_pair(ctx, oracle, N, 0.2, n, eps, rho, nu, 1e-3, "table", (), "table", (),
                                coal_table=np.ascontiguousarray(tab), br_table=g)
        assert_parity(S, So, scale)


def test_constant_kernel_equals_constant_table(ctx, oracle):
    rng = np.random.default_rng(22)
    N = 17
    n, eps, rho, nu = rand_fields(rng, N, 64, 0.0, 2.0)
    a = run_pair(ctx, oracle, N, 0.3, n, eps, rho, nu, 1e-3, "constantCoalescence", (0.37,), "binaryBreakup", ())[0]
    xi = 0.3 * np.arange(1, N + 1)
    b = run_pair(ctx, oracle, N, 0.3, n, eps, rho, nu, 1e-3, "table", (), "table", (),
                 coal_table=np.full((N, N), 0.37), br_table=xi * xi)[0]
    np.testing.assert_array_equal(a, b)


# ---- committed golden fixtures ---------------------------------------------------------------
@pytest.mark.parametrize("name", ["moc30_ct_257", "moc100_ct_64", "moc50_c01_96", "moc100_const_40"])
def test_golden_fixture(ctx, name):
    z = np.load(os.path.join(GOLD, name + ".npz"))
    w = synthetic.workload(str(z["workload"]))
    spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    moc = engine.MOC(ctx, int(z["N"]), float(z["xi1"]), spec, float(z["relax"]))
    n = T(z["n"])
    e, r, v = T(z["eps"]), T(z["rho"]), T(z["nu"])
    S = moc.source(n, e, r, v, float(z["dt"])).cpu().numpy()
    assert_parity(S, z["S"], z["scale"])
    d = torch.empty(n.shape[1], dtype=torch.float64, device="cuda")
    moc.step(n, e, r, v, float(z["dt"]), d=d)
    assert np.all(np.abs(n.cpu().numpy() - z["n_after"]) <= float(z["dt"]) * TOL * z["scale"] + 1e-15 * z["n_after"])
    np.testing.assert_allclose(d.cpu().numpy(), z["d_after"], rtol=1e-12)


# ---- scalar functors on the device vs the reference's unit-test goldens --------------------------
def test_reference_unit_test_goldens_on_device(ctx):
    g = json.load(open(os.path.join(GOLD, "reference_unit_tests.json")))
    mk = engine.make_spec
    assert ctx.eval_breakup(mk(breakup="binaryBreakup"), g["G1_binary_breakup"]["xi"], 1.0, 1.0)[0] == 36.0
    assert ctx.eval_breakup(mk(breakup="noBreakup"), 6.0, 1.0, 1.0)[0] == 0.0
    g3 = g["G3_ct_breakup"]
    r = ctx.eval_breakup(mk(breakup="CoulaloglouTavlarides", br_coeffs=g3["coeffs"]), g3["xi"], g3["rho_d"], g3["eps"])[0]
    assert r == pytest.approx(g3["expected"], rel=1e-13)
    assert ctx.eval_coalescence(mk("constantCoalescence", (5.0,)), 1.0, 2.0, 0.1, 1e3, 1e-6)[0] == 5.0
    assert ctx.eval_coalescence(mk("noCoalescence"), 1.0, 2.0, 0.1, 1e3, 1e-6)[0] == 0.0
    g6 = g["G6_ct_coalescence"]
    r = ctx.eval_coalescence(mk("CoulaloglouTavlaridesC", g6["coeffs"]), g6["xi1"], g6["xi2"], g6["eps"], g6["rho"], g6["nu"])[0]
    assert r == pytest.approx(g6["head_expected"], rel=1e-13)
    assert r / g6["V"] == pytest.approx(g6["file_expected"], rel=1e-13)   # the file's stale "/ Vcell" literal
    g9 = g["G9_integrate_exp"]
    # the reference's own assertion: ASSERT_DOUBLE_EQ(integrate(exp(-x), 0), 1.0), i.e. 4 ulp
    # (tests/unit_tests/integration.cpp:6-13); GSL's request is only epsabs 1e-7 (Integrator.H:21-22)
    got = ctx.eval_integrate_exp(1.0, g9["lower"])[0]
    assert abs(got - g9["expected"]) <= 4 * np.spacing(g9["expected"]), got - g9["expected"]
    # the family exp(-rate x) from other lower limits: closed form exp(-rate lower) / rate
    rate = np.array([1.0, 0.5, 3.0, 10.0, 1.0, 2.0])
    lower = np.array([2.0, 0.0, 0.25, 0.0, 10.0, 1.5])
    got = ctx.eval_integrate_exp(rate, lower)
    assert np.allclose(got, np.exp(-rate * lower) / rate, rtol=1e-13, atol=0.0), got * rate / np.exp(-rate * lower) - 1
    assert list(ctx.eval_daughter_beta([1.0, 4.0, 5.0], 4.0)) == [0.5, 0.5, 0.0]   # pos(0) = 1
    assert list(ctx.eval_daughter_moment([3.0, 3.0, 0.0], [2, 0, 0])) == [6.0, 2.0, 2.0]


def test_functor_batch_vs_oracle(ctx, oracle):
    rng = np.random.default_rng(5)
    k = 4000
    xi1 = 10 ** rng.uniform(-13, -9, k)
    xi2 = 10 ** rng.uniform(-13, -9, k)
    eps = 10 ** rng.uniform(-3, 0, k)
    got = ctx.eval_coalescence(engine.make_spec("CoulaloglouTavlaridesC", CT), xi1, xi2, eps, 972.0, 1e-3)
    L = oracle.lib()
    ref = np.array([L.orc_coal_ct(*CT, xi1[i], xi2[i], eps[i], 972.0, 1e-3, 1.0, 0) for i in range(k)])
    ok = ref > 1e-300
    tol = 1e-13 + 1e-15 * np.abs(np.log(ref[ok]))
    assert np.all(np.abs(got[ok] - ref[ok]) <= tol * ref[ok]), float(np.max(np.abs(got[ok] - ref[ok]) / ref[ok]))
    assert np.all(got[~ok] < 1e-299)
    gotb = ctx.eval_breakup(engine.make_spec(breakup="CoulaloglouTavlarides", br_coeffs=CT), xi1, 972.0, eps)
    refb = np.array([L.orc_break_ct(*CT, xi1[i], 972.0, eps[i]) for i in range(k)])
    # exp(-x) turns an argument error of a few ulp (cbrt chains vs pow) into |x| times that:
    # the relative tolerance grows with |ln g|
    okb = refb > 1e-300
    tolb = 1e-13 + 1e-15 * np.abs(np.log(refb[okb]))
    assert np.all(np.abs(gotb[okb] - refb[okb]) <= tolb * refb[okb]), \
        float(np.max(np.abs(gotb[okb] - refb[okb]) / refb[okb]))
    assert np.all(gotb[~okb] < 1e-299)


# ---- error behaviour ---------------------------------------------------------------------------
def test_create_errors(ctx):
    spec = engine.make_spec("constantCoalescence", (0.1,), "binaryBreakup")
    with pytest.raises(capi.PbeError) as e:
        engine.MOC(ctx, 0, 0.1, spec)
    assert e.value.code == capi.PBE_E_INVALID
    with pytest.raises(capi.PbeError) as e:
        engine.MOC(ctx, 5000, 0.1, spec)
    assert e.value.code == capi.PBE_E_UNSUPPORTED
    bad = engine.make_spec("constantCoalescence", (0.1,), "binaryBreakup")
    bad.coalescence = 17
    with pytest.raises(capi.PbeError) as e:
        engine.MOC(ctx, 10, 0.1, bad)
    assert e.value.code == capi.PBE_E_UNKNOWN_TYPE
    with pytest.raises(capi.PbeError):
        engine.make_spec("Madsen")
    moc = engine.MOC(ctx, 10, 0.1, spec)
    n = torch.ones((10, 8), dtype=torch.float64, device="cuda")
    e1 = torch.ones(8, dtype=torch.float64, device="cuda")
    with pytest.raises(capi.PbeError):
        moc.source(n, e1, e1, e1, 0.0)   # dt must be positive


# ---- host-buffer entry point (what the OpenFOAM adapter calls) ----------------------------------
def test_source_host_matches_device_entry(ctx, oracle):
    w = synthetic.workload("moc30_ct", 50_001)
    f = synthetic.moc_fields(w)
    spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    moc = engine.MOC(ctx, w.N, w.xi1, spec)
    S_host = moc.source_host(f["n"], f["eps"], f["rho"], f["nu"], w.dt).numpy()
    S_dev = moc.source(T(f["n"]), T(f["eps"]), T(f["rho"]), T(f["nu"]), w.dt).cpu().numpy()
    np.testing.assert_array_equal(S_host, S_dev)
    sl = slice(0, 3000)
    ospec = oracle.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    So, scale = oracle.moc_source(ospec, w.N, w.xi1, 1.0, np.ascontiguousarray(f["n"][:, sl]), f["eps"][sl],
                                  f["rho"][sl], f["nu"][sl], w.dt, nthreads=8)
    assert_parity(S_host[:, sl], So, scale)
