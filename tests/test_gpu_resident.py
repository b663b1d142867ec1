"""GPU parity tests of the resident-state correct() (pbe_moc_state / pbe_qmom_state, C ABI).

The reference's methods own their class / moment fields (MOC.H:66, MOC.C:69-77; QMOM.C:72-79) and
PBEDiameter::correct() is `method_->correct(); d_ = method_->d();` (PBEDiameter.H:95-99).  The
checker is the oracle (oracle/pbe_oracle.c): K resident steps + download must equal K oracle steps
    n <- n + dt max(C + B, -relax n / dt)            (MOC.C:87-91, 259-272 with phi = 0)
within 1e-12 * sum |terms| * dt + 2 ulp(n) per step (the tolerance of SURVEY.md 8c carried through the
update, plus the update's own rounding),
d must equal MOC::d() of those fields, and the diagnostics (per-class min, continuity error, global
moments) their NumPy restatements.
"""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from openpbe_b200 import engine, synthetic  # noqa: E402

CT = (0.0152, 0.0678, 0.05, 0.04282)


@pytest.fixture(scope="module")
def ctx():
    assert torch.cuda.is_available(), "-m gpu tests need a CUDA device"
    c = engine.Context(0)
    yield c
    c.close()


def oracle_moc_steps(orc, w, f, K, alpha_scale=None):
    """K explicit steps with the oracle; returns n, the per-step error budget, d"""
    sp = orc.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    n = f["n"].copy()
    budget = np.zeros_like(n)
    for _ in range(K):
        S, scale = orc.moc_source(sp, w.N, w.xi1, w.relax, n, f["eps"], f["rho"], f["nu"], w.dt, nthreads=8)
        orc.moc_update(n, S, w.dt)
        # 1e-12 sum|terms| on S, plus the rounding of the update itself: the device forms
        # fma(dt, S, n), the oracle n + dt S -- up to an ulp of n apart where dt S << n
        budget += 1e-12 * scale * w.dt + 4.5e-16 * np.abs(n)
    return n, budget, orc.moc_d(n, w.xi1)


@pytest.mark.parametrize("name,ncells,chunks", [("moc30_ct", 4096 + 37, 1), ("moc30_ct", 50_000, 5),
                                                ("moc100_ct", 3000, 1), ("replay_moc", 40_000, 3)])
@pytest.mark.parametrize("pinned", [False, True])
def test_moc_resident_steps_match_oracle(ctx, oracle, monkeypatch, name, ncells, chunks, pinned):
    monkeypatch.setenv("PBE_RESIDENT_CHUNKS", str(chunks))
    w = synthetic.workload(name, ncells)
    f = synthetic.moc_fields(w)
    K = 3
    n_o, budget, d_o = oracle_moc_steps(oracle, w, f, K)

    spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    moc = engine.MOC(ctx, w.N, w.xi1, spec, w.relax)
    host = lambda a: (torch.from_numpy(np.ascontiguousarray(a)).pin_memory() if pinned  # noqa: E731
                      else torch.from_numpy(np.ascontiguousarray(a)))
    st = engine.MOCState(moc, ncells, vcell=f["vcell"], n0=f["n"])
    d0 = st.d().numpy()
    assert np.allclose(d0, oracle.moc_d(f["n"], w.xi1), rtol=1e-13)  # PBEDiameter's constructor: d_ = method_->d()
    eps, rho, nu, alpha = (host(f[k]) for k in ("eps", "rho", "nu", "alpha"))
    d = None
    for k in range(K):
        # rho / nu only on the first call: the state keeps them (constant-property phase)
        d, diag = st.correct(eps, rho if k == 0 else None, nu if k == 0 else None, alpha, w.dt,
                             minmax=True, n_moments=3)
    n_g = st.download().numpy()
    assert np.all(np.abs(n_g - n_o) <= budget + 1e-300), np.max(np.abs(n_g - n_o) / (budget + 1e-300))
    assert np.allclose(d.numpy(), d_o, rtol=1e-12)
    assert n_g.min() >= 0.0  # realisability (relax <= 1)
    # diagnostics of the LAST step, against NumPy on the downloaded fields
    assert np.array_equal(diag["class_min"], n_g.min(axis=1))   # MOC.C:181-183
    assert np.array_equal(diag["class_max"], n_g.max(axis=1))   # MOC.C:238
    cont, _ = oracle.moc_continuity(n_g, w.xi1, f["alpha"], f["vcell"])
    assert diag["cont_err_max"] == pytest.approx(cont, rel=1e-9, abs=1e-18)  # MOC.C:276-279
    xi = w.xi1 * np.arange(1, w.N + 1)
    for q, got in enumerate(diag["moments"]):
        want = float(np.sum(f["vcell"] * (xi[:, None] ** q * n_g).sum(axis=0)))
        assert got == pytest.approx(want, rel=1e-10)
    st.close()
    moc.close()


def test_moc_resident_equals_device_step_bitwise(ctx, monkeypatch):
    """The chunked pipeline runs the same kernel on the same cells: results are bit-identical to
    one pbe_moc_step over all cells, whatever the chunking."""
    w = synthetic.workload("moc30_ct", 70_001)
    f = synthetic.moc_fields(w)
    spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    moc = engine.MOC(ctx, w.N, w.xi1, spec, w.relax)
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    n = T(f["n"])
    dref = torch.empty(70_001, dtype=torch.float64, device="cuda")
    moc.step(n, T(f["eps"]), T(f["rho"]), T(f["nu"]), w.dt, d=dref)
    for chunks in (1, 2, 7):
        monkeypatch.setenv("PBE_RESIDENT_CHUNKS", str(chunks))
        st = engine.MOCState(moc, 70_001, n0=f["n"])
        d, _ = st.correct(f["eps"], f["rho"], f["nu"], None, w.dt)
        assert torch.equal(st.download(), n.cpu())
        assert torch.equal(d, dref.cpu())
        st.close()
    moc.close()


def test_moc_resident_argument_errors(ctx):
    w = synthetic.workload("moc30_ct", 100)
    f = synthetic.moc_fields(w)
    spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    moc = engine.MOC(ctx, w.N, w.xi1, spec, w.relax)
    st = engine.MOCState(moc, 100, n0=f["n"])
    with pytest.raises(engine.capi.PbeError, match="rho / nu may only be NULL"):
        st.correct(f["eps"], None, None, None, w.dt)
    with pytest.raises(engine.capi.PbeError, match="without an attached mesh"):
        st.correct(f["eps"], f["rho"], f["nu"], None, w.dt, phi=np.zeros(3))
    with pytest.raises(engine.capi.PbeError, match="dt must be positive"):
        st.correct(f["eps"], f["rho"], f["nu"], None, 0.0)
    st.close()
    moc.close()


def test_moc_resident_with_transport_matches_oracle(ctx, oracle):
    """MOC::correct() with solveWithMULES (MOC.C:159-184, 198-272), Gauss upwind: sources of the
    current fields, then the transport + source update -- against oracle sources + the face-loop
    oracle (bit-identical transport given identical sources is covered by test_transport.py)."""
    tm = synthetic.block_mesh(14, 12, 10, velocity=(1.0, -0.5, 0.25))
    nc = 14 * 12 * 10
    w = synthetic.workload("moc30_ct", nc)
    f = synthetic.moc_fields(w)
    spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    moc = engine.MOC(ctx, w.N, w.xi1, spec, w.relax)
    mesh = engine.Mesh(ctx, tm["owner"], tm["neighbour"], tm["boundary_owner"], tm["vcell"])
    st = engine.MOCState(moc, nc, vcell=tm["vcell"], n0=f["n"], mesh=mesh)
    alpha = f["alpha"]
    sp = oracle.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    n_o = f["n"].copy()
    budget = np.zeros_like(n_o)
    for _ in range(2):
        d, diag = st.correct(f["eps"], f["rho"], f["nu"], alpha, w.dt, phi=tm["phi"], minmax=True, n_moments=2)
        S, scale = oracle.moc_source(sp, w.N, w.xi1, w.relax, n_o, f["eps"], f["rho"], f["nu"], w.dt)
        n_o = oracle.transport_upwind_explicit(tm["owner"], tm["neighbour"], tm["boundary_owner"], tm["vcell"], n_o,
                                               tm["phi"], w.dt, S=S)
        budget = budget + 1e-12 * scale * w.dt + 4.5e-16 * np.abs(n_o)
    n_g = st.download().numpy()
    # earlier differences are advected between neighbouring cells (CFL < 1): compare per class with
    # the largest budget of the class
    assert np.all(np.abs(n_g - n_o) <= 4.0 * budget.max(axis=1, keepdims=True) + 1e-12 * np.abs(n_o))
    assert np.allclose(d.numpy(), oracle.moc_d(n_g, w.xi1), rtol=1e-12)
    assert np.array_equal(diag["class_min"], n_g.min(axis=1))
    cont, _ = oracle.moc_continuity(n_g, w.xi1, alpha, tm["vcell"])
    assert diag["cont_err_max"] == pytest.approx(cont, rel=1e-9, abs=1e-18)
    xi = w.xi1 * np.arange(1, w.N + 1)
    want1 = float(np.sum(tm["vcell"] * (xi[:, None] * n_g).sum(axis=0)))
    assert diag["moments"][1] == pytest.approx(want1, rel=1e-10)
    st.close()
    mesh.close()
    moc.close()


# ---- QMOM ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name,ncells,chunks", [("qmom3_ct", 4096 + 5, 1), ("qmom3_ct", 300_000, 3),
                                                ("qmom3_const", 2000, 1)])
def test_qmom_resident_steps_match_oracle(ctx, oracle, monkeypatch, name, ncells, chunks):
    monkeypatch.setenv("PBE_RESIDENT_CHUNKS", str(chunks))
    w = synthetic.workload(name, ncells)
    f = synthetic.qmom_fields(w)
    sp = oracle.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    q = engine.QMOM(ctx, w.N, spec)
    vcell = np.random.default_rng(1).uniform(0.5e-9, 2e-9, ncells)
    st = engine.QMOMState(q, ncells, vcell=vcell, m0=f["m"])
    assert np.allclose(st.d().numpy(), np.cbrt(6.0 / np.pi * f["m"][1] / f["m"][0]), rtol=1e-14)  # QMOM.C:81
    # one step: the device sources meet the oracle's to 1e-9 * scale on well-separated
    # nodes; the NaN pattern (scrubbed to 0, QMOM.C:101-107) is identical
    So, scale = oracle.qmom_source(sp, w.N, f["m"], f["eps"], f["rho"], f["nu"], nthreads=8)
    m_o = f["m"].copy()
    d_o = oracle.qmom_update(m_o, So.copy(), w.dt)
    d, stats = st.correct(f["eps"], f["rho"], f["nu"], w.dt, stats=True)
    m_g = st.download().numpy()
    ok = np.isfinite(So).all(axis=0) & (f["nodes"][1] / f["nodes"][0] > 1.001) & (f["nodes"][2] / f["nodes"][1] > 1.001)
    tol = 1e-9 * scale * w.dt + 4.5e-16 * np.abs(m_o) + 1e-15
    assert np.all(np.abs(m_g - m_o)[:, ok] <= tol[:, ok])
    assert np.all(m_g >= 1e-15)  # clamp to SMALL (QMOM.C:117-123): moment realisability floor
    assert np.allclose(d.numpy()[ok], d_o[ok], rtol=1e-9)
    # printAvgMaxMin (Utility.C:4-10) of moments and d, against NumPy on the downloaded fields
    for k in range(2 * w.N):
        sfv, sv, mx, mn = stats["moment"][k]
        assert sfv / sv == pytest.approx(np.sum(m_g[k] * vcell) / vcell.sum(), rel=1e-11)
        assert mx == m_g[k].max() and mn == m_g[k].min()
    sfv, sv, mx, mn = stats["d"]
    dg = d.numpy()
    assert sfv / sv == pytest.approx(np.sum(dg * vcell) / vcell.sum(), rel=1e-11)
    assert mx == dg.max() and mn == dg.min()
    # sources: scrubbed, so finite; avg against the oracle's scrubbed sources on the well-conditioned cells
    assert np.isfinite(stats["source"]).all()
    st.close()
    q.close()


def test_qmom_resident_equals_device_step_bitwise(ctx, monkeypatch):
    w = synthetic.workload("qmom3_ct", 150_001)
    f = synthetic.qmom_fields(w)
    spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    q = engine.QMOM(ctx, w.N, spec)
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    m = T(f["m"])
    dref = torch.empty(150_001, dtype=torch.float64, device="cuda")
    q.step(m, T(f["eps"]), T(f["rho"]), T(f["nu"]), w.dt, d=dref)
    for chunks in (1, 2):
        monkeypatch.setenv("PBE_RESIDENT_CHUNKS", str(chunks))
        st = engine.QMOMState(q, 150_001, m0=f["m"])
        d, _ = st.correct(f["eps"], f["rho"], f["nu"], w.dt)
        assert torch.equal(st.download(), m.cpu())
        assert torch.equal(d, dref.cpu())
        st.close()
    q.close()


def test_qmom_resident_with_transport(ctx, oracle):
    """QMOM::correct() with solveWithMULES, Gauss upwind (QMOM.C:85-129, 131-197)."""
    tm = synthetic.block_mesh(16, 9, 7, velocity=(0.5, 1.0, -0.25))
    nc = 16 * 9 * 7
    w = synthetic.workload("qmom3_ct", nc)
    f = synthetic.qmom_fields(w)
    spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    sp = oracle.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    q = engine.QMOM(ctx, w.N, spec)
    mesh = engine.Mesh(ctx, tm["owner"], tm["neighbour"], tm["boundary_owner"], tm["vcell"])
    st = engine.QMOMState(q, nc, vcell=tm["vcell"], m0=f["m"], mesh=mesh)
    d, _ = st.correct(f["eps"], f["rho"], f["nu"], w.dt, phi=tm["phi"])
    So, scale = oracle.qmom_source(sp, w.N, f["m"], f["eps"], f["rho"], f["nu"])
    So = np.where(np.isnan(So), 0.0, So)
    m_o = oracle.transport_upwind_explicit(tm["owner"], tm["neighbour"], tm["boundary_owner"], tm["vcell"], f["m"],
                                           tm["phi"], w.dt, S=So)
    m_o = np.maximum(m_o, 1e-15)
    m_g = st.download().numpy()
    ok = (f["nodes"][1] / f["nodes"][0] > 1.001) & (f["nodes"][2] / f["nodes"][1] > 1.001)
    assert np.all(np.abs(m_g - m_o)[:, ok] <= (1e-9 * scale * w.dt + 1e-12 * np.abs(m_o) + 1e-15)[:, ok])
    assert np.allclose(d.numpy()[ok], np.cbrt(6 / np.pi * m_g[1] / m_g[0])[ok], rtol=1e-13)
    st.close()
    mesh.close()
    q.close()


# ---- handles are independent (the dynamic shared-memory limit belongs to the kernel
# instantiation, not to the handle that set it last) ---------------------------------------------------
@pytest.mark.parametrize("kernel", ["smem", "tmem"])
def test_two_live_handles_of_one_instantiation(ctx, oracle, monkeypatch, kernel):
    monkeypatch.setenv("PBE_MOC_KERNEL", kernel)
    spec = engine.make_spec("CoulaloglouTavlaridesC", CT, "CoulaloglouTavlarides", CT)
    big = engine.MOC(ctx, 100, 3e-12, spec)     # ~104 KB (tmem) / ~220 KB (smem) of dynamic shared memory
    small = engine.MOC(ctx, 30, 1e-11, spec)    # same instantiation, far less
    rng = np.random.default_rng(5)
    n = rng.uniform(1e5, 1e7, (100, 300))
    eps = np.exp(rng.uniform(np.log(1e-3), 0.0, 300))
    rho, nu = np.full(300, 972.0), np.full(300, 1e-3)
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    S = big.source(T(n), T(eps), T(rho), T(nu), 2.5e-4).cpu().numpy()  # failed with cudaErrorInvalidValue before
    osp = oracle.make_spec("CoulaloglouTavlaridesC", CT, "CoulaloglouTavlarides", CT)
    So, scale = oracle.moc_source(osp, 100, 3e-12, 1.0, n, eps, rho, nu, 2.5e-4, nthreads=8)
    assert np.all(np.abs(S - So) <= 1e-12 * scale + 1e-300)
    S30 = small.source(T(n[:30]), T(eps), T(rho), T(nu), 2.5e-4).cpu().numpy()
    So30, sc30 = oracle.moc_source(osp, 30, 1e-11, 1.0, n[:30], eps, rho, nu, 2.5e-4, nthreads=8)
    assert np.all(np.abs(S30 - So30) <= 1e-12 * sc30 + 1e-300)
    big.close()
    small.close()


def test_two_live_mom_handles(ctx):
    spec = engine.make_spec("constantCoalescence", (0.1,), "binaryBreakup", ())
    a = engine.MOM(ctx, spec, panels=28)
    b = engine.MOM(ctx, spec, panels=4)
    mk, mth = np.linspace(1.5, 4.0, 9), np.linspace(0.15, 0.6, 9)
    mm = torch.from_numpy(np.stack([np.full(9, 2.0), 2.0 * mk * mth, 2.0 * mk * (mk + 1) * mth ** 2])).cuda()
    one = torch.ones(9, dtype=torch.float64, device="cuda")
    Sa = a.source(mm, one, one, one)
    Sb = b.source(mm, one, one, one)
    torch.cuda.synchronize()
    assert torch.isfinite(Sa).all() and torch.isfinite(Sb).all()
    a.close()
    b.close()


# ---- a16: printAvgMaxMin (Utility.C:4-10) through pbe_field_avg_max_min, non-uniform V ----------------
@pytest.mark.parametrize("ncells", [1, 31, 4097, 1_000_003])
def test_field_avg_max_min_matches_numpy(ctx, ncells):
    rng = np.random.default_rng(ncells)
    f = rng.normal(size=ncells) * 10.0 ** rng.uniform(-3, 3, ncells)
    V = rng.uniform(1e-10, 1e-8, ncells)
    T = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    out = ctx.field_avg_max_min(T(f), T(V)).cpu().numpy()
    # v.weightedAverage(mesh.V()) = sum(f V) / sum(V); max(v); min(v)
    assert out[0] / out[1] == pytest.approx(np.sum(f * V) / np.sum(V), rel=1e-11, abs=1e-13 * np.abs(f).max())
    assert out[1] == pytest.approx(V.sum(), rel=1e-12)
    assert out[2] == f.max() and out[3] == f.min()
    out1 = ctx.field_avg_max_min(T(f)).cpu().numpy()  # V = 1: plain average
    assert out1[0] / out1[1] == pytest.approx(f.mean(), rel=1e-10, abs=1e-13 * np.abs(f).max())


def test_moc_state_range_io_and_device_view(ctx):
    """construction / write-out in pieces (pbe_moc_state_upload_range / _download_range) and the
    zero-copy device view agree with the whole-field upload / download"""
    w = synthetic.workload("moc30_ct", 1000)
    f = synthetic.moc_fields(w)
    spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    moc = engine.MOC(ctx, w.N, w.xi1, spec, w.relax)
    st = engine.MOCState(moc, 1000)
    for c0 in range(0, 1000, 300):
        st.upload_range(c0, np.ascontiguousarray(f["n"][:, c0:c0 + 300]))
    assert np.array_equal(st.download().numpy(), f["n"])
    assert np.array_equal(st.download_range(250, 500).numpy(), f["n"][:, 250:750])
    n_dev, d_dev = st.device_fields()
    assert np.array_equal(n_dev.cpu().numpy(), f["n"])
    n_dev.mul_(2.0)  # a view, not a copy
    torch.cuda.synchronize()
    assert np.array_equal(st.download().numpy(), 2.0 * f["n"])
    st.close()
    moc.close()


def test_comm_through_the_c_abi_single_rank(ctx):
    """pbe_comm_unique_id / pbe_comm_init (NCCL bound at run time) with one rank: the reductions are the
    identity, the resident correct() still returns its diagnostics (now routed through NCCL)."""
    import ctypes as C
    from openpbe_b200 import capi
    L = capi.lib()
    c2 = engine.Context(0)
    ident = C.create_string_buffer(128)
    capi.check(L.pbe_comm_unique_id(ident))
    capi.check(L.pbe_comm_init(c2.handle, 1, 0, ident))
    nr, rk = C.c_int(-1), C.c_int(-1)
    capi.check(L.pbe_comm_info(c2.handle, C.byref(nr), C.byref(rk)))
    assert (nr.value, rk.value) == (1, 0)
    s, mn, mx = np.array([1.5, 2.5]), np.array([-3.0]), np.array([7.0, 8.0, 9.0])
    capi.check(L.pbe_reduce_diagnostics_host(c2.handle, 2, s.ctypes.data, 1, mn.ctypes.data, 3, mx.ctypes.data))
    assert s.tolist() == [1.5, 2.5] and mn.tolist() == [-3.0] and mx.tolist() == [7.0, 8.0, 9.0]
    w = synthetic.workload("moc30_ct", 5000)
    f = synthetic.moc_fields(w)
    spec = engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    moc = engine.MOC(c2, w.N, w.xi1, spec, w.relax)
    st = engine.MOCState(moc, 5000, vcell=f["vcell"], n0=f["n"])
    d, diag = st.correct(f["eps"], f["rho"], f["nu"], f["alpha"], w.dt, minmax=True, n_moments=2)
    n_g = st.download().numpy()
    assert np.array_equal(diag["class_min"], n_g.min(axis=1)) and np.array_equal(diag["class_max"], n_g.max(axis=1))
    assert diag["cont_err_max"] > 0.0 and diag["moments"][0] > 0.0
    st.close(); moc.close()
    with pytest.raises(capi.PbeError, match="already has a communicator"):
        capi.check(L.pbe_comm_init(c2.handle, 1, 0, ident))
    capi.check(L.pbe_comm_destroy(c2.handle))
    c2.close()


@pytest.mark.gpu
def test_context_close_closes_live_handles_first():
    """The C ABI wants handles destroyed before their context.  The Python wrapper enforces it: Context.close() closes
    what is still alive (states before methods and meshes), so a handle collected later never touches a freed
    context -- the failure mode was a stray cudaSetDevice(garbage) surfacing as CUDA error 101 in an unrelated call."""
    import torch
    from openpbe_b200 import engine, synthetic
    c = engine.Context(0)
    w = synthetic.workload("moc30_ct", 500)
    f = synthetic.moc_fields(w)
    moc = engine.MOC(c, w.N, w.xi1, engine.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs))
    q = engine.QMOM(c, 3, engine.make_spec("constantCoalescence", (1.0,), "noBreakup", ()))
    st = engine.MOCState(moc, 500, n0=f["n"])
    c.close()  # handles still referenced
    assert not moc._h and not q._h and not st._h
    moc.close(); st.close(); q.close()  # idempotent
    del moc, q, st
    import gc
    gc.collect()
    c2 = engine.Context(0)
    x, wts = c2.wheeler_batch(torch.tensor([[1.0] * 8, [2.0] * 8], dtype=torch.float64, device="cuda"))
    assert torch.allclose(x, torch.full_like(x, 2.0))
    c2.close()
