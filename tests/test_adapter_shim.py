"""The OpenFOAM-side adapter (adapter/PBEMethodB200.{H,C}: PBEMethod / MOC / QMOM / PBEDiameter in
OpenFOAM's idiom, the code a maintainer drops into src/PBESystem) compiled against the API shim
adapter/foam_shim/foamShim.H -- SURVEY.md 8(f) row f1.  OpenFOAM itself is not in the image, so
this is the compile check plus, on a GPU box, a run through `pbe_moc_source_host` /
`pbe_qmom_source_host` compared with the oracle.

CPU: builds with -Wall -Wextra; run-time selection and kernel selection fail like
PBEMethod::New (PBEMethod.C:97-105) and coalescenceKernel::New (coalescenceKernel.C:57-65);
construction without a device fails loudly (no CPU path)."""
import os
import subprocess

import numpy as np
import pytest

from openpbe_b200 import synthetic

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "adapter", "shimMain")

C01 = """
coalescenceKernel CoulaloglouTavlaridesC; breakupKernel CoulaloglouTavlarides;
daughterParticleDistribution uniformBinaryBreakup;
CoulaloglouTavlaridesCCoeffs { c1 %r; c2 %r; gamma %r; sigma %r; }
CoulaloglouTavlaridesCoeffs  { c1 %r; c2 %r; gamma %r; sigma %r; }
uniformBinaryBreakupCoeffs { }
"""


NOKERNELS = ("coalescenceKernel noCoalescence; breakupKernel noBreakup; daughterParticleDistribution uniformBinaryBreakup; "
             "noCoalescenceCoeffs { } noBreakupCoeffs { } uniformBinaryBreakupCoeffs { } ")


@pytest.fixture(scope="module")
def shim():
    if not os.path.exists(os.path.join(ROOT, "openpbe_b200", "lib", "libpbe_b200.so")):
        pytest.skip("libpbe_b200.so not built")
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "adapter"), "-s"])
    return EXE


def run(shim, tmp_path, dict_text, state, eps, rho, nu, alpha, dt, steps=1, vcell=1e-9):
    K, nc = state.shape
    (tmp_path / "dict").write_text(dict_text)
    np.concatenate([[K, nc, dt, vcell], state.ravel(), eps, rho, nu, alpha]).astype(np.float64).tofile(tmp_path / "in.bin")
    p = subprocess.run([shim, str(tmp_path / "dict"), str(tmp_path / "in.bin"), str(tmp_path / "out.bin"), str(steps)],
                       capture_output=True, text=True)
    out = None
    if p.returncode == 0:
        r = np.fromfile(tmp_path / "out.bin")
        out = (r[:K * nc].reshape(K, nc), r[K * nc:K * nc + nc], r[-1])
    return p, out


def test_selection_errors_read_like_openfoam(shim, tmp_path):
    one = np.ones((2, 4))
    v = np.ones(4)
    p, _ = run(shim, tmp_path, "method bogus; " + NOKERNELS, one, v, v, v, v, 1e-3)
    assert p.returncode == 1
    assert "Unknown pbeModelType type bogus" in p.stdout and "Valid PBEMethod types are" in p.stdout
    # all four methods of the reference are registered (MOC.C:38-39, QMOM.C:50-51, MOM.C:51-52, dummy.C:37-38)
    assert "(dummy methodOfClasses methodOfMoments quadratureMethodOfMoments)" in p.stdout and "FOAM exiting" in p.stdout
    p, _ = run(shim, tmp_path, "method methodOfClasses; MOCCoeffs { numberOfClasses 2; xi1 1; } "
               + NOKERNELS.replace("coalescenceKernel noCoalescence", "coalescenceKernel Madsen"), one, v, v, v, v, 1e-3)
    assert p.returncode == 1 and "Madsen" in p.stdout and "CoulaloglouTavlaridesC" in p.stdout
    p, _ = run(shim, tmp_path, "coalescenceKernel noCoalescence;", one, v, v, v, v, 1e-3)
    assert p.returncode == 1 and "keyword method is undefined in dictionary" in p.stdout
    # <Type>Coeffs is looked up with subDict: fatal when absent, as coalescenceKernel.C:66-70
    p, _ = run(shim, tmp_path, "method methodOfClasses; MOCCoeffs { numberOfClasses 2; xi1 1; } coalescenceKernel "
               "noCoalescence; breakupKernel noBreakup; daughterParticleDistribution uniformBinaryBreakup;", one, v, v, v, v, 1e-3)
    assert p.returncode == 1 and "keyword noCoalescenceCoeffs is undefined" in p.stdout


def test_no_device_is_fatal_not_a_fallback(shim, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    one = np.ones((2, 4))
    v = np.ones(4)
    p, _ = run(shim, tmp_path, "method methodOfClasses; MOCCoeffs { numberOfClasses 2; xi1 1; } " + NOKERNELS,
               one, v, v, v, v, 1e-3)
    assert p.returncode == 1 and "FOAM FATAL ERROR" in p.stdout and "no CPU path" in p.stdout


@pytest.mark.gpu
def test_moc_correct_matches_oracle(shim, tmp_path, oracle):
    w = synthetic.workload("moc30_ct", 3001)
    f = synthetic.moc_fields(w)
    text = ("method methodOfClasses; MOCCoeffs { numberOfClasses %d; xi1 %r; usingMULES on; }" % (w.N, w.xi1)
            + C01 % (tuple(w.coal_coeffs) + tuple(w.br_coeffs)))
    p, out = run(shim, tmp_path, text, f["n"], f["eps"], f["rho"], f["nu"], f["alpha"], w.dt, steps=2)
    assert p.returncode == 0, p.stdout + p.stderr
    n, d, cerr = out
    sp = oracle.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    n_o = f["n"].copy()
    for _ in range(2):
        S, scale = oracle.moc_source(sp, w.N, w.xi1, w.relax, n_o, f["eps"], f["rho"], f["nu"], w.dt)
        oracle.moc_update(n_o, S, w.dt)
    # |dS| <= 1e-12 scale per step (tests/test_gpu_moc.py) -> |dn| <= 2 dt 1e-12 scale, plus rounding of n
    assert np.all(np.abs(n - n_o) <= 2 * w.dt * 1e-12 * scale + 4e-16 * np.abs(n_o))
    np.testing.assert_allclose(d, oracle.moc_d(n_o, w.xi1), rtol=1e-12)
    assert np.isclose(cerr, oracle.moc_continuity(n_o, w.xi1, f["alpha"], f["vcell"])[0], rtol=1e-9)
    assert n.min() >= 0.0


@pytest.mark.gpu
def test_moc_geometric_grid_through_the_adapter(shim, tmp_path):
    """MOCCoeffs { classGridRatio r; } -- the fixed-pivot extension (row f4, not a reference key) selected by
    dictionary through diameterModel::New -> PBEMethod::New -> MOC, against the CPU restatement."""
    from oracle import fixed_pivot as fp
    w = synthetic.workload("moc30_ct", 2001)
    f = synthetic.moc_fields(w)
    ratio = 1.25
    x = fp.geometric_pivots(w.N, w.xi1, ratio)
    n0 = f["n"] * ((w.xi1 * (np.arange(w.N) + 1)) / x)[:, None]  # same volume per class on the new pivots
    text = ("method methodOfClasses; MOCCoeffs { numberOfClasses %d; xi1 %r; classGridRatio %r; }" % (w.N, w.xi1, ratio)
            + C01 % (tuple(w.coal_coeffs) + tuple(w.br_coeffs)))
    p, out = run(shim, tmp_path, text, n0, f["eps"], f["rho"], f["nu"], f["alpha"], w.dt, steps=1)
    assert p.returncode == 0, p.stdout + p.stderr
    n, d, _ = out
    spec = {"coalescence": w.coalescence, "coal_c": tuple(w.coal_coeffs), "breakup": w.breakup, "br_c": tuple(w.br_coeffs)}
    S, scale = fp.moc_source(spec, x, w.relax, n0, f["eps"], f["rho"], f["nu"], w.dt)
    n1 = n0 + w.dt * S
    assert np.all(np.abs(n - n1) <= w.dt * 1e-12 * scale + 4e-16 * np.abs(n1))
    np.testing.assert_allclose(d, ((6.0 / np.pi * x) ** (1.0 / 3.0) @ n1) / n1.sum(axis=0), rtol=1e-12)
    p, _ = run(shim, tmp_path, text.replace("classGridRatio %r" % ratio, "classGridRatio 0.9"), n0, f["eps"], f["rho"],
               f["nu"], f["alpha"], w.dt)
    assert p.returncode == 1 and "classGridRatio must be > 1" in p.stdout


@pytest.mark.gpu
def test_qmom_correct_matches_oracle(shim, tmp_path, oracle):
    w = synthetic.workload("qmom3_ct", 2003)
    f = synthetic.qmom_fields(w)
    text = ("method quadratureMethodOfMoments; QMOMCoeffs { quadratureOrder 3; }"
            + C01 % (tuple(w.coal_coeffs) + tuple(w.br_coeffs)))
    dt = 1e-9   # small enough that no moment reaches the SMALL clamp: the update itself is compared
    p, out = run(shim, tmp_path, text, f["m"], f["eps"], f["rho"], f["nu"], f["alpha"], dt, steps=1)
    assert p.returncode == 0, p.stdout + p.stderr
    m, d, _ = out
    sp = oracle.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    S, scale = oracle.qmom_source(sp, 3, f["m"], f["eps"], f["rho"], f["nu"])
    m_o = f["m"].copy()
    d_o = oracle.qmom_update(m_o, S.copy(), dt)
    well = (f["nodes"][1] / f["nodes"][0] > 1.001) & np.isfinite(S).all(axis=0)
    assert np.all(np.abs(m - m_o)[:, well] <= dt * 1e-9 * scale[:, well] + 4e-16 * np.abs(m_o[:, well]))
    np.testing.assert_allclose(d[well], d_o[well], rtol=1e-9)
    assert np.all(m >= 1e-15)


def _write_mesh(path, m, geometry=False):
    with open(path, "wb") as fh:
        np.array([m["owner"].size, m["boundary_owner"].size], dtype=np.float64).tofile(fh)
        for k in ("owner", "neighbour", "boundary_owner"):
            m[k].astype(np.int32).tofile(fh)
        m["phi"].astype(np.float64).tofile(fh)
        if geometry:  # fvMesh::Sf(), C(), weights() for the limited schemes
            for k in ("Sf", "C", "weights"):
                np.ascontiguousarray(m[k], dtype=np.float64).tofile(fh)


@pytest.mark.gpu
def test_moc_correct_with_transport_matches_oracle(shim, tmp_path, oracle):
    """MOC::correct() with solveWithMULES (MOC.C:198-272, Gauss upwind) through the adapter: sources
    and the explicit transport update on the device, against oracle sources + the face-loop oracle."""
    tm = synthetic.block_mesh(14, 11, 9, velocity=(1.0, -0.5, 0.25))
    nc = 14 * 11 * 9
    w = synthetic.workload("moc30_ct", nc)
    f = synthetic.moc_fields(w)
    vcell = float(tm["vcell"][0])
    text = ("method methodOfClasses; MOCCoeffs { numberOfClasses %d; xi1 %r; usingMULES on; divScheme Gauss upwind; }"
            % (w.N, w.xi1) + C01 % (tuple(w.coal_coeffs) + tuple(w.br_coeffs)))
    K = w.N
    (tmp_path / "dict").write_text(text)
    np.concatenate([[K, nc, w.dt, vcell], f["n"].ravel(), f["eps"], f["rho"], f["nu"], f["alpha"]]).tofile(tmp_path / "in.bin")
    _write_mesh(tmp_path / "mesh.bin", tm)
    p = subprocess.run([shim, str(tmp_path / "dict"), str(tmp_path / "in.bin"), str(tmp_path / "out.bin"), "2",
                        str(tmp_path / "mesh.bin")], capture_output=True, text=True)
    assert p.returncode == 0, p.stdout + p.stderr
    r = np.fromfile(tmp_path / "out.bin")
    n, d = r[:K * nc].reshape(K, nc), r[K * nc:K * nc + nc]
    sp = oracle.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    n_o = f["n"].copy()
    for _ in range(2):
        S, scale = oracle.moc_source(sp, w.N, w.xi1, w.relax, n_o, f["eps"], f["rho"], f["nu"], w.dt)
        n_o = oracle.transport_upwind_explicit(tm["owner"], tm["neighbour"], tm["boundary_owner"], tm["vcell"], n_o,
                                               tm["phi"], w.dt, S=S)
    assert np.all(np.abs(n - n_o) <= 4 * w.dt * 1e-12 * scale + 1e-14 * np.abs(n_o))
    np.testing.assert_allclose(d, oracle.moc_d(n_o, w.xi1), rtol=1e-12)
    # a scheme without a device path is a FatalError that lists the valid ones
    (tmp_path / "dict").write_text(text.replace("Gauss upwind", "Gauss QUICK"))
    p = subprocess.run([shim, str(tmp_path / "dict"), str(tmp_path / "in.bin"), str(tmp_path / "out.bin"), "1",
                        str(tmp_path / "mesh.bin")], capture_output=True, text=True)
    assert p.returncode == 1 and "QUICK" in p.stdout and "Valid schemes are" in p.stdout and "FOAM exiting" in p.stdout


@pytest.mark.gpu
def test_moc_correct_default_branch_is_the_implicit_solve(shim, tmp_path, oracle):
    """usingMULES absent = false, the reference's default (MOC.C:58): MOC::correct() through the adapter runs
    solveWithFVM on the device -- oracle sources + the direct solve of the restated fvMatrix; fvSolution's controls come
    from the dictionary in the shim build."""
    from oracle import fvm_implicit as fvm
    tm = synthetic.block_mesh(14, 11, 9, velocity=(1.0, -0.5, 0.25))
    nc = 14 * 11 * 9
    w = synthetic.workload("moc30_ct", nc)
    f = synthetic.moc_fields(w)
    vcell = float(tm["vcell"][0])
    dt = 0.02  # Courant ~0.5: far beyond what the case's 2.5e-4 would give, still cheap
    text = ("method methodOfClasses; MOCCoeffs { numberOfClasses %d; xi1 %r; tolerance 1e-13; minIter 1; }" % (w.N, w.xi1)
            + C01 % (tuple(w.coal_coeffs) + tuple(w.br_coeffs)))
    K = w.N
    (tmp_path / "dict").write_text(text)
    np.concatenate([[K, nc, dt, vcell], f["n"].ravel(), f["eps"], f["rho"], f["nu"], f["alpha"]]).tofile(tmp_path / "in.bin")
    _write_mesh(tmp_path / "mesh.bin", tm)
    p = subprocess.run([shim, str(tmp_path / "dict"), str(tmp_path / "in.bin"), str(tmp_path / "out.bin"), "2",
                        str(tmp_path / "mesh.bin")], capture_output=True, text=True)
    assert p.returncode == 0, p.stdout + p.stderr
    r = np.fromfile(tmp_path / "out.bin")
    n, d = r[:K * nc].reshape(K, nc), r[K * nc:K * nc + nc]
    sp = oracle.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    n_o = f["n"].copy()
    for _ in range(2):
        S, scale = oracle.moc_source(sp, w.N, w.xi1, w.relax, n_o, f["eps"], f["rho"], f["nu"], dt)
        n_o = fvm.solve(tm, tm["phi"], dt, n_o, S=S)
    assert np.all(np.abs(n - n_o) <= 1e-9 * np.abs(n_o).max(axis=1, keepdims=True))
    np.testing.assert_allclose(d, oracle.moc_d(n_o, w.xi1), rtol=1e-9)
    # the scheme of the reference's MOC cases for the implicit term (tests/integration/MOC/template/system/fvSchemes:44)
    (tmp_path / "dict").write_text(text.replace("minIter 1;", "minIter 1; implicitDivScheme Gauss limitedLinear 1;"))
    _write_mesh(tmp_path / "mesh.bin", tm, geometry=True)
    p = subprocess.run([shim, str(tmp_path / "dict"), str(tmp_path / "in.bin"), str(tmp_path / "out.bin"), "1",
                        str(tmp_path / "mesh.bin")], capture_output=True, text=True)
    assert p.returncode == 0, p.stdout + p.stderr
    n1 = np.fromfile(tmp_path / "out.bin")[:K * nc].reshape(K, nc)
    S, _ = oracle.moc_source(sp, w.N, w.xi1, w.relax, f["n"], f["eps"], f["rho"], f["nu"], dt)
    want = fvm.solve(tm, tm["phi"], dt, f["n"], S=S, scheme="Gauss limitedLinear 1")
    assert np.all(np.abs(n1 - want) <= 1e-9 * np.abs(want).max(axis=1, keepdims=True))
    # a scheme without a device path: FatalError that lists the valid ones, not a silent change of scheme
    (tmp_path / "dict").write_text(text.replace("minIter 1;", "minIter 1; implicitDivScheme Gauss QUICK;"))
    p = subprocess.run([shim, str(tmp_path / "dict"), str(tmp_path / "in.bin"), str(tmp_path / "out.bin"), "1",
                        str(tmp_path / "mesh.bin")], capture_output=True, text=True)
    assert p.returncode == 1 and "QUICK" in p.stdout and "Valid schemes are" in p.stdout and "FOAM exiting" in p.stdout


@pytest.mark.gpu
def test_moc_correct_with_vanleer_mules_matches_port(shim, tmp_path, oracle):
    """The scheme the reference's validation case selects (c01/system/fvSchemes:34) through the adapter:
    MOC::correct() with solveWithMULES = sources, MULES::limit against max(alpha V / xi_k), limitSum,
    explicitSolve -- against oracle sources + the port of the call sequence (oracle/pbe_oracle_mules.c)."""
    tm = synthetic.block_mesh(12, 10, 8, velocity=(1.0, -0.5, 0.25))
    nc = 12 * 10 * 8
    w = synthetic.workload("moc30_ct", nc)
    f = synthetic.moc_fields(w)
    f["n"] *= 1.0 + 5.0 * np.sin(np.arange(nc) * 0.7)[None, :] ** 2
    vcell = float(tm["vcell"][0])
    dt = 0.01
    text = ("method methodOfClasses; MOCCoeffs { numberOfClasses %d; xi1 %r; usingMULES on; divScheme Gauss vanLeer; }"
            % (w.N, w.xi1) + C01 % (tuple(w.coal_coeffs) + tuple(w.br_coeffs)))
    (tmp_path / "dict").write_text(text)
    np.concatenate([[w.N, nc, dt, vcell], f["n"].ravel(), f["eps"], f["rho"], f["nu"], f["alpha"]]).tofile(tmp_path / "in.bin")
    _write_mesh(tmp_path / "mesh.bin", tm, geometry=True)
    p = subprocess.run([shim, str(tmp_path / "dict"), str(tmp_path / "in.bin"), str(tmp_path / "out.bin"), "1",
                        str(tmp_path / "mesh.bin")], capture_output=True, text=True)
    assert p.returncode == 0, p.stdout + p.stderr
    r = np.fromfile(tmp_path / "out.bin")
    n = r[:w.N * nc].reshape(w.N, nc)
    sp = oracle.make_spec(w.coalescence, w.coal_coeffs, w.breakup, w.br_coeffs)
    S, scale = oracle.moc_source(sp, w.N, w.xi1, w.relax, f["n"], f["eps"], f["rho"], f["nu"], dt)
    xi = w.xi1 * np.arange(1, w.N + 1)
    pmax = np.array([np.max(f["alpha"] * tm["vcell"] / x) for x in xi])
    n_o = oracle.transport_mules(tm, "Gauss vanLeer", f["n"], tm["phi"], dt, pmax, S=S)
    assert np.all(np.abs(n - n_o) <= 4 * dt * 1e-12 * scale.max(axis=1, keepdims=True) + 1e-13 * np.abs(n_o))
    assert "Max continuity error between advection and PBE" in p.stdout and "min(n0)" in p.stdout
    # without the geometry the limited scheme cannot be built: FatalError, not a silent upwind
    _write_mesh(tmp_path / "mesh.bin", tm, geometry=False)
    p = subprocess.run([shim, str(tmp_path / "dict"), str(tmp_path / "in.bin"), str(tmp_path / "out.bin"), "1",
                        str(tmp_path / "mesh.bin")], capture_output=True, text=True)
    assert p.returncode == 1 and "needs fvMesh::Sf()" in p.stdout


@pytest.mark.gpu
def test_all_four_methods_select_by_dictionary(shim, tmp_path):
    """method methodOfClasses | quadratureMethodOfMoments | methodOfMoments | dummy through
    diameterModel::New -> PBEDiameter -> PBEMethod::New (PBEDiameter.C:36-81, PBEMethod.C:81-108)."""
    nc = 64
    v = np.ones(nc)
    kern = ("coalescenceKernel constantCoalescence; constantCoalescenceCoeffs { constant 0.1; } breakupKernel binaryBreakup; "
            "binaryBreakupCoeffs { } daughterParticleDistribution uniformBinaryBreakup; uniformBinaryBreakupCoeffs { } ")
    # dummy: constant diameter, no fields
    p, out = run(shim, tmp_path, "method dummy; dummyCoeffs { constantDiameter 0.00123; } " + kern, np.ones((1, nc)), v, v, v, v, 1e-3)
    assert p.returncode == 0, p.stdout + p.stderr
    assert np.all(out[1] == 0.00123) and "Selecting population balance model: dummy" in p.stdout
    # methodOfMoments: three moments of a gamma density, closed forms for constant coalescence + binary breakup
    from oracle import mom_oracle as mo
    k, th, m0 = 2.5, 0.3, 2.0
    m = np.stack([np.full(nc, m0), np.full(nc, m0 * k * th), np.full(nc, m0 * k * (k + 1) * th ** 2)])
    dt = 1e-3
    p, out = run(shim, tmp_path, "method methodOfMoments; MOMCoeffs { } " + kern, m, v, v, v, v, dt)
    assert p.returncode == 0, p.stdout + p.stderr
    S = np.sum(mo.closed_form_constant_binary(m[:, 0], 0.1), axis=0)
    assert np.allclose(out[0][:, 0], m[:, 0] + dt * S, rtol=1e-9)
    assert np.allclose(out[1], np.cbrt(6 / np.pi * out[0][1] / out[0][0]), rtol=1e-12)
    # the two hot-path methods select as well (their arithmetic is covered above)
    p, out = run(shim, tmp_path, "method methodOfClasses; MOCCoeffs { numberOfClasses 4; xi1 0.25; } " + kern,
                 np.ones((4, nc)), v, v, v, 0.05 * v, dt)
    assert p.returncode == 0 and "Selecting diameterModel for phase air: PBEDiameter" in p.stdout
    p, out = run(shim, tmp_path, "method quadratureMethodOfMoments; QMOMCoeffs { quadratureOrder 2; } " + kern,
                 np.array([[1.0], [1.0], [1.5], [3.0]]) * np.ones((4, nc)), v, v, v, v, dt)
    assert p.returncode == 0 and "moment sources:" in p.stdout and "diameter: avg, max,min" in p.stdout


def test_real_openfoam_branch_has_no_error_directive():
    """-DPBE_B200_REAL_OPENFOAM selects real OpenFOAM headers (absent here, so it cannot be compiled) but
    no longer stops at an #error: both transport branches are written out."""
    for fn in ("PBEMethodB200.H", "PBEMethodB200.C"):
        assert "#error" not in open(os.path.join(ROOT, "adapter", fn)).read()
