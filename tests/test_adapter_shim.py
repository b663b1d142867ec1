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
    p, _ = run(shim, tmp_path, "method bogus; coalescenceKernel noCoalescence; breakupKernel noBreakup;"
               " daughterParticleDistribution uniformBinaryBreakup;", one, v, v, v, v, 1e-3)
    assert p.returncode == 1
    assert "Unknown pbeModelType type bogus" in p.stdout and "Valid PBEMethod types are" in p.stdout
    assert "(methodOfClasses quadratureMethodOfMoments)" in p.stdout and "FOAM exiting" in p.stdout
    p, _ = run(shim, tmp_path, "method methodOfClasses; MOCCoeffs { numberOfClasses 2; xi1 1; } coalescenceKernel Madsen;"
               " breakupKernel noBreakup; daughterParticleDistribution uniformBinaryBreakup;", one, v, v, v, v, 1e-3)
    assert p.returncode == 1 and "Madsen" in p.stdout and "CoulaloglouTavlaridesC" in p.stdout
    p, _ = run(shim, tmp_path, "coalescenceKernel noCoalescence;", one, v, v, v, v, 1e-3)
    assert p.returncode == 1 and "keyword method is undefined in dictionary" in p.stdout


def test_no_device_is_fatal_not_a_fallback(shim, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    one = np.ones((2, 4))
    v = np.ones(4)
    p, _ = run(shim, tmp_path, "method methodOfClasses; MOCCoeffs { numberOfClasses 2; xi1 1; } coalescenceKernel "
               "noCoalescence; breakupKernel noBreakup; daughterParticleDistribution uniformBinaryBreakup;",
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


def _write_mesh(path, m):
    with open(path, "wb") as fh:
        np.array([m["owner"].size, m["boundary_owner"].size], dtype=np.float64).tofile(fh)
        for k in ("owner", "neighbour", "boundary_owner"):
            m[k].astype(np.int32).tofile(fh)
        m["phi"].astype(np.float64).tofile(fh)


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
    (tmp_path / "dict").write_text(text.replace("Gauss upwind", "Gauss vanLeer"))
    p = subprocess.run([shim, str(tmp_path / "dict"), str(tmp_path / "in.bin"), str(tmp_path / "out.bin"), "1",
                        str(tmp_path / "mesh.bin")], capture_output=True, text=True)
    assert p.returncode == 1 and "vanLeer" in p.stdout and "Valid schemes are" in p.stdout and "FOAM exiting" in p.stdout
