"""The C++ host layer above the C ABI: include/pbe_method.hpp (Foam-free mirror of PBEMethod /
MOC / QMOM and the kernel factories) and apps/pbe_replay (the PBEFoam-shaped driver).

CPU part: the header compiles with plain g++ against libpbe_b200.so, the dictionary parser reads
the reference's phaseProperties syntax, run-time selection fails like PBEMethod::New
(PBEMethod.C:97-105), and without a GPU construction fails loudly (no CPU fallback).
GPU part: the replay driver runs MOC and QMOM side by side and its MOC result matches the oracle.
"""
import json
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "openpbe_b200", "lib")

PROBE = r'''
#include "pbe_method.hpp"
#include <cstdio>
int main(int argc, char**) {
    // syntax of tests/integration/MOC/diffs/phaseProperties.coalescence:19-48
    auto d = pbe::Dictionary::parse(R"(
        method methodOfClasses;   // comment
        MOCCoeffs { numberOfClasses 10; xi1 1.0; usingMULES on; }
        coalescenceKernel constantCoalescence; /* block */ breakupKernel noBreakup;
        daughterParticleDistribution uniformBinaryBreakup;
        constantCoalescenceCoeffs { constant 0.1; } noBreakupCoeffs { } uniformBinaryBreakupCoeffs { }
    )", "PBEDiameterCoeffs");
    std::printf("method=%s N=%d xi1=%g mules=%d relax=%g c=%g\n", d.lookup("method").c_str(),
                d.subDict("MOCCoeffs").readLabel("numberOfClasses"), d.subDict("MOCCoeffs").readScalar("xi1"),
                (int)d.subDict("MOCCoeffs").lookupOrDefault<bool>("usingMULES", false),
                d.subDict("MOCCoeffs").lookupOrDefault<double>("sourceRelaxation", 1.0),
                d.subDict("constantCoalescenceCoeffs").readScalar("constant"));
    pbe_kernel_spec s = pbe::kernelSpecFromDict(d);
    std::printf("spec=%d,%d,%d,%g\n", s.coalescence, s.breakup, s.daughter, s.coal_c[0]);
    pbe::Phase ph; ph.deltaT = 1e-3;
    const char* cases[] = {"method bogus;", "method methodOfMoments;", "nomethod 1;"};
    for (const char* c : cases) {
        try { pbe::PBEMethod::New(pbe::Dictionary::parse(c), ph); std::printf("no error\n"); }
        catch (const pbe::FatalError& e) { std::printf("FATAL<<%s>>\n", e.what()); }
    }
    try { auto bad = d; bad.set("coalescenceKernel", "Madsen"); pbe::kernelSpecFromDict(bad); }
    catch (const pbe::FatalError& e) { std::printf("FATAL<<%s>>\n", e.what()); }
    if (argc > 1) {  // construction needs a device: must fail loudly without one
        try { auto m = pbe::PBEMethod::New(d, ph); std::printf("constructed\n"); }
        catch (const pbe::FatalError& e) { std::printf("FATAL<<%s>>\n", e.what()); }
    }
    return 0;
}
'''


@pytest.fixture(scope="module")
def probe(tmp_path_factory):
    if not os.path.exists(os.path.join(LIBDIR, "libpbe_b200.so")):
        pytest.skip("libpbe_b200.so not built")
    d = tmp_path_factory.mktemp("hostcpp")
    src = d / "probe.cpp"
    src.write_text(PROBE)
    exe = d / "probe"
    subprocess.check_call(["/usr/bin/g++", "-std=c++17", "-Wall", "-Wextra", "-Werror", "-I",
                           os.path.join(ROOT, "include"), str(src), "-o", str(exe), "-L", LIBDIR,
                           "-lpbe_b200", f"-Wl,-rpath,{LIBDIR}"])
    return str(exe)


def test_dictionary_and_runtime_selection(probe):
    out = subprocess.check_output([probe], text=True)
    assert "method=methodOfClasses N=10 xi1=1 mules=1 relax=1 c=0.1" in out
    assert "spec=1,0,0,0.1" in out                      # constantCoalescence, noBreakup, uniformBinaryBreakup
    assert "Unknown pbeModelType type bogus" in out and "Valid PBEMethod types are" in out
    assert "(dummy methodOfClasses methodOfMoments quadratureMethodOfMoments)" in out
    assert 'keyword method is undefined in dictionary' in out
    assert "Unknown coalescenceKernel type Madsen" in out and "CoulaloglouTavlaridesC" in out


def test_construction_fails_loudly_without_gpu(probe):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    out = subprocess.check_output([probe, "construct"], text=True)
    assert "constructed" not in out
    assert "no CUDA device" in out and "no CPU path" in out


def test_replay_driver_rejects_ranks_with_mesh():
    """--ranks forks one process per GPU before CUDA is touched; the decomposed transport is not wired into the
    driver, and it says so instead of running a wrong case (checked before any device is needed)."""
    exe = os.path.join(ROOT, "apps", "pbe_replay")
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "apps"), "-s"])
    r = subprocess.run([exe, "--ranks", "2", "--mesh", "8", "8", "8"], capture_output=True, text=True)
    assert r.returncode != 0 and "--ranks with --mesh" in r.stderr


@pytest.mark.gpu
def test_replay_driver_moc_and_qmom(oracle):
    exe = os.path.join(ROOT, "apps", "pbe_replay")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "apps"), "-s"])
    out = subprocess.check_output([exe, "--cells", "20000", "--steps", "3", "--classes", "30"], text=True)
    rows = [json.loads(ln) for ln in out.splitlines() if ln.startswith("{")]
    moc = [r for r in rows if r["method"] == "methodOfClasses" and "step" in r]
    qm = [r for r in rows if r["method"] == "quadratureMethodOfMoments" and "step" in r]
    assert len(moc) == 3 and len(qm) == 3
    for r in moc + qm:   # diameters stay in the physical range of the log-normal initial condition
        assert 1e-5 < r["d_min"] <= r["d_avg"] <= r["d_max"] < 2e-3 and np.isfinite(r["ms"])
    # same initial condition in both methods: the volume-mean diameter QMOM reports (QMOM.C:125)
    # sits within a few per cent of MOC's number-mean diameter (MOC.C:282-290) for this narrow pdf
    assert abs(qm[-1]["d_avg"] / moc[-1]["d_avg"] - 1.0) < 0.05
    summ = [r for r in rows if r.get("summary")]
    assert {r["method"] for r in summ} == {"methodOfClasses", "quadratureMethodOfMoments"}
    tail = [r for r in rows if "continuity_error" in r][0]
    assert tail["min_n0"] >= 0.0 and np.isfinite(tail["continuity_error"])


@pytest.mark.gpu
def test_replay_driver_with_transport(oracle):
    """MOC::correct() through solveWithMULES (source + explicit upwind transport on a hex block):
    a uniform velocity field with zero-gradient patches carries a spatially random field around
    without creating new extrema, and with zero velocity the step is the zero-flux update."""
    exe = os.path.join(ROOT, "apps", "pbe_replay")
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "apps"), "-s"])

    def run(*extra):
        out = subprocess.check_output([exe, "--mesh", "30", "20", "24", "--steps", "3", "--classes", "30", *extra],
                                      text=True)
        rows = [json.loads(ln) for ln in out.splitlines() if ln.startswith("{")]
        return [r for r in rows if r["method"] == "methodOfClasses" and "step" in r], rows

    still, _ = run("--velocity", "0", "0", "0")
    moving, rows = run("--velocity", "1", "0.5", "0.25")
    assert len(still) == 3 and len(moving) == 3
    for a, b in zip(still, moving):
        for r in (a, b):
            assert 1e-5 < r["d_min"] <= r["d_avg"] <= r["d_max"] < 2e-3
        # mixing neighbouring cells does not widen the spread of the per-cell mean diameter (up to
        # the coalescence rate's dependence on the mixed densities)
        assert b["d_max"] <= a["d_max"] * (1 + 1e-3) and b["d_min"] >= a["d_min"] * (1 - 1e-3)
    # zero velocity == the no-mesh (zero-flux) driver on the same number of cells
    out = subprocess.check_output([exe, "--cells", str(30 * 20 * 24), "--steps", "3", "--classes", "30"], text=True)
    flat = [json.loads(ln) for ln in out.splitlines() if ln.startswith("{")]
    flat = [r for r in flat if r["method"] == "methodOfClasses" and "step" in r]
    for a, b in zip(still, flat):
        assert abs(a["d_avg"] / b["d_avg"] - 1) < 1e-9 and abs(a["d_max"] / b["d_max"] - 1) < 1e-9
    tail = [r for r in rows if "min_n0" in r][0]
    assert tail["min_n0"] >= 0.0


@pytest.mark.gpu
def test_replay_driver_with_the_implicit_solve(oracle):
    """usingMULES left at the reference's default (false): both methods run solveWithFVM.  At a small Courant number the
    implicit and the explicit upwind update agree to first order in dt; at Courant 5 -- where the explicit update would
    blow up -- the implicit one stays bounded; zero velocity is the zero-flux update."""
    exe = os.path.join(ROOT, "apps", "pbe_replay")
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "apps"), "-s"])

    def run(*extra):
        out = subprocess.check_output([exe, "--mesh", "30", "20", "24", "--steps", "3", "--classes", "30", *extra], text=True)
        rows = [json.loads(ln) for ln in out.splitlines() if ln.startswith("{")]
        return [r for r in rows if r["method"] == "methodOfClasses" and "step" in r], rows

    expl, _ = run("--courant", "0.05")
    impl, rows = run("--implicit", "--courant", "0.05")
    for a, b in zip(expl, impl):
        assert abs(a["d_avg"] / b["d_avg"] - 1) < 1e-4 and abs(a["d_max"] / b["d_max"] - 1) < 1e-2
    perf = [r for r in rows if r.get("transport") == "solveWithFVM"][0]
    assert 1 <= perf["iterations"][0] <= perf["iterations"][1] <= 10 and perf["final_residual_max"] < 1e-5
    big, rows = run("--implicit", "--courant", "5")
    still, _ = run("--implicit", "--velocity", "0", "0", "0")
    for b in big:  # deltaT is 400 x the default here: the populations evolve much further, and stay physical
        assert 1e-5 < b["d_min"] <= b["d_avg"] <= b["d_max"] < 2e-3
    tail = [r for r in rows if "min_n0" in r][0]
    assert tail["min_n0"] >= 0.0  # M-matrix: no class goes negative at any Courant number
    perf = [r for r in rows if r.get("transport") == "solveWithFVM"][0]
    assert perf["iterations"][1] > 10 and perf["final_residual_max"] < 1e-5
    qm = [r for r in rows if r["method"] == "quadratureMethodOfMoments" and "step" in r]
    assert len(qm) == 3 and all(1e-5 < r["d_min"] <= r["d_max"] < 2e-3 for r in qm)
    out = subprocess.check_output([exe, "--cells", str(30 * 20 * 24), "--steps", "3", "--classes", "30"], text=True)
    flat = [json.loads(ln) for ln in out.splitlines() if ln.startswith("{")]
    flat = [r for r in flat if r["method"] == "methodOfClasses" and "step" in r]
    for a, b in zip(still, flat):
        assert abs(a["d_avg"] / b["d_avg"] - 1) < 1e-9 and abs(a["d_max"] / b["d_max"] - 1) < 1e-9


MOM_PROBE = r'''
#include "pbe_method.hpp"
#include <cstdio>
// PBEMethod::New -> methodOfMoments (MOM.C:60-236) on a few cells with O(1) moments; prints the
// moments, scrubbed sources, gamma parameters and d after one correct().
int main() {
    const int nc = 6;
    std::vector<double> m(3 * nc), eps(nc, 1.0), rho(nc, 1.0), nu(nc, 1.0);
    for (int c = 0; c < nc; ++c) {
        const double m0 = 1.0 + 0.3 * c, k = 1.5 + 0.4 * c, th = 0.2 + 0.05 * c;
        m[c] = m0; m[nc + c] = m0 * k * th; m[2 * nc + c] = m0 * k * (k + 1) * th * th;
    }
    pbe::Phase ph; ph.ncells = nc; ph.deltaT = 1e-2; ph.epsilon = eps.data(); ph.rho = rho.data(); ph.nu = nu.data();
    auto dict = pbe::Dictionary::parse(R"(
        method methodOfMoments; MOMCoeffs { }
        coalescenceKernel constantCoalescence; breakupKernel binaryBreakup;
        daughterParticleDistribution uniformBinaryBreakup;
        constantCoalescenceCoeffs { constant 0.1; } binaryBreakupCoeffs { } uniformBinaryBreakupCoeffs { }
    )");
    try {
        pbe::PBEMethods::MOM mom(dict, ph, m.data());
        std::printf("d0 %.17g\n", mom.d()[2]);
        mom.correct();
        for (int c = 0; c < nc; ++c) {
            std::printf("cell");
            for (int k = 0; k < 3; ++k) std::printf(" %.17g", m[k * nc + c]);
            for (auto f : {pbe::PBEMethods::MOM::Moment, pbe::PBEMethods::MOM::Source, pbe::PBEMethods::MOM::Gamma})
                for (int k = 0; k < 3; ++k) std::printf(" %.17g", mom.field(f, k)[c]);
            std::printf(" %.17g\n", mom.d()[c]);
        }
        auto viaNew = pbe::PBEMethod::New(dict, ph);
        viaNew->correct();
        std::printf("new ok %zu\n", viaNew->d().size());
    } catch (const pbe::FatalError &e) { std::printf("FATAL<<%s>>\n", e.what()); return 1; }
    return 0;
}
'''


@pytest.mark.gpu
def test_method_of_moments_host_mirror(tmp_path):
    """pbe::PBEMethods::MOM (the MOM class of MOM.H:55-96 over pbe_mom_step) against the closed
    forms of the gamma closure for constantCoalescence + binaryBreakup."""
    from oracle import mom_oracle as mo
    src = tmp_path / "mom_probe.cpp"
    src.write_text(MOM_PROBE)
    exe = tmp_path / "mom_probe"
    subprocess.check_call(["/usr/bin/g++", "-std=c++17", "-Wall", "-Wextra", "-Werror", "-I",
                           os.path.join(ROOT, "include"), str(src), "-o", str(exe), "-L", LIBDIR,
                           "-lpbe_b200", f"-Wl,-rpath,{LIBDIR}"])
    out = subprocess.check_output([str(exe)], text=True)
    assert "new ok 6" in out
    rows = [np.array(ln.split()[1:], dtype=float) for ln in out.splitlines() if ln.startswith("cell")]
    assert len(rows) == 6
    for r in rows:
        m_in, m_out, S, gam, d = r[0:3], r[3:6], r[6:9], r[9:12], r[12]
        coal, br = mo.closed_form_constant_binary(m_in, 0.1)
        np.testing.assert_allclose(S, coal + br, rtol=0, atol=1e-10 * max(1.0, np.max(np.abs(coal + br))))
        np.testing.assert_allclose(gam, mo.gamma_params(*m_in), rtol=1e-14)
        _, want, dwant = mo.mom_step_cell(S, m_in, 1e-2)
        np.testing.assert_allclose(m_out, want, rtol=1e-15)
        np.testing.assert_allclose(d, dwant, rtol=1e-14)
    d0 = float([ln for ln in out.splitlines() if ln.startswith("d0")][0].split()[1])
    r2 = rows[2]
    assert d0 == (6 / np.pi * r2[1] / r2[0]) ** (1 / 3) or abs(d0 / (6 / np.pi * r2[1] / r2[0]) ** (1 / 3) - 1) < 1e-15


MOM_MESH_PROBE = r'''
#include "pbe_method.hpp"
#include <cstdio>
// methodOfMoments on a chain of 6 cells (5 internal faces, an inlet and an outlet face) with a uniform flux: MOM::correct()
// then ends in the implicit solve of MOM.C:211-222.  Prints moments before / after and the scrubbed sources.
int main(int argc, char **argv) {
    const int nc = 6;
    const double flux = argc > 1 ? std::atof(argv[1]) : 0.0;
    std::vector<double> m(3 * nc), eps(nc, 1.0), rho(nc, 1.0), nu(nc, 1.0), V(nc, 0.5), alpha(nc, 0.1);
    for (int c = 0; c < nc; ++c) {
        const double m0 = 1.0 + 0.3 * c, k = 1.5 + 0.4 * c, th = 0.2 + 0.05 * c;
        m[c] = m0; m[nc + c] = m0 * k * th; m[2 * nc + c] = m0 * k * (k + 1) * th * th;
    }
    std::vector<int32_t> own = {0, 1, 2, 3, 4}, nei = {1, 2, 3, 4, 5}, bown = {0, 5};
    std::vector<double> phi = {flux, flux, flux, flux, flux, -flux, flux};
    pbe::FaceMesh fm;
    fm.nInternalFaces = 5; fm.nBoundaryFaces = 2;
    fm.owner = own.data(); fm.neighbour = nei.data(); fm.boundaryOwner = bown.data(); fm.phi = phi.data();
    fm.tolerance = 1e-14;
    pbe::Phase ph; ph.ncells = nc; ph.deltaT = 1e-2; ph.epsilon = eps.data(); ph.rho = rho.data(); ph.nu = nu.data();
    ph.V = V.data(); ph.alpha = alpha.data(); ph.mesh = &fm;
    auto dict = pbe::Dictionary::parse(R"(
        method methodOfMoments; MOMCoeffs { }
        coalescenceKernel constantCoalescence; breakupKernel binaryBreakup;
        daughterParticleDistribution uniformBinaryBreakup;
        constantCoalescenceCoeffs { constant 0.1; } binaryBreakupCoeffs { } uniformBinaryBreakupCoeffs { }
    )");
    try {
        pbe::PBEMethods::MOM mom(dict, ph, m.data());
        mom.correct();
        for (int c = 0; c < nc; ++c) {
            std::printf("cell");
            for (int k = 0; k < 3; ++k) std::printf(" %.17g", m[k * nc + c]);
            for (auto f : {pbe::PBEMethods::MOM::Moment, pbe::PBEMethods::MOM::Source})
                for (int k = 0; k < 3; ++k) std::printf(" %.17g", mom.field(f, k)[c]);
            std::printf(" %.17g\n", mom.d()[c]);
        }
    } catch (const pbe::FatalError &e) { std::printf("FATAL<<%s>>\n", e.what()); return 1; }
    return 0;
}
'''


@pytest.mark.gpu
def test_method_of_moments_with_the_implicit_transport(tmp_path):
    """MOM::correct() with face fluxes (MOM.C:211-222: fvm::ddt + fvm::div == S for the three moments): the host mirror's
    result against the direct solve of the restated matrix with the device's own sources; zero flux = the zero-flux
    update."""
    from oracle import fvm_implicit as fvm
    src = tmp_path / "mom_mesh_probe.cpp"
    src.write_text(MOM_MESH_PROBE)
    exe = tmp_path / "mom_mesh_probe"
    subprocess.check_call(["/usr/bin/g++", "-std=c++17", "-Wall", "-Wextra", "-Werror", "-I",
                           os.path.join(ROOT, "include"), str(src), "-o", str(exe), "-L", LIBDIR,
                           "-lpbe_b200", f"-Wl,-rpath,{LIBDIR}"])
    mesh = {"owner": np.arange(5), "neighbour": np.arange(1, 6), "boundary_owner": np.array([0, 5]), "vcell": np.full(6, 0.5)}
    for flux in (0.0, 3.0):
        out = subprocess.check_output([str(exe), repr(flux)], text=True)
        rows = np.array([ln.split()[1:] for ln in out.splitlines() if ln.startswith("cell")], dtype=float)
        assert rows.shape == (6, 10), out
        m_in, m_out, S, d = rows[:, 0:3].T, rows[:, 3:6].T, rows[:, 6:9].T, rows[:, 9]
        phi = np.array([flux] * 5 + [-flux, flux])
        want = np.maximum(fvm.solve(mesh, phi, 1e-2, m_in, S=S), 1e-15)
        np.testing.assert_allclose(m_out, want, rtol=1e-12)
        np.testing.assert_allclose(d, np.cbrt(6 / np.pi * m_out[1] / m_out[0]), rtol=1e-14)
        if flux == 0.0:
            np.testing.assert_allclose(m_out, m_in + 1e-2 * S, rtol=1e-14)
        else:
            assert np.abs(m_out - (m_in + 1e-2 * S)).max() > 1e-3  # the flux did move something
