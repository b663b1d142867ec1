// ref_driver.C -- TEST INFRASTRUCTURE: C entry points over the reference's OWN classes, compiled
// unmodified from /root/reference/src/PBESystem against oracle/ref_shim (see oracle/Makefile.ref).
// Used by tests/ to pin oracle/pbe_oracle.c on the reference itself, and by bench.py as the
// `--impl reference` arm and the `cpu_baseline` (kind "reference").  Never linked into the product.
//
// What runs here is the reference: PBEMethod::New -> MOC / QMOM construction from a
// PBEDiameterCoeffs-style dictionary (PBEMethod.C:81-108, MOC.C:49-83, QMOM.C:59-82), then
// method->correct() and method->d() per time step (the loop of applications/PBEFoam/PBEFoam.C:56-67).
// The explicit sources S_k the method hands to its transport step are read back from the shim's
// MULES::explicitSolve / fvMatrix probe.
//
// One mesh per process (the reference's kernels cache `static const volScalarField&`,
// coalescenceKernels/CoulaloglouTavlarides.C:52-53): the mesh, the phase and epsilon.<phase> are
// process-lifetime singletons, resized per call.
#include <chrono>
#include <cstring>
#include <string>

#include "fvCFD.H"
#include "PBEMethod.H"
#include "MomentInversion.H"
#include "coalescenceKernels/CoulaloglouTavlarides.H"
#include "coalescenceKernels/constantCoalescence.H"
#include "coalescenceKernels/noCoalescence.H"
#include "breakupKernels/CoulaloglouTavlarides.H"
#include "breakupKernels/binaryBreakup.H"
#include "breakupKernels/noBreakup.H"
#include "daughterParticleDistributions/uniformBinaryBreakup.H"

using namespace Foam;

namespace {

const char *const kPhase = "dispersed";
std::string g_error;

struct World {
    fvMesh mesh;
    phaseModel phase;
    volScalarField epsilon;
    World()
        : mesh(), phase(mesh, kPhase),
          epsilon(IOobject(std::string("epsilon.") + kPhase, "0", mesh), mesh, dimensionedScalar("eps", dimless, 0.0)) {}
};
World &world() {
    static World *w = new World(); // never destroyed: static references into it live in the reference's kernels
    return *w;
}

void set_cells(int64_t nc, const double *eps, const double *rho, const double *nu, const double *alpha,
               const double *V, double dt) {
    World &w = world();
    if (w.mesh.nFaces() == 0) w.mesh.resize((label)nc, 1.0); // 0-D cells; a face mesh was sized by ref_set_mesh
    w.mesh.time().setDeltaT(dt);
    w.phase.resizeAll((label)nc);
    w.epsilon.resize((label)nc, 0.0);
    for (label c = 0; c < (label)nc; ++c) {
        w.epsilon[c] = eps ? eps[c] : 0.0;
        w.phase.rhoRef()[c] = rho ? rho[c] : 0.0;
        w.phase.nuRef()[c] = nu ? nu[c] : 0.0;
        w.phase[c] = alpha ? alpha[c] : 0.0;
        if (V) w.mesh.Vref()[c] = V[c];
    }
}

} // namespace

extern "C" {

const char *ref_last_error() { return g_error.c_str(); }
void ref_set_quiet(int q) { refShimControl::quiet() = (q != 0); }

// Face addressing and geometry of the (single) mesh; nint = nbnd = 0 restores the 0-D mesh.
// Sf [nfaces][3], C [ncells][3], weights [nint] may be NULL when only `Gauss upwind` is used.
int ref_set_mesh(int64_t ncells, int64_t nint, const int32_t *owner, const int32_t *neighbour, int64_t nbnd,
                 const int32_t *boundary_owner, const double *Sf, const double *C, const double *weights,
                 const char *div_scheme, int n_limiter_iter) {
    World &w = world();
    fvMesh &m = w.mesh;
    m.resize((label)ncells, 1.0);
    m.owner_.assign(owner, owner + nint);
    m.neighbour_.assign(neighbour, neighbour + nint);
    m.boundaryFaceCells_.assign(boundary_owner, boundary_owner + nbnd);
    m.Sf_ = vectorField(Sf ? (label)(nint + nbnd) : 0);
    for (size_t f = 0; f < m.Sf_.size(); ++f) m.Sf_[f] = vector{Sf[3 * f], Sf[3 * f + 1], Sf[3 * f + 2]};
    if (C)
        for (label c = 0; c < (label)ncells; ++c) m.Cref()[c] = vector{C[3 * c], C[3 * c + 1], C[3 * c + 2]};
    m.weights_ = scalarField(weights ? (label)nint : 0);
    for (label f = 0; f < m.weights_.size(); ++f) m.weights_[f] = weights[f];
    m.divSchemes_.clear();
    m.divSchemes_["default"] = div_scheme ? div_scheme : "Gauss upwind";
    m.nLimiterIter_ = n_limiter_iter > 0 ? n_limiter_iter : 3;
    w.phase.phiRef().sync();
    return 0;
}

// nsteps x { method->correct(); d = method->d(); } of the method the dictionary selects.
//   dict_text : PBEDiameterCoeffs-style dictionary (method, MOCCoeffs / QMOMCoeffs, kernels, <Type>Coeffs)
//   fields    : [nfields][ncells] -- n0..n{N-1} (MOC) or m0..m{2 order - 1} (QMOM): read as the MUST_READ
//               fields, updated in place (what runTime.write() would write)
//   phi       : [faces] flux of the phase or NULL (zero); S_out [nfields][ncells] (sources of the LAST
//               step) and d_out [ncells] may be NULL.  secs: time inside the correct() + d() loop.
int ref_method_steps(const char *dict_text, const char *field_stem, int nfields, int64_t ncells, double *fields,
                     const double *eps, const double *rho, const double *nu, const double *alpha, const double *V,
                     const double *phi, double dt, int nsteps, double *S_out, double *d_out, double *secs) {
    try {
        World &w = world();
        if (w.mesh.nFaces() != 0 && (int64_t)w.mesh.nCells() != ncells) {
            g_error = "ref_method_steps: ncells differs from the mesh set by ref_set_mesh";
            return 1;
        }
        set_cells(ncells, eps, rho, nu, alpha, V, dt);
        w.phase.phiRef().sync();
        for (label f = 0; f < w.phase.phiRef().size(); ++f) w.phase.phiRef()[f] = phi ? phi[f] : 0.0;
        w.mesh.disk.clear();
        for (int k = 0; k < nfields; ++k)
            w.mesh.disk[field_stem + std::to_string(k)].assign(fields + (size_t)k * ncells, fields + (size_t)(k + 1) * ncells);
        refProbe::sources().clear();
        refShimControl::probeSources() = (S_out != nullptr);
        const dictionary dict = dictionary::fromString(dict_text, "PBEDiameterCoeffs");
        autoPtr<PBEMethod> method = PBEMethod::New(dict, w.phase);
        w.mesh.disk.clear();
        const auto t0 = std::chrono::steady_clock::now();
        for (int s = 0; s < nsteps; ++s) {
            method->correct();
            const volScalarField d(method->d());
            if (s == nsteps - 1 && d_out)
                for (label c = 0; c < (label)ncells; ++c) d_out[c] = d[c];
        }
        const auto t1 = std::chrono::steady_clock::now();
        if (secs) *secs = std::chrono::duration<double>(t1 - t0).count();
        for (int k = 0; k < nfields; ++k) {
            const volScalarField &f = w.mesh.lookupObject<volScalarField>(field_stem + std::to_string(k));
            for (label c = 0; c < (label)ncells; ++c) fields[(size_t)k * ncells + c] = f[c];
            if (S_out) {
                const std::vector<scalar> &S = refProbe::sources()[field_stem + std::to_string(k)];
                if ((int64_t)S.size() != ncells) {
                    g_error = "ref_method_steps: no source was recorded for field " + std::string(field_stem) + std::to_string(k);
                    return 1;
                }
                std::memcpy(S_out + (size_t)k * ncells, S.data(), sizeof(double) * (size_t)ncells);
            }
        }
        return 0;
    } catch (const std::exception &e) {
        g_error = e.what();
        return 2;
    }
}

// ---- the Foam-free *Impl functors (what tests/unit_tests and bindings/kernels.cpp exercise) ------------------
double ref_ct_coalescence_S(double c1, double c2, double gamma, double sigma, double xi1, double xi2, double eps,
                            double rho, double nu, double V) {
    return coalescenceKernels::CoulaloglouTavlaridesCImpl(c1, c2, gamma, sigma).S(xi1, xi2, eps, rho, nu, V);
}
double ref_constant_coalescence_S(double c) { return coalescenceKernels::constantCoalescenceImpl(c).S(); }
double ref_no_coalescence_S() { return coalescenceKernels::noCoalescenceImpl().S(); }
double ref_ct_breakup_S(double c1, double c2, double gamma, double sigma, double xi, double rho_d, double eps) {
    return breakupKernels::CoulaloglouTavlaridesImp(c1, c2, gamma, sigma).S(xi, rho_d, eps);
}
double ref_binary_breakup_S(double xi) { return breakupKernels::binaryBreakupImpl().S(xi); }
double ref_no_breakup_S(double xi) { return breakupKernels::noBreakupImpl().S(xi); }
double ref_uniform_binary_beta(double xi1, double xi2) {
    return daughterParticleDistributions::uniformBinaryBreakup(dictionary("uniformBinaryBreakupCoeffs")).beta(xi1, xi2);
}
double ref_uniform_binary_moment(double xi, int k) {
    return daughterParticleDistributions::uniformBinaryBreakup(dictionary("uniformBinaryBreakupCoeffs"))
        .moment(dimensionedScalar(xi), k)
        .value();
}
// wheeler_inversion (MomentInversion.C:9-51) of n2 = 2 N moments -> N abscissas, N weights
int ref_wheeler(int n2, const double *moments, double *x, double *wts) {
    Eigen::VectorXd m(n2);
    for (int i = 0; i < n2; ++i) m[i] = moments[i];
    const Quadrature q = wheeler_inversion(m);
    for (int i = 0; i < n2 / 2; ++i) {
        x[i] = q.abcissas[i];
        wts[i] = q.weights[i];
    }
    return 0;
}

} // extern "C"
