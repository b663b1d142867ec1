// refFoam.C -- TEST INFRASTRUCTURE: the out-of-line part of oracle/ref_shim/refFoam.H.
// Restatements of OpenFOAM 3.0.x algorithms the reference's transport steps call (MOC.C:186-280,
// QMOM.C:131-209).  OpenFOAM itself is absent from the reference tree and from this image: parity
// of these functions is UNPINNED (they are written from OpenFOAM 3.0.x's published sources:
// finiteVolume/fvMatrices/solvers/MULES/{MULES.C,MULESTemplates.C},
// interpolation/surfaceInterpolation/{schemes/upwind/upwind.H,limitedSchemes/{limitedScheme.C,
// vanLeer/vanLeer.H,LimitedScheme/NVDTVD.H}}, finiteVolume/fvc/fvcSurfaceIntegrate.C,
// finiteVolume/gradSchemes/gaussGrad).
#include "refFoam.H"

namespace Foam {

extern "C" {
#include "../pbe_oracle.h"
}

// The OpenFOAM algorithms themselves live in oracle/pbe_oracle_mules.c (plain C on plain arrays) so that
// ONE restatement serves both the reference's own call sequence (through this shim) and the port's
// (orc_transport_mules); the functions below only unpack the shim's containers.
static orc_mesh as_orc(const fvMesh &mesh) {
    orc_mesh m;
    m.ncells = mesh.nCells();
    m.nint = mesh.nInternalFaces();
    m.nbnd = (int64_t)mesh.boundaryFaceCells_.size();
    m.owner = mesh.owner_.data();
    m.neighbour = mesh.neighbour_.data();
    m.bown = mesh.boundaryFaceCells_.data();
    m.V = mesh.V().data();
    static_assert(sizeof(Vec3) == 3 * sizeof(double), "Vec3 is three packed doubles");
    m.Sf = ((label)mesh.Sf_.size() == mesh.nFaces() && mesh.nFaces() > 0) ? &mesh.Sf_[0].x : nullptr;
    m.C = (mesh.nCells() > 0) ? &mesh.C()[0].x : nullptr;
    m.weights = (mesh.weights_.size() == mesh.nInternalFaces() && mesh.nInternalFaces() > 0) ? mesh.weights_.data() : nullptr;
    return m;
}
static const double *patch_values(const volScalarField &vf) { return vf.bf_.empty() ? nullptr : vf.bf_.data(); }

namespace fvc {
// fvc::flux(phi, vf, name) = scheme(mesh, phi, divScheme(name)).dotInterpolate(phi, vf)
surfaceScalarField flux(const surfaceScalarField &phi, const volScalarField &vf, const word &name) {
    const fvMesh &mesh = vf.mesh();
    const word scheme = mesh.divScheme(name);
    int id = -1;
    if (scheme == "Gauss upwind" || scheme == "upwind") id = ORC_DIV_UPWIND;
    if (scheme == "Gauss vanLeer" || scheme == "vanLeer") id = ORC_DIV_VANLEER;
    if (id < 0)
        FatalErrorIn("fvc::flux") << "unknown discretisation scheme " << scheme
                                  << "\n\nValid schemes are :\n2\n(\nupwind\nvanLeer\n)" << exit(FatalError);
    surfaceScalarField r("flux(" + vf.name() + ")", mesh);
    const orc_mesh m = as_orc(mesh);
    if (orc_div_flux(&m, id, phi.f_.data(), vf.internalField().data(), patch_values(vf), r.f_.data()) != 0)
        FatalErrorIn("fvc::flux") << "scheme " << scheme << " needs the mesh geometry (Sf, C, weights)" << exit(FatalError);
    return r;
}
} // namespace fvc

namespace MULES {

void limiter(scalarField &allLambda, const scalar rDeltaT, const volScalarField &psi, const surfaceScalarField &phiBD,
             const surfaceScalarField &phiCorr, const volScalarField &Su, const scalar psiMax, const scalar psiMin) {
    const orc_mesh m = as_orc(psi.mesh());
    orc_mules_limiter(&m, allLambda.data(), rDeltaT, psi.internalField().data(), patch_values(psi), phiBD.f_.data(),
                      phiCorr.f_.data(), Su.internalField().data(), psiMax, psiMin, psi.mesh().nLimiterIter_);
}

// MULESTemplates.C: limit(): phiBD = upwind flux; phiCorr = phiPsi - phiBD; limiter; returnCorr
void limit(const scalar rDeltaT, const geometricOneField &, const volScalarField &psi, const surfaceScalarField &phi,
           surfaceScalarField &phiPsi, const zeroField &, const volScalarField &Su, const scalar psiMax,
           const scalar psiMin, const bool returnCorr) {
    const fvMesh &mesh = psi.mesh();
    surfaceScalarField phiBD(upwind<scalar>(mesh, phi).flux(psi));
    surfaceScalarField &phiCorr = phiPsi;
    phiCorr -= phiBD;
    scalarField allLambda(mesh.nFaces(), 1.0);
    limiter(allLambda, rDeltaT, psi, phiBD, phiCorr, Su, psiMax, psiMin);
    if (returnCorr) {
        for (label f = 0; f < phiCorr.size(); ++f) phiCorr[f] *= allLambda[f];
    } else {
        for (label f = 0; f < phiCorr.size(); ++f) phiPsi[f] = phiBD[f] + allLambda[f] * phiCorr[f];
    }
}

// MULES.C: limitSum() over the fields, internal faces
void limitSum(PtrList<surfaceScalarField> &phiPsiCorrs) {
    const label K = phiPsiCorrs.size();
    if (K == 0) return;
    const label nf = phiPsiCorrs[0].size(), ni = phiPsiCorrs[0].mesh().nInternalFaces();
    std::vector<double> all((size_t)K * (size_t)(nf > 0 ? nf : 1));
    for (label k = 0; k < K; ++k)
        for (label f = 0; f < nf; ++f) all[(size_t)k * nf + f] = phiPsiCorrs[k][f];
    orc_mules_limit_sum(K, ni, nf, all.data());
    for (label k = 0; k < K; ++k)
        for (label f = 0; f < nf; ++f) phiPsiCorrs[k][f] = all[(size_t)k * nf + f];
}

// MULESTemplates.C: explicitSolve() with rho = 1, Sp = 0
void explicitSolve(const geometricOneField &, volScalarField &psi, const surfaceScalarField &phiPsi, const zeroField &,
                   const volScalarField &Su) {
    Info << "MULES: Solving for " << psi.name() << endl;
    refProbe::record(psi.name(), Su.internalField());
    const orc_mesh m = as_orc(psi.mesh());
    orc_mules_explicit_solve(&m, 1.0 / psi.mesh().time().deltaTValue(), psi.internalField().data(), phiPsi.f_.data(),
                             Su.internalField().data());
}

} // namespace MULES

// fvm::ddt(psi) + fvm::div(phi, psi) == Su  (Euler implicit, `Gauss upwind` for the implicit div as the
// reference cases select, bounded by construction).  Assembled as fvMatrix does (EulerDdtScheme::fvmDdt:
// diag = rDeltaT V, source = rDeltaT psi0 V; gaussConvectionScheme::fvmDiv: lower = -w phi, upper =
// lower + phi, diag -= sums; zero-gradient patches add phi_b to the diagonal), solved by Gauss-Seidel.
void fvScalarMatrix::solve() {
    volScalarField &psi = *psi_;
    const fvMesh &mesh = psi.mesh();
    const label nc = psi.size(), ni = mesh.nInternalFaces();
    const scalar rDeltaT = 1.0 / mesh.time().deltaTValue();
    scalarField diag(nc, 0.0), b(nc, 0.0), lower(ni, 0.0), upper(ni, 0.0);
    for (label c = 0; c < nc; ++c) {
        if (hasDdt_) {
            diag[c] = rDeltaT * mesh.V()[c];
            b[c] = rDeltaT * psi[c] * mesh.V()[c];
        }
        b[c] += source_.size() ? source_[c] : 0.0;
    }
    if (phi_) {
        const surfaceScalarField &phi = *phi_;
        for (label f = 0; f < ni; ++f) {
            const scalar w = pos(phi[f]);
            lower[f] = -w * phi[f];
            upper[f] = lower[f] + phi[f];
            diag[mesh.owner_[f]] -= lower[f];
            diag[mesh.neighbour_[f]] -= upper[f];
        }
        for (label bf = 0; bf < (label)mesh.boundaryFaceCells_.size(); ++bf) {
            const label c = mesh.boundaryFaceCells_[bf];
            if (psi.bf_.empty())
                diag[c] += phi[ni + bf]; // zero-gradient: internalCoeffs = phi_b
            else
                b[c] -= phi[ni + bf] * psi.bf_[bf]; // fixed value: boundaryCoeffs = -phi_b psi_b
        }
    }
    if (ni == 0 || !phi_) {
        for (label c = 0; c < nc; ++c) psi[c] = b[c] / diag[c];
        return;
    }
    // cell -> faces
    std::vector<std::vector<std::pair<label, scalar>>> offd((size_t)nc);
    for (label f = 0; f < ni; ++f) {
        offd[mesh.owner_[f]].push_back(std::make_pair(mesh.neighbour_[f], upper[f])); // row own, column nei
        offd[mesh.neighbour_[f]].push_back(std::make_pair(mesh.owner_[f], lower[f])); // row nei, column own
    }
    scalarField x(psi.internalField());
    for (int it = 0; it < 1000; ++it) {
        scalar change = 0, norm = 0;
        for (label c = 0; c < nc; ++c) {
            scalar s = b[c];
            for (size_t k = 0; k < offd[c].size(); ++k) s -= offd[c][k].second * x[offd[c][k].first];
            const scalar xn = s / diag[c];
            change += mag(xn - x[c]);
            norm += mag(xn);
            x[c] = xn;
        }
        if (change <= 1e-14 * norm) break;
    }
    psi.internalField() = x;
}

} // namespace Foam
