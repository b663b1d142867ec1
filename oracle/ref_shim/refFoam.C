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

// ---- Gauss linear gradient (gradSchemes default of the reference's cases) -----------------------------
static std::vector<vector> gaussLinearGrad(const volScalarField &vf) {
    const fvMesh &mesh = vf.mesh();
    const label ni = mesh.nInternalFaces();
    std::vector<vector> g((size_t)mesh.nCells(), vector{0, 0, 0});
    for (label f = 0; f < ni; ++f) {
        const label own = mesh.owner_[f], nei = mesh.neighbour_[f];
        const scalar w = mesh.weights_[f];
        const scalar ssf = w * (vf[own] - vf[nei]) + vf[nei]; // linear interpolate: lambda*(P - N) + N
        const vector Sfssf = ssf * mesh.Sf_[f];
        g[own] = g[own] + Sfssf;
        g[nei] = g[nei] - Sfssf;
    }
    for (label b = 0; b < (label)mesh.boundaryFaceCells_.size(); ++b) {
        const label c = mesh.boundaryFaceCells_[b];
        g[c] = g[c] + vf.patchValue(b) * mesh.Sf_[ni + b];
    }
    for (label c = 0; c < mesh.nCells(); ++c) g[c] = (1.0 / mesh.V()[c]) * g[c];
    return g;
}

// NVDTVD::r (LimitedScheme/NVDTVD.H) and the vanLeer limiter (vanLeer.H)
static inline scalar nvdtvd_r(const scalar faceFlux, const scalar phiP, const scalar phiN, const vector &gradcP,
                              const vector &gradcN, const vector &d) {
    const scalar gradf = phiN - phiP;
    scalar gradcf;
    if (faceFlux > 0)
        gradcf = d & gradcP;
    else
        gradcf = d & gradcN;
    if (mag(gradcf) >= 1000 * mag(gradf))
        return 2 * 1000 * sign(gradcf) * sign(gradf) - 1;
    return 2 * (gradcf / gradf) - 1;
}

namespace fvc {
// fvc::flux(phi, vf, name) = scheme(mesh, phi, divScheme(name)).dotInterpolate(phi, vf):
// phi * (weights * P + (1 - weights) * N) with the scheme's weights; patch values on boundary faces.
surfaceScalarField flux(const surfaceScalarField &phi, const volScalarField &vf, const word &name) {
    const fvMesh &mesh = vf.mesh();
    const word scheme = mesh.divScheme(name);
    if (scheme == "Gauss upwind" || scheme == "upwind") return upwind<scalar>(mesh, phi).flux(vf);
    if (scheme != "Gauss vanLeer" && scheme != "vanLeer")
        FatalErrorIn("fvc::flux") << "unknown discretisation scheme " << scheme
                                  << "\n\nValid schemes are :\n2\n(\nupwind\nvanLeer\n)" << exit(FatalError);
    // limitedScheme<vanLeer>: limiter(r), weights = limiter * w_linear + (1 - limiter) * pos(phi)
    const label ni = mesh.nInternalFaces();
    if ((label)mesh.Sf_.size() != mesh.nFaces() || mesh.weights_.size() != ni)
        FatalErrorIn("fvc::flux") << "scheme " << scheme << " needs the mesh geometry (Sf, C, weights)" << exit(FatalError);
    const std::vector<vector> gradc = gaussLinearGrad(vf);
    surfaceScalarField r("flux(" + vf.name() + ")", mesh);
    for (label f = 0; f < ni; ++f) {
        const label own = mesh.owner_[f], nei = mesh.neighbour_[f];
        const vector d = mesh.C()[nei] - mesh.C()[own];
        const scalar rr = nvdtvd_r(phi[f], vf[own], vf[nei], gradc[own], gradc[nei], d);
        const scalar limiter = (rr + mag(rr)) / (1 + mag(rr)); // vanLeer
        const scalar w = limiter * mesh.weights_[f] + (1.0 - limiter) * pos(phi[f]);
        r[f] = phi[f] * (w * (vf[own] - vf[nei]) + vf[nei]);
    }
    // uncoupled patches: the limiter is 1 and the face value the patch value
    for (label b = 0; b < (label)mesh.boundaryFaceCells_.size(); ++b) r[ni + b] = phi[ni + b] * vf.patchValue(b);
    return r;
}
} // namespace fvc

namespace MULES {

// MULESTemplates.C: limiter().  rho = geometricOneField, Sp = zeroField (what MOC.C:240-249 and
// QMOM.C:173-182 pass); no coupled patches, no moving mesh, smoothLimiter = 0.
void limiter(scalarField &allLambda, const scalar rDeltaT, const volScalarField &psi, const surfaceScalarField &phiBD,
             const surfaceScalarField &phiCorr, const volScalarField &Su, const scalar psiMax, const scalar psiMin) {
    const fvMesh &mesh = psi.mesh();
    const scalarField &psiIf = psi.internalField();
    const label nLimiterIter = mesh.nLimiterIter_;
    const volScalarField &psi0f = psi.oldTime();
    const std::vector<label> &owner = mesh.owner_, &neighb = mesh.neighbour_, &bfc = mesh.boundaryFaceCells_;
    const scalarField &V = mesh.Vsc();
    const label ni = mesh.nInternalFaces(), nb = (label)bfc.size(), nc = psiIf.size();

    scalarField psiMaxn(nc, psiMin), psiMinn(nc, psiMax);
    scalarField sumPhiBD(nc, 0.0), sumPhip(nc, VSMALL), mSumPhim(nc, VSMALL);

    for (label facei = 0; facei < ni; ++facei) {
        const label own = owner[facei], nei = neighb[facei];
        psiMaxn[own] = max(psiMaxn[own], psiIf[nei]);
        psiMinn[own] = min(psiMinn[own], psiIf[nei]);
        psiMaxn[nei] = max(psiMaxn[nei], psiIf[own]);
        psiMinn[nei] = min(psiMinn[nei], psiIf[own]);
        sumPhiBD[own] += phiBD[facei];
        sumPhiBD[nei] -= phiBD[facei];
        const scalar phiCorrf = phiCorr[facei];
        if (phiCorrf > 0.0) {
            sumPhip[own] += phiCorrf;
            mSumPhim[nei] += phiCorrf;
        } else {
            mSumPhim[own] -= phiCorrf;
            sumPhip[nei] -= phiCorrf;
        }
    }
    for (label b = 0; b < nb; ++b) { // uncoupled patches: the patch value bounds the cell
        const label pfCelli = bfc[b];
        const scalar psiPf = psi.patchValue(b);
        psiMaxn[pfCelli] = max(psiMaxn[pfCelli], psiPf);
        psiMinn[pfCelli] = min(psiMinn[pfCelli], psiPf);
    }
    for (label b = 0; b < nb; ++b) {
        const label pfCelli = bfc[b];
        sumPhiBD[pfCelli] += phiBD[ni + b];
        const scalar phiCorrf = phiCorr[ni + b];
        if (phiCorrf > 0.0)
            sumPhip[pfCelli] += phiCorrf;
        else
            mSumPhim[pfCelli] -= phiCorrf;
    }
    for (label c = 0; c < nc; ++c) {
        psiMaxn[c] = min(psiMaxn[c], psiMax);
        psiMinn[c] = max(psiMinn[c], psiMin);
    }
    // psiMaxn = V*((rho*rDeltaT - Sp)*psiMaxn - Su - rho*rDeltaT*psi0) + sumPhiBD, rho = 1, Sp = 0
    for (label c = 0; c < nc; ++c) {
        psiMaxn[c] = V[c] * ((rDeltaT - 0.0) * psiMaxn[c] - Su[c] - rDeltaT * psi0f[c]) + sumPhiBD[c];
        psiMinn[c] = V[c] * (Su[c] - (rDeltaT - 0.0) * psiMinn[c] + rDeltaT * psi0f[c]) - sumPhiBD[c];
    }

    scalarField sumlPhip(nc), mSumlPhim(nc);
    for (int j = 0; j < nLimiterIter; j++) {
        sumlPhip = 0.0;
        mSumlPhim = 0.0;
        for (label facei = 0; facei < ni; ++facei) {
            const label own = owner[facei], nei = neighb[facei];
            const scalar lambdaPhiCorrf = allLambda[facei] * phiCorr[facei];
            if (lambdaPhiCorrf > 0.0) {
                sumlPhip[own] += lambdaPhiCorrf;
                mSumlPhim[nei] += lambdaPhiCorrf;
            } else {
                mSumlPhim[own] -= lambdaPhiCorrf;
                sumlPhip[nei] -= lambdaPhiCorrf;
            }
        }
        for (label b = 0; b < nb; ++b) {
            const label pfCelli = bfc[b];
            const scalar lambdaPhiCorrf = allLambda[ni + b] * phiCorr[ni + b];
            if (lambdaPhiCorrf > 0.0)
                sumlPhip[pfCelli] += lambdaPhiCorrf;
            else
                mSumlPhim[pfCelli] -= lambdaPhiCorrf;
        }
        for (label celli = 0; celli < nc; ++celli) {
            sumlPhip[celli] = max(min((sumlPhip[celli] + psiMaxn[celli]) / (mSumPhim[celli] - SMALL), 1.0), 0.0);
            mSumlPhim[celli] = max(min((mSumlPhim[celli] + psiMinn[celli]) / (sumPhip[celli] + SMALL), 1.0), 0.0);
        }
        const scalarField &lambdam = sumlPhip;
        const scalarField &lambdap = mSumlPhim;
        for (label facei = 0; facei < ni; ++facei) {
            if (phiCorr[facei] > 0.0)
                allLambda[facei] = min(allLambda[facei], min(lambdap[owner[facei]], lambdam[neighb[facei]]));
            else
                allLambda[facei] = min(allLambda[facei], min(lambdam[owner[facei]], lambdap[neighb[facei]]));
        }
        for (label b = 0; b < nb; ++b) { // uncoupled patch: limit outlet faces only
            if ((phiBD[ni + b] + phiCorr[ni + b]) > SMALL * SMALL) {
                const label pfCelli = bfc[b];
                if (phiCorr[ni + b] > 0.0)
                    allLambda[ni + b] = min(allLambda[ni + b], lambdap[pfCelli]);
                else
                    allLambda[ni + b] = min(allLambda[ni + b], lambdam[pfCelli]);
            }
        }
    }
}

// MULESTemplates.C: limit()
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

// MULES.C: limitSum() on the internal faces (boundary patches only when coupled: none here)
void limitSum(PtrList<surfaceScalarField> &phiPsiCorrs) {
    if (phiPsiCorrs.size() == 0) return;
    const label ni = phiPsiCorrs[0].mesh().nInternalFaces();
    for (label facei = 0; facei < ni; ++facei) {
        scalar sumPos = 0, sumNeg = 0;
        for (int phasei = 0; phasei < phiPsiCorrs.size(); phasei++) {
            if (phiPsiCorrs[phasei][facei] > 0)
                sumPos += phiPsiCorrs[phasei][facei];
            else
                sumNeg += phiPsiCorrs[phasei][facei];
        }
        const scalar sum = sumPos + sumNeg;
        if (sum > 0 && sumPos > VSMALL) {
            const scalar lambda = -sumNeg / sumPos;
            for (int phasei = 0; phasei < phiPsiCorrs.size(); phasei++)
                if (phiPsiCorrs[phasei][facei] > 0) phiPsiCorrs[phasei][facei] *= lambda;
        } else if (sum < 0 && sumNeg < -VSMALL) {
            const scalar lambda = -sumPos / sumNeg;
            for (int phasei = 0; phasei < phiPsiCorrs.size(); phasei++)
                if (phiPsiCorrs[phasei][facei] < 0) phiPsiCorrs[phasei][facei] *= lambda;
        }
    }
}

// MULESTemplates.C: explicitSolve() with rho = 1, Sp = 0:
//   psi = (psi0 rDeltaT + Su - surfaceIntegrate(phiPsi)) / rDeltaT,  rDeltaT = 1 / deltaTValue
void explicitSolve(const geometricOneField &, volScalarField &psi, const surfaceScalarField &phiPsi, const zeroField &,
                   const volScalarField &Su) {
    Info << "MULES: Solving for " << psi.name() << endl;
    refProbe::record(psi.name(), Su.internalField());
    const fvMesh &mesh = psi.mesh();
    const scalar rDeltaT = 1.0 / mesh.time().deltaTValue();
    scalarField psi0(psi.internalField());
    scalarField &psiIf = psi.internalField();
    psiIf = 0.0;
    fvc::surfaceIntegrate(psiIf, phiPsi);
    for (label c = 0; c < psiIf.size(); ++c) psiIf[c] = (psi0[c] * rDeltaT + Su[c] - psiIf[c]) / rDeltaT;
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
