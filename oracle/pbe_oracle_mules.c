/* pbe_oracle_mules.c -- TEST INFRASTRUCTURE: CPU restatement of the OpenFOAM 3.0.x pieces that
 * MOC::solveWithMULES (src/PBESystem/MOC.C:198-280) and QMOM::solveWithMULES (QMOM.C:131-197) call:
 *   fvc::flux(phi, psi, "div(phi,psi)") for `Gauss upwind` and `Gauss vanLeer`  (MOC.C:209-212)
 *   MULES::limit -> limiter (nLimiterIter sweeps), returnCorr = true              (MOC.C:240-249)
 *   MULES::limitSum over the fields                                                (MOC.C:252)
 *   upwind flux + MULES::explicitSolve(1, psi, phiPsi, 0, Su)                       (MOC.C:263-270)
 * OpenFOAM is a third-party dependency absent from the reference tree and from this image, so this
 * is written from OpenFOAM 3.0.x's published sources (finiteVolume/fvMatrices/solvers/MULES/
 * {MULES.C,MULESTemplates.C}, interpolation/surfaceInterpolation/{schemes/upwind/upwind.H,
 * limitedSchemes/{limitedScheme/limitedScheme.C,vanLeer/vanLeer.H,LimitedScheme/NVDTVD.H}},
 * finiteVolume/{fvc/fvcSurfaceIntegrate.C,gradSchemes/gaussGrad/gaussGrad.C}) and its parity is
 * UNPINNED: no assertion in the reference exercises a non-zero flux.  The SAME functions serve
 * (a) the OpenFOAM API shim that the reference's own MOC.C / QMOM.C are compiled against
 * (oracle/ref_shim/refFoam.C forwards MULES::*, upwind, fvc::flux here), so the reference's call
 * sequence drives them, and (b) orc_transport_mules below, the port of that call sequence which the
 * CUDA path is compared with.  tests/test_ref_oracle.py checks (a) == (b) bit for bit.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may use anything under oracle/.
 *
 * Uncoupled patches only (no processor / cyclic patches), static mesh, rho = 1, Sp = 0,
 * smoothLimiter = 0 -- what the reference passes and what its cases select. */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "pbe_oracle.h"

#define ORC_SMALL 1.0e-15   /* OpenFOAM SMALL, double precision */
#define ORC_VSMALL 1.0e-300 /* OpenFOAM VSMALL */

static inline double dmax(double a, double b) { return a > b ? a : b; }
static inline double dmin(double a, double b) { return a < b ? a : b; }
static inline double pos_(double s) { return (s >= 0) ? 1.0 : 0.0; }
static inline double sign_(double s) { return (s >= 0) ? 1.0 : -1.0; }

/* patch value of boundary face b: psi_b (fixed value) or the owner cell's (zero gradient) */
static inline double patch_value(const orc_mesh *m, const double *psi, const double *psi_b, int64_t b) {
    return psi_b ? psi_b[b] : psi[m->bown[b]];
}

/* upwind<scalar>(mesh, phi).flux(psi): phi_f (lambda (P - N) + N), lambda = pos(phi_f); patch value
 * on boundary faces (surfaceInterpolationScheme::interpolate; upwind.H) */
void orc_upwind_flux(const orc_mesh *m, const double *phi, const double *psi, const double *psi_b, double *out) {
    for (int64_t f = 0; f < m->nint; ++f) {
        const double lambda = pos_(phi[f]);
        const double P = psi[m->owner[f]], N = psi[m->neighbour[f]];
        out[f] = phi[f] * (lambda * (P - N) + N);
    }
    for (int64_t b = 0; b < m->nbnd; ++b) out[m->nint + b] = phi[m->nint + b] * patch_value(m, psi, psi_b, b);
}

/* gaussGrad with linear interpolation: grad_c = (sum_f Sf psi_f) / V */
void orc_gauss_linear_grad(const orc_mesh *m, const double *psi, const double *psi_b, double *grad /* [ncells][3] */) {
    memset(grad, 0, sizeof(double) * 3 * (size_t)m->ncells);
    for (int64_t f = 0; f < m->nint; ++f) {
        const int32_t own = m->owner[f], nei = m->neighbour[f];
        const double ssf = m->weights[f] * (psi[own] - psi[nei]) + psi[nei];
        for (int k = 0; k < 3; ++k) {
            const double Sfssf = m->Sf[3 * f + k] * ssf;
            grad[3 * (size_t)own + k] += Sfssf;
            grad[3 * (size_t)nei + k] -= Sfssf;
        }
    }
    for (int64_t b = 0; b < m->nbnd; ++b) {
        const int32_t c = m->bown[b];
        const double pv = patch_value(m, psi, psi_b, b);
        for (int k = 0; k < 3; ++k) grad[3 * (size_t)c + k] += m->Sf[3 * (m->nint + b) + k] * pv;
    }
    for (int64_t c = 0; c < m->ncells; ++c)
        for (int k = 0; k < 3; ++k) grad[3 * (size_t)c + k] /= m->V[c];
}

/* NVDTVD::r and the vanLeer limiter (r + |r|) / (1 + |r|) */
static inline double vanleer_limiter(double faceFlux, double phiP, double phiN, const double *gradcP,
                                     const double *gradcN, const double *d) {
    const double gradf = phiN - phiP;
    const double *g = (faceFlux > 0) ? gradcP : gradcN;
    const double gradcf = d[0] * g[0] + d[1] * g[1] + d[2] * g[2];
    double r;
    if (fabs(gradcf) >= 1000 * fabs(gradf))
        r = 2 * 1000 * sign_(gradcf) * sign_(gradf) - 1;
    else
        r = 2 * (gradcf / gradf) - 1;
    return (r + fabs(r)) / (1 + fabs(r));
}

/* fvc::flux(phi, psi, scheme) for `Gauss vanLeer`: phi_f (w (P - N) + N),
 * w = limiter w_linear + (1 - limiter) pos(phi_f); limiter = 1 on uncoupled patches (patch value) */
void orc_vanleer_flux(const orc_mesh *m, const double *phi, const double *psi, const double *psi_b, double *out) {
    double *grad = (double *)malloc(sizeof(double) * 3 * (size_t)(m->ncells > 0 ? m->ncells : 1));
    orc_gauss_linear_grad(m, psi, psi_b, grad);
    for (int64_t f = 0; f < m->nint; ++f) {
        const int32_t own = m->owner[f], nei = m->neighbour[f];
        const double d[3] = {m->C[3 * (size_t)nei] - m->C[3 * (size_t)own], m->C[3 * (size_t)nei + 1] - m->C[3 * (size_t)own + 1],
                             m->C[3 * (size_t)nei + 2] - m->C[3 * (size_t)own + 2]};
        const double lim = vanleer_limiter(phi[f], psi[own], psi[nei], grad + 3 * (size_t)own, grad + 3 * (size_t)nei, d);
        const double w = lim * m->weights[f] + (1.0 - lim) * pos_(phi[f]);
        out[f] = phi[f] * (w * (psi[own] - psi[nei]) + psi[nei]);
    }
    for (int64_t b = 0; b < m->nbnd; ++b) out[m->nint + b] = phi[m->nint + b] * patch_value(m, psi, psi_b, b);
    free(grad);
}

int orc_div_flux(const orc_mesh *m, int scheme, const double *phi, const double *psi, const double *psi_b, double *out) {
    if (scheme == ORC_DIV_UPWIND) {
        orc_upwind_flux(m, phi, psi, psi_b, out);
        return 0;
    }
    if (scheme == ORC_DIV_VANLEER) {
        if (!m->Sf || !m->C || !m->weights) return 2;
        orc_vanleer_flux(m, phi, psi, psi_b, out);
        return 0;
    }
    return 1;
}

/* MULESTemplates.C limiter(): lambda [nfaces] in (all 1 on entry), limited on exit */
void orc_mules_limiter(const orc_mesh *m, double *lambda, double rDeltaT, const double *psi, const double *psi_b,
                       const double *phiBD, const double *phiCorr, const double *Su, double psiMax, double psiMin,
                       int nLimiterIter) {
    const int64_t nc = m->ncells, ni = m->nint, nb = m->nbnd;
    const double *psi0 = psi; /* oldTime(): nothing touched the field since the time step began */
    double *w = (double *)malloc(sizeof(double) * 7 * (size_t)(nc > 0 ? nc : 1));
    double *psiMaxn = w, *psiMinn = w + nc, *sumPhiBD = w + 2 * nc, *sumPhip = w + 3 * nc, *mSumPhim = w + 4 * nc,
           *sumlPhip = w + 5 * nc, *mSumlPhim = w + 6 * nc;
    for (int64_t c = 0; c < nc; ++c) {
        psiMaxn[c] = psiMin;
        psiMinn[c] = psiMax;
        sumPhiBD[c] = 0.0;
        sumPhip[c] = ORC_VSMALL;
        mSumPhim[c] = ORC_VSMALL;
    }
    for (int64_t f = 0; f < ni; ++f) {
        const int32_t own = m->owner[f], nei = m->neighbour[f];
        psiMaxn[own] = dmax(psiMaxn[own], psi[nei]);
        psiMinn[own] = dmin(psiMinn[own], psi[nei]);
        psiMaxn[nei] = dmax(psiMaxn[nei], psi[own]);
        psiMinn[nei] = dmin(psiMinn[nei], psi[own]);
        sumPhiBD[own] += phiBD[f];
        sumPhiBD[nei] -= phiBD[f];
        const double phiCorrf = phiCorr[f];
        if (phiCorrf > 0.0) {
            sumPhip[own] += phiCorrf;
            mSumPhim[nei] += phiCorrf;
        } else {
            mSumPhim[own] -= phiCorrf;
            sumPhip[nei] -= phiCorrf;
        }
    }
    for (int64_t b = 0; b < nb; ++b) {
        const int32_t c = m->bown[b];
        const double pv = patch_value(m, psi, psi_b, b);
        psiMaxn[c] = dmax(psiMaxn[c], pv);
        psiMinn[c] = dmin(psiMinn[c], pv);
    }
    for (int64_t b = 0; b < nb; ++b) {
        const int32_t c = m->bown[b];
        sumPhiBD[c] += phiBD[ni + b];
        const double phiCorrf = phiCorr[ni + b];
        if (phiCorrf > 0.0)
            sumPhip[c] += phiCorrf;
        else
            mSumPhim[c] -= phiCorrf;
    }
    for (int64_t c = 0; c < nc; ++c) {
        psiMaxn[c] = dmin(psiMaxn[c], psiMax);
        psiMinn[c] = dmax(psiMinn[c], psiMin);
    }
    /* V ((rho rDeltaT - Sp) psiMaxn - Su - rho rDeltaT psi0) + sumPhiBD with rho = 1, Sp = 0 */
    for (int64_t c = 0; c < nc; ++c) {
        psiMaxn[c] = m->V[c] * ((rDeltaT - 0.0) * psiMaxn[c] - Su[c] - rDeltaT * psi0[c]) + sumPhiBD[c];
        psiMinn[c] = m->V[c] * (Su[c] - (rDeltaT - 0.0) * psiMinn[c] + rDeltaT * psi0[c]) - sumPhiBD[c];
    }
    for (int it = 0; it < nLimiterIter; ++it) {
        for (int64_t c = 0; c < nc; ++c) sumlPhip[c] = mSumlPhim[c] = 0.0;
        for (int64_t f = 0; f < ni; ++f) {
            const int32_t own = m->owner[f], nei = m->neighbour[f];
            const double lambdaPhiCorrf = lambda[f] * phiCorr[f];
            if (lambdaPhiCorrf > 0.0) {
                sumlPhip[own] += lambdaPhiCorrf;
                mSumlPhim[nei] += lambdaPhiCorrf;
            } else {
                mSumlPhim[own] -= lambdaPhiCorrf;
                sumlPhip[nei] -= lambdaPhiCorrf;
            }
        }
        for (int64_t b = 0; b < nb; ++b) {
            const int32_t c = m->bown[b];
            const double lambdaPhiCorrf = lambda[ni + b] * phiCorr[ni + b];
            if (lambdaPhiCorrf > 0.0)
                sumlPhip[c] += lambdaPhiCorrf;
            else
                mSumlPhim[c] -= lambdaPhiCorrf;
        }
        for (int64_t c = 0; c < nc; ++c) {
            sumlPhip[c] = dmax(dmin((sumlPhip[c] + psiMaxn[c]) / (mSumPhim[c] - ORC_SMALL), 1.0), 0.0);
            mSumlPhim[c] = dmax(dmin((mSumlPhim[c] + psiMinn[c]) / (sumPhip[c] + ORC_SMALL), 1.0), 0.0);
        }
        const double *lambdam = sumlPhip, *lambdap = mSumlPhim;
        for (int64_t f = 0; f < ni; ++f) {
            if (phiCorr[f] > 0.0)
                lambda[f] = dmin(lambda[f], dmin(lambdap[m->owner[f]], lambdam[m->neighbour[f]]));
            else
                lambda[f] = dmin(lambda[f], dmin(lambdam[m->owner[f]], lambdap[m->neighbour[f]]));
        }
        for (int64_t b = 0; b < nb; ++b) { /* uncoupled patch: limit outlet faces only */
            if ((phiBD[ni + b] + phiCorr[ni + b]) > ORC_SMALL * ORC_SMALL) {
                const int32_t c = m->bown[b];
                if (phiCorr[ni + b] > 0.0)
                    lambda[ni + b] = dmin(lambda[ni + b], lambdap[c]);
                else
                    lambda[ni + b] = dmin(lambda[ni + b], lambdam[c]);
            }
        }
    }
    free(w);
}

/* MULESTemplates.C limit(..., returnCorr = true): phiPsi holds the high-order flux on entry and the
 * limited correction lambda (phiHO - phiBD) on exit */
void orc_mules_limit(const orc_mesh *m, double rDeltaT, const double *psi, const double *psi_b, const double *phi,
                     double *phiPsi, const double *Su, double psiMax, double psiMin, int nLimiterIter) {
    const int64_t nf = m->nint + m->nbnd;
    double *phiBD = (double *)malloc(sizeof(double) * 2 * (size_t)(nf > 0 ? nf : 1));
    double *lambda = phiBD + nf;
    orc_upwind_flux(m, phi, psi, psi_b, phiBD);
    for (int64_t f = 0; f < nf; ++f) {
        phiPsi[f] -= phiBD[f];
        lambda[f] = 1.0;
    }
    orc_mules_limiter(m, lambda, rDeltaT, psi, psi_b, phiBD, phiPsi, Su, psiMax, psiMin, nLimiterIter);
    for (int64_t f = 0; f < nf; ++f) phiPsi[f] *= lambda[f];
    free(phiBD);
}

/* MULES.C limitSum(): internal faces (patches only when coupled: none here); corr [nfields][stride_f] */
void orc_mules_limit_sum(int nfields, int64_t nint, int64_t stride_f, double *corr) {
    for (int64_t f = 0; f < nint; ++f) {
        double sumPos = 0, sumNeg = 0;
        for (int k = 0; k < nfields; ++k) {
            const double v = corr[(size_t)k * stride_f + f];
            if (v > 0)
                sumPos += v;
            else
                sumNeg += v;
        }
        const double sum = sumPos + sumNeg;
        if (sum > 0 && sumPos > ORC_VSMALL) {
            const double lambda = -sumNeg / sumPos;
            for (int k = 0; k < nfields; ++k)
                if (corr[(size_t)k * stride_f + f] > 0) corr[(size_t)k * stride_f + f] *= lambda;
        } else if (sum < 0 && sumNeg < -ORC_VSMALL) {
            const double lambda = -sumPos / sumNeg;
            for (int k = 0; k < nfields; ++k)
                if (corr[(size_t)k * stride_f + f] < 0) corr[(size_t)k * stride_f + f] *= lambda;
        }
    }
}

/* MULESTemplates.C explicitSolve() with rho = 1, Sp = 0: psi <- (psi0 rDeltaT + Su - surfaceIntegrate(phiPsi)) / rDeltaT */
void orc_mules_explicit_solve(const orc_mesh *m, double rDeltaT, double *psi, const double *phiPsi, const double *Su) {
    const int64_t nc = m->ncells;
    double *div = (double *)calloc((size_t)(nc > 0 ? nc : 1), sizeof(double));
    for (int64_t f = 0; f < m->nint; ++f) {
        div[m->owner[f]] += phiPsi[f];
        div[m->neighbour[f]] -= phiPsi[f];
    }
    for (int64_t b = 0; b < m->nbnd; ++b) div[m->bown[b]] += phiPsi[m->nint + b];
    for (int64_t c = 0; c < nc; ++c) {
        div[c] /= m->V[c];
        psi[c] = (psi[c] * rDeltaT + (Su ? Su[c] : 0.0) - div[c]) / rDeltaT;
    }
    free(div);
}

/* The call sequence of MOC::solveWithMULES (MOC.C:198-272) / QMOM::solveWithMULES (QMOM.C:131-197) on
 * nfields fields psi [nfields][stride] (updated in place), S [nfields][stride], psi_b [nfields][stride_b]
 * patch values or NULL, psiMax [nfields] (MOC: max(alpha V / xi_k), MOC.C:238; QMOM: 1.2 max(m_k),
 * QMOM.C:171), psiMin = 0.  Returns 0, or non-zero for an unknown scheme / missing geometry. */
int orc_transport_mules(const orc_mesh *m, int scheme, int nfields, int64_t stride, double *psi, const double *phi,
                        const double *psi_b, int64_t stride_b, const double *S, const double *psiMax, double psiMin,
                        int nLimiterIter, double dt) {
    const int64_t nf = m->nint + m->nbnd;
    const double rDeltaT = 1.0 / dt;
    double *corr = (double *)malloc(sizeof(double) * (size_t)nfields * (size_t)(nf > 0 ? nf : 1));
    double *bd = (double *)malloc(sizeof(double) * (size_t)(nf > 0 ? nf : 1));
    int rc = 0;
    for (int k = 0; k < nfields && rc == 0; ++k) {
        double *pk = psi + (size_t)k * stride;
        const double *pb = psi_b ? psi_b + (size_t)k * stride_b : NULL;
        double *ck = corr + (size_t)k * nf;
        rc = orc_div_flux(m, scheme, phi, pk, pb, ck);                                       /* MOC.C:209-212 */
        if (rc) break;
        orc_mules_limit(m, rDeltaT, pk, pb, phi, ck, S + (size_t)k * stride, psiMax[k], psiMin, nLimiterIter); /* :240-249 */
    }
    if (rc == 0) {
        orc_mules_limit_sum(nfields, m->nint, nf, corr); /* MOC.C:252 */
        for (int k = 0; k < nfields; ++k) {
            double *pk = psi + (size_t)k * stride;
            const double *pb = psi_b ? psi_b + (size_t)k * stride_b : NULL;
            double *ck = corr + (size_t)k * nf;
            orc_upwind_flux(m, phi, pk, pb, bd);
            for (int64_t f = 0; f < nf; ++f) ck[f] += bd[f];                                  /* MOC.C:263 */
            orc_mules_explicit_solve(m, rDeltaT, pk, ck, S + (size_t)k * stride);           /* MOC.C:265-269 */
        }
    }
    free(corr);
    free(bd);
    return rc;
}
