// pbe_mules.cu -- the MULES-limited transport of the class / moment fields (SURVEY.md section 8 row
// f2, second slice): what MOC::solveWithMULES (src/PBESystem/MOC.C:198-280) and
// QMOM::solveWithMULES (src/PBESystem/QMOM.C:131-197) do for a limited divergence scheme -- the
// reference's validation case runs `div(phi,n.*) Gauss vanLeer`
// (tests/validation/simmons_and_azzopardi/c01/system/fvSchemes:34):
//
//   per field k:  phiCorr_k = fvc::flux(phi, psi_k, scheme)                           MOC.C:209-212
//                 MULES::limit(1/dt, 1, psi_k, phi, phiCorr_k, 0, S_k, psiMax_k, 0, true)   :240-249
//                    = phiCorr_k -= upwind flux; lambda by nLimiterIter limiter sweeps; phiCorr_k *= lambda
//   all fields:   MULES::limitSum(phiCorr)                                             MOC.C:252
//   per field k:  phiCorr_k += upwind flux; MULES::explicitSolve(1, psi_k, phiCorr_k, 0, S_k)  :263-269
//
// OpenFOAM's MULES / limitedScheme / gaussGrad are third-party code absent from the reference tree:
// the checker is the CPU restatement of OpenFOAM 3.0.x in the test infrastructure (PARITY UNPINNED),
// driven by the reference's own MOC.C through an API shim and by the port of its call sequence;
// this file follows the same operation order so that the device result is BIT-IDENTICAL to it
// (this translation unit is compiled with -fmad=false: no multiply-add contraction).
//
// Parallel form.  The reference's loops run over faces and scatter into cells; here every cell-sized
// quantity is gathered by one thread per cell walking the cell's faces in ascending face order
// (pbe_mesh's CSR: internal faces, then boundary faces) -- the order in which the face loops add --
// and every face-sized quantity by one thread per internal face.  On an uncoupled boundary face the
// high-order flux IS the upwind flux (the patch value), so its correction is exactly 0 and only the
// internal faces carry phiCorr / lambda.  grid.y runs over a chunk of fields.
//
//   HBM per cell-field (hex mesh, 3 internal faces per cell): phiCorr 24 B written once and read /
//   rewritten by the 2 x nLimiterIter sweeps, lambda 24 B likewise, 9 cell-sized temporaries:
//   ~20 x the 24 B of the upwind update -- the price of the limiter's iterations.
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "pbe_mesh.h"

using namespace pbe;

namespace {

constexpr double kSMALL = 1.0e-15;   // OpenFOAM SMALL, double precision
constexpr double kVSMALL = 1.0e-300; // OpenFOAM VSMALL

struct MulesParams {
    int64_t nc, nint, nbnd, stride, stride_b;
    const int32_t *csr_ptr, *csr_other, *csr_face, *owner, *neigh;
    const double *V, *Sf, *C, *w;
    const double *psi, *psi_b, *phi, *S, *psi_max;
    double psi_min, rdt;
    double *grad;                // [KC][3][nc]
    double *corr;                // [K][nint]
    double *lambda;              // [KC][nint]
    double *bmax, *bmin, *sp, *sm, *lm, *lp; // [KC][nc]
    double *out;
    int k0, scheme, last;
};

__device__ __forceinline__ double dmax_(double a, double b) { return a > b ? a : b; }
__device__ __forceinline__ double dmin_(double a, double b) { return a < b ? a : b; }
__device__ __forceinline__ double pos_(double s) { return (s >= 0) ? 1.0 : 0.0; }
__device__ __forceinline__ double sign_(double s) { return (s >= 0) ? 1.0 : -1.0; }

__device__ __forceinline__ double patch_value(const MulesParams &p, const double *psi_k, int k, int64_t b, int64_t c) {
    return p.psi_b ? __ldg(p.psi_b + (size_t)k * p.stride_b + b) : psi_k[c];
}

// gaussGrad, linear interpolation: grad_c = (sum_f Sf psi_f) / V
__global__ void __launch_bounds__(256) mules_grad_kernel(const MulesParams p) {
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= p.nc) return;
    const int kc = blockIdx.y, k = p.k0 + kc;
    const double *psi = p.psi + (size_t)k * p.stride;
    const int64_t nf = p.nint + p.nbnd;
    double g0 = 0.0, g1 = 0.0, g2 = 0.0;
    const double mine = psi[c];
    for (int e = p.csr_ptr[c]; e < p.csr_ptr[c + 1]; ++e) {
        const int32_t oth = p.csr_other[e], fc = p.csr_face[e];
        if (oth >= 0) {
            const bool own = fc >= 0;
            const int64_t f = own ? fc : ~fc;
            const double theirs = __ldg(psi + oth);
            const double P = own ? mine : theirs, N = own ? theirs : mine;
            const double ssf = __ldg(p.w + f) * (P - N) + N;
            const double t0 = __ldg(p.Sf + f) * ssf, t1 = __ldg(p.Sf + nf + f) * ssf, t2 = __ldg(p.Sf + 2 * nf + f) * ssf;
            g0 = own ? g0 + t0 : g0 - t0;
            g1 = own ? g1 + t1 : g1 - t1;
            g2 = own ? g2 + t2 : g2 - t2;
        } else {
            const int64_t b = -(int64_t)oth - 2;
            const double pv = patch_value(p, psi, k, b, c);
            g0 += __ldg(p.Sf + p.nint + b) * pv;
            g1 += __ldg(p.Sf + nf + p.nint + b) * pv;
            g2 += __ldg(p.Sf + 2 * nf + p.nint + b) * pv;
        }
    }
    const double V = __ldg(p.V + c);
    double *g = p.grad + (size_t)kc * 3 * p.nc;
    g[c] = g0 / V;
    g[p.nc + c] = g1 / V;
    g[2 * p.nc + c] = g2 / V;
}

__device__ __forceinline__ double upwind_flux(double ph, double P, double N) { return ph * (pos_(ph) * (P - N) + N); }

// per internal face: phiCorr = fvc::flux(phi, psi, scheme) - upwind flux; lambda = 1
__global__ void __launch_bounds__(256) mules_corr_kernel(const MulesParams p) {
    const int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= p.nint) return;
    const int kc = blockIdx.y, k = p.k0 + kc;
    const double *psi = p.psi + (size_t)k * p.stride;
    const int32_t own = __ldg(p.owner + f), nei = __ldg(p.neigh + f);
    const double P = __ldg(psi + own), N = __ldg(psi + nei), ph = __ldg(p.phi + f);
    const double bd = upwind_flux(ph, P, N);
    double ho = bd;
    if (p.scheme == PBE_DIV_VANLEER) { // limitedScheme<vanLeer>: NVDTVD::r, limiter (r + |r|) / (1 + |r|)
        const double *g = p.grad + (size_t)kc * 3 * p.nc;
        const double d0 = __ldg(p.C + nei) - __ldg(p.C + own), d1 = __ldg(p.C + p.nc + nei) - __ldg(p.C + p.nc + own),
                     d2 = __ldg(p.C + 2 * p.nc + nei) - __ldg(p.C + 2 * p.nc + own);
        const int32_t up = (ph > 0) ? own : nei;
        const double gradcf = d0 * g[up] + d1 * g[p.nc + up] + d2 * g[2 * p.nc + up];
        const double gradf = N - P;
        double r;
        if (fabs(gradcf) >= 1000 * fabs(gradf))
            r = 2 * 1000 * sign_(gradcf) * sign_(gradf) - 1;
        else
            r = 2 * (gradcf / gradf) - 1;
        const double lim = (r + fabs(r)) / (1 + fabs(r));
        const double wgt = lim * __ldg(p.w + f) + (1.0 - lim) * pos_(ph);
        ho = ph * (wgt * (P - N) + N);
    }
    p.corr[(size_t)k * p.nint + f] = ho - bd;
    p.lambda[(size_t)kc * p.nint + f] = 1.0;
}

// MULESTemplates.C limiter(), the part before the sweeps: local extrema, sumPhiBD, sumPhip, mSumPhim
__global__ void __launch_bounds__(256) mules_bounds_kernel(const MulesParams p) {
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= p.nc) return;
    const int kc = blockIdx.y, k = p.k0 + kc;
    const double *psi = p.psi + (size_t)k * p.stride;
    const double *corr = p.corr + (size_t)k * p.nint;
    const double psiMax = __ldg(p.psi_max + k), psiMin = p.psi_min;
    const double mine = psi[c];
    double pmax = psiMin, pmin = psiMax, sBD = 0.0, sp = kVSMALL, sm = kVSMALL;
    for (int e = p.csr_ptr[c]; e < p.csr_ptr[c + 1]; ++e) {
        const int32_t oth = p.csr_other[e], fc = p.csr_face[e];
        if (oth >= 0) {
            const bool own = fc >= 0;
            const int64_t f = own ? fc : ~fc;
            const double theirs = __ldg(psi + oth);
            pmax = dmax_(pmax, theirs);
            pmin = dmin_(pmin, theirs);
            const double P = own ? mine : theirs, N = own ? theirs : mine;
            const double bd = upwind_flux(__ldg(p.phi + f), P, N);
            sBD = own ? sBD + bd : sBD - bd;
            const double cf = __ldg(corr + f);
            if (own) {
                if (cf > 0.0) sp += cf; else sm -= cf;
            } else {
                if (cf > 0.0) sm += cf; else sp -= cf;
            }
        } else {
            const int64_t b = -(int64_t)oth - 2;
            const double pv = patch_value(p, psi, k, b, c);
            pmax = dmax_(pmax, pv);
            pmin = dmin_(pmin, pv);
            sBD += __ldg(p.phi + p.nint + b) * pv; // the patch face's correction is exactly 0
        }
    }
    pmax = dmin_(pmax, psiMax);
    pmin = dmax_(pmin, psiMin);
    const double V = __ldg(p.V + c), Su = p.S ? __ldg(p.S + (size_t)k * p.stride + c) : 0.0, rdt = p.rdt;
    const size_t o = (size_t)kc * p.nc + c;
    p.bmax[o] = V * ((rdt - 0.0) * pmax - Su - rdt * mine) + sBD;
    p.bmin[o] = V * (Su - (rdt - 0.0) * pmin + rdt * mine) - sBD;
    p.sp[o] = sp;
    p.sm[o] = sm;
}

// one limiter sweep, cell half: lambdam, lambdap from the currently limited corrections
__global__ void __launch_bounds__(256) mules_sweep_cells_kernel(const MulesParams p) {
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= p.nc) return;
    const int kc = blockIdx.y, k = p.k0 + kc;
    const double *corr = p.corr + (size_t)k * p.nint;
    const double *lam = p.lambda + (size_t)kc * p.nint;
    double slp = 0.0, mslm = 0.0;
    for (int e = p.csr_ptr[c]; e < p.csr_ptr[c + 1]; ++e) {
        const int32_t oth = p.csr_other[e], fc = p.csr_face[e];
        if (oth < 0) continue; // lambda * 0
        const bool own = fc >= 0;
        const int64_t f = own ? fc : ~fc;
        const double l = __ldg(lam + f) * __ldg(corr + f);
        if (own) {
            if (l > 0.0) slp += l; else mslm -= l;
        } else {
            if (l > 0.0) mslm += l; else slp -= l;
        }
    }
    const size_t o = (size_t)kc * p.nc + c;
    p.lm[o] = dmax_(dmin_((slp + p.bmax[o]) / (p.sm[o] - kSMALL), 1.0), 0.0);
    p.lp[o] = dmax_(dmin_((mslm + p.bmin[o]) / (p.sp[o] + kSMALL), 1.0), 0.0);
}

// one limiter sweep, face half; after the last sweep phiCorr *= lambda (returnCorr = true)
__global__ void __launch_bounds__(256) mules_sweep_faces_kernel(const MulesParams p) {
    const int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= p.nint) return;
    const int kc = blockIdx.y, k = p.k0 + kc;
    const int32_t own = __ldg(p.owner + f), nei = __ldg(p.neigh + f);
    const double *lm = p.lm + (size_t)kc * p.nc, *lp = p.lp + (size_t)kc * p.nc;
    double *corr = p.corr + (size_t)k * p.nint + f;
    double *lam = p.lambda + (size_t)kc * p.nint + f;
    const double cf = *corr;
    double l = *lam;
    if (cf > 0.0)
        l = dmin_(l, dmin_(lp[own], lm[nei]));
    else
        l = dmin_(l, dmin_(lm[own], lp[nei]));
    if (p.last)
        *corr = cf * l;
    else
        *lam = l;
}

// MULES.C limitSum(): per internal face over all fields
__global__ void __launch_bounds__(256) mules_limit_sum_kernel(int nfields, int64_t nint, double *__restrict__ corr) {
    const int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= nint) return;
    double sumPos = 0, sumNeg = 0;
    for (int k = 0; k < nfields; ++k) {
        const double v = corr[(size_t)k * nint + f];
        if (v > 0) sumPos += v; else sumNeg += v;
    }
    const double sum = sumPos + sumNeg;
    if (sum > 0 && sumPos > kVSMALL) {
        const double lambda = -sumNeg / sumPos;
        for (int k = 0; k < nfields; ++k) {
            const double v = corr[(size_t)k * nint + f];
            if (v > 0) corr[(size_t)k * nint + f] = v * lambda;
        }
    } else if (sum < 0 && sumNeg < -kVSMALL) {
        const double lambda = -sumPos / sumNeg;
        for (int k = 0; k < nfields; ++k) {
            const double v = corr[(size_t)k * nint + f];
            if (v < 0) corr[(size_t)k * nint + f] = v * lambda;
        }
    }
}

// phiPsi = phiCorr + upwind flux; MULES::explicitSolve: psi <- (psi rDeltaT + Su - surfaceIntegrate(phiPsi)) / rDeltaT
__global__ void __launch_bounds__(256) mules_solve_kernel(const MulesParams p) {
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= p.nc) return;
    const int k = p.k0 + blockIdx.y;
    const double *psi = p.psi + (size_t)k * p.stride;
    const double *corr = p.corr + (size_t)k * p.nint;
    const double mine = psi[c];
    double acc = 0.0;
    for (int e = p.csr_ptr[c]; e < p.csr_ptr[c + 1]; ++e) {
        const int32_t oth = p.csr_other[e], fc = p.csr_face[e];
        if (oth >= 0) {
            const bool own = fc >= 0;
            const int64_t f = own ? fc : ~fc;
            const double theirs = __ldg(psi + oth);
            const double P = own ? mine : theirs, N = own ? theirs : mine;
            const double v = __ldg(corr + f) + upwind_flux(__ldg(p.phi + f), P, N);
            acc = own ? acc + v : acc - v;
        } else {
            const int64_t b = -(int64_t)oth - 2;
            acc += __ldg(p.phi + p.nint + b) * patch_value(p, psi, k, b, c);
        }
    }
    const double div = acc / __ldg(p.V + c);
    const double Su = p.S ? __ldg(p.S + (size_t)k * p.stride + c) : 0.0;
    p.out[(size_t)k * p.stride + c] = (mine * p.rdt + Su - div) / p.rdt;
}


// ======================================================================================================
// Fused form (default; PBE_MULES_PATH=0 selects the per-field kernels above, which are kept as the
// in-library cross-check).  ncu on the per-field kernels (profiles/r02m_mules_launches.csv): they run at
// 3-5.6 TB/s of DRAM traffic but move 1 410 B per cell-field -- 330 B of it CSR / owner / neighbour indices
// and 170 B geometry re-read for every field, 450 B the per-face lambda array read and rewritten by every
// sweep.  Here (a) every kernel loops over a chunk of fields with the cell's / face's addressing in
// registers, (b) lambda is never stored: it only ever decreases, lambda_f after sweep s = min over the sweeps
// s' <= s of min(lp_s'[.], lm_s'[.]) at the face's two cells, so a sweep's cell half rebuilds the lambdas of
// its faces from the cell-sized lp / lm rows of the previous sweeps (same min sequence: bit-identical),
// (c) the first sweep's cell half (lambda = 1) is fused with the bounds, (d) limitSum only READS the limited
// corrections and leaves two scaling factors per face, which the solve kernel applies -- the limited
// corrections are written once.
// Thread mapping of the fused kernels: blockDim = (32, 8): a warp is 32 consecutive cells (or faces) of ONE
// field (coalesced rows), a CTA covers 8 fields of the chunk, so the CSR / owner / neighbour / geometry words of
// a cell are requested by 8 warps of the same CTA at about the same time and come out of L1 / L2 -- the
// addressing is amortised over the fields without holding it in registers.  (Measured alternatives at 160^3 x
// 100 fields: per-thread loops over 4-8 fields with the addressing in registers, faces walked from the CSR,
// 106-111 ms; the same with every per-face load made unconditional, 110 registers, 23 % occupancy: 109-116 ms;
// the per-field kernels above: 135 ms.)
// resident 256-thread CTAs per SM the fused kernels are compiled for (40 registers): 124 -> 110 ms at 160^3 x 100 fields
#ifndef PBE_MULES_MINB
#define PBE_MULES_MINB 6
#endif
#ifndef PBE_MULES_BX
#define PBE_MULES_BX 32
#endif
constexpr int kFusedBx = PBE_MULES_BX, kFusedBy = 256 / PBE_MULES_BX;

struct FusedParams {
    MulesParams m;
    int kn;          // fields in this chunk (m.k0 ..)
    int nsweeps;     // nLimiterIter
    int sweep;       // 1-based sweep whose cell half runs (fused_sweep_kernel)
    double2 *lml;    // [nsweeps][KC][nc]: {lm, lp} of every sweep (one 16-byte gather gives both)
    int KC;
    double *sums;    // [2][nint]: limitSum's sumPos, sumNeg; turned into its two scaling factors before the solve
};
__device__ __forceinline__ double2 *lml_row(const FusedParams &q, int s, int kc) {
    return q.lml + ((size_t)s * q.KC + kc) * q.m.nc;
}
// one replayed face-half update (mules_sweep_faces_kernel): o / n = {lm, lp} of the face's owner / neighbour
__device__ __forceinline__ double lambda_step(double l, double cf, double2 o, double2 n) {
    return (cf > 0.0) ? dmin_(l, dmin_(o.y, n.x)) : dmin_(l, dmin_(o.x, n.y));
}
// lambda of a face after `upto` sweeps: cf = the unlimited correction, own / nei = the face's cells.  The
// gathers do not depend on cf; with the sweep count a compile-time constant they are all issued together.
template <int UPTO>
__device__ __forceinline__ double face_lambda_n(const FusedParams &q, int kc, double cf, int64_t own, int64_t nei) {
    double2 o[UPTO > 0 ? UPTO : 1], n[UPTO > 0 ? UPTO : 1];
#pragma unroll
    for (int s = 0; s < UPTO; ++s) {
        const double2 *row = lml_row(q, s, kc);
        o[s] = __ldg(row + own), n[s] = __ldg(row + nei);
    }
    double l = 1.0;
#pragma unroll
    for (int s = 0; s < UPTO; ++s) l = lambda_step(l, cf, o[s], n[s]);
    return l;
}
__device__ __forceinline__ double face_lambda(const FusedParams &q, int kc, int upto, double cf, int64_t own, int64_t nei) {
    switch (upto) {
    case 0: return 1.0;
    case 1: return face_lambda_n<1>(q, kc, cf, own, nei);
    case 2: return face_lambda_n<2>(q, kc, cf, own, nei);
    case 3: return face_lambda_n<3>(q, kc, cf, own, nei);
    default: break;
    }
    double l = 1.0;
    for (int s = 0; s < upto; ++s) {
        const double2 *row = lml_row(q, s, kc);
        l = lambda_step(l, cf, __ldg(row + own), __ldg(row + nei));
    }
    return l;
}
#define FUSED_CELL_THREAD()                                                  \
    const MulesParams &p = q.m;                                              \
    const int64_t c = (int64_t)blockIdx.x * kFusedBx + threadIdx.x;          \
    const int kc = blockIdx.y * kFusedBy + threadIdx.y;                      \
    if (c >= p.nc || kc >= q.kn) return;                                     \
    const int k = p.k0 + kc
#define FUSED_FACE_THREAD()                                                  \
    const MulesParams &p = q.m;                                              \
    const int64_t f = (int64_t)blockIdx.x * kFusedBx + threadIdx.x;          \
    const int kc = blockIdx.y * kFusedBy + threadIdx.y;                      \
    if (f >= p.nint || kc >= q.kn) return;                                   \
    const int k = p.k0 + kc

__global__ void __launch_bounds__(256, PBE_MULES_MINB) fused_grad_kernel(const FusedParams q) {
    FUSED_CELL_THREAD();
    const double *psi = p.psi + (size_t)k * p.stride;
    const int64_t nf = p.nint + p.nbnd;
    double g0 = 0.0, g1 = 0.0, g2 = 0.0;
    const double mine = psi[c];
#pragma unroll 3
    for (int e = p.csr_ptr[c]; e < p.csr_ptr[c + 1]; ++e) {
        const int32_t oth = p.csr_other[e], fc = p.csr_face[e];
        if (oth >= 0) {
            const bool own = fc >= 0;
            const int64_t f = own ? fc : ~fc;
            const double theirs = __ldg(psi + oth);
            const double P = own ? mine : theirs, N = own ? theirs : mine;
            const double ssf = __ldg(p.w + f) * (P - N) + N;
            const double t0 = __ldg(p.Sf + f) * ssf, t1 = __ldg(p.Sf + nf + f) * ssf, t2 = __ldg(p.Sf + 2 * nf + f) * ssf;
            g0 = own ? g0 + t0 : g0 - t0;
            g1 = own ? g1 + t1 : g1 - t1;
            g2 = own ? g2 + t2 : g2 - t2;
        } else {
            const int64_t b = -(int64_t)oth - 2;
            const double pv = patch_value(p, psi, k, b, c);
            g0 += __ldg(p.Sf + p.nint + b) * pv;
            g1 += __ldg(p.Sf + nf + p.nint + b) * pv;
            g2 += __ldg(p.Sf + 2 * nf + p.nint + b) * pv;
        }
    }
    const double V = __ldg(p.V + c);
    double *g = p.grad + (size_t)kc * 3 * p.nc;
    g[c] = g0 / V;
    g[p.nc + c] = g1 / V;
    g[2 * p.nc + c] = g2 / V;
}

// per internal face: the unlimited correction phiCorr = fvc::flux(phi, psi, scheme) - upwind flux
__global__ void __launch_bounds__(256, PBE_MULES_MINB) fused_corr_kernel(const FusedParams q) {
    FUSED_FACE_THREAD();
    const double *psi = p.psi + (size_t)k * p.stride;
    const int32_t own = __ldg(p.owner + f), nei = __ldg(p.neigh + f);
    const double P = __ldg(psi + own), N = __ldg(psi + nei), ph = __ldg(p.phi + f);
    const double bd = upwind_flux(ph, P, N);
    double ho = bd;
    if (p.scheme == PBE_DIV_VANLEER) { // limitedScheme<vanLeer>: NVDTVD::r, limiter (r + |r|) / (1 + |r|)
        const double *g = p.grad + (size_t)kc * 3 * p.nc;
        const double d0 = __ldg(p.C + nei) - __ldg(p.C + own), d1 = __ldg(p.C + p.nc + nei) - __ldg(p.C + p.nc + own),
                     d2 = __ldg(p.C + 2 * p.nc + nei) - __ldg(p.C + 2 * p.nc + own);
        const int32_t up = (ph > 0) ? own : nei;
        const double gradcf = d0 * g[up] + d1 * g[p.nc + up] + d2 * g[2 * p.nc + up];
        const double gradf = N - P;
        double r;
        if (fabs(gradcf) >= 1000 * fabs(gradf))
            r = 2 * 1000 * sign_(gradcf) * sign_(gradf) - 1;
        else
            r = 2 * (gradcf / gradf) - 1;
        const double lim = (r + fabs(r)) / (1 + fabs(r));
        const double wgt = lim * __ldg(p.w + f) + (1.0 - lim) * pos_(ph);
        ho = ph * (wgt * (P - N) + N);
    }
    p.corr[(size_t)k * p.nint + f] = ho - bd;
}

// per cell: the limiter's bounds (mules_bounds_kernel) and the cell half of sweep 1, where lambda is 1
// (1.0 * phiCorr is phiCorr)
__global__ void __launch_bounds__(256, PBE_MULES_MINB) fused_bounds_kernel(const FusedParams q) {
    FUSED_CELL_THREAD();
    const double *psi = p.psi + (size_t)k * p.stride;
    const double *corr = p.corr + (size_t)k * p.nint;
    const double psiMax = __ldg(p.psi_max + k), psiMin = p.psi_min;
    const double mine = psi[c];
    double pmax = psiMin, pmin = psiMax, sBD = 0.0, sp = kVSMALL, sm = kVSMALL, slp = 0.0, mslm = 0.0;
#pragma unroll 3
    for (int e = p.csr_ptr[c]; e < p.csr_ptr[c + 1]; ++e) {
        const int32_t oth = p.csr_other[e], fc = p.csr_face[e];
        if (oth >= 0) {
            const bool own = fc >= 0;
            const int64_t f = own ? fc : ~fc;
            const double theirs = __ldg(psi + oth);
            pmax = dmax_(pmax, theirs);
            pmin = dmin_(pmin, theirs);
            const double P = own ? mine : theirs, N = own ? theirs : mine;
            const double bd = upwind_flux(__ldg(p.phi + f), P, N);
            sBD = own ? sBD + bd : sBD - bd;
            const double cf = __ldg(corr + f);
            const double l = 1.0 * cf;
            if (own) {
                if (cf > 0.0) sp += cf; else sm -= cf;
                if (l > 0.0) slp += l; else mslm -= l;
            } else {
                if (cf > 0.0) sm += cf; else sp -= cf;
                if (l > 0.0) mslm += l; else slp -= l;
            }
        } else {
            const int64_t b = -(int64_t)oth - 2;
            const double pv = patch_value(p, psi, k, b, c);
            pmax = dmax_(pmax, pv);
            pmin = dmin_(pmin, pv);
            sBD += __ldg(p.phi + p.nint + b) * pv; // the patch face's correction is exactly 0
        }
    }
    pmax = dmin_(pmax, psiMax);
    pmin = dmax_(pmin, psiMin);
    const double V = __ldg(p.V + c), Su = p.S ? __ldg(p.S + (size_t)k * p.stride + c) : 0.0, rdt = p.rdt;
    const size_t o = (size_t)kc * p.nc + c;
    const double bmax = V * ((rdt - 0.0) * pmax - Su - rdt * mine) + sBD;
    const double bmin = V * (Su - (rdt - 0.0) * pmin + rdt * mine) - sBD;
    p.bmax[o] = bmax;
    p.bmin[o] = bmin;
    p.sp[o] = sp;
    p.sm[o] = sm;
    lml_row(q, 0, kc)[c] = make_double2(dmax_(dmin_((slp + bmax) / (sm - kSMALL), 1.0), 0.0),
                                        dmax_(dmin_((mslm + bmin) / (sp + kSMALL), 1.0), 0.0));
}

// per cell: the cell half of sweep q.sweep >= 2, the lambdas of its faces rebuilt from sweeps 1 .. sweep-1
__global__ void __launch_bounds__(256, PBE_MULES_MINB) fused_sweep_kernel(const FusedParams q) {
    FUSED_CELL_THREAD();
    const double *corr = p.corr + (size_t)k * p.nint;
    const int sw = q.sweep - 1; // row this sweep writes
    double slp = 0.0, mslm = 0.0;
#pragma unroll 3
    for (int e = p.csr_ptr[c]; e < p.csr_ptr[c + 1]; ++e) {
        const int32_t oth = p.csr_other[e], fc = p.csr_face[e];
        if (oth < 0) continue; // lambda * 0
        const bool own = fc >= 0;
        const int64_t f = own ? fc : ~fc;
        const double cf = __ldg(corr + f);
        const double l = face_lambda(q, kc, sw, cf, own ? c : (int64_t)oth, own ? (int64_t)oth : c) * cf;
        if (own) {
            if (l > 0.0) slp += l; else mslm -= l;
        } else {
            if (l > 0.0) mslm += l; else slp -= l;
        }
    }
    const size_t o = (size_t)kc * p.nc + c;
    lml_row(q, sw, kc)[c] = make_double2(dmax_(dmin_((slp + p.bmax[o]) / (p.sm[o] - kSMALL), 1.0), 0.0),
                                         dmax_(dmin_((mslm + p.bmin[o]) / (p.sp[o] + kSMALL), 1.0), 0.0));
}

// per internal face and field: the face half of the last sweep, phiCorr *= lambda
__global__ void __launch_bounds__(256, PBE_MULES_MINB) fused_final_kernel(const FusedParams q) {
    FUSED_FACE_THREAD();
    const int32_t own = __ldg(p.owner + f), nei = __ldg(p.neigh + f);
    double *corr = p.corr + (size_t)k * p.nint + f;
    const double cf = *corr;
    *corr = cf * face_lambda(q, kc, q.nsweeps, cf, own, nei);
}

// MULES.C limitSum(), first half: per internal face the sums over all fields (ascending k) and from them the
// two scaling factors -- for the positive and for the negative corrections (1 = untouched: v * 1.0 is v)
__global__ void __launch_bounds__(256) fused_limit_sum_kernel(int nfields, int64_t nint, const double *__restrict__ corr,
                                                              double *__restrict__ fac) {
    const int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= nint) return;
    double sumPos = 0, sumNeg = 0;
#pragma unroll 4
    for (int k = 0; k < nfields; ++k) {
        const double v = __ldg(corr + (size_t)k * nint + f);
        if (v > 0) sumPos += v; else sumNeg += v;
    }
    double fpos = 1.0, fneg = 1.0;
    const double sum = sumPos + sumNeg;
    if (sum > 0 && sumPos > kVSMALL)
        fpos = -sumNeg / sumPos;
    else if (sum < 0 && sumNeg < -kVSMALL)
        fneg = -sumPos / sumNeg;
    fac[f] = fpos;
    fac[nint + f] = fneg;
}

// per cell and field: limitSum's scaling applied, phiPsi = phiCorr + upwind flux, MULES::explicitSolve
__global__ void __launch_bounds__(256, PBE_MULES_MINB) fused_solve_kernel(const FusedParams q) {
    FUSED_CELL_THREAD();
    const double *psi = p.psi + (size_t)k * p.stride;
    const double *corr = p.corr + (size_t)k * p.nint;
    const double mine = psi[c];
    double acc = 0.0;
#pragma unroll 3
    for (int e = p.csr_ptr[c]; e < p.csr_ptr[c + 1]; ++e) {
        const int32_t oth = p.csr_other[e], fc = p.csr_face[e];
        if (oth >= 0) {
            const bool own = fc >= 0;
            const int64_t f = own ? fc : ~fc;
            const double theirs = __ldg(psi + oth);
            const double P = own ? mine : theirs, N = own ? theirs : mine;
            double cv = __ldg(corr + f);
            const double fp = __ldg(q.sums + f), fn = __ldg(q.sums + p.nint + f);
            cv = (cv > 0) ? cv * fp : ((cv < 0) ? cv * fn : cv);
            const double v = cv + upwind_flux(__ldg(p.phi + f), P, N);
            acc = own ? acc + v : acc - v;
        } else {
            const int64_t b = -(int64_t)oth - 2;
            acc += __ldg(p.phi + p.nint + b) * patch_value(p, psi, k, b, c);
        }
    }
    const double div = acc / __ldg(p.V + c);
    const double Su = p.S ? __ldg(p.S + (size_t)k * p.stride + c) : 0.0;
    p.out[(size_t)k * p.stride + c] = (mine * p.rdt + Su - div) / p.rdt;
}

// max over cells of a(c) * b(c) (b may be NULL): psiMax of MOC.C:238 before the division by xi_k
__global__ void __launch_bounds__(256) max_product_kernel(int64_t n, const double *__restrict__ a,
                                                          const double *__restrict__ b, unsigned long long *out) {
    double m = 0.0;
    const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
    for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c < n; c += nthreads) {
        const double v = b ? a[c] * b[c] : a[c];
        m = v > m ? v : m;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    // non-negative doubles order like their bit patterns
    if ((threadIdx.x & 31) == 0) atomicMax(out, (unsigned long long)__double_as_longlong(m));
}
__global__ void moc_psi_max_kernel(int N, const double *__restrict__ xi_k, const double *maxAV, double *out) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < N) out[k] = maxAV[0] / xi_k[k]; // max(alpha V / xi_k) = max(alpha V) / xi_k (monotone rounding)
}
__global__ void qmom_psi_max_kernel(int nm, int64_t ncells, int64_t stride, const double *__restrict__ m,
                                    int64_t nbnd, int64_t stride_b, const double *__restrict__ psi_b, double *out) {
    // 1.2 max(m_k) over the internal field and the patch values (QMOM.C:171); one CTA per moment
    const int k = blockIdx.x;
    double mx = -INFINITY;
    for (int64_t c = threadIdx.x; c < ncells; c += blockDim.x) mx = fmax(mx, m[(size_t)k * stride + c]);
    if (psi_b)
        for (int64_t b = threadIdx.x; b < nbnd; b += blockDim.x) mx = fmax(mx, psi_b[(size_t)k * stride_b + b]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    __shared__ double sh[32];
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = mx;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (int)(blockDim.x >> 5); ++w) mx = fmax(mx, sh[w]);
        out[k] = 1.2 * mx;
    }
}

} // namespace

extern "C" {

int pbe_mesh_set_geometry(pbe_mesh *m, const double *Sf, const double *C, const double *weights) {
    PBE_REQUIRE(m && Sf && C && (weights || m->nint == 0), PBE_E_INVALID, "pbe_mesh_set_geometry: NULL argument");
    DeviceGuard g(m->ctx->device);
    const int64_t nf = m->nint + m->nbnd, nc = m->ncells;
    std::vector<double> sf((size_t)3 * std::max<int64_t>(nf, 1)), cc((size_t)3 * nc);
    for (int64_t f = 0; f < nf; ++f)
        for (int d = 0; d < 3; ++d) sf[(size_t)d * nf + f] = Sf[3 * f + d];
    for (int64_t c = 0; c < nc; ++c)
        for (int d = 0; d < 3; ++d) cc[(size_t)d * nc + c] = C[3 * c + d];
    for (double **p : {&m->d_Sf, &m->d_C, &m->d_w})
        if (*p) {
            cudaFree(*p);
            *p = nullptr;
        }
    PBE_CUDA(cudaMalloc((void **)&m->d_Sf, sf.size() * 8));
    PBE_CUDA(cudaMalloc((void **)&m->d_C, cc.size() * 8));
    PBE_CUDA(cudaMalloc((void **)&m->d_w, (size_t)std::max<int64_t>(m->nint, 1) * 8));
    PBE_CUDA(cudaMemcpy(m->d_Sf, sf.data(), sf.size() * 8, cudaMemcpyHostToDevice));
    PBE_CUDA(cudaMemcpy(m->d_C, cc.data(), cc.size() * 8, cudaMemcpyHostToDevice));
    if (m->nint) PBE_CUDA(cudaMemcpy(m->d_w, weights, (size_t)m->nint * 8, cudaMemcpyHostToDevice));
    return PBE_OK;
}

int pbe_transport_mules(pbe_mesh *m, int scheme, int nfields, int64_t stride, const double *psi, const double *phi,
                        const double *psi_b, int64_t stride_b, const double *S, const double *psi_max, double psi_min,
                        int n_limiter_iter, double dt, double *psi_out, void *cuda_stream) {
    PBE_REQUIRE(m != nullptr, PBE_E_INVALID, "pbe_transport_mules: NULL mesh");
    PBE_REQUIRE(scheme != PBE_DIV_LIMITEDLINEAR, PBE_E_UNSUPPORTED,
                "pbe_transport_mules: Gauss limitedLinear is built for the implicit convection term only "
                "(pbe_transport_implicit); the explicit MULES path has Gauss upwind and Gauss vanLeer");
    PBE_REQUIRE(scheme == PBE_DIV_UPWIND || scheme == PBE_DIV_VANLEER, PBE_E_UNKNOWN_TYPE,
                "pbe_transport_mules: unknown divergence scheme id %d", scheme);
    PBE_REQUIRE(nfields >= 0 && stride >= m->ncells && dt > 0.0 && n_limiter_iter >= 1, PBE_E_INVALID,
                "pbe_transport_mules: bad sizes");
    if (nfields == 0) return PBE_OK;
    PBE_REQUIRE(psi && psi_out && psi_max && (phi || m->nint + m->nbnd == 0), PBE_E_INVALID,
                "pbe_transport_mules: NULL field pointer");
    PBE_REQUIRE(psi != psi_out, PBE_E_INVALID, "pbe_transport_mules: psi_out must not alias psi");
    PBE_REQUIRE(psi_b == nullptr || stride_b >= m->nbnd, PBE_E_INVALID, "pbe_transport_mules: stride_b < boundary faces");
    PBE_REQUIRE(m->ncoupled == 0, PBE_E_UNSUPPORTED,
                "pbe_transport_mules: processor patches are not supported by the MULES path yet (the limiter "
                "needs lambda synchronised across the patch, syncTools::syncFaceList); use Gauss upwind on "
                "decomposed meshes");
    PBE_REQUIRE(scheme != PBE_DIV_VANLEER || (m->d_Sf && m->d_C && m->d_w), PBE_E_INVALID,
                "pbe_transport_mules: Gauss vanLeer needs the mesh geometry (pbe_mesh_set_geometry)");
    DeviceGuard g(m->ctx->device);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    const int64_t nc = m->ncells, ni = m->nint;
    const bool fused = !(getenv("PBE_MULES_PATH") && atoi(getenv("PBE_MULES_PATH")) == 0);
    if (fused) {
        // fields per chunk: the chunk's cell rows (gradient 3, bounds 4, lm / lp of every sweep) within ~4 GB
        const size_t rows = 7 + 2 * (size_t)n_limiter_iter;
        const size_t nie = (size_t)std::max<int64_t>(ni, 1);
        int KC = (int)std::max<size_t>(1, std::min<size_t>((size_t)nfields, ((size_t)1 << 32) / (rows * (size_t)nc * 8) + 1));
        if (const char *e = getenv("PBE_MULES_CHUNK")) KC = std::max(1, std::min(nfields, atoi(e)));
        const size_t need = (size_t)nfields * nie + 2 * nie + (size_t)KC * rows * (size_t)nc + 2;
        if (m->mules_ws_doubles < need) {
            cudaFree(m->mules_ws);
            m->mules_ws = nullptr;
            m->mules_ws_doubles = 0;
            PBE_CUDA(cudaMalloc((void **)&m->mules_ws, need * sizeof(double)));
            m->mules_ws_doubles = need;
        }
        FusedParams q{};
        MulesParams &p = q.m;
        p.nc = nc, p.nint = ni, p.nbnd = m->nbnd, p.stride = stride, p.stride_b = stride_b;
        p.csr_ptr = m->d_csr_ptr, p.csr_other = m->d_csr_other, p.csr_face = m->d_csr_face;
        p.owner = m->d_owner, p.neigh = m->d_neigh;
        p.V = m->d_vcell, p.Sf = m->d_Sf, p.C = m->d_C, p.w = m->d_w;
        p.psi = psi, p.psi_b = psi_b, p.phi = phi, p.S = S, p.psi_max = psi_max;
        p.psi_min = psi_min, p.rdt = 1.0 / dt;
        p.scheme = scheme;
        double *w = m->mules_ws;
        p.corr = w, w += (size_t)nfields * nie;
        q.sums = w, w += 2 * nie;
        p.grad = w, w += (size_t)KC * 3 * nc;
        p.bmax = w, w += (size_t)KC * nc;
        p.bmin = w, w += (size_t)KC * nc;
        p.sp = w, w += (size_t)KC * nc;
        p.sm = w, w += (size_t)KC * nc;
        if ((w - m->mules_ws) & 1) ++w; // 16-byte alignment of the {lm, lp} rows
        q.lml = reinterpret_cast<double2 *>(w);
        q.KC = KC;
        q.nsweeps = n_limiter_iter;
        p.out = psi_out;
        int launches = 0;
        const dim3 blk(kFusedBx, kFusedBy);
        const unsigned gcx = (unsigned)((nc + kFusedBx - 1) / kFusedBx), gfx = (unsigned)((nie + kFusedBx - 1) / kFusedBx);
        for (int k0 = 0; k0 < nfields; k0 += KC) {
            q.kn = std::min(KC, nfields - k0);
            p.k0 = k0;
            const unsigned gy = (unsigned)((q.kn + kFusedBy - 1) / kFusedBy);
            if (ni) {
                if (scheme == PBE_DIV_VANLEER) {
                    fused_grad_kernel<<<dim3(gcx, gy), blk, 0, st>>>(q);
                    ++launches;
                }
                fused_corr_kernel<<<dim3(gfx, gy), blk, 0, st>>>(q);
                ++launches;
            }
            fused_bounds_kernel<<<dim3(gcx, gy), blk, 0, st>>>(q);
            ++launches;
            for (int it = 2; it <= n_limiter_iter; ++it) {
                q.sweep = it;
                fused_sweep_kernel<<<dim3(gcx, gy), blk, 0, st>>>(q);
                ++launches;
            }
            if (ni) {
                fused_final_kernel<<<dim3(gfx, gy), blk, 0, st>>>(q);
                ++launches;
            }
        }
        if (ni) {
            fused_limit_sum_kernel<<<(unsigned)((nie + 255) / 256), 256, 0, st>>>(nfields, ni, p.corr, q.sums);
            ++launches;
        }
        p.k0 = 0;
        q.kn = nfields;
        fused_solve_kernel<<<dim3(gcx, (unsigned)((nfields + kFusedBy - 1) / kFusedBy)), blk, 0, st>>>(q);
        ++launches;
        PBE_CUDA(cudaGetLastError());
        m->ctx->launches += launches;
        return PBE_OK;
    }
    // fields per chunk: the chunk's temporaries (lambda + 9 cell rows) within ~1.5 GB
    const size_t per_field = (size_t)std::max<int64_t>(ni, 1) + 9 * (size_t)nc;
    int KC = (int)std::max<size_t>(1, std::min<size_t>((size_t)nfields, ((size_t)3 << 28) / (per_field * 8) + 1));
    if (const char *e = getenv("PBE_MULES_CHUNK")) KC = std::max(1, std::min(nfields, atoi(e)));
    const size_t need = (size_t)nfields * std::max<int64_t>(ni, 1) + (size_t)KC * per_field;
    if (m->mules_ws_doubles < need) {
        cudaFree(m->mules_ws);
        m->mules_ws = nullptr;
        m->mules_ws_doubles = 0;
        PBE_CUDA(cudaMalloc((void **)&m->mules_ws, need * sizeof(double)));
        m->mules_ws_doubles = need;
    }
    MulesParams p{};
    p.nc = nc, p.nint = ni, p.nbnd = m->nbnd, p.stride = stride, p.stride_b = stride_b;
    p.csr_ptr = m->d_csr_ptr, p.csr_other = m->d_csr_other, p.csr_face = m->d_csr_face;
    p.owner = m->d_owner, p.neigh = m->d_neigh;
    p.V = m->d_vcell, p.Sf = m->d_Sf, p.C = m->d_C, p.w = m->d_w;
    p.psi = psi, p.psi_b = psi_b, p.phi = phi, p.S = S, p.psi_max = psi_max;
    p.psi_min = psi_min, p.rdt = 1.0 / dt;
    p.scheme = scheme;
    double *w = m->mules_ws;
    p.corr = w, w += (size_t)nfields * std::max<int64_t>(ni, 1);
    p.lambda = w, w += (size_t)KC * std::max<int64_t>(ni, 1);
    p.grad = w, w += (size_t)KC * 3 * nc;
    p.bmax = w, w += (size_t)KC * nc;
    p.bmin = w, w += (size_t)KC * nc;
    p.sp = w, w += (size_t)KC * nc;
    p.sm = w, w += (size_t)KC * nc;
    p.lm = w, w += (size_t)KC * nc;
    p.lp = w;
    p.out = psi_out;
    const unsigned gc = (unsigned)((nc + 255) / 256), gf = (unsigned)((std::max<int64_t>(ni, 1) + 255) / 256);
    int launches = 0;
    for (int k0 = 0; k0 < nfields; k0 += KC) {
        const int kn = std::min(KC, nfields - k0);
        p.k0 = k0;
        if (scheme == PBE_DIV_VANLEER) {
            mules_grad_kernel<<<dim3(gc, kn), 256, 0, st>>>(p);
            ++launches;
        }
        if (ni) mules_corr_kernel<<<dim3(gf, kn), 256, 0, st>>>(p);
        mules_bounds_kernel<<<dim3(gc, kn), 256, 0, st>>>(p);
        launches += 2;
        for (int it = 0; it < n_limiter_iter; ++it) {
            p.last = (it == n_limiter_iter - 1);
            mules_sweep_cells_kernel<<<dim3(gc, kn), 256, 0, st>>>(p);
            if (ni) mules_sweep_faces_kernel<<<dim3(gf, kn), 256, 0, st>>>(p);
            launches += 2;
        }
    }
    if (ni) {
        mules_limit_sum_kernel<<<gf, 256, 0, st>>>(nfields, ni, p.corr);
        ++launches;
    }
    p.k0 = 0;
    mules_solve_kernel<<<dim3(gc, nfields), 256, 0, st>>>(p);
    ++launches;
    PBE_CUDA(cudaGetLastError());
    m->ctx->launches += launches;
    return PBE_OK;
}

int pbe_moc_psi_max(pbe_moc *moc, int64_t ncells, const double *alpha, const double *vcell, double *psi_max,
                    void *cuda_stream) {
    PBE_REQUIRE(moc && psi_max && (ncells == 0 || (alpha && vcell)), PBE_E_INVALID, "pbe_moc_psi_max: NULL argument");
    int N = 0;
    pbe_ctx *ctx = nullptr;
    double xi1 = 0.0;
    int rc = pbe_moc_num_classes(moc, &N, &ctx);
    if (rc) return rc;
    pbe_moc_coeffs(moc, &xi1, nullptr);
    const double *xi_dev = nullptr;
    pbe_moc_grid_device(moc, &xi_dev);
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    double *tmp = nullptr;
    PBE_CUDA(cudaMallocAsync((void **)&tmp, sizeof(double), st));
    PBE_CUDA(cudaMemsetAsync(tmp, 0, sizeof(double), st));
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((ncells + 255) / 256, (int64_t)ctx->sm_count * 8));
    max_product_kernel<<<grid, 256, 0, st>>>(ncells, alpha, vcell, (unsigned long long *)tmp);
    moc_psi_max_kernel<<<(N + 127) / 128, 128, 0, st>>>(N, xi_dev, tmp, psi_max);
    PBE_CUDA(cudaGetLastError());
    PBE_CUDA(cudaFreeAsync(tmp, st));
    ctx->launches += 2;
    return PBE_OK;
}

int pbe_qmom_psi_max(pbe_ctx *ctx, int nmoments, int64_t ncells, int64_t stride, const double *m, int64_t nbnd,
                     int64_t stride_b, const double *psi_b, double *psi_max, void *cuda_stream) {
    PBE_REQUIRE(ctx && psi_max && (ncells == 0 || m) && nmoments >= 1, PBE_E_INVALID, "pbe_qmom_psi_max: bad argument");
    DeviceGuard g(ctx->device);
    qmom_psi_max_kernel<<<nmoments, 1024, 0, (cudaStream_t)cuda_stream>>>(nmoments, ncells, stride, m, nbnd, stride_b, psi_b,
                                                                         psi_max);
    PBE_CUDA(cudaGetLastError());
    ctx->launches++;
    return PBE_OK;
}

} // extern "C"
