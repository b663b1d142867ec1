// pbe_qmom_kernel.cuh -- QMOM moment inversion + quadrature source kernels for sm_100a.
//
// Replaces wheeler_inversion() (src/PBESystem/MomentInversion.C:9-51), QMOM::momentSourceTerm
// (src/PBESystem/QMOM.C:215-267) for all 2*order moments at once, and the element-wise part of
// QMOM::correct() (QMOM.C:98-128: NaN scrub, zero-flux update, clamp, d).
//
// One thread per cell, everything in registers: the (order+1) x 2*order sigma table of the
// Wheeler recurrence, the order x order symmetric tridiagonal Jacobi matrix, its eigenvalues
// and the first row of its eigenvectors.  No Eigen, no heap, no local memory: all loops are
// fully unrolled over the compile-time order with predication where the QL iteration needs a
// run-time bound.  The inversion is done ONCE per cell (the reference redoes it for every
// moment index, QMOM.C:231) and the kernel values S(xi_i, xi_j), g(xi_i) are evaluated once
// per node pair / node and reused for every moment.
#pragma once

#include <cuda_runtime.h>
#include <float.h>
#include <stdint.h>

#include "pbe_eig.cuh"
#include "pbe_kernels.cuh"

namespace pbe {

// ---- all 2N moment sources of one cell, QMOM.C:226-264 ---------------------------------------
// out_k = sum_i w_i [ 1/2 sum_j w_j (x_i+x_j)^k S_ij - x_i^k sum_j w_j S_ij ]          (QMOM.C:237-255)
//       + sum_i w_i g_i (2 x_i^k/(k+1) - x_i^k)                                        (QMOM.C:257-260)
// S is symmetric for every kernel of the plug-in surface (coalescenceKernels/*.C), so the
// birth double sum is folded over unordered node pairs: W_ij = w_i w_j S_ij (i < j),
// W_ii = w_i^2 S_ii / 2, and the kernel values are evaluated once per pair / node and reused for
// every moment index (the reference re-evaluates them, and re-inverts, for every k).
// Node roots by ninth_roots() (pbe_math.cuh): r = x^(1/9), u = x^(-1/9), no division; the
// harmonic-mean radius of a pair is 1 / (x_i^(-1/3) + x_j^(-1/3)) by rcp_fast().
// COAL / BR are the compile-time kernel ids (PBE_COAL_*, PBE_BR_*): no run-time dispatch and no
// dead code inside the cell loop.
template <int N, int COAL, int BR>
__device__ __forceinline__ void qmom_cell_sources(const DevKernelSpec &ks, const double (&x)[N],
                                                  const double (&w)[N], double eps, double rho,
                                                  double nu, const double *__restrict__ T128,
                                                  double (&out)[2 * N]) {
    constexpr bool ct_c = (COAL == PBE_COAL_CT), ct_b = (BR == PBE_BR_CT);
    CellCoef cc;
    cc.A = cc.negK = cc.Ab = cc.negKb = 0.0;
#if PBE_QMOM_FASTMATH
    if (ct_c || ct_b) cc = cell_coef_fast(ks, eps, rho, nu);
#else
    if (ct_c || ct_b) cc = cell_coef(ks, eps, rho, nu);
#endif
    // exponentials as 2^(K log2(e) .) through the shared-memory table (exp2_neg_tab_safe)
    const double K2 = cc.negK * 1.4426950408889634074, Kb2 = cc.negKb * 1.4426950408889634074;
    double ra[N], rp[N], ia[N], r1[N], gw[N]; // x^(1/3), x^(2/9), x^(-1/3), x^(1/9), w_i g_i
#pragma unroll
    for (int i = 0; i < N; ++i) {
        ra[i] = rp[i] = ia[i] = r1[i] = 0.0;
        double g = 0.0;
        if (ct_c || ct_b) {
            double r9, u9;
            ninth_roots(x[i], r9, u9);
            const double r2 = r9 * r9, u2 = u9 * u9;
            ra[i] = r2 * r9;
            rp[i] = r2;
            r1[i] = r9;
            ia[i] = u2 * u9;
            if (ct_b) // breakupKernels/CoulaloglouTavlarides.C:83-92
                g = (cc.Ab * u2) * exp2_neg_tab_safe(Kb2 * (ia[i] * u2), T128);
        }
        if (BR == PBE_BR_BINARY) g = x[i] * x[i]; // binaryBreakup.C:51-53
        gw[i] = w[i] * g;
    }
    // unordered pairs: W (birth weight) and the death sums
    double W[N][N], dw[N];
#pragma unroll
    for (int i = 0; i < N; ++i) dw[i] = 0.0;
#pragma unroll
    for (int i = 0; i < N; ++i)
#pragma unroll
        for (int j = i; j < N; ++j) {
            double S = 0.0;
            if (ct_c && j == i) {
                // equal volumes: a + b = 2a, sqrt(2 x^(2/9)) = sqrt(2) x^(1/9), a b/(a + b) = a/2;
                // t = x^(1/3) x^(-1/3) is 1, or NaN exactly where 0*0/(0+0) is (xi = 0, inf)
                const double t = ra[i] * ia[i];
                const double G = (4.0 * 1.41421356237309504880) * (ra[i] * ra[i]) * r1[i];
                const double h = (0.5 * ra[i]) * t;
                const double h2 = h * h;
                S = (cc.A * G) * exp2_neg_tab_safe(K2 * (h2 * h2), T128);
            } else if (ct_c) { // coalescenceKernels/CoulaloglouTavlarides.C:69-90
                const double sa = ra[i] + ra[j];
#if PBE_QMOM_FASTMATH
                const double G = (sa * sa) * sqrt_pos_fast(rp[i] + rp[j]);
#else
                const double G = (sa * sa) * sqrt(rp[i] + rp[j]);
#endif
                const double h = rcp_fast(ia[i] + ia[j]); // a b / (a + b)
                const double h2 = h * h;
                S = (cc.A * G) * exp2_neg_tab_safe(K2 * (h2 * h2), T128);
            } else if (COAL == PBE_COAL_CONSTANT) {
                S = ks.coal_const;
            }
            dw[i] = fma_(w[j], S, dw[i]); // QMOM.C:240-243
            if (j != i) dw[j] = fma_(w[i], S, dw[j]);
            W[i][j] = (j == i) ? 0.5 * (w[i] * w[i]) * S : (w[i] * w[j]) * S;
        }
    double xs[N][N], pw_s[N][N], pw_x[N];
#pragma unroll
    for (int i = 0; i < N; ++i) {
        dw[i] *= w[i];
        pw_x[i] = 1.0; // pow(x, 0) = 1 for every x
#pragma unroll
        for (int j = i; j < N; ++j) {
            xs[i][j] = x[i] + x[j];
            pw_s[i][j] = 1.0;
        }
    }
#pragma unroll
    for (int k = 0; k < 2 * N; ++k) {
        double birth = 0.0, death = 0.0, br = 0.0;
#pragma unroll
        for (int i = 0; i < N; ++i) {
            if (COAL != PBE_COAL_NONE) {
#pragma unroll
                for (int j = i; j < N; ++j) birth = fma_(W[i][j], pw_s[i][j], birth); // QMOM.C:245-249
                death = fma_(dw[i], pw_x[i], death);                                   // QMOM.C:252-255
            }
            if (BR != PBE_BR_NONE) br = fma_(gw[i], pw_x[i], br); // QMOM.C:257-260
        }
        // uniformBinaryBreakup::moment (uniformBinaryBreakup.C:62-69): 2 x^k/(k+1) - x^k
        out[k] = (birth - death) + br * (2.0 / (double)(k + 1) - 1.0); // QMOM.C:263
        if (k + 1 < 2 * N) {
#pragma unroll
            for (int i = 0; i < N; ++i) {
                pw_x[i] *= x[i];
                if (COAL != PBE_COAL_NONE) {
#pragma unroll
                    for (int j = i; j < N; ++j) pw_s[i][j] *= xs[i][j];
                }
            }
        }
    }
}

struct QmomParams {
    int64_t ncells, stride;
    const double *m_in;
    double *m_out; // do_step: may alias m_in
    double *S_out; // may be NULL
    double *d_out; // may be NULL
    const double *eps, *rho, *nu;
    double dt;
    int do_step;
    DevKernelSpec ks;
};

#ifdef PBE_QMOM_MIN_CTAS // A/B builds: resident CTAs per SM the register allocation must allow
#define PBE_QMOM_BOUNDS __launch_bounds__(128, PBE_QMOM_MIN_CTAS)
#else
#define PBE_QMOM_BOUNDS __launch_bounds__(128)
#endif
template <int N, int COAL, int BR>
__global__ void PBE_QMOM_BOUNDS qmom_kernel(const __grid_constant__ QmomParams p) {
    __shared__ double T128[128]; // 2^(j/128), exp2_neg_tab<7>
    T128[threadIdx.x] = PBE_EXP2_TAB128[threadIdx.x];
    __syncthreads();
    const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
    for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c < p.ncells; c += nthreads) {
        double m[2 * N], x[N], w[N], S[2 * N];
#pragma unroll
        for (int k = 0; k < 2 * N; ++k) m[k] = __ldg(p.m_in + (size_t)k * p.stride + c);
        const double eps = __ldg(p.eps + c), rho = __ldg(p.rho + c), nu = __ldg(p.nu + c);
        // pull the next grid-stride cell's lines into L2 meanwhile (one request per 32-B sector):
        // the loads of the next trip then see L2, not HBM, latency
        if ((threadIdx.x & 3) == 0 && c + nthreads < p.ncells) {
            const int64_t cn = c + nthreads;
#pragma unroll
            for (int k = 0; k < 2 * N; ++k)
                asm volatile("prefetch.global.L2 [%0];" ::"l"(p.m_in + (size_t)k * p.stride + cn));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(p.eps + cn));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(p.rho + cn));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(p.nu + cn));
        }
        wheeler_dev<N>(m, x, w);
        qmom_cell_sources<N, COAL, BR>(p.ks, x, w, eps, rho, nu, T128, S);
        if (p.do_step) {
#pragma unroll
            for (int k = 0; k < 2 * N; ++k) {
                if (S[k] != S[k]) S[k] = 0.0;                       // QMOM.C:101-107
                const double v = __dadd_rn(m[k], __dmul_rn(p.dt, S[k])); // QMOM.C:193-197, phi = 0
                m[k] = (v > 1.0e-15) ? v : 1.0e-15;                 // QMOM.C:117-123, SMALL
                p.m_out[(size_t)k * p.stride + c] = m[k];
            }
            if (p.d_out != nullptr && N >= 1)
#if PBE_QMOM_FASTMATH
                p.d_out[c] = cbrt_pos_fast((6.0 / 3.14159265358979323846 * m[1]) * rcp_fast(m[0])); // QMOM.C:125
#else
                p.d_out[c] = cbrt(6.0 / 3.14159265358979323846 * m[1] / m[0]); // QMOM.C:125
#endif
        }
        if (p.S_out != nullptr) {
#pragma unroll
            for (int k = 0; k < 2 * N; ++k) p.S_out[(size_t)k * p.stride + c] = S[k];
        }
    }
}

struct WheelerParams {
    int64_t ncells, stride;
    const double *m;
    double *x, *w;
};

template <int N>
__global__ void __launch_bounds__(128) wheeler_kernel(const __grid_constant__ WheelerParams p) {
    const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
    for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c < p.ncells; c += nthreads) {
        double m[2 * N], x[N], w[N];
#pragma unroll
        for (int k = 0; k < 2 * N; ++k) m[k] = __ldg(p.m + (size_t)k * p.stride + c);
        wheeler_dev<N>(m, x, w);
#pragma unroll
        for (int i = 0; i < N; ++i) {
            p.x[(size_t)i * p.stride + c] = x[i];
            p.w[(size_t)i * p.stride + c] = w[i];
        }
    }
}

} // namespace pbe
