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

#include "pbe_kernels.cuh"

namespace pbe {

// ---- symmetric tridiagonal eigen-solve, first eigenvector row only --------------------------
// Implicit-shift QL with Wilkinson-type shift.  d: diagonal (in) / eigenvalues ascending (out);
// e[0..N-2]: off-diagonal (destroyed, e[N-1] scratch); z: first components of the normalised
// eigenvectors.  Non-finite input or no convergence in 60 sweeps -> all NaN (returns false).
template <int N>
__device__ __forceinline__ bool tridiag_ql_firstrow(double (&d)[N], double (&e)[N], double (&z)[N]) {
    bool bad = false;
#pragma unroll
    for (int i = 0; i < N; ++i) {
        z[i] = (i == 0) ? 1.0 : 0.0;
        bad |= !isfinite(d[i]);
        if (i < N - 1) bad |= !isfinite(e[i]);
    }
    e[N - 1] = 0.0;
    if (!bad) {
#pragma unroll
        for (int l = 0; l < N - 1; ++l) {
            int iter = 0;
            while (true) {
                // smallest m >= l whose off-diagonal is negligible (m = N-1 if none)
                int m = N - 1;
#pragma unroll
                for (int mm = N - 2; mm >= l; --mm) {
                    const double dd = fabs(d[mm]) + fabs(d[mm + 1]);
                    if (fabs(e[mm]) <= DBL_EPSILON * dd) m = mm;
                }
                if (m == l) break;
                if (++iter > 60) {
                    bad = true;
                    break;
                }
                double dm = d[N - 1];
#pragma unroll
                for (int mm = 0; mm < N - 1; ++mm)
                    if (mm == m) dm = d[mm];
                double g = (d[l + 1] - d[l]) / (2.0 * e[l]);
                double r = sqrt(g * g + 1.0);
                g = dm - d[l] + e[l] / (g + copysign(r, g));
                double s = 1.0, c = 1.0, p = 0.0;
                bool early = false;
#pragma unroll
                for (int i = N - 2; i >= l; --i) {
                    if (i < m && !early) {
                        double f = s * e[i];
                        const double b = c * e[i];
                        r = sqrt(f * f + g * g);
                        e[i + 1] = r;
                        if (r == 0.0) {
                            d[i + 1] -= p;
                            early = true;
                        } else {
                            s = f / r;
                            c = g / r;
                            g = d[i + 1] - p;
                            r = (d[i] - g) * s + 2.0 * c * b;
                            p = s * r;
                            d[i + 1] = g + p;
                            g = c * r - b;
                            f = z[i + 1];
                            z[i + 1] = s * z[i] + c * f;
                            z[i] = c * z[i] - s * f;
                        }
                    }
                }
                if (!early) {
                    d[l] -= p;
                    e[l] = g;
                }
#pragma unroll
                for (int mm = 0; mm < N; ++mm)
                    if (mm == m) e[mm] = 0.0;
            }
            if (bad) break;
        }
    }
    if (bad) {
#pragma unroll
        for (int i = 0; i < N; ++i) {
            d[i] = NAN;
            z[i] = NAN;
        }
        return false;
    }
    // ascending eigenvalues (Eigen's documented order): odd-even transposition network
#pragma unroll
    for (int pass = 0; pass < N; ++pass) {
#pragma unroll
        for (int i = (pass & 1); i + 1 < N; i += 2) {
            if (d[i] > d[i + 1]) {
                const double td = d[i], tz = z[i];
                d[i] = d[i + 1];
                z[i] = z[i + 1];
                d[i + 1] = td;
                z[i + 1] = tz;
            }
        }
    }
    return true;
}

// ---- Wheeler (1974) moment inversion, MomentInversion.C:9-51 ---------------------------------
// The sigma recurrence is written with explicit round-to-nearest multiplies and subtracts (no
// FMA contraction) so that a, b -- and with them the NaN / Inf pattern of degenerate moment
// sets (0/0 in MomentInversion.C:30-33) -- are bit-identical to un-fused host arithmetic.
template <int N>
__device__ __forceinline__ void wheeler_dev(const double (&m)[2 * N], double (&x)[N], double (&w)[N]) {
    if (N == 1) {
        x[0] = m[1] / m[0];
        w[0] = m[0];
        return;
    }
    double a[N], b[N];
    double sp[2 * N], sc[2 * N], sn[2 * N]; // sigma rows k-1, k, k+1
#pragma unroll
    for (int j = 0; j < 2 * N; ++j) {
        sp[j] = 0.0;
        sc[j] = m[j];
        sn[j] = 0.0;
    }
    a[0] = m[1] / m[0];
    b[0] = m[0];
#pragma unroll
    for (int k = 1; k < N; ++k) {
#pragma unroll
        for (int col = k; col < 2 * N - k; ++col) {
            const double t = __dsub_rn(sc[col + 1], __dmul_rn(a[k - 1], sc[col]));
            sn[col] = __dsub_rn(t, __dmul_rn(b[k - 1], sp[col]));
        }
        a[k] = __dadd_rn(-(sc[k] / sc[k - 1]), sn[k + 1] / sn[k]);
        b[k] = sn[k] / sc[k - 1];
#pragma unroll
        for (int j = 0; j < 2 * N; ++j) {
            sp[j] = sc[j];
            sc[j] = sn[j];
        }
    }
    double e[N], z[N];
#pragma unroll
    for (int i = 0; i < N; ++i) {
        x[i] = a[i];
        e[i] = (i < N - 1) ? -sqrt(fabs(b[i + 1])) : 0.0; // MomentInversion.C:35-41
    }
    tridiag_ql_firstrow<N>(x, e, z);
#pragma unroll
    for (int i = 0; i < N; ++i) w[i] = m[0] * (z[i] * z[i]); // MomentInversion.C:48
}

// ---- all 2N moment sources of one cell, QMOM.C:226-264 ---------------------------------------
template <int N>
__device__ __forceinline__ void qmom_cell_sources(const DevKernelSpec &ks, const double (&x)[N],
                                                  const double (&w)[N], double eps, double rho,
                                                  double nu, double (&out)[2 * N]) {
    const CellCoef cc = cell_coef(ks, eps, rho, nu);
    NodeRoots rt[N];
    double g[N];
#pragma unroll
    for (int i = 0; i < N; ++i) {
        rt[i].a = rt[i].p = 0.0;
        if (ks.coal == PBE_COAL_CT || ks.br == PBE_BR_CT) rt[i] = node_roots(x[i]);
        g[i] = break_S(ks, cc, x[i], rt[i]);
    }
    // symmetric kernel matrix, each unordered node pair evaluated once
    double Sm[N][N];
#pragma unroll
    for (int i = 0; i < N; ++i)
#pragma unroll
        for (int j = i; j < N; ++j) {
            Sm[i][j] = coal_S(ks, cc, rt[i], rt[j]);
            Sm[j][i] = Sm[i][j];
        }
    double death[N], xs[N][N], pw_sum[N][N], pw_x[N];
#pragma unroll
    for (int i = 0; i < N; ++i) {
        double dsum = 0.0;
#pragma unroll
        for (int j = 0; j < N; ++j) {
            dsum += w[j] * Sm[i][j]; // QMOM.C:240-243
            xs[i][j] = x[i] + x[j];
            pw_sum[i][j] = 1.0; // pow(x, 0) = 1 for every x
        }
        death[i] = dsum;
        pw_x[i] = 1.0;
    }
#pragma unroll
    for (int k = 0; k < 2 * N; ++k) {
        double coal = 0.0, br = 0.0;
#pragma unroll
        for (int i = 0; i < N; ++i) {
            double birth = 0.0;
#pragma unroll
            for (int j = 0; j < N; ++j) birth += (w[j] * pw_sum[i][j]) * Sm[i][j]; // QMOM.C:245-249
            coal += w[i] * (birth / 2. - pw_x[i] * death[i]);                      // QMOM.C:252-255
            br += (w[i] * g[i]) * (daughter_moment_from_pow(pw_x[i], k) - pw_x[i]); // QMOM.C:257-260
        }
        out[k] = coal + br; // QMOM.C:263
#pragma unroll
        for (int i = 0; i < N; ++i) {
            pw_x[i] *= x[i];
#pragma unroll
            for (int j = 0; j < N; ++j) pw_sum[i][j] *= xs[i][j];
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

template <int N>
__global__ void __launch_bounds__(128) qmom_kernel(const __grid_constant__ QmomParams p) {
    const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
    for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c < p.ncells; c += nthreads) {
        double m[2 * N], x[N], w[N], S[2 * N];
#pragma unroll
        for (int k = 0; k < 2 * N; ++k) m[k] = __ldg(p.m_in + (size_t)k * p.stride + c);
        const double eps = __ldg(p.eps + c), rho = __ldg(p.rho + c), nu = __ldg(p.nu + c);
        wheeler_dev<N>(m, x, w);
        qmom_cell_sources<N>(p.ks, x, w, eps, rho, nu, S);
        if (p.do_step) {
#pragma unroll
            for (int k = 0; k < 2 * N; ++k) {
                if (S[k] != S[k]) S[k] = 0.0;                       // QMOM.C:101-107
                const double v = __dadd_rn(m[k], __dmul_rn(p.dt, S[k])); // QMOM.C:193-197, phi = 0
                m[k] = (v > 1.0e-15) ? v : 1.0e-15;                 // QMOM.C:117-123, SMALL
                p.m_out[(size_t)k * p.stride + c] = m[k];
            }
            if (p.d_out != nullptr && N >= 1)
                p.d_out[c] = cbrt(6.0 / 3.14159265358979323846 * m[1] / m[0]); // QMOM.C:125
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
