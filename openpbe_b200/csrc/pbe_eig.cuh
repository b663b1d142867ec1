// pbe_eig.cuh -- symmetric tridiagonal eigen-solves for the per-cell Jacobi matrix of the QMOM
// inversion (MomentInversion.C:35-48: Eigen::SelfAdjointEigenSolver on the order x order matrix;
// only the eigenvalues and the FIRST ROW of the eigenvector matrix are used, :47-48).
//
//   tridiag_ql_firstrow<N>   generic implicit-shift QL (any order; data-dependent iteration count)
//   jacobi2_firstrow         order 2: one closed-form Jacobi rotation
//   jacobi3_firstrow         order 3: fixed work, no data-dependent loop --
//        1. the eigenvalue that is guaranteed well separated (the largest if det(J - q I) >= 0,
//           else the smallest) from the trigonometric form of the characteristic cubic
//           4 t^3 - 3 t = r, t in [sqrt(3)/2, 1]: FP32 seed + two FP64 Newton steps (h' >= 6);
//        2. ONE implicit QL sweep with that eigenvalue as the shift ("perfect shift"): the
//           orthogonal similarity deflates it into d[0] and leaves a 2 x 2 block;
//        3. the 2 x 2 block by a closed-form rotation, which resolves a clustered node pair to
//           full relative accuracy (a root-of-the-cubic formula cannot: its roots are
//           conditioned by 1/gap);
//        4. a residual test on the deflated off-diagonal; cells that fail it (or produce a
//           non-finite value from finite input) re-run through the generic QL -- rare, so the
//           divergence costs nothing measurable.
//
// Host-compilable (PBE_HD) so that tests/test_eig_host.py can run the arithmetic against the
// oracle on the CPU; the product only runs the device instantiation.
#pragma once

#include <float.h>

#include "pbe_math.cuh"

namespace pbe {

PBE_HD bool finite_(double x) {
#if defined(__CUDA_ARCH__)
    return isfinite(x);
#else
    return std::isfinite(x);
#endif
}
// reciprocal without the IEEE-division slow path on the device (<= 1 ulp; pbe_math.cuh)
PBE_HD double rcp_(double x) {
#if defined(__CUDA_ARCH__)
    return rcp_fast(x);
#else
    return 1.0 / x;
#endif
}
PBE_HD double rsqrt_(double x) {
#if defined(__CUDA_ARCH__)
    return rsqrt(x);
#else
    return 1.0 / sqrt(x);
#endif
}

// ---- generic: implicit-shift QL with Wilkinson-type shift, first eigenvector row only --------
// d: diagonal (in) / eigenvalues ascending (out); e[0..N-2]: off-diagonal (destroyed, e[N-1]
// scratch); z: first components of the normalised eigenvectors.  Non-finite input or no
// convergence in 60 sweeps -> all NaN (returns false).
template <int N>
PBE_HD bool tridiag_ql_firstrow(double (&d)[N], double (&e)[N], double (&z)[N]) {
    bool bad = false;
#pragma unroll
    for (int i = 0; i < N; ++i) {
        z[i] = (i == 0) ? 1.0 : 0.0;
        bad |= !finite_(d[i]);
        if (i < N - 1) bad |= !finite_(e[i]);
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

// ---- closed-form Jacobi rotation of the symmetric 2 x 2 block [[p, e], [e, q]] ----------------
// p, q become the eigenvalues; (zp, zq) -- the first-row components carried in this basis --
// are rotated with it.  e == 0 leaves everything untouched.
PBE_HD void rot2_firstrow(double &p, double &q, double e, double &zp, double &zq) {
    const double theta = (q - p) * rcp_(2.0 * e);
    const double at = fabs(theta);
    double t = rcp_(at + sqrt(fma_(theta, theta, 1.0))); // smaller root of t^2 + 2 theta t - 1
    t = (theta < 0.0) ? -t : t;
    t = (e == 0.0) ? 0.0 : t; // also covers theta = 0/0
    const double c = rsqrt_(fma_(t, t, 1.0));
    const double s = t * c;
    const double te = t * e;
    p -= te;
    q += te;
    const double a = zp, b = zq;
    zp = c * a - s * b;
    zq = s * a + c * b;
}

PBE_HD void cswap_firstrow(double &da, double &db, double &za, double &zb) {
    if (da > db) {
        const double t = da, u = za;
        da = db;
        za = zb;
        db = t;
        zb = u;
    }
}

// order 2: d[2] diagonal, e[0] off-diagonal -> ascending eigenvalues, first-row components
PBE_HD bool jacobi2_firstrow(double (&d)[2], double (&e)[2], double (&z)[2]) {
    if (!(finite_(d[0]) && finite_(d[1]) && finite_(e[0]))) {
        d[0] = d[1] = z[0] = z[1] = NAN;
        return false;
    }
    z[0] = 1.0;
    z[1] = 0.0;
    rot2_firstrow(d[0], d[1], e[0], z[0], z[1]);
    cswap_firstrow(d[0], d[1], z[0], z[1]);
    return true;
}

#if defined(PBE_EIG_COUNT_FALLBACKS)
static long jacobi3_fallbacks = 0; // host test instrumentation only
#endif
// out-of-line generic solve for the rare cells the fixed-work path hands back: keeps ~600
// instructions of unrolled QL out of the straight-line cell code (instruction-cache pressure)
#if defined(__CUDACC__)
__host__ __device__ __noinline__
#else
inline
#endif
bool jacobi3_generic(double *d3, double *e3, double *z3) {
    double d[3] = {d3[0], d3[1], d3[2]}, e[3] = {e3[0], e3[1], 0.0}, z[3];
    const bool ok = tridiag_ql_firstrow<3>(d, e, z);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        d3[i] = d[i];
        z3[i] = z[i];
    }
    return ok;
}
// order 3, fixed work (see the header).  d: diagonal in / ascending eigenvalues out;
// e[0], e[1]: off-diagonals (e[2] scratch); z: first-row components.
PBE_HD bool jacobi3_firstrow(double (&d)[3], double (&e)[3], double (&z)[3]) {
    const double a0 = d[0], a1 = d[1], a2 = d[2], e0 = e[0], e1 = e[1];
    // 1. well-separated eigenvalue of the characteristic cubic, trigonometric form
    const double q = (a0 + a1 + a2) * (1.0 / 3.0);
    const double c0 = a0 - q, c1 = a1 - q, c2 = a2 - q;
    const double b1 = e0 * e0, b2 = e1 * e1;
    const double pp2 = (fma_(c0, c0, fma_(c1, c1, c2 * c2)) + 2.0 * (b1 + b2)) * (1.0 / 6.0);
    const double ipp = rsqrt_(pp2);
    const double pp = pp2 * ipp;
    const double det = c0 * fma_(c1, c2, -b2) - b1 * c2;
    const double r = 0.5 * det * (ipp * ipp * ipp); // cos(3 phi)
    double ra = fabs(r);
    ra = (ra > 1.0) ? 1.0 : ra; // NaN stays NaN
    // largest root of 4T^3 - 3T = ra; the seed only has to be good to ~1e-5 (MUFU.COS is)
#if defined(__CUDA_ARCH__)
    double T = (double)__cosf(acosf((float)ra) * (1.0f / 3.0f));
#else
    double T = (double)cosf(acosf((float)ra) * (1.0f / 3.0f));
#endif
    {
        const double T2 = T * T;
        const double hp = fma_(12.0, T2, -3.0);         // in [6, 9]
        double R = (double)(1.0f / (float)hp);
        R = R * fma_(-hp, R, 2.0);                      // 1/h' to ~1e-14
        T = fma_(-fma_(fma_(4.0, T2, -3.0), T, -ra), R, T);
        const double U2 = T * T;
        T = fma_(-fma_(fma_(4.0, U2, -3.0), T, -ra), R, T);
    }
    const double sigma = fma_((r < 0.0) ? -2.0 * pp : 2.0 * pp, T, q);

    // 2. one implicit QL sweep with shift sigma (z starts as (1, 0, 0))
    double g = a2 - sigma;
    double r2 = fma_(e1, e1, g * g);
    double ir = rsqrt_(r2);
    double s = e1 * ir, c = g * ir;
    double rr = fma_(a1 - a2, s, 2.0 * c * e1);
    double p = s * rr;
    const double d2 = a2 + p;
    g = fma_(c, rr, -e1);
    const double f = s * e0, b = c * e0;
    r2 = fma_(f, f, g * g);
    ir = rsqrt_(r2);
    const double e1n = r2 * ir;
    s = f * ir;
    c = g * ir;
    g = a1 - p;
    rr = fma_(a0 - g, s, 2.0 * c * b);
    p = s * rr;
    double d1 = g + p;
    const double e0n = fma_(c, rr, -b); // residual off-diagonal: ~0 when sigma is an eigenvalue
    double d0 = a0 - p;
    double z0 = c, z1 = s, z2 = 0.0;
    double dd2 = d2;

    // 3. the remaining 2 x 2 block in closed form
    rot2_firstrow(d1, dd2, e1n, z1, z2);

    // 4. accept, or redo with the generic solver
    const bool ok = finite_(d0) && finite_(d1) && finite_(dd2) && finite_(z0) && finite_(z1) &&
                    finite_(z2) && (fabs(e0n) <= 1.0e-13 * pp);
    if (!ok) {
#if !defined(__CUDA_ARCH__) && defined(PBE_EIG_COUNT_FALLBACKS)
        ++jacobi3_fallbacks;
#endif
        // copies made HERE: handing the caller's arrays to the out-of-line function would pin them in local memory
        // on the common path as well (STL / LDL per cell in the SASS of qmom_kernel)
        double dg[3] = {a0, a1, a2}, eg[3] = {e0, e1, 0.0}, zg[3];
        const bool okg = jacobi3_generic(dg, eg, zg);
        d[0] = dg[0], d[1] = dg[1], d[2] = dg[2];
        z[0] = zg[0], z[1] = zg[1], z[2] = zg[2];
        return okg;
    }
    cswap_firstrow(d0, d1, z0, z1);
    cswap_firstrow(d1, dd2, z1, z2);
    cswap_firstrow(d0, d1, z0, z1);
    d[0] = d0, d[1] = d1, d[2] = dd2;
    z[0] = z0, z[1] = z1, z[2] = z2;
    return true;
}

// dispatch: fixed-work solvers for the orders that have one
template <int N>
PBE_HD bool jacobi_firstrow(double (&d)[N], double (&e)[N], double (&z)[N]) {
    if constexpr (N == 2)
        return jacobi2_firstrow(d, e, z);
    else if constexpr (N == 3)
        return jacobi3_firstrow(d, e, z);
    else
        return tridiag_ql_firstrow<N>(d, e, z);
}

// un-fused FP64 multiply / subtract / add (no FMA contraction on either side of the build)
PBE_HD double mul_rn_(double a, double b) {
#if defined(__CUDA_ARCH__)
    return __dmul_rn(a, b);
#else
    volatile double r = a * b; // host builds also pass -ffp-contract=off
    return r;
#endif
}
PBE_HD double sub_rn_(double a, double b) {
#if defined(__CUDA_ARCH__)
    return __dsub_rn(a, b);
#else
    volatile double r = a - b;
    return r;
#endif
}
PBE_HD double add_rn_(double a, double b) {
#if defined(__CUDA_ARCH__)
    return __dadd_rn(a, b);
#else
    volatile double r = a + b;
    return r;
#endif
}

// ---- Wheeler (1974) moment inversion, MomentInversion.C:9-51 ---------------------------------
// The sigma recurrence is written with explicit round-to-nearest multiplies and subtracts (no
// FMA contraction) so that a, b -- and with them the NaN / Inf pattern of degenerate moment
// sets (0/0 in MomentInversion.C:30-33) -- are bit-identical to un-fused host arithmetic.
template <int N>
PBE_HD void wheeler_dev(const double (&m)[2 * N], double (&x)[N], double (&w)[N]) {
    if constexpr (N == 1) {
        x[0] = m[1] / m[0];
        w[0] = m[0];
    } else {
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
                const double t = sub_rn_(sc[col + 1], mul_rn_(a[k - 1], sc[col]));
                sn[col] = sub_rn_(t, mul_rn_(b[k - 1], sp[col]));
            }
            // sigma row 1 is the moment row: sc[1] / sc[0] at k = 1 is a[0], bit for bit
            const double prev = (k == 1) ? a[0] : sc[k] / sc[k - 1];
            a[k] = add_rn_(-prev, sn[k + 1] / sn[k]);
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
        jacobi_firstrow<N>(x, e, z);
#pragma unroll
        for (int i = 0; i < N; ++i) w[i] = m[0] * (z[i] * z[i]); // MomentInversion.C:48
    }
}

} // namespace pbe
