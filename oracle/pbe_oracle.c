/*
 * pbe_oracle.c -- CPU restatement of OpenPBE's per-cell PBE source path (see pbe_oracle.h).
 *
 * TEST INFRASTRUCTURE ONLY: the checker for tests/ and the timed CPU baseline of bench.py.
 * Loop order and operation order follow the reference so that the single-thread run is the
 * "reference-shaped" baseline: class-major outer loops, one pow/exp kernel evaluation per
 * (ordered pair, cell), a breakup-rate cache, QMOM inversion once per moment index.
 *
 * Build: make -C oracle   (gcc -O3 -std=c11 -fopenmp, no other dependency)
 */
#include "pbe_oracle.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_PI 3.14159265358979323846
#define ORC_SMALL 1.0e-15 /* OpenFOAM SMALL for WM_DP (QMOM.C:120) */
#define ORC_MAX_ORDER 16

int orc_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* ------------------------------------------------------------------------------------ */
/* scalar kernels                                                                         */
/* ------------------------------------------------------------------------------------ */

/* coalescenceKernels/CoulaloglouTavlarides.C:69-90.  The "/ Vcell" factor is commented out at
 * HEAD (:78-83); div_vcell = 1 restores it to reproduce the stale golden of
 * tests/unit_tests/coalescence.cpp:61-63. */
double orc_coal_ct(double c1, double c2, double gamma, double sigma, double xi1, double xi2,
                   double eps, double rhoc, double nuc, double vcell, int div_vcell) {
    double xi1pow13 = pow(xi1, 1.0 / 3.0);
    double xi2pow13 = pow(xi2, 1.0 / 3.0);
    double frequency = c1 * pow(eps, 1.0 / 3.0) * pow(xi1pow13 + xi2pow13, 2) *
                       pow(pow(xi1, 2.0 / 9.0) + pow(xi2, 2.0 / 9.0), 0.5) / (1 + gamma);
    if (div_vcell) frequency = frequency / vcell;
    double rate = exp((-c2 * rhoc * eps * (rhoc * nuc) / (pow(sigma, 2) * pow(1 + gamma, 3)) *
                       pow(xi1pow13 * xi2pow13 / (xi1pow13 + xi2pow13), 4.0)));
    return frequency * rate;
}

/* coalescenceKernels/constantCoalescence.H:51-54 */
double orc_coal_constant(double c) { return c; }
/* coalescenceKernels/noCoalescence.H:49-52 */
double orc_coal_none(void) { return 0.0; }

/* breakupKernels/CoulaloglouTavlarides.C:77-93 */
double orc_break_ct(double c1, double c2, double gamma, double sigma, double xi, double rho_d,
                    double eps) {
    return c1 * pow(eps, 1.0 / 3.0) *
           exp(-c2 * sigma * pow(1 + gamma, 2) /
               (rho_d * pow(eps, 2.0 / 3.0) * pow(xi, 5.0 / 9.0))) /
           ((1 + gamma) * pow(xi, 2.0 / 9.0));
}

/* breakupKernels/binaryBreakup.C:51-53 */
double orc_break_binary(double xi) { return xi * xi; }
/* breakupKernels/noBreakup.C:51-54 */
double orc_break_none(double xi) {
    (void)xi;
    return 0.0;
}

/* daughterParticleDistributions/uniformBinaryBreakup.C:55-60; OpenFOAM pos(x) = x >= 0 */
double orc_daughter_beta(double xi1, double xi2) {
    return (2.0 / xi2) * ((xi2 - xi1) >= 0 ? 1.0 : 0.0);
}
/* daughterParticleDistributions/uniformBinaryBreakup.C:62-69 */
double orc_daughter_moment(double xi, int k) { return 2 * pow(xi, (double)k) / (k + 1); }

/* dispatch on the run-time selected kernel (coalescenceKernel.C:46-71, breakupKernel.C:46-68).
 * ia / ib are class indices, only meaningful for the TABLE kernels (MOC). */
static inline double coal_S(const orc_kernel_spec *s, int N, int ia, int ib, double xa, double xb,
                            double eps, double rho, double nu, double vcell) {
    switch (s->coalescence) {
    case ORC_COAL_CONSTANT:
        return orc_coal_constant(s->coal_c[0]);
    case ORC_COAL_CT:
        return orc_coal_ct(s->coal_c[0], s->coal_c[1], s->coal_c[2], s->coal_c[3], xa, xb, eps,
                           rho, nu, vcell, s->coal_div_vcell);
    case ORC_COAL_TABLE:
        return s->coal_table[(size_t)ia * N + ib];
    default:
        return orc_coal_none();
    }
}

static inline double break_S(const orc_kernel_spec *s, int i, double xi, double rho, double eps) {
    switch (s->breakup) {
    case ORC_BR_BINARY:
        return orc_break_binary(xi);
    case ORC_BR_CT:
        return orc_break_ct(s->br_c[0], s->br_c[1], s->br_c[2], s->br_c[3], xi, rho, eps);
    case ORC_BR_TABLE:
        return s->br_table[i];
    default:
        return orc_break_none(xi);
    }
}

/* ------------------------------------------------------------------------------------ */
/* method of classes                                                                      */
/* ------------------------------------------------------------------------------------ */

int orc_moc_source(const orc_kernel_spec *spec, int N, double xi1, double relax, int64_t ncells,
                   int64_t stride, const double *n, const double *eps, const double *rho,
                   const double *nu, const double *vcell, double dt, double *S, double *scale,
                   double *coal_only, double *br_only, int nthreads) {
    if (N < 1 || ncells < 0 || stride < ncells) return 1;
    if (nthreads < 1) nthreads = 1;
    /* xi_[i] = deltaXi_ * (i + 1)   (MOC.C:79) */
    double *xi = (double *)malloc(sizeof(double) * (size_t)N);
    /* breakupCache_[i * phaseSize + j] = breakup_->S(xi_[i], j)   (MOC.C:165-171) */
    double *cache = (double *)malloc(sizeof(double) * (size_t)N * (size_t)(ncells > 0 ? ncells : 1));
    if (!xi || !cache) {
        free(xi);
        free(cache);
        return 2;
    }
    for (int i = 0; i < N; ++i) xi[i] = xi1 * (i + 1);

    for (int i = 0; i < N; ++i) {
#pragma omp parallel for schedule(static) num_threads(nthreads) if (nthreads > 1)
        for (int64_t c = 0; c < ncells; ++c)
            cache[(size_t)i * ncells + c] = break_S(spec, i, xi[i], rho[c], eps[c]);
    }

    /* forAll(classNumberDensity_, k) S.set(k, classSourceTerm(k))   (MOC.C:174-175) */
    for (int i = 0; i < N; ++i) {
#pragma omp parallel for schedule(static) num_threads(nthreads) if (nthreads > 1)
        for (int64_t c = 0; c < ncells; ++c) {
            const double vc = vcell ? vcell[c] : 1.0;
            const double ni = n[(size_t)i * stride + c];
            double sc = 0.0;
            /* coalescenceSourceTerm(i), MOC.C:106-122 */
            double f = 0.0, fs = 0.0;
            for (int j = 0; j < N - i - 1; ++j) {
                double t = coal_S(spec, N, i, j, xi[i], xi[j], eps[c], rho[c], nu[c], vc) *
                           n[(size_t)j * stride + c];
                f -= t;
                fs += fabs(t);
            }
            f *= ni;
            sc += fs * fabs(ni);
            for (int j = 0; j < i; ++j) {
                double t = 0.5 *
                           coal_S(spec, N, i - j - 1, j, xi[i - j - 1], xi[j], eps[c], rho[c],
                                  nu[c], vc) *
                           n[(size_t)(i - j - 1) * stride + c] * n[(size_t)j * stride + c];
                f += t;
                sc += fabs(t);
            }
            /* breakupSourceTerm(i), MOC.C:142-155 */
            double h = 0.0;
            {
                double t = cache[(size_t)i * ncells + c] * ni;
                h -= t;
                sc += fabs(t);
            }
            for (int j = i + 1; j < N; ++j) {
                double t = orc_daughter_beta(xi[i], xi[j]) * cache[(size_t)j * ncells + c] *
                           n[(size_t)j * stride + c] * xi1;
                h += t;
                sc += fabs(t);
            }
            /* classSourceTerm(i), MOC.C:87-91 */
            double lim = -relax * ni / dt;
            double s = f + h;
            S[(size_t)i * stride + c] = s > lim ? s : lim; /* Foam max(a,b) */
            if (scale) scale[(size_t)i * stride + c] = sc;
            if (coal_only) coal_only[(size_t)i * stride + c] = f;
            if (br_only) br_only[(size_t)i * stride + c] = h;
        }
    }
    free(xi);
    free(cache);
    return 0;
}

void orc_moc_update(int N, int64_t ncells, int64_t stride, double *n, const double *S, double dt) {
    for (int k = 0; k < N; ++k)
        for (int64_t c = 0; c < ncells; ++c)
            n[(size_t)k * stride + c] += dt * S[(size_t)k * stride + c];
}

/* volumeToDiameter, MOC.C:43-45 */
static inline double volume_to_diameter(double v) { return pow(6.0 / ORC_PI * v, 1.0 / 3.0); }

void orc_moc_d(int N, double xi1, int64_t ncells, int64_t stride, const double *n, double *d) {
    for (int64_t c = 0; c < ncells; ++c) {
        double sum = volume_to_diameter(xi1 * 1) * n[c];
        double norm = 1.0 * n[c];
        for (int i = 1; i < N; ++i) {
            sum += volume_to_diameter(xi1 * (i + 1)) * n[(size_t)i * stride + c];
            norm += n[(size_t)i * stride + c];
        }
        d[c] = sum / norm;
    }
}

double orc_moc_continuity(int N, double xi1, int64_t ncells, int64_t stride, const double *n,
                          const double *alpha, const double *vcell, double *err) {
    double mx = 0.0;
    for (int64_t c = 0; c < ncells; ++c) {
        double vs = 0.0;
        for (int k = 0; k < N; ++k) vs += (xi1 * (k + 1)) * n[(size_t)k * stride + c];
        double e = fabs(vs / vcell[c] - alpha[c]);
        if (err) err[c] = e;
        if (e > mx) mx = e;
    }
    return mx;
}

/* ------------------------------------------------------------------------------------ */
/* Wheeler inversion                                                                      */
/* ------------------------------------------------------------------------------------ */

/* Implicit-shift QL on a symmetric tridiagonal matrix, carrying only the first row of the
 * eigenvector matrix (all the Golub-Welsch weights need).  Stands in for
 * Eigen::SelfAdjointEigenSolver at MomentInversion.C:46-48. */
int orc_tridiag_eig_firstrow(int n, double *d, double *e, double *z) {
    int bad = 0;
    for (int i = 0; i < n; ++i) {
        z[i] = (i == 0) ? 1.0 : 0.0;
        if (!isfinite(d[i])) bad = 1;
        if (i < n - 1 && !isfinite(e[i])) bad = 1;
    }
    e[n - 1] = 0.0; /* sentinel: e has n slots, the last one is scratch */
    if (!bad) {
        for (int l = 0; l < n && !bad; ++l) {
            int iter = 0;
            for (;;) {
                int m = l;
                for (; m < n - 1; ++m) {
                    double dd = fabs(d[m]) + fabs(d[m + 1]);
                    if (fabs(e[m]) <= DBL_EPSILON * dd) break;
                }
                if (m == l) break;
                if (++iter > 60) {
                    bad = 1;
                    break;
                }
                double g = (d[l + 1] - d[l]) / (2.0 * e[l]);
                double r = hypot(g, 1.0);
                g = d[m] - d[l] + e[l] / (g + copysign(r, g));
                double s = 1.0, c = 1.0, p = 0.0;
                int i = m - 1;
                for (; i >= l; --i) {
                    double f = s * e[i];
                    double b = c * e[i];
                    r = hypot(f, g);
                    e[i + 1] = r;
                    if (r == 0.0) {
                        d[i + 1] -= p;
                        e[m] = 0.0;
                        break;
                    }
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
                if (r == 0.0 && i >= l) continue;
                d[l] -= p;
                e[l] = g;
                e[m] = 0.0;
            }
        }
    }
    if (bad) {
        for (int i = 0; i < n; ++i) {
            d[i] = NAN;
            z[i] = NAN;
        }
        return 1;
    }
    /* eigenvalues ascending (Eigen's documented order) */
    for (int i = 1; i < n; ++i) {
        double dv = d[i], zv = z[i];
        int j = i - 1;
        while (j >= 0 && d[j] > dv) {
            d[j + 1] = d[j];
            z[j + 1] = z[j];
            --j;
        }
        d[j + 1] = dv;
        z[j + 1] = zv;
    }
    return 0;
}

/* MomentInversion.C:9-51 */
void orc_wheeler(int order, const double *m, double *x, double *w) {
    const int N = order;
    if (N < 1 || N > ORC_MAX_ORDER) return;
    double sigma[ORC_MAX_ORDER + 1][2 * ORC_MAX_ORDER];
    double a[ORC_MAX_ORDER], b[ORC_MAX_ORDER], e[ORC_MAX_ORDER], z[ORC_MAX_ORDER];
    memset(sigma, 0, sizeof(sigma));
    for (int j = 0; j < 2 * N; ++j) sigma[1][j] = m[j];
    for (int k = 0; k < N; ++k) a[k] = b[k] = 0.0;
    a[0] = m[1] / m[0];
    b[0] = m[0];
    for (int k = 1; k < N; ++k) {
        int Ncol = 2 * (N - k);
        for (int j = 0; j < Ncol; ++j)
            sigma[k + 1][k + j] =
                sigma[k][k + 1 + j] - a[k - 1] * sigma[k][k + j] - b[k - 1] * sigma[k - 1][k + j];
        a[k] = -sigma[k][k] / sigma[k][k - 1] + sigma[k + 1][k + 1] / sigma[k + 1][k];
        b[k] = sigma[k + 1][k] / sigma[k][k - 1];
    }
    /* jacobi(i, i+1) = jacobi(i+1, i) = -sqrt(|b(i+1)|)   (MomentInversion.C:35-41) */
    for (int i = 0; i < N - 1; ++i) e[i] = -sqrt(fabs(b[i + 1]));
    for (int i = 0; i < N; ++i) x[i] = a[i];
    orc_tridiag_eig_firstrow(N, x, e, z);
    for (int i = 0; i < N; ++i) w[i] = m[0] * (z[i] * z[i]);
}

/* ------------------------------------------------------------------------------------ */
/* QMOM                                                                                   */
/* ------------------------------------------------------------------------------------ */

/* momentSourceTerm(momenti) for one cell, QMOM.C:226-264 */
static inline void qmom_cell_source(const orc_kernel_spec *spec, int order, int k, const double *x,
                                    const double *w, double eps, double rho, double nu, double vc,
                                    double *out, double *scale) {
    const int N = order;
    double coalescenceSource = 0.0, breakupSource = 0.0, sc = 0.0;
    for (int i = 0; i < N; ++i) {
        double xi_i = x[i];
        double death = 0.0, birth = 0.0, sdeath = 0.0, sbirth = 0.0;
        for (int j = 0; j < N; ++j) {
            double xi_j = x[j];
            double Sd = coal_S(spec, N, i, j, xi_i, xi_j, eps, rho, nu, vc);
            death += w[j] * Sd;
            sdeath += fabs(w[j] * Sd);
            double Sb = coal_S(spec, N, i, j, xi_i, xi_j, eps, rho, nu, vc);
            double tb = w[j] * pow(xi_i + xi_j, (double)k) * Sb;
            birth += tb;
            sbirth += fabs(tb);
        }
        coalescenceSource += w[i] * (birth / 2. - pow(xi_i, (double)k) * death);
        sc += fabs(w[i]) * (sbirth / 2. + fabs(pow(xi_i, (double)k)) * sdeath);
        double g = break_S(spec, i, xi_i, rho, eps);
        double dm = orc_daughter_moment(xi_i, k);
        double pk = pow(xi_i, (double)k);
        breakupSource += w[i] * g * (dm - pk);
        sc += fabs(w[i] * g) * (fabs(dm) + fabs(pk));
    }
    *out = coalescenceSource + breakupSource;
    if (scale) *scale = sc;
}

int orc_qmom_source(const orc_kernel_spec *spec, int order, int64_t ncells, int64_t stride,
                    const double *m, const double *eps, const double *rho, const double *nu,
                    const double *vcell, double *S, double *scale, int reference_shaped,
                    int nthreads) {
    if (order < 1 || order > ORC_MAX_ORDER || stride < ncells) return 1;
    if (spec->coalescence == ORC_COAL_TABLE || spec->breakup == ORC_BR_TABLE) return 3;
    if (nthreads < 1) nthreads = 1;
    const int nm = 2 * order;
    if (reference_shaped) {
        /* for k: mSources[k] = momentSourceTerm(k)  (QMOM.C:88-95): moment-major, inversion
         * recomputed for every k (QMOM.C:231) */
        for (int k = 0; k < nm; ++k) {
#pragma omp parallel for schedule(static) num_threads(nthreads) if (nthreads > 1)
            for (int64_t c = 0; c < ncells; ++c) {
                double mv[2 * ORC_MAX_ORDER], x[ORC_MAX_ORDER], w[ORC_MAX_ORDER];
                for (int i = 0; i < nm; ++i) mv[i] = m[(size_t)i * stride + c];
                orc_wheeler(order, mv, x, w);
                qmom_cell_source(spec, order, k, x, w, eps[c], rho[c], nu[c],
                                 vcell ? vcell[c] : 1.0, &S[(size_t)k * stride + c],
                                 scale ? &scale[(size_t)k * stride + c] : NULL);
            }
        }
    } else {
#pragma omp parallel for schedule(static) num_threads(nthreads) if (nthreads > 1)
        for (int64_t c = 0; c < ncells; ++c) {
            double mv[2 * ORC_MAX_ORDER], x[ORC_MAX_ORDER], w[ORC_MAX_ORDER];
            for (int i = 0; i < nm; ++i) mv[i] = m[(size_t)i * stride + c];
            orc_wheeler(order, mv, x, w);
            for (int k = 0; k < nm; ++k)
                qmom_cell_source(spec, order, k, x, w, eps[c], rho[c], nu[c],
                                 vcell ? vcell[c] : 1.0, &S[(size_t)k * stride + c],
                                 scale ? &scale[(size_t)k * stride + c] : NULL);
        }
    }
    return 0;
}

void orc_qmom_update(int order, int64_t ncells, int64_t stride, double *m, double *S, double dt,
                     double *d) {
    const int nm = 2 * order;
    for (int k = 0; k < nm; ++k)
        for (int64_t c = 0; c < ncells; ++c) {
            double *s = &S[(size_t)k * stride + c];
            if (isnan(*s)) *s = 0.0;                /* QMOM.C:101-107 */
            double v = m[(size_t)k * stride + c] + dt * (*s); /* zero-flux explicitSolve */
            m[(size_t)k * stride + c] = v > ORC_SMALL ? v : ORC_SMALL; /* QMOM.C:117-123 */
        }
    if (d && nm >= 2)
        for (int64_t c = 0; c < ncells; ++c)
            d[c] = pow(6.0 / ORC_PI * m[(size_t)1 * stride + c] / m[c], 1.0 / 3.0); /* QMOM.C:125 */
}

/* ------------------------------------------------------------------------------------ */
/* semi-infinite integral: QUADPACK QAGI restated (Integrator.H:12-27 calls                */
/* gsl_integration_qagiu(f, lower, 1e-7, 0, 1000, ...)).  Map x = lower + (1-t)/t, t in    */
/* (0,1]; 15-point Gauss-Kronrod transformed rule on each sub-interval; bisect the         */
/* sub-interval with the largest error estimate until sum(err) <= max(epsabs, epsrel|I|).  */
/* (QUADPACK additionally applies the epsilon algorithm; it never triggers for the smooth  */
/* integrands on this path and is omitted.)                                                */
/* ------------------------------------------------------------------------------------ */

static const double GK15_XGK[8] = {0.991455371120812639206854697526329, 0.949107912342758524526189684047851,
                                   0.864864423359769072789712788640926, 0.741531185599394439863864773280788,
                                   0.586087235467691130294144838258730, 0.405845151377397166906606412076961,
                                   0.207784955007898467600689403773245, 0.000000000000000000000000000000000};
static const double GK15_WGK[8] = {0.022935322010529224963732008058970, 0.063092092629978553290700663189204,
                                   0.104790010322250183839876322541518, 0.140653259715525918745189590510238,
                                   0.169004726639267902826583426598550, 0.190350578064785409913256402421014,
                                   0.204432940075298892414161999234649, 0.209482141084727828012999174891714};
/* 7-point Gauss weights, aligned to xgk[1], xgk[3], xgk[5], xgk[7] */
static const double GK15_WG[4] = {0.129484966168869693270611432679082, 0.279705391489276667901467771423780,
                                  0.381830050505118944950369775488975, 0.417959183673469387755102040816327};

static double qagiu_transformed(orc_integrand f, void *p, double lower, double t) {
    double x = lower + (1.0 - t) / t;
    return f(x, p) / (t * t);
}

static void qk15_transformed(orc_integrand f, void *p, double lower, double a, double b,
                             double *result, double *abserr) {
    const double center = 0.5 * (a + b), half = 0.5 * (b - a);
    double fv1[7], fv2[7];
    const double fc = qagiu_transformed(f, p, lower, center);
    double resk = GK15_WGK[7] * fc;
    double resg = GK15_WG[3] * fc;
    double resabs = fabs(resk);
    for (int j = 0; j < 7; ++j) {
        double dx = half * GK15_XGK[j];
        double f1 = qagiu_transformed(f, p, lower, center - dx);
        double f2 = qagiu_transformed(f, p, lower, center + dx);
        fv1[j] = f1;
        fv2[j] = f2;
        resk += GK15_WGK[j] * (f1 + f2);
        resabs += GK15_WGK[j] * (fabs(f1) + fabs(f2));
        if (j & 1) resg += GK15_WG[j / 2] * (f1 + f2);
    }
    const double mean = 0.5 * resk;
    double resasc = GK15_WGK[7] * fabs(fc - mean);
    for (int j = 0; j < 7; ++j) resasc += GK15_WGK[j] * (fabs(fv1[j] - mean) + fabs(fv2[j] - mean));
    double err = fabs((resk - resg) * half);
    resabs *= fabs(half);
    resasc *= fabs(half);
    if (resasc != 0.0 && err != 0.0) {
        double scale = pow(200.0 * err / resasc, 1.5);
        err = scale < 1.0 ? resasc * scale : resasc;
    }
    if (resabs > DBL_MIN / (50.0 * DBL_EPSILON)) {
        double min_err = 50.0 * DBL_EPSILON * resabs;
        if (min_err > err) err = min_err;
    }
    *result = resk * half;
    *abserr = err;
}

double orc_integrate_qagiu(orc_integrand f, void *params, double lower, double epsabs,
                           double epsrel, int limit, double *abserr, int *neval) {
    if (limit < 1) limit = 1;
    double *a = (double *)malloc(sizeof(double) * 4 * (size_t)limit);
    double *b = a + limit, *r = b + limit, *e = r + limit;
    int n = 1, evals = 15;
    a[0] = 0.0;
    b[0] = 1.0;
    qk15_transformed(f, params, lower, 0.0, 1.0, &r[0], &e[0]);
    double result = r[0], errsum = e[0];
    while (n < limit) {
        double tol = fmax(epsabs, epsrel * fabs(result));
        if (errsum <= tol) break;
        int worst = 0;
        for (int i = 1; i < n; ++i)
            if (e[i] > e[worst]) worst = i;
        double a1 = a[worst], b2 = b[worst], mid = 0.5 * (a1 + b2);
        double r1, e1, r2, e2;
        qk15_transformed(f, params, lower, a1, mid, &r1, &e1);
        qk15_transformed(f, params, lower, mid, b2, &r2, &e2);
        evals += 30;
        b[worst] = mid;
        r[worst] = r1;
        e[worst] = e1;
        a[n] = mid;
        b[n] = b2;
        r[n] = r2;
        e[n] = e2;
        ++n;
        result = 0.0;
        errsum = 0.0;
        for (int i = 0; i < n; ++i) {
            result += r[i];
            errsum += e[i];
        }
    }
    if (abserr) *abserr = errsum;
    if (neval) *neval = evals;
    free(a);
    return result;
}

static double exp_neg(double x, void *p) {
    (void)p;
    return exp(-x);
}
/* tests/unit_tests/integration.cpp:6-13 via Integrator.H:21-22 (epsabs 1e-7, epsrel 0, 1000) */
double orc_integrate_exp_neg(double lower) {
    return orc_integrate_qagiu(exp_neg, NULL, lower, 1e-7, 0.0, 1000, NULL, NULL);
}

/* ---- transport (row f2, first slice): see pbe_oracle.h for the OpenFOAM functions restated --- */
void orc_transport_upwind_explicit(int64_t ncells, int64_t nint, const int32_t *owner,
                                   const int32_t *neighbour, int64_t nbnd, const int32_t *bowner,
                                   const double *vcell, int nfields, int64_t stride,
                                   const double *psi, const double *phi, const double *psi_b,
                                   int64_t stride_b, const double *S, double dt, double *out) {
    orc_transport_upwind_explicit_coupled(ncells, nint, owner, neighbour, nbnd, bowner, vcell, nfields,
                                          stride, psi, phi, psi_b, stride_b, 0, 0, NULL, 0, S, dt, out);
}

void orc_transport_upwind_explicit_coupled(int64_t ncells, int64_t nint, const int32_t *owner,
                                           const int32_t *neighbour, int64_t nbnd,
                                           const int32_t *bowner, const double *vcell, int nfields,
                                           int64_t stride, const double *psi, const double *phi,
                                           const double *psi_b, int64_t stride_b, int64_t coupled0,
                                           int64_t ncoupled, const double *psi_halo,
                                           int64_t stride_h, const double *S, double dt,
                                           double *out) {
    const double rDeltaT = 1.0 / dt;
    double *ivf = (double *)malloc(sizeof(double) * (size_t)(ncells > 0 ? ncells : 1));
    for (int k = 0; k < nfields; ++k) {
        const double *f = psi + (size_t)k * stride;
        for (int64_t c = 0; c < ncells; ++c) ivf[c] = 0.0;
        for (int64_t fi = 0; fi < nint; ++fi) { /* upwind<scalar>(mesh, phi).flux(psi) */
            const double lambda = phi[fi] >= 0.0 ? 1.0 : 0.0; /* pos() */
            const double P = f[owner[fi]], N = f[neighbour[fi]];
            const double face = lambda * (P - N) + N;
            const double flux = phi[fi] * face;
            ivf[owner[fi]] += flux;
            ivf[neighbour[fi]] -= flux;
        }
        for (int64_t b = 0; b < nbnd; ++b) {
            double face;
            if (b >= coupled0 && b < coupled0 + ncoupled) {
                /* coupled (processor) patch: pLambda patchInternalField + (1 - pLambda)
                 * patchNeighbourField, pLambda = pos(phi_b) */
                const double w = phi[nint + b] >= 0.0 ? 1.0 : 0.0;
                face = w * f[bowner[b]] + (1.0 - w) * psi_halo[(size_t)k * stride_h + (b - coupled0)];
            } else {
                face = psi_b ? psi_b[(size_t)k * stride_b + b] : f[bowner[b]];
            }
            ivf[bowner[b]] += phi[nint + b] * face;
        }
        for (int64_t c = 0; c < ncells; ++c) {
            const double div = ivf[c] / vcell[c];
            const double Su = S ? S[(size_t)k * stride + c] : 0.0;
            out[(size_t)k * stride + c] = (f[c] * rDeltaT + Su - div) / rDeltaT;
        }
    }
    free(ivf);
}
