/*
 * pbe_oracle.h -- CPU restatement of OpenPBE's per-cell population-balance source path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it,
 * and there only as the checker / the timed CPU baseline.  The product (libpbe_b200.so) never
 * links, loads or calls this file and has no CPU fallback.
 *
 * Parity status: the reference (robertsawko/OpenPBE) cannot be compiled here -- every source
 * file on the path includes OpenFOAM headers and MomentInversion.C needs Eigen3; neither is in
 * the image.  This restatement is therefore pinned on the reference's own
 * golden vectors: tests/unit_tests/breakup.cpp:6-59, coalescence.cpp:6-78,
 * moment_inversion.cpp:15-49, integration.cpp:6-13, and on the 0-D analytical solutions of
 * tests/integration/MOC/testConvergence.py:10-47 (tests/test_oracle_*.py).  The Eigen
 * SelfAdjointEigenSolver call (MomentInversion.C:46-48; Eigen3 un-vendored, unpinned) is
 * restated as an implicit-shift QL iteration on the symmetric tridiagonal Jacobi matrix and
 * cross-checked against numpy.linalg.eigh.  The GSL QAGIU call (Integrator.H:21-22; GSL
 * un-vendored, unpinned) is restated from the published QUADPACK algorithm (15-point
 * Gauss-Kronrod on the (0,1] transform with adaptive bisection).
 *
 * All citations are file:line relative to the reference root.
 */
#ifndef PBE_ORACLE_H
#define PBE_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* kernel selectors: names follow the reference TypeNames */
enum { ORC_COAL_NONE = 0, ORC_COAL_CONSTANT = 1, ORC_COAL_CT = 2, ORC_COAL_TABLE = 3 };
enum { ORC_BR_NONE = 0, ORC_BR_BINARY = 1, ORC_BR_CT = 2, ORC_BR_TABLE = 3 };
enum { ORC_DAUGHTER_UNIFORM_BINARY = 0 };

typedef struct {
    int coalescence;     /* ORC_COAL_* */
    double coal_c[4];    /* constant: c[0]; CT: c1, c2, gamma, sigma */
    int breakup;         /* ORC_BR_* */
    double br_c[4];      /* CT: c1, c2, gamma, sigma */
    int daughter;        /* ORC_DAUGHTER_* */
    int coal_div_vcell;  /* 1 = historical "/ Vcell" variant (the stale golden of
                            tests/unit_tests/coalescence.cpp:61-63); HEAD = 0 */
    const double *coal_table; /* ORC_COAL_TABLE: [N][N] cell-independent S(xi_i, xi_j) */
    const double *br_table;   /* ORC_BR_TABLE:   [N]    cell-independent g(xi_i)       */
} orc_kernel_spec;

/* ---- scalar kernels (the reference's Foam-free *Impl functors) ---------------------- */
double orc_coal_ct(double c1, double c2, double gamma, double sigma, double xi1, double xi2,
                   double eps, double rhoc, double nuc, double vcell, int div_vcell);
double orc_coal_constant(double c);
double orc_coal_none(void);
double orc_break_ct(double c1, double c2, double gamma, double sigma, double xi, double rho_d,
                    double eps);
double orc_break_binary(double xi);
double orc_break_none(double xi);
double orc_daughter_beta(double xi1, double xi2);
double orc_daughter_moment(double xi, int k);

/* ---- method of classes -------------------------------------------------------------- */
/* Fields are class-major SoA: n[i*stride + c], i < N, c < ncells (one contiguous field per
 * class, as MOC.H:66).  S receives max(C_i + B_i, -relax n_i/dt) (MOC.C:87-91).  scale (may
 * be NULL) receives sum|terms| of every product that entered S_i: the parity tolerance base.
 * coal_only / br_only (may be NULL) receive C_i and B_i before the limiter.  vcell may be
 * NULL (only read when spec->coal_div_vcell).  nthreads <= 1: the reference's single thread. */
int orc_moc_source(const orc_kernel_spec *spec, int N, double xi1, double relax, int64_t ncells,
                   int64_t stride, const double *n, const double *eps, const double *rho,
                   const double *nu, const double *vcell, double dt, double *S, double *scale,
                   double *coal_only, double *br_only, int nthreads);
/* zero-flux explicit update, MOC.C:259-272 with phi = 0: n_k += dt S_k */
void orc_moc_update(int N, int64_t ncells, int64_t stride, double *n, const double *S, double dt);
/* number-mean diameter, MOC.C:282-290 */
void orc_moc_d(int N, double xi1, int64_t ncells, int64_t stride, const double *n, double *d);
/* continuity check MOC.C:254-279: err[c] = | sum_k xi_k n_k / V - alpha | ; returns max */
double orc_moc_continuity(int N, double xi1, int64_t ncells, int64_t stride, const double *n,
                          const double *alpha, const double *vcell, double *err);

/* ---- Wheeler inversion (MomentInversion.C:9-51) -------------------------------------- */
/* order = number of nodes; m has 2*order entries; x ascending, w = m0 * v0^2 */
void orc_wheeler(int order, const double *m, double *x, double *w);
/* symmetric tridiagonal eigen-solve: diag d[n], off-diag in e[0..n-2] (e must have n slots,
 * the last is scratch); on exit d ascending and
 * z[i] = first component of the i-th normalised eigenvector.  Returns 0, or 1 when the
 * iteration limit was hit / the input was not finite (outputs are then NaN). */
int orc_tridiag_eig_firstrow(int n, double *d, double *e, double *z);

/* ---- QMOM (QMOM.C:215-267, 98-128) ---------------------------------------------------- */
/* m[k*stride + c], k < 2*order.  S[k*stride + c] = coalescence + breakup moment source (NOT yet
 * NaN-scrubbed).  scale as above.  reference_shaped != 0 recomputes the inversion for every
 * k exactly as QMOM.C:231 does (timed baseline); 0 inverts once per cell (same results). */
int orc_qmom_source(const orc_kernel_spec *spec, int order, int64_t ncells, int64_t stride,
                    const double *m, const double *eps, const double *rho, const double *nu,
                    const double *vcell, double *S, double *scale, int reference_shaped,
                    int nthreads);
/* NaN -> 0 scrub (QMOM.C:101-107), m += dt S (QMOM.C:193-197 with phi = 0), clamp to SMALL
 * (QMOM.C:117-123), d = (6/pi m1/m0)^(1/3) (QMOM.C:125).  d may be NULL. */
void orc_qmom_update(int order, int64_t ncells, int64_t stride, double *m, double *S, double dt,
                     double *d);

/* ---- semi-infinite integral (Integrator.H:12-27) -------------------------------------- */
typedef double (*orc_integrand)(double x, void *params);
double orc_integrate_qagiu(orc_integrand f, void *params, double lower, double epsabs,
                           double epsrel, int limit, double *abserr, int *neval);
/* convenience for tests: integral of exp(-x) from `lower` (tests/unit_tests/integration.cpp) */
double orc_integrate_exp_neg(double lower);

/* ---- transport, SURVEY 8 row f2 (first slice): MULES::explicitSolve with the upwind face flux,
 * i.e. MOC::solveWithMULES (MOC.C:198-272) / QMOM::solveWithMULES (QMOM.C:131-197) when the
 * divergence scheme of the case is `Gauss upwind` (the limited correction phiCorr is then zero).
 * PARITY UNPINNED: OpenFOAM is a third-party dependency absent from the reference tree and from
 * this image; the loops restate OpenFOAM 3.0.x from its published sources:
 *   face value   lambda (psi_P - psi_N) + psi_N, lambda = pos(phi_f)    (surfaceInterpolationScheme
 *                ::interpolate with upwind::weights = pos(faceFlux)); boundary faces take the
 *                patch value psi_b (NULL = zero-gradient = the owner cell's value)
 *   divergence   fvc::surfaceIntegrate: ivf[own] += flux, ivf[nei] -= flux over internal faces
 *                in face order, then boundary faces, then ivf /= V
 *   update       psi = (psi0 rDeltaT + Su - div) / rDeltaT, rDeltaT = 1/dt  (rho = 1, Sp = 0)
 * Faces: OpenFOAM addressing, internal faces 0..nint-1 (owner, neighbour), boundary faces
 * nint..nint+nbnd-1 with their owner cell; phi has nint + nbnd entries. */
void orc_transport_upwind_explicit(int64_t ncells, int64_t nint, const int32_t *owner,
                                   const int32_t *neighbour, int64_t nbnd, const int32_t *bowner,
                                   const double *vcell, int nfields, int64_t stride,
                                   const double *psi, const double *phi, const double *psi_b,
                                   int64_t stride_b, const double *S, double dt, double *out);
/* The same on one sub-domain of a decomposed mesh: boundary faces coupled0 .. coupled0+ncoupled-1
 * are processor-patch faces; psi_halo[k*stride_h + j] is the value of the cell across coupled face
 * j (patchNeighbourField) and the face value is OpenFOAM's coupled-patch interpolation
 * pLambda psi_own + (1 - pLambda) psi_halo with pLambda = pos(phi_b) (surfaceInterpolationScheme::
 * interpolate, coupled branch).  A decomposed run therefore differs from the serial one in the last
 * bit on the cells next to a processor patch -- in OpenFOAM and here alike. */
void orc_transport_upwind_explicit_coupled(int64_t ncells, int64_t nint, const int32_t *owner,
                                           const int32_t *neighbour, int64_t nbnd,
                                           const int32_t *bowner, const double *vcell, int nfields,
                                           int64_t stride, const double *psi, const double *phi,
                                           const double *psi_b, int64_t stride_b, int64_t coupled0,
                                           int64_t ncoupled, const double *psi_halo,
                                           int64_t stride_h, const double *S, double dt,
                                           double *out);

/* ---- transport, SURVEY 8 row f2 (second slice): the MULES-limited path of MOC::solveWithMULES
 * (MOC.C:198-280) / QMOM::solveWithMULES (QMOM.C:131-197) for the schemes the reference's cases
 * select (`Gauss vanLeer`, tests/validation/simmons_and_azzopardi/c01/system/fvSchemes:34; and
 * `Gauss upwind`).  pbe_oracle_mules.c; PARITY UNPINNED (OpenFOAM 3.0.x restated, see that file). */
typedef struct orc_mesh {
    int64_t ncells, nint, nbnd;
    const int32_t *owner, *neighbour, *bown; /* OpenFOAM face addressing; bown = faceCells of all patches */
    const double *V;                         /* mesh.V() */
    const double *Sf;                        /* [nint + nbnd][3] face area vectors (limited schemes) or NULL */
    const double *C;                         /* [ncells][3] cell centres or NULL */
    const double *weights;                   /* [nint] linear interpolation weights (mesh.weights()) or NULL */
} orc_mesh;
enum { ORC_DIV_UPWIND = 0, ORC_DIV_VANLEER = 1 };
void orc_upwind_flux(const orc_mesh *m, const double *phi, const double *psi, const double *psi_b, double *out);
void orc_gauss_linear_grad(const orc_mesh *m, const double *psi, const double *psi_b, double *grad);
void orc_vanleer_flux(const orc_mesh *m, const double *phi, const double *psi, const double *psi_b, double *out);
int orc_div_flux(const orc_mesh *m, int scheme, const double *phi, const double *psi, const double *psi_b, double *out);
void orc_mules_limiter(const orc_mesh *m, double *lambda, double rDeltaT, const double *psi, const double *psi_b,
                       const double *phiBD, const double *phiCorr, const double *Su, double psiMax, double psiMin,
                       int nLimiterIter);
void orc_mules_limit(const orc_mesh *m, double rDeltaT, const double *psi, const double *psi_b, const double *phi,
                     double *phiPsi, const double *Su, double psiMax, double psiMin, int nLimiterIter);
void orc_mules_limit_sum(int nfields, int64_t nint, int64_t stride_f, double *corr);
void orc_mules_explicit_solve(const orc_mesh *m, double rDeltaT, double *psi, const double *phiPsi, const double *Su);
int orc_transport_mules(const orc_mesh *m, int scheme, int nfields, int64_t stride, double *psi, const double *phi,
                        const double *psi_b, int64_t stride_b, const double *S, const double *psiMax, double psiMin,
                        int nLimiterIter, double dt);

int orc_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
