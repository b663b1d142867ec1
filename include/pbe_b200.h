/*
 * pbe_b200.h -- C ABI of libpbe_b200.so, the B200 (sm_100a) population-balance source engine.
 *
 * This is the drop-in boundary for the per-cell hot path of robertsawko/OpenPBE: what an
 * OpenFOAM-side `libPBE.so` adapter binds instead of running the reference's
 * CPU loops.  Plain pointers and sizes only; every function returns 0 on success or a
 * PBE_E_* code, with a per-thread message behind pbe_last_error().  There is NO CPU fallback:
 * every compute entry point launches sm_100a kernels and fails with PBE_E_CUDA without a device.
 *
 * Field layout (device and host entry points alike): class-major / moment-major
 * structure-of-arrays FP64, `f[k*stride + c]`, cell index fastest -- the layout of the
 * reference's PtrList<volScalarField> (src/PBESystem/MOC.H:66, QMOM.H:62) so an adapter
 * stages OpenFOAM fields without transposition.
 *
 * Citations are file:line relative to the reference root.
 */
#ifndef PBE_B200_H
#define PBE_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PBE_B200_VERSION 100 /* 0.1.0 */

/* ---- status codes (reference convention: FatalErrorIn(...) << exit(FatalError),
 * PBEMethod.C:97-105; the adapter turns a non-zero status into FatalErrorIn) ------------- */
enum {
    PBE_OK = 0,
    PBE_E_INVALID = 1,   /* bad argument (NULL handle, N < 1, stride < ncells, ...) */
    PBE_E_UNKNOWN_TYPE = 2, /* unknown kernel / method TypeName (coalescenceKernel.C:57-65) */
    PBE_E_CUDA = 3,      /* CUDA runtime error or no device: message has the cudaError string */
    PBE_E_UNSUPPORTED = 4, /* valid request outside the device path's limits (e.g. N too big) */
    PBE_E_NOMEM = 5
};

/* ---- run-time selectable kernels.  Values follow the reference TypeNames:
 * coalescenceKernel <- noCoalescence | constantCoalescence | CoulaloglouTavlaridesC
 *   (coalescenceKernels/{noCoalescence.H:49-52, constantCoalescence.H:51-54,
 *    CoulaloglouTavlarides.C:69-90}),
 * breakupKernel <- noBreakup | binaryBreakup | CoulaloglouTavlarides
 *   (breakupKernels/{noBreakup.C:51-54, binaryBreakup.C:51-53, CoulaloglouTavlarides.C:77-93}),
 * daughterParticleDistribution <- uniformBinaryBreakup
 *   (daughterParticleDistributions/uniformBinaryBreakup.C:55-69).
 * *_TABLE is the plug-in path for user kernels that do not depend on the cell
 * (coalescenceKernel.H:85-90 virtual S()): the host evaluates the user functor once per
 * class pair / class and uploads the table (MOC only). */
enum { PBE_COAL_NONE = 0, PBE_COAL_CONSTANT = 1, PBE_COAL_CT = 2, PBE_COAL_TABLE = 3 };
enum { PBE_BR_NONE = 0, PBE_BR_BINARY = 1, PBE_BR_CT = 2, PBE_BR_TABLE = 3 };
enum { PBE_DAUGHTER_UNIFORM_BINARY = 0 };

typedef struct pbe_kernel_spec {
    int coalescence;          /* PBE_COAL_* */
    double coal_c[4];         /* constant: c[0] (dict key `constant`); CT: c1, c2, gamma, sigma */
    int breakup;              /* PBE_BR_* */
    double br_c[4];           /* CT: c1, c2, gamma, sigma */
    int daughter;             /* PBE_DAUGHTER_* */
    const double *coal_table; /* PBE_COAL_TABLE: HOST [N][N], S(xi_i, xi_j); may be asymmetric */
    const double *br_table;   /* PBE_BR_TABLE:   HOST [N],    g(xi_i) */
} pbe_kernel_spec;

/* TypeName -> enum, as the reference's run-time selection tables do.  Unknown names return
 * PBE_E_UNKNOWN_TYPE and pbe_last_error() lists the valid ones (PBEMethod.C:97-105). */
int pbe_coalescence_type(const char *type_name, int *out);
int pbe_breakup_type(const char *type_name, int *out);
int pbe_daughter_type(const char *type_name, int *out);

const char *pbe_last_error(void);
int pbe_version(void);

/* ---- context: one per (process, device).  Owns streams, staging buffers, workspaces. ---
 * Every handle created on a context (pbe_moc, pbe_qmom, pbe_mom, pbe_mesh, pbe_*_state) keeps a pointer to it:
 * destroy the handles BEFORE the context, as the reference's members are destroyed before the object that owns
 * them (PBEMethod.H:69-73).  A handle destroyed after its context reads freed memory. */
typedef struct pbe_ctx pbe_ctx;
int pbe_ctx_create(int device, pbe_ctx **out);
int pbe_ctx_destroy(pbe_ctx *ctx);
/* name / SM count / bytes of HBM of the context's device */
int pbe_ctx_info(pbe_ctx *ctx, char *name, int name_len, int *sm_count, int64_t *hbm_bytes);
/* count of this library's kernel launches since the context was created (bench `gpu_launches`) */
int64_t pbe_ctx_launch_count(pbe_ctx *ctx);
/* Roofline denominator not in MEASURED_PEAKS.json: a register-resident DFMA-chain kernel,
 * timed with CUDA events.  Returns TFLOP/s (2 flops per DFMA). */
int pbe_measure_fp64_peak(pbe_ctx *ctx, double *tflops, double *ms);

/* ======================================================================================= */
/* Method of classes -- replaces MOC::correct()'s source path, src/PBESystem/MOC.C:87-184   */
/* ======================================================================================= */
typedef struct pbe_moc pbe_moc;

/* MOCCoeffs{numberOfClasses, xi1, sourceRelaxation} (MOC.C:49-83): xi_i = xi1 (i+1). */
int pbe_moc_create(pbe_ctx *ctx, int N, double xi1, const pbe_kernel_spec *spec,
                   double source_relaxation, pbe_moc **out);
int pbe_moc_destroy(pbe_moc *moc);
/* Two device kernels implement the same source path: the shared-memory kernel (any N the
 * shared memory holds) and the tensor-memory kernel (accumulators in TMEM, 256 cells per CTA;
 * N <= 118).  pbe_moc_create picks by occupancy; these calls let tests and benches pin one.
 * Environment variable PBE_MOC_KERNEL=smem|tmem overrides the automatic choice at create. */
enum { PBE_MOC_KERNEL_SMEM = 1, PBE_MOC_KERNEL_TMEM = 2, PBE_MOC_KERNEL_PIVOT = 3 /* fixed-pivot grids only */ };
int pbe_moc_set_kernel(pbe_moc *moc, int which); /* PBE_E_UNSUPPORTED if not available */
int pbe_moc_get_kernel(pbe_moc *moc, int *which, int *cells_per_cta, int *ctas_per_sm);

/* Optional fused outputs of one MOC launch; every pointer is a DEVICE pointer or NULL. */
typedef struct pbe_moc_io {
    const double *alpha;  /* [ncells] phase fraction: continuity check (MOC.C:254-279) */
    const double *vcell;  /* [ncells] cell volumes (continuity check, global moments) */
    double *S;            /* [N][stride] limited source max(C+B, -relax n/dt)  (MOC.C:87-91) */
    double *d;            /* [ncells] number-mean diameter of the (updated) n   (MOC.C:282-290) */
    double *moments;      /* [n_moments] += sum_c V_c sum_i xi_i^k n_i[c] (partial: this rank) */
    int n_moments;        /* <= 8 */
    double *cont_err_max; /* [1] max_c | sum_k xi_k n_k / V - alpha |  (MOC.C:276-279); caller
                             zeroes it */
} pbe_moc_io;

/* S = limited source of the CURRENT n (read-only).  a1-a4 of SURVEY.md section 8. */
int pbe_moc_source(pbe_moc *moc, int64_t ncells, int64_t stride, const double *n,
                   const double *eps, const double *rho, const double *nu, double dt, double *S,
                   void *cuda_stream);
/* One zero-flux explicit step in place: n_k += dt S_k (MOC.C:259-272 with phi = 0), plus the
 * fused outputs requested in io (may be NULL). */
int pbe_moc_step(pbe_moc *moc, int64_t ncells, int64_t stride, double *n, const double *eps,
                 const double *rho, const double *nu, double dt, const pbe_moc_io *io,
                 void *cuda_stream);
/* d[c] = sum_i (6 xi_i/pi)^(1/3) n_i / sum_i n_i   (MOC.C:282-290, 43-45) */
int pbe_moc_d(pbe_moc *moc, int64_t ncells, int64_t stride, const double *n, double *d,
              void *cuda_stream);
/* per-class global min / max over cells (MOC.C:181-183, 238): out_min/out_max DEVICE [N] */
int pbe_moc_class_minmax(pbe_moc *moc, int64_t ncells, int64_t stride, const double *n,
                         double *out_min, double *out_max, void *cuda_stream);
/* The same result without any allocation (pbe_moc_class_minmax takes its per-CTA partials from
 * cudaMallocAsync).  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The partials live in ONE file-scope device array of PBE_NOALLOC_MAX_CLASSES x PBE_NOALLOC_MAX_PARTS x 2
 * doubles (1 MiB, held by every GPU that loads the library until exit), whose device address is resolved
 * once per device when a MOC handle is created on it.  numberOfClasses > PBE_NOALLOC_MAX_CLASSES returns
 * PBE_E_UNSUPPORTED; ncells is not limited (at most PBE_NOALLOC_MAX_PARTS partials per class).  Every call
 * on a device shares the array: calls that overlap on the GPU (different streams, no dependency) RACE --
 * order them.  n, out_min, out_max: DEVICE pointers as above; calls no cudaMalloc / cudaFree. */
#define PBE_NOALLOC_MAX_CLASSES 128
#define PBE_NOALLOC_MAX_PARTS 512
int pbe_moc_class_minmax_noalloc(pbe_moc *moc, int64_t ncells, int64_t stride, const double *n,
                                 double *out_min, double *out_max, void *cuda_stream);

/* HOST-buffer entry point: what the OpenFOAM adapter calls from MOC::correct().  Copies the
 * fields host->device in cell chunks, runs pbe_moc_source per chunk, copies S back; copies and
 * kernels of successive chunks overlap on the context's streams.  Synchronous on return.
 * Pinned host memory (pbe_host_alloc) gives full PCIe rate; pageable memory works. */
int pbe_moc_source_host(pbe_moc *moc, int64_t ncells, int64_t stride, const double *n_host,
                        const double *eps_host, const double *rho_host, const double *nu_host,
                        double dt, double *S_host);

/* ======================================================================================= */
/* QMOM -- replaces QMOM::correct()'s source path, src/PBESystem/QMOM.C:85-129, 215-267,    */
/* and wheeler_inversion(), src/PBESystem/MomentInversion.C:9-51                            */
/* ======================================================================================= */
typedef struct pbe_qmom pbe_qmom;
#define PBE_QMOM_MAX_ORDER 4 /* QMOMCoeffs.quadratureOrder 1..4 (tests/integration/QMOM/setupCases.py:92) */

int pbe_qmom_create(pbe_ctx *ctx, int order, const pbe_kernel_spec *spec, pbe_qmom **out);
int pbe_qmom_destroy(pbe_qmom *q);
int pbe_qmom_order(pbe_qmom *q, int *order, pbe_ctx **ctx); /* QMOMCoeffs.quadratureOrder */

/* S[k][c] for k < 2*order, NOT NaN-scrubbed (as momentSourceTerm returns it, QMOM.C:215-267).
 * The inversion is done once per cell and reused for every k. */
int pbe_qmom_source(pbe_qmom *q, int64_t ncells, int64_t stride, const double *m,
                    const double *eps, const double *rho, const double *nu, double *S,
                    void *cuda_stream);
/* NaN->0 scrub (QMOM.C:101-107), m_k += dt S_k (QMOM.C:193-197 with phi = 0), clamp to SMALL =
 * 1e-15 (QMOM.C:117-123), d = (6/pi m1/m0)^(1/3) (QMOM.C:125).  S (scrubbed) and d optional. */
int pbe_qmom_step(pbe_qmom *q, int64_t ncells, int64_t stride, double *m, const double *eps,
                  const double *rho, const double *nu, double dt, double *S, double *d,
                  void *cuda_stream);
/* batched wheeler_inversion: m [2*order][stride] -> x, w [order][stride]; x ascending. */
int pbe_wheeler_batch(pbe_ctx *ctx, int order, int64_t ncells, int64_t stride, const double *m,
                      double *x, double *w, void *cuda_stream);
int pbe_qmom_source_host(pbe_qmom *q, int64_t ncells, int64_t stride, const double *m_host,
                         const double *eps_host, const double *rho_host, const double *nu_host,
                         double *S_host);

/* ======================================================================================= */
/* methodOfMoments (first slice) -- replaces the per-cell part of          */
/* MOM::correct(), src/PBESystem/MOM.C:148-236, and its nested-integral source terms,       */
/* MOM.C:259-311 (coalescence) and MOM.C:313-362 (breakup).  Three moments m0, m1, m2;       */
/* presumed gamma density with k = m1^2 / (m2 m0 - m1^2 + SMALL), theta = (m0 m2 - m1^2) /   */
/* max(m0 m1, SMALL) (MOM.C:151-161).  integrate() (Integrator.H:12-27, GSL QAGIU) becomes a */
/* fixed-depth 15-point Gauss-Kronrod rule on the same x = lower + (1-t)/t transform with    */
/* `panels` equal panels of (0,1] (0 = default), the end ones graded towards x = 0 and       */
/* x = inf (the macros below).  The implicit fvm::ddt + fvm::div solve of                    */
/* MOM.C:211-222 is out of scope (zero-flux update only).  Parity unpinned in the reference. */
/* ======================================================================================= */
typedef struct pbe_mom pbe_mom;
#define PBE_MOM_DEFAULT_PANELS 12
#define PBE_MOM_MAX_PANELS 28
#define PBE_MOM_GRADED_LEVELS 15 /* bisections of the panel next to the lower limit (x -> 0) */
#define PBE_MOM_TAIL_LEVELS 5    /* bisections of the panel next to t = 0 (x -> inf) */
#define PBE_MOM_ORIGIN_POWER 8   /* u = a s^8 on the panel that touches x = 0 (density ~ x^(k-1)) */
#define PBE_MOM_EXTRA_PANELS (PBE_MOM_GRADED_LEVELS + PBE_MOM_TAIL_LEVELS) /* nodes = 15 (panels + this) */
/* The rule is FIXED where QAGIU adapts.  With the default 12 panels it gives back m0 and m1 of a gamma
 * density to better than 1e-6 for mean volumes m1/m0 between 1e-5 and 30 and shapes k between 0.3 and
 * 20 (the reference's own 0-D cases have mean volumes of O(1): tests/integration/MOC/setupCases.py:41-54;
 * map and closed-form checks in tests/test_mom.py).
 * Every call checks, per cell, that the rule gives back the density's own m0 and m1; cells off by more
 * than this relative tolerance are counted (pbe_mom_coverage) instead of being integrated silently
 * wrong -- rescale the volume unit of such a case (the kernels are homogeneous in xi). */
#define PBE_MOM_COVERAGE_TOL 1.0e-6

int pbe_mom_create(pbe_ctx *ctx, int panels, const pbe_kernel_spec *spec, pbe_mom **out);
int pbe_mom_destroy(pbe_mom *h);
/* uncovered cells since the last reset and the worst relative deviation seen (see
 * PBE_MOM_COVERAGE_TOL); synchronises the device; reset != 0 clears the counters */
int pbe_mom_coverage(pbe_mom *h, int64_t *uncovered_cells, double *worst_deviation, int reset);
/* S[k][c] = coalescenceSourceTerm(k) + breakupSourceTerm(k), k < 3, NOT NaN-scrubbed
 * (MOM::momentSourceTerm, MOM.C:244-254).  m, S: [3][stride] device arrays.  gamma (optional,
 * [3][stride]) receives gamma_c0, gamma_k, gamma_theta (MOM.C:151-161).  Cells whose moments give
 * k <= 0 or theta <= 0 (Boost.Math raises a domain error there) get NaN sources. */
int pbe_mom_source(pbe_mom *h, int64_t ncells, int64_t stride, const double *m, const double *eps,
                   const double *rho, const double *nu, double *S, double *gamma,
                   void *cuda_stream);
/* NaN->0 scrub (MOM.C:198-207), m_k += dt S_k (MOM.C:211-222 with phi = 0), clamp to SMALL =
 * 1e-15 (MOM.C:224-231), d = (6/pi m1/m0)^(1/3) (MOM.C:233).  S (scrubbed), d, gamma optional. */
int pbe_mom_step(pbe_mom *h, int64_t ncells, int64_t stride, double *m, const double *eps,
                 const double *rho, const double *nu, double dt, double *S, double *d, double *gamma,
                 void *cuda_stream);

/* ======================================================================================= */
/* Transport of the class / moment fields (first slice) --       */
/* replaces the explicit update of MOC::solveWithMULES, src/PBESystem/MOC.C:198-272, and     */
/* QMOM::solveWithMULES, src/PBESystem/QMOM.C:131-197, for the divergence scheme             */
/* `Gauss upwind` (the MULES-limited high-order correction is then identically zero):        */
/*   psi_k <- (psi_k rDeltaT + S_k - (1/V) sum_f phi_f psi_k,f) / rDeltaT                     */
/* with OpenFOAM's upwind face value and MULES::explicitSolve (rho = 1, Sp = 0).             */
/* ======================================================================================= */
typedef struct pbe_mesh pbe_mesh;
/* OpenFOAM face addressing (fvMesh::owner() / neighbour(), boundary patches' faceCells()):
 * internal faces 0 .. n_internal_faces-1, then the boundary faces of all patches in patch
 * order, each with its owner cell.  HOST arrays, copied; vcell = mesh.V() (MOC.C:276).
 * Builds the device-side cell -> faces table once per mesh. */
int pbe_mesh_create(pbe_ctx *ctx, int64_t ncells, int64_t n_internal_faces, const int32_t *owner,
                    const int32_t *neighbour, int64_t n_boundary_faces,
                    const int32_t *boundary_owner, const double *vcell, pbe_mesh **out);
int pbe_mesh_destroy(pbe_mesh *mesh);
int pbe_mesh_info(pbe_mesh *mesh, int64_t *ncells, int64_t *n_internal_faces,
                  int64_t *n_boundary_faces, int *max_faces_per_cell);
/* How the device tables split the cells: "regular" cells (exactly *faces_per_regular_cell faces,
 * internal or on a physical patch -- the mesh's most common count among 4, 5, 6 -- and none on a
 * processor patch) go through the register-resident kernel, all others through the general one.
 * pbe_mesh_set_coupled re-splits: cells on processor patches need halo values and leave the set. */
int pbe_mesh_regular_cells(pbe_mesh *mesh, int *faces_per_regular_cell, int64_t *n_regular);

/* fvSchemes divSchemes entry "div(phi,n.*)" -> enum ("Gauss upwind" | "upwind" | "Gauss vanLeer" |
 * "vanLeer" | "Gauss limitedLinear 1" | "limitedLinear 1"); anything else returns PBE_E_UNKNOWN_TYPE with the valid names listed (the
 * reference's FatalError convention).  The reference's validation case selects Gauss vanLeer
 * (tests/validation/simmons_and_azzopardi/c01/system/fvSchemes:34). */
enum { PBE_DIV_UPWIND = 0, PBE_DIV_VANLEER = 1,
       PBE_DIV_LIMITEDLINEAR = 2 /* "Gauss limitedLinear 1": the implicit term only (pbe_transport_implicit), the
                                    scheme of "div(U,nk)" in the reference's MOC cases (tests/integration/MOC/template/
                                    system/fvSchemes:44) */ };
int pbe_div_scheme_type(const char *scheme_name, int *out);

/* Mesh geometry the limited schemes need (fvMesh::Sf(), C(), weights()): HOST arrays, copied.
 * Sf [n_internal + n_boundary][3] face area vectors (owner -> neighbour / outward), C [ncells][3]
 * cell centres, weights [n_internal] linear interpolation weights. */
int pbe_mesh_set_geometry(pbe_mesh *mesh, const double *Sf, const double *C, const double *weights);

/* The MULES-limited update of MOC::solveWithMULES (MOC.C:198-272) / QMOM::solveWithMULES
 * (QMOM.C:131-197) on nfields DEVICE fields psi [nfields][stride]:
 *   phiCorr_k = fvc::flux(phi, psi_k, scheme) - upwind flux, limited by MULES::limit against
 *   [psi_min, psi_max[k]] with n_limiter_iter sweeps (OpenFOAM's nLimiterIter, default 3), then
 *   MULES::limitSum over the fields, then psi_k <- (psi_k rDeltaT + S_k - div(phiCorr_k + upwind)) / rDeltaT.
 * psi_max: DEVICE [nfields] (pbe_moc_psi_max / pbe_qmom_psi_max); psi_b, S as pbe_transport_explicit.
 * With PBE_DIV_UPWIND the correction is identically zero and the result equals
 * pbe_transport_explicit's.  Uncoupled patches only (no processor patches yet: PBE_E_UNSUPPORTED).
 * OpenFOAM's MULES is third-party code: parity unpinned. */
int pbe_transport_mules(pbe_mesh *mesh, int scheme, int nfields, int64_t stride, const double *psi,
                        const double *phi, const double *psi_b, int64_t stride_b, const double *S,
                        const double *psi_max, double psi_min, int n_limiter_iter, double dt,
                        double *psi_out, void *cuda_stream);
/* psi_max[k] = max_c(alpha V / xi_k) (MOC.C:238), DEVICE in / out */
int pbe_moc_psi_max(pbe_moc *moc, int64_t ncells, const double *alpha, const double *vcell,
                    double *psi_max, void *cuda_stream);
/* psi_max[k] = 1.2 max(m_k) over cells and patch values (QMOM.C:171), DEVICE in / out */
int pbe_qmom_psi_max(pbe_ctx *ctx, int nmoments, int64_t ncells, int64_t stride, const double *m,
                     int64_t n_boundary_faces, int64_t stride_b, const double *psi_b,
                     double *psi_max, void *cuda_stream);

/* One explicit transport + source update of nfields fields psi [nfields][stride] (DEVICE).
 * phi: DEVICE [n_internal_faces + n_boundary_faces] volumetric face flux of the phase
 * (phase_.phi(), MOC.C:200), positive from owner to neighbour / out of the domain.
 * psi_b: DEVICE [nfields][stride_b] boundary-face values (Nk.boundaryField()), or NULL for
 * zero-gradient on every boundary face.  S: DEVICE [nfields][stride] explicit source or NULL.
 * psi_out must not alias psi.  Per-cell sums run in ascending face order, as
 * fvc::surfaceIntegrate's loops do. */
int pbe_transport_explicit(pbe_mesh *mesh, int scheme, int nfields, int64_t stride,
                           const double *psi, const double *phi, const double *psi_b,
                           int64_t stride_b, const double *S, double dt, double *psi_out,
                           void *cuda_stream);

/* ---- the implicit path: MOC::solveWithFVM (MOC.C:186-198), QMOM::solveWithFVM (QMOM.C:201-209) -- the
 * reference's DEFAULT transport (usingMULES false, MOC.C:58, QMOM.C:71):
 *     fvScalarMatrix eqn(fvm::ddt(psi_k) + fvm::div(phi, psi_k) == S_k);  eqn.relax();  eqn.solve();
 * Euler implicit ddt (diag = rDeltaT V, source = rDeltaT psi0 V), Gauss convection with the scheme's weights
 * (lower = -w phi, upper = lower + phi, diag = -(sums), gaussConvectionScheme::fvmDiv); zero-gradient patches add
 * phi_b to the diagonal, fixed-value patches -phi_b psi_b to the source; fvMatrix::relax(factor) (diagonal dominance,
 * D / factor, source += (D - D0) psi0) when relaxation > 0.  One matrix serves all nfields fields for
 * PBE_DIV_UPWIND; the limited schemes (PBE_DIV_VANLEER, PBE_DIV_LIMITEDLINEAR; they need pbe_mesh_set_geometry) take
 * their per-field weights from the current field, as limitedScheme::weights does: one matrix per field, Jacobi only.  OpenFOAM solves field after field with the solver fvSolution names (PBiCG + DILU or
 * smoothSolver symGaussSeidel in the reference's cases -- sequential sweeps); here ALL fields are solved together by
 * Jacobi sweeps or a diagonally preconditioned BiCGStab whose vector updates and matrix-vector products run over
 * [nfields][ncells] at once, each field with its own scalars, stopped by OpenFOAM's own criterion: normalised residual
 * sum|b - Ax| / (sum(|Ax - xbar A1| + |b - xbar A1|) + 1e-20) < tolerance, or < rel_tol x initial, after >= min_iter
 * iterations (lduMatrix::solver::normFactor, solverPerformance::checkConvergence).  The result is the matrix's
 * solution to that tolerance -- as the reference's is: the two differ by what the tolerance allows (1e-5 in the
 * reference's fvSolution files); parity unpinned against OpenFOAM itself, checked against a direct sparse solve of the
 * restated matrix and against the reference's own MOC.C / QMOM.C driven through oracle/_ref. */
typedef struct pbe_fvm_controls {
    double tolerance;  /* fvSolution solvers { "n.*" { tolerance 1e-5; } }                                        */
    double rel_tol;    /* relTol (0 = off)                                                                        */
    int min_iter;      /* minIter (the reference's cases: 1)                                                      */
    int max_iter;      /* maxIter (OpenFOAM default 1000)                                                         */
    double relaxation; /* fvSolution relaxationFactors { equations { ".*" f; } }; <= 0: eqn.relax() does nothing  */
    int solver;        /* PBE_FVM_SOLVER_*: which iteration; AUTO picks by the matrix (below)                     */
} pbe_fvm_controls;
/* JACOBI: x += z, z <- -D^-1 R z (the residual is D z: no extra pass), one kernel and 4 vector passes per matrix-vector
 * product; BICGSTAB: diagonally preconditioned, ~9.5 passes per product.  Upwind convection makes the matrix triangular
 * along the flow, so either needs about as many products as the solution travels cells in dt; Jacobi was the faster of
 * the two at every Courant number measured (0.4 .. 150): AUTO = JACOBI. */
enum { PBE_FVM_SOLVER_AUTO = 0, PBE_FVM_SOLVER_BICGSTAB = 1, PBE_FVM_SOLVER_JACOBI = 2 };
/* OpenFOAM's SolverPerformance line per field: "Solving for n7, Initial residual = .., Final residual = .., No Iterations .." */
typedef struct pbe_fvm_perf {
    double initial_residual, final_residual;
    int iterations;
    int state; /* 1 converged, 2 breakdown (stagnated short of the tolerance), 3 max_iter reached */
} pbe_fvm_perf;
/* psi, S, psi_out DEVICE [nfields][stride] (psi_out must not alias psi); phi DEVICE [internal + boundary faces];
 * psi_b DEVICE [nfields][stride_b] fixed patch values, or NULL = zero-gradient on every boundary face; controls NULL =
 * {1e-5, 0, 1, 1000, 0, AUTO}; perf HOST [nfields] or NULL.  Synchronous on return (the iteration count is data dependent).
 * Processor patches: see pbe_mesh_set_halo_plan below (PBE_E_UNSUPPORTED without a plan).  The solver's vectors live on
 * the mesh handle: calls on one mesh must not overlap (use one mesh handle per concurrent solve). */
int pbe_transport_implicit(pbe_mesh *mesh, int scheme, int nfields, int64_t stride, const double *psi,
                           const double *phi, const double *psi_b, int64_t stride_b, const double *S, double dt,
                           const pbe_fvm_controls *controls, double *psi_out, pbe_fvm_perf *perf, void *cuda_stream);

/* The implicit solve on a DECOMPOSED mesh (one sub-domain per GPU, processor patches marked by pbe_mesh_set_coupled,
 * a communicator on the context: pbe_comm_init).  OpenFOAM treats a processor patch as a coupled patch: w phi_b on the
 * diagonal, (1 - w) phi_b times the value ACROSS the face as an off-diagonal coefficient that every solver iteration
 * refreshes through the interface update, and global sums (gSum / gAverage) in the residual.  Here a coupled face is
 * one more column of the cell's row, read from a halo row; before every Jacobi sweep the rank's boundary values of z
 * are packed (pbe_gather_cells) and exchanged (ncclSend / ncclRecv in one group), and the per-field sums are
 * all-reduced before the fields' states are updated -- all on the caller's stream, so all ranks take the same decisions
 * and the host still looks at the states one sweep behind.  The plan: neighbour q = rank peer[q] shares the coupled
 * faces [off[q], off[q] + cnt[q]) (indices within the coupled range, same order on both sides as for
 * pbe_halo_exchange); send_cells HOST [ncoupled]: this rank's cell on each coupled face; global_ncells: cells of the
 * whole mesh (gAverage).  pbe_transport_implicit then solves the GLOBAL system (Gauss upwind, Jacobi); every rank must
 * call it with the same nfields and controls. */
int pbe_mesh_set_halo_plan(pbe_mesh *mesh, int nneigh, const int *peer, const int64_t *off, const int64_t *cnt,
                           const int32_t *send_cells, int64_t global_ncells);

/* HOST-buffer MOC::correct() with solveWithMULES (MOC.C:159-184, 198-272): what the OpenFOAM-side
 * adapter calls when the case's div(phi,n.*) scheme is on the device.  n_host [N][stride] is
 * updated in place (N = the method's numberOfClasses, ncells = the mesh's); phi_host
 * [internal + boundary faces]; psi_b_host [N][stride_b] patch values or NULL (zero-gradient);
 * d_host [ncells] receives MOC::d() of the new fields or is NULL.  Uploads overlap the source
 * kernel (cell chunks), downloads overlap the transport kernels (field chunks); the fields stay
 * on the device in between.  Synchronous on return. */
int pbe_moc_num_classes(pbe_moc *moc, int *N, pbe_ctx **ctx);
/* MOCCoeffs of the handle: xi1 (MOC.C:56) and sourceRelaxation (MOC.C:60); either may be NULL */
int pbe_moc_coeffs(pbe_moc *moc, double *xi1, double *source_relaxation);

/* ---- geometric / arbitrary class grid (EXTENSION, parity unpinned) ------------------
 * The reference's MOC only has xi_i = (i + 1) xi1 (MOC.C:56,79); its comment at MOC.C:147-148 names Kumar
 * & Ramkrishna (1996), whose fixed-pivot technique this is, in the reference's conventions: a pair (i, j)
 * with x_i + x_j <= x_{N-1} (death bound of MOC.C:107-109) dies in both classes and is born on the two
 * pivots that bracket x_i + x_j with the number- and volume-preserving weights; breakup daughters
 * (uniformBinaryBreakup) are assigned through the same hat functions (volume exact).  On the uniform
 * volumes the coalescence terms are the reference's own.  pivots [N], strictly increasing, positive.
 * The handle works with every pbe_moc_* entry, the resident state and its transport (continuity check, MULES bound
 * maxNk = max(alpha V) / x_k) included. */
int pbe_moc_create_pivots(pbe_ctx *ctx, int N, const double *pivots, const pbe_kernel_spec *spec,
                          double source_relaxation, pbe_moc **out);
/* class volumes of the handle's grid ([N], may be NULL) and whether it is the reference's uniform one */
int pbe_moc_grid(pbe_moc *moc, double *volumes, int *is_uniform);
/* the same N volumes as a DEVICE array owned by the handle (read by the transport path's continuity check and bound) */
int pbe_moc_grid_device(pbe_moc *moc, const double **volumes_dev);
int pbe_moc_correct_host(pbe_moc *moc, pbe_mesh *mesh, int scheme, int64_t stride, double *n_host,
                         const double *eps_host, const double *rho_host, const double *nu_host,
                         const double *phi_host, const double *psi_b_host, int64_t stride_b,
                         double dt, double *d_host);

/* QMOM::correct() with transport (QMOM.C:85-129): pbe_qmom_source -> pbe_field_scrub_nan(S)
 * (NaN -> 0, QMOM.C:101-107) -> pbe_transport_explicit on the 2*order moment fields
 * (QMOM.C:131-197) -> pbe_qmom_clamp_d: m_k = max(m_k, SMALL = 1e-15) (QMOM.C:117-123) and
 * d = (6/pi m1/m0)^(1/3) (QMOM.C:125; d may be NULL).  All pointers DEVICE. */
int pbe_field_scrub_nan(pbe_ctx *ctx, int64_t count, double *f, void *cuda_stream);
int pbe_qmom_clamp_d(pbe_ctx *ctx, int nmoments, int64_t ncells, int64_t stride, double *m, double *d,
                     void *cuda_stream);

/* Decomposed meshes (one sub-domain per GPU; cells block-partitioned as on the source path).
 * Boundary faces [first, first + count) of this sub-domain are processor-patch faces; the face
 * value there is OpenFOAM's coupled interpolation pos(phi_b) psi_own + (1 - pos(phi_b)) psi_across,
 * with psi_across delivered per step in psi_halo [nfields][stride_h] (DEVICE; entry j = the cell
 * across coupled face first + j).  pbe_gather_cells packs the values the neighbouring ranks need
 * (out[k][j] = psi[k][cells[j]], cells DEVICE int32); the exchange itself is the caller's
 * (NCCL send / recv over NVLink in openpbe_b200/sharding.py). */
int pbe_mesh_set_coupled(pbe_mesh *mesh, int64_t first_boundary_face, int64_t count);
int pbe_gather_cells(pbe_ctx *ctx, int nfields, int64_t stride, const double *psi, int64_t n,
                     const int32_t *cells, double *out, int64_t stride_out, void *cuda_stream);
int pbe_transport_explicit_halo(pbe_mesh *mesh, int scheme, int nfields, int64_t stride,
                                const double *psi, const double *phi, const double *psi_b,
                                int64_t stride_b, const double *psi_halo, int64_t stride_h,
                                const double *S, double dt, double *psi_out, void *cuda_stream);
/* The same update restricted to one of the two cell sets of pbe_mesh_regular_cells, so that the
 * halo exchange overlaps the bulk of the work: launch PBE_CELLS_REGULAR (they read
 * no halo value, psi_halo may be NULL) while the exchange is in flight, then PBE_CELLS_IRREGULAR
 * (boundary and processor-patch cells) once psi_halo has arrived.  Both together = PBE_CELLS_ALL. */
enum { PBE_CELLS_ALL = 0, PBE_CELLS_REGULAR = 1, PBE_CELLS_IRREGULAR = 2 };
int pbe_transport_explicit_part(pbe_mesh *mesh, int scheme, int which, int nfields, int64_t stride,
                                const double *psi, const double *phi, const double *psi_b,
                                int64_t stride_b, const double *psi_halo, int64_t stride_h,
                                const double *S, double dt, double *psi_out, void *cuda_stream);

/* ======================================================================================= */
/* Resident state -- correct() the way the reference's solver sees it.  The reference's     */
/* methods OWN their class / moment fields (PtrList<volScalarField> classNumberDensity_,     */
/* src/PBESystem/MOC.H:66, read MUST_READ at construction MOC.C:69-77, written by            */
/* runTime.write(); QMOM.H:62, QMOM.C:72-79); once per time step the solver calls            */
/* PBEDiameter::correct() { method_->correct(); d_ = method_->d(); } (PBEDiameter.H:95-99,   */
/* applications/PBEFoam/PBEFoam.C:56-67) and reads only d.  A pbe_*_state is that owned     */
/* state on the device: per step only this step's phase fields go in (epsilon, rho, nu,      */
/* alpha: 32 B/cell, plus the face flux when a mesh is attached) and d comes out (8 B/cell); */
/* the class / moment fields cross the bus at construction and at write time only.           */
/* ======================================================================================= */
typedef struct pbe_moc_state pbe_moc_state;

/* vcell_host [ncells] = mesh.V() (continuity check, global moments) or NULL.  The state
 * allocates n [N][ncells] and the per-cell rows on the method's device, plus one event pool. */
int pbe_moc_state_create(pbe_moc *moc, int64_t ncells, const double *vcell_host,
                         pbe_moc_state **out);
int pbe_moc_state_destroy(pbe_moc_state *st);
/* construction: the MUST_READ fields n0..n{N-1} (MOC.C:69-77), host [N][stride] */
int pbe_moc_state_upload(pbe_moc_state *st, int64_t stride, const double *n_host);
/* runTime.write(): the AUTO_WRITE fields back to the host, [N][stride] */
int pbe_moc_state_download(pbe_moc_state *st, int64_t stride, double *n_host);
/* the same for the cell range [cell0, cell0 + count) only: n_host [N][stride] holds those cells
 * (construction / write-out in pieces when the host cannot hold all class fields at once) */
int pbe_moc_state_upload_range(pbe_moc_state *st, int64_t cell0, int64_t count, int64_t stride,
                               const double *n_host);
int pbe_moc_state_download_range(pbe_moc_state *st, int64_t cell0, int64_t count, int64_t stride,
                                 double *n_host);
/* MOC::d() of the CURRENT fields (PBEDiameter's constructor, PBEDiameter.C:64-75) -> host */
int pbe_moc_state_d(pbe_moc_state *st, double *d_host);
/* device views for callers that chain their own device work: n [N][*stride], d [ncells] */
int pbe_moc_state_device(pbe_moc_state *st, double **n_dev, int64_t *stride, double **d_dev);
/* MOC::solveWithMULES (MOC.C:198-272) instead of the zero-flux update: the mesh's face
 * addressing and the div(phi,n.*) scheme; correct_resident then wants phi_host. */
int pbe_moc_state_attach_mesh(pbe_moc_state *st, pbe_mesh *mesh, int scheme);
/* usingMULES false -- the reference's default (MOC.C:58, 177-180; QMOM.C:71, 112-115): correct_resident then runs
 * solveWithFVM (pbe_transport_implicit with these controls) instead of solveWithMULES.  controls NULL switches back.
 * Needs an attached mesh (its scheme is then the implicit term's).  solver_performance: what OpenFOAM prints per field for the LAST step,
 * HOST perf [N] ([2 * order] for QMOM). */
int pbe_moc_state_set_implicit(pbe_moc_state *st, const pbe_fvm_controls *controls);
int pbe_moc_state_solver_performance(pbe_moc_state *st, pbe_fvm_perf *perf);

/* What MOC::correct() prints every step (MOC.C:181-183, 276-279) plus configs[3]'s global
 * moments.  HOST memory; the caller all-reduces across ranks (pbe_comm_*). */
typedef struct pbe_moc_diag {
    double *class_min;   /* [N] or NULL: min over this rank's cells of n_k        (MOC.C:181-183) */
    double *class_max;   /* [N] or NULL                                           (MOC.C:238)     */
    double cont_err_max; /* out: max_c |sum_k xi_k n_k / V - alpha| (needs alpha) (MOC.C:276-279) */
    int n_moments;       /* in: 0..8 */
    double moments[8];   /* out: sum_c V_c sum_i xi_i^k n_i[c] over this rank's cells (needs vcell) */
} pbe_moc_diag;

/* One MOC::correct() + d(): HOST eps, rho, nu, alpha [ncells] in, HOST d [ncells] out.
 * rho_host / nu_host may be NULL after the first call: the values of the previous call are kept
 * (constant-property phases).  alpha_host NULL: no continuity check.  phi_host [internal +
 * boundary faces] and psi_b_host [N][stride_b] (NULL = zero-gradient patches) only with an
 * attached mesh.  d_host NULL leaves d on the device.  Cells are processed in chunks: the
 * uploads of chunk k+1 and the download of d for chunk k-1 overlap the kernel of chunk k.
 * Pinned host memory (pbe_host_alloc / pbe_host_register) gives full PCIe rate and true overlap.
 * Synchronous on return. */
int pbe_moc_correct_resident(pbe_moc_state *st, const double *eps_host, const double *rho_host,
                             const double *nu_host, const double *alpha_host,
                             const double *phi_host, const double *psi_b_host, int64_t stride_b,
                             double dt, double *d_host, pbe_moc_diag *diag);

typedef struct pbe_qmom_state pbe_qmom_state;
int pbe_qmom_state_create(pbe_qmom *q, int64_t ncells, const double *vcell_host,
                          pbe_qmom_state **out);
int pbe_qmom_state_destroy(pbe_qmom_state *st);
int pbe_qmom_state_upload(pbe_qmom_state *st, int64_t stride, const double *m_host);  /* QMOM.C:72-79 */
int pbe_qmom_state_download(pbe_qmom_state *st, int64_t stride, double *m_host);
int pbe_qmom_state_d(pbe_qmom_state *st, double *d_host); /* d = (6/pi m1/m0)^(1/3), QMOM.C:81 */
int pbe_qmom_state_device(pbe_qmom_state *st, double **m_dev, int64_t *stride, double **d_dev,
                          double **S_dev);
int pbe_qmom_state_attach_mesh(pbe_qmom_state *st, pbe_mesh *mesh, int scheme);
int pbe_qmom_state_set_implicit(pbe_qmom_state *st, const pbe_fvm_controls *controls);
int pbe_qmom_state_solver_performance(pbe_qmom_state *st, pbe_fvm_perf *perf);

/* printAvgMaxMin (Utility.C:4-10) of what QMOM::correct() prints each step: per moment source
 * (QMOM.C:109), per moment (QMOM.C:122) and d (QMOM.C:128).  Each entry is {sum f V, sum V, max,
 * min} over this rank's cells (V = 1 without vcell); avg = [0] / [1] after the cross-rank
 * reduction.  Pointers are HOST arrays or NULL (not wanted). */
typedef struct pbe_qmom_diag {
    double *source; /* [2*order][4] */
    double *moment; /* [2*order][4] */
    double *d;      /* [4] */
} pbe_qmom_diag;

/* One QMOM::correct() (QMOM.C:85-129): sources, NaN scrub, update (zero flux, or transport with
 * an attached mesh), clamp to SMALL, d.  Argument conventions as pbe_moc_correct_resident. */
int pbe_qmom_correct_resident(pbe_qmom_state *st, const double *eps_host, const double *rho_host,
                              const double *nu_host, const double *phi_host,
                              const double *psi_b_host, int64_t stride_b, double dt,
                              double *d_host, pbe_qmom_diag *diag);

/* ======================================================================================= */
/* Scalar kernel functors evaluated ON THE DEVICE (batched) -- the reference's Foam-free     */
/* *Impl::S surface (coalescenceKernels/CoulaloglouTavlarides.H:44-63, breakupKernels/       */
/* CoulaloglouTavlarides.H:46-60; exported to Python by bindings/kernels.cpp:9-40).          */
/* All arrays are HOST pointers of length count; results come back in out[].                 */
/* ======================================================================================= */
int pbe_eval_coalescence(pbe_ctx *ctx, const pbe_kernel_spec *spec, int64_t count,
                         const double *xi1, const double *xi2, const double *eps,
                         const double *rho, const double *nu, double *out);
int pbe_eval_breakup(pbe_ctx *ctx, const pbe_kernel_spec *spec, int64_t count, const double *xi,
                     const double *rho, const double *eps, double *out);
/* beta(xi1, xi2) and moment(xi, k) of the daughter distribution */
int pbe_eval_daughter_beta(pbe_ctx *ctx, int daughter, int64_t count, const double *xi1,
                           const double *xi2, double *out);
int pbe_eval_daughter_moment(pbe_ctx *ctx, int daughter, int64_t count, const double *xi,
                             const int *k, double *out);
/* integrate(f, lower) of Integrator.H:12-27 for the integrand family the reference's test
 * pins (tests/unit_tests/integration.cpp:6-13): f(x) = exp(-rate x).  Device 15-point
 * Gauss-Kronrod on the QAGIU transform with fixed-depth bisection, one thread per integral. */
int pbe_eval_integrate_exp(pbe_ctx *ctx, int64_t count, const double *rate, const double *lower,
                           double *out);

/* ---- global diagnostics, Utility.C:4-10 printAvgMaxMin: out[0] = sum(f V), out[1] = sum V,
 * out[2] = max f, out[3] = min f over this rank's cells (DEVICE out[4]); the caller
 * all-reduces (sum, sum, max, min) across ranks: pbe_reduce_diagnostics. */
int pbe_field_avg_max_min(pbe_ctx *ctx, int64_t ncells, const double *f, const double *vcell,
                          double *out4, void *cuda_stream);

/* ======================================================================================= */
/* Multi-GPU: one process per GPU, cells block-partitioned (they are independent on the      */
/* source path: no data-path collective).  The library owns an NCCL communicator over the   */
/* ranks and carries what the reference gets from OpenFOAM's implicit MPI all-reduces        */
/* (SURVEY.md 2.3): min(n_k) MOC.C:181-183, max(alpha V / xi_k) MOC.C:238, max continuity      */
/* error MOC.C:276-279, max(m_k) QMOM.C:171, printAvgMaxMin Utility.C:4-10 -- and the halo     */
/* exchange of the transport step.  NCCL is bound at run time (dlopen libnccl.so.2, or        */
/* $PBE_NCCL_LIB): a single-GPU host needs none.  With a communicator on the context,          */
/* pbe_moc_correct_resident / pbe_qmom_correct_resident return GLOBAL diagnostics.             */
/* ======================================================================================= */
#define PBE_COMM_UNIQUE_ID_BYTES 128
/* rank 0 creates the id (ncclGetUniqueId); the host hands it to the other ranks (MPI_Bcast, a file) */
int pbe_comm_unique_id(void *id128);
/* ncclCommInitRank on the context's device; collective over the nranks processes */
int pbe_comm_init(pbe_ctx *ctx, int nranks, int rank, const void *id128);
/* adopt a communicator the host already has (an ncclComm_t; not destroyed with the context) */
int pbe_comm_init_from(pbe_ctx *ctx, void *nccl_comm);
int pbe_comm_destroy(pbe_ctx *ctx);
int pbe_comm_info(pbe_ctx *ctx, int *nranks, int *rank); /* 1, 0 without a communicator */
/* in-place all-reduces of DEVICE vectors (sum / min / max) as one NCCL group on the stream */
int pbe_reduce_diagnostics(pbe_ctx *ctx, int nsum, double *sums, int nmin, double *mins, int nmax,
                           double *maxs, void *cuda_stream);
/* the same for HOST vectors, synchronous (staged on the communicator's side stream) */
int pbe_reduce_diagnostics_host(pbe_ctx *ctx, int nsum, double *sums, int nmin, double *mins,
                                int nmax, double *maxs);
/* Halo exchange of a decomposed transport step: send [nfields][stride_s] = this rank's cell values
 * along its processor-patch faces (pbe_gather_cells), recv [nfields][stride_r] = the values across
 * them (psi_halo of pbe_transport_explicit_halo); neighbour q = rank peer[q] shares coupled faces
 * [off[q], off[q] + cnt[q]) (same order on both sides).  One NCCL group of sends / receives. */
int pbe_halo_exchange(pbe_ctx *ctx, int nneigh, const int *peer, const int64_t *off,
                      const int64_t *cnt, int nfields, const double *send, int64_t stride_s,
                      double *recv, int64_t stride_r, void *cuda_stream);

/* ---- memory helpers so a C/C++ host needs no CUDA headers ------------------------------ */
int pbe_device_alloc(pbe_ctx *ctx, int64_t bytes, void **out);
int pbe_device_free(pbe_ctx *ctx, void *p);
int pbe_host_alloc(pbe_ctx *ctx, int64_t bytes, void **out); /* pinned */
int pbe_host_free(pbe_ctx *ctx, void *p);
/* page-lock memory the caller already owns (an OpenFOAM field's storage) so that the *_resident
 * and *_host entries copy it at full PCIe rate and asynchronously */
int pbe_host_register(pbe_ctx *ctx, void *p, int64_t bytes);
int pbe_host_unregister(pbe_ctx *ctx, void *p);
int pbe_memcpy_h2d(pbe_ctx *ctx, void *dst, const void *src, int64_t bytes);
int pbe_memcpy_d2h(pbe_ctx *ctx, void *dst, const void *src, int64_t bytes);
int pbe_memset(pbe_ctx *ctx, void *dst, int value, int64_t bytes);
int pbe_sync(pbe_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif /* PBE_B200_H */
