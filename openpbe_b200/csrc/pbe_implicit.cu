// pbe_implicit.cu -- the implicit transport of the class / moment fields: what MOC::solveWithFVM
// (src/PBESystem/MOC.C:186-198) and QMOM::solveWithFVM (src/PBESystem/QMOM.C:201-209) do, the reference's DEFAULT
// branch (usingMULES false: MOC.C:58, QMOM.C:71):
//
//     fvScalarMatrix eqn(fvm::ddt(psi_k) + fvm::div(phi, psi_k) == S_k);   eqn.relax();   eqn.solve();
//
// OpenFOAM's fvMatrix / lduMatrix solvers are third-party code absent from the reference tree: PARITY UNPINNED.  The
// matrix is assembled as OpenFOAM 3.0.x assembles it (EulerDdtScheme::fvmDdt, gaussConvectionScheme::fvmDiv,
// fvMatrix::relax) -- the restatement the test infrastructure drives the reference's own sources through -- and solved to
// OpenFOAM's own stopping criterion, but by a solver made for this machine:
//
//   * the reference solves field after field (100 classes = 100 matrices assembled and solved one by one) with PBiCG +
//     DILU or symGaussSeidel sweeps: sequential recurrences along the cell numbering;
//   * for `Gauss upwind` ALL fields share ONE matrix (it depends on phi, V and dt only), so the fields are solved
//     together, every field with its own scalars, residual and stopping point.  One thread per cell keeps the cell's
//     row (neighbour columns and coefficients) in registers and sweeps a chunk of fields: the row is read once per
//     chunk, the vectors stream through HBM.  Two iterations: JACOBI in correction form (x += z, z <- -D^-1 R z; D z is
//     the current residual, so the stopping test costs nothing; the matrix is an M-matrix, column-dominant by V / dt:
//     the residual's L1 norm contracts every sweep) at 32 B per cell-field and sweep, and a diagonally preconditioned
//     BiCGStab at ~150 B per iteration.  Upwind convection makes the matrix triangular along the flow, so either needs
//     about as many matrix-vector products as the solution travels cells in dt: Jacobi won at every Courant number
//     measured and is what AUTO takes.
//   * the limited schemes of the reference's cases (Gauss vanLeer, Gauss limitedLinear 1) take their weights from the
//     field being solved for: one matrix per field, rebuilt per sweep from phi and the field's face weights.
//   * reductions (L1 norms, dot products) are two-stage and ordered: per-CTA partials, then one CTA per field adds them
//     in a fixed order and updates the field's state on the device.  The host reads the states back one sweep BEHIND
//     what it has queued, so the device never waits for it; converged fields are frozen on the device.
//   * on a decomposed mesh the ranks solve ONE system: a processor-patch face is one more column of the row, read from
//     a halo row that is exchanged before every sweep (one NCCL message per neighbour), and the sums are all-reduced
//     before the states are updated -- all on the caller's stream.
#include <algorithm>
#include <cstdint>
#include <cstdlib>
#include <cstring>

#include "pbe_mesh.h"

using namespace pbe;

namespace {

// Where the next field's gathered values wait (sweep of the shared upwind matrix on one rank): 1 = in registers
// (6 LDG per cell-field), 0 = in a per-thread shared-memory column filled by cp.async (6 LDGSTS + 6 LDS).  The sweep is
// bound by the load-store path: 160^3 x 100 fields 3.74 ms per sweep with 0, 3.40 ms with 1 (50 fields per CTA).
#ifndef PBE_IMP_REGSTAGE
#define PBE_IMP_REGSTAGE 1
#endif

constexpr int kBlock = 256;
constexpr int kPW = 4;    // cells per thread of the pointwise kernels
constexpr int kNS = 3;    // sums per field and kernel, at most
constexpr int kMaxChunk = 128; // fields per CTA of the row-holding kernels, at most
constexpr double kNormSmall = 1.0e-20; // lduMatrix::small_

struct FieldSc { // one per field, on the device
    double rho, alpha, omega, beta, norm, res0, res, xbar;
    int iters, state, half, pad;
};

enum { PH_XSUM = 0, PH_INIT, PH_A, PH_B, PH_C, PH_D, PH_J };

struct ImpParams {
    int64_t nc, stride, stride_b, nint, nbnd;
    int W, nfields, chunk, G;
    const int32_t *other, *face; // [W][nc]
    const double *phi, *V;
    double *diag, *dinv, *a1, *drel; // [nc]
    const double *psi, *psi_b, *S;
    double *x, *r, *r0, *p, *v, *t;
    FieldSc *sc;
    double *partial; // [nfields][kNS][G]
    // decomposed mesh: boundary faces [coupled0, coupled0 + ncpl) are processor-patch faces; halo_x / halo_z
    // [nfields][ncpl]: psi0 / the current z across them; gsum [nfields][kNS]: the all-reduced sums (null: one rank)
    int64_t coupled0, ncpl, nc_global;
    const double *halo_x, *halo_z;
    double *gsum;
    const double *wface; // limited schemes: limitedScheme::weights of every internal face, [nfields][nint]
    double *dinvK;       // limited schemes: the fields' own inverse diagonals, [nfields][nc]
    const double *Sf, *Cc, *wlin; // mesh geometry (component-major Sf [3][faces], C [3][nc]; linear weights [nint])
    const int32_t *owner, *neigh;
    double *grad;   // [chunk of fields][3][nc] scratch
    double *wout;   // = wface, written by the weights kernel
    int k0, kn, scheme;
    unsigned long long *est; // limited schemes: bits of max over fields and cells of sum|off-diagonal| / diagonal
    double rdt, relax, tol, rel_tol;
    int min_iter, max_iter;
};

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// all kBlock threads call; thread 0 writes the CTA's NS sums of field k to its column of the partials
template <int NS>
__device__ __forceinline__ void block_sums(const ImpParams &q, int k, double (&v)[NS]) {
    __shared__ double sm[kNS][kBlock / 32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int j = 0; j < NS; ++j) {
        const double s = warp_sum(v[j]);
        if (lane == 0) sm[j][w] = s;
    }
    __syncthreads();
    if (threadIdx.x < NS) {
        double s = 0.0;
#pragma unroll
        for (int i = 0; i < kBlock / 32; ++i) s += sm[threadIdx.x][i];
        q.partial[((size_t)k * kNS + threadIdx.x) * q.G + blockIdx.x] = s;
    }
    __syncthreads();
}

// The per-field sums of a CTA that sweeps a chunk of fields: every warp parks its sums in shared memory, ONE barrier at
// the end of the chunk, then thread j adds the warps' values of field k0 + j in a fixed order (instead of two barriers
// per field as block_sums needs).
template <int NS>
struct ChunkSums {
    double (*w)[NS][kBlock / 32];
    __device__ __forceinline__ void put(int kk, const double (&v)[NS]) {
        const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
        for (int j = 0; j < NS; ++j) {
            const double s = warp_sum(v[j]);
            if (lane == 0) w[kk][j][wid] = s;
        }
    }
    __device__ __forceinline__ void flush(const ImpParams &q, int k0, int kend) {
        __syncthreads();
        const int k = k0 + threadIdx.x;
        if (k < kend) {
#pragma unroll
            for (int j = 0; j < NS; ++j) {
                double s = 0.0;
#pragma unroll
                for (int i = 0; i < kBlock / 32; ++i) s += w[threadIdx.x][j][i];
                q.partial[((size_t)k * kNS + j) * q.G + blockIdx.x] = s;
            }
        }
    }
};

// gaussConvectionScheme::fvmDiv with the upwind weights w = pos(phi): lower = -w phi, upper = lower + phi
__device__ __forceinline__ void face_coeffs(double ph, double &lower, double &upper) {
    const double w = ph >= 0.0 ? 1.0 : 0.0;
    lower = -w * ph;
    upper = lower + ph;
}

// Assembly of the part of the matrix all fields share: diagonal (after relax()), its inverse, the row sum A 1 (for the
// residual's normalisation) and relax()'s source factor.
__global__ void __launch_bounds__(kBlock) imp_diag_kernel(const ImpParams q) {
    const int64_t c = (int64_t)blockIdx.x * kBlock + threadIdx.x;
    if (c >= q.nc) return;
    const double V = __ldg(q.V + c);
    double D0 = q.rdt * V, sumOff = 0.0, offsum = 0.0, ic = 0.0, icmag = 0.0, icc = 0.0;
    for (int s = 0; s < q.W; ++s) {
        const int32_t oth = __ldg(q.other + (size_t)s * q.nc + c), fc = __ldg(q.face + (size_t)s * q.nc + c);
        if (oth >= 0) {
            const bool own = fc >= 0;
            double lower, upper;
            face_coeffs(__ldg(q.phi + (own ? fc : ~fc)), lower, upper);
            D0 -= own ? lower : upper; // negSumDiag
            const double coef = own ? upper : lower;
            sumOff += fabs(coef);
            offsum += coef;
        } else if (oth <= -2) {
            const int64_t bf = -(int64_t)oth - 2;
            const double phb = __ldg(q.phi + q.nint + bf);
            if (bf >= q.coupled0 && bf < q.coupled0 + q.ncpl) {
                // processor patch = coupled: internalCoeffs = w phi_b on the diagonal, boundaryCoeffs = -(1 - w) phi_b
                // times the value across: an off-diagonal coefficient like an internal face's upper
                double lower, upper;
                face_coeffs(phb, lower, upper);
                icc -= lower;
                sumOff += fabs(upper);
                offsum += upper;
            } else if (q.psi_b == nullptr) { // zero-gradient patch face: internalCoeffs = phi_b
                ic += phb;
                icmag += fabs(phb);
            }
        }
    }
    double drel = 0.0;
    if (q.relax > 0.0) { // fvMatrix::relax(alpha): coupled patches add their (signed) diagonal, uncoupled the magnitude
        const double Dr = fmax(fabs(D0 + icc + icmag), sumOff) / q.relax;
        const double Dfin = Dr - ic - icc;
        drel = Dfin - D0;
        D0 = Dfin;
    }
    const double diag = D0 + ic + icc; // addBoundaryDiag
    q.diag[c] = diag;
    q.dinv[c] = 1.0 / diag;
    q.a1[c] = diag + offsum;
    q.drel[c] = drel;
}

// The cell's row for the field sweeps: columns and off-diagonal coefficients (x the column's inverse diagonal when the
// operator is A M^-1), W slots in registers; W == 0: the tables are walked per field (cells with more than 8 faces).
template <int W>
struct Row {
    int32_t col[W > 0 ? W : 1];
    double cf[W > 0 ? W : 1];
    double diag, dinv;
};

template <int W>
__device__ __forceinline__ void slot(const ImpParams &q, int64_t c, int s, bool precond, int32_t &col, double &cf) {
    const int32_t oth = __ldg(q.other + (size_t)s * q.nc + c), fc = __ldg(q.face + (size_t)s * q.nc + c);
    col = (int32_t)c;
    cf = 0.0;
    if (oth >= 0) {
        const bool own = fc >= 0;
        double lower, upper;
        face_coeffs(__ldg(q.phi + (own ? fc : ~fc)), lower, upper);
        col = oth;
        cf = (own ? upper : lower) * (precond ? __ldg(q.dinv + oth) : 1.0);
    } else if (oth <= -2) {
        const int64_t bf = -(int64_t)oth - 2;
        if (bf >= q.coupled0 && bf < q.coupled0 + q.ncpl) { // column -(j + 1): entry j of the halo row
            double lower, upper;
            face_coeffs(__ldg(q.phi + q.nint + bf), lower, upper);
            col = -(int32_t)(bf - q.coupled0) - 1;
            cf = upper; // Jacobi only on decomposed meshes: no preconditioned operator
        }
    }
}

// the value a column refers to: a cell of this sub-domain, or the cell across a processor-patch face (halo row)
__device__ __forceinline__ const double *col_src(const double *field, const double *halo, int32_t col) {
    return col >= 0 ? field + col : halo + (-(int64_t)col - 1);
}

template <int W>
__device__ __forceinline__ void load_row(const ImpParams &q, int64_t c, bool precond, Row<W> &R) {
    R.diag = __ldg(q.diag + c);
    R.dinv = __ldg(q.dinv + c);
    if constexpr (W > 0) {
#pragma unroll
        for (int s = 0; s < W; ++s) slot<W>(q, c, s, precond, R.col[s], R.cf[s]);
    }
}

// sum_s cf_s in[col_s]
template <int W>
__device__ __forceinline__ double offdiag(const ImpParams &q, int64_t c, bool precond, const Row<W> &R,
                                          const double *__restrict__ in, const double *__restrict__ halo = nullptr) {
    double acc = 0.0;
    if constexpr (W > 0) {
#pragma unroll
        for (int s = 0; s < W; ++s) acc = fma(R.cf[s], __ldg(col_src(in, halo, R.col[s])), acc);
    } else {
        for (int s = 0; s < q.W; ++s) {
            int32_t col;
            double cf;
            slot<0>(q, c, s, precond, col, cf);
            acc = fma(cf, __ldg(col_src(in, halo, col)), acc);
        }
    }
    return acc;
}

// gAverage(psi) per field, the xbar of lduMatrix::solver::normFactor
__global__ void __launch_bounds__(kBlock) imp_xsum_kernel(const ImpParams q) {
    const int k = blockIdx.y;
    const double *x = q.psi + (size_t)k * q.stride;
    double s[1] = {0.0};
#pragma unroll
    for (int j = 0; j < kPW; ++j) {
        const int64_t i = ((int64_t)blockIdx.x * kPW + j) * kBlock + threadIdx.x;
        if (i < q.nc) s[0] += x[i];
    }
    block_sums<1>(q, k, s);
}

// x = psi0;  b = rDeltaT psi0 V + S V (+ relax()'s (D - D0) psi0) - sum_fixed-value-faces phi_b psi_b;
// r = r0 = p = b - A x;  sums: normFactor, |r|_1, (r0, r)
template <int W, bool JAC>
__global__ void __launch_bounds__(kBlock) imp_init_kernel(const ImpParams q) {
    const int64_t c0 = (int64_t)blockIdx.x * kBlock + threadIdx.x;
    const bool valid = c0 < q.nc;
    const int64_t c = valid ? c0 : q.nc - 1;
    Row<W> R;
    load_row<W>(q, c, false, R);
    const double V = __ldg(q.V + c), a1 = __ldg(q.a1 + c), drel = __ldg(q.drel + c);
    __shared__ double csum[kMaxChunk][3][kBlock / 32];
    ChunkSums<3> sums{csum};
    const int kbeg = blockIdx.y * q.chunk;
    const int kend = min(q.nfields, (int)(blockIdx.y + 1) * q.chunk);
    for (int k = kbeg; k < kend; ++k) {
        const double *x = q.psi + (size_t)k * q.stride;
        const double mine = x[c];
        double b = q.rdt * mine * V;
        if (q.S) b += __ldg(q.S + (size_t)k * q.stride + c) * V;
        b += drel * mine;
        if (q.psi_b) {
            for (int s = 0; s < q.W; ++s) {
                const int32_t oth = __ldg(q.other + (size_t)s * q.nc + c);
                if (oth <= -2) {
                    const int64_t bf = -(int64_t)oth - 2;
                    if (bf >= q.coupled0 && bf < q.coupled0 + q.ncpl) continue; // processor patch: in the matrix
                    b -= __ldg(q.phi + q.nint + bf) * __ldg(q.psi_b + (size_t)k * q.stride_b + bf);
                }
            }
        }
        const double Ax = fma(R.diag, mine, offdiag<W>(q, c, false, R, x, q.halo_x ? q.halo_x + (size_t)k * q.ncpl : nullptr));
        const double res = b - Ax;
        if (valid) {
            const size_t o = (size_t)k * q.nc + c;
            q.x[(size_t)k * q.stride + c] = mine;
            if (JAC) {
                q.p[o] = R.dinv * res; // z0 = M^-1 r0
            } else {
                q.r[o] = res;
                q.r0[o] = res;
                q.p[o] = res;
            }
        }
        const double pA = __ldg(&q.sc[k].xbar) * a1;
        double s[3] = {valid ? fabs(Ax - pA) + fabs(b - pA) : 0.0, valid ? fabs(res) : 0.0, valid ? res * res : 0.0};
        sums.put(k - kbeg, s);
    }
    sums.flush(q, kbeg, kend);
}

// out = A M^-1 in.  MODE 0: in = p, out = v, sum (r0, v).  MODE 1: in = s (held in r), out = t, sums (t, s), (t, t).
template <int W, int MODE>
__global__ void __launch_bounds__(kBlock) imp_spmv_kernel(const ImpParams q) {
    const int64_t c0 = (int64_t)blockIdx.x * kBlock + threadIdx.x;
    const bool valid = c0 < q.nc;
    const int64_t c = valid ? c0 : q.nc - 1;
    Row<W> R;
    load_row<W>(q, c, true, R);
    const int kend = min(q.nfields, (int)(blockIdx.y + 1) * q.chunk);
    for (int k = blockIdx.y * q.chunk; k < kend; ++k) {
        const FieldSc *sc = q.sc + k;
        if (sc->state != 0 || (MODE == 1 && sc->half)) continue; // uniform over the CTA
        const size_t o = (size_t)k * q.nc;
        const double *in = (MODE == 0 ? q.p : q.r) + o;
        const double mine = in[c];
        const double y = fma(R.diag, R.dinv * mine, offdiag<W>(q, c, true, R, in));
        if (MODE == 0) {
            if (valid) q.v[o + c] = y;
            double s[1] = {valid ? __ldg(q.r0 + o + c) * y : 0.0};
            block_sums<1>(q, k, s);
        } else {
            if (valid) q.t[o + c] = y;
            double s[2] = {valid ? y * mine : 0.0, valid ? y * y : 0.0};
            block_sums<2>(q, k, s);
        }
    }
}

// s = r - alpha v (in place); sum |s|_1
__global__ void __launch_bounds__(kBlock) imp_s_kernel(const ImpParams q) {
    const int k = blockIdx.y;
    const FieldSc *sc = q.sc + k;
    if (sc->state != 0) return;
    const double alpha = sc->alpha;
    const size_t o = (size_t)k * q.nc;
    double s[1] = {0.0};
#pragma unroll
    for (int j = 0; j < kPW; ++j) {
        const int64_t i = ((int64_t)blockIdx.x * kPW + j) * kBlock + threadIdx.x;
        if (i < q.nc) {
            const double sv = fma(-alpha, q.v[o + i], q.r[o + i]);
            q.r[o + i] = sv;
            s[0] += fabs(sv);
        }
    }
    block_sums<1>(q, k, s);
}

// x += alpha M^-1 p + omega M^-1 s;  r = s - omega t;  sums |r|_1, (r0, r)
__global__ void __launch_bounds__(kBlock) imp_update_kernel(const ImpParams q) {
    const int k = blockIdx.y;
    const FieldSc *sc = q.sc + k;
    if (sc->state != 0) return;
    const double alpha = sc->alpha, omega = sc->omega;
    const size_t o = (size_t)k * q.nc;
    double *x = q.x + (size_t)k * q.stride;
    double s[2] = {0.0, 0.0};
#pragma unroll
    for (int j = 0; j < kPW; ++j) {
        const int64_t i = ((int64_t)blockIdx.x * kPW + j) * kBlock + threadIdx.x;
        if (i < q.nc) {
            const double di = q.dinvK ? __ldg(q.dinvK + o + i) : __ldg(q.dinv + i), sv = q.r[o + i];
            x[i] += di * fma(alpha, q.p[o + i], omega * sv);
            const double rn = omega != 0.0 ? fma(-omega, q.t[o + i], sv) : sv;
            q.r[o + i] = rn;
            s[0] += fabs(rn);
            s[1] += __ldg(q.r0 + o + i) * rn;
        }
    }
    block_sums<2>(q, k, s);
}

// p = r + beta (p - omega v)
__global__ void __launch_bounds__(kBlock) imp_p_kernel(const ImpParams q) {
    const int k = blockIdx.y;
    const FieldSc *sc = q.sc + k;
    if (sc->state != 0) return;
    const double beta = sc->beta, omega = sc->omega;
    const size_t o = (size_t)k * q.nc;
#pragma unroll
    for (int j = 0; j < kPW; ++j) {
        const int64_t i = ((int64_t)blockIdx.x * kPW + j) * kBlock + threadIdx.x;
        if (i < q.nc) q.p[o + i] = fma(beta, fma(-omega, q.v[o + i], q.p[o + i]), q.r[o + i]);
    }
}

// Jacobi in correction form: x_{k+1} = x_k + z_k, z_{k+1} = -D^-1 R z_k (R = the off-diagonal part), so that
// D z_{k+1} = b - A x_{k+1} is the CURRENT iterate's residual: sum |D z_{k+1}|.  The matrix is column-dominant (column
// sums = V / dt + boundary outflow), hence the residual's L1 norm contracts at every step.  z_k in p, z_{k+1} in v.
template <int W, bool HALO>
__global__ void __launch_bounds__(kBlock) imp_jacobi_kernel(const ImpParams q) {
    __shared__ double wsum[kMaxChunk][kBlock / 32];
    const int k0 = blockIdx.y * q.chunk, kend = min(q.nfields, k0 + q.chunk);
    // converged fields are skipped: uniform over the CTA, a chunk with none left costs a few loads
    auto next_active = [&](int k) {
        while (k < kend && q.sc[k].state != 0) ++k;
        return k;
    };
    int k = next_active(k0);
    if (k >= kend) return;
    const int64_t c0 = (int64_t)blockIdx.x * kBlock + threadIdx.x;
    const bool valid = c0 < q.nc;
    const int64_t c = valid ? c0 : q.nc - 1;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    Row<W> R;
    load_row<W>(q, c, false, R);
    if constexpr (W > 0) {
        // The W gathered values of the NEXT active field travel global -> shared (LDGSTS) into this thread's own column
        // while the current field is summed: no registers held by loads in flight, no barrier (nothing is shared
        // between threads) -- the pipeline of the explicit update (pbe_transport.cu, sweep_async).  The coalesced
        // streams (own z, x) go through registers one field ahead.
        __shared__ double stage[2][W][kBlock];
        auto cp8 = [](double *dst, const double *src) {
            asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src)
                         : "memory");
        };
        // Addresses: one base pointer per field (uniform) + per-thread BYTE offsets computed once -- the compiler
        // otherwise rebuilds k nc + col in 64 bits for every slot and field (6 instructions per gather: the sweep was
        // issue-bound, ncu: issue slots 79 % busy, DRAM 49 %).
        long long off[W];
#pragma unroll
        for (int s = 0; s < W; ++s) off[s] = (long long)R.col[s] * 8;
        const long long c8 = (long long)c * 8;
#if PBE_IMP_REGSTAGE
        // decomposed meshes: only the warps that hold a processor-patch cell need the halo row (and the per-slot choice of
        // address); all others -- the bulk -- run the same loop as on one rank
        bool warp_halo = false;
        if constexpr (HALO) {
            bool mine_halo = false;
#pragma unroll
            for (int s = 0; s < W; ++s) mine_halo |= R.col[s] < 0;
            warp_halo = __any_sync(0xffffffffu, mine_halo);
        }
        if (!HALO || !warp_halo) {
            // the next field's gathers wait in registers instead of shared memory (6 LDG instead of 6 LDGSTS + 6 LDS
            // per cell-field on the load-store pipe); the loads have a whole field's work to land
            double gn[W], mine_n = 0.0, x_n = 0.0;
            auto issue = [&](int kk) {
                if (kk < kend) {
                    const char *zb = reinterpret_cast<const char *>(q.p + (size_t)kk * q.nc);
                    const char *xb = reinterpret_cast<const char *>(q.x + (size_t)kk * q.stride);
#pragma unroll
                    for (int s = 0; s < W; ++s) gn[s] = __ldg(reinterpret_cast<const double *>(zb + off[s]));
                    mine_n = __ldg(reinterpret_cast<const double *>(zb + c8));
                    x_n = *reinterpret_cast<const double *>(xb + c8);
                }
            };
            issue(k);
            while (k < kend) {
                const int kn = next_active(k + 1);
                double g[W];
#pragma unroll
                for (int s = 0; s < W; ++s) g[s] = gn[s];
                const double mine = mine_n, xv = x_n;
                char *xo = reinterpret_cast<char *>(q.x + (size_t)k * q.stride), *vo = reinterpret_cast<char *>(q.v + (size_t)k * q.nc);
                issue(kn);
                double acc = 0.0;
#pragma unroll
                for (int s = 0; s < W; ++s) acc = fma(R.cf[s], g[s], acc);
                const double rz = -acc;
                if (valid) {
                    __stcs(reinterpret_cast<double *>(xo + c8), xv + mine);
                    __stcs(reinterpret_cast<double *>(vo + c8), R.dinv * rz);
                }
                const double sum = warp_sum(valid ? fabs(rz) : 0.0);
                if (lane == 0) wsum[k - k0][wid] = sum;
                k = kn;
            }
        } else
#endif
        {
        double mine_n = 0.0, x_n = 0.0;
        auto issue = [&](int kk, int par) {
            if (kk < kend) {
                const char *zb = reinterpret_cast<const char *>(q.p + (size_t)kk * q.nc);
                const char *xb = reinterpret_cast<const char *>(q.x + (size_t)kk * q.stride);
                // (the choice between the field and the halo row costs ~10 instructions per slot and field: only
                // decomposed meshes pay it)
                const double *zh = HALO ? q.halo_z + (size_t)kk * q.ncpl : nullptr;
#pragma unroll
                for (int s = 0; s < W; ++s)
                    cp8(&stage[par][s][threadIdx.x], HALO ? col_src(reinterpret_cast<const double *>(zb), zh, R.col[s])
                                                          : reinterpret_cast<const double *>(zb + off[s]));
                mine_n = __ldg(reinterpret_cast<const double *>(zb + c8));
                x_n = *reinterpret_cast<const double *>(xb + c8);
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        };
        int par = 0;
        issue(k, par);
        while (k < kend) {
            const int kn = next_active(k + 1);
            const double mine = mine_n, xv = x_n;
            char *xo = reinterpret_cast<char *>(q.x + (size_t)k * q.stride), *vo = reinterpret_cast<char *>(q.v + (size_t)k * q.nc);
            issue(kn, par ^ 1);
            asm volatile("cp.async.wait_group 1;" ::: "memory");
            double acc = 0.0;
#pragma unroll
            for (int s = 0; s < W; ++s) acc = fma(R.cf[s], stage[par][s][threadIdx.x], acc);
            const double rz = -acc;
            if (valid) {
                __stcs(reinterpret_cast<double *>(xo + c8), xv + mine);
                __stcs(reinterpret_cast<double *>(vo + c8), R.dinv * rz);
            }
            const double sum = warp_sum(valid ? fabs(rz) : 0.0);
            if (lane == 0) wsum[k - k0][wid] = sum;
            k = kn;
            par ^= 1;
        }
        }
    } else {
        for (; k < kend; k = next_active(k + 1)) {
            const size_t o = (size_t)k * q.nc;
            const double *zin = q.p + o;
            const double rz = -offdiag<W>(q, c, false, R, zin, HALO ? q.halo_z + (size_t)k * q.ncpl : nullptr);
            if (valid) {
                q.x[(size_t)k * q.stride + c] += zin[c];
                q.v[o + c] = R.dinv * rz;
            }
            const double sum = warp_sum(valid ? fabs(rz) : 0.0);
            if (lane == 0) wsum[k - k0][wid] = sum;
        }
    }
    __syncthreads(); // one barrier per chunk, not per field: the warps' sums wait in shared memory
    const int kt = k0 + threadIdx.x;
    if (kt < kend && q.sc[kt].state == 0) {
        double sum = 0.0;
#pragma unroll
        for (int i = 0; i < kBlock / 32; ++i) sum += wsum[threadIdx.x][i];
        q.partial[(size_t)kt * kNS * q.G + blockIdx.x] = sum;
    }
}

// ---- limited schemes in the implicit term (Gauss vanLeer, Gauss limitedLinear 1) ------------------------------------
// The weights come from the CURRENT field (limitedScheme::weights(vf) inside gaussConvectionScheme::fvmDiv), so every
// field has its own matrix: lower = -w phi, upper = lower + phi with w = wface[k][f].  Nothing but the inverse diagonal
// is stored per field; a sweep rebuilds a cell's coefficients from phi (registers) and the field's face weights.
template <int W>
struct LimRow {
    int32_t col[W], fidx[W];
    double ph[W], addc[W]; // coefficient of slot s = addc - w ph: owner side (1 - w) phi, neighbour side -w phi
};

template <int W>
__device__ __forceinline__ void load_lim_row(const ImpParams &q, int64_t c, LimRow<W> &R) {
#pragma unroll
    for (int s = 0; s < W; ++s) {
        const int32_t oth = __ldg(q.other + (size_t)s * q.nc + c), fc = __ldg(q.face + (size_t)s * q.nc + c);
        R.col[s] = (int32_t)c, R.fidx[s] = 0, R.ph[s] = 0.0, R.addc[s] = 0.0;
        if (oth >= 0) {
            const bool own = fc >= 0;
            const int32_t f = own ? fc : ~fc;
            const double ph = __ldg(q.phi + f);
            R.col[s] = oth, R.fidx[s] = f, R.ph[s] = ph, R.addc[s] = own ? ph : 0.0;
        }
    }
}

// per field: the diagonal (negSumDiag, patches, relax()), b, r0 = b - A x0, z0 = r0 / D; sums: normFactor, |r0|_1
template <int W, bool JAC>
__global__ void __launch_bounds__(kBlock) imp_lim_init_kernel(const ImpParams q) {
    const int64_t c0 = (int64_t)blockIdx.x * kBlock + threadIdx.x;
    const bool valid = c0 < q.nc;
    const int64_t c = valid ? c0 : q.nc - 1;
    LimRow<W> R;
    load_lim_row<W>(q, c, R);
    const double V = __ldg(q.V + c);
    // zero-gradient patch faces: internalCoeffs = phi_b (the same for every field)
    double ic = 0.0, icmag = 0.0;
    for (int s = 0; s < q.W; ++s) {
        const int32_t oth = __ldg(q.other + (size_t)s * q.nc + c);
        if (oth <= -2 && q.psi_b == nullptr) {
            const double phb = __ldg(q.phi + q.nint + (-(int64_t)oth - 2));
            ic += phb;
            icmag += fabs(phb);
        }
    }
    __shared__ double csum[kMaxChunk][3][kBlock / 32];
    ChunkSums<3> sums{csum};
    const int kbeg = blockIdx.y * q.chunk;
    const int kend = min(q.nfields, (int)(blockIdx.y + 1) * q.chunk);
    for (int k = kbeg; k < kend; ++k) {
        const double *x = q.psi + (size_t)k * q.stride;
        const double *wk = q.wface + (size_t)k * q.nint;
        const double mine = x[c];
        double D0 = q.rdt * V, sumOff = 0.0, offsum = 0.0, offx = 0.0;
#pragma unroll
        for (int s = 0; s < W; ++s) {
            const double w = __ldg(wk + R.fidx[s]);
            const double cf = fma(-w, R.ph[s], R.addc[s]);
            D0 += fma(w, R.ph[s], R.addc[s] - R.ph[s]); // -lower on the owner side, -upper on the neighbour side
            sumOff += fabs(cf);
            offsum += cf;
            offx = fma(cf, __ldg(x + R.col[s]), offx);
        }
        double b = q.rdt * mine * V;
        if (q.S) b += __ldg(q.S + (size_t)k * q.stride + c) * V;
        if (q.relax > 0.0) { // fvMatrix::relax(alpha)
            const double Dfin = fmax(fabs(D0 + icmag), sumOff) / q.relax - ic;
            b += (Dfin - D0) * mine;
            D0 = Dfin;
        }
        const double diag = D0 + ic;
        if (q.psi_b) {
            for (int s = 0; s < q.W; ++s) {
                const int32_t oth = __ldg(q.other + (size_t)s * q.nc + c);
                if (oth <= -2) {
                    const int64_t bf = -(int64_t)oth - 2;
                    b -= __ldg(q.phi + q.nint + bf) * __ldg(q.psi_b + (size_t)k * q.stride_b + bf);
                }
            }
        }
        const double Ax = fma(diag, mine, offx);
        const double res = b - Ax, dinv = 1.0 / diag;
        if (valid) {
            const size_t o = (size_t)k * q.nc + c;
            q.x[(size_t)k * q.stride + c] = mine;
            q.dinvK[o] = dinv;
            if (JAC) {
                q.p[o] = dinv * res;
            } else {
                q.r[o] = res;
                q.r0[o] = res;
                q.p[o] = res;
            }
        }
        const double pA = __ldg(&q.sc[k].xbar) * (diag + offsum);
        double sm[3] = {valid ? fabs(Ax - pA) + fabs(b - pA) : 0.0, valid ? fabs(res) : 0.0, valid ? res * res : 0.0};
        sums.put(k - kbeg, sm);
    }
    sums.flush(q, kbeg, kend);
}

// gaussGrad with linear interpolation for a chunk of fields: grad_c = (sum_f Sf psi_f) / V, faces in ascending order.
// The cell's geometry (per slot: +-Sf, linear weight) lives in registers, a field costs its own value and W gathers.
template <int W>
__global__ void __launch_bounds__(kBlock) imp_lim_grad_kernel(const ImpParams q) {
    const int64_t c = (int64_t)blockIdx.x * kBlock + threadIdx.x;
    if (c >= q.nc) return;
    const int64_t nf = q.nint + q.nbnd;
    int32_t col[W], kind[W]; // kind: 0 padding, 1 owner side, 2 neighbour side, 3 + b boundary face b
    double sx[W], sy[W], sz[W], wl[W];
#pragma unroll
    for (int s = 0; s < W; ++s) {
        const int32_t oth = __ldg(q.other + (size_t)s * q.nc + c), fc = __ldg(q.face + (size_t)s * q.nc + c);
        col[s] = (int32_t)c, kind[s] = 0, sx[s] = sy[s] = sz[s] = 0.0, wl[s] = 0.0;
        if (oth >= 0) {
            const bool own = fc >= 0;
            const int64_t f = own ? fc : ~fc;
            const double sg = own ? 1.0 : -1.0;
            col[s] = oth, kind[s] = own ? 1 : 2, wl[s] = __ldg(q.wlin + f);
            sx[s] = sg * __ldg(q.Sf + f), sy[s] = sg * __ldg(q.Sf + nf + f), sz[s] = sg * __ldg(q.Sf + 2 * nf + f);
        } else if (oth <= -2) {
            const int64_t b = -(int64_t)oth - 2, f = q.nint + b;
            kind[s] = 3 + (int32_t)b;
            sx[s] = __ldg(q.Sf + f), sy[s] = __ldg(q.Sf + nf + f), sz[s] = __ldg(q.Sf + 2 * nf + f);
        }
    }
    const double rV = 1.0 / __ldg(q.V + c);
    const int kend = min(q.kn, (int)(blockIdx.y + 1) * q.chunk);
    for (int kc = blockIdx.y * q.chunk; kc < kend; ++kc) {
        const int k = q.k0 + kc;
        const double *psi = q.psi + (size_t)k * q.stride;
        const double mine = psi[c];
        double g0 = 0.0, g1 = 0.0, g2 = 0.0;
#pragma unroll
        for (int s = 0; s < W; ++s) {
            double fv;
            if (kind[s] >= 3)
                fv = q.psi_b ? __ldg(q.psi_b + (size_t)k * q.stride_b + (kind[s] - 3)) : mine;
            else {
                const double theirs = __ldg(psi + col[s]);
                const double P = kind[s] == 1 ? mine : theirs, N = kind[s] == 1 ? theirs : mine;
                fv = fma(wl[s], P - N, N);
            }
            g0 = fma(sx[s], fv, g0), g1 = fma(sy[s], fv, g1), g2 = fma(sz[s], fv, g2);
        }
        double *g = q.grad + (size_t)kc * 3 * q.nc;
        g[c] = g0 * rV, g[q.nc + c] = g1 * rV, g[2 * q.nc + c] = g2 * rV;
    }
}

// limitedScheme::weights(vf) of the internal faces for the chunk's fields: limiter CDweight + (1 - limiter) pos(phi),
// limiter(r) of vanLeer or limitedLinear(1), r = NVDTVD::r from the upwind cell's gradient.  One thread per face keeps
// the face's geometry in registers.
__global__ void __launch_bounds__(kBlock) imp_lim_weights_kernel(const ImpParams q) {
    const int64_t f = (int64_t)blockIdx.x * kBlock + threadIdx.x;
    if (f >= q.nint) return;
    const int32_t own = __ldg(q.owner + f), nei = __ldg(q.neigh + f);
    const double ph = __ldg(q.phi + f), wl = __ldg(q.wlin + f), up_w = ph >= 0.0 ? 1.0 : 0.0;
    const int32_t up = ph > 0.0 ? own : nei;
    const double d0 = __ldg(q.Cc + nei) - __ldg(q.Cc + own), d1 = __ldg(q.Cc + q.nc + nei) - __ldg(q.Cc + q.nc + own),
                 d2 = __ldg(q.Cc + 2 * q.nc + nei) - __ldg(q.Cc + 2 * q.nc + own);
    const int kend = min(q.kn, (int)(blockIdx.y + 1) * q.chunk);
    for (int kc = blockIdx.y * q.chunk; kc < kend; ++kc) {
        const int k = q.k0 + kc;
        const double *psi = q.psi + (size_t)k * q.stride;
        const double *g = q.grad + (size_t)kc * 3 * q.nc;
        const double P = __ldg(psi + own), N = __ldg(psi + nei);
        const double gradcf = d0 * g[up] + d1 * g[q.nc + up] + d2 * g[2 * q.nc + up];
        const double gradf = N - P;
        double r;
        if (fabs(gradcf) >= 1000 * fabs(gradf))
            r = 2 * 1000 * (gradcf >= 0 ? 1.0 : -1.0) * (gradf >= 0 ? 1.0 : -1.0) - 1;
        else
            r = 2 * (gradcf / gradf) - 1;
        const double lim = q.scheme == PBE_DIV_VANLEER ? (r + fabs(r)) / (1 + fabs(r)) : fmax(fmin(2.0 * r, 1.0), 0.0);
        q.wout[(size_t)k * q.nint + f] = lim * wl + (1.0 - lim) * up_w;
    }
}

// How far from diagonally dominant the fields' matrices are: max over fields and cells of sum|off-diagonal| / diagonal
// (< 1: Jacobi contracts).  Downwind weights put positive coefficients off the diagonal, so unlike the upwind matrix
// these are M-matrices only while the cell Courant number is small.
template <int W>
__global__ void __launch_bounds__(kBlock) imp_lim_est_kernel(const ImpParams q) {
    const int64_t c0 = (int64_t)blockIdx.x * kBlock + threadIdx.x;
    const bool valid = c0 < q.nc;
    const int64_t c = valid ? c0 : q.nc - 1;
    LimRow<W> R;
    load_lim_row<W>(q, c, R);
    const double V = __ldg(q.V + c);
    double ic = 0.0;
    for (int s = 0; s < q.W; ++s) {
        const int32_t oth = __ldg(q.other + (size_t)s * q.nc + c);
        if (oth <= -2 && q.psi_b == nullptr) ic += __ldg(q.phi + q.nint + (-(int64_t)oth - 2));
    }
    double worst = 0.0;
    const int kend = min(q.nfields, (int)(blockIdx.y + 1) * q.chunk);
    for (int k = blockIdx.y * q.chunk; k < kend; ++k) {
        const double *wk = q.wface + (size_t)k * q.nint;
        double D0 = q.rdt * V + ic, sumOff = 0.0;
#pragma unroll
        for (int s = 0; s < W; ++s) {
            const double w = __ldg(wk + R.fidx[s]);
            D0 += fma(w, R.ph[s], R.addc[s] - R.ph[s]);
            sumOff += fabs(fma(-w, R.ph[s], R.addc[s]));
        }
        const double e = D0 > 0.0 ? sumOff / D0 : 1.0e30;
        worst = fmax(worst, e);
    }
    if (!valid) worst = 0.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) worst = fmax(worst, __shfl_xor_sync(0xffffffffu, worst, o));
    if ((threadIdx.x & 31) == 0 && worst > 0.0) atomicMax(q.est, (unsigned long long)__double_as_longlong(worst));
}

// out = A_k M_k^-1 in for the fields' own matrices (BiCGStab on a limited scheme); modes as imp_spmv_kernel
template <int W, int MODE>
__global__ void __launch_bounds__(kBlock) imp_lim_spmv_kernel(const ImpParams q) {
    const int64_t c0 = (int64_t)blockIdx.x * kBlock + threadIdx.x;
    const bool valid = c0 < q.nc;
    const int64_t c = valid ? c0 : q.nc - 1;
    LimRow<W> R;
    load_lim_row<W>(q, c, R);
    const int kend = min(q.nfields, (int)(blockIdx.y + 1) * q.chunk);
    for (int k = blockIdx.y * q.chunk; k < kend; ++k) {
        const FieldSc *sc = q.sc + k;
        if (sc->state != 0 || (MODE == 1 && sc->half)) continue;
        const size_t o = (size_t)k * q.nc;
        const double *in = (MODE == 0 ? q.p : q.r) + o;
        const double *wk = q.wface + (size_t)k * q.nint, *di = q.dinvK + o;
        const double mine = in[c];
        double y = mine; // D (D^-1 in)
#pragma unroll
        for (int s = 0; s < W; ++s)
            y = fma(fma(-__ldg(wk + R.fidx[s]), R.ph[s], R.addc[s]) * __ldg(di + R.col[s]), __ldg(in + R.col[s]), y);
        if (MODE == 0) {
            if (valid) q.v[o + c] = y;
            double sm[1] = {valid ? __ldg(q.r0 + o + c) * y : 0.0};
            block_sums<1>(q, k, sm);
        } else {
            if (valid) q.t[o + c] = y;
            double sm[2] = {valid ? y * mine : 0.0, valid ? y * y : 0.0};
            block_sums<2>(q, k, sm);
        }
    }
}

template <int W>
__global__ void __launch_bounds__(kBlock) imp_lim_jacobi_kernel(const ImpParams q) {
    extern __shared__ double lim_stage[]; // [2][2 W][kBlock]: the gathered z and face weights of two fields
    __shared__ double wsum[kMaxChunk][kBlock / 32];
    const int k0 = blockIdx.y * q.chunk, kend = min(q.nfields, k0 + q.chunk);
    auto next_active = [&](int k) {
        while (k < kend && q.sc[k].state != 0) ++k;
        return k;
    };
    int k = next_active(k0);
    if (k >= kend) return;
    const int64_t c0 = (int64_t)blockIdx.x * kBlock + threadIdx.x;
    const bool valid = c0 < q.nc;
    const int64_t c = valid ? c0 : q.nc - 1;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    LimRow<W> R;
    load_lim_row<W>(q, c, R);
    // the pipeline of imp_jacobi_kernel, with the field's face weights travelling beside its z values
    auto cp8 = [](double *dst, const double *src) {
        asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src)
                     : "memory");
    };
    double mine_n = 0.0, x_n = 0.0, di_n = 0.0;
    auto issue = [&](int kk, int par) {
        if (kk < kend) {
            const size_t o = (size_t)kk * q.nc;
            const double *zin = q.p + o, *wk = q.wface + (size_t)kk * q.nint;
            double *st = lim_stage + (size_t)par * 2 * W * kBlock + threadIdx.x;
#pragma unroll
            for (int s = 0; s < W; ++s) {
                cp8(st + (size_t)s * kBlock, zin + R.col[s]);
                cp8(st + (size_t)(W + s) * kBlock, wk + R.fidx[s]);
            }
            mine_n = __ldg(zin + c);
            x_n = q.x[(size_t)kk * q.stride + c];
            di_n = __ldg(q.dinvK + o + c);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    int par = 0;
    issue(k, par);
    while (k < kend) {
        const int kn = next_active(k + 1);
        const double mine = mine_n, xv = x_n, di = di_n;
        issue(kn, par ^ 1);
        asm volatile("cp.async.wait_group 1;" ::: "memory");
        const double *st = lim_stage + (size_t)par * 2 * W * kBlock + threadIdx.x;
        double acc = 0.0;
#pragma unroll
        for (int s = 0; s < W; ++s)
            acc = fma(fma(-st[(size_t)(W + s) * kBlock], R.ph[s], R.addc[s]), st[(size_t)s * kBlock], acc);
        const double rz = -acc;
        if (valid) {
            __stcs(q.x + (size_t)k * q.stride + c, xv + mine);
            __stcs(q.v + (size_t)k * q.nc + c, di * rz);
        }
        const double sum = warp_sum(valid ? fabs(rz) : 0.0);
        if (lane == 0) wsum[k - k0][wid] = sum;
        k = kn;
        par ^= 1;
    }
    __syncthreads();
    const int kt = k0 + threadIdx.x;
    if (kt < kend && q.sc[kt].state == 0) {
        double sum = 0.0;
#pragma unroll
        for (int i = 0; i < kBlock / 32; ++i) sum += wsum[threadIdx.x][i];
        q.partial[(size_t)kt * kNS * q.G + blockIdx.x] = sum;
    }
}

// One CTA per field: the CTAs' partial sums added in a fixed order, then the field's scalars.
// mode 0: one rank, sums and states in one go.  Decomposed meshes split it around the all-reduce: mode 1 writes this
// rank's sums to gsum [nfields][kNS], mode 2 takes the all-reduced gsum and updates the states (every rank computes the
// same states from the same sums: the ranks' loops stay in step without talking about it).
__global__ void __launch_bounds__(kBlock) imp_finalize_kernel(const ImpParams q, int phase, int G, int mode) {
    const int k = blockIdx.x;
    FieldSc *sc = q.sc + k;
    if (phase >= PH_A && (sc->state != 0 || (phase == PH_C && sc->half))) return;
    __shared__ double sm[kNS][kBlock];
    if (mode != 2) {
        const double *part = q.partial + (size_t)k * kNS * q.G;
        for (int j = 0; j < kNS; ++j) {
            double s = 0.0;
            for (int i = threadIdx.x; i < G; i += kBlock) s += part[(size_t)j * q.G + i];
            sm[j][threadIdx.x] = s;
        }
        __syncthreads();
        for (int w = kBlock / 2; w > 0; w >>= 1) {
            if ((int)threadIdx.x < w)
                for (int j = 0; j < kNS; ++j) sm[j][threadIdx.x] += sm[j][threadIdx.x + w];
            __syncthreads();
        }
    }
    if (threadIdx.x != 0) return;
    if (mode == 1) {
        for (int j = 0; j < kNS; ++j) q.gsum[(size_t)k * kNS + j] = sm[j][0];
        return;
    }
    const double s0 = mode == 2 ? q.gsum[(size_t)k * kNS] : sm[0][0], s1 = mode == 2 ? q.gsum[(size_t)k * kNS + 1] : sm[1][0],
                 s2 = mode == 2 ? q.gsum[(size_t)k * kNS + 2] : sm[2][0];
    auto conv = [&](double res) { return res < q.tol || (q.rel_tol > 0.0 && res < q.rel_tol * sc->res0); };
    switch (phase) {
    case PH_XSUM: sc->xbar = s0 / (double)q.nc_global; break;
    case PH_INIT:
        sc->norm = s0 + kNormSmall;
        sc->res0 = sc->res = s1 / sc->norm;
        sc->rho = s2;
        sc->alpha = sc->omega = sc->beta = 0.0;
        sc->iters = 0;
        sc->half = 0;
        sc->state = (s2 == 0.0 || (q.min_iter <= 0 && conv(sc->res0))) ? 1 : 0;
        break;
    case PH_A: // alpha = rho / (r0, v)
        if (s0 == 0.0 || !isfinite(s0)) sc->state = 2;
        else sc->alpha = sc->rho / s0;
        break;
    case PH_B: { // PBiCGStab's mid-iteration exit: s already small enough
        const double sres = s0 / sc->norm;
        sc->half = (conv(sres) && sc->iters + 1 >= q.min_iter) ? 1 : 0;
        if (sc->half) sc->omega = 0.0;
    } break;
    case PH_C: sc->omega = s1 > 0.0 ? s0 / s1 : 0.0; break; // (t, s) / (t, t)
    case PH_J:
        sc->res = s0 / sc->norm;
        sc->iters += 1;
        if (conv(sc->res) && sc->iters >= q.min_iter) sc->state = 1;
        else if (sc->iters >= q.max_iter) sc->state = 3;
        else if (!(sc->res <= 1.0e6 * fmax(sc->res0, 1.0e-30))) sc->state = 2; // diverging (or NaN): stop before it overflows
        break;
    case PH_D: {
        sc->res = s0 / sc->norm;
        sc->iters += 1;
        if (sc->half || (conv(sc->res) && sc->iters >= q.min_iter)) sc->state = 1;
        else if (sc->iters >= q.max_iter) sc->state = 3;
        else if (sc->omega == 0.0 || s1 == 0.0 || !isfinite(s1)) sc->state = 2;
        else {
            sc->beta = (s1 / sc->rho) * (sc->alpha / sc->omega);
            sc->rho = s1;
        }
    } break;
    }
}

} // namespace

// Halo traffic of a sweep in ONE message per neighbour: the values along the processor patches are packed
// neighbour-major ([neighbour][field][face]: contiguous per neighbour) and unpacked into the field-major halo rows the
// sweeps read.  Face j of neighbour q (first face off_q, cnt_q faces) sits at nfields off_q + k cnt_q + (j - off_q).
__global__ void __launch_bounds__(256) imp_halo_pack_kernel(int nfields, int64_t fstride, const double *__restrict__ field,
                                                            int64_t ncpl, const int32_t *__restrict__ cells,
                                                            const int64_t *__restrict__ qoff, const int64_t *__restrict__ qcnt,
                                                            double *__restrict__ out) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= ncpl) return;
    const int64_t o = qoff[j], n = qcnt[j], base = (int64_t)nfields * o + (j - o);
    const int32_t c = cells[j];
    for (int k = blockIdx.y; k < nfields; k += gridDim.y) out[base + (int64_t)k * n] = __ldg(field + (size_t)k * fstride + c);
}
__global__ void __launch_bounds__(256) imp_halo_unpack_kernel(int nfields, int64_t ncpl, const int64_t *__restrict__ qoff,
                                                              const int64_t *__restrict__ qcnt, const double *__restrict__ in,
                                                              double *__restrict__ halo) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= ncpl) return;
    const int64_t o = qoff[j], n = qcnt[j], base = (int64_t)nfields * o + (j - o);
    for (int k = blockIdx.y; k < nfields; k += gridDim.y) halo[(size_t)k * ncpl + j] = in[base + (int64_t)k * n];
}

// every cell's faces as [W][ncells] slots in ascending face order (internal faces, then the patches' faces)
static int build_slots(pbe_mesh *m) {
    if (m->imp_W >= 0) return PBE_OK;
    const int64_t nc = m->ncells, nint = m->nint, nbnd = m->nbnd;
    std::vector<int32_t> cnt((size_t)nc, 0);
    for (int64_t f = 0; f < nint; ++f) {
        cnt[m->h_owner[f]]++;
        cnt[m->h_neigh[f]]++;
    }
    for (int64_t b = 0; b < nbnd; ++b) cnt[m->h_bown[b]]++;
    int maxf = 1;
    for (int64_t c = 0; c < nc; ++c) maxf = std::max(maxf, (int)cnt[c]);
    const int W = maxf <= 4 ? 4 : maxf <= 6 ? 6 : maxf <= 8 ? 8 : maxf;
    std::vector<int32_t> other((size_t)W * nc, -1), face((size_t)W * nc, 0), fill((size_t)nc, 0);
    auto put = [&](int32_t c, int32_t oth, int32_t fc) {
        const int32_t k = fill[c]++;
        other[(size_t)k * nc + c] = oth;
        face[(size_t)k * nc + c] = fc;
    };
    for (int64_t f = 0; f < nint; ++f) {
        put(m->h_owner[f], m->h_neigh[f], (int32_t)f);
        put(m->h_neigh[f], m->h_owner[f], ~(int32_t)f);
    }
    for (int64_t b = 0; b < nbnd; ++b) put(m->h_bown[b], -(int32_t)b - 2, (int32_t)(nint + b));
    PBE_CUDA(cudaMalloc((void **)&m->d_imp_other, other.size() * sizeof(int32_t)));
    PBE_CUDA(cudaMalloc((void **)&m->d_imp_face, face.size() * sizeof(int32_t)));
    PBE_CUDA(cudaMemcpy(m->d_imp_other, other.data(), other.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
    PBE_CUDA(cudaMemcpy(m->d_imp_face, face.data(), face.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
    m->imp_W = W;
    return PBE_OK;
}

template <int W>
static void launch_spmv(const ImpParams &q, int mode, dim3 grid, cudaStream_t st) {
    if (mode == 0) imp_spmv_kernel<W, 0><<<grid, kBlock, 0, st>>>(q);
    else imp_spmv_kernel<W, 1><<<grid, kBlock, 0, st>>>(q);
}

int pbe_transport_implicit(pbe_mesh *m, int scheme, int nfields, int64_t stride, const double *psi, const double *phi,
                           const double *psi_b, int64_t stride_b, const double *S, double dt,
                           const pbe_fvm_controls *controls, double *psi_out, pbe_fvm_perf *perf, void *cuda_stream) {
    PBE_REQUIRE(m != nullptr, PBE_E_INVALID, "pbe_transport_implicit: NULL mesh");
    PBE_REQUIRE(nfields >= 1 && psi && phi && psi_out, PBE_E_INVALID, "pbe_transport_implicit: NULL field or nfields < 1");
    PBE_REQUIRE(stride >= m->ncells, PBE_E_INVALID, "pbe_transport_implicit: stride %lld < ncells %lld", (long long)stride,
                (long long)m->ncells);
    PBE_REQUIRE(psi_out != psi && psi_out != S, PBE_E_INVALID, "pbe_transport_implicit: psi_out must not alias psi or S");
    PBE_REQUIRE(psi_b == nullptr || stride_b >= m->nbnd, PBE_E_INVALID, "pbe_transport_implicit: stride_b %lld < boundary faces %lld",
                (long long)stride_b, (long long)m->nbnd);
    PBE_REQUIRE(dt > 0.0, PBE_E_INVALID, "pbe_transport_implicit: dt must be positive");
    PBE_REQUIRE(scheme == PBE_DIV_UPWIND || scheme == PBE_DIV_VANLEER || scheme == PBE_DIV_LIMITEDLINEAR, PBE_E_UNKNOWN_TYPE,
                "pbe_transport_implicit: unknown divergence scheme id %d", scheme);
    const bool decomposed = m->global_ncells > 0; // pbe_mesh_set_halo_plan: the ranks solve ONE system together
    PBE_REQUIRE(m->ncoupled == 0 || decomposed, PBE_E_UNSUPPORTED,
                "pbe_transport_implicit: the mesh has processor patches but no halo plan (pbe_mesh_set_halo_plan)");
    PBE_REQUIRE(!decomposed || m->ctx->comm != nullptr, PBE_E_INVALID,
                "pbe_transport_implicit: decomposed mesh, but the context has no communicator (pbe_comm_init)");
    PBE_REQUIRE(!decomposed || scheme == PBE_DIV_UPWIND, PBE_E_UNSUPPORTED,
                "pbe_transport_implicit: decomposed meshes are solved for Gauss upwind only (the limited schemes' weights "
                "on processor patches need the neighbour's gradient)");
    pbe_fvm_controls ctl = {1.0e-5, 0.0, 1, 1000, 0.0, PBE_FVM_SOLVER_AUTO};
    if (controls) ctl = *controls;
    PBE_REQUIRE(ctl.tolerance >= 0.0 && ctl.rel_tol >= 0.0 && ctl.max_iter >= 1 && ctl.min_iter >= 0, PBE_E_INVALID,
                "pbe_transport_implicit: controls out of range");
    PBE_REQUIRE(ctl.solver >= PBE_FVM_SOLVER_AUTO && ctl.solver <= PBE_FVM_SOLVER_JACOBI, PBE_E_INVALID,
                "pbe_transport_implicit: controls out of range (solver %d)", ctl.solver);
    DeviceGuard g(m->ctx->device);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    int rc = build_slots(m);
    if (rc != PBE_OK) return rc;
    const int64_t nc = m->ncells;
    const int Gs = (int)((nc + kBlock - 1) / kBlock), Gp = (int)((nc + (int64_t)kBlock * kPW - 1) / ((int64_t)kBlock * kPW));
    // Which solver.  Upwind convection makes the matrix triangular along the flow: ANY iteration needs about as many
    // matrix-vector products as the solution travels cells in dt, Jacobi at 4 vector passes per product, BiCGStab at
    // ~9.5.  Measured, 160^3 x 100 fields to 1e-5: Courant 0.4 40 / 69 ms, 4 196 / 326 ms, 40 1 304 / 2 535 ms,
    // 150 1 984 / 5 238 ms (Jacobi / BiCGStab): for the upwind matrix AUTO is Jacobi.
    // A limited scheme gives one matrix per field (weights from the field itself) that is diagonally dominant only at
    // small Courant numbers: AUTO measures the dominance once the weights are known and takes Jacobi below 0.9,
    // BiCGStab above (Jacobi diverges beyond 1).  Without internal faces there is no coefficient to weight.
    const bool lim = scheme != PBE_DIV_UPWIND && m->nint > 0;
    int solver = ctl.solver;
    if (!lim && solver == PBE_FVM_SOLVER_AUTO) solver = PBE_FVM_SOLVER_JACOBI;
    const bool may_bicg = solver != PBE_FVM_SOLVER_JACOBI;
    PBE_REQUIRE(!decomposed || !may_bicg, PBE_E_UNSUPPORTED,
                "pbe_transport_implicit: decomposed meshes are solved by the Jacobi iteration only (solver %d asked)", solver);
    PBE_REQUIRE(!lim || m->imp_W <= 8, PBE_E_UNSUPPORTED,
                "pbe_transport_implicit: limited schemes on cells with more than 8 faces (%d) are not built", m->imp_W);
    PBE_REQUIRE(!lim || (m->d_Sf && m->d_C && m->d_w), PBE_E_INVALID,
                "pbe_transport_implicit: a limited scheme needs the mesh geometry (pbe_mesh_set_geometry)");
    // workspace: Jacobi z, z'; BiCGStab r, r0, p, v, t; for a limited scheme the fields' inverse diagonals, the face
    // weights and a gradient scratch; 4 per-cell rows
    const int KCg = lim ? (int)std::max<size_t>(1, std::min<size_t>((size_t)nfields, ((size_t)1 << 31) / ((size_t)3 * nc * 8) + 1)) : 0;
    const size_t nvec = (may_bicg ? 5 : 2) + (lim ? 1 : 0);
    const size_t need = nvec * nfields * nc + (size_t)4 * nc +
                        (lim ? (size_t)nfields * m->nint + (size_t)KCg * 3 * nc : 0);
    if (m->imp_ws_doubles < need) {
        PBE_CUDA(cudaStreamSynchronize(st));
        cudaFree(m->imp_ws);
        m->imp_ws = nullptr;
        m->imp_ws_doubles = 0;
        cudaError_t e = cudaMalloc((void **)&m->imp_ws, need * sizeof(double));
        if (e != cudaSuccess) {
            cudaGetLastError();
            set_error("pbe_transport_implicit: %zu bytes of solver vectors: %s", need * sizeof(double), cudaGetErrorString(e));
            return PBE_E_NOMEM;
        }
        m->imp_ws_doubles = need;
    }
    const size_t small = (size_t)nfields * sizeof(FieldSc) + (size_t)nfields * kNS * Gs * sizeof(double) + 16;
    static_assert(sizeof(FieldSc) % 8 == 0, "the partial sums follow the per-field scalars");
    if (m->imp_small_bytes < small) {
        PBE_CUDA(cudaStreamSynchronize(st));
        cudaFree(m->imp_small);
        m->imp_small = nullptr;
        m->imp_small_bytes = 0;
        PBE_CUDA(cudaMalloc(&m->imp_small, small));
        m->imp_small_bytes = small;
    }
    if (m->imp_pinned_bytes < (size_t)2 * nfields * sizeof(FieldSc)) { // two slots: the look-ahead below
        if (m->imp_pinned) cudaFreeHost(m->imp_pinned);
        m->imp_pinned = nullptr;
        m->imp_pinned_bytes = 0;
        PBE_CUDA(cudaMallocHost(&m->imp_pinned, (size_t)2 * nfields * sizeof(FieldSc)));
        m->imp_pinned_bytes = (size_t)2 * nfields * sizeof(FieldSc);
    }
    for (int i = 0; i < 2; ++i)
        if (!m->imp_ev[i]) PBE_CUDA(cudaEventCreateWithFlags(&m->imp_ev[i], cudaEventDisableTiming));
    ImpParams q;
    memset(&q, 0, sizeof(q));
    q.nc = nc, q.stride = stride, q.stride_b = stride_b, q.nint = m->nint, q.nbnd = m->nbnd;
    q.W = m->imp_W, q.nfields = nfields, q.G = Gs;
    q.other = m->d_imp_other, q.face = m->d_imp_face;
    q.phi = phi, q.V = m->d_vcell;
    double *w = m->imp_ws;
    const size_t vec = (size_t)nfields * nc;
    double *grad_ws = nullptr, *wface = nullptr;
    if (may_bicg)
        q.r = w, q.r0 = w + vec, q.p = w + 2 * vec, q.v = w + 3 * vec, q.t = w + 4 * vec;
    else
        q.p = w, q.v = w + vec;
    if (lim) q.dinvK = w + (nvec - 1) * vec;
    double *rows = w + nvec * vec;
    if (lim) {
        wface = rows + 4 * nc;
        grad_ws = wface + (size_t)nfields * m->nint;
        q.wface = wface;
    }
    q.diag = rows, q.dinv = rows + nc, q.a1 = rows + 2 * nc, q.drel = rows + 3 * nc;
    q.psi = psi, q.psi_b = psi_b, q.S = S, q.x = psi_out;
    q.sc = (FieldSc *)m->imp_small;
    q.partial = (double *)((char *)m->imp_small + (size_t)nfields * sizeof(FieldSc));
    q.est = (unsigned long long *)(q.partial + (size_t)nfields * kNS * Gs);
    q.nc_global = decomposed ? m->global_ncells : nc;
    q.coupled0 = m->coupled0, q.ncpl = m->ncoupled;
    int64_t launches = 0;
    double *halo_send = nullptr, *halo_stage = nullptr, *halo_recv = nullptr;
    std::vector<int64_t> blk_off, blk_cnt; // the neighbours' blocks of the packed buffers
    if (decomposed) {
        const size_t hneed = (size_t)3 * nfields * std::max<int64_t>(m->ncoupled, 1) + (size_t)nfields * kNS;
        if (m->imp_halo_doubles < hneed) {
            PBE_CUDA(cudaStreamSynchronize(st));
            cudaFree(m->imp_halo);
            m->imp_halo = nullptr;
            m->imp_halo_doubles = 0;
            PBE_CUDA(cudaMalloc((void **)&m->imp_halo, hneed * sizeof(double)));
            m->imp_halo_doubles = hneed;
        }
        halo_send = m->imp_halo;
        halo_stage = halo_send + (size_t)nfields * std::max<int64_t>(m->ncoupled, 1);
        halo_recv = halo_stage + (size_t)nfields * std::max<int64_t>(m->ncoupled, 1);
        q.gsum = halo_recv + (size_t)nfields * std::max<int64_t>(m->ncoupled, 1);
        for (size_t i = 0; i < m->halo_peer.size(); ++i) {
            blk_off.push_back(m->halo_off[i] * nfields);
            blk_cnt.push_back(m->halo_cnt[i] * nfields);
        }
    }
    // this rank's values along its processor patches -> the neighbours' halo rows, and theirs into ours
    auto exchange = [&](const double *field, int64_t fstride) -> int {
        if (!decomposed || m->ncoupled == 0) return PBE_OK;
        const dim3 hg((unsigned)((m->ncoupled + 255) / 256), (unsigned)std::min(nfields, 16));
        imp_halo_pack_kernel<<<hg, 256, 0, st>>>(nfields, fstride, field, m->ncoupled, m->d_send_cells, m->d_halo_qoff,
                                                 m->d_halo_qcnt, halo_send);
        // one ncclSend / ncclRecv per neighbour (a "field" of pbe_halo_exchange is then the neighbour's whole block)
        const int e = pbe_halo_exchange(m->ctx, (int)m->halo_peer.size(), m->halo_peer.data(), blk_off.data(), blk_cnt.data(), 1,
                                        halo_send, 0, halo_stage, 0, st);
        if (e != PBE_OK) return e;
        imp_halo_unpack_kernel<<<hg, 256, 0, st>>>(nfields, m->ncoupled, m->d_halo_qoff, m->d_halo_qcnt, halo_stage, halo_recv);
        launches += 2;
        return PBE_OK;
    };
    q.rdt = 1.0 / dt, q.relax = ctl.relaxation, q.tol = ctl.tolerance, q.rel_tol = ctl.rel_tol;
    q.min_iter = ctl.min_iter, q.max_iter = ctl.max_iter;
    // fields per CTA of the row-holding kernels: enough CTAs to fill the machine, rows amortised over many fields
    static const int env_chunk = [] { const char *e = getenv("PBE_IMPLICIT_CHUNK"); return e ? atoi(e) : 0; }();
    // 160^3 x 100, upwind sweep: 32 -> 3.67, 50 / 64 -> 3.40, 100 -> 3.42 ms; the limited sweeps (more registers) like 32
    int chunk = env_chunk > 0 ? env_chunk : (lim ? 32 : 50);
    while (chunk > 4 && (int64_t)Gs * ((nfields + chunk - 1) / chunk) < 4 * (int64_t)m->ctx->sm_count) chunk /= 2;
    q.chunk = std::min(std::min(chunk, kMaxChunk), nfields);
    const dim3 grid_s((unsigned)Gs, (unsigned)((nfields + q.chunk - 1) / q.chunk)), grid_p((unsigned)Gp, (unsigned)nfields);
    const int W = q.W <= 8 ? q.W : 0;
    int frc = PBE_OK;
    auto finalize = [&](int phase, int G) {
        if (!decomposed) {
            imp_finalize_kernel<<<nfields, kBlock, 0, st>>>(q, phase, G, 0);
            ++launches;
            return;
        }
        imp_finalize_kernel<<<nfields, kBlock, 0, st>>>(q, phase, G, 1);
        const int e = pbe_reduce_diagnostics(m->ctx, nfields * kNS, q.gsum, 0, nullptr, 0, nullptr, st); // gSum
        if (e != PBE_OK && frc == PBE_OK) frc = e;
        imp_finalize_kernel<<<nfields, kBlock, 0, st>>>(q, phase, G, 2);
        launches += 2;
    };
    auto spmv = [&](int mode) {
        if (q.wface) {
            switch (W) {
            case 4: mode == 0 ? imp_lim_spmv_kernel<4, 0><<<grid_s, kBlock, 0, st>>>(q) : imp_lim_spmv_kernel<4, 1><<<grid_s, kBlock, 0, st>>>(q); break;
            case 6: mode == 0 ? imp_lim_spmv_kernel<6, 0><<<grid_s, kBlock, 0, st>>>(q) : imp_lim_spmv_kernel<6, 1><<<grid_s, kBlock, 0, st>>>(q); break;
            default: mode == 0 ? imp_lim_spmv_kernel<8, 0><<<grid_s, kBlock, 0, st>>>(q) : imp_lim_spmv_kernel<8, 1><<<grid_s, kBlock, 0, st>>>(q); break;
            }
            ++launches;
            return;
        }
        switch (W) {
        case 4: launch_spmv<4>(q, mode, grid_s, st); break;
        case 6: launch_spmv<6>(q, mode, grid_s, st); break;
        case 8: launch_spmv<8>(q, mode, grid_s, st); break;
        default: launch_spmv<0>(q, mode, grid_s, st); break;
        }
        ++launches;
    };
    FieldSc *host = (FieldSc *)m->imp_pinned;
    auto active_fields = [&](int *n) -> int {
        PBE_CUDA(cudaMemcpyAsync(host, q.sc, (size_t)nfields * sizeof(FieldSc), cudaMemcpyDeviceToHost, st));
        PBE_CUDA(cudaStreamSynchronize(st));
        *n = 0;
        for (int k = 0; k < nfields; ++k) *n += host[k].state == 0;
        return PBE_OK;
    };

#define PBE_IMP_W(KERNEL, ...)                                               \
    switch (W) {                                                             \
    case 4: KERNEL<4 __VA_ARGS__><<<grid_s, kBlock, 0, st>>>(q); break;      \
    case 6: KERNEL<6 __VA_ARGS__><<<grid_s, kBlock, 0, st>>>(q); break;      \
    case 8: KERNEL<8 __VA_ARGS__><<<grid_s, kBlock, 0, st>>>(q); break;      \
    default: KERNEL<0 __VA_ARGS__><<<grid_s, kBlock, 0, st>>>(q); break;     \
    }
#define PBE_IMP_WL(KERNEL)                                                   \
    switch (W) {                                                             \
    case 4: KERNEL<4><<<grid_s, kBlock, 0, st>>>(q); break;                  \
    case 6: KERNEL<6><<<grid_s, kBlock, 0, st>>>(q); break;                  \
    default: KERNEL<8><<<grid_s, kBlock, 0, st>>>(q); break;                 \
    }
#define PBE_IMP_W8(KERNEL, ...)                                              \
    switch (W) {                                                             \
    case 4: KERNEL<4 __VA_ARGS__><<<grid_s, kBlock, 0, st>>>(q); break;      \
    case 6: KERNEL<6 __VA_ARGS__><<<grid_s, kBlock, 0, st>>>(q); break;      \
    default: KERNEL<8 __VA_ARGS__><<<grid_s, kBlock, 0, st>>>(q); break;     \
    }
#define PBE_IMP_COMMA ,
    if (lim) {
        // gradient and face weights of all fields, KCg fields at a time (the gradient scratch holds one chunk)
        q.Sf = m->d_Sf, q.Cc = m->d_C, q.wlin = m->d_w, q.owner = m->d_owner, q.neigh = m->d_neigh;
        q.grad = grad_ws, q.wout = wface, q.scheme = scheme;
        const int chunk_keep = q.chunk;
        for (int k0 = 0; k0 < nfields; k0 += KCg) {
            q.k0 = k0, q.kn = std::min(KCg, nfields - k0);
            q.chunk = std::min(chunk_keep, q.kn);
            const unsigned gy = (unsigned)((q.kn + q.chunk - 1) / q.chunk);
            switch (W) {
            case 4: imp_lim_grad_kernel<4><<<dim3((unsigned)Gs, gy), kBlock, 0, st>>>(q); break;
            case 6: imp_lim_grad_kernel<6><<<dim3((unsigned)Gs, gy), kBlock, 0, st>>>(q); break;
            default: imp_lim_grad_kernel<8><<<dim3((unsigned)Gs, gy), kBlock, 0, st>>>(q); break;
            }
            imp_lim_weights_kernel<<<dim3((unsigned)((m->nint + kBlock - 1) / kBlock), gy), kBlock, 0, st>>>(q);
            launches += 2;
        }
        q.chunk = chunk_keep;
        if (solver == PBE_FVM_SOLVER_AUTO) {
            double est = 0.0;
            PBE_CUDA(cudaMemsetAsync(q.est, 0, sizeof(unsigned long long), st));
            PBE_IMP_WL(imp_lim_est_kernel)
            ++launches;
            PBE_CUDA(cudaMemcpyAsync(&est, q.est, sizeof(double), cudaMemcpyDeviceToHost, st));
            PBE_CUDA(cudaStreamSynchronize(st));
            solver = est <= 0.9 ? PBE_FVM_SOLVER_JACOBI : PBE_FVM_SOLVER_BICGSTAB;
        }
    } else {
        imp_diag_kernel<<<Gs, kBlock, 0, st>>>(q);
        ++launches;
    }
    const bool jac = solver == PBE_FVM_SOLVER_JACOBI;
    imp_xsum_kernel<<<grid_p, kBlock, 0, st>>>(q);
    ++launches;
    finalize(PH_XSUM, Gp);
    rc = exchange(psi, stride); // psi0 across the processor patches, for b - A x0
    if (rc != PBE_OK) return rc;
    q.halo_x = (decomposed && m->ncoupled) ? halo_recv : nullptr;
    if (lim && jac) {
        PBE_IMP_W8(imp_lim_init_kernel, PBE_IMP_COMMA true)
    } else if (lim) {
        PBE_IMP_W8(imp_lim_init_kernel, PBE_IMP_COMMA false)
    } else if (jac) {
        PBE_IMP_W(imp_init_kernel, PBE_IMP_COMMA true)
    } else {
        PBE_IMP_W(imp_init_kernel, PBE_IMP_COMMA false)
    }
    ++launches;
    finalize(PH_INIT, Gs);
    int nactive = 0;
    rc = active_fields(&nactive);
    if (jac) {
        // One iteration ahead: iteration k + 1 is queued before the host looks at the states iteration k left, so the
        // device never waits for the host.  Converged fields are frozen on the device (their state is not 0), hence the
        // one iteration queued past the end changes nothing.
        auto issue = [&](int slot) -> int {
            const int e = exchange(q.p, nc); // z across the processor patches
            if (e != PBE_OK) return e;
            q.halo_z = (decomposed && m->ncoupled) ? halo_recv : nullptr;
            if (lim) {
                const size_t smem = (size_t)2 * 2 * W * kBlock * sizeof(double);
                const void *fn = W == 4 ? (const void *)imp_lim_jacobi_kernel<4>
                                        : W == 6 ? (const void *)imp_lim_jacobi_kernel<6> : (const void *)imp_lim_jacobi_kernel<8>;
                PBE_CUDA(raise_dyn_smem(m->ctx, fn, smem));
                switch (W) {
                case 4: imp_lim_jacobi_kernel<4><<<grid_s, kBlock, smem, st>>>(q); break;
                case 6: imp_lim_jacobi_kernel<6><<<grid_s, kBlock, smem, st>>>(q); break;
                default: imp_lim_jacobi_kernel<8><<<grid_s, kBlock, smem, st>>>(q); break;
                }
            } else if (q.halo_z) {
                PBE_IMP_W(imp_jacobi_kernel, PBE_IMP_COMMA true)
            } else {
                PBE_IMP_W(imp_jacobi_kernel, PBE_IMP_COMMA false)
            }
            ++launches;
            finalize(PH_J, Gs);
            std::swap(q.p, q.v);
            PBE_CUDA(cudaMemcpyAsync(host + (size_t)slot * nfields, q.sc, (size_t)nfields * sizeof(FieldSc),
                                     cudaMemcpyDeviceToHost, st));
            PBE_CUDA(cudaEventRecord(m->imp_ev[slot], st));
            return PBE_OK;
        };
        int slot = 0;
        if (rc == PBE_OK && nactive > 0) rc = issue(0);
        while (rc == PBE_OK && nactive > 0) {
            rc = issue(slot ^ 1);
            if (rc != PBE_OK) break;
            PBE_CUDA(cudaEventSynchronize(m->imp_ev[slot]));
            nactive = 0;
            for (int k = 0; k < nfields; ++k) nactive += host[(size_t)slot * nfields + k].state == 0;
            if (nactive == 0) break;
            slot ^= 1;
        }
        PBE_CUDA(cudaStreamSynchronize(st));
        if (slot != 0) memcpy(host, host + (size_t)slot * nfields, (size_t)nfields * sizeof(FieldSc));
    }
    while (!jac && rc == PBE_OK && nactive > 0) {
        spmv(0);
        finalize(PH_A, Gs);
        imp_s_kernel<<<grid_p, kBlock, 0, st>>>(q);
        ++launches;
        finalize(PH_B, Gp);
        spmv(1);
        finalize(PH_C, Gs);
        imp_update_kernel<<<grid_p, kBlock, 0, st>>>(q);
        ++launches;
        finalize(PH_D, Gp);
        imp_p_kernel<<<grid_p, kBlock, 0, st>>>(q);
        ++launches;
        rc = active_fields(&nactive);
    }
#undef PBE_IMP_W
#undef PBE_IMP_COMMA
#undef PBE_IMP_WL
#undef PBE_IMP_W8
    m->ctx->launches += launches;
    if (rc == PBE_OK) rc = frc;
    if (rc != PBE_OK) return rc;
    PBE_CUDA(cudaGetLastError());
    if (perf)
        for (int k = 0; k < nfields; ++k)
            perf[k] = pbe_fvm_perf{host[k].res0, host[k].res, host[k].iters, host[k].state};
    for (int k = 0; k < nfields; ++k) // a field left worse than it came in is an error, not a result
        PBE_REQUIRE(host[k].state != 2 || host[k].res <= host[k].res0, PBE_E_UNSUPPORTED,
                    "pbe_transport_implicit: field %d diverged (%s, residual %g -> %g after %d iterations): the matrix is "
                    "not diagonally dominant at this time step; ask for PBE_FVM_SOLVER_BICGSTAB or reduce deltaT",
                    k, jac ? "Jacobi" : "BiCGStab", host[k].res0, host[k].res, host[k].iters);
    return PBE_OK;
}

extern "C" int pbe_mesh_set_halo_plan(pbe_mesh *m, int nneigh, const int *peer, const int64_t *off, const int64_t *cnt,
                                      const int32_t *send_cells, int64_t global_ncells) {
    PBE_REQUIRE(m != nullptr && nneigh >= 0 && global_ncells >= m->ncells, PBE_E_INVALID,
                "pbe_mesh_set_halo_plan: bad argument (global cells %lld, local %lld)", (long long)global_ncells,
                (long long)(m ? m->ncells : 0));
    PBE_REQUIRE(nneigh == 0 || (peer && off && cnt), PBE_E_INVALID, "pbe_mesh_set_halo_plan: NULL plan arrays");
    PBE_REQUIRE(m->ncoupled == 0 || send_cells != nullptr, PBE_E_INVALID, "pbe_mesh_set_halo_plan: NULL send_cells");
    int64_t covered = 0;
    for (int q = 0; q < nneigh; ++q) {
        PBE_REQUIRE(off[q] >= 0 && cnt[q] >= 0 && off[q] + cnt[q] <= m->ncoupled, PBE_E_INVALID,
                    "pbe_mesh_set_halo_plan: neighbour %d covers faces [%lld, %lld) of %lld coupled faces", q, (long long)off[q],
                    (long long)(off[q] + cnt[q]), (long long)m->ncoupled);
        covered += cnt[q];
    }
    PBE_REQUIRE(covered == m->ncoupled, PBE_E_INVALID,
                "pbe_mesh_set_halo_plan: the neighbours cover %lld of the %lld coupled faces (pbe_mesh_set_coupled first)",
                (long long)covered, (long long)m->ncoupled);
    for (int64_t j = 0; j < m->ncoupled; ++j)
        PBE_REQUIRE(send_cells[j] >= 0 && send_cells[j] < m->ncells, PBE_E_INVALID, "pbe_mesh_set_halo_plan: send cell out of range");
    DeviceGuard g(m->ctx->device);
    cudaFree(m->d_send_cells);
    m->d_send_cells = nullptr;
    if (m->ncoupled) {
        PBE_CUDA(cudaMalloc((void **)&m->d_send_cells, (size_t)m->ncoupled * sizeof(int32_t)));
        PBE_CUDA(cudaMemcpy(m->d_send_cells, send_cells, (size_t)m->ncoupled * sizeof(int32_t), cudaMemcpyHostToDevice));
    }
    cudaFree(m->d_halo_qoff);
    cudaFree(m->d_halo_qcnt);
    m->d_halo_qoff = m->d_halo_qcnt = nullptr;
    if (m->ncoupled) {
        std::vector<int64_t> qo((size_t)m->ncoupled, -1), qc((size_t)m->ncoupled, 0);
        for (int q = 0; q < nneigh; ++q)
            for (int64_t j = off[q]; j < off[q] + cnt[q]; ++j) {
                PBE_REQUIRE(qo[(size_t)j] < 0, PBE_E_INVALID, "pbe_mesh_set_halo_plan: coupled face %lld belongs to two neighbours",
                            (long long)j);
                qo[(size_t)j] = off[q], qc[(size_t)j] = cnt[q];
            }
        PBE_CUDA(cudaMalloc((void **)&m->d_halo_qoff, qo.size() * sizeof(int64_t)));
        PBE_CUDA(cudaMalloc((void **)&m->d_halo_qcnt, qc.size() * sizeof(int64_t)));
        PBE_CUDA(cudaMemcpy(m->d_halo_qoff, qo.data(), qo.size() * sizeof(int64_t), cudaMemcpyHostToDevice));
        PBE_CUDA(cudaMemcpy(m->d_halo_qcnt, qc.data(), qc.size() * sizeof(int64_t), cudaMemcpyHostToDevice));
    }
    m->halo_peer.assign(peer, peer + nneigh);
    m->halo_off.assign(off, off + nneigh);
    m->halo_cnt.assign(cnt, cnt + nneigh);
    m->global_ncells = global_ncells;
    return PBE_OK;
}
