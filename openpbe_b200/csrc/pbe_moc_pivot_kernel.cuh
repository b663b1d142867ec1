// pbe_moc_pivot_kernel.cuh -- method of classes on an ARBITRARY class grid (fixed-pivot technique of
// Kumar & Ramkrishna 1996, the paper the reference names at src/PBESystem/MOC.C:147-148).
//
// SURVEY 8 row f4, an EXTENSION with parity unpinned: the reference only has the uniform grid
// xi_i = (i + 1) xi1, where the coalescence of classes i and j lands exactly on class i + j + 1
// (MOC.C:115-121) -- the structure the two kernels of pbe_moc_kernel.cuh / pbe_moc_tmem_kernel.cuh
// are built around (anti-diagonal birth windows).  On a geometric grid the new drop x_i + x_j falls
// BETWEEN two pivots and is split over them so that number and volume are preserved; the target
// class is a table entry instead of an index rule, so the births are scattered through shared
// memory.  Conventions are the reference's (death bound of MOC.C:107-109, the 0.5 of MOC.C:118 on
// the diagonal, limiter of MOC.C:87-91); formulas and the CPU restatement the tests compare with:
// tests/test_fixed_pivot.py.
//
// One thread per cell (lane = cell) as in the other kernels; per-cell class densities and source
// accumulators are columns of shared-memory arrays; the pair list is cell-independent and read
// through uniform (broadcast) loads in i-major order, so the death of class i is a register sum
// flushed when i changes.
#pragma once

#include "pbe_moc_kernel.cuh"

namespace pbe {

struct FpPair {
    double a, b; // CT: G_ij (halved on the diagonal), H_ij; otherwise S_ij (halved on the diagonal), unused
    double eta;  // share of the birth that goes to pivot t (the rest to t + 1)
    int i, j, t, pad;
};

struct MocFpExtra {
    const FpPair *pairs;
    int npairs;
    const double *hh;    // [N]: (x_{i+1} - x_{i-1}) / 2, x_{-1} = 0: hat area of pivot i
    const double *cself; // [N]: (x_i - x_{i-1}) / x_i: daughters of class i that stay in class i
    const double *wk;    // [N]: 2 / x_k: beta(., x_k) of uniformBinaryBreakup.C:55-60
};

constexpr int MOCP_BLOCK = 128;

__host__ __device__ inline size_t mocp_smem_bytes(int N) {
    return ((size_t)(2 * N + 1) * MOCP_BLOCK + (size_t)(CLS_ROWS + 3) * (N + 1)) * sizeof(double);
}

template <int COAL, int BR>
__global__ void __launch_bounds__(MOCP_BLOCK) moc_pivot_kernel(const __grid_constant__ MocParams p,
                                                                const __grid_constant__ MocFpExtra x) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int BLOCK = MOCP_BLOCK;
    const int N = p.N;
    const int tid = threadIdx.x;
    double *n_s = reinterpret_cast<double *>(smem_raw);     // [N][BLOCK]
    double *src_s = n_s + (size_t)N * BLOCK;                // [N + 1][BLOCK]: row N takes the (1 - eta) = 0 share of v == x_{N-1}
    double *cls_s = src_s + (size_t)(N + 1) * BLOCK;        // CLS_ROWS rows of the common tables, then hh, cself, wk
    volatile double *sv = src_s;
    for (int k = tid; k < CLS_ROWS * (N + 1); k += BLOCK) cls_s[k] = p.cls[k];
    double *hh_s = cls_s + CLS_ROWS * (N + 1), *cs_s = hh_s + (N + 1), *wk_s = cs_s + (N + 1);
    for (int k = tid; k < N; k += BLOCK) {
        hh_s[k] = x.hh[k];
        cs_s[k] = x.cself[k];
        wk_s[k] = x.wk[k];
    }
    __syncthreads();
    const double *cBP = cls_s + CLS_BR_P * (N + 1);
    const double *cBQ = cls_s + CLS_BR_Q * (N + 1);
    const double *cDV = cls_s + CLS_DVOL * (N + 1);
    const double *cXI = cls_s + CLS_XI * (N + 1);

    double macc[MOC_MAX_MOMENTS];
#pragma unroll
    for (int k = 0; k < MOC_MAX_MOMENTS; ++k) macc[k] = 0.0;
    double cerr = 0.0;
    const bool want_diag = (p.moments != nullptr) || (p.cont_err_max != nullptr);

    const int64_t nbatches = (p.ncells + BLOCK - 1) / BLOCK;
    for (int64_t batch = blockIdx.x; batch < nbatches; batch += gridDim.x) {
        const int64_t cell = batch * BLOCK + tid;
        const bool live = cell < p.ncells;
        const int64_t cc = live ? cell : p.ncells - 1;
        {
            const double *src = p.n_in + cc;
#pragma unroll 4
            for (int i = 0; i < N; ++i) {
                n_s[i * BLOCK + tid] = __ldg(src + (size_t)i * p.stride);
                sv[i * BLOCK + tid] = 0.0;
            }
            sv[N * BLOCK + tid] = 0.0;
        }
        const double eps = __ldg(p.eps + cc), rho = __ldg(p.rho + cc), nu = __ldg(p.nu + cc);
        const double eps13 = cbrt(eps);
        double A = 1.0, negK = 0.0;
        if (COAL == DCOAL_CT) {
            A = p.coalA * eps13;
            negK = -(p.coalK * rho * eps * (rho * nu));
        }
        if (COAL != DCOAL_OFF) {
            int irow = -1;
            double ni = 0.0, di = 0.0;
#pragma unroll 2
            for (int q = 0; q < x.npairs; ++q) {
                const FpPair pr = x.pairs[q]; // uniform address: one broadcast load per warp
                if (pr.i != irow) {           // warp-uniform
                    if (irow >= 0) sv[irow * BLOCK + tid] -= di;
                    irow = pr.i;
                    ni = n_s[irow * BLOCK + tid];
                    di = 0.0;
                }
                const double w = ni * n_s[pr.j * BLOCK + tid];
                double tt;
                if (COAL == DCOAL_CT)
                    tt = (exp_neg(negK * pr.b) * pr.a) * w; // the operation order of moc_tile (pbe_moc_kernel.cuh)
                else
                    tt = pr.a * w;
                di += tt;
                sv[pr.j * BLOCK + tid] -= tt;
                const double up = pr.eta * tt;
                sv[pr.t * BLOCK + tid] += up;
                sv[(pr.t + 1) * BLOCK + tid] += tt - up;
            }
            if (irow >= 0) sv[irow * BLOCK + tid] -= di;
        }

        // ---- epilogue: breakup through the pivots' hat functions, limiter, update, d, diagnostics ----
        double Ab = 0.0, negKb = 0.0;
        if (BR == DBR_CT) {
            Ab = p.brA * eps13;
            negKb = -(p.brK / (rho * (eps13 * eps13)));
        }
        double R = 0.0, sum_d = 0.0, sum_n = 0.0;
        double mloc[MOC_MAX_MOMENTS];
#pragma unroll
        for (int k = 0; k < MOC_MAX_MOMENTS; ++k) mloc[k] = 0.0;
#pragma unroll 2
        for (int i = N - 1; i >= 0; --i) {
            const double ni = n_s[i * BLOCK + tid];
            double g = cBP[i];
            if (BR == DBR_CT) g = (Ab * g) * exp_neg(negKb * cBQ[i]);
            const double gn = g * ni;
            // daughters of the classes above (suffix sum R) and of class i itself, minus its death
            const double B = fma_(hh_s[i], R, (cs_s[i] - 1.0) * gn);
            R = fma_(wk_s[i], gn, R);
            const double C = A * sv[i * BLOCK + tid];
            const double lim = __dmul_rn(-p.relax, ni) / p.dt; // MOC.C:87-91
            double s = C + B;
            s = (s > lim) ? s : lim;
            if (p.S_out != nullptr && live) p.S_out[(size_t)i * p.stride + cell] = s;
            double nn = ni;
            if (p.do_step) {
                nn = __dadd_rn(ni, __dmul_rn(p.dt, s));
                if (live) p.n_out[(size_t)i * p.stride + cell] = nn;
            }
            sum_d = fma_(cDV[i], nn, sum_d);
            sum_n += nn;
            if (want_diag) {
                const double xv = cXI[i];
                double pw = nn;
#pragma unroll
                for (int k = 0; k < MOC_MAX_MOMENTS; ++k) {
                    if (k < p.n_moments || k == 1) mloc[k] += pw;
                    pw *= xv;
                }
            }
        }
        if (p.d_out != nullptr && live) p.d_out[cell] = sum_d / sum_n;
        if (want_diag && live) {
            const double V = (p.vcell != nullptr) ? __ldg(p.vcell + cell) : 1.0;
#pragma unroll
            for (int k = 0; k < MOC_MAX_MOMENTS; ++k) macc[k] = fma_(V, mloc[k], macc[k]);
            if (p.cont_err_max != nullptr && p.alpha != nullptr) {
                const double e = fabs(mloc[1] / V - __ldg(p.alpha + cell));
                cerr = (e > cerr) ? e : cerr;
            }
        }
    }

    if (want_diag) {
        __syncthreads();
        double *red = n_s;
        for (int k = 0; k < MOC_MAX_MOMENTS + 1; ++k) {
            const bool is_err = (k == MOC_MAX_MOMENTS);
            if (!is_err && (p.moments == nullptr || k >= p.n_moments)) continue;
            if (is_err && p.cont_err_max == nullptr) continue;
            double v = is_err ? cerr : macc[k];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const double u = __shfl_xor_sync(0xffffffffu, v, o);
                v = is_err ? ((u > v) ? u : v) : (v + u);
            }
            if ((tid & 31) == 0) red[tid >> 5] = v;
            __syncthreads();
            if (tid == 0) {
                double t = red[0];
                for (int w = 1; w < BLOCK / 32; ++w) t = is_err ? ((red[w] > t) ? red[w] : t) : t + red[w];
                if (is_err)
                    atomic_max_nonneg(p.cont_err_max, t);
                else
                    atomicAdd(p.moments + k, t);
            }
            __syncthreads();
        }
    }
}

} // namespace pbe
