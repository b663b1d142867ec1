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
// accumulators are columns of shared-memory arrays.  The pair list is cell-independent and cut into TILES of
// up to four pairs (i, j0 .. j0+3) whose births all land inside a window of MOCP_WIN consecutive pivots
// starting at tb (for a geometric grid the targets of four consecutive j span at most five pivots; any grid
// works: a tile is simply cut shorter when its window would overflow).  A tile is read through uniform
// (broadcast) loads; its four kernel values are independent chains (the exponentials overlap), the births are
// formed in a register window as weighted sums (weights = the two-pivot shares, zero elsewhere) and retired
// with one batch of independent shared-memory read-modify-writes, the deaths of the four j likewise; the death
// of class i is a register sum flushed when i changes.  (First version: one pair at a time, four dependent
// read-modify-writes each: 78.8 ms per 1.2e6 cells at N = 100)
#pragma once

#include "pbe_moc_kernel.cuh"

namespace pbe {

#ifndef PBE_MOCP_UNROLL
#define PBE_MOCP_UNROLL 2
#endif
constexpr int MOCP_UNROLL = PBE_MOCP_UNROLL; // tiles in flight per trip of the slot loop
constexpr int MOCP_TILE = 4;        // pairs per tile
constexpr int MOCP_WIN = 8;         // birth window, pivots
constexpr int MOCP_SLOT_TILES = 16; // tiles per ring slot: the tile list streams through two shared-memory slots
                                    // filled by cp.async.bulk (UBLKCP) on mbarriers, as the pair table of the
                                    // uniform-grid kernels does (read straight from global memory a tile cost
                                    // 42 L2 round trips at the head of its dependency chain: 111 ms)

struct alignas(16) FpTile {
    int i, j0, tb, npairs;
    double a[MOCP_TILE];  // CT: G_ij (halved on the diagonal); otherwise S_ij (halved on the diagonal); 0 = padding
    double b[MOCP_TILE];  // CT: H_ij
    double cw[MOCP_TILE][MOCP_WIN]; // share of pair k's birth that goes to pivot tb + w
};

struct MocFpExtra {
    const FpTile *tiles; // ntiles entries, a multiple of MOCP_SLOT_TILES (padded with empty tiles)
    int ntiles;
    const double *hh;    // [N]: (x_{i+1} - x_{i-1}) / 2, x_{-1} = 0: hat area of pivot i
    const double *cself; // [N]: (x_i - x_{i-1}) / x_i: daughters of class i that stay in class i
    const double *wk;    // [N]: 2 / x_k: beta(., x_k) of uniformBinaryBreakup.C:55-60
};

constexpr int MOCP_BLOCK = 128;

__host__ __device__ inline size_t mocp_smem_bytes(int N) {
    return ((size_t)(2 * N + 2) * MOCP_BLOCK + (size_t)(CLS_ROWS + 3) * (N + 1)) * sizeof(double) +
           2 * MOCP_SLOT_TILES * sizeof(FpTile) + 2 * sizeof(uint64_t) + 128 * sizeof(double);
}

template <int COAL, int BR>
__global__ void __launch_bounds__(MOCP_BLOCK, 1) moc_pivot_kernel(const __grid_constant__ MocParams p,
                                                                const __grid_constant__ MocFpExtra x) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int BLOCK = MOCP_BLOCK;
    const int N = p.N;
    const int tid = threadIdx.x;
    FpTile *ring = reinterpret_cast<FpTile *>(smem_raw);    // [2][MOCP_SLOT_TILES]
    uint64_t *bars = reinterpret_cast<uint64_t *>(ring + 2 * MOCP_SLOT_TILES);
    double *T128 = reinterpret_cast<double *>(bars + 2);    // 2^(j/128) table of exp2_neg_tab_safe (pbe_math.cuh)
    double *n_s = T128 + 128;                               // [N + 1][BLOCK]: row N = 0, the class of padding pairs
    double *src_s = n_s + (size_t)(N + 1) * BLOCK;          // [N + 1][BLOCK]: row N absorbs window slots past the grid
    double *cls_s = src_s + (size_t)(N + 1) * BLOCK;        // CLS_ROWS rows of the common tables, then hh, cself, wk
    volatile double *sv = src_s;
    if (tid == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        mbar_fence_init();
    }
    static_assert(MOCP_BLOCK == 128, "one table entry per thread");
    T128[tid] = PBE_EXP2_TAB128[tid];
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
    const int nslots = (COAL == DCOAL_OFF) ? 0 : x.ntiles / MOCP_SLOT_TILES; // ring fills per batch
    const int64_t my_batches =
        (nbatches > (int64_t)blockIdx.x) ? (nbatches - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const int64_t total_q = my_batches * nslots;
    auto issue = [&](int64_t qq) {
        const int sl = (int)(qq % nslots);
        uint64_t *bar = &bars[qq & 1];
        const unsigned bytes = (unsigned)(MOCP_SLOT_TILES * sizeof(FpTile));
        mbar_expect_tx(bar, bytes);
        bulk_g2s(ring + (qq & 1) * MOCP_SLOT_TILES, x.tiles + (size_t)sl * MOCP_SLOT_TILES, bytes, bar);
    };
    if (tid == 0) {
        if (0 < total_q) issue(0);
        if (1 < total_q) issue(1);
    }
    int64_t qq = 0;
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
            n_s[N * BLOCK + tid] = 0.0;
        }
        const double eps = __ldg(p.eps + cc), rho = __ldg(p.rho + cc), nu = __ldg(p.nu + cc);
        const double eps13 = cbrt(eps);
        double A = 1.0, negK = 0.0;
        if (COAL == DCOAL_CT) {
            A = p.coalA * eps13;
            negK = -(p.coalK * rho * eps * (rho * nu)) * 1.4426950408889634074; // x log2(e): exp(x) = 2^(x log2 e)
        }
        if (COAL != DCOAL_OFF) {
            int irow = -1;
            double ni = 0.0, di = 0.0;
#pragma unroll 1
            for (int sl = 0; sl < nslots; ++sl, ++qq) {
              mbar_wait(&bars[qq & 1], (unsigned)((qq >> 1) & 1));
              const FpTile *slot = ring + (qq & 1) * MOCP_SLOT_TILES;
#pragma unroll MOCP_UNROLL
              for (int q = 0; q < MOCP_SLOT_TILES; ++q) {
                const FpTile &T = slot[q]; // uniform shared-memory address: broadcast loads
                const int ti = T.i, j0 = T.j0, tb = T.tb;
                if (ti != irow) { // warp-uniform
                    if (irow >= 0) sv[irow * BLOCK + tid] -= di;
                    irow = ti;
                    ni = n_s[irow * BLOCK + tid];
                    di = 0.0;
                }
                double tt[MOCP_TILE];
                int jrow[MOCP_TILE];
#pragma unroll
                for (int k = 0; k < MOCP_TILE; ++k) {
                    jrow[k] = min(j0 + k, N) * BLOCK + tid; // padding pairs: the zero class
                    const double w = ni * n_s[jrow[k]];
                    if (COAL == DCOAL_CT)
                        tt[k] = (exp2_neg_tab_safe(negK * T.b[k], T128) * T.a[k]) * w; // 8 FP64 instructions (exp_neg: ~25)
                    else
                        tt[k] = T.a[k] * w;
                }
                // deaths of the four j: independent targets (padding pairs all hit the dummy row with tt = 0)
                double dj[MOCP_TILE];
#pragma unroll
                for (int k = 0; k < MOCP_TILE; ++k) dj[k] = sv[jrow[k]];
#pragma unroll
                for (int k = 0; k < MOCP_TILE; ++k) {
                    di += tt[k];
                    sv[jrow[k]] = dj[k] - tt[k];
                }
                // births: weighted sums into the window tb .. tb + WIN - 1, then one batch of read-modify-writes
                double bw[MOCP_WIN];
#pragma unroll
                for (int w = 0; w < MOCP_WIN; ++w) {
                    double acc = T.cw[0][w] * tt[0];
#pragma unroll
                    for (int k = 1; k < MOCP_TILE; ++k) acc = fma_(T.cw[k][w], tt[k], acc);
                    bw[w] = acc;
                }
                double cur[MOCP_WIN];
                int wrow[MOCP_WIN];
#pragma unroll
                for (int w = 0; w < MOCP_WIN; ++w) {
                    wrow[w] = min(tb + w, N) * BLOCK + tid; // past the grid: weights are 0, the dummy row takes them
                    cur[w] = sv[wrow[w]];
                }
#pragma unroll
                for (int w = 0; w < MOCP_WIN; ++w) sv[wrow[w]] = cur[w] + bw[w];
              }
              __syncthreads(); // every warp is done with this ring slot
              if (tid == 0 && qq + 2 < total_q) issue(qq + 2);
            }
            if (irow >= 0) sv[irow * BLOCK + tid] -= di;
        }

        // ---- epilogue: breakup through the pivots' hat functions, limiter, update, d, diagnostics ----
        double Ab = 0.0, negKb = 0.0;
        if (BR == DBR_CT) {
            Ab = p.brA * eps13;
            negKb = -(p.brK / (rho * (eps13 * eps13))) * 1.4426950408889634074;
        }
        double R = 0.0, sum_d = 0.0, sum_n = 0.0;
        double mloc[MOC_MAX_MOMENTS];
#pragma unroll
        for (int k = 0; k < MOC_MAX_MOMENTS; ++k) mloc[k] = 0.0;
#pragma unroll 2
        for (int i = N - 1; i >= 0; --i) {
            const double ni = n_s[i * BLOCK + tid];
            double g = cBP[i];
            if (BR == DBR_CT) g = (Ab * g) * exp2_neg_tab_safe(negKb * cBQ[i], T128);
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
