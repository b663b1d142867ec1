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
// MOCP_TI x MOCP_TJ class pairs (rows i0 .. i0+3, columns j0 .. j0+3, as the uniform-grid kernels walk them) whose
// births all land inside a window of MOCP_WIN consecutive pivots starting at tb: on the uniform grid a 4 x 4 tile
// spans 7 pivots, on a geometric grid 4-5; any increasing grid works, because a tile whose window would overflow is
// cut into row tiles and those into single pairs.  Invalid slots (i > j, sum beyond the grid, cut away) carry a
// zero kernel value and zero weights, so the loop has no predicates.  A tile is read through uniform (broadcast)
// shared-memory loads from a ring that cp.async.bulk fills; its 16 kernel values are independent chains, row sums
// (death of i) stay in registers for the whole row block, column sums (death of j) and the births -- formed in a
// register window as weighted sums, weights = the two-pivot shares -- retire with one batch of independent
// read-modify-writes each per TILE.  History (N = 100, ratio 1.08, 1.2e6 cells): one pair at a time 78.8 ms; 1 x 4
// tiles from global memory 111 ms, through the ring 29.7 ms, + table exponential 26.5 ms; 4 x 4 tiles: DESIGN.md 4.8.
#pragma once

#include "pbe_moc_kernel.cuh"
#include "pbe_moc_tmem_kernel.cuh" // tensor-memory helpers (tmem_alloc, tmem_ld4, ...)

namespace pbe {

#ifndef PBE_MOCP_UNROLL
#define PBE_MOCP_UNROLL 1
#endif
constexpr int MOCP_UNROLL = PBE_MOCP_UNROLL; // tiles in flight per trip of the slot loop
constexpr int MOCP_TI = 4, MOCP_TJ = 4;      // rows x columns of a tile
constexpr int MOCP_PAIRS = MOCP_TI * MOCP_TJ;
constexpr int MOCP_WIN = 8;                  // birth window, pivots
constexpr int MOCP_SLOT_TILES = 4;           // tiles per ring slot: the tile list streams through two shared-memory slots
                                             // filled by cp.async.bulk (UBLKCP) on mbarriers, as the pair table of the
                                             // uniform-grid kernels does

struct alignas(16) FpTile {
    int i0, j0, tb, npairs;
    double a[MOCP_PAIRS];  // [r][c]: CT: G_ij (halved on the diagonal); otherwise S_ij (halved on the diagonal); 0 = no pair
    double b[MOCP_PAIRS];  // CT: H_ij
    double cw[MOCP_PAIRS][MOCP_WIN]; // share of the pair's birth that goes to pivot tb + w (two non-zeros per pair)
};

struct MocFpExtra {
    const FpTile *tiles; // ntiles entries, a multiple of MOCP_SLOT_TILES (padded with empty tiles)
    int ntiles;
    const double *hh;    // [N]: (x_{i+1} - x_{i-1}) / 2, x_{-1} = 0: hat area of pivot i
    const double *cself; // [N]: (x_i - x_{i-1}) / x_i: daughters of class i that stay in class i
    const double *wk;    // [N]: 2 / x_k: beta(., x_k) of uniformBinaryBreakup.C:55-60
};

constexpr int MOCP_BLOCK = 128;

// ring geometry: the shared-memory variant has room for two slots of four tiles; the tensor-memory variant
// (accumulators in TMEM, two CTAs per SM) keeps three single-tile slots
template <bool TM>
struct MocpRing {
    static constexpr int slot_tiles = TM ? 1 : MOCP_SLOT_TILES;
    static constexpr int nslots = TM ? 3 : 2;
};
__host__ __device__ inline size_t mocp_smem_bytes(int N, bool tm = false) {
    const size_t rows = tm ? (size_t)(N + 1) : (size_t)(2 * N + 2); // densities (+ accumulators)
    const size_t ring = tm ? (size_t)MocpRing<true>::nslots * MocpRing<true>::slot_tiles
                           : (size_t)MocpRing<false>::nslots * MocpRing<false>::slot_tiles;
    return (rows * MOCP_BLOCK + (size_t)(CLS_ROWS + 3) * (N + 1)) * sizeof(double) + ring * sizeof(FpTile) +
           4 * sizeof(uint64_t) + 128 * sizeof(double);
}

// TM: the per-cell source accumulators live in TENSOR MEMORY (class c in columns 2c, 2c+1 of the thread's lane, as
// in pbe_moc_tmem_kernel.cuh) instead of shared memory: the CTA's shared memory halves and two CTAs share an SM --
// two warps per scheduler for a loop that one warp runs at half rate.  tmem_cols: columns to allocate (power of 2).
// WIN: the part of the MOCP_WIN-wide birth window the grid's tiles actually use (6 on geometric grids of ratio
// >= ~1.06, where a 4 x 4 tile spans at most five pivots; 8 on the uniform grid): unused weight columns are neither
// loaded nor multiplied.
template <int COAL, int BR, bool TM, int WIN>
__global__ void __launch_bounds__(MOCP_BLOCK, TM ? 2 : 1)
    moc_pivot_kernel(const __grid_constant__ MocParams p, const __grid_constant__ MocFpExtra x, const int tmem_cols) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int BLOCK = MOCP_BLOCK;
    const int N = p.N;
    const int tid = threadIdx.x;
    constexpr int SLOT_TILES = MocpRing<TM>::slot_tiles, NSLOT = MocpRing<TM>::nslots;
    FpTile *ring = reinterpret_cast<FpTile *>(smem_raw);    // [NSLOT][SLOT_TILES]
    uint64_t *bars = reinterpret_cast<uint64_t *>(ring + NSLOT * SLOT_TILES); // NSLOT mbarriers, then the TMEM base address
    uint32_t *tmem_base_s = reinterpret_cast<uint32_t *>(bars + 3);
    double *T128 = reinterpret_cast<double *>(bars + 4);    // 2^(j/128) table of exp2_neg_tab_safe (pbe_math.cuh)
    double *n_s = T128 + 128;                               // [N + 1][BLOCK]: row N = 0, the class of padding pairs
    double *src_s = n_s + (size_t)(N + 1) * BLOCK;          // !TM: [N + 1][BLOCK], row N absorbs window slots past the grid
    double *cls_s = src_s + (TM ? 0 : (size_t)(N + 1) * BLOCK); // CLS_ROWS rows of the common tables, then hh, cself, wk
    volatile double *sv = src_s;
    if (TM && tid < 32) tmem_alloc(tmem_base_s, (uint32_t)tmem_cols);
    if (tid == 0) {
        for (int j = 0; j < NSLOT; ++j) mbar_init(&bars[j], 1);
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
    if (TM) tmem_fence_before_sync();
    __syncthreads();
    if (TM) tmem_fence_after_sync();
    // lane quarter of this warp in bits 31:16; the thread's accumulators start at column 0
    const uint32_t tacc = TM ? *tmem_base_s + ((uint32_t)((tid >> 5) * 32) << 16) : 0u;
    const int n_acc_groups = (N + MOCT_PAD_CLASSES + 3) / 4;
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
    const int nslots = (COAL == DCOAL_OFF) ? 0 : x.ntiles / SLOT_TILES; // ring fills per batch
    const int64_t my_batches =
        (nbatches > (int64_t)blockIdx.x) ? (nbatches - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const int64_t total_q = my_batches * nslots;
    auto issue = [&](int64_t qq) {
        const int sl = (int)(qq % nslots);
        const int rs = (int)(qq % NSLOT);
        uint64_t *bar = &bars[rs];
        const unsigned bytes = (unsigned)(SLOT_TILES * sizeof(FpTile));
        mbar_expect_tx(bar, bytes);
        bulk_g2s(ring + rs * SLOT_TILES, x.tiles + (size_t)sl * SLOT_TILES, bytes, bar);
    };
    if (tid == 0)
        for (int j = 0; j < NSLOT; ++j)
            if (j < total_q) issue(j);
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
                if (!TM) sv[i * BLOCK + tid] = 0.0;
            }
            if (!TM) sv[N * BLOCK + tid] = 0.0;
            n_s[N * BLOCK + tid] = 0.0;
            if (TM && COAL != DCOAL_OFF) { // zero the accumulators (classes 0 .. N + pad, whole x8 groups)
                Tm4 z;
#pragma unroll
                for (int k = 0; k < 8; ++k) z.r[k] = 0u;
                for (int g = 0; g < n_acc_groups; ++g) tmem_st4(tacc + 8u * (uint32_t)g, z);
            }
        }
        const double eps = __ldg(p.eps + cc), rho = __ldg(p.rho + cc), nu = __ldg(p.nu + cc);
        const double eps13 = cbrt(eps);
        double A = 1.0, negK = 0.0;
        if (COAL == DCOAL_CT) {
            A = p.coalA * eps13;
            negK = -(p.coalK * rho * eps * (rho * nu)) * 1.4426950408889634074; // x log2(e): exp(x) = 2^(x log2 e)
        }
        if (COAL != DCOAL_OFF) {
            int iblk = -1;
            double ni[MOCP_TI], di[MOCP_TI];
            int irow[MOCP_TI];
#pragma unroll
            for (int r = 0; r < MOCP_TI; ++r) ni[r] = di[r] = 0.0, irow[r] = N * BLOCK + tid;
            auto flush_rows = [&]() { // death of the row block's classes: distinct rows, one batch
                if (TM) {
                    Tm4 v;
                    const uint32_t a = tacc + 2u * (uint32_t)iblk;
                    tmem_wait_st();
                    tmem_ld4(a, v);
                    tmem_wait_ld(v);
                    tm_axpy4(v, -1.0, di);
                    tmem_st4(a, v);
                    return;
                }
                double cur[MOCP_TI];
#pragma unroll
                for (int r = 0; r < MOCP_TI; ++r) cur[r] = sv[irow[r]];
#pragma unroll
                for (int r = 0; r < MOCP_TI; ++r) sv[irow[r]] = cur[r] - di[r];
            };
#pragma unroll 1
            for (int sl = 0; sl < nslots; ++sl, ++qq) {
              const int rs = (int)(qq % NSLOT);
              mbar_wait(&bars[rs], (unsigned)((qq / NSLOT) & 1));
              const FpTile *slot = ring + rs * SLOT_TILES;
#pragma unroll MOCP_UNROLL
              for (int q = 0; q < SLOT_TILES; ++q) {
                const FpTile &T = slot[q]; // uniform shared-memory address: broadcast loads
                const int i0 = T.i0, j0 = T.j0, tb = T.tb;
                if (i0 != iblk) { // warp-uniform: a new row block
                    if (iblk >= 0) flush_rows();
                    iblk = i0;
#pragma unroll
                    for (int r = 0; r < MOCP_TI; ++r) {
                        irow[r] = min(i0 + r, N) * BLOCK + tid; // rows past the grid: the zero class
                        ni[r] = n_s[irow[r]];
                        di[r] = 0.0;
                    }
                }
                double nj[MOCP_TJ], dj[MOCP_TJ], bw[MOCP_WIN];
                int jrow[MOCP_TJ];
#pragma unroll
                for (int c = 0; c < MOCP_TJ; ++c) {
                    jrow[c] = min(j0 + c, N) * BLOCK + tid;
                    nj[c] = n_s[jrow[c]];
                    dj[c] = 0.0;
                }
#pragma unroll
                for (int w = 0; w < MOCP_WIN; ++w) bw[w] = 0.0;
                // TM: the targets -- column sums at classes j0 .. j0+3, births at tb .. tb+7 (scratch classes above
                // N-1 take what falls past the grid: no clamping).  tb >= j0 always.  Near the diagonal of the pair
                // list (tb < j0 + 4, most tiles of a geometric grid) the two ranges overlap: ONE 12-class window
                // starting at j0 takes both updates in registers.  The loads are issued between the two halves of
                // the tile's arithmetic: the first half covers the previous tile's stores, the second the loads.
                Tm4 tv[3];
                const int dwin = tb - j0;              // warp-uniform
                const bool merged = dwin < MOCP_TJ;
                const uint32_t a_dj = tacc + 2u * (uint32_t)j0, a_bw = tacc + 2u * (uint32_t)tb;
#pragma unroll
                for (int r = 0; r < MOCP_TI; ++r) {
                    if (TM && r == MOCP_TI / 2) {
                        tmem_wait_st();
                        tmem_ld4(a_dj, tv[0]);
                        tmem_ld4(merged ? a_dj + 8u : a_bw, tv[1]);
                        tmem_ld4(merged ? a_dj + 16u : a_bw + 8u, tv[2]);
                    }
#pragma unroll
                    for (int c = 0; c < MOCP_TJ; ++c) {
                        const int k = r * MOCP_TJ + c;
                        const double w = ni[r] * nj[c];
                        double tt;
                        if (COAL == DCOAL_CT)
                            tt = (exp2_neg_tab_safe(negK * T.b[k], T128) * T.a[k]) * w; // 8 FP64 instructions (exp_neg: ~25)
                        else
                            tt = T.a[k] * w;
                        di[r] += tt;
                        dj[c] += tt;
                        // births: a weight row per pair (the two-pivot shares, zero elsewhere) keeps the loop free of
                        // branches.  Tried instead: one table word per pair (the window offset, the same for every
                        // cell) and a warp-uniform switch that puts the two adds on compile-time registers -- inside
                        // this loop it cuts the tile into 16 basic blocks and serialises the exponentials (25.2 ms),
                        // after the loop 16 branch round trips per tile remain (21.4 ms); the weights: 14.8 ms.
#pragma unroll
                        for (int ww = 0; ww < WIN; ++ww) bw[ww] = fma_(T.cw[k][ww], tt, bw[ww]);
                    }
                }
                if (TM) {
                    tmem_wait_ld(tv[0], tv[1]);
                    tmem_wait_ld(tv[2]);
                    tm_axpy4(tv[0], -1.0, dj); // death of the four j
                    if (!merged) {
                        tm_axpy4(tv[1], 1.0, bw);
                        tm_axpy4(tv[2], 1.0, bw + 4);
                        tmem_st4(a_dj, tv[0]);
                        tmem_st4(a_bw, tv[1]);
                        tmem_st4(a_bw + 8u, tv[2]);
                    } else {
                        switch (dwin) { // births at window offset dwin .. dwin + 7 of the 12 loaded classes
                        case 0: tm_window_add<0, WIN, 3>(tv, bw); break;
                        case 1: tm_window_add<1, WIN, 3>(tv, bw); break;
                        case 2: tm_window_add<2, WIN, 3>(tv, bw); break;
                        default: tm_window_add<3, WIN, 3>(tv, bw); break;
                        }
                        tmem_st4(a_dj, tv[0]);
                        tmem_st4(a_dj + 8u, tv[1]);
                        tmem_st4(a_dj + 16u, tv[2]);
                    }
                    continue;
                }
                // deaths of the four j: independent targets (columns past the grid hit the dummy row with 0)
                double cj[MOCP_TJ];
#pragma unroll
                for (int c = 0; c < MOCP_TJ; ++c) cj[c] = sv[jrow[c]];
#pragma unroll
                for (int c = 0; c < MOCP_TJ; ++c) sv[jrow[c]] = cj[c] - dj[c];
                // births: the window tb .. tb + WIN - 1, one batch of read-modify-writes
                double cur[WIN];
                int wrow[WIN];
#pragma unroll
                for (int w = 0; w < WIN; ++w) {
                    wrow[w] = min(tb + w, N) * BLOCK + tid; // past the grid: weights are 0, the dummy row takes them
                    cur[w] = sv[wrow[w]];
                }
#pragma unroll
                for (int w = 0; w < WIN; ++w) sv[wrow[w]] = cur[w] + bw[w];
              }
              __syncthreads(); // every warp is done with this ring slot
              if (tid == 0 && qq + NSLOT < total_q) issue(qq + NSLOT);
            }
            if (iblk >= 0) flush_rows();
            if (TM) tmem_wait_st();
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
        Tm4 accv; // TM: the group of four accumulators that holds class i
#pragma unroll
        for (int k = 0; k < 8; ++k) accv.r[k] = 0u;
#pragma unroll 2
        for (int i = N - 1; i >= 0; --i) {
            double acc_i;
            if (TM) {
                if (COAL != DCOAL_OFF && ((i & 3) == 3 || i == N - 1)) { // warp-uniform: a new group
                    tmem_ld4(tacc + 2u * (uint32_t)(i & ~3), accv);
                    tmem_wait_ld(accv);
                }
                const int r = i & 3;
                acc_i = (r == 0) ? tm_get(accv, 0) : (r == 1) ? tm_get(accv, 1) : (r == 2) ? tm_get(accv, 2) : tm_get(accv, 3);
            } else {
                acc_i = sv[i * BLOCK + tid];
            }
            const double ni = n_s[i * BLOCK + tid];
            double g = cBP[i];
            if (BR == DBR_CT) g = (Ab * g) * exp2_neg_tab_safe(negKb * cBQ[i], T128);
            const double gn = g * ni;
            // daughters of the classes above (suffix sum R) and of class i itself, minus its death
            const double B = fma_(hh_s[i], R, (cs_s[i] - 1.0) * gn);
            R = fma_(wk_s[i], gn, R);
            const double C = A * acc_i;
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
    if (TM) {
        tmem_fence_before_sync();
        __syncthreads();
        if (tid < 32) tmem_dealloc(*tmem_base_s, (uint32_t)tmem_cols);
    }
}

} // namespace pbe
