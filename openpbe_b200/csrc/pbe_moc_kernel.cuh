// pbe_moc_kernel.cuh -- the method-of-classes source kernel for sm_100a.
//
// Replaces, in ONE launch, the reference's MOC::correct() source path
// (src/PBESystem/MOC.C:87-184): breakup-rate cache fill (:162-171), N calls of
// coalescenceSourceTerm (:92-124) and breakupSourceTerm (:126-157), the limiter of
// classSourceTerm (:87-91), optionally the zero-flux explicit update (:259-272 with phi = 0),
// the continuity check (:254-279), d() (:282-290) and the global-moment partial sums.
//
// Design:
//  * one THREAD per cell, lane = cell: every global access is a coalesced 256-B warp
//    transaction on the class-major SoA fields; no cross-lane traffic at all;
//  * per-thread state n[0..N) and src[0..N) lives in shared memory as column `tid` of
//    [N+1][BLOCK] FP64 arrays (conflict-free: consecutive lanes, consecutive 8-B words);
//    row N is a zero "dummy" class that absorbs the padded table slots;
//  * the coalescence kernel S(xi_i, xi_j, c) is symmetric, so each UNORDERED class pair is
//    evaluated once and scattered to death(i), death(j) and birth(i+j+1);
//  * Coulaloglou-Tavlarides factorises into (cell scalars) x (pair geometry):
//    S = A_c G_ij exp(-K_c H_ij); G and H are host-built tables streamed in traversal order
//    through a double-buffered shared-memory ring by cp.async.bulk (TMA 1-D bulk copy,
//    SASS UBLKCP) completing on an mbarrier -- one FP64 exp per pair is all that is left;
//  * pairs are walked in TI x TJ register tiles: row sums (death_i) stay in registers for the
//    whole row block, column sums (death_j) and the sliding anti-diagonal window (birth_k)
//    are retired to shared memory once per tile;
//  * invalid slots of edge tiles carry G = 0 (and the diagonal carries G/2), so the inner
//    loop has no predicates at all;
//  * breakup birth with uniformBinaryBreakup is a suffix sum (O(N) instead of O(N^2)), fused
//    with the limiter / update / d / moments epilogue.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "pbe_math.cuh"

namespace pbe {

constexpr int MOC_TI = 4;
constexpr int MOC_TJ = 4;
constexpr int MOC_MAX_BLOCKS = 64; // row blocks: N <= 2 * TI * 64
constexpr int MOC_MAX_MOMENTS = 8;

// device-side coalescence / breakup evaluation modes
enum { DCOAL_OFF = 0, DCOAL_CT = 1, DCOAL_SYM = 2, DCOAL_ASYM = 3 };
enum { DBR_TAB = 0, DBR_CT = 1 };

// per-class table rows in MocParams::cls (each row has N+1 doubles, the last is the dummy 0)
enum { CLS_BR_P = 0, CLS_BR_Q = 1, CLS_W = 2, CLS_DVOL = 3, CLS_XI = 4, CLS_ROWS = 5 };

struct MocParams {
    int N;
    int nblocks;            // row blocks of TI classes that own at least one valid pair
    int chunk_max_pairs;    // largest chunk, in pairs (ring slot size)
    int do_step;            // 1: n_out = n + dt S (in place allowed)
    int64_t ncells, stride;
    const double *n_in;
    double *n_out;          // do_step only
    double *S_out;          // may be NULL
    double *d_out;          // may be NULL
    const double *eps, *rho, *nu;
    const double *alpha, *vcell; // may be NULL
    const double2 *pair_tab;     // [total_pairs] {G,H} / {S_ij,S_ji} in traversal order
    const double *cls;           // [CLS_ROWS][N+1]
    double coalA, coalK;         // CT: A_c = coalA cbrt(eps); K_c = coalK rho eps (rho nu)
    double brA, brK;             // CT: Ab_c = brA cbrt(eps);  Kb_c = brK / (rho eps^(2/3))
    double relax, dt;            // limiter -relax n / dt (MOC.C:87-91)
    double negK_floor;           // TMEM kernel: lower clamp of -K_c (pbe_moc_tmem_kernel.cuh)
    double negKb_floor;          // TMEM kernel: lower clamp of -Kb_c
    double *moments;             // may be NULL: [n_moments] atomicAdd
    int n_moments;
    double *cont_err_max;        // may be NULL: [1] atomic max (values >= 0)
    int chunk_off[MOC_MAX_BLOCKS + 1]; // pair offset of each row block's chunk
};

// ---- mbarrier / bulk-copy helpers (PTX ISA 8.x; SASS: SYNCS.*, UBLKCP) ---------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// 1-D bulk copy global -> shared, completion counted in bytes on the mbarrier
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, unsigned bytes,
                                         uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
            "r"(smem_u32(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

__device__ __forceinline__ void atomic_max_nonneg(double *addr, double v) {
    // v >= 0: IEEE ordering equals unsigned-integer ordering of the bit patterns
    atomicMax((unsigned long long *)addr, (unsigned long long)__double_as_longlong(v));
}

// shared-memory footprint of one CTA
__host__ __device__ inline size_t moc_smem_bytes(int N, int block, int chunk_max_pairs) {
    size_t b = 0;
    b += (size_t)2 * (N + 1) * block * sizeof(double);       // n_s, src_s
    b += (size_t)2 * chunk_max_pairs * sizeof(double2);      // table ring
    b += (size_t)CLS_ROWS * (N + 1) * sizeof(double);        // per-class tables
    b += 2 * sizeof(uint64_t);                               // mbarriers
    return b;
}

// ---- one TI x TJ tile of class pairs -------------------------------------------------------
// tab points at the tile's TI*TJ {G,H} entries (uniform address: LDS.128 broadcast).
template <int COAL>
__device__ __forceinline__ void moc_tile(const double2 *__restrict__ tab, double negK,
                                         const double (&ni)[MOC_TI], const double (&nj)[MOC_TJ],
                                         double (&di)[MOC_TI], double (&dj)[MOC_TJ],
                                         double (&bw)[MOC_TI + MOC_TJ - 1]) {
#pragma unroll
    for (int r = 0; r < MOC_TI; ++r) {
#pragma unroll
        for (int c = 0; c < MOC_TJ; ++c) {
            const double2 gh = tab[r * MOC_TJ + c];
            const double w = ni[r] * nj[c];
            if (COAL == DCOAL_CT) {
                // S/A_c = G exp(-K_c H)   (coalescenceKernels/CoulaloglouTavlarides.C:75-89)
                const double e = exp_neg(negK * gh.y);
                const double tt = (e * gh.x) * w;
                di[r] += tt;
                dj[c] += tt;
                bw[r + c] += tt;
            } else if (COAL == DCOAL_SYM) {
                const double tt = gh.x * w;
                di[r] += tt;
                dj[c] += tt;
                bw[r + c] += tt;
            } else { // DCOAL_ASYM: x = S(xi_i, xi_j), y = S(xi_j, xi_i)
                const double t1 = gh.x * w, t2 = gh.y * w;
                di[r] += t1;
                dj[c] += t2;
                bw[r + c] += 0.5 * (t1 + t2);
            }
        }
    }
}

// ---- all pairs owned by row block blk (rows i0 .. i0+TI-1) ----------------------------------
template <int COAL, int BLOCK>
__device__ __forceinline__ void moc_row_block(const MocParams &p, int blk,
                                              const double2 *__restrict__ tab, double negK,
                                              const double *__restrict__ n_s,
                                              volatile double *sv, int tid) {
    const int N = p.N;
    const int i0 = blk * MOC_TI;
    const int ntiles = (p.chunk_off[blk + 1] - p.chunk_off[blk]) / (MOC_TI * MOC_TJ);
    double ni[MOC_TI], di[MOC_TI], bw[MOC_TI + MOC_TJ - 1];
#pragma unroll
    for (int r = 0; r < MOC_TI; ++r) {
        ni[r] = n_s[min(i0 + r, N) * BLOCK + tid];
        di[r] = 0.0;
    }
#pragma unroll
    for (int q = 0; q < MOC_TI + MOC_TJ - 1; ++q) bw[q] = 0.0;

    int j0 = i0;
#pragma unroll 1
    for (int t = 0; t < ntiles; ++t, j0 += MOC_TJ, tab += MOC_TI * MOC_TJ) {
        double nj[MOC_TJ], dj[MOC_TJ];
        int jidx[MOC_TJ], kidx[MOC_TJ];
#pragma unroll
        for (int c = 0; c < MOC_TJ; ++c) {
            jidx[c] = min(j0 + c, N) * BLOCK + tid;
            kidx[c] = min(i0 + j0 + 1 + c, N) * BLOCK + tid;
            nj[c] = n_s[jidx[c]];
            dj[c] = 0.0;
        }
        moc_tile<COAL>(tab, negK, ni, nj, di, dj, bw);
        // retire column sums (death of class j): batched load / store, distinct targets
        double a[MOC_TJ];
#pragma unroll
        for (int c = 0; c < MOC_TJ; ++c) a[c] = sv[jidx[c]];
#pragma unroll
        for (int c = 0; c < MOC_TJ; ++c) sv[jidx[c]] = a[c] - dj[c];
        // retire the TJ lowest anti-diagonals (birth of class i+j+1); may alias the column
        // targets when i0 + 1 < TJ, hence a second batch after the first
#pragma unroll
        for (int c = 0; c < MOC_TJ; ++c) a[c] = sv[kidx[c]];
#pragma unroll
        for (int c = 0; c < MOC_TJ; ++c) sv[kidx[c]] = a[c] + bw[c];
#pragma unroll
        for (int q = 0; q < MOC_TI + MOC_TJ - 1; ++q)
            bw[q] = (q + MOC_TJ < MOC_TI + MOC_TJ - 1) ? bw[q + MOC_TJ] : 0.0;
    }
    // flush the TI-1 anti-diagonals still in the window, then the row sums (death of class i)
    {
        double a[MOC_TI];
        int idx[MOC_TI];
#pragma unroll
        for (int q = 0; q < MOC_TI - 1; ++q) {
            idx[q] = min(i0 + j0 + 1 + q, N) * BLOCK + tid;
            a[q] = sv[idx[q]];
        }
#pragma unroll
        for (int q = 0; q < MOC_TI - 1; ++q) sv[idx[q]] = a[q] + bw[q];
#pragma unroll
        for (int r = 0; r < MOC_TI; ++r) {
            idx[r] = min(i0 + r, N) * BLOCK + tid;
            a[r] = sv[idx[r]];
        }
#pragma unroll
        for (int r = 0; r < MOC_TI; ++r) sv[idx[r]] = a[r] - di[r];
    }
}

// ---- the kernel: persistent CTAs, BLOCK cells per batch ------------------------------------
template <int COAL, int BR, int BLOCK>
__global__ void __launch_bounds__(BLOCK, 1) moc_source_kernel(const __grid_constant__ MocParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int N = p.N;
    const int tid = threadIdx.x;
    double2 *ring = reinterpret_cast<double2 *>(smem_raw);
    double *n_s = reinterpret_cast<double *>(ring + 2 * (size_t)p.chunk_max_pairs);
    double *src_s = n_s + (size_t)(N + 1) * BLOCK;
    double *cls_s = src_s + (size_t)(N + 1) * BLOCK;
    uint64_t *bars = reinterpret_cast<uint64_t *>(cls_s + CLS_ROWS * (N + 1));
    volatile double *sv = src_s;

    const int64_t nbatches = (p.ncells + BLOCK - 1) / BLOCK;
    const int64_t my_batches =
        (nbatches > (int64_t)blockIdx.x) ? (nbatches - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const int64_t total_q = (COAL == DCOAL_OFF) ? 0 : my_batches * p.nblocks;

    if (tid == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        mbar_fence_init();
    }
    for (int k = tid; k < CLS_ROWS * (N + 1); k += BLOCK) cls_s[k] = p.cls[k];
    n_s[N * BLOCK + tid] = 0.0; // dummy class: absorbs padded table slots
    sv[N * BLOCK + tid] = 0.0;
    __syncthreads();

    auto issue = [&](int64_t q) {
        const int blk = (int)(q % p.nblocks);
        const int off = p.chunk_off[blk];
        const unsigned bytes = (unsigned)(p.chunk_off[blk + 1] - off) * (unsigned)sizeof(double2);
        uint64_t *bar = &bars[q & 1];
        mbar_expect_tx(bar, bytes);
        bulk_g2s(ring + (q & 1) * (size_t)p.chunk_max_pairs, p.pair_tab + off, bytes, bar);
    };
    if (tid == 0) {
        if (0 < total_q) issue(0);
        if (1 < total_q) issue(1);
    }

    const double *cBP = cls_s + CLS_BR_P * (N + 1);
    const double *cBQ = cls_s + CLS_BR_Q * (N + 1);
    const double *cW = cls_s + CLS_W * (N + 1);
    const double *cDV = cls_s + CLS_DVOL * (N + 1);
    const double *cXI = cls_s + CLS_XI * (N + 1);

    double macc[MOC_MAX_MOMENTS];
#pragma unroll
    for (int k = 0; k < MOC_MAX_MOMENTS; ++k) macc[k] = 0.0;
    double cerr = 0.0;
    const bool want_diag = (p.moments != nullptr) || (p.cont_err_max != nullptr);

    int64_t q = 0;
    for (int64_t batch = blockIdx.x; batch < nbatches; batch += gridDim.x) {
        const int64_t cell = batch * BLOCK + tid;
        const bool live = cell < p.ncells;
        const int64_t cc = live ? cell : p.ncells - 1;

        // stage this cell's class densities (coalesced: lane = cell)
        {
            const double *src = p.n_in + cc;
#pragma unroll 10
            for (int i = 0; i < N; ++i) {
                n_s[i * BLOCK + tid] = __ldg(src + (size_t)i * p.stride);
                sv[i * BLOCK + tid] = 0.0;
            }
        }
        const double eps = __ldg(p.eps + cc), rho = __ldg(p.rho + cc), nu = __ldg(p.nu + cc);
        const double eps13 = cbrt(eps);
        double A = 1.0, negK = 0.0;
        if (COAL == DCOAL_CT) {
            A = p.coalA * eps13;
            negK = -(p.coalK * rho * eps * (rho * nu));
        }

        if (COAL != DCOAL_OFF) {
#pragma unroll 1
            for (int blk = 0; blk < p.nblocks; ++blk, ++q) {
                mbar_wait(&bars[q & 1], (unsigned)((q >> 1) & 1));
                moc_row_block<COAL, BLOCK>(p, blk, ring + (q & 1) * (size_t)p.chunk_max_pairs,
                                           negK, n_s, sv, tid);
                __syncthreads(); // every warp is done with ring slot q&1
                if (tid == 0 && q + 2 < total_q) issue(q + 2);
            }
        }

        // ---- epilogue: breakup (suffix sum), limiter, update, d, diagnostics ----------------
        double Ab = 0.0, negKb = 0.0;
        if (BR == DBR_CT) {
            Ab = p.brA * eps13;
            negKb = -(p.brK / (rho * (eps13 * eps13)));
        }
        double R = 0.0, sum_d = 0.0, sum_n = 0.0;
        double mloc[MOC_MAX_MOMENTS];
#pragma unroll
        for (int k = 0; k < MOC_MAX_MOMENTS; ++k) mloc[k] = 0.0;
#pragma unroll 4
        for (int i = N - 1; i >= 0; --i) {
            const double ni = n_s[i * BLOCK + tid];
            // breakupCache_ (MOC.C:165-171): CT = Ab xi^(-2/9) exp(-Kb xi^(-5/9))
            double g = cBP[i];
            if (BR == DBR_CT) g = (Ab * g) * exp_neg(negKb * cBQ[i]);
            const double gn = g * ni;
            const double B = R - gn;       // -g_i n_i + sum_{j>i} beta g_j n_j dxi (MOC.C:142-155)
            R = fma_(cW[i], gn, R);
            const double C = A * sv[i * BLOCK + tid];
            // MOC.C:87-91: -relax n_i / dt with the reference's operation order (a true
            // division), so that limited cells carry bit-identical values
            const double lim = __dmul_rn(-p.relax, ni) / p.dt;
            double s = C + B;
            s = (s > lim) ? s : lim;
            if (p.S_out != nullptr && live) p.S_out[(size_t)i * p.stride + cell] = s;
            double nn = ni;
            if (p.do_step) {
                nn = __dadd_rn(ni, __dmul_rn(p.dt, s)); // n += dt S  (MOC.C:259-272, phi = 0)
                if (live) p.n_out[(size_t)i * p.stride + cell] = nn;
            }
            sum_d = fma_(cDV[i], nn, sum_d);
            sum_n += nn;
            if (want_diag) {
                const double x = cXI[i];
                double pw = nn;
#pragma unroll
                for (int k = 0; k < MOC_MAX_MOMENTS; ++k) {
                    if (k < p.n_moments || k == 1) mloc[k] += pw;
                    pw *= x;
                }
            }
        }
        if (p.d_out != nullptr && live) p.d_out[cell] = sum_d / sum_n; // MOC.C:282-290
        if (want_diag && live) {
            const double V = (p.vcell != nullptr) ? __ldg(p.vcell + cell) : 1.0;
#pragma unroll
            for (int k = 0; k < MOC_MAX_MOMENTS; ++k) macc[k] = fma_(V, mloc[k], macc[k]);
            if (p.cont_err_max != nullptr && p.alpha != nullptr) {
                // | sum_k xi_k n_k / V - alpha |   (MOC.C:254-279)
                const double e = fabs(mloc[1] / V - __ldg(p.alpha + cell));
                cerr = (e > cerr) ? e : cerr;
            }
        }
    }

    if (want_diag) {
        // block reduction through the (now idle) n_s columns, one atomic per CTA per value
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
