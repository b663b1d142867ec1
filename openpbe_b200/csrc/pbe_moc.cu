// pbe_moc.cu -- host side of the method-of-classes path: table construction at create time,
// launch configuration, the host-buffer pipeline.  Kernels: pbe_moc_kernel.cuh.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <cmath>

#include <algorithm>

#include "pbe_common.h"
#include "pbe_moc_kernel.cuh"
#include "pbe_moc_pivot_kernel.cuh"
#include "pbe_moc_tmem_kernel.cuh"

using namespace pbe;

namespace {
cudaError_t resolve_minmax_scratch(int device); // static scratch of pbe_moc_class_minmax_noalloc (end of file)
}

struct pbe_moc {
    pbe_ctx *ctx = nullptr;
    int N = 0;
    double xi1 = 0, relax = 1.0;
    pbe_kernel_spec spec{};
    int dcoal = DCOAL_OFF, dbr = DBR_TAB;
    int block = 128;       // cells per CTA of the selected kernel
    int smem_block = 128;  // shared-memory kernel: cells per CTA, CTAs per SM, bytes per CTA
    int smem_ctas_per_sm = 1;
    size_t smem_bytes = 0;
    bool tmem_ok = false;  // tensor-memory kernel available for this N / table
    bool tmem_wide = false; // tensor-memory kernel walking two table tiles per step (more registers)
    bool use_tmem = false; // Blackwell fast path: accumulators in tensor memory, 256 cells per CTA
    int tmem_cols = 0;     // TMEM columns per thread (= per CTA: one lane per thread)
    int tmem_ctas_per_sm = 1;
    size_t tmem_bytes = 0;
    MoctSched sched{};     // ring schedule of the tensor-memory kernel
    int ctas_per_sm = 1;
    size_t smem = 0;
    MocParams base{};      // table pointers, chunk offsets, folded coefficients
    double2 *d_pair = nullptr;
    double2 *d_pair_log = nullptr; // CT, tensor-memory kernel: {log2 G_ij, H_ij}
    double *d_cls = nullptr;
    // fixed-pivot grid (pbe_moc_create_pivots): arbitrary class volumes, pbe_moc_pivot_kernel.cuh
    bool pivot = false;
    std::vector<double> pivots;
    FpTile *d_fp = nullptr;
    double *d_fpw = nullptr; // [3][N]: hat areas, self shares, 2 / x_k
    MocFpExtra fpx{};
    int pivot_win = 8; // birth-window width the pivot kernel is instantiated for (6 or MOCP_WIN)
};

namespace {

typedef void (*moc_kernel_fn)(const MocParams);

template <int BLOCK>
moc_kernel_fn pick_kernel(int dcoal, int dbr) {
#define PBE_PICK(C, B) \
    if (dcoal == C && dbr == B) return (moc_kernel_fn)moc_source_kernel<C, B, BLOCK>;
    PBE_PICK(DCOAL_OFF, DBR_TAB)
    PBE_PICK(DCOAL_OFF, DBR_CT)
    PBE_PICK(DCOAL_CT, DBR_TAB)
    PBE_PICK(DCOAL_CT, DBR_CT)
    PBE_PICK(DCOAL_SYM, DBR_TAB)
    PBE_PICK(DCOAL_SYM, DBR_CT)
    PBE_PICK(DCOAL_ASYM, DBR_TAB)
    PBE_PICK(DCOAL_ASYM, DBR_CT)
#undef PBE_PICK
    return nullptr;
}

typedef void (*moct_kernel_fn)(const MocParams, const MoctSched);
moct_kernel_fn pick_tmem_kernel(int dcoal, int dbr, bool wide) {
#define PBE_PICK(C, B)                                                                             \
    if (dcoal == C && dbr == B)                                                                    \
        return wide ? (moct_kernel_fn)moc_tmem_kernel<C, B, true> : (moct_kernel_fn)moc_tmem_kernel<C, B, false>;
    PBE_PICK(DCOAL_OFF, DBR_TAB)
    PBE_PICK(DCOAL_OFF, DBR_CT)
    PBE_PICK(DCOAL_CT, DBR_TAB)
    PBE_PICK(DCOAL_CT, DBR_CT)
    PBE_PICK(DCOAL_SYM, DBR_TAB)
    PBE_PICK(DCOAL_SYM, DBR_CT)
    PBE_PICK(DCOAL_ASYM, DBR_TAB)
    PBE_PICK(DCOAL_ASYM, DBR_CT)
#undef PBE_PICK
    return nullptr;
}

typedef void (*mocp_kernel_fn)(const MocParams, const MocFpExtra, const int);
mocp_kernel_fn pick_pivot_kernel(int dcoal, int dbr, bool tm, int win = MOCP_WIN) {
#define PBE_PICK(C, B)                                                                                                  \
    if (dcoal == C && dbr == B) {                                                                                       \
        if (win <= 6)                                                                                                   \
            return tm ? (mocp_kernel_fn)moc_pivot_kernel<C, B, true, 6> : (mocp_kernel_fn)moc_pivot_kernel<C, B, false, 6>; \
        return tm ? (mocp_kernel_fn)moc_pivot_kernel<C, B, true, MOCP_WIN>                                              \
                  : (mocp_kernel_fn)moc_pivot_kernel<C, B, false, MOCP_WIN>;                                            \
    }
    PBE_PICK(DCOAL_OFF, DBR_TAB)
    PBE_PICK(DCOAL_OFF, DBR_CT)
    PBE_PICK(DCOAL_CT, DBR_TAB)
    PBE_PICK(DCOAL_CT, DBR_CT)
    PBE_PICK(DCOAL_SYM, DBR_TAB)
    PBE_PICK(DCOAL_SYM, DBR_CT)
#undef PBE_PICK
    return nullptr;
}

moc_kernel_fn kernel_for(const pbe_moc *m) {
    return m->smem_block == 128 ? pick_kernel<128>(m->dcoal, m->dbr) : pick_kernel<64>(m->dcoal, m->dbr);
}

void select_kernel(pbe_moc *m, int which) {
    m->use_tmem = (which == PBE_MOC_KERNEL_TMEM);
    if (m->use_tmem) {
        m->block = MOCT_BLOCK;
        m->ctas_per_sm = m->tmem_ctas_per_sm;
        m->smem = m->tmem_bytes;
    } else {
        m->block = m->smem_block;
        m->ctas_per_sm = m->smem_ctas_per_sm;
        m->smem = m->smem_bytes;
    }
}

// Pair geometry of the Coulaloglou-Tavlarides coalescence kernel, with the reference's own
// libm calls (coalescenceKernels/CoulaloglouTavlarides.C:75-89): everything that does not
// depend on the cell.
void ct_pair_geometry(double xa, double xb, double *G, double *H) {
    const double a = pow(xa, 1.0 / 3.0), b = pow(xb, 1.0 / 3.0);
    *G = pow(a + b, 2) * pow(pow(xa, 2.0 / 9.0) + pow(xb, 2.0 / 9.0), 0.5);
    *H = pow(a * b / (a + b), 4.0);
}

// d = sum_i dvol_i n_i / sum_i n_i  (MOC.C:282-290); one thread per cell, coalesced class rows
__global__ void __launch_bounds__(256) moc_d_kernel(int N, int64_t ncells, int64_t stride,
                                                    const double *__restrict__ n,
                                                    const double *__restrict__ dvol,
                                                    double *__restrict__ d) {
    const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
    for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c < ncells; c += nthreads) {
        double sum = 0.0, norm = 0.0;
#pragma unroll 4
        for (int i = 0; i < N; ++i) {
            const double v = __ldg(n + (size_t)i * stride + c);
            sum = fma_(__ldg(dvol + i), v, sum);
            norm += v;
        }
        d[c] = sum / norm;
    }
}

// per-class min / max over cells (MOC.C:181-183, :238): stage 1 = per-CTA partials
__global__ void __launch_bounds__(256) moc_minmax_partial_kernel(int64_t ncells, int64_t stride,
                                                                 const double *__restrict__ n,
                                                                 double *__restrict__ part) {
    const int k = blockIdx.y;
    const double *row = n + (size_t)k * stride;
    double mn = INFINITY, mx = -INFINITY;
    const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
    for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c < ncells; c += nthreads) {
        const double v = __ldg(row + c);
        mn = fmin(mn, v);
        mx = fmax(mx, v);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    __shared__ double smn[8], smx[8];
    if ((threadIdx.x & 31) == 0) {
        smn[threadIdx.x >> 5] = mn;
        smx[threadIdx.x >> 5] = mx;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) {
            mn = fmin(mn, smn[w]);
            mx = fmax(mx, smx[w]);
        }
        part[((size_t)k * gridDim.x + blockIdx.x) * 2 + 0] = mn;
        part[((size_t)k * gridDim.x + blockIdx.x) * 2 + 1] = mx;
    }
}
__global__ void moc_minmax_final_kernel(int nparts, const double *__restrict__ part,
                                        double *__restrict__ out_min, double *__restrict__ out_max) {
    const int k = blockIdx.x;
    double mn = INFINITY, mx = -INFINITY;
    for (int b = threadIdx.x; b < nparts; b += 32) {
        mn = fmin(mn, part[((size_t)k * nparts + b) * 2 + 0]);
        mx = fmax(mx, part[((size_t)k * nparts + b) * 2 + 1]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    if (threadIdx.x == 0) {
        if (out_min) out_min[k] = mn;
        if (out_max) out_max[k] = mx;
    }
}

int launch(pbe_moc *m, MocParams &p, cudaStream_t st) {
    if (p.ncells == 0) return PBE_OK;
    const int64_t nb = (p.ncells + m->block - 1) / m->block;
    const int64_t cap = (int64_t)m->ctx->sm_count * m->ctas_per_sm;
    const int grid = (int)std::min<int64_t>(nb, cap);
    if (m->pivot) {
        mocp_kernel_fn fn = pick_pivot_kernel(m->dcoal, m->dbr, m->use_tmem, m->pivot_win);
        int cols = m->tmem_cols;
        void *args[] = {(void *)&p, (void *)&m->fpx, (void *)&cols};
        PBE_CUDA(raise_dyn_smem(m->ctx, (const void *)fn, m->smem));
        PBE_CUDA(cudaLaunchKernel((const void *)fn, dim3(grid), dim3(m->block), args, m->smem, st));
        m->ctx->launches++;
        return PBE_OK;
    }
    if (m->use_tmem) {
        moct_kernel_fn fn = pick_tmem_kernel(m->dcoal, m->dbr, m->tmem_wide);
        if (m->dcoal == DCOAL_CT) p.pair_tab = m->d_pair_log;
        void *args[] = {(void *)&p, (void *)&m->sched};
        // another handle of this instantiation may have been created since with a smaller table
        PBE_CUDA(raise_dyn_smem(m->ctx, (const void *)fn, m->smem));
        PBE_CUDA(cudaLaunchKernel((const void *)fn, dim3(grid), dim3(m->block), args, m->smem, st));
        m->ctx->launches++;
        return PBE_OK;
    }
    moc_kernel_fn fn = kernel_for(m);
    void *args[] = {(void *)&p};
    PBE_CUDA(raise_dyn_smem(m->ctx, (const void *)fn, m->smem));
    PBE_CUDA(cudaLaunchKernel((const void *)fn, dim3(grid), dim3(m->block), args, m->smem, st));
    m->ctx->launches++;
    return PBE_OK;
}

int check_fields(const char *fn, const pbe_moc *m, int64_t ncells, int64_t stride, const void *n,
                 const void *eps, const void *rho, const void *nu, double dt) {
    PBE_REQUIRE(m != nullptr, PBE_E_INVALID, "%s: NULL handle", fn);
    PBE_REQUIRE(ncells >= 0 && stride >= ncells, PBE_E_INVALID,
                "%s: need 0 <= ncells <= stride (got %lld, %lld)", fn, (long long)ncells,
                (long long)stride);
    PBE_REQUIRE(ncells == 0 || (n && eps && rho && nu), PBE_E_INVALID, "%s: NULL field pointer", fn);
    PBE_REQUIRE(dt > 0.0, PBE_E_INVALID, "%s: dt must be positive (got %g)", fn, dt);
    return PBE_OK;
}

} // namespace

extern "C" {

int pbe_moc_create(pbe_ctx *ctx, int N, double xi1, const pbe_kernel_spec *spec,
                   double source_relaxation, pbe_moc **out) {
    PBE_REQUIRE(ctx && spec && out, PBE_E_INVALID, "pbe_moc_create: NULL argument");
    *out = nullptr;
    PBE_REQUIRE(N >= 1, PBE_E_INVALID, "pbe_moc_create: numberOfClasses must be >= 1 (got %d)", N);
    PBE_REQUIRE(xi1 > 0.0, PBE_E_INVALID, "pbe_moc_create: xi1 must be positive (got %g)", xi1);
    PBE_REQUIRE(spec->coalescence >= PBE_COAL_NONE && spec->coalescence <= PBE_COAL_TABLE,
                PBE_E_UNKNOWN_TYPE, "pbe_moc_create: unknown coalescenceKernel id %d",
                spec->coalescence);
    PBE_REQUIRE(spec->breakup >= PBE_BR_NONE && spec->breakup <= PBE_BR_TABLE, PBE_E_UNKNOWN_TYPE,
                "pbe_moc_create: unknown breakupKernel id %d", spec->breakup);
    PBE_REQUIRE(spec->daughter == PBE_DAUGHTER_UNIFORM_BINARY, PBE_E_UNKNOWN_TYPE,
                "pbe_moc_create: unknown daughterParticleDistribution id %d", spec->daughter);
    PBE_REQUIRE(spec->coalescence != PBE_COAL_TABLE || spec->coal_table, PBE_E_INVALID,
                "pbe_moc_create: PBE_COAL_TABLE needs coal_table");
    PBE_REQUIRE(spec->breakup != PBE_BR_TABLE || spec->br_table, PBE_E_INVALID,
                "pbe_moc_create: PBE_BR_TABLE needs br_table");
    DeviceGuard g(ctx->device);

    pbe_moc *m = new pbe_moc();
    m->ctx = ctx;
    m->N = N;
    m->xi1 = xi1;
    m->relax = source_relaxation;
    m->spec = *spec;
    m->spec.coal_table = nullptr;
    m->spec.br_table = nullptr;

    std::vector<double> xi(N);
    for (int i = 0; i < N; ++i) xi[i] = xi1 * (i + 1); // MOC.C:79

    // ---- device evaluation mode ----------------------------------------------------------
    switch (spec->coalescence) {
    case PBE_COAL_NONE: m->dcoal = DCOAL_OFF; break;
    case PBE_COAL_CONSTANT: m->dcoal = DCOAL_SYM; break;
    case PBE_COAL_CT: m->dcoal = DCOAL_CT; break;
    case PBE_COAL_TABLE: {
        bool sym = true;
        for (int i = 0; i < N && sym; ++i)
            for (int j = 0; j < i; ++j)
                if (spec->coal_table[(size_t)i * N + j] != spec->coal_table[(size_t)j * N + i]) {
                    sym = false;
                    break;
                }
        m->dcoal = sym ? DCOAL_SYM : DCOAL_ASYM;
    } break;
    }
    m->dbr = (spec->breakup == PBE_BR_CT) ? DBR_CT : DBR_TAB;
    if (N < 2) m->dcoal = DCOAL_OFF; // one class: no pair keeps its product inside the grid

    // ---- pair table in traversal order: row blocks of TI classes, TI x TJ tiles ----------
    MocParams &b = m->base;
    b.N = N;
    const int nrows = N / 2; // rows i with 2 i <= N - 2
    b.nblocks = (m->dcoal == DCOAL_OFF) ? 0 : (nrows + MOC_TI - 1) / MOC_TI;
    if (b.nblocks > MOC_MAX_BLOCKS) {
        set_error("pbe_moc_create: numberOfClasses %d exceeds the device path's limit %d", N,
                  2 * MOC_TI * MOC_MAX_BLOCKS);
        delete m;
        return PBE_E_UNSUPPORTED;
    }
    std::vector<double2> tab;
    b.chunk_off[0] = 0;
    b.chunk_max_pairs = 1;
    for (int blk = 0; blk < b.nblocks; ++blk) {
        const int i0 = blk * MOC_TI;
        const int ntiles = (N - 1 - 2 * i0 + MOC_TJ - 1) / MOC_TJ;
        for (int t = 0; t < ntiles; ++t)
            for (int r = 0; r < MOC_TI; ++r)
                for (int c = 0; c < MOC_TJ; ++c) {
                    const int i = i0 + r, j = i0 + t * MOC_TJ + c;
                    double2 v = make_double2(0.0, 0.0);
                    if (i <= j && i + j <= N - 2) { // MOC.C:109 death bound, MOC.C:116 birth
                        switch (spec->coalescence) {
                        case PBE_COAL_CONSTANT: v.x = v.y = spec->coal_c[0]; break;
                        case PBE_COAL_CT: ct_pair_geometry(xi[i], xi[j], &v.x, &v.y); break;
                        case PBE_COAL_TABLE:
                            v.x = spec->coal_table[(size_t)i * N + j];
                            v.y = spec->coal_table[(size_t)j * N + i];
                            break;
                        }
                        if (i == j) { // the diagonal pair feeds death once and birth with 1/2
                            v.x *= 0.5;
                            if (m->dcoal != DCOAL_CT) v.y *= 0.5;
                        }
                    }
                    tab.push_back(v);
                }
        b.chunk_off[blk + 1] = (int)tab.size();
        b.chunk_max_pairs = std::max(b.chunk_max_pairs, b.chunk_off[blk + 1] - b.chunk_off[blk]);
    }

    // ---- per-class tables ------------------------------------------------------------------
    std::vector<double> cls((size_t)CLS_ROWS * (N + 1), 0.0);
    for (int i = 0; i < N; ++i) {
        double P = 0.0, Q = 0.0;
        switch (spec->breakup) {
        case PBE_BR_BINARY: P = xi[i] * xi[i]; break;            // binaryBreakup.C:51-53
        case PBE_BR_CT:                                          // CoulaloglouTavlarides.C:83-92
            P = 1.0 / pow(xi[i], 2.0 / 9.0);
            Q = 1.0 / pow(xi[i], 5.0 / 9.0);
            break;
        case PBE_BR_TABLE: P = spec->br_table[i]; break;
        default: break;                                          // noBreakup.C:51-54
        }
        cls[(size_t)CLS_BR_P * (N + 1) + i] = P;
        cls[(size_t)CLS_BR_Q * (N + 1) + i] = Q;
        cls[(size_t)CLS_W * (N + 1) + i] = (2.0 / xi[i]) * xi1; // beta(.,xi_i) dxi, MOC.C:146-154
        cls[(size_t)CLS_DVOL * (N + 1) + i] = pow(6.0 / M_PI * xi[i], 1.0 / 3.0); // MOC.C:43-45
        cls[(size_t)CLS_XI * (N + 1) + i] = xi[i];
    }

    // ---- folded coefficients (pbe_kernels.cuh uses the same algebra) --------------------------
    {
        const double c1 = spec->coal_c[0], c2 = spec->coal_c[1], gm = spec->coal_c[2],
                     sg = spec->coal_c[3];
        b.coalA = c1 / (1 + gm);
        b.coalK = c2 / (sg * sg * ((1 + gm) * (1 + gm) * (1 + gm)));
    }
    {
        const double c1 = spec->br_c[0], c2 = spec->br_c[1], gm = spec->br_c[2], sg = spec->br_c[3];
        b.brA = c1 / (1 + gm);
        b.brK = c2 * sg * ((1 + gm) * (1 + gm));
    }

    // ---- launch shape, shared-memory kernel: 128 cells per CTA if the state fits, else 64 -----
    m->smem_block = 128;
    m->smem_bytes = moc_smem_bytes(N, 128, b.chunk_max_pairs);
    if (m->smem_bytes > ctx->smem_optin) {
        m->smem_block = 64;
        m->smem_bytes = moc_smem_bytes(N, 64, b.chunk_max_pairs);
    }
    if (m->smem_bytes > ctx->smem_optin) {
        set_error("pbe_moc_create: numberOfClasses %d needs %zu B of shared memory per CTA, device "
                  "offers %zu",
                  N, m->smem_bytes, ctx->smem_optin);
        delete m;
        return PBE_E_UNSUPPORTED;
    }
    moc_kernel_fn fn = kernel_for(m);
    cudaError_t e = raise_dyn_smem(ctx, (const void *)fn, m->smem_bytes);
    if (e == cudaSuccess) {
        int occ = 0;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, (const void *)fn, m->smem_block,
                                                          m->smem_bytes);
        m->smem_ctas_per_sm = std::max(1, occ);
    }
    // ---- tensor-memory kernel (pbe_moc_tmem_kernel.cuh): 128 cells per CTA, accumulators in
    // TMEM, two or more CTAs per SM.  Available when the accumulators fit 256 columns per
    // thread, n[N+1][128] fits in shared memory and (CT only) the H table spans less than 1e6, so
    // that the clamp on K_c cannot change a result (see the kernel's header).  Its table ring
    // holds sub-chunks of at most MOCT_SUB_TILES tiles of one row block.
    {
        MoctSched &sc = m->sched;
        sc.nsubs = 0;
        sc.slot_pairs = 1;
        sc.sub_off[0] = 0;
        bool fits = true;
        for (int blk = 0; blk < b.nblocks && fits; ++blk) {
            sc.blk_sub0[blk] = (unsigned short)sc.nsubs;
            const int ntiles = (b.chunk_off[blk + 1] - b.chunk_off[blk]) / (MOC_TI * MOC_TJ);
            for (int t0 = 0; t0 < ntiles; t0 += MOCT_SUB_TILES) {
                if (sc.nsubs >= MOCT_MAX_SUBS) {
                    fits = false;
                    break;
                }
                const int nt = std::min(MOCT_SUB_TILES, ntiles - t0);
                sc.sub_off[sc.nsubs + 1] = sc.sub_off[sc.nsubs] + nt * MOC_TI * MOC_TJ;
                sc.slot_pairs = std::max(sc.slot_pairs, nt * MOC_TI * MOC_TJ);
                ++sc.nsubs;
            }
        }
        sc.blk_sub0[b.nblocks] = (unsigned short)sc.nsubs;
        m->tmem_cols = fits ? moct_cols_per_thread(N) : 0;
        sc.tmem_cols = m->tmem_cols;
        m->tmem_bytes = moct_smem_bytes(N, sc.slot_pairs);
    }
    double hmax = 0.0, hmin = INFINITY;
    if (m->dcoal == DCOAL_CT)
        for (const double2 &v : tab)
            if (v.x != 0.0) {
                hmax = std::max(hmax, v.y);
                hmin = std::min(hmin, v.y);
            }
    // exp2_neg_tab<LOGT> needs |2^LOGT y| < 2^31: |K log2(e) H| <= 1.45 * 2^(29 - LOGT)
    const double exp_arg_cap = (double)(1 << (29 - MOCT_LOGT));
    b.negK_floor = (hmax > 0.0) ? -exp_arg_cap / hmax : -INFINITY;
    double qmax = 0.0, qmin = INFINITY; // CT breakup exponent table xi^(-5/9)
    if (m->dbr == DBR_CT)
        for (int i = 0; i < N; ++i) {
            qmax = std::max(qmax, cls[(size_t)CLS_BR_Q * (N + 1) + i]);
            qmin = std::min(qmin, cls[(size_t)CLS_BR_Q * (N + 1) + i]);
        }
    b.negKb_floor = (qmax > 0.0) ? -exp_arg_cap / qmax : -INFINITY;
    m->tmem_ok = m->tmem_cols > 0 && m->tmem_bytes <= ctx->smem_optin &&
                 (m->dcoal != DCOAL_CT || (hmax > 0.0 && hmin >= hmax * (1024.0 / exp_arg_cap) && std::isfinite(hmax))) &&
                 (m->dbr != DBR_CT || (qmax > 0.0 && qmin >= qmax * (1024.0 / exp_arg_cap) && std::isfinite(qmax)));
    if (e == cudaSuccess && m->tmem_ok) {
        // Two instantiations: the wide one (two table tiles per step, ~190 registers) is faster
        // per CTA but caps residency at 2 CTAs per SM; it is used where shared memory or tensor
        // memory cap it there anyway (N >~ 60), the narrow one (~160 registers) elsewhere.
        const void *tfn = nullptr;
        cudaFuncAttributes fa{};
        for (int wide = 1; wide >= 0 && e == cudaSuccess; --wide) {
            tfn = (const void *)pick_tmem_kernel(m->dcoal, m->dbr, wide != 0);
            e = raise_dyn_smem(ctx, tfn, m->tmem_bytes);
            if (e == cudaSuccess) e = cudaFuncGetAttributes(&fa, tfn);
            if (e != cudaSuccess) break;
            // cudaOccupancyMaxActiveBlocksPerMultiprocessor answers 1 for a kernel that holds a
            // tcgen05.alloc (it cannot see the column count), so the residency limit is computed
            // here: registers (allocated per warp in units of 256), shared memory (+1 KB the
            // driver reserves per CTA) and the SM's 512 TMEM columns.
            const int regs_per_warp = ((fa.numRegs * 32 + 255) / 256) * 256;
            const int by_regs = 65536 / (regs_per_warp * (MOCT_BLOCK / 32));
            const int by_smem = (int)(ctx->smem_per_sm / (m->tmem_bytes + fa.sharedSizeBytes + 1024));
            const int by_tmem = 512 / m->tmem_cols;
            const int other = std::min(by_smem, std::min(by_tmem, 4));
            m->tmem_ctas_per_sm = std::max(1, std::min(by_regs, other));
            m->tmem_wide = (wide != 0);
            if (by_regs >= other) break; // registers do not cost a CTA: keep this instantiation
        }
        if (e == cudaSuccess) {
            if (const char *f = getenv("PBE_MOC_TMEM_CTAS")) m->tmem_ctas_per_sm = std::max(1, atoi(f));
            // stagger = 1/k of the batch time of a CTA running alone: FP64-pipe instructions per
            // cell (19 per pair, 45 per class) x 2 cycles at the ~60 % pipe use of one warp per
            // scheduler (profiles/r01a)
            double frac = 1.0;
            if (const char *f = getenv("PBE_MOC_STAGGER")) frac = atof(f);
            const double solo = (19.0 * (double)(tab.size()) + 45.0 * N) * 2.0 / 0.6;
            m->sched.stagger_cycles =
                (m->tmem_ctas_per_sm > 1) ? (int)std::min(2.0e9, frac * solo / m->tmem_ctas_per_sm) : 0;
        }
    }
    // auto rule: the tensor-memory kernel wherever it keeps at least as many warps per SM as the
    // shared-memory kernel (it has the shorter instruction stream per pair)
    select_kernel(m, (m->tmem_ok && m->tmem_ctas_per_sm * MOCT_BLOCK >= m->smem_ctas_per_sm * m->smem_block)
                         ? PBE_MOC_KERNEL_TMEM
                         : PBE_MOC_KERNEL_SMEM);
    if (const char *force = getenv("PBE_MOC_KERNEL")) {
        if (strcmp(force, "smem") == 0) select_kernel(m, PBE_MOC_KERNEL_SMEM);
        if (strcmp(force, "tmem") == 0 && m->tmem_ok) select_kernel(m, PBE_MOC_KERNEL_TMEM);
    }
    if (e == cudaSuccess && !tab.empty()) {
        e = cudaMalloc(&m->d_pair, tab.size() * sizeof(double2));
        if (e == cudaSuccess)
            e = cudaMemcpy(m->d_pair, tab.data(), tab.size() * sizeof(double2), cudaMemcpyHostToDevice);
    }
    if (e == cudaSuccess && m->tmem_ok && m->dcoal == DCOAL_CT) {
        // the tensor-memory kernel folds G into the exponent: S = A 2^(log2 G - K log2(e) H).
        // Zero-weight padding slots get {log2 G, H} = {-4096, 0}: 2^y is then 2^-1022 (the clamp of
        // exp2_neg_tab), so a padding slot between two real classes (i > j or i + j > N - 2 inside an
        // edge tile) adds A 2^-1022 n_i n_j < 1e-290 at physical densities instead of exactly 0 as
        // in the shared-memory kernel: 250 orders of magnitude below an ulp of any sum it joins.
        // (An exact 0 needs exponent field 0, i.e. a clamp at -1023, which is unsafe: see pbe_math.cuh.)
        std::vector<double2> tl(tab.size());
        for (size_t k = 0; k < tab.size(); ++k)
            tl[k] = make_double2(tab[k].x > 0.0 ? std::log2(tab[k].x) : -4096.0, tab[k].y);
        e = cudaMalloc(&m->d_pair_log, tl.size() * sizeof(double2));
        if (e == cudaSuccess)
            e = cudaMemcpy(m->d_pair_log, tl.data(), tl.size() * sizeof(double2), cudaMemcpyHostToDevice);
    }
    if (e == cudaSuccess) e = cudaMalloc(&m->d_cls, cls.size() * sizeof(double));
    if (e == cudaSuccess)
        e = cudaMemcpy(m->d_cls, cls.data(), cls.size() * sizeof(double), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        int rc = cuda_fail(e, "pbe_moc_create device setup", __FILE__, __LINE__);
        pbe_moc_destroy(m);
        return rc;
    }
    b.pair_tab = m->d_pair;
    b.cls = m->d_cls;
    (void)resolve_minmax_scratch(ctx->device); // set-up of pbe_moc_class_minmax_noalloc; a failure is reported there
    *out = m;
    return PBE_OK;
}

int pbe_moc_set_kernel(pbe_moc *m, int which) {
    PBE_REQUIRE(m != nullptr, PBE_E_INVALID, "pbe_moc_set_kernel: NULL handle");
    PBE_REQUIRE(which == PBE_MOC_KERNEL_SMEM || which == PBE_MOC_KERNEL_TMEM, PBE_E_INVALID,
                "pbe_moc_set_kernel: unknown kernel id %d", which);
    PBE_REQUIRE(!m->pivot, PBE_E_UNSUPPORTED,
                "pbe_moc_set_kernel: a fixed-pivot grid (pbe_moc_create_pivots) has one kernel");
    PBE_REQUIRE(which != PBE_MOC_KERNEL_TMEM || m->tmem_ok, PBE_E_UNSUPPORTED,
                "pbe_moc_set_kernel: the tensor-memory kernel is not available for numberOfClasses %d "
                "(needs N <= %d, n[N+1][256] in shared memory and an H table spanning < 1e6)",
                m->N, 128 - MOCT_PAD_CLASSES);
    select_kernel(m, which);
    return PBE_OK;
}

// Fixed-pivot method of classes on the class volumes pivots[0] < ... < pivots[N-1] (SURVEY 8 row f4;
// pbe_moc_pivot_kernel.cuh).  Parity unpinned: the reference has no such grid.
int pbe_moc_create_pivots(pbe_ctx *ctx, int N, const double *pivots, const pbe_kernel_spec *spec,
                          double source_relaxation, pbe_moc **out) {
    PBE_REQUIRE(ctx && spec && out && pivots, PBE_E_INVALID, "pbe_moc_create_pivots: NULL argument");
    *out = nullptr;
    PBE_REQUIRE(N >= 1, PBE_E_INVALID, "pbe_moc_create_pivots: numberOfClasses must be >= 1 (got %d)", N);
    PBE_REQUIRE(pivots[0] > 0.0, PBE_E_INVALID, "pbe_moc_create_pivots: class volumes must be positive (got %g)",
                pivots[0]);
    for (int i = 1; i < N; ++i)
        PBE_REQUIRE(pivots[i] > pivots[i - 1] && std::isfinite(pivots[i]), PBE_E_INVALID,
                    "pbe_moc_create_pivots: class volumes must increase (pivot %d = %g after %g)", i, pivots[i],
                    pivots[i - 1]);
    PBE_REQUIRE(spec->coalescence >= PBE_COAL_NONE && spec->coalescence <= PBE_COAL_TABLE, PBE_E_UNKNOWN_TYPE,
                "pbe_moc_create_pivots: unknown coalescenceKernel id %d", spec->coalescence);
    PBE_REQUIRE(spec->breakup >= PBE_BR_NONE && spec->breakup <= PBE_BR_TABLE, PBE_E_UNKNOWN_TYPE,
                "pbe_moc_create_pivots: unknown breakupKernel id %d", spec->breakup);
    PBE_REQUIRE(spec->daughter == PBE_DAUGHTER_UNIFORM_BINARY, PBE_E_UNKNOWN_TYPE,
                "pbe_moc_create_pivots: unknown daughterParticleDistribution id %d", spec->daughter);
    PBE_REQUIRE(spec->coalescence != PBE_COAL_TABLE || spec->coal_table, PBE_E_INVALID,
                "pbe_moc_create_pivots: PBE_COAL_TABLE needs coal_table");
    PBE_REQUIRE(spec->breakup != PBE_BR_TABLE || spec->br_table, PBE_E_INVALID,
                "pbe_moc_create_pivots: PBE_BR_TABLE needs br_table");
    if (spec->coalescence == PBE_COAL_TABLE)
        for (int i = 0; i < N; ++i)
            for (int j = 0; j < i; ++j)
                PBE_REQUIRE(spec->coal_table[(size_t)i * N + j] == spec->coal_table[(size_t)j * N + i],
                            PBE_E_UNSUPPORTED, "pbe_moc_create_pivots: the tabulated coalescence kernel must be symmetric");
    size_t smem = mocp_smem_bytes(N);
    PBE_REQUIRE(smem <= ctx->smem_optin, PBE_E_UNSUPPORTED,
                "pbe_moc_create_pivots: numberOfClasses %d needs %zu B of shared memory per CTA, device offers %zu", N,
                smem, ctx->smem_optin);
    DeviceGuard g(ctx->device);
    pbe_moc *m = new pbe_moc();
    m->ctx = ctx;
    m->N = N;
    m->xi1 = pivots[0];
    m->relax = source_relaxation;
    m->spec = *spec;
    m->spec.coal_table = nullptr;
    m->spec.br_table = nullptr;
    m->pivot = true;
    m->pivots.assign(pivots, pivots + N);
    const std::vector<double> &xi = m->pivots;
    switch (spec->coalescence) {
    case PBE_COAL_NONE: m->dcoal = DCOAL_OFF; break;
    case PBE_COAL_CT: m->dcoal = DCOAL_CT; break;
    default: m->dcoal = DCOAL_SYM; break;
    }
    m->dbr = (spec->breakup == PBE_BR_CT) ? DBR_CT : DBR_TAB;

    // pairs that stay on the grid (the death bound of MOC.C:107-109), i-major; the new drop goes to the
    // pivots that bracket it with the number- and volume-preserving weights
    struct HostPair {
        int i, j, t;
        double a, b, eta;
    };
    std::vector<HostPair> pairs;
    if (m->dcoal != DCOAL_OFF)
        for (int i = 0; i < N; ++i)
            for (int j = i; j < N; ++j) {
                // a sum within a few ulp of a pivot is ON that pivot (so that volumes (i + 1) xi1 that are not
                // exact in binary still reproduce the reference's i + j + 1 rule and death bound)
                const double snap = 8.0 * 2.220446049250313e-16;
                const double v = xi[i] + xi[j], vt = v * (1.0 + snap);
                if (v > xi[N - 1] * (1.0 + snap)) break;
                HostPair pr{};
                pr.i = i;
                pr.j = j;
                pr.t = std::min(N - 1, (int)(std::upper_bound(xi.begin(), xi.end(), vt) - xi.begin()) - 1);
                pr.eta = (pr.t == N - 1) ? 1.0
                                         : std::min(1.0, std::max(0.0, (xi[pr.t + 1] - v) / (xi[pr.t + 1] - xi[pr.t])));
                switch (spec->coalescence) {
                case PBE_COAL_CONSTANT: pr.a = spec->coal_c[0]; break;
                case PBE_COAL_CT: ct_pair_geometry(xi[i], xi[j], &pr.a, &pr.b); break;
                case PBE_COAL_TABLE: pr.a = spec->coal_table[(size_t)i * N + j]; break;
                }
                if (i == j) pr.a *= 0.5; // MOC.C:118
                pairs.push_back(pr);
            }
    // tiles of MOCP_TI x MOCP_TJ pairs (rows i0.., columns j0..) whose births fit one window of MOCP_WIN pivots; a tile
    // whose window would overflow is cut into row tiles, and those into single pairs (any increasing grid works)
    std::vector<FpTile> tiles;
    int win_used = 2; // widest birth window any tile needs
    {
        std::vector<int> pidx((size_t)N * N, -1); // (i, j) -> index into pairs
        for (size_t q = 0; q < pairs.size(); ++q) pidx[(size_t)pairs[q].i * N + pairs[q].j] = (int)q;
        auto emit = [&](int i0, int j0, int r_lo, int r_hi, int c_lo, int c_hi) -> bool { // false: window overflow
            int tmin = N, tmax = -1, cnt = 0;
            for (int r = r_lo; r < r_hi; ++r)
                for (int c = c_lo; c < c_hi; ++c) {
                    const int i = i0 + r, j = j0 + c;
                    if (i >= N || j >= N || pidx[(size_t)i * N + j] < 0) continue;
                    const HostPair &pr = pairs[pidx[(size_t)i * N + j]];
                    tmin = std::min(tmin, pr.t);
                    tmax = std::max(tmax, pr.t);
                    ++cnt;
                }
            if (cnt == 0) return true;
            if (tmax - tmin > MOCP_WIN - 2) return false;
            FpTile T{};
            T.i0 = i0, T.j0 = j0, T.tb = tmin, T.npairs = cnt;
            win_used = std::max(win_used, tmax - tmin + 2);
            for (int r = r_lo; r < r_hi; ++r)
                for (int c = c_lo; c < c_hi; ++c) {
                    const int i = i0 + r, j = j0 + c;
                    if (i >= N || j >= N || pidx[(size_t)i * N + j] < 0) continue;
                    const HostPair &pr = pairs[pidx[(size_t)i * N + j]];
                    const int k = r * MOCP_TJ + c;
                    T.a[k] = pr.a;
                    T.b[k] = pr.b;
                    T.cw[k][pr.t - tmin] = pr.eta;
                    if (pr.t < N - 1) T.cw[k][pr.t - tmin + 1] = 1.0 - pr.eta;
                }
            tiles.push_back(T);
            return true;
        };
        for (int i0 = 0; i0 < N; i0 += MOCP_TI)
            for (int j0 = i0; j0 < N; j0 += MOCP_TJ) {
                if (emit(i0, j0, 0, MOCP_TI, 0, MOCP_TJ)) continue;
                for (int r = 0; r < MOCP_TI; ++r) {
                    if (emit(i0, j0, r, r + 1, 0, MOCP_TJ)) continue;
                    for (int c = 0; c < MOCP_TJ; ++c) emit(i0, j0, r, r + 1, c, c + 1); // a single pair always fits
                }
            }
    }
    while (!tiles.empty() && tiles.size() % MOCP_SLOT_TILES != 0) { // whole ring slots: empty tiles that touch the dummy rows only
        FpTile T{};
        T.i0 = tiles.back().i0;
        T.j0 = N;
        T.tb = N;
        tiles.push_back(T);
    }
    if (pairs.empty()) m->dcoal = DCOAL_OFF;
    m->pivot_win = (win_used <= 6 && !(getenv("PBE_MOC_PIVOT_WIN8"))) ? 6 : MOCP_WIN;

    MocParams &b = m->base;
    b.N = N;
    b.nblocks = 0;
    b.chunk_max_pairs = 1;
    std::vector<double> cls((size_t)CLS_ROWS * (N + 1), 0.0), fpw((size_t)3 * N, 0.0);
    for (int i = 0; i < N; ++i) {
        double P = 0.0, Q = 0.0;
        switch (spec->breakup) {
        case PBE_BR_BINARY: P = xi[i] * xi[i]; break;
        case PBE_BR_CT:
            P = 1.0 / pow(xi[i], 2.0 / 9.0);
            Q = 1.0 / pow(xi[i], 5.0 / 9.0);
            break;
        case PBE_BR_TABLE: P = spec->br_table[i]; break;
        default: break;
        }
        cls[(size_t)CLS_BR_P * (N + 1) + i] = P;
        cls[(size_t)CLS_BR_Q * (N + 1) + i] = Q;
        cls[(size_t)CLS_DVOL * (N + 1) + i] = pow(6.0 / M_PI * xi[i], 1.0 / 3.0);
        cls[(size_t)CLS_XI * (N + 1) + i] = xi[i];
        const double xm = (i > 0) ? xi[i - 1] : 0.0, xp = (i + 1 < N) ? xi[i + 1] : xi[i];
        fpw[i] = 0.5 * (xp - xm);
        fpw[(size_t)N + i] = (xi[i] - xm) / xi[i];
        fpw[(size_t)2 * N + i] = 2.0 / xi[i];
    }
    {
        const double c1 = spec->coal_c[0], c2 = spec->coal_c[1], gm = spec->coal_c[2], sg = spec->coal_c[3];
        b.coalA = c1 / (1 + gm);
        b.coalK = c2 / (sg * sg * ((1 + gm) * (1 + gm) * (1 + gm)));
    }
    {
        const double c1 = spec->br_c[0], c2 = spec->br_c[1], gm = spec->br_c[2], sg = spec->br_c[3];
        b.brA = c1 / (1 + gm);
        b.brK = c2 * sg * ((1 + gm) * (1 + gm));
    }
    b.negK_floor = b.negKb_floor = -INFINITY;
    // accumulators in tensor memory (two CTAs per SM) when they fit 256 columns and two CTAs' shared memory fits the SM
    // (cudaOccupancyMaxActiveBlocksPerMultiprocessor cannot see tcgen05.alloc: the limit is computed here)
    m->tmem_cols = moct_cols_per_thread(N);
    const size_t smem_tm = mocp_smem_bytes(N, true);
    m->use_tmem = m->dcoal != DCOAL_OFF && m->tmem_cols > 0 && 2 * (smem_tm + 1024) <= ctx->smem_per_sm &&
                  2 * m->tmem_cols <= 512;
    if (const char *f = getenv("PBE_MOC_PIVOT_KERNEL")) {
        if (strcmp(f, "smem") == 0) m->use_tmem = false;
    }
    if (m->use_tmem) smem = smem_tm;
    m->block = m->smem_block = MOCP_BLOCK;
    m->smem = m->smem_bytes = smem;
    const void *fn = (const void *)pick_pivot_kernel(m->dcoal, m->dbr, m->use_tmem, m->pivot_win);
    cudaError_t e = raise_dyn_smem(ctx, fn, smem);
    if (e == cudaSuccess) {
        if (m->use_tmem) {
            cudaFuncAttributes fa{};
            e = cudaFuncGetAttributes(&fa, fn);
            const int regs_per_warp = ((fa.numRegs * 32 + 255) / 256) * 256;
            const int by_regs = 65536 / (regs_per_warp * (MOCP_BLOCK / 32));
            m->ctas_per_sm = m->smem_ctas_per_sm = std::max(1, std::min(2, by_regs));
        } else {
            int occ = 0;
            e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, MOCP_BLOCK, smem);
            m->ctas_per_sm = m->smem_ctas_per_sm = std::max(1, occ);
        }
    }
    if (e == cudaSuccess && !tiles.empty()) {
        e = cudaMalloc(&m->d_fp, tiles.size() * sizeof(FpTile));
        if (e == cudaSuccess) e = cudaMemcpy(m->d_fp, tiles.data(), tiles.size() * sizeof(FpTile), cudaMemcpyHostToDevice);
    }
    if (e == cudaSuccess) e = cudaMalloc(&m->d_fpw, fpw.size() * sizeof(double));
    if (e == cudaSuccess) e = cudaMemcpy(m->d_fpw, fpw.data(), fpw.size() * sizeof(double), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMalloc(&m->d_cls, cls.size() * sizeof(double));
    if (e == cudaSuccess) e = cudaMemcpy(m->d_cls, cls.data(), cls.size() * sizeof(double), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        int rc = cuda_fail(e, "pbe_moc_create_pivots device setup", __FILE__, __LINE__);
        pbe_moc_destroy(m);
        return rc;
    }
    b.pair_tab = nullptr;
    b.cls = m->d_cls;
    m->fpx = MocFpExtra{m->d_fp, (int)tiles.size(), m->d_fpw, m->d_fpw + N, m->d_fpw + 2 * N};
    (void)resolve_minmax_scratch(ctx->device); // set-up of pbe_moc_class_minmax_noalloc; a failure is reported there
    *out = m;
    return PBE_OK;
}

// class volumes of the handle's grid: (i + 1) xi1 or the pivots it was created with
int pbe_moc_grid(pbe_moc *m, double *volumes, int *is_uniform) {
    PBE_REQUIRE(m != nullptr, PBE_E_INVALID, "pbe_moc_grid: NULL handle");
    if (volumes)
        for (int i = 0; i < m->N; ++i) volumes[i] = m->pivot ? m->pivots[i] : m->xi1 * (i + 1);
    if (is_uniform) *is_uniform = m->pivot ? 0 : 1;
    return PBE_OK;
}

// DEVICE pointer to the N class volumes (the CLS_XI row of the handle's per-class table): what the continuity check
// and the MULES bound of the transport path read, uniform grid or pivots alike
int pbe_moc_grid_device(pbe_moc *m, const double **volumes_dev) {
    PBE_REQUIRE(m != nullptr && volumes_dev != nullptr, PBE_E_INVALID, "pbe_moc_grid_device: NULL argument");
    *volumes_dev = m->d_cls + (size_t)CLS_XI * (m->N + 1);
    return PBE_OK;
}

int pbe_moc_get_kernel(pbe_moc *m, int *which, int *cells_per_cta, int *ctas_per_sm) {
    PBE_REQUIRE(m != nullptr, PBE_E_INVALID, "pbe_moc_get_kernel: NULL handle");
    if (which) *which = m->pivot ? PBE_MOC_KERNEL_PIVOT : (m->use_tmem ? PBE_MOC_KERNEL_TMEM : PBE_MOC_KERNEL_SMEM);
    if (cells_per_cta) *cells_per_cta = m->block;
    if (ctas_per_sm) *ctas_per_sm = m->ctas_per_sm;
    return PBE_OK;
}

int pbe_moc_num_classes(pbe_moc *m, int *N, pbe_ctx **ctx) {
    PBE_REQUIRE(m != nullptr, PBE_E_INVALID, "pbe_moc_num_classes: NULL handle");
    if (N) *N = m->N;
    if (ctx) *ctx = m->ctx;
    return PBE_OK;
}

int pbe_moc_coeffs(pbe_moc *m, double *xi1, double *source_relaxation) {
    PBE_REQUIRE(m != nullptr, PBE_E_INVALID, "pbe_moc_coeffs: NULL handle");
    if (xi1) *xi1 = m->xi1;
    if (source_relaxation) *source_relaxation = m->relax;
    return PBE_OK;
}

int pbe_moc_destroy(pbe_moc *m) {
    if (!m) return PBE_OK;
    DeviceGuard g(m->ctx->device);
    if (m->d_pair) cudaFree(m->d_pair);
    if (m->d_pair_log) cudaFree(m->d_pair_log);
    if (m->d_cls) cudaFree(m->d_cls);
    if (m->d_fp) cudaFree(m->d_fp);
    if (m->d_fpw) cudaFree(m->d_fpw);
    delete m;
    return PBE_OK;
}

int pbe_moc_source(pbe_moc *m, int64_t ncells, int64_t stride, const double *n, const double *eps,
                   const double *rho, const double *nu, double dt, double *S, void *cuda_stream) {
    int rc = check_fields("pbe_moc_source", m, ncells, stride, n, eps, rho, nu, dt);
    if (rc) return rc;
    PBE_REQUIRE(ncells == 0 || S != nullptr, PBE_E_INVALID, "pbe_moc_source: S is NULL");
    DeviceGuard g(m->ctx->device);
    MocParams p = m->base;
    p.ncells = ncells;
    p.stride = stride;
    p.n_in = n;
    p.n_out = nullptr;
    p.S_out = S;
    p.d_out = nullptr;
    p.eps = eps;
    p.rho = rho;
    p.nu = nu;
    p.alpha = p.vcell = nullptr;
    p.relax = m->relax;
    p.dt = dt;
    p.do_step = 0;
    p.moments = nullptr;
    p.n_moments = 0;
    p.cont_err_max = nullptr;
    return launch(m, p, (cudaStream_t)cuda_stream);
}

int pbe_moc_step(pbe_moc *m, int64_t ncells, int64_t stride, double *n, const double *eps,
                 const double *rho, const double *nu, double dt, const pbe_moc_io *io,
                 void *cuda_stream) {
    int rc = check_fields("pbe_moc_step", m, ncells, stride, n, eps, rho, nu, dt);
    if (rc) return rc;
    DeviceGuard g(m->ctx->device);
    MocParams p = m->base;
    p.ncells = ncells;
    p.stride = stride;
    p.n_in = n;
    p.n_out = n;
    p.eps = eps;
    p.rho = rho;
    p.nu = nu;
    p.relax = m->relax;
    p.dt = dt;
    p.do_step = 1;
    p.S_out = nullptr;
    p.d_out = nullptr;
    p.alpha = p.vcell = nullptr;
    p.moments = nullptr;
    p.n_moments = 0;
    p.cont_err_max = nullptr;
    if (io) {
        PBE_REQUIRE(io->n_moments >= 0 && io->n_moments <= MOC_MAX_MOMENTS, PBE_E_INVALID,
                    "pbe_moc_step: n_moments must be in [0,%d]", MOC_MAX_MOMENTS);
        PBE_REQUIRE(io->cont_err_max == nullptr || (io->alpha && io->vcell), PBE_E_INVALID,
                    "pbe_moc_step: the continuity check needs alpha and vcell");
        p.S_out = io->S;
        p.d_out = io->d;
        p.alpha = io->alpha;
        p.vcell = io->vcell;
        p.moments = (io->n_moments > 0) ? io->moments : nullptr;
        p.n_moments = io->n_moments;
        p.cont_err_max = io->cont_err_max;
    }
    return launch(m, p, (cudaStream_t)cuda_stream);
}

int pbe_moc_d(pbe_moc *m, int64_t ncells, int64_t stride, const double *n, double *d,
              void *cuda_stream) {
    PBE_REQUIRE(m != nullptr, PBE_E_INVALID, "pbe_moc_d: NULL handle");
    PBE_REQUIRE(ncells >= 0 && stride >= ncells, PBE_E_INVALID, "pbe_moc_d: bad ncells/stride");
    if (ncells == 0) return PBE_OK;
    PBE_REQUIRE(n && d, PBE_E_INVALID, "pbe_moc_d: NULL field pointer");
    DeviceGuard g(m->ctx->device);
    const int grid = (int)std::min<int64_t>((ncells + 255) / 256, (int64_t)m->ctx->sm_count * 8);
    moc_d_kernel<<<grid, 256, 0, (cudaStream_t)cuda_stream>>>(
        m->N, ncells, stride, n, m->d_cls + (size_t)CLS_DVOL * (m->N + 1), d);
    PBE_CUDA(cudaGetLastError());
    m->ctx->launches++;
    return PBE_OK;
}

int pbe_moc_class_minmax(pbe_moc *m, int64_t ncells, int64_t stride, const double *n,
                         double *out_min, double *out_max, void *cuda_stream) {
    PBE_REQUIRE(m != nullptr, PBE_E_INVALID, "pbe_moc_class_minmax: NULL handle");
    PBE_REQUIRE(ncells >= 0 && stride >= ncells, PBE_E_INVALID,
                "pbe_moc_class_minmax: bad ncells/stride");
    PBE_REQUIRE(n != nullptr || ncells == 0, PBE_E_INVALID, "pbe_moc_class_minmax: NULL field");
    DeviceGuard g(m->ctx->device);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    const int nparts =
        (int)std::max<int64_t>(1, std::min<int64_t>((ncells + 2047) / 2048, (int64_t)m->ctx->sm_count * 2));
    double *part = nullptr;
    PBE_CUDA(cudaMallocAsync(&part, (size_t)m->N * nparts * 2 * sizeof(double), st));
    moc_minmax_partial_kernel<<<dim3(nparts, m->N), 256, 0, st>>>(ncells, stride, n, part);
    moc_minmax_final_kernel<<<m->N, 32, 0, st>>>(nparts, part, out_min, out_max);
    PBE_CUDA(cudaGetLastError());
    PBE_CUDA(cudaFreeAsync(part, st));
    m->ctx->launches += 2;
    return PBE_OK;
}

// Host-buffer entry point: chunked three-slot pipeline H2D | kernel | D2H.
int pbe_moc_source_host(pbe_moc *m, int64_t ncells, int64_t stride, const double *n_host,
                        const double *eps_host, const double *rho_host, const double *nu_host,
                        double dt, double *S_host) {
    int rc = check_fields("pbe_moc_source_host", m, ncells, stride, n_host, eps_host, rho_host,
                          nu_host, dt);
    if (rc) return rc;
    PBE_REQUIRE(ncells == 0 || S_host != nullptr, PBE_E_INVALID, "pbe_moc_source_host: S is NULL");
    if (ncells == 0) return PBE_OK;
    pbe_ctx *ctx = m->ctx;
    DeviceGuard g(ctx->device);
    const int N = m->N;
    // chunk: whole waves of the persistent grid, about 512 MB of fields per slot
    const int64_t wave = (int64_t)m->block * ctx->sm_count * m->ctas_per_sm;
    const int64_t per_cell = (int64_t)(2 * N + 3) * sizeof(double);
    int64_t C = std::max<int64_t>(1, ((int64_t)512 << 20) / per_cell / wave) * wave;
    C = std::min(C, ((ncells + wave - 1) / wave) * wave);
    rc = ensure_stage(ctx, (size_t)C * per_cell);
    if (rc) return rc;
    // one chunk; a failure leaves the loop so that the streams are drained before returning (the
    // copies already queued still read / write the caller's buffers)
    auto chunk = [&](int64_t c0, int64_t k) -> int {
        const int64_t nc = std::min(C, ncells - c0);
        const int s = (int)(k % kHostSlots);
        cudaStream_t st = ctx->slot_stream[s];
        double *dn = (double *)ctx->stage[s];
        double *dS = dn + (size_t)N * C;
        double *de = dS + (size_t)N * C;
        double *dr = de + C;
        double *dv = dr + C;
        PBE_CUDA(cudaMemcpy2DAsync(dn, (size_t)C * 8, n_host + c0, (size_t)stride * 8, (size_t)nc * 8,
                                   N, cudaMemcpyHostToDevice, st));
        PBE_CUDA(cudaMemcpyAsync(de, eps_host + c0, (size_t)nc * 8, cudaMemcpyHostToDevice, st));
        PBE_CUDA(cudaMemcpyAsync(dr, rho_host + c0, (size_t)nc * 8, cudaMemcpyHostToDevice, st));
        PBE_CUDA(cudaMemcpyAsync(dv, nu_host + c0, (size_t)nc * 8, cudaMemcpyHostToDevice, st));
        int rc2 = pbe_moc_source(m, nc, C, dn, de, dr, dv, dt, dS, st);
        if (rc2) return rc2;
        PBE_CUDA(cudaMemcpy2DAsync(S_host + c0, (size_t)stride * 8, dS, (size_t)C * 8, (size_t)nc * 8,
                                   N, cudaMemcpyDeviceToHost, st));
        return PBE_OK;
    };
    int64_t k = 0;
    for (int64_t c0 = 0; c0 < ncells && rc == PBE_OK; c0 += C, ++k) rc = chunk(c0, k);
    if (rc) {
        drain_slot_streams(ctx);
        return rc;
    }
    for (int s = 0; s < kHostSlots; ++s) PBE_CUDA(cudaStreamSynchronize(ctx->slot_stream[s]));
    return PBE_OK;
}

} // extern "C"

// ---- allocation-free variant of pbe_moc_class_minmax -------------------------------------------------
// No counterpart in the reference: added for callers that must not
// allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
// pbe_moc_class_minmax takes its per-CTA partials from cudaMallocAsync; this variant keeps them in a
// file-scope device array instead: PBE_NOALLOC_MAX_CLASSES x PBE_NOALLOC_MAX_PARTS x {min, max} doubles
// = 1 MiB, held by every GPU that loads the library until the process exits.
//   * In host code `g_minmax_scratch` is a host-side stand-in, NOT a device address: the address is
//     resolved with cudaGetSymbolAddress once per device, with that device current, when a handle is
//     first created on it (pbe_moc_create / pbe_moc_create_pivots), and kept in g_minmax_scratch_dev[].
//   * Every call on a device shares the array and finds the last call's contents in it.  Nothing has to
//     start at zero: the first kernel writes every partial [k][b], b < nparts, that the second one reads.
//   * Calls that overlap on the GPU (two streams, no dependency) race on the array: order them.
static __device__ __align__(256) double g_minmax_scratch[(size_t)PBE_NOALLOC_MAX_CLASSES * PBE_NOALLOC_MAX_PARTS * 2];

namespace {
constexpr int kMaxDevices = 64;
double *g_minmax_scratch_dev[kMaxDevices]; // per device: the array's DEVICE address (nullptr: not resolved)
std::mutex g_minmax_scratch_mu;

// set-up: once per device, that device current (the caller's DeviceGuard)
cudaError_t resolve_minmax_scratch(int device) {
    if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lk(g_minmax_scratch_mu);
    if (g_minmax_scratch_dev[device] != nullptr) return cudaSuccess;
    double *ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void **)&ptr, g_minmax_scratch);
    if (e != cudaSuccess) return e;
    g_minmax_scratch_dev[device] = ptr;
    return cudaSuccess;
}
} // namespace

extern "C" int pbe_moc_class_minmax_noalloc(pbe_moc *m, int64_t ncells, int64_t stride, const double *n,
                                            double *out_min, double *out_max, void *cuda_stream) {
    PBE_REQUIRE(m != nullptr, PBE_E_INVALID, "pbe_moc_class_minmax_noalloc: NULL handle");
    PBE_REQUIRE(ncells >= 0 && stride >= ncells, PBE_E_INVALID, "pbe_moc_class_minmax_noalloc: bad ncells/stride");
    PBE_REQUIRE(n != nullptr || ncells == 0, PBE_E_INVALID, "pbe_moc_class_minmax_noalloc: NULL field");
    PBE_REQUIRE(m->N <= PBE_NOALLOC_MAX_CLASSES, PBE_E_UNSUPPORTED,
                "pbe_moc_class_minmax_noalloc: numberOfClasses %d exceeds the static scratch's limit %d "
                "(use pbe_moc_class_minmax, which allocates)",
                m->N, PBE_NOALLOC_MAX_CLASSES);
    const int dev = m->ctx->device;
    PBE_REQUIRE(dev >= 0 && dev < kMaxDevices && g_minmax_scratch_dev[dev] != nullptr, PBE_E_CUDA,
                "pbe_moc_class_minmax_noalloc: the static scratch of device %d was not resolved at set-up "
                "(cudaGetSymbolAddress failed in pbe_moc_create)", dev);
    DeviceGuard g(dev);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    double *part = g_minmax_scratch_dev[dev];
    const int nparts = (int)std::max<int64_t>(
        1, std::min<int64_t>(std::min<int64_t>((ncells + 2047) / 2048, (int64_t)m->ctx->sm_count * 2),
                             PBE_NOALLOC_MAX_PARTS));
    moc_minmax_partial_kernel<<<dim3(nparts, m->N), 256, 0, st>>>(ncells, stride, n, part);
    moc_minmax_final_kernel<<<m->N, 32, 0, st>>>(nparts, part, out_min, out_max);
    PBE_CUDA(cudaGetLastError());
    m->ctx->launches += 2;
    return PBE_OK;
}
