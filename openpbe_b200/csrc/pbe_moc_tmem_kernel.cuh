// pbe_moc_tmem_kernel.cuh -- method-of-classes source kernel, Blackwell fast path (sm_100a).
//
// Same arithmetic and the same traversal as pbe_moc_kernel.cuh (one thread per cell, each
// unordered class pair evaluated once from the TMA-streamed {G,H} table, TI x TJ register
// tiles), but the per-cell SOURCE ACCUMULATORS live in TENSOR MEMORY instead of shared memory:
//
//   * ncu on the shared-memory kernel (profiles/r01a_moc_ncu_details.txt) shows the FP64 pipe
//     53 % busy with ONE warp per scheduler: 206 KB of shared memory per 128 cells is what caps
//     the occupancy, and every fixed-latency dependency stall is exposed.
//   * TMEM is 128 lanes x 512 32-bit columns per SM and idle on this path (no tensor-core
//     work).  With the 32x32b access shape every thread of a 128-thread CTA owns one lane; a
//     CTA that allocates 256 columns gives each thread 128 FP64 accumulators -- enough for
//     N <= 118 classes -- and TWO such CTAs fit one SM's 512 columns.
//   * Shared memory then only holds the class densities n[N+1][128] (103 KB at N = 100) and the
//     table ring (2 x 3.3 KB): two CTAs = 256 cells = 8 warps per SM, two per scheduler.  The
//     two CTAs are independent, so one is usually inside the FP64-dense pair sweep while the
//     other runs its latency-bound staging loads or epilogue (ncu r01b: with one 256-thread
//     CTA per SM the phases were in lockstep and prologue + epilogue cost 30 % of the time).
//
// Accumulator class c of a thread sits in columns 2c, 2c+1 of its lane.  A tile retires its 4
// column sums and 4 anti-diagonal sums with one tcgen05.ld.32x32b.x8 / tcgen05.st.32x32b.x8
// pair each (SASS LDTM / STTM).  Classes N .. N+9 are scratch bins for the zero-weight padding
// slots of edge tiles, so no index is clamped on the accumulator side.
#pragma once

#include "pbe_moc_kernel.cuh"

namespace pbe {

constexpr int MOCT_BLOCK = 128;       // cells (threads) per CTA
constexpr int MOCT_PAD_CLASSES = 10;  // scratch bins above class N-1 (see header)
constexpr int MOCT_SUB_TILES = 12;    // ring slot = at most 12 tiles (3072 B) of one row block
#ifndef PBE_MOCT_LOGT
#define PBE_MOCT_LOGT 7
#endif
// log2 of the 2^y table size (exp2_neg_tab: 4, 5 or 7).  Measured at N = 100, 6.06e6 cells:
// 16 entries 22.80 ms, 32 entries 21.96 ms, 128 entries 21.08 ms -- the two FP64 instructions the
// shorter polynomial saves per pair outweigh the bank conflicts of the 1 KB table.
constexpr int MOCT_LOGT = PBE_MOCT_LOGT;
#ifndef PBE_MOCT_BIASED_TAB
#define PBE_MOCT_BIASED_TAB 1
#endif
// 2^y through the table staged in shared memory (pbe_math.cuh)
__device__ __forceinline__ double moct_exp2(double y, const double *T) {
#if PBE_MOCT_BIASED_TAB
    return exp2_neg_tabb<MOCT_LOGT>(y, T);
#else
    return exp2_neg_tab<MOCT_LOGT>(y, T);
#endif
}
constexpr int MOCT_MAX_SUBS = 2 * MOC_MAX_BLOCKS;




// ---- TMEM helpers (PTX ISA 8.6 tcgen05; SASS UTCALLOC / LDTM / STTM) -------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t *smem_dst, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(smem_dst)),
                 "r"(ncols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
                 : "memory");
}
__device__ __forceinline__ void tmem_fence_before_sync() {
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tmem_fence_after_sync() {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tmem_wait_st() {
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

// four consecutive FP64 accumulators of this thread's lane
struct Tm4 {
    uint32_t r[8];
};
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, Tm4 &v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(v.r[0]), "=r"(v.r[1]), "=r"(v.r[2]), "=r"(v.r[3]), "=r"(v.r[4]),
                   "=r"(v.r[5]), "=r"(v.r[6]), "=r"(v.r[7])
                 : "r"(taddr)
                 : "memory");
}
// tcgen05.wait::ld with the loaded registers as in/out operands: a true dependency, so no
// consumer of v can be scheduled above the wait
__device__ __forceinline__ void tmem_wait_ld(Tm4 &v) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(v.r[0]), "+r"(v.r[1]), "+r"(v.r[2]), "+r"(v.r[3]), "+r"(v.r[4]),
                   "+r"(v.r[5]), "+r"(v.r[6]), "+r"(v.r[7])::"memory");
}
__device__ __forceinline__ void tmem_wait_ld(Tm4 &a, Tm4 &b) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(a.r[0]), "+r"(a.r[1]), "+r"(a.r[2]), "+r"(a.r[3]), "+r"(a.r[4]),
                   "+r"(a.r[5]), "+r"(a.r[6]), "+r"(a.r[7]), "+r"(b.r[0]), "+r"(b.r[1]),
                   "+r"(b.r[2]), "+r"(b.r[3]), "+r"(b.r[4]), "+r"(b.r[5]), "+r"(b.r[6]),
                   "+r"(b.r[7])::"memory");
}
__device__ __forceinline__ void tmem_st4(uint32_t taddr, const Tm4 &v) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::
                     "r"(taddr),
                 "r"(v.r[0]), "r"(v.r[1]), "r"(v.r[2]), "r"(v.r[3]), "r"(v.r[4]), "r"(v.r[5]),
                 "r"(v.r[6]), "r"(v.r[7])
                 : "memory");
}
__device__ __forceinline__ double tm_get(const Tm4 &v, int k) {
    return __hiloint2double((int)v.r[2 * k + 1], (int)v.r[2 * k]);
}
__device__ __forceinline__ void tm_set(Tm4 &v, int k, double x) {
    v.r[2 * k] = (uint32_t)__double2loint(x);
    v.r[2 * k + 1] = (uint32_t)__double2hiint(x);
}
// acc[c0 .. c0+3] += sgn * add[0..3]   (v holds the loaded accumulators)
__device__ __forceinline__ void tm_axpy4(Tm4 &v, double sgn, const double *add) {
#pragma unroll
    for (int k = 0; k < 4; ++k) tm_set(v, k, fma_(sgn, add[k], tm_get(v, k)));
}

// v[0..NV-1] hold 4 NV consecutive accumulators (a window starting at class j0): window[OFF + w] +=
// add[w] for w < NW, with every index a compile-time constant once unrolled
template <int OFF, int NW, int NV>
__device__ __forceinline__ void tm_window_add(Tm4 (&v)[NV], const double *add) {
    static_assert(OFF + NW <= 4 * NV, "window too small");
#pragma unroll
    for (int w = 0; w < NW; ++w) {
        const int idx = OFF + w;
        tm_set(v[idx >> 2], idx & 3, tm_get(v[idx >> 2], idx & 3) + add[w]);
    }
}

// launch-time schedule of the table ring: row blocks cut into sub-chunks of <= MOCT_SUB_TILES tiles
struct MoctSched {
    int nsubs;                                   // sub-chunks per cell batch
    int slot_pairs;                              // ring slot size in pairs
    int tmem_cols;                               // TMEM columns this CTA allocates (power of 2)
    int stagger_cycles;                          // start-up delay per co-resident CTA slot
    int sub_off[MOCT_MAX_SUBS + 1];              // pair offset of every sub-chunk in pair_tab
    unsigned short blk_sub0[MOC_MAX_BLOCKS + 1]; // first sub-chunk of every row block
};

// shared-memory footprint of one CTA of the TMEM kernel
__host__ __device__ inline size_t moct_smem_bytes(int N, int slot_pairs) {
    size_t b = 0;
    b += (size_t)(N + 1) * MOCT_BLOCK * sizeof(double); // n_s (row N = zero dummy class)
    b += (size_t)2 * slot_pairs * sizeof(double2);      // table ring
    b += (size_t)CLS_ROWS * (N + 1) * sizeof(double);   // per-class tables
    b += 2 * sizeof(uint64_t) + 16;                     // mbarriers, TMEM base address
    b += (size_t)(1 << MOCT_LOGT) * sizeof(double);     // 2^(j/2^LOGT) table of exp2_neg_tab
    return b;
}
// TMEM columns per thread: 2 per accumulator, N + pad classes rounded up to whole x8 groups,
// then to a power of two >= 32 (allocation granularity); 0 if it does not fit
__host__ __device__ inline int moct_cols_per_thread(int N) {
    const int need = ((2 * (N + MOCT_PAD_CLASSES) + 7) / 8) * 8;
    int c = 32;
    while (c < need) c *= 2;
    return c <= 256 ? c : 0;
}

__device__ __forceinline__ double ldg_stream(const double *p) {
    // streaming load that does not allocate in L1
    double v;
    asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
    return v;
}

// ---- one TI x TJ tile of class pairs (register tile; identical sums to moc_tile) ---------------
// K2 = -K_c log2(e): the Coulaloglou-Tavlarides kernel value G_ij exp(-K_c H_ij) is
// 2^(log2 G_ij + K2 H_ij) (coalescenceKernels/CoulaloglouTavlarides.C:77-88); the CT table holds
// {log2 G, H} and T16 is the 2^(j/2^LOGT) table of exp2_neg_tab: 13 FP64-pipe instructions per pair.
// Rows R0 .. R1-1 of the tile only: the kernel issues its tensor-memory traffic between the two
// halves of a tile (see the pair sweep).
template <int COAL, int R0 = 0, int R1 = MOC_TI, int C0 = 0, int NJ = MOC_TJ>
__device__ __forceinline__ void moct_tile(const double2 *__restrict__ tab, double K2,
                                          const double *__restrict__ T16,
                                          const double (&ni)[MOC_TI], const double (&nj)[NJ],
                                          double (&di)[MOC_TI], double (&dj)[NJ],
                                          double (&bw)[MOC_TI + NJ - 1]) {
    // C0: column offset of this table tile inside a double-width register tile (NJ = 2 TJ)
#pragma unroll
    for (int r = R0; r < R1; ++r) {
#pragma unroll
        for (int c = 0; c < MOC_TJ; ++c) {
            const double2 gh = tab[r * MOC_TJ + c];
            const double w = ni[r] * nj[C0 + c];
            if (COAL == DCOAL_CT) {
                const double eg = moct_exp2(fma_(K2, gh.y, gh.x), T16);
                di[r] = fma_(eg, w, di[r]);
                dj[C0 + c] = fma_(eg, w, dj[C0 + c]);
                bw[r + C0 + c] = fma_(eg, w, bw[r + C0 + c]);
            } else if (COAL == DCOAL_SYM) {
                di[r] = fma_(gh.x, w, di[r]);
                dj[C0 + c] = fma_(gh.x, w, dj[C0 + c]);
                bw[r + C0 + c] = fma_(gh.x, w, bw[r + C0 + c]);
            } else { // DCOAL_ASYM: x = S(xi_i, xi_j), y = S(xi_j, xi_i)
                const double t1 = gh.x * w, t2 = gh.y * w;
                di[r] += t1;
                dj[C0 + c] += t2;
                bw[r + C0 + c] += 0.5 * (t1 + t2);
            }
        }
    }
}

// ---- the kernel: two persistent CTAs per SM, 128 cells per batch ----------------------------------
// p.negK_floor (MocParams) = -2^(29-LOGT) / max H: with it 2^LOGT y of exp2_neg_tab stays inside
// an int32; the host only selects this kernel when every H_ij >= 2^(LOGT-19) max H, so a clamped
// cell has every efficiency below 2^-1024 either way.
// The narrow instantiation is the one used where three CTAs fit an SM: its registers are capped for that.
template <int COAL, int BR, bool WIDE>
__global__ void __launch_bounds__(MOCT_BLOCK, WIDE ? 2 : 3)
    moc_tmem_kernel(const __grid_constant__ MocParams p, const __grid_constant__ MoctSched sc) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int BLOCK = MOCT_BLOCK;
    const int N = p.N;
    const int tid = threadIdx.x;
    double2 *ring = reinterpret_cast<double2 *>(smem_raw);
    double *n_s = reinterpret_cast<double *>(ring + 2 * (size_t)sc.slot_pairs);
    double *cls_s = n_s + (size_t)(N + 1) * BLOCK;
    uint64_t *bars = reinterpret_cast<uint64_t *>(cls_s + CLS_ROWS * (N + 1));
    uint32_t *tmem_base_s = reinterpret_cast<uint32_t *>(bars + 2);
    double *T16 = reinterpret_cast<double *>(bars + 4); // 16-byte aligned: bars + 2 is, + 16 B

    const int64_t nbatches = (p.ncells + BLOCK - 1) / BLOCK;
    const int64_t my_batches =
        (nbatches > (int64_t)blockIdx.x) ? (nbatches - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const int64_t total_q = (COAL == DCOAL_OFF) ? 0 : my_batches * sc.nsubs;

    // ---- tensor-memory accumulators: warp 0 allocates, everybody reads the base address ---------
    if (tid < 32) tmem_alloc(tmem_base_s, (uint32_t)sc.tmem_cols);
    if (tid == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        mbar_fence_init();
    }
    for (int k = tid; k < CLS_ROWS * (N + 1); k += BLOCK) cls_s[k] = p.cls[k];
    n_s[N * BLOCK + tid] = 0.0; // dummy class: absorbs padded table slots
#if PBE_MOCT_BIASED_TAB
    if (tid < (1 << MOCT_LOGT)) T16[tid] = exp2_tab_bias(PBE_EXP2_TAB128[tid << (7 - MOCT_LOGT)], tid, MOCT_LOGT);
#else
    if (tid < (1 << MOCT_LOGT)) T16[tid] = PBE_EXP2_TAB128[tid << (7 - MOCT_LOGT)];
#endif
    tmem_fence_before_sync();
    __syncthreads();
    tmem_fence_after_sync();
    const uint32_t tmem_base = *tmem_base_s;
    // lane quarter of this warp in bits 31:16; the thread's accumulators start at column 0
    const uint32_t tacc = tmem_base + ((uint32_t)((tid >> 5) * 32) << 16);
    const int n_acc_groups = (N + MOCT_PAD_CLASSES + 3) / 4;

    // Co-resident CTAs start together and every batch costs the same, so left alone they stay
    // in lockstep: staging loads and epilogues of both would coincide and leave the FP64 pipe
    // idle.  The TMEM allocator hands the first CTA of an SM columns [0, cols), the second
    // [cols, 2 cols) ...: that slot index staggers the start by a fraction of a batch, once.
    if (sc.stagger_cycles > 0 && my_batches >= 2) {
        const uint32_t slot = (tmem_base & 0xFFFFu) / (uint32_t)sc.tmem_cols;
        if (slot != 0 && tid == 0) {
            const long long t0 = clock64();
            const long long wait = (long long)sc.stagger_cycles * (long long)slot;
            while (clock64() - t0 < wait) __nanosleep(200);
        }
        __syncthreads();
    }

    auto issue = [&](int64_t q) {
        const int s = (int)(q % sc.nsubs);
        const int off = sc.sub_off[s];
        const unsigned bytes = (unsigned)(sc.sub_off[s + 1] - off) * (unsigned)sizeof(double2);
        uint64_t *bar = &bars[q & 1];
        mbar_expect_tx(bar, bytes);
        bulk_g2s(ring + (q & 1) * (size_t)sc.slot_pairs, p.pair_tab + off, bytes, bar);
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
    // per-class moment work: none / sum and volume only (continuity check) / all eight powers
    const int mom_mode = !want_diag ? 0 : ((p.moments == nullptr || p.n_moments <= 2) ? 1 : 2);
    const double inv_dt = 1.0 / p.dt;

    uint64_t pol_ef; // L2 policy of the staging loads: a line is read once per step
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_ef));
    int64_t q = 0;
    for (int64_t batch = blockIdx.x; batch < nbatches; batch += gridDim.x) {
#ifdef PBE_PHASE_TIMING // scratch builds only (tools/moc_phase_time.py): cycles per phase
        const long long tp0 = clock64();
#endif
        const int64_t cell = batch * BLOCK + tid;
        const bool live = cell < p.ncells;
        const int64_t cc = live ? cell : p.ncells - 1;

        // stage this cell's class densities (coalesced: lane = cell) with asynchronous 8-byte
        // global->shared copies (LDGSTS): all N are in flight at once, no register round trip;
        // they are waited for just before the pair sweep, behind the accumulator zeroing and the
        // per-cell coefficients.  The next batch's lines are pulled into L2 meanwhile (one
        // request per 32-B sector); they have to survive one batch period there, so everything
        // else this kernel moves through L2 is marked dead on arrival: these loads carry an
        // evict-first policy and the epilogue's stores are streaming (st.global.cs).  Measured at
        // 6.06e6 cells (algorithmic reads 5.04 GB): plain loads and stores 6.46 GB read, 20.69 ms;
        // streaming stores 6.06 GB; both 5.19 GB, 20.60 ms; no prefetch at all 5.05 GB, 20.83 ms.
        {
            const double *src = p.n_in + cc;
            const int64_t ncell = cell + (int64_t)gridDim.x * BLOCK;
            const bool pf = ((tid & 3) == 0) && (ncell < p.ncells);
            uint32_t dst = smem_u32(n_s + tid);
#pragma unroll 4
            for (int i = 0; i < N; ++i, dst += BLOCK * (uint32_t)sizeof(double)) {
                asm volatile("cp.async.ca.shared.global.L2::cache_hint [%0], [%1], 8, %2;" ::"r"(dst),
                             "l"(src + (size_t)i * p.stride), "l"(pol_ef)
                             : "memory");
                if (pf)
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(p.n_in + (size_t)i * p.stride + ncell));
            }
        }
        // zero the accumulators (classes 0 .. N+pad, whole x8 groups)
        if (COAL != DCOAL_OFF) {
            Tm4 z;
#pragma unroll
            for (int k = 0; k < 8; ++k) z.r[k] = 0u;
            for (int g = 0; g < n_acc_groups; ++g) tmem_st4(tacc + 8u * (uint32_t)g, z);
        }
        const double eps = __ldg(p.eps + cc), rho = __ldg(p.rho + cc), nu = __ldg(p.nu + cc);
        const double eps13 = cbrt(eps);
        double A = 1.0, K2 = 0.0;
        if (COAL == DCOAL_CT) {
            A = p.coalA * eps13;
            double negK = -(p.coalK * rho * eps * (rho * nu));
            negK = (negK < p.negK_floor) ? p.negK_floor : negK; // NaN stays NaN
            K2 = negK * 1.4426950408889634074;
        }

        asm volatile("cp.async.wait_all;" ::: "memory"); // n_s column of this thread is complete
#ifdef PBE_PHASE_TIMING
        const long long tp1 = clock64();
#endif
        // ---- pair sweep: row blocks of TI classes, each in <= MOCT_SUB_TILES-tile ring slots ------
        if (COAL != DCOAL_OFF) {
#pragma unroll 1
            for (int blk = 0; blk < p.nblocks; ++blk) {
                const int i0 = blk * MOC_TI;
                // row block 0: the anti-diagonal targets i0+j0+1.. overlap the column targets j0..
                const bool alias = (i0 + 1 < MOC_TJ);
                static_assert(MOC_TI == MOC_TJ, "alias <=> i0 == 0 (window offset 1) needs TI == TJ");
                // The wide instantiation walks the table two tiles at a time (a TI x 2TJ register
                // tile: 32 independent pair chains, one tensor-memory round trip and one loop trip
                // per 32 pairs).  From i0 = 2 TJ - 1 on, the 2TJ column targets j0.. and the 2TJ
                // anti-diagonal targets i0+j0+1.. are disjoint (four x8 loads at two places);
                // for the first two row blocks (i0 = 0, TI) they overlap, so ONE 16-class window
                // starting at j0 is loaded and both updates are applied to it in registers.
                const bool wide = WIDE;
                const int wmode = (i0 + 1 >= 2 * MOC_TJ) ? 0 : (i0 == 0 ? 1 : 2);
                static_assert(!WIDE || (MOC_TI == 4 && MOC_TJ == 4), "window offsets below assume 4 x 4 tiles");
                constexpr int NJ2 = 2 * MOC_TJ, NB2 = MOC_TI + 2 * MOC_TJ - 1;
                double ni[MOC_TI], di[MOC_TI], bw[NB2];
#pragma unroll
                for (int r = 0; r < MOC_TI; ++r) {
                    ni[r] = n_s[min(i0 + r, N) * BLOCK + tid];
                    di[r] = 0.0;
                }
#pragma unroll
                for (int w = 0; w < NB2; ++w) bw[w] = 0.0;
                int j0 = i0;
#pragma unroll 1
                for (int s = sc.blk_sub0[blk]; s < (int)sc.blk_sub0[blk + 1]; ++s, ++q) {
                    mbar_wait(&bars[q & 1], (unsigned)((q >> 1) & 1));
                    const double2 *__restrict__ tab = ring + (q & 1) * (size_t)sc.slot_pairs;
                    const int ntiles = (sc.sub_off[s + 1] - sc.sub_off[s]) / (MOC_TI * MOC_TJ);
                    int t = 0;
                    if (wide) {
#pragma unroll 1
                        for (; t + 1 < ntiles; t += 2, j0 += NJ2, tab += 2 * MOC_TI * MOC_TJ) {
                            double nj[NJ2], dj[NJ2];
#pragma unroll
                            for (int c = 0; c < NJ2; ++c) {
                                nj[c] = n_s[min(j0 + c, N) * BLOCK + tid];
                                dj[c] = 0.0;
                            }
                            const uint32_t a_dj = tacc + 2u * (uint32_t)j0;
                            const uint32_t a_bw = tacc + 2u * (uint32_t)(i0 + j0 + 1);
                            // four x8 groups: columns + anti-diagonals (wmode 0) or one window
                            const uint32_t a2 = (wmode == 0) ? a_bw : a_dj + 16u;
                            Tm4 v[4];
                            // the first table tile covers the latency of the previous step's
                            // stores, the second that of this step's loads
                            moct_tile<COAL, 0, MOC_TI, 0, NJ2>(tab, K2, T16, ni, nj, di, dj, bw);
                            tmem_wait_st();
                            tmem_ld4(a_dj, v[0]);
                            tmem_ld4(a_dj + 8u, v[1]);
                            tmem_ld4(a2, v[2]);
                            tmem_ld4(a2 + 8u, v[3]);
                            moct_tile<COAL, 0, MOC_TI, MOC_TJ, NJ2>(tab + MOC_TI * MOC_TJ, K2, T16, ni, nj,
                                                                    di, dj, bw);
                            tmem_wait_ld(v[0], v[1]);
                            tmem_wait_ld(v[2], v[3]);
                            tm_axpy4(v[0], -1.0, dj); // death of class j
                            tm_axpy4(v[1], -1.0, dj + MOC_TJ);
                            // birth of class i+j+1: the 2TJ lowest anti-diagonals of the window
                            if (wmode == 0) {
                                tm_axpy4(v[2], 1.0, bw);
                                tm_axpy4(v[3], 1.0, bw + MOC_TJ);
                            } else if (wmode == 1) {
                                tm_window_add<1, 2 * MOC_TJ, 4>(v, bw); // i0 = 0: classes j0+1 ..
                            } else {
                                tm_window_add<MOC_TI + 1, 2 * MOC_TJ, 4>(v, bw); // i0 = TI: classes j0+5 ..
                            }
                            tmem_st4(a_dj, v[0]);
                            tmem_st4(a_dj + 8u, v[1]);
                            tmem_st4(a2, v[2]);
                            tmem_st4(a2 + 8u, v[3]);
#pragma unroll
                            for (int w = 0; w < NB2; ++w) bw[w] = (w + NJ2 < NB2) ? bw[w + NJ2] : 0.0;
                        }
                    }
#pragma unroll 1
                    for (; t < ntiles; ++t, j0 += MOC_TJ, tab += MOC_TI * MOC_TJ) {
                        double nj[NJ2], dj[NJ2];
#pragma unroll
                        for (int c = 0; c < NJ2; ++c) {
                            nj[c] = (c < MOC_TJ) ? n_s[min(j0 + c, N) * BLOCK + tid] : 0.0;
                            dj[c] = 0.0;
                        }
                        const uint32_t a_dj = tacc + 2u * (uint32_t)j0;
                        const uint32_t a_bw = tacc + 2u * (uint32_t)(i0 + j0 + 1);
                        // row block 0: the TJ anti-diagonal targets j0+1.. overlap the column
                        // targets j0..: one 2TJ-class window takes both updates
                        const uint32_t a1 = alias ? a_dj + 8u : a_bw;
                        Tm4 v[2];
                        // first half of the tile covers the latency of the previous tile's
                        // stores, the second half that of this tile's loads
                        moct_tile<COAL, 0, MOC_TI / 2, 0, NJ2>(tab, K2, T16, ni, nj, di, dj, bw);
                        tmem_wait_st(); // the previous tile's stores have landed
                        tmem_ld4(a_dj, v[0]);
                        tmem_ld4(a1, v[1]);
                        moct_tile<COAL, MOC_TI / 2, MOC_TI, 0, NJ2>(tab, K2, T16, ni, nj, di, dj, bw);
                        tmem_wait_ld(v[0], v[1]);
                        tm_axpy4(v[0], -1.0, dj); // death of class j
                        if (!alias)
                            tm_axpy4(v[1], 1.0, bw); // birth of class i+j+1: TJ lowest anti-diagonals
                        else
                            tm_window_add<1, MOC_TJ, 2>(v, bw);
                        tmem_st4(a_dj, v[0]);
                        tmem_st4(a1, v[1]);
#pragma unroll
                        for (int w = 0; w < NB2; ++w) bw[w] = (w + MOC_TJ < NB2) ? bw[w + MOC_TJ] : 0.0;
                    }
                    // every warp is done with ring slot q&1.  (Measured: releasing the slot per warp through
                    // an "empty" mbarrier instead -- no CTA-wide barrier, warps free to drift by one
                    // sub-chunk -- is 6 % SLOWER, 57.9 vs 54.4 ms at 1.6e7 cells.)
                    __syncthreads();
                    if (tid == 0 && q + 2 < total_q) issue(q + 2);
                }
                // flush the TI-1 anti-diagonals still in the window and the row sums (death of
                // class i); the targets never overlap (j0 >= i0 + TJ here)
                {
                    const uint32_t a_k = tacc + 2u * (uint32_t)(i0 + j0 + 1);
                    const uint32_t a_i = tacc + 2u * (uint32_t)i0;
                    Tm4 vk, vi;
                    tmem_wait_st();
                    tmem_ld4(a_k, vk);
                    tmem_ld4(a_i, vi);
                    tmem_wait_ld(vk, vi);
                    tm_axpy4(vk, 1.0, bw); // bw[TI-1 ..] are zero after the shift
                    tm_axpy4(vi, -1.0, di);
                    tmem_st4(a_k, vk);
                    tmem_st4(a_i, vi);
                }
            }
            tmem_wait_st();
        }

#ifdef PBE_PHASE_TIMING
        const long long tp2 = clock64();
#endif
        // ---- epilogue: breakup (suffix sum), limiter, update, d, diagnostics ----------------
        double Ab = 0.0, Kb2 = 0.0;
        if (BR == DBR_CT) {
            Ab = p.brA * eps13;
            double negKb = -(p.brK / (rho * (eps13 * eps13)));
            negKb = (negKb < p.negKb_floor) ? p.negKb_floor : negKb; // NaN stays NaN
            Kb2 = negKb * 1.4426950408889634074;
        }
        double R = 0.0, sum_d = 0.0, sum_n = 0.0;
        double mloc[MOC_MAX_MOMENTS];
#pragma unroll
        for (int k = 0; k < MOC_MAX_MOMENTS; ++k) mloc[k] = 0.0;
        const bool wS = live && (p.S_out != nullptr);
        const bool wN = live && (p.do_step != 0);
        const int c4top = ((N - 1) >> 2) << 2;
        // output pointers walk down one class row per class (no 64-bit multiply per store)
        double *pS = (p.S_out != nullptr ? p.S_out : p.n_out) + (size_t)(c4top + 3) * p.stride + cc;
        double *pN = (p.do_step ? p.n_out : pS) + (p.do_step ? (size_t)(c4top + 3) * p.stride + cc : 0);

        // Four classes per tensor-memory load, written as straight-line code over the group: the
        // four breakup exponentials, limiter tests and moment chains are independent and
        // interleave (two warps per scheduler leave little else to hide latency); only the suffix
        // sum R runs serially, and the limiter's true division sits behind ONE rare branch per
        // group instead of one per class.
        auto group_body = [&](const int c4, const Tm4 &v, const bool full) {
            double ni4[4], s4[4], num4[4];
            bool need = false, need4[4];
#pragma unroll
            for (int r = 3; r >= 0; --r) {
                const int i = c4 + r;
                const bool on = full || (i < N);
                const int ii = on ? i : N; // row N: zero dummy class, zero table entries
                const double ni = n_s[ii * BLOCK + tid];
                // breakupCache_ (MOC.C:165-171): CT = Ab xi^(-2/9) exp(-Kb xi^(-5/9))
                double g = cBP[ii];
                if (BR == DBR_CT) g = (Ab * g) * moct_exp2(Kb2 * cBQ[ii], T16);
                const double gn = g * ni;
                const double B = R - gn; // -g_i n_i + sum_{j>i} beta g_j n_j dxi (MOC.C:142-155)
                R = fma_(cW[ii], gn, R);
                const double s = fma_(A, tm_get(v, r), B);
                // MOC.C:87-91: max(s, -relax n_i / dt).  The reference divides; a product with
                // 1/dt is within 2 ulp of that, so the true division (bit-identical limited
                // values) is only evaluated when s is within 1e-15 of the limit.
                const double num = __dmul_rn(-p.relax, ni);
                const double lim0 = num * inv_dt;
                need4[r] = on && !(s > fma_(fabs(lim0), 1e-15, lim0));
                need |= need4[r];
                ni4[r] = ni;
                s4[r] = s;
                num4[r] = num;
            }
            if (need) {
#pragma unroll
                for (int r = 0; r < 4; ++r)
                    if (need4[r]) {
                        const double lim = num4[r] / p.dt;
                        s4[r] = (s4[r] > lim) ? s4[r] : lim;
                    }
            }
#pragma unroll
            for (int r = 3; r >= 0; --r) {
                const int i = c4 + r;
                const bool on = full || (i < N);
                const int ii = on ? i : N;
                const double s = s4[r];
                if (wS && on) __stcs(pS, s);
                // n += dt S (MOC.C:259-272, phi = 0)
                double nn = p.do_step ? __dadd_rn(ni4[r], __dmul_rn(p.dt, s)) : ni4[r];
                nn = on ? nn : 0.0; // classes past N-1 of the last group do not exist
                if (wN && on) __stcs(pN, nn);
                pS -= p.stride;
                pN -= p.stride;
                sum_d = fma_(cDV[ii], nn, sum_d);
                sum_n += nn;
                ni4[r] = nn; // reused: the updated density, for the moment sums below
            }
            // moment sums of the four classes behind ONE uniform branch (a branch per class
            // would cut the group into basic blocks and serialise the classes)
            if (mom_mode == 1) {
#pragma unroll
                for (int r = 3; r >= 0; --r) {
                    const int ii = (full || (c4 + r < N)) ? c4 + r : N;
                    mloc[0] += ni4[r];
                    mloc[1] = fma_(cXI[ii], ni4[r], mloc[1]);
                }
            } else if (mom_mode == 2) {
#pragma unroll
                for (int r = 3; r >= 0; --r) {
                    // xi^k n for k = 0..7 as two chains (even / odd powers) of depth 3
                    const int ii = (full || (c4 + r < N)) ? c4 + r : N;
                    const double x = cXI[ii], x2 = x * x;
                    double pe = ni4[r], po = ni4[r] * x;
#pragma unroll
                    for (int k = 0; k < MOC_MAX_MOMENTS; k += 2) {
                        mloc[k] += pe;
                        mloc[k + 1] += po;
                        pe *= x2;
                        po *= x2;
                    }
                }
            }
        };
        // the accumulators of the next group are fetched from tensor memory while this group is
        // being worked on
        Tm4 v, vn;
#pragma unroll
        for (int k = 0; k < 8; ++k) v.r[k] = vn.r[k] = 0u;
        if (COAL != DCOAL_OFF) tmem_ld4(tacc + 2u * (uint32_t)c4top, v);
#pragma unroll 1
        for (int c4 = c4top; c4 >= 0; c4 -= 4) {
            if (COAL != DCOAL_OFF) {
                tmem_wait_ld(v);
                if (c4 >= 4) tmem_ld4(tacc + 2u * (uint32_t)(c4 - 4), vn);
            }
            if (c4 + 3 < N)
                group_body(c4, v, true);
            else
                group_body(c4, v, false);
            v = vn;
        }
#ifdef PBE_PHASE_TIMING
        if (tid == 0 && p.moments != nullptr) { // overwrites moments 5..7 with phase cycles
            const long long tp3 = clock64();
            atomicAdd(p.moments + 5, (double)(tp1 - tp0));
            atomicAdd(p.moments + 6, (double)(tp2 - tp1));
            atomicAdd(p.moments + 7, (double)(tp3 - tp2));
        }
#endif
        if (p.d_out != nullptr && live) p.d_out[cell] = sum_d / sum_n; // MOC.C:282-290
        if (want_diag && live) {
            const double V = (p.vcell != nullptr) ? __ldg(p.vcell + cell) : 1.0;
#pragma unroll
            for (int k = 0; k < MOC_MAX_MOMENTS; ++k) macc[k] = fma_(V, mloc[k], macc[k]);
            if (p.cont_err_max != nullptr && p.alpha != nullptr) {
                const double e = fabs(mloc[1] / V - __ldg(p.alpha + cell)); // MOC.C:254-279
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
    tmem_fence_before_sync();
    __syncthreads();
    if (tid < 32) tmem_dealloc(tmem_base, (uint32_t)sc.tmem_cols);
}

} // namespace pbe
