// pbe_transport.cu -- explicit transport of class / moment fields on a face-addressed mesh
// (first slice).
//
// What it replaces: the update at the end of MOC::solveWithMULES (src/PBESystem/MOC.C:198-272)
// and QMOM::solveWithMULES (src/PBESystem/QMOM.C:131-197),
//     phiNk = lambda (phiHO - phiUpwind) + upwind(mesh, phi).flux(Nk);
//     MULES::explicitSolve(geometricOneField(), Nk, phiNk, zeroField(), S[k]);
// for the divergence scheme `Gauss upwind`, where the limited correction is identically zero.
// The vanLeer / MULES-limited correction is the next slice.
//
// Device layout.  pbe_mesh_create turns OpenFOAM's face addressing (owner / neighbour lists)
// into cell-centred tables whose entries are in ascending face order per cell -- the order
// fvc::surfaceIntegrate's face loops add them in -- so per-cell sums are bit-identical to a face
// loop.  Cells are split in two sets:
//   * regular cells: exactly F faces (F = the mesh's most common count: 6 for hexahedra, 4 for
//     tetrahedra, 5 for prisms), none of them on a processor patch.  ELL table [F][ncells], cell
//     fastest; one thread per cell keeps its F (neighbour, signed flux, lambda) triples in
//     registers and sweeps a chunk of fields with F + 2 loads, 4 F + 9 FP64 instructions and no
//     branch per field.  A boundary face is a slot like any other whose value comes from the
//     patch row psi_b (or from the cell itself: zero gradient); warps without such a slot -- the
//     bulk -- run a loop that does not know about them.  (Until r02 boundary cells went to the
//     list below: scattered 8-byte reads of 32-byte sectors, 1.9 GB of DRAM reads for 0.24 GB of
//     cell data and 0.55 of the 3.3 ms of a 160^3 x 100 update.)
//   * everything else (other face counts, processor-patch cells): a compact CSR list, one thread
//     per listed cell, any number of faces.
// HBM-bound: 24 B per cell-field (psi and S in, psi out); neighbour gathers are L1 / L2 hits.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <new>

#include "pbe_common.h"

// Resident 256-thread CTAs per SM the regular-cell kernel is compiled for (64 registers).  Measured
// on B200, 160^3 cells x 100 fields: 4 CTAs 3.70 ms, 3 CTAs (80 registers) 3.80 ms; an L2 prefetch
// of the streamed rows four fields ahead changes nothing (the loop is issue- and latency-bound at
// ~150 instructions per warp and field, not waiting on DRAM alone).
#ifndef PBE_TR_CTAS
#define PBE_TR_CTAS 4
#endif
#ifndef PBE_TR_DIRECT_OWN
#define PBE_TR_DIRECT_OWN 1 // own value and source through registers instead of the shared-memory stage: 2.66 -> 2.55 ms
#endif
#ifndef PBE_TR_STAGES
#define PBE_TR_STAGES 2 // 0: register-staged loop; 2..4: cp.async pipeline depth (measured: 2, 3, 4 alike)
#endif

#include "pbe_mesh.h"

namespace pbe {

struct TransportParams {
    int64_t ncells, stride, stride_b, nint, nirr;
    int nfields, chunk;
    const int32_t *other, *face, *irr_cell, *irr_ptr, *irr_other, *irr_face;
    const double *vcell, *psi, *phi, *psi_b, *S;
    const double *psi_halo; // [nfields][stride_h]: cell values across the coupled faces
    int64_t stride_h, coupled0, ncoupled;
    double *out;
    double rdt;  // 1/dt, as MULES::explicitSolve forms it
    double rrdt; // RN(1/rdt): reciprocal for div_by
};

// Correctly rounded a / b from y = RN(1/b) (Markstein): q0 = RN(a y), r = a - b q0 (exact in an
// FMA), q = RN(q0 + r y).  Outside a comfortable exponent range (results that could be subnormal,
// overflow, non-finite operands) the IEEE division is used, so the quotient is bit-identical to
// `a / b` everywhere; three FP64 instructions instead of the ~30 of the division's full sequence.
__device__ __forceinline__ double div_by(double a, double b, double y) {
    const double q0 = __dmul_rn(a, y);
    const int hi = __double2hiint(q0) & 0x7ff00000;
    if (hi < 0x10000000 || hi > 0x6ff00000) return __ddiv_rn(a, b); // |q0| outside [2^-767, 2^768): rare
    return __fma_rn(__fma_rn(-q0, b, a), y, q0);
}

// One face seen from one of its two cells.  The face loop adds the flux on the owner side and
// subtracts it on the neighbour side; -(phi f) == (-phi) f exactly, so the signed flux
// phs = +-phi is stored and the sum is a plain addition.  The face value
// lambda (psi_P - psi_N) + psi_N (surfaceInterpolationScheme::interpolate, lambda = pos(phi)) is
// fma(lam, mine - theirs, N) with lam = +-lambda: the product is exact (lambda is 0 or 1), hence
// the FMA rounds once, exactly like the reference's multiply-then-add.
struct Side {
    double phs; // own side ? phi : -phi
    double lam; // own side ? pos(phi) : -pos(phi): +-1 or +-0; the sign bit says "neighbour side"
};

__device__ __forceinline__ Side make_side(int32_t face, const double *phi) {
    const bool own = face >= 0;
    const double ph = __ldg(phi + (own ? face : ~face));
    const double lambda = ph >= 0.0 ? 1.0 : 0.0;
    return Side{own ? ph : -ph, own ? lambda : -lambda};
}

__device__ __forceinline__ double add_internal(double acc, const Side &t, double mine, double theirs) {
    const double N = (__double2hiint(t.lam) >= 0) ? theirs : mine;
    const double fv = __fma_rn(t.lam, __dsub_rn(mine, theirs), N);
    return __dadd_rn(acc, __dmul_rn(t.phs, fv));
}

// psi_new = ((psi rDeltaT + Su) - acc / V) / rDeltaT with both quotients correctly rounded (div_by's
// sequence; one range test for the two of them)
__device__ __forceinline__ double updated(const TransportParams &p, double mine, double Su, double acc,
                                          double V, double rV) {
    const double a = __dadd_rn(__dmul_rn(mine, p.rdt), Su);
    const double q0 = __dmul_rn(acc, rV);
    const double div = __fma_rn(__fma_rn(-q0, V, acc), rV, q0);
    const double num = __dsub_rn(a, div);
    const double r0 = __dmul_rn(num, p.rrdt);
    const double res = __fma_rn(__fma_rn(-r0, p.rdt, num), p.rrdt, r0);
    const unsigned h0 = (unsigned)__double2hiint(q0) & 0x7ff00000u, h1 = (unsigned)__double2hiint(r0) & 0x7ff00000u;
    // exponents of both first guesses inside [2^-767, 2^768): otherwise (zero, subnormal, huge,
    // non-finite) take the IEEE divisions
    if (min(h0, h1) < 0x10000000u || max(h0, h1) > 0x6ff00000u)
        return div_by(__dsub_rn(a, div_by(acc, V, rV)), p.rdt, p.rrdt);
    return res;
}

// The F faces of an ELL cell, decoded once per thread.  A boundary face b is a slot whose value is the
// patch value psi_b[k][b] (bit s of `patch` set: read from the patch rows) or, without patch values,
// the cell's own (zero gradient); its flux term is phi_b * value as it stands (bit s of `bnd`),
// fvc::surfaceIntegrate's boundary loop.
template <int F>
struct CellFaces {
    int idx[F];
    Side sd[F];
    unsigned bnd, patch;
};

template <int F>
__device__ __forceinline__ bool load_cell_faces(const TransportParams &p, int c, CellFaces<F> &cf) {
    cf.idx[0] = __ldg(p.other + c);
    if (cf.idx[0] == -1) return false; // irregular cell: transport_upwind_irregular_kernel
    cf.bnd = cf.patch = 0u;
#pragma unroll
    for (int s = 0; s < F; ++s) {
        if (s > 0) cf.idx[s] = __ldg(p.other + (size_t)s * p.ncells + c);
        cf.sd[s] = make_side(__ldg(p.face + (size_t)s * p.ncells + c), p.phi); // boundary: face = nint + b >= 0
        if (cf.idx[s] < 0) {
            cf.bnd |= 1u << s;
            if (p.psi_b != nullptr) {
                cf.patch |= 1u << s;
                cf.idx[s] = -cf.idx[s] - 2;
            } else {
                cf.idx[s] = c;
            }
        }
    }
    return true;
}

template <int F, bool BND>
__device__ __forceinline__ const double *slot_addr(const CellFaces<F> &cf, int s, const double *row,
                                                   const double *brow) {
    return (BND && ((cf.patch >> s) & 1u)) ? brow + cf.idx[s] : row + cf.idx[s];
}
template <int F, bool BND>
__device__ __forceinline__ double add_slot(double acc, const CellFaces<F> &cf, int s, double mine, double theirs) {
    if (BND && ((cf.bnd >> s) & 1u)) return __dadd_rn(acc, __dmul_rn(cf.sd[s].phs, theirs));
    return add_internal(acc, cf.sd[s], mine, theirs);
}

template <int F, bool BND>
__device__ __forceinline__ void sweep_regs(const TransportParams &p, int c, const CellFaces<F> &cf, double V,
                                           double rV) {
    const int k0 = blockIdx.y * p.chunk;
    const int k1 = min(p.nfields, k0 + p.chunk);
    const double *row = p.psi + (size_t)k0 * p.stride;
    const double *brow = BND ? p.psi_b + (size_t)k0 * p.stride_b : nullptr; // only dereferenced for patch slots
    const double *srow = (p.S != nullptr) ? p.S + (size_t)k0 * p.stride : nullptr;
    double *orow = p.out + (size_t)k0 * p.stride;
#pragma unroll 2
    for (int k = k0; k < k1; ++k, row += p.stride, orow += p.stride) {
        const double mine = __ldg(row + c);
        double th[F];
#pragma unroll
        for (int s = 0; s < F; ++s) th[s] = __ldg(slot_addr<F, BND>(cf, s, row, brow));
        double Su = 0.0;
        if (srow != nullptr) {
            Su = __ldg(srow + c);
            srow += p.stride;
        }
        double acc = 0.0;
#pragma unroll
        for (int s = 0; s < F; ++s) acc = add_slot<F, BND>(acc, cf, s, mine, th[s]);
        __stcs(orow + c, updated(p, mine, Su, acc, V, rV));
        if (BND) brow += p.stride_b;
    }
}

// The same sweep with the NEXT field's values (own, source, F gathered) already on their way into registers while the
// current field is summed: the loads have a whole field's arithmetic to land, and nothing goes through shared memory
// (F + 2 LDG per cell-field instead of F LDGSTS + F LDS + 2 LDG of the cp.async form below).  MEASURED SLOWER here --
// 160^3 x 100: 3.19 ms against 2.55 ms (PBE_TRANSPORT_KERNEL=1): the update's arithmetic needs 80 registers with the
// second set of values (three CTAs per SM), whereas the implicit sweep (pbe_implicit.cu), which gains 9 % from this
// form, stays at 57.  Kept selectable for that comparison.
template <int F, bool BND>
__device__ __forceinline__ void sweep_ahead(const TransportParams &p, int c, const CellFaces<F> &cf, double V, double rV) {
    const int k0 = blockIdx.y * p.chunk;
    const int nk = min(p.nfields, k0 + p.chunk) - k0;
    const double *row = p.psi + (size_t)k0 * p.stride;
    const double *brow = BND ? p.psi_b + (size_t)k0 * p.stride_b : nullptr;
    const double *srow = (p.S != nullptr) ? p.S + (size_t)k0 * p.stride : nullptr;
    double *orow = p.out + (size_t)k0 * p.stride;
    double thn[F], mine_n = 0.0, su_n = 0.0;
    auto fetch = [&](int k) {
        if (k < nk) {
#pragma unroll
            for (int s = 0; s < F; ++s) thn[s] = __ldg(slot_addr<F, BND>(cf, s, row, brow));
            mine_n = __ldg(row + c);
            if (srow != nullptr) {
                su_n = __ldg(srow + c);
                srow += p.stride;
            }
            row += p.stride;
            if (BND) brow += p.stride_b;
        }
    };
    fetch(0);
    for (int k = 0; k < nk; ++k, orow += p.stride) {
        double th[F];
#pragma unroll
        for (int s = 0; s < F; ++s) th[s] = thn[s];
        const double mine = mine_n, Su = su_n;
        fetch(k + 1);
        double acc = 0.0;
#pragma unroll
        for (int s = 0; s < F; ++s) acc = add_slot<F, BND>(acc, cf, s, mine, th[s]);
        __stcs(orow + c, updated(p, mine, Su, acc, V, rV));
    }
}

template <int F>
__global__ void __launch_bounds__(256, 3) transport_upwind_ahead_kernel(TransportParams p) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= p.ncells) return;
    CellFaces<F> cf;
    if (!load_cell_faces<F>(p, c, cf)) return;
    const double V = __ldg(p.vcell + c);
    const double rV = __ddiv_rn(1.0, V);
    if (__any_sync(__activemask(), cf.bnd != 0u))
        sweep_ahead<F, true>(p, c, cf, V, rV);
    else
        sweep_ahead<F, false>(p, c, cf, V, rV);
}

template <int F>
__global__ void __launch_bounds__(256, PBE_TR_CTAS) transport_upwind_kernel(TransportParams p) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x; // 32-bit labels (pbe_mesh_create checks)
    if (c >= p.ncells) return;
    CellFaces<F> cf;
    if (!load_cell_faces<F>(p, c, cf)) return;
    const double V = __ldg(p.vcell + c);
    const double rV = __ddiv_rn(1.0, V);
    // warps with a boundary slot (the cells next to a patch) take the loop that knows about them
    if (__any_sync(__activemask(), cf.bnd != 0u))
        sweep_regs<F, true>(p, c, cf, V, rV);
    else
        sweep_regs<F, false>(p, c, cf, V, rV);
}

// The same sweep with the loads of the next PBE_TR_STAGES-1 fields in flight as asynchronous
// global->shared copies (LDGSTS): every thread copies ITS OWN F + 2 values of a field into its own
// column of a shared-memory stage and reads them back after cp.async.wait_group -- no data is
// shared between threads, so no barrier is needed; the copies hold no registers while they are in
// flight, which is what the register-staged loop above runs out of (ncu: 7.2 of 13.5 stall cycles
// per issue are long-scoreboard waits).
template <int F, int D, bool BND>
__device__ __forceinline__ void sweep_async(const TransportParams &p, int c, int tid, const CellFaces<F> &cf,
                                            double V, double rV, double *tr_sm) {
    const int k0 = blockIdx.y * p.chunk;
    const int nk = min(p.nfields, k0 + p.chunk) - k0;
    const double *row = p.psi + (size_t)k0 * p.stride;       // row of the field being ISSUED
    const double *brow = BND ? p.psi_b + (size_t)k0 * p.stride_b : nullptr;
    const double *srow = (p.S != nullptr) ? p.S + (size_t)k0 * p.stride : nullptr;
    double *orow = p.out + (size_t)k0 * p.stride;
    const uint32_t sm0 = (uint32_t)__cvta_generic_to_shared(tr_sm + tid);
    constexpr uint32_t kStageBytes = (F + 2) * 256 * sizeof(double);
    auto cp8 = [](uint32_t dst, const double *src) {
        asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src) : "memory");
    };
    if constexpr (PBE_TR_DIRECT_OWN != 0 && D == 2) {
    // The cell's own value and its source are coalesced streams that nobody else reads: they go through registers
    // (one field ahead), only the F gathered values through the shared-memory stage -- two copies and two read-backs
    // fewer per cell-field on the load-store pipe, which is what bounds this kernel (ncu: L1/TEX 80 % busy).
    double mine_n = 0.0, su_n = 0.0;
    auto issue = [&](int k) {
        if (k < nk) {
            const uint32_t d = sm0 + (uint32_t)(k % D) * kStageBytes;
#pragma unroll
            for (int s = 0; s < F; ++s) cp8(d + (uint32_t)(1 + s) * 256u * 8u, slot_addr<F, BND>(cf, s, row, brow));
            mine_n = __ldg(row + c);
            if (srow != nullptr) {
                su_n = __ldg(srow + c);
                srow += p.stride;
            }
            row += p.stride;
            if (BND) brow += p.stride_b;
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    issue(0);
    for (int k = 0; k < nk; ++k, orow += p.stride) {
        const double mine = mine_n, Su = su_n;
        issue(k + 1);
        asm volatile("cp.async.wait_group %0;" ::"n"(D - 1) : "memory");
        const double *st = tr_sm + (size_t)(k % D) * ((F + 2) * 256) + tid;
        double acc = 0.0;
#pragma unroll
        for (int s = 0; s < F; ++s) acc = add_slot<F, BND>(acc, cf, s, mine, st[(1 + s) * 256]);
        __stcs(orow + c, updated(p, mine, Su, acc, V, rV));
    }
    return;
    }
    auto issue = [&](int k) {
        if (k < nk) {
            const uint32_t d = sm0 + (uint32_t)(k % D) * kStageBytes;
            cp8(d, row + c);
#pragma unroll
            for (int s = 0; s < F; ++s) cp8(d + (uint32_t)(1 + s) * 256u * 8u, slot_addr<F, BND>(cf, s, row, brow));
            if (srow != nullptr) {
                cp8(d + (uint32_t)(F + 1) * 256u * 8u, srow + c);
                srow += p.stride;
            }
            row += p.stride;
            if (BND) brow += p.stride_b;
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
#pragma unroll
    for (int k = 0; k < D - 1; ++k) issue(k);
    for (int k = 0; k < nk; ++k, orow += p.stride) {
        issue(k + D - 1);
        asm volatile("cp.async.wait_group %0;" ::"n"(D - 1) : "memory");
        const double *st = tr_sm + (size_t)(k % D) * ((F + 2) * 256) + tid;
        const double mine = st[0];
        double acc = 0.0;
#pragma unroll
        for (int s = 0; s < F; ++s) acc = add_slot<F, BND>(acc, cf, s, mine, st[(1 + s) * 256]);
        const double Su = (p.S != nullptr) ? st[(F + 1) * 256] : 0.0;
        __stcs(orow + c, updated(p, mine, Su, acc, V, rV));
    }
}

template <int F, int D>
__global__ void __launch_bounds__(256) transport_upwind_async_kernel(TransportParams p) {
    extern __shared__ double tr_sm[]; // [D][F + 2][256]
    const int tid = threadIdx.x;
    const int c = blockIdx.x * blockDim.x + tid;
    if (c >= p.ncells) return;
    CellFaces<F> cf;
    if (!load_cell_faces<F>(p, c, cf)) return;
    const double V = __ldg(p.vcell + c);
    const double rV = __ddiv_rn(1.0, V);
    if (__any_sync(__activemask(), cf.bnd != 0u))
        sweep_async<F, D, true>(p, c, tid, cf, V, rV, tr_sm);
    else
        sweep_async<F, D, false>(p, c, tid, cf, V, rV, tr_sm);
}

// cells with another face count and cells on processor patches: one thread per listed cell.  Up to kIrrRegs
// faces are decoded once into registers (like the regular cells'), so a field costs independent
// loads only; cells with more faces re-read their list per field.
constexpr int kIrrRegs = 8;

struct IrrFace {
    int32_t oth; // >= 0 neighbour cell; -(b + 2) boundary face b; -1 unused
    Side sd;     // internal: signed flux / lambda; boundary: phs = phi_b, lam = 1 if coupled
};

__device__ __forceinline__ IrrFace load_irr(const TransportParams &p, int e) {
    IrrFace t;
    t.oth = __ldg(p.irr_other + e);
    const int face = __ldg(p.irr_face + e);
    if (t.oth >= 0) {
        t.sd = make_side(face, p.phi);
    } else {
        const int64_t b = -(int64_t)t.oth - 2;
        t.sd.phs = __ldg(p.phi + face);
        t.sd.lam = (b >= p.coupled0 && b < p.coupled0 + p.ncoupled) ? 1.0 : 0.0;
    }
    return t;
}

// address of the value across the face (own value for zero-gradient patches)
__device__ __forceinline__ const double *irr_addr(const TransportParams &p, const IrrFace &t, const double *row,
                                                  int k, int c) {
    if (t.oth >= 0) return row + t.oth;
    const int64_t b = -(int64_t)t.oth - 2;
    if (t.sd.lam != 0.0) return p.psi_halo + (size_t)k * p.stride_h + (b - p.coupled0);
    return (p.psi_b != nullptr) ? p.psi_b + (size_t)k * p.stride_b + b : row + c;
}

__device__ __forceinline__ double add_irr(double acc, const IrrFace &t, double mine, double theirs) {
    if (t.oth >= 0) return add_internal(acc, t.sd, mine, theirs);
    double pv = theirs; // patch value
    if (t.sd.lam != 0.0) {  // processor patch: pLambda psi_own + (1 - pLambda) psi_across, pLambda = pos(phi_b)
        const double w = t.sd.phs >= 0.0 ? 1.0 : 0.0;
        pv = __dadd_rn(__dmul_rn(w, mine), __dmul_rn(__dsub_rn(1.0, w), theirs));
    }
    return __dadd_rn(acc, __dmul_rn(t.sd.phs, pv));
}

__global__ void __launch_bounds__(128) transport_upwind_irregular_kernel(TransportParams p) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.nirr) return;
    const int c = __ldg(p.irr_cell + i);
    const int e0 = __ldg(p.irr_ptr + i), e1 = __ldg(p.irr_ptr + i + 1);
    const int ne = e1 - e0;
    const double V = __ldg(p.vcell + c);
    const double rV = __ddiv_rn(1.0, V);
    const int k0 = blockIdx.y * p.chunk;
    const int k1 = min(p.nfields, k0 + p.chunk);
    IrrFace fc[kIrrRegs];
#pragma unroll
    for (int s = 0; s < kIrrRegs; ++s) {
        fc[s].oth = -1;
        fc[s].sd = Side{0.0, 0.0};
        if (s < ne) fc[s] = load_irr(p, e0 + s);
    }
    for (int k = k0; k < k1; ++k) {
        const double *row = p.psi + (size_t)k * p.stride;
        const double mine = __ldg(row + c);
        double th[kIrrRegs];
#pragma unroll
        for (int s = 0; s < kIrrRegs; ++s) th[s] = (s < ne) ? __ldg(irr_addr(p, fc[s], row, k, c)) : 0.0;
        double acc = 0.0;
#pragma unroll
        for (int s = 0; s < kIrrRegs; ++s)
            if (s < ne) acc = add_irr(acc, fc[s], mine, th[s]);
        for (int e = e0 + kIrrRegs; e < e1; ++e) { // polyhedral cells with many faces
            const IrrFace t = load_irr(p, e);
            acc = add_irr(acc, t, mine, __ldg(irr_addr(p, t, row, k, c)));
        }
        const double Su = (p.S != nullptr) ? __ldg(p.S + (size_t)k * p.stride + c) : 0.0;
        p.out[(size_t)k * p.stride + c] = updated(p, mine, Su, acc, V, rV);
    }
}

// halo packing: out[k][j] = psi[k][cells[j]]
__global__ void __launch_bounds__(256) gather_cells_kernel(int nfields, int64_t stride, const double *psi,
                                                           int64_t n, const int32_t *cells, double *out,
                                                           int64_t stride_out) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const int c = __ldg(cells + j);
    for (int k = blockIdx.y; k < nfields; k += gridDim.y) out[(size_t)k * stride_out + j] = __ldg(psi + (size_t)k * stride + c);
}

// QMOM::correct()'s steps either side of the moment transport (QMOM.C:98-128)
__global__ void __launch_bounds__(256) scrub_nan_kernel(int64_t count, double *f) {
    const int64_t n = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += n) {
        const double v = f[i];
        if (v != v) f[i] = 0.0; // QMOM.C:101-107
    }
}

__global__ void __launch_bounds__(256) clamp_d_kernel(int nmoments, int64_t ncells, int64_t stride, double *m,
                                                      double *d) {
    const int64_t n = (int64_t)gridDim.x * blockDim.x;
    for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c < ncells; c += n) {
        double m0 = 0.0, m1 = 0.0;
        for (int k = 0; k < nmoments; ++k) {
            double v = m[(size_t)k * stride + c];
            // max(moment, SMALL), QMOM.C:117-123; written like the fused step's clamp (NaN -> SMALL)
            if (!(v > 1.0e-15)) m[(size_t)k * stride + c] = v = 1.0e-15;
            if (k == 0) m0 = v;
            if (k == 1) m1 = v;
        }
        if (d != nullptr && nmoments >= 2) d[c] = cbrt(6.0 / 3.14159265358979323846 * m1 / m0); // QMOM.C:125
    }
}

} // namespace pbe

using namespace pbe;

extern "C" {

int pbe_div_scheme_type(const char *name, int *out) {
    PBE_REQUIRE(name != nullptr && out != nullptr, PBE_E_INVALID, "pbe_div_scheme_type: NULL argument");
    if (!strcmp(name, "upwind") || !strcmp(name, "Gauss upwind")) {
        *out = PBE_DIV_UPWIND;
        return PBE_OK;
    }
    if (!strcmp(name, "vanLeer") || !strcmp(name, "Gauss vanLeer")) {
        *out = PBE_DIV_VANLEER;
        return PBE_OK;
    }
    if (!strcmp(name, "limitedLinear 1") || !strcmp(name, "Gauss limitedLinear 1") || !strcmp(name, "limitedLinear 1.0") ||
        !strcmp(name, "Gauss limitedLinear 1.0")) {
        *out = PBE_DIV_LIMITEDLINEAR; // implicit convection term only
        return PBE_OK;
    }
    set_error("Unknown or not yet supported divergence scheme \"%s\" for the device transport\n\n"
              "Valid schemes are :\n3\n(\nupwind\nvanLeer\nlimitedLinear 1\n)",
              name);
    return PBE_E_UNKNOWN_TYPE;
}

// Splits the cells into the ELL set (exactly F faces, none coupled) and the CSR list of the rest, and
// (re)builds the device tables of both; pbe_mesh_set_coupled calls it again, because which boundary
// faces are processor patches is only known then.
static int mesh_split_cells(pbe_mesh *m) {
    const int64_t ncells = m->ncells, nint = m->nint, nbnd = m->nbnd;
    const int32_t *owner = m->h_owner.data(), *neighbour = m->h_neigh.data(), *bown = m->h_bown.data();
    std::vector<int32_t> cnt((size_t)ncells, 0), bcnt((size_t)ncells, 0);
    std::vector<char> coupled((size_t)ncells, 0);
    for (int64_t f = 0; f < nint; ++f) {
        cnt[owner[f]]++;
        cnt[neighbour[f]]++;
    }
    for (int64_t b = 0; b < nbnd; ++b) {
        bcnt[bown[b]]++;
        if (b >= m->coupled0 && b < m->coupled0 + m->ncoupled) coupled[bown[b]] = 1;
    }
    // F: the most common face count among the cells the ELL kernel can take, if a kernel exists for it
    int64_t hist[7] = {0, 0, 0, 0, 0, 0, 0};
    for (int64_t c = 0; c < ncells; ++c) {
        const int t = cnt[c] + bcnt[c];
        if (!coupled[c] && t >= 4 && t <= 6) hist[t]++;
    }
    int F = 0;
    for (int f = 4; f <= 6; ++f)
        if (hist[f] > 0 && (F == 0 || hist[f] > hist[F])) F = f;
    auto regular = [&](int64_t c) { return F > 0 && !coupled[c] && cnt[c] + bcnt[c] == F; };
    // irregular cells -> CSR
    std::vector<int32_t> irr_cell, irr_ptr(1, 0), slot_of((size_t)ncells, -1);
    for (int64_t c = 0; c < ncells; ++c)
        if (!regular(c)) {
            slot_of[c] = (int32_t)irr_cell.size();
            irr_cell.push_back((int32_t)c);
            irr_ptr.push_back(irr_ptr.back() + cnt[c] + bcnt[c]);
        }
    std::vector<int32_t> irr_other((size_t)irr_ptr.back()), irr_face((size_t)irr_ptr.back());
    const int Fa = std::max(F, 1);
    std::vector<int32_t> other((size_t)Fa * ncells, -1), face((size_t)Fa * ncells, 0), fill((size_t)ncells, 0);
    // ascending face order per cell = the order of fvc::surfaceIntegrate's face loops (internal
    // faces, then the patches' faces)
    auto put = [&](int32_t c, int32_t oth, int32_t fc) {
        const int32_t k = fill[c]++;
        if (slot_of[c] < 0) {
            other[(size_t)k * ncells + c] = oth;
            face[(size_t)k * ncells + c] = fc;
        } else {
            irr_other[(size_t)irr_ptr[slot_of[c]] + k] = oth;
            irr_face[(size_t)irr_ptr[slot_of[c]] + k] = fc;
        }
    };
    for (int64_t f = 0; f < nint; ++f) {
        put(owner[f], neighbour[f], (int32_t)f);
        put(neighbour[f], owner[f], ~(int32_t)f);
    }
    for (int64_t b = 0; b < nbnd; ++b) put(bown[b], -(int32_t)b - 2, (int32_t)(nint + b));
    for (int64_t c = 0; c < ncells; ++c)
        if (slot_of[c] >= 0) other[c] = -1; // row 0: "not an ELL cell"

    DeviceGuard g(m->ctx->device);
    cudaError_t e = cudaDeviceSynchronize(); // a re-split must not pull tables from under a running kernel
    auto upload = [&](int32_t **dst, const std::vector<int32_t> &v) {
        if (e != cudaSuccess) return;
        cudaFree(*dst);
        *dst = nullptr;
        e = cudaMalloc((void **)dst, std::max<size_t>(v.size(), 1) * sizeof(int32_t));
        if (e == cudaSuccess && !v.empty())
            e = cudaMemcpy(*dst, v.data(), v.size() * sizeof(int32_t), cudaMemcpyHostToDevice);
    };
    upload(&m->d_other, other);
    upload(&m->d_face, face);
    upload(&m->d_irr_cell, irr_cell);
    upload(&m->d_irr_ptr, irr_ptr);
    upload(&m->d_irr_other, irr_other);
    upload(&m->d_irr_face, irr_face);
    if (e != cudaSuccess) return cuda_fail(e, "pbe_mesh: cell tables", __FILE__, __LINE__);
    m->nirr = (int64_t)irr_cell.size();
    m->F = F;
    return PBE_OK;
}

int pbe_mesh_create(pbe_ctx *ctx, int64_t ncells, int64_t n_internal_faces, const int32_t *owner,
                    const int32_t *neighbour, int64_t n_boundary_faces,
                    const int32_t *boundary_owner, const double *vcell, pbe_mesh **out) {
    PBE_REQUIRE(ctx != nullptr && out != nullptr, PBE_E_INVALID, "pbe_mesh_create: NULL argument");
    *out = nullptr;
    PBE_REQUIRE(ncells >= 1 && n_internal_faces >= 0 && n_boundary_faces >= 0, PBE_E_INVALID,
                "pbe_mesh_create: bad sizes (ncells %lld, internal faces %lld, boundary faces %lld)",
                (long long)ncells, (long long)n_internal_faces, (long long)n_boundary_faces);
    PBE_REQUIRE(ncells < INT32_MAX - 256 && 2 * n_internal_faces + n_boundary_faces < INT32_MAX - 2,
                PBE_E_UNSUPPORTED, "pbe_mesh_create: 32-bit labels (WM_LABEL_SIZE=32) exceeded");
    PBE_REQUIRE(vcell != nullptr && (n_internal_faces == 0 || (owner && neighbour)) &&
                    (n_boundary_faces == 0 || boundary_owner),
                PBE_E_INVALID, "pbe_mesh_create: NULL array");
    for (int64_t f = 0; f < n_internal_faces; ++f) {
        const int32_t o = owner[f], n = neighbour[f];
        PBE_REQUIRE(o >= 0 && o < ncells && n >= 0 && n < ncells && o != n, PBE_E_INVALID,
                    "pbe_mesh_create: internal face %lld has owner %d, neighbour %d (ncells %lld)",
                    (long long)f, o, n, (long long)ncells);
    }
    for (int64_t b = 0; b < n_boundary_faces; ++b) {
        const int32_t o = boundary_owner[b];
        PBE_REQUIRE(o >= 0 && o < ncells, PBE_E_INVALID,
                    "pbe_mesh_create: boundary face %lld has owner %d (ncells %lld)", (long long)b, o,
                    (long long)ncells);
    }
    std::vector<int32_t> cnt((size_t)ncells, 0), bcnt((size_t)ncells, 0);
    for (int64_t f = 0; f < n_internal_faces; ++f) {
        cnt[owner[f]]++;
        cnt[neighbour[f]]++;
    }
    for (int64_t b = 0; b < n_boundary_faces; ++b) bcnt[boundary_owner[b]]++;
    int maxf = 0;
    for (int64_t c = 0; c < ncells; ++c) maxf = std::max(maxf, (int)(cnt[c] + bcnt[c]));

    // one CSR over ALL cells, ascending face order per cell (internal faces, then boundary faces): the
    // MULES kernels (pbe_mules.cu) walk it; entries as in the irregular-cell list
    std::vector<int32_t> csr_ptr((size_t)ncells + 1, 0);
    for (int64_t c = 0; c < ncells; ++c) csr_ptr[c + 1] = csr_ptr[c] + cnt[c] + bcnt[c];
    std::vector<int32_t> csr_other((size_t)csr_ptr[ncells]), csr_face((size_t)csr_ptr[ncells]), cfill((size_t)ncells, 0);
    auto cput = [&](int32_t c, int32_t oth, int32_t fc) {
        const size_t k = (size_t)csr_ptr[c] + cfill[c]++;
        csr_other[k] = oth;
        csr_face[k] = fc;
    };
    for (int64_t f = 0; f < n_internal_faces; ++f) {
        cput(owner[f], neighbour[f], (int32_t)f);
        cput(neighbour[f], owner[f], ~(int32_t)f);
    }
    for (int64_t b = 0; b < n_boundary_faces; ++b)
        cput(boundary_owner[b], -(int32_t)b - 2, (int32_t)(n_internal_faces + b));

    pbe_mesh *m = new (std::nothrow) pbe_mesh;
    PBE_REQUIRE(m != nullptr, PBE_E_NOMEM, "pbe_mesh_create: out of host memory");
    m->ctx = ctx;
    m->ncells = ncells;
    m->nint = n_internal_faces;
    m->nbnd = n_boundary_faces;
    m->maxf = maxf;
    m->h_owner.assign(owner, owner + n_internal_faces);
    m->h_neigh.assign(neighbour, neighbour + n_internal_faces);
    m->h_bown.assign(boundary_owner, boundary_owner + n_boundary_faces);
    DeviceGuard g(ctx->device);
    cudaError_t e = cudaSuccess;
    auto upload = [&](auto **dst, const auto &v) {
        if (e != cudaSuccess) return;
        const size_t bytes = std::max<size_t>(v.size(), 1) * sizeof(v[0]);
        e = cudaMalloc((void **)dst, bytes);
        if (e == cudaSuccess && !v.empty()) e = cudaMemcpy(*dst, v.data(), v.size() * sizeof(v[0]), cudaMemcpyHostToDevice);
    };
    upload(&m->d_csr_ptr, csr_ptr);
    upload(&m->d_csr_other, csr_other);
    upload(&m->d_csr_face, csr_face);
    upload(&m->d_owner, m->h_owner);
    upload(&m->d_neigh, m->h_neigh);
    upload(&m->d_bown, m->h_bown);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&m->side, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&m->ev_fork, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&m->ev_join, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaMalloc((void **)&m->d_vcell, (size_t)ncells * sizeof(double));
    if (e == cudaSuccess)
        e = cudaMemcpy(m->d_vcell, vcell, (size_t)ncells * sizeof(double), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        pbe_mesh_destroy(m);
        return cuda_fail(e, "pbe_mesh_create: device tables", __FILE__, __LINE__);
    }
    const int rc = mesh_split_cells(m);
    if (rc != PBE_OK) {
        pbe_mesh_destroy(m);
        return rc;
    }
    *out = m;
    return PBE_OK;
}

int pbe_mesh_destroy(pbe_mesh *m) {
    if (m == nullptr) return PBE_OK;
    DeviceGuard g(m->ctx->device);
    cudaFree(m->d_other);
    cudaFree(m->d_face);
    cudaFree(m->d_vcell);
    cudaFree(m->d_irr_cell);
    cudaFree(m->d_irr_ptr);
    cudaFree(m->d_irr_other);
    cudaFree(m->d_irr_face);
    cudaFree(m->d_csr_ptr);
    cudaFree(m->d_csr_other);
    cudaFree(m->d_csr_face);
    cudaFree(m->d_owner);
    cudaFree(m->d_neigh);
    cudaFree(m->d_bown);
    cudaFree(m->d_Sf);
    cudaFree(m->d_C);
    cudaFree(m->d_w);
    cudaFree(m->mules_ws);
    cudaFree(m->d_imp_other);
    cudaFree(m->d_imp_face);
    cudaFree(m->imp_ws);
    cudaFree(m->d_send_cells);
    cudaFree(m->d_halo_qoff);
    cudaFree(m->d_halo_qcnt);
    cudaFree(m->imp_halo);
    cudaFree(m->imp_small);
    if (m->imp_pinned) cudaFreeHost(m->imp_pinned);
    for (cudaEvent_t e : m->imp_ev)
        if (e) cudaEventDestroy(e);
    cudaFree(m->res);
    if (m->ev_fork) cudaEventDestroy(m->ev_fork);
    if (m->ev_join) cudaEventDestroy(m->ev_join);
    if (m->side) cudaStreamDestroy(m->side);
    delete m;
    return PBE_OK;
}

int pbe_mesh_info(pbe_mesh *m, int64_t *ncells, int64_t *n_internal_faces,
                  int64_t *n_boundary_faces, int *max_faces_per_cell) {
    PBE_REQUIRE(m != nullptr, PBE_E_INVALID, "pbe_mesh_info: NULL mesh");
    if (ncells) *ncells = m->ncells;
    if (n_internal_faces) *n_internal_faces = m->nint;
    if (n_boundary_faces) *n_boundary_faces = m->nbnd;
    if (max_faces_per_cell) *max_faces_per_cell = m->maxf;
    return PBE_OK;
}

int pbe_mesh_regular_cells(pbe_mesh *m, int *faces_per_regular_cell, int64_t *n_regular) {
    PBE_REQUIRE(m != nullptr, PBE_E_INVALID, "pbe_mesh_regular_cells: NULL mesh");
    if (faces_per_regular_cell) *faces_per_regular_cell = m->F;
    if (n_regular) *n_regular = m->ncells - m->nirr;
    return PBE_OK;
}

int pbe_mesh_set_coupled(pbe_mesh *m, int64_t first_boundary_face, int64_t count) {
    PBE_REQUIRE(m != nullptr, PBE_E_INVALID, "pbe_mesh_set_coupled: NULL mesh");
    PBE_REQUIRE(first_boundary_face >= 0 && count >= 0 && first_boundary_face + count <= m->nbnd,
                PBE_E_INVALID, "pbe_mesh_set_coupled: faces [%lld, %lld) outside the %lld boundary faces",
                (long long)first_boundary_face, (long long)(first_boundary_face + count), (long long)m->nbnd);
    if (m->coupled0 == first_boundary_face && m->ncoupled == count) return PBE_OK;
    m->coupled0 = first_boundary_face;
    m->ncoupled = count;
    return mesh_split_cells(m); // cells on processor patches leave the ELL set
}

int pbe_gather_cells(pbe_ctx *ctx, int nfields, int64_t stride, const double *psi, int64_t n,
                     const int32_t *cells, double *out, int64_t stride_out, void *cuda_stream) {
    PBE_REQUIRE(ctx != nullptr, PBE_E_INVALID, "pbe_gather_cells: NULL context");
    PBE_REQUIRE(nfields >= 0 && n >= 0 && stride_out >= n, PBE_E_INVALID, "pbe_gather_cells: bad sizes");
    if (nfields == 0 || n == 0) return PBE_OK;
    PBE_REQUIRE(psi && cells && out, PBE_E_INVALID, "pbe_gather_cells: NULL pointer");
    DeviceGuard g(ctx->device);
    dim3 grid((unsigned)((n + 255) / 256), (unsigned)std::min(nfields, 64));
    gather_cells_kernel<<<grid, 256, 0, (cudaStream_t)cuda_stream>>>(nfields, stride, psi, n, cells, out, stride_out);
    PBE_CUDA(cudaGetLastError());
    ctx->launches++;
    return PBE_OK;
}

int pbe_field_scrub_nan(pbe_ctx *ctx, int64_t count, double *f, void *cuda_stream) {
    PBE_REQUIRE(ctx != nullptr && count >= 0, PBE_E_INVALID, "pbe_field_scrub_nan: bad argument");
    if (count == 0) return PBE_OK;
    PBE_REQUIRE(f != nullptr, PBE_E_INVALID, "pbe_field_scrub_nan: NULL field");
    DeviceGuard g(ctx->device);
    const int grid = (int)std::min<int64_t>((count + 255) / 256, (int64_t)ctx->sm_count * 16);
    scrub_nan_kernel<<<grid, 256, 0, (cudaStream_t)cuda_stream>>>(count, f);
    PBE_CUDA(cudaGetLastError());
    ctx->launches++;
    return PBE_OK;
}

int pbe_qmom_clamp_d(pbe_ctx *ctx, int nmoments, int64_t ncells, int64_t stride, double *m, double *d,
                     void *cuda_stream) {
    PBE_REQUIRE(ctx != nullptr && nmoments >= 0 && ncells >= 0 && stride >= ncells, PBE_E_INVALID,
                "pbe_qmom_clamp_d: bad argument");
    if (nmoments == 0 || ncells == 0) return PBE_OK;
    PBE_REQUIRE(m != nullptr, PBE_E_INVALID, "pbe_qmom_clamp_d: NULL moments");
    DeviceGuard g(ctx->device);
    const int grid = (int)std::min<int64_t>((ncells + 255) / 256, (int64_t)ctx->sm_count * 16);
    clamp_d_kernel<<<grid, 256, 0, (cudaStream_t)cuda_stream>>>(nmoments, ncells, stride, m, d);
    PBE_CUDA(cudaGetLastError());
    ctx->launches++;
    return PBE_OK;
}

int pbe_transport_explicit(pbe_mesh *m, int scheme, int nfields, int64_t stride, const double *psi,
                           const double *phi, const double *psi_b, int64_t stride_b,
                           const double *S, double dt, double *psi_out, void *cuda_stream) {
    return pbe_transport_explicit_halo(m, scheme, nfields, stride, psi, phi, psi_b, stride_b, nullptr, 0, S,
                                       dt, psi_out, cuda_stream);
}

int pbe_transport_explicit_halo(pbe_mesh *m, int scheme, int nfields, int64_t stride,
                                const double *psi, const double *phi, const double *psi_b,
                                int64_t stride_b, const double *psi_halo, int64_t stride_h,
                                const double *S, double dt, double *psi_out, void *cuda_stream) {
    return pbe_transport_explicit_part(m, scheme, PBE_CELLS_ALL, nfields, stride, psi, phi, psi_b, stride_b,
                                       psi_halo, stride_h, S, dt, psi_out, cuda_stream);
}

int pbe_transport_explicit_part(pbe_mesh *m, int scheme, int which, int nfields, int64_t stride,
                                const double *psi, const double *phi, const double *psi_b,
                                int64_t stride_b, const double *psi_halo, int64_t stride_h,
                                const double *S, double dt, double *psi_out, void *cuda_stream) {
    PBE_REQUIRE(which == PBE_CELLS_ALL || which == PBE_CELLS_REGULAR || which == PBE_CELLS_IRREGULAR,
                PBE_E_INVALID, "pbe_transport_explicit_part: which must be PBE_CELLS_ALL / _REGULAR / _IRREGULAR");
    PBE_REQUIRE(m != nullptr, PBE_E_INVALID, "pbe_transport_explicit: NULL mesh");
    PBE_REQUIRE(scheme == PBE_DIV_UPWIND, PBE_E_UNSUPPORTED,
                "pbe_transport_explicit: only Gauss upwind (PBE_DIV_UPWIND) takes the unlimited explicit update; "
                "limited schemes (Gauss vanLeer, id %d) go through pbe_transport_mules", scheme);
    PBE_REQUIRE(nfields >= 0 && stride >= m->ncells, PBE_E_INVALID,
                "pbe_transport_explicit: bad nfields/stride (%d, %lld; mesh has %lld cells)", nfields,
                (long long)stride, (long long)m->ncells);
    PBE_REQUIRE(dt > 0.0, PBE_E_INVALID, "pbe_transport_explicit: dt must be positive");
    if (nfields == 0) return PBE_OK;
    PBE_REQUIRE(psi && psi_out && (phi || m->nint + m->nbnd == 0), PBE_E_INVALID,
                "pbe_transport_explicit: NULL pointer");
    PBE_REQUIRE(psi != psi_out, PBE_E_INVALID,
                "pbe_transport_explicit: psi_out must not alias psi (neighbours read old values)");
    PBE_REQUIRE(psi_b == nullptr || stride_b >= m->nbnd, PBE_E_INVALID,
                "pbe_transport_explicit: stride_b < boundary faces");
    PBE_REQUIRE(m->ncoupled == 0 || which == PBE_CELLS_REGULAR || (psi_halo != nullptr && stride_h >= m->ncoupled),
                PBE_E_INVALID,
                "pbe_transport_explicit: the mesh has %lld coupled (processor) faces: "
                "pbe_transport_explicit_halo needs psi_halo [nfields][stride_h >= %lld]",
                (long long)m->ncoupled, (long long)m->ncoupled);
    pbe_ctx *ctx = m->ctx;
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    TransportParams p{m->ncells, stride, stride_b, m->nint, m->nirr, nfields, 1,
                      m->d_other, m->d_face, m->d_irr_cell, m->d_irr_ptr, m->d_irr_other, m->d_irr_face,
                      m->d_vcell, psi, phi, psi_b, S, psi_halo, stride_h, m->coupled0, m->ncoupled,
                      psi_out, 1.0 / dt, 1.0 / (1.0 / dt)};
    // fields per CTA (grid.y = field chunks): few enough that the rows a wave of CTAs gathers
    // neighbours from stay in L2, many enough to amortise the per-cell face data; small meshes
    // get more chunks so that every SM has work
    auto chunk_for = [&](int64_t blocks, int start) {
        int chunk = std::min(start, nfields);
        while (chunk > 1 && blocks * ((nfields + chunk - 1) / chunk) < (int64_t)ctx->sm_count * 8)
            chunk = (chunk + 1) / 2;
        return chunk;
    };
    const bool have_regular = m->F > 0 && m->nirr < m->ncells && which != PBE_CELLS_IRREGULAR;
    if (m->nirr > 0 && which != PBE_CELLS_REGULAR) {
        // beside the regular cells' kernel when there is one (few, latency-bound threads)
        cudaStream_t is = have_regular ? m->side : st;
        if (have_regular) {
            PBE_CUDA(cudaEventRecord(m->ev_fork, st));
            PBE_CUDA(cudaStreamWaitEvent(is, m->ev_fork, 0));
        }
        TransportParams q = p;
        const int64_t blocks = (m->nirr + 127) / 128;
        q.chunk = chunk_for(blocks, 10);
        dim3 grid((unsigned)blocks, (unsigned)((nfields + q.chunk - 1) / q.chunk));
        transport_upwind_irregular_kernel<<<grid, 128, 0, is>>>(q);
        PBE_CUDA(cudaGetLastError());
        ctx->launches++;
        if (have_regular) PBE_CUDA(cudaEventRecord(m->ev_join, is));
    }
    if (have_regular) {
        int start = 50; // 160^3 x 100 fields: 25 -> 2.61 ms, 50 -> 2.55 ms, 100 -> slower (fewer CTAs per wave of rows)
        if (const char *e = getenv("PBE_TRANSPORT_CHUNK")) start = std::max(1, atoi(e));
        const int64_t blocks = (m->ncells + 255) / 256;
        p.chunk = chunk_for(blocks, start);
        dim3 grid((unsigned)blocks, (unsigned)((nfields + p.chunk - 1) / p.chunk));
        const int variant = getenv("PBE_TRANSPORT_KERNEL") ? atoi(getenv("PBE_TRANSPORT_KERNEL")) : PBE_TR_STAGES;
        auto launch_async = [&](auto kern, int F, int D) {
            const size_t smem = (size_t)D * (F + 2) * 256 * sizeof(double);
            cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            kern<<<grid, 256, smem, st>>>(p);
        };
#define PBE_TR_ASYNC(F)                                                                      \
    switch (variant) {                                                                       \
    case 1: transport_upwind_ahead_kernel<F><<<grid, 256, 0, st>>>(p); break;                \
    case 2: launch_async(transport_upwind_async_kernel<F, 2>, F, 2); break;                  \
    case 3: launch_async(transport_upwind_async_kernel<F, 3>, F, 3); break;                  \
    case 4: launch_async(transport_upwind_async_kernel<F, 4>, F, 4); break;                  \
    default: transport_upwind_kernel<F><<<grid, 256, 0, st>>>(p); break;                     \
    }
        switch (m->F) {
        case 4: PBE_TR_ASYNC(4); break;
        case 5: PBE_TR_ASYNC(5); break;
        default: PBE_TR_ASYNC(6); break;
        }
#undef PBE_TR_ASYNC
        PBE_CUDA(cudaGetLastError());
        ctx->launches++;
        if (m->nirr > 0 && which != PBE_CELLS_REGULAR) PBE_CUDA(cudaStreamWaitEvent(st, m->ev_join, 0));
    }
    return PBE_OK;
}

// MOC::correct() with solveWithMULES, host buffers in and out: what the OpenFOAM-side adapter
// calls.  The class fields are uploaded in cell chunks and the source kernel runs on each chunk as
// it arrives (three streams: copies overlap kernels); n and S stay on the device; the transport
// update then runs in field chunks, each chunk's new values copied back while the next chunk is
// being computed.
static int moc_correct_host_impl(pbe_moc *moc, pbe_mesh *m, int scheme, int64_t stride, double *n_host,
                                 const double *eps_host, const double *rho_host, const double *nu_host,
                                 const double *phi_host, const double *psi_b_host, int64_t stride_b, double dt,
                                 double *d_host);
int pbe_moc_correct_host(pbe_moc *moc, pbe_mesh *m, int scheme, int64_t stride, double *n_host,
                         const double *eps_host, const double *rho_host, const double *nu_host,
                         const double *phi_host, const double *psi_b_host, int64_t stride_b, double dt,
                         double *d_host) {
    const int rc = moc_correct_host_impl(moc, m, scheme, stride, n_host, eps_host, rho_host, nu_host, phi_host,
                                         psi_b_host, stride_b, dt, d_host);
    // a failure mid-pipeline leaves copies queued that still touch the caller's buffers
    if (rc != PBE_OK && m != nullptr) drain_slot_streams(m->ctx);
    return rc;
}
static int moc_correct_host_impl(pbe_moc *moc, pbe_mesh *m, int scheme, int64_t stride, double *n_host,
                                 const double *eps_host, const double *rho_host, const double *nu_host,
                                 const double *phi_host, const double *psi_b_host, int64_t stride_b, double dt,
                                 double *d_host) {
    PBE_REQUIRE(moc != nullptr && m != nullptr, PBE_E_INVALID, "pbe_moc_correct_host: NULL handle");
    int N = 0;
    pbe_ctx *mctx = nullptr;
    int rc = pbe_moc_num_classes(moc, &N, &mctx);
    if (rc) return rc;
    pbe_ctx *ctx = m->ctx;
    PBE_REQUIRE(mctx == ctx, PBE_E_INVALID, "pbe_moc_correct_host: method and mesh belong to different contexts");
    PBE_REQUIRE(m->ncoupled == 0, PBE_E_UNSUPPORTED,
                "pbe_moc_correct_host: decomposed meshes go through pbe_transport_explicit_halo");
    PBE_REQUIRE(scheme == PBE_DIV_UPWIND, PBE_E_UNSUPPORTED,
                "pbe_moc_correct_host: Gauss upwind only; limited schemes (Gauss vanLeer) need the phase fraction for "
                "the MULES bound: use pbe_moc_state_attach_mesh + pbe_moc_correct_resident");
    const int64_t nc = m->ncells, nf = m->nint + m->nbnd;
    PBE_REQUIRE(stride >= nc && n_host && eps_host && rho_host && nu_host && (phi_host || nf == 0) && dt > 0.0,
                PBE_E_INVALID, "pbe_moc_correct_host: bad argument");
    PBE_REQUIRE(psi_b_host == nullptr || stride_b >= m->nbnd, PBE_E_INVALID,
                "pbe_moc_correct_host: stride_b < boundary faces");
    DeviceGuard g(ctx->device);
    const size_t field = (size_t)N * nc;
    const size_t nb = psi_b_host ? (size_t)N * m->nbnd : 0;
    const size_t need = 3 * field + 4 * (size_t)nc + (size_t)nf + nb;
    if (m->res_doubles < need) {
        cudaFree(m->res);
        m->res = nullptr;
        m->res_doubles = 0;
        PBE_CUDA(cudaMalloc((void **)&m->res, need * sizeof(double)));
        m->res_doubles = need;
    }
    double *dn = m->res, *dS = dn + field, *dnew = dS + field, *de = dnew + field, *dr = de + nc, *dv = dr + nc,
           *dd = dv + nc, *dphi = dd + nc, *dpb = dphi + nf;
    // ---- upload + sources, cell chunks ----------------------------------------------------------
    int block = 128, ctas = 1;
    pbe_moc_get_kernel(moc, nullptr, &block, &ctas);
    const int64_t wave = (int64_t)block * ctx->sm_count * ctas;
    int64_t C = std::max<int64_t>(1, ((int64_t)256 << 20) / ((int64_t)N * 8) / wave) * wave;
    int64_t k = 0;
    for (int64_t c0 = 0; c0 < nc; c0 += C, ++k) {
        const int64_t cn = std::min(C, nc - c0);
        cudaStream_t st = ctx->slot_stream[k % kHostSlots];
        PBE_CUDA(cudaMemcpy2DAsync(dn + c0, (size_t)nc * 8, n_host + c0, (size_t)stride * 8, (size_t)cn * 8, N,
                                   cudaMemcpyHostToDevice, st));
        PBE_CUDA(cudaMemcpyAsync(de + c0, eps_host + c0, (size_t)cn * 8, cudaMemcpyHostToDevice, st));
        PBE_CUDA(cudaMemcpyAsync(dr + c0, rho_host + c0, (size_t)cn * 8, cudaMemcpyHostToDevice, st));
        PBE_CUDA(cudaMemcpyAsync(dv + c0, nu_host + c0, (size_t)cn * 8, cudaMemcpyHostToDevice, st));
        rc = pbe_moc_source(moc, cn, nc, dn + c0, de + c0, dr + c0, dv + c0, dt, dS + c0, st);
        if (rc) return rc;
    }
    if (nf) PBE_CUDA(cudaMemcpyAsync(dphi, phi_host, (size_t)nf * 8, cudaMemcpyHostToDevice, ctx->slot_stream[0]));
    if (nb)
        PBE_CUDA(cudaMemcpy2DAsync(dpb, (size_t)m->nbnd * 8, psi_b_host, (size_t)stride_b * 8, (size_t)m->nbnd * 8, N,
                                   cudaMemcpyHostToDevice, ctx->slot_stream[0]));
    for (int s = 0; s < kHostSlots; ++s) PBE_CUDA(cudaStreamSynchronize(ctx->slot_stream[s]));
    // ---- transport + download, field chunks -----------------------------------------------------
    const int KC = std::max(1, std::min(N, (int)(((int64_t)256 << 20) / (nc * 8) + 1)));
    k = 0;
    for (int k0 = 0; k0 < N; k0 += KC, ++k) {
        const int kn = std::min(KC, N - k0);
        cudaStream_t st = ctx->slot_stream[k % kHostSlots];
        rc = pbe_transport_explicit(m, scheme, kn, nc, dn + (size_t)k0 * nc, dphi,
                                    nb ? dpb + (size_t)k0 * m->nbnd : nullptr, m->nbnd, dS + (size_t)k0 * nc, dt,
                                    dnew + (size_t)k0 * nc, st);
        if (rc) return rc;
        PBE_CUDA(cudaMemcpy2DAsync(n_host + (size_t)k0 * stride, (size_t)stride * 8, dnew + (size_t)k0 * nc,
                                   (size_t)nc * 8, (size_t)nc * 8, kn, cudaMemcpyDeviceToHost, st));
    }
    for (int s = 0; s < kHostSlots; ++s) PBE_CUDA(cudaStreamSynchronize(ctx->slot_stream[s]));
    if (d_host != nullptr) { // MOC::d(), MOC.C:282-290, of the updated fields
        rc = pbe_moc_d(moc, nc, nc, dnew, dd, ctx->slot_stream[0]);
        if (rc) return rc;
        PBE_CUDA(cudaMemcpyAsync(d_host, dd, (size_t)nc * 8, cudaMemcpyDeviceToHost, ctx->slot_stream[0]));
        PBE_CUDA(cudaStreamSynchronize(ctx->slot_stream[0]));
    }
    return PBE_OK;
}

} // extern "C"
