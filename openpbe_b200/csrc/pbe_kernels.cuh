// pbe_kernels.cuh -- device functors for the run-time selectable coalescence / breakup /
// daughter kernels (the reference's Foam-free *Impl classes), shared by the QMOM kernel and
// the batched scalar evaluators behind pbe_eval_*.
//
// The MOC kernel does not call these per pair: it uses the factorised table form
// S = A_c G_ij exp(-K_c H_ij) (pbe_moc_kernel.cuh).  Both forms take the SAME per-cell
// scalars from DevKernelSpec so that MOC and QMOM agree on what a kernel means.
#pragma once

#include "pbe_math.cuh"
#include "../../include/pbe_b200.h"

// 1: division- and call-free device math in the per-cell scalars of the QMOM kernel (cbrt_pos_fast,
// rcp_fast, sqrt_pos_fast; <= 3 ulp each, inside the 1e-9 sum|terms| tolerance of SURVEY.md 8c)
#ifndef PBE_QMOM_FASTMATH
#define PBE_QMOM_FASTMATH 1
#endif

namespace pbe {

// Device view of pbe_kernel_spec with the cell-independent algebra folded on the host.
struct DevKernelSpec {
    int coal;          // PBE_COAL_*
    int br;            // PBE_BR_*
    int daughter;      // PBE_DAUGHTER_*
    double coal_const; // constantCoalescence: S = c
    double coalA;      // CT: c1 / (1 + gamma)
    double coalK;      // CT: c2 / (sigma^2 (1 + gamma)^3)
    double brA;        // CT: c1 / (1 + gamma)
    double brK;        // CT: c2 sigma (1 + gamma)^2
};

inline DevKernelSpec make_dev_spec(const pbe_kernel_spec &s) {
    DevKernelSpec d{};
    d.coal = s.coalescence;
    d.br = s.breakup;
    d.daughter = s.daughter;
    d.coal_const = s.coal_c[0];
    {
        const double c1 = s.coal_c[0], c2 = s.coal_c[1], gamma = s.coal_c[2], sigma = s.coal_c[3];
        d.coalA = c1 / (1 + gamma);
        d.coalK = c2 / (sigma * sigma * ((1 + gamma) * (1 + gamma) * (1 + gamma)));
    }
    {
        const double c1 = s.br_c[0], c2 = s.br_c[1], gamma = s.br_c[2], sigma = s.br_c[3];
        d.brA = c1 / (1 + gamma);
        d.brK = c2 * sigma * ((1 + gamma) * (1 + gamma));
    }
    return d;
}

// per-cell scalars of the Coulaloglou-Tavlarides kernels
struct CellCoef {
    double A, negK;   // coalescence: S = A G exp(negK H)
    double Ab, negKb; // breakup:     g = Ab xi^(-2/9) exp(negKb xi^(-5/9))
};

PBE_HD CellCoef cell_coef(const DevKernelSpec &k, double eps, double rho, double nu) {
    CellCoef c;
    const double e13 = pow_third(eps);
    // coalescenceKernels/CoulaloglouTavlarides.C:77-88 (wrapper :52-60 feeds the DISPERSED
    // phase rho and nu into the parameters named rhoc, nuc)
    c.A = k.coalA * e13;
    c.negK = -(k.coalK * rho * eps * (rho * nu));
    // breakupKernels/CoulaloglouTavlarides.C:83-92
    c.Ab = k.brA * e13;
    c.negKb = -(k.brK / (rho * (e13 * e13)));
    return c;
}
#if defined(__CUDACC__)
// the same with call- and division-free device math (QMOM kernel only; the batched functor
// evaluators behind pbe_eval_* keep the libm forms above and their 1e-13 tolerance)
__device__ __forceinline__ CellCoef cell_coef_fast(const DevKernelSpec &k, double eps, double rho, double nu) {
    CellCoef c;
    const double e13 = cbrt_pos_fast(eps);
    c.A = k.coalA * e13;
    c.negK = -(k.coalK * rho * eps * (rho * nu));
    c.Ab = k.brA * e13;
    c.negKb = -(k.brK * rcp_fast(rho * (e13 * e13)));
    return c;
}
#endif

// node-dependent roots: a = xi^(1/3), p = xi^(2/9)
struct NodeRoots {
    double a, p;
};
PBE_HD NodeRoots node_roots(double xi) {
    NodeRoots r;
    r.a = pow_third(xi);
    const double a13 = pow_third(r.a);
    r.p = a13 * a13;
    return r;
}

// coalescenceKernels/CoulaloglouTavlarides.C:69-90 with the roots precomputed
PBE_HD double coal_ct(const CellCoef &c, const NodeRoots &x, const NodeRoots &y) {
    const double s = x.a + y.a;
    const double G = (s * s) * sqrt(x.p + y.p);
    const double h = x.a * y.a / s;
    const double h2 = h * h;
    return (c.A * G) * exp(c.negK * (h2 * h2));
}

PBE_HD double coal_S(const DevKernelSpec &k, const CellCoef &c, const NodeRoots &x,
                     const NodeRoots &y) {
    switch (k.coal) {
    case PBE_COAL_CONSTANT: return k.coal_const; // constantCoalescence.H:51-54
    case PBE_COAL_CT: return coal_ct(c, x, y);
    default: return 0.0; // noCoalescence.H:49-52
    }
}

// breakupKernels: CoulaloglouTavlarides.C:77-93, binaryBreakup.C:51-53, noBreakup.C:51-54
PBE_HD double break_S(const DevKernelSpec &k, const CellCoef &c, double xi, const NodeRoots &x) {
    switch (k.br) {
    case PBE_BR_BINARY: return xi * xi;
    case PBE_BR_CT: return c.Ab * exp(c.negKb / (x.a * x.p)) / x.p;
    default: return 0.0;
    }
}

// daughterParticleDistributions/uniformBinaryBreakup.C:55-60; OpenFOAM pos(x) = (x >= 0)
PBE_HD double daughter_beta(double xi1, double xi2) {
    return (2.0 / xi2) * ((xi2 - xi1) >= 0.0 ? 1.0 : 0.0);
}
// daughterParticleDistributions/uniformBinaryBreakup.C:62-69: 2 xi^k / (k + 1)
PBE_HD double daughter_moment_from_pow(double xi_pow_k, int k) {
    return 2.0 * xi_pow_k / (double)(k + 1);
}

} // namespace pbe
