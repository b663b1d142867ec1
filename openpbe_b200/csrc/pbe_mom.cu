// pbe_mom.cu -- methodOfMoments (first slice): the per-cell part of
// MOM::correct() (src/PBESystem/MOM.C:148-236) -- presumed gamma number density from three
// moments, the nested integrals of coalescenceSourceTerm / breakupSourceTerm
// (MOM.C:259-362) and the scrub / update / clamp / diameter tail.
//
// The reference evaluates every integral with integrate() = GSL QAGIU (Integrator.H:12-27:
// x = lower + (1-t)/t on t in (0,1], adaptive 15-point Gauss-Kronrod, epsabs 1e-7).  The device
// version keeps the transform and the 15-point rule on a FIXED set of panels (`panels` equal
// panels of (0,1], the one next to t = 1 bisected PBE_MOM_GRADED_LEVELS times towards the lower
// limit, where the gamma density is x^(k-1)): with a fixed rule the kernel evaluations
// are shared by the three moment indices and by the birth and death integrals, so a cell
// costs (15 panels)^2 / 2 coalescence-kernel evaluations instead of 6 nested adaptive runs.  Parity is UNPINNED in the reference (no MOM test exists; Boost.Math / GSL absent
// here): the tests compare against nested adaptive QUADPACK QAGI integrals in the reference's
// shape and against closed forms for the cell-independent kernels (tests/test_mom.py).
//
// One warp per cell: lanes split the outer nodes, the inner loop runs over the warp's
// shared-memory copy of the cell's weighted density at the nodes.
#include <algorithm>

#include "pbe_common.h"
#include "pbe_kernels.cuh"

using namespace pbe;

struct pbe_mom {
    pbe_ctx *ctx = nullptr;
    int panels = 0;
    DevKernelSpec ks{};
};

namespace {

constexpr int kWarpsPerCta = 4;
constexpr double kSmall = 1.0e-15; // OpenFOAM SMALL (double precision)

// signed Kronrod abscissae / weights of the 15-point rule on [-1, 1]
__constant__ double GK15_X[15] = {
    -0.991455371120812639206854697526329, -0.949107912342758524526189684047851,
    -0.864864423359769072789712788640926, -0.741531185599394439863864773280788,
    -0.586087235467691130294144838258730, -0.405845151377397166906606412076961,
    -0.207784955007898467600689403773245, 0.0,
    0.207784955007898467600689403773245,  0.405845151377397166906606412076961,
    0.586087235467691130294144838258730,  0.741531185599394439863864773280788,
    0.864864423359769072789712788640926,  0.949107912342758524526189684047851,
    0.991455371120812639206854697526329};
__constant__ double GK15_W[15] = {
    0.022935322010529224963732008058970, 0.063092092629978553290700663189204,
    0.104790010322250183839876322541518, 0.140653259715525918745189590510238,
    0.169004726639267902826583426598550, 0.190350578064785409913256402421014,
    0.204432940075298892414161999234649, 0.209482141084727828012999174891714,
    0.204432940075298892414161999234649, 0.190350578064785409913256402421014,
    0.169004726639267902826583426598550, 0.140653259715525918745189590510238,
    0.104790010322250183839876322541518, 0.063092092629978553290700663189204,
    0.022935322010529224963732008058970};

struct MomParams {
    int64_t ncells, stride;
    const double *m_in; // [3][stride]
    double *m_out;      // in-place update when do_step
    double *S_out;      // [3][stride] or NULL
    double *d_out;      // [stride] or NULL
    double *gamma_out;  // [3][stride]: c0, k, theta (MOM.C:151-161) or NULL
    const double *eps, *rho, *nu;
    double dt;
    int do_step;
    int panels;
    DevKernelSpec ks;
};

// boost::math::pdf(gamma_distribution(k, theta), x) for x > 0 (MOM.C:256-257): through logs so
// that large shapes do not overflow
struct GammaPdf {
    double km1, inv_theta, lognorm;
    __device__ GammaPdf(double k, double theta) {
        km1 = k - 1.0;
        inv_theta = 1.0 / theta;
        lognorm = -lgamma(k) - k * log(theta);
        if (!(k > 0.0) || !(theta > 0.0)) lognorm = NAN; // Boost raises a domain error here
    }
    __device__ double operator()(double x) const {
        return exp(km1 * log(x) - x * inv_theta + lognorm);
    }
};

__global__ void __launch_bounds__(kWarpsPerCta * 32) mom_kernel(MomParams p) {
    extern __shared__ double sh[];
    const int P = p.panels, npan = P + PBE_MOM_GRADED_LEVELS;
    const int nq = 15 * npan;
    double *xq = sh;            // node position x = u / (1 - u), u = 1 - t (lower limit 0)
    double *wq = xq + nq;       // Kronrod weight * panel half-width / t^2
    double *ra = wq + nq;       // x^(1/3)
    double *rp = ra + nq;       // x^(2/9)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    double *fq = rp + nq + (size_t)warp * nq; // this warp's cell: wq * m0 * pdf(x)

    // panels in u = 1 - t: [h, 1) in P - 1 equal panels, [0, h) bisected PBE_MOM_GRADED_LEVELS
    // times towards u = 0 (x = 0), where the gamma density behaves like x^(k-1) -- the panels an
    // adaptive QAGIU run ends up with for non-integer shapes
    const double h = 1.0 / P;
    for (int q = threadIdx.x; q < nq; q += blockDim.x) {
        const int panel = q / 15, j = q - 15 * panel;
        double lo, hi;
        if (panel < P - 1) {
            lo = (panel + 1) * h;
            hi = (panel + 2) * h;
        } else {
            const int l = panel - (P - 1);
            hi = ldexp(h, -l);
            lo = (l < PBE_MOM_GRADED_LEVELS) ? ldexp(h, -(l + 1)) : 0.0;
        }
        const double uc = 0.5 * (lo + hi), uh = 0.5 * (hi - lo);
        const double u = uc + uh * GK15_X[j];
        const double t = 1.0 - u;
        const double x = u / t;
        xq[q] = x;
        wq[q] = GK15_W[j] * uh / (t * t);
        const NodeRoots r = node_roots(x);
        ra[q] = r.a;
        rp[q] = r.p;
    }
    __syncthreads();

    const int64_t nwarps = (int64_t)gridDim.x * kWarpsPerCta;
    for (int64_t c = (int64_t)blockIdx.x * kWarpsPerCta + warp; c < p.ncells; c += nwarps) {
        const double m0 = p.m_in[c], m1 = p.m_in[p.stride + c], m2 = p.m_in[2 * p.stride + c];
        // MOM.C:151-161
        const double var = m2 * m0 - m1 * m1;
        const double gk = (m1 * m1) / (var + kSmall);
        const double gtheta = (m0 * m2 - m1 * m1) / fmax(m0 * m1, kSmall);
        const GammaPdf pdf(gk, gtheta);
        const CellCoef cc = cell_coef(p.ks, p.eps[c], p.rho[c], p.nu[c]);

        __syncwarp();
        for (int q = lane; q < nq; q += 32) fq[q] = wq[q] * (m0 * pdf(xq[q]));
        __syncwarp();

        // Coalescence (MOM.C:278-308), birth / 2 - death.  Every built-in kernel is symmetric in
        // (xi, xi'), so the two double integrals fold into one sum over unordered node pairs:
        //   k = 0: -1/2 sum f_a f_b S_ab;  k = 1: 0 exactly (volume is conserved);
        //   k = 2: 1/2 (xa + xb)^2 - 1/2 (xa^2 + xb^2) = xa xb.
        // Breakup (MOM.C:338-361): exchanging the order of the two integrals turns
        //   int_0^inf xi^k int_xi^inf beta(xi, xi') g(xi') n(xi') dxi' dxi   into
        //   int_0^inf moment(xi', k) g(xi') n(xi') dxi',  moment = int_0^xi' xi^k beta dxi = 2 xi'^k / (k+1)
        // (daughterParticleDistribution::moment, uniformBinaryBreakup.C:62-69 -- the form
        // QMOM.C:252-262 uses), one pass over the nodes instead of a nested integral per node.
        double s0 = 0.0, s1 = 0.0, s2 = 0.0;
        for (int a = lane; a < nq; a += 32) {
            const double xa = xq[a], fa = fq[a];
            const NodeRoots na{ra[a], rp[a]};
            if (p.ks.coal != PBE_COAL_NONE) {
                double c0 = 0.0, c2 = 0.0;
                for (int b = 0; b < a; ++b) {
                    const NodeRoots nb{ra[b], rp[b]};
                    const double t = fq[b] * coal_S(p.ks, cc, na, nb);
                    c0 += t;
                    c2 = fma_(t, xq[b], c2);
                }
                const double tdiag = 0.5 * (fa * coal_S(p.ks, cc, na, na)); // pair (a, a) counts once
                c0 += tdiag;
                c2 = fma_(tdiag, xa, c2);
                s0 = fma_(-fa, c0, s0);             // 2 * (-1/2) * fa * sum_b
                s2 = fma_(2.0 * (fa * xa), c2, s2);
            }
            if (p.ks.br != PBE_BR_NONE) {
                const double gn = break_S(p.ks, cc, xa, na) * fa;
                s0 += gn * (daughter_moment_from_pow(1.0, 0) - 1.0);
                s1 = fma_(xa * gn, daughter_moment_from_pow(1.0, 1) - 1.0, s1);
                s2 = fma_(xa * xa * gn, daughter_moment_from_pow(1.0, 2) - 1.0, s2);
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            s0 += __shfl_xor_sync(0xffffffffu, s0, o);
            s1 += __shfl_xor_sync(0xffffffffu, s1, o);
            s2 += __shfl_xor_sync(0xffffffffu, s2, o);
        }
        if (lane == 0) {
            double S[3] = {s0, s1, s2};
            if (p.gamma_out) {
                p.gamma_out[c] = m0;
                p.gamma_out[p.stride + c] = gk;
                p.gamma_out[2 * p.stride + c] = gtheta;
            }
            if (p.do_step) {
                double m[3] = {m0, m1, m2};
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    if (isnan(S[k])) S[k] = 0.0;                   // MOM.C:198-207
                    const double v = fma_(p.dt, S[k], m[k]);       // ddt(m) = S, zero flux
                    m[k] = (v > kSmall) ? v : kSmall;              // MOM.C:224-231
                    p.m_out[k * p.stride + c] = m[k];
                }
                if (p.d_out) p.d_out[c] = cbrt(6.0 / 3.14159265358979323846 * m[1] / m[0]); // MOM.C:233
            }
            if (p.S_out) {
#pragma unroll
                for (int k = 0; k < 3; ++k) p.S_out[k * p.stride + c] = S[k];
            }
        }
    }
}

int check_m(const char *fn, const pbe_mom *h, int64_t ncells, int64_t stride, const void *m,
            const void *eps, const void *rho, const void *nu) {
    PBE_REQUIRE(h != nullptr, PBE_E_INVALID, "%s: NULL handle", fn);
    PBE_REQUIRE(ncells >= 0 && stride >= ncells, PBE_E_INVALID,
                "%s: need 0 <= ncells <= stride (got %lld, %lld)", fn, (long long)ncells,
                (long long)stride);
    PBE_REQUIRE(ncells == 0 || (m && eps && rho && nu), PBE_E_INVALID, "%s: NULL field pointer", fn);
    return PBE_OK;
}

int launch_mom(pbe_mom *h, MomParams &p, cudaStream_t st) {
    if (p.ncells == 0) return PBE_OK;
    p.panels = h->panels;
    p.ks = h->ks;
    const int nq = 15 * (h->panels + PBE_MOM_GRADED_LEVELS);
    const size_t smem = (size_t)(4 + kWarpsPerCta) * nq * sizeof(double);
    // one warp per cell; 8 CTAs of 4 warps per SM, grid-stride beyond that
    const int64_t ctas = (p.ncells + kWarpsPerCta - 1) / kWarpsPerCta;
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>(ctas, (int64_t)h->ctx->sm_count * 8));
    mom_kernel<<<grid, kWarpsPerCta * 32, smem, st>>>(p);
    PBE_CUDA(cudaGetLastError());
    h->ctx->launches++;
    return PBE_OK;
}

} // namespace

extern "C" {

int pbe_mom_create(pbe_ctx *ctx, int panels, const pbe_kernel_spec *spec, pbe_mom **out) {
    PBE_REQUIRE(ctx && spec && out, PBE_E_INVALID, "pbe_mom_create: NULL argument");
    *out = nullptr;
    if (panels == 0) panels = PBE_MOM_DEFAULT_PANELS;
    PBE_REQUIRE(panels >= 1 && panels <= PBE_MOM_MAX_PANELS, PBE_E_UNSUPPORTED,
                "pbe_mom_create: panels must be in [1,%d] (got %d)", PBE_MOM_MAX_PANELS, panels);
    PBE_REQUIRE(spec->coalescence >= PBE_COAL_NONE && spec->coalescence <= PBE_COAL_CT,
                spec->coalescence == PBE_COAL_TABLE ? PBE_E_UNSUPPORTED : PBE_E_UNKNOWN_TYPE,
                "pbe_mom_create: coalescenceKernel id %d is not available for methodOfMoments",
                spec->coalescence);
    PBE_REQUIRE(spec->breakup >= PBE_BR_NONE && spec->breakup <= PBE_BR_CT,
                spec->breakup == PBE_BR_TABLE ? PBE_E_UNSUPPORTED : PBE_E_UNKNOWN_TYPE,
                "pbe_mom_create: breakupKernel id %d is not available for methodOfMoments",
                spec->breakup);
    PBE_REQUIRE(spec->daughter == PBE_DAUGHTER_UNIFORM_BINARY, PBE_E_UNKNOWN_TYPE,
                "pbe_mom_create: unknown daughterParticleDistribution id %d", spec->daughter);
    pbe_mom *h = new pbe_mom();
    h->ctx = ctx;
    h->panels = panels;
    h->ks = make_dev_spec(*spec);
    *out = h;
    return PBE_OK;
}

int pbe_mom_destroy(pbe_mom *h) {
    delete h;
    return PBE_OK;
}

int pbe_mom_source(pbe_mom *h, int64_t ncells, int64_t stride, const double *m, const double *eps,
                   const double *rho, const double *nu, double *S, double *gamma,
                   void *cuda_stream) {
    int rc = check_m("pbe_mom_source", h, ncells, stride, m, eps, rho, nu);
    if (rc) return rc;
    PBE_REQUIRE(ncells == 0 || S, PBE_E_INVALID, "pbe_mom_source: S is NULL");
    DeviceGuard g(h->ctx->device);
    MomParams p{};
    p.ncells = ncells;
    p.stride = stride;
    p.m_in = m;
    p.S_out = S;
    p.gamma_out = gamma;
    p.eps = eps;
    p.rho = rho;
    p.nu = nu;
    return launch_mom(h, p, (cudaStream_t)cuda_stream);
}

int pbe_mom_step(pbe_mom *h, int64_t ncells, int64_t stride, double *m, const double *eps,
                 const double *rho, const double *nu, double dt, double *S, double *d, double *gamma,
                 void *cuda_stream) {
    int rc = check_m("pbe_mom_step", h, ncells, stride, m, eps, rho, nu);
    if (rc) return rc;
    DeviceGuard g(h->ctx->device);
    MomParams p{};
    p.ncells = ncells;
    p.stride = stride;
    p.m_in = m;
    p.m_out = m;
    p.S_out = S;
    p.d_out = d;
    p.gamma_out = gamma;
    p.eps = eps;
    p.rho = rho;
    p.nu = nu;
    p.dt = dt;
    p.do_step = 1;
    return launch_mom(h, p, (cudaStream_t)cuda_stream);
}

} // extern "C"
