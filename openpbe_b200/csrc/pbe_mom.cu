// pbe_mom.cu -- methodOfMoments (first slice): the per-cell part of
// MOM::correct() (src/PBESystem/MOM.C:148-236) -- presumed gamma number density from three
// moments, the nested integrals of coalescenceSourceTerm / breakupSourceTerm
// (MOM.C:259-362) and the scrub / update / clamp / diameter tail.
//
// The reference evaluates every integral with integrate() = GSL QAGIU (Integrator.H:12-27:
// x = lower + (1-t)/t on t in (0,1], adaptive 15-point Gauss-Kronrod, epsabs 1e-7).  The device
// version keeps the transform and the 15-point rule on a FIXED set of panels (`panels` equal
// panels of (0,1], the two end ones graded towards the lower limit, where the gamma density is
// x^(k-1), and towards infinity: mom_node below): with a fixed rule the kernel evaluations
// are shared by the three moment indices and by the birth and death integrals, so a cell
// costs (15 panels)^2 / 2 coalescence-kernel evaluations instead of 6 nested adaptive runs.  Parity is UNPINNED in the reference (no MOM test exists; Boost.Math / GSL absent
// here): the tests compare against nested adaptive QUADPACK QAGI integrals in the reference's
// shape and against closed forms for the cell-independent kernels (tests/test_mom.py).
//
// One warp per two cells: lanes split the outer nodes, the inner loop runs over the warp's
// shared-memory copies of the cells' weighted density at the nodes; the cell-independent factors
// of the CT coalescence kernel come from a per-handle pair table.
#include <algorithm>

#include <string.h>

#include "pbe_common.h"
#include "pbe_kernels.cuh"

using namespace pbe;

struct pbe_mom {
    pbe_ctx *ctx = nullptr;
    int panels = 0;
    DevKernelSpec ks{};
    double *pair_tab = nullptr; // device, owned
    unsigned long long *d_cov = nullptr; // device [2]: uncovered-cell count, worst deviation (bit pattern)
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
    const double *pair_tab; // CT coalescence: G, H of every node pair, [2][nq][nq], or NULL
    unsigned long long *cov; // [2]: cells whose density the fixed rule does not cover; worst deviation (bits)
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

// node q of the rule, panels in u = 1 - t (x = u / (1 - u)):
//   [h, 1 - h)   P - 2 equal panels;
//   [1 - h, 1)   bisected PBE_MOM_TAIL_LEVELS times towards u = 1 (x -> inf), the refinement QAGIU
//                makes for densities whose mass sits beyond x = P - 1;
//   [a, h)       bisected PBE_MOM_GRADED_LEVELS times towards u = 0, a = h 2^-GRADED;
//   [0, a)       one panel under u = a s^PBE_MOM_ORIGIN_POWER: the gamma density is x^(k-1) there, and
//                s^(8k-1) is smooth enough for the 15-point rule down to k = 0.3 where plain bisection
//                converges like a^k (the panels an adaptive run keeps splitting for shapes below 1).
__device__ __forceinline__ void mom_node(int P, int q, double &x, double &w) {
    const double h = 1.0 / P;
    const int panel = q / 15, j = q - 15 * panel;
    const int n_eq = P - 2, n_tail = PBE_MOM_TAIL_LEVELS + 1;
    double u, wu;
    if (panel == P + PBE_MOM_GRADED_LEVELS + PBE_MOM_TAIL_LEVELS - 1) {
        const double a = ldexp(h, -PBE_MOM_GRADED_LEVELS);
        const double s = 0.5 + 0.5 * GK15_X[j];
        double sp = s; // s^(POWER-1)
        for (int i = 2; i < PBE_MOM_ORIGIN_POWER; ++i) sp *= s;
        u = a * sp * s;
        wu = 0.5 * GK15_W[j] * a * PBE_MOM_ORIGIN_POWER * sp;
    } else {
        double lo, hi;
        if (panel < n_eq) {
            lo = (panel + 1) * h;
            hi = (panel + 2) * h;
        } else if (panel < n_eq + n_tail) {
            const int l = panel - n_eq;
            lo = 1.0 - ldexp(h, -l);
            hi = (l < PBE_MOM_TAIL_LEVELS) ? 1.0 - ldexp(h, -(l + 1)) : 1.0;
        } else {
            const int l = panel - n_eq - n_tail;
            hi = ldexp(h, -l);
            lo = ldexp(h, -(l + 1));
        }
        const double uc = 0.5 * (lo + hi), uh = 0.5 * (hi - lo);
        u = uc + uh * GK15_X[j];
        wu = GK15_W[j] * uh;
    }
    const double t = 1.0 - u;
    x = u / t;
    w = wu / (t * t);
}

// The CT coalescence kernel factorises into per-cell and per-pair parts, S = A_c G_ab
// exp(negK_c H_ab) (pbe_kernels.cuh coal_ct; the form the MOC kernel tabulates per class pair):
// G and H of every node pair are cell-independent and built once per handle, laid out [b][a]
// so that the lanes of a warp (consecutive a) read consecutive addresses.  3.7 MB at 480
// nodes: resident in the 126 MB L2.
__global__ void mom_pair_table_kernel(int P, int nq, double *__restrict__ tab) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (int64_t)nq * nq) return;
    const int a = (int)(idx % nq), b = (int)(idx / nq);
    double xa, xb, w;
    mom_node(P, a, xa, w);
    mom_node(P, b, xb, w);
    const NodeRoots na = node_roots(xa), nb = node_roots(xb);
    const double sab = na.a + nb.a;
    const double hh = na.a * nb.a / sab;
    const double h2 = hh * hh;
    tab[idx] = (sab * sab) * sqrt(na.p + nb.p);
    tab[(int64_t)nq * nq + idx] = h2 * h2;
}

// NC cells per warp share every table / node load of the pair sweep
template <int NC>
__global__ void __launch_bounds__(kWarpsPerCta * 32, NC == 1 ? 6 : 3) mom_kernel(MomParams p) {
    extern __shared__ double sh[];
    const int P = p.panels, npan = P + PBE_MOM_EXTRA_PANELS;
    const int nq = 15 * npan;
    double *xq = sh;            // node position x = u / (1 - u), u = 1 - t (lower limit 0)
    double *wq = xq + nq;       // Kronrod weight * panel half-width / t^2
    double *ra = wq + nq;       // x^(1/3)
    double *rp = ra + nq;       // x^(2/9)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    double *fq = rp + nq + (size_t)warp * NC * nq; // this warp's cells: wq * m0 * pdf(x)

    for (int q = threadIdx.x; q < nq; q += blockDim.x) {
        double x, w;
        mom_node(P, q, x, w);
        xq[q] = x;
        wq[q] = w;
        const NodeRoots r = node_roots(x);
        ra[q] = r.a;
        rp[q] = r.p;
    }
    __syncthreads();

    const bool use_tab = (p.ks.coal == PBE_COAL_CT) && p.pair_tab != nullptr;
    const int64_t nwarps = (int64_t)gridDim.x * kWarpsPerCta;
    for (int64_t cb = ((int64_t)blockIdx.x * kWarpsPerCta + warp) * NC; cb < p.ncells; cb += nwarps * NC) {
        double m0[NC], m1[NC], m2[NC], gk[NC], gtheta[NC];
        CellCoef cc[NC];
        __syncwarp();
#pragma unroll
        for (int i = 0; i < NC; ++i) {
            const int64_t c = (cb + i < p.ncells) ? cb + i : p.ncells - 1; // odd tail: recomputed, not written
            m0[i] = p.m_in[c], m1[i] = p.m_in[p.stride + c], m2[i] = p.m_in[2 * p.stride + c];
            // MOM.C:151-161
            const double var = m2[i] * m0[i] - m1[i] * m1[i];
            gk[i] = (m1[i] * m1[i]) / (var + kSmall);
            gtheta[i] = (m0[i] * m2[i] - m1[i] * m1[i]) / fmax(m0[i] * m1[i], kSmall);
            const GammaPdf pdf(gk[i], gtheta[i]);
            cc[i] = cell_coef(p.ks, p.eps[c], p.rho[c], p.nu[c]);
            // coverage of the presumed density by the FIXED rule: sum_q f_q must give back the density's
            // own zeroth and first moments, m0 and m0 k theta (= m1 up to the SMALL of MOM.C:154).  QAGIU would refine where this rule cannot (mass beyond the last
            // equal panel, or below the last graded one): such cells are counted, not silently wrong.
            double cv0 = 0.0, cv1 = 0.0;
            for (int q = lane; q < nq; q += 32) {
                const double fv = wq[q] * (m0[i] * pdf(xq[q]));
                fq[i * nq + q] = fv;
                cv0 += fv;
                cv1 = fma_(fv, xq[q], cv1);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                cv0 += __shfl_xor_sync(0xffffffffu, cv0, o);
                cv1 += __shfl_xor_sync(0xffffffffu, cv1, o);
            }
            if (lane == 0 && cb + i < p.ncells && p.cov != nullptr) {
                const double dev = fmax(fabs(cv0 / m0[i] - 1.0), fabs(cv1 / (m0[i] * gk[i] * gtheta[i]) - 1.0));
                if (dev > PBE_MOM_COVERAGE_TOL) { // NaN (invalid gamma parameters) compares false: scrubbed as in the reference
                    atomicAdd(p.cov, 1ull);
                    atomicMax(p.cov + 1, (unsigned long long)__double_as_longlong(dev));
                }
            }
        }
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
        double s0[NC], s1[NC], s2[NC];
#pragma unroll
        for (int i = 0; i < NC; ++i) s0[i] = s1[i] = s2[i] = 0.0;
        for (int a = lane; a < nq; a += 32) {
            const double xa = xq[a];
            const NodeRoots na{ra[a], rp[a]};
            if (p.ks.coal != PBE_COAL_NONE) {
                double c0[NC], c2[NC];
#pragma unroll
                for (int i = 0; i < NC; ++i) c0[i] = c2[i] = 0.0;
                if (use_tab) {
                    const double *Gt = p.pair_tab + a, *Ht = Gt + (size_t)nq * nq;
#pragma unroll 4
                    for (int b = 0; b < a; ++b) {
                        const double G = __ldg(Gt + (size_t)b * nq), H = __ldg(Ht + (size_t)b * nq);
                        const double xb = xq[b];
#pragma unroll
                        for (int i = 0; i < NC; ++i) {
                            const double t = (fq[i * nq + b] * G) * exp(cc[i].negK * H);
                            c0[i] += t;
                            c2[i] = fma_(t, xb, c2[i]);
                        }
                    }
                    const double G = __ldg(Gt + (size_t)a * nq), H = __ldg(Ht + (size_t)a * nq);
#pragma unroll
                    for (int i = 0; i < NC; ++i) { // pair (a, a) counts once
                        const double t = 0.5 * ((fq[i * nq + a] * G) * exp(cc[i].negK * H));
                        c0[i] = (c0[i] + t) * cc[i].A;
                        c2[i] = fma_(t, xa, c2[i]) * cc[i].A;
                    }
                } else {
                    for (int b = 0; b < a; ++b) {
                        const NodeRoots nb{ra[b], rp[b]};
                        const double xb = xq[b];
#pragma unroll
                        for (int i = 0; i < NC; ++i) {
                            const double t = fq[i * nq + b] * coal_S(p.ks, cc[i], na, nb);
                            c0[i] += t;
                            c2[i] = fma_(t, xb, c2[i]);
                        }
                    }
#pragma unroll
                    for (int i = 0; i < NC; ++i) {
                        const double t = 0.5 * (fq[i * nq + a] * coal_S(p.ks, cc[i], na, na));
                        c0[i] += t;
                        c2[i] = fma_(t, xa, c2[i]);
                    }
                }
#pragma unroll
                for (int i = 0; i < NC; ++i) {
                    const double fa = fq[i * nq + a];
                    s0[i] = fma_(-fa, c0[i], s0[i]); // 2 * (-1/2) * fa * sum_b
                    s2[i] = fma_(2.0 * (fa * xa), c2[i], s2[i]);
                }
            }
            if (p.ks.br != PBE_BR_NONE) {
#pragma unroll
                for (int i = 0; i < NC; ++i) {
                    const double gn = break_S(p.ks, cc[i], xa, na) * fq[i * nq + a];
                    s0[i] += gn * (daughter_moment_from_pow(1.0, 0) - 1.0);
                    s1[i] = fma_(xa * gn, daughter_moment_from_pow(1.0, 1) - 1.0, s1[i]);
                    s2[i] = fma_(xa * xa * gn, daughter_moment_from_pow(1.0, 2) - 1.0, s2[i]);
                }
            }
        }
#pragma unroll
        for (int i = 0; i < NC; ++i) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                s0[i] += __shfl_xor_sync(0xffffffffu, s0[i], o);
                s1[i] += __shfl_xor_sync(0xffffffffu, s1[i], o);
                s2[i] += __shfl_xor_sync(0xffffffffu, s2[i], o);
            }
        }
        if (lane == 0) {
#pragma unroll
            for (int i = 0; i < NC; ++i) {
                const int64_t c = cb + i;
                if (c >= p.ncells) break;
                double S[3] = {s0[i], s1[i], s2[i]};
                if (p.gamma_out) {
                    p.gamma_out[c] = m0[i];
                    p.gamma_out[p.stride + c] = gk[i];
                    p.gamma_out[2 * p.stride + c] = gtheta[i];
                }
                if (p.do_step) {
                    double m[3] = {m0[i], m1[i], m2[i]};
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
    p.pair_tab = h->pair_tab;
    p.cov = h->d_cov;
    const int nq = 15 * (h->panels + PBE_MOM_EXTRA_PANELS);
    // CT coalescence (pair table): two cells per warp share every table load; the cell-independent
    // kernels have nothing to share and run one cell per warp (measured: 13.8 vs 85 ms per 4e5 cells)
    const int NC = h->pair_tab ? 2 : 1;
    const size_t smem = (size_t)(4 + NC * kWarpsPerCta) * nq * sizeof(double);
    // up to 8 CTAs of 4 warps per SM, grid-stride beyond that
    const int64_t ctas = (p.ncells + NC * kWarpsPerCta - 1) / (NC * kWarpsPerCta);
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>(ctas, (int64_t)h->ctx->sm_count * 8));
    // the limit belongs to the instantiation: a handle with fewer panels may have been created since
    PBE_CUDA(raise_dyn_smem(h->ctx, NC == 2 ? (const void *)mom_kernel<2> : (const void *)mom_kernel<1>, smem));
    if (NC == 2)
        mom_kernel<2><<<grid, kWarpsPerCta * 32, smem, st>>>(p);
    else
        mom_kernel<1><<<grid, kWarpsPerCta * 32, smem, st>>>(p);
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
    PBE_REQUIRE(panels >= 2 && panels <= PBE_MOM_MAX_PANELS, PBE_E_UNSUPPORTED,
                "pbe_mom_create: panels must be in [2,%d] (got %d)", PBE_MOM_MAX_PANELS, panels);
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
    DeviceGuard g(ctx->device);
    const int nq = 15 * (panels + PBE_MOM_EXTRA_PANELS);
    // 12 rows of nq doubles per CTA (4 node tables + 2 cells x 4 warps): above 48 KB from 33 panels on
    cudaError_t e = raise_dyn_smem(ctx, (const void *)mom_kernel<2>, (4 + 2 * kWarpsPerCta) * nq * sizeof(double));
    if (e == cudaSuccess)
        e = raise_dyn_smem(ctx, (const void *)mom_kernel<1>, (4 + kWarpsPerCta) * nq * sizeof(double));
    if (e == cudaSuccess && spec->coalescence == PBE_COAL_CT) {
        e = cudaMalloc(&h->pair_tab, (size_t)2 * nq * nq * sizeof(double));
        if (e == cudaSuccess) {
            const int64_t n = (int64_t)nq * nq;
            mom_pair_table_kernel<<<(unsigned)((n + 255) / 256), 256>>>(panels, nq, h->pair_tab);
            e = cudaGetLastError();
            if (e == cudaSuccess) e = cudaDeviceSynchronize();
            ctx->launches++;
        }
    }
    if (e == cudaSuccess) e = cudaMalloc((void **)&h->d_cov, 2 * sizeof(unsigned long long));
    if (e == cudaSuccess) e = cudaMemset(h->d_cov, 0, 2 * sizeof(unsigned long long));
    if (e != cudaSuccess) {
        if (h->pair_tab) cudaFree(h->pair_tab);
        if (h->d_cov) cudaFree(h->d_cov);
        delete h;
        return cuda_fail(e, "pbe_mom_create", __FILE__, __LINE__);
    }
    *out = h;
    return PBE_OK;
}

int pbe_mom_destroy(pbe_mom *h) {
    if (h) {
        DeviceGuard g(h->ctx->device);
        if (h->pair_tab) cudaFree(h->pair_tab);
        if (h->d_cov) cudaFree(h->d_cov);
    }
    delete h;
    return PBE_OK;
}

// Cells (since the last reset) whose presumed gamma density the fixed quadrature rule did not cover:
// |sum_q f_q / m0 - 1| or |sum_q f_q x_q / m1 - 1| above PBE_MOM_COVERAGE_TOL.  Synchronises the device.
int pbe_mom_coverage(pbe_mom *h, int64_t *uncovered_cells, double *worst_deviation, int reset) {
    PBE_REQUIRE(h != nullptr, PBE_E_INVALID, "pbe_mom_coverage: NULL handle");
    DeviceGuard g(h->ctx->device);
    unsigned long long v[2] = {0, 0};
    PBE_CUDA(cudaDeviceSynchronize());
    PBE_CUDA(cudaMemcpy(v, h->d_cov, sizeof(v), cudaMemcpyDeviceToHost));
    if (uncovered_cells) *uncovered_cells = (int64_t)v[0];
    if (worst_deviation) {
        double d;
        memcpy(&d, &v[1], sizeof(d));
        *worst_deviation = d;
    }
    if (reset) PBE_CUDA(cudaMemset(h->d_cov, 0, sizeof(v)));
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
