// pbe_qmom.cu -- host side of the QMOM path (launch dispatch on the quadrature order, the
// host-buffer pipeline) and the batched device evaluators of the scalar kernel functors.
#include <stdlib.h>

#include <algorithm>

#include "pbe_common.h"
#include "pbe_qmom_kernel.cuh"

using namespace pbe;

struct pbe_qmom {
    pbe_ctx *ctx = nullptr;
    int order = 0;
    DevKernelSpec ks{};
};

namespace {

int grid_for(pbe_ctx *ctx, int64_t n, int threads, int per_sm) {
    return (int)std::max<int64_t>(
        1, std::min<int64_t>((n + threads - 1) / threads, (int64_t)ctx->sm_count * per_sm));
}

// 16 CTAs of 128 cells per SM, grid-stride over the cells.  Measured on B200 (1e7 cells, CT + CT):
// a persistent grid of exactly SMs x resident CTAs is 6 % slower (co-resident CTAs stay in
// lockstep: loads and FP64 phases coincide), more CTAs change nothing.
// Grid = `waves` x the resident CTAs (occupancy x SMs), grid-stride over the cells.  Measured on B200
// (1e7 cells, CT + CT, fused step, L2 flushed; profiles/r02c_qmom_ab.jsonl): 1 wave (a persistent
// grid) 0.678 ms -- co-resident CTAs stay in lockstep, loads and FP64 phases coincide --, 3 waves
// 0.647, the old fixed 16 CTAs per SM 0.643 (3.2 waves: ends in a partial wave), 8 waves 0.625:
// shorter CTAs de-phase and shrink the tail.  PBE_QMOM_WAVES overrides (0 = 16 CTAs per SM).
template <int N, int C, int B>
void launch_q(pbe_qmom *q, const QmomParams &p, cudaStream_t st) {
    static int resident = 0; // per instantiation, same on every B200 of the node
    if (resident == 0) {
        int occ = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, qmom_kernel<N, C, B>, 128, 0) != cudaSuccess || occ < 1)
            occ = 4;
        resident = occ;
    }
    const char *we = getenv("PBE_QMOM_WAVES");
    const int waves = we ? atoi(we) : 8;
    const int grid = waves > 0 ? grid_for(q->ctx, p.ncells, 128, resident * waves) : grid_for(q->ctx, p.ncells, 128, 16);
    qmom_kernel<N, C, B><<<grid, 128, 0, st>>>(p);
}

int launch_qmom(pbe_qmom *q, const QmomParams &p, cudaStream_t st) {
    if (p.ncells == 0) return PBE_OK;
    // one instantiation per (order, coalescenceKernel, breakupKernel): pbe_qmom_create has
    // already restricted the ids to {none, constant, CT} x {none, binary, CT}
#define PBE_Q_LAUNCH(N, C, B) launch_q<N, C, B>(q, p, st)
#define PBE_Q_BR(N, C)                                                                             \
    switch (q->ks.br) {                                                                            \
    case PBE_BR_NONE: PBE_Q_LAUNCH(N, C, PBE_BR_NONE); break;                                      \
    case PBE_BR_BINARY: PBE_Q_LAUNCH(N, C, PBE_BR_BINARY); break;                                  \
    default: PBE_Q_LAUNCH(N, C, PBE_BR_CT); break;                                                 \
    }
#define PBE_Q_COAL(N)                                                                              \
    switch (q->ks.coal) {                                                                          \
    case PBE_COAL_NONE: PBE_Q_BR(N, PBE_COAL_NONE) break;                                          \
    case PBE_COAL_CONSTANT: PBE_Q_BR(N, PBE_COAL_CONSTANT) break;                                  \
    default: PBE_Q_BR(N, PBE_COAL_CT) break;                                                       \
    }
    switch (q->order) {
    case 1: PBE_Q_COAL(1) break;
    case 2: PBE_Q_COAL(2) break;
    case 3: PBE_Q_COAL(3) break;
    case 4: PBE_Q_COAL(4) break;
    default: set_error("pbe_qmom: order %d not instantiated", q->order); return PBE_E_UNSUPPORTED;
    }
#undef PBE_Q_COAL
#undef PBE_Q_BR
#undef PBE_Q_LAUNCH
    PBE_CUDA(cudaGetLastError());
    q->ctx->launches++;
    return PBE_OK;
}

int check_q(const char *fn, const pbe_qmom *q, int64_t ncells, int64_t stride, const void *m,
            const void *eps, const void *rho, const void *nu) {
    PBE_REQUIRE(q != nullptr, PBE_E_INVALID, "%s: NULL handle", fn);
    PBE_REQUIRE(ncells >= 0 && stride >= ncells, PBE_E_INVALID,
                "%s: need 0 <= ncells <= stride (got %lld, %lld)", fn, (long long)ncells,
                (long long)stride);
    PBE_REQUIRE(ncells == 0 || (m && eps && rho && nu), PBE_E_INVALID, "%s: NULL field pointer", fn);
    return PBE_OK;
}

// ---- batched scalar functors ---------------------------------------------------------------
__global__ void eval_coal_kernel(DevKernelSpec ks, int64_t count, const double *xi1,
                                 const double *xi2, const double *eps, const double *rho,
                                 const double *nu, double *out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const CellCoef cc = cell_coef(ks, eps[i], rho[i], nu[i]);
    out[i] = coal_S(ks, cc, node_roots(xi1[i]), node_roots(xi2[i]));
}
__global__ void eval_break_kernel(DevKernelSpec ks, int64_t count, const double *xi,
                                  const double *rho, const double *eps, double *out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const CellCoef cc = cell_coef(ks, eps[i], rho[i], 0.0);
    out[i] = break_S(ks, cc, xi[i], node_roots(xi[i]));
}
__global__ void eval_beta_kernel(int64_t count, const double *xi1, const double *xi2, double *out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) out[i] = daughter_beta(xi1[i], xi2[i]);
}
__global__ void eval_dmoment_kernel(int64_t count, const double *xi, const int *k, double *out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) out[i] = daughter_moment_from_pow(pow_int(xi[i], k[i]), k[i]);
}

// 15-point Gauss-Kronrod on the QAGIU transform x = lower + (1-t)/t, t in (0,1]
// (Integrator.H:12-27 calls gsl_integration_qagiu(f, lower, 1e-7, 0, 1000, ...)).  The device
// version bisects (0,1] to a FIXED depth (64 equal panels): for the smooth, exponentially
// decaying integrands of this path that meets epsabs 1e-7 with a wide margin.
__device__ __forceinline__ double gk15_panel_exp(double rate, double lower, double a, double b) {
    const double xgk[8] = {0.991455371120812639206854697526329, 0.949107912342758524526189684047851,
                           0.864864423359769072789712788640926, 0.741531185599394439863864773280788,
                           0.586087235467691130294144838258730, 0.405845151377397166906606412076961,
                           0.207784955007898467600689403773245, 0.0};
    const double wgk[8] = {0.022935322010529224963732008058970, 0.063092092629978553290700663189204,
                           0.104790010322250183839876322541518, 0.140653259715525918745189590510238,
                           0.169004726639267902826583426598550, 0.190350578064785409913256402421014,
                           0.204432940075298892414161999234649, 0.209482141084727828012999174891714};
    const double center = 0.5 * (a + b), half = 0.5 * (b - a);
    auto f = [&](double t) {
        const double x = lower + (1.0 - t) / t;
        return exp(-rate * x) / (t * t);
    };
    double res = wgk[7] * f(center);
#pragma unroll
    for (int j = 0; j < 7; ++j) {
        const double dx = half * xgk[j];
        res += wgk[j] * (f(center - dx) + f(center + dx));
    }
    return res * half;
}
__global__ void eval_integrate_exp_kernel(int64_t count, const double *rate, const double *lower,
                                          double *out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const int panels = 64;
    double s = 0.0;
    for (int k = 0; k < panels; ++k)
        s += gk15_panel_exp(rate[i], lower[i], (double)k / panels, (double)(k + 1) / panels);
    out[i] = s;
}

// Utility.C:4-10 printAvgMaxMin partials: sum(f V), sum V, max f, min f
__global__ void __launch_bounds__(256) avg_max_min_kernel(int64_t ncells, const double *__restrict__ f,
                                                          const double *__restrict__ vcell,
                                                          double *__restrict__ part) {
    double sfv = 0.0, sv = 0.0, mx = -INFINITY, mn = INFINITY;
    const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
    for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c < ncells; c += nthreads) {
        const double v = __ldg(f + c), V = vcell ? __ldg(vcell + c) : 1.0;
        sfv = fma_(v, V, sfv);
        sv += V;
        mx = fmax(mx, v);
        mn = fmin(mn, v);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sfv += __shfl_xor_sync(0xffffffffu, sfv, o);
        sv += __shfl_xor_sync(0xffffffffu, sv, o);
        mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    }
    __shared__ double sh[8][4];
    if ((threadIdx.x & 31) == 0) {
        sh[threadIdx.x >> 5][0] = sfv;
        sh[threadIdx.x >> 5][1] = sv;
        sh[threadIdx.x >> 5][2] = mx;
        sh[threadIdx.x >> 5][3] = mn;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) {
            sfv += sh[w][0];
            sv += sh[w][1];
            mx = fmax(mx, sh[w][2]);
            mn = fmin(mn, sh[w][3]);
        }
        double *o = part + (size_t)blockIdx.x * 4;
        o[0] = sfv;
        o[1] = sv;
        o[2] = mx;
        o[3] = mn;
    }
}
__global__ void avg_max_min_final_kernel(int nparts, const double *__restrict__ part,
                                         double *__restrict__ out4) {
    double sfv = 0.0, sv = 0.0, mx = -INFINITY, mn = INFINITY;
    for (int b = threadIdx.x; b < nparts; b += 32) {
        sfv += part[(size_t)b * 4 + 0];
        sv += part[(size_t)b * 4 + 1];
        mx = fmax(mx, part[(size_t)b * 4 + 2]);
        mn = fmin(mn, part[(size_t)b * 4 + 3]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sfv += __shfl_xor_sync(0xffffffffu, sfv, o);
        sv += __shfl_xor_sync(0xffffffffu, sv, o);
        mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    }
    if (threadIdx.x == 0) {
        out4[0] = sfv;
        out4[1] = sv;
        out4[2] = mx;
        out4[3] = mn;
    }
}

// upload `narr` host arrays of `count` doubles, run fn, download out
struct HostBatch {
    pbe_ctx *ctx;
    std::vector<void *> dev;
    ~HostBatch() {
        for (void *p : dev)
            if (p) cudaFree(p);
    }
    int up(const void *host, size_t bytes, void **out) {
        void *d = nullptr;
        PBE_CUDA(cudaMalloc(&d, bytes ? bytes : 8));
        dev.push_back(d);
        if (host && bytes) PBE_CUDA(cudaMemcpy(d, host, bytes, cudaMemcpyHostToDevice));
        *out = d;
        return PBE_OK;
    }
};

} // namespace

extern "C" {

int pbe_qmom_create(pbe_ctx *ctx, int order, const pbe_kernel_spec *spec, pbe_qmom **out) {
    PBE_REQUIRE(ctx && spec && out, PBE_E_INVALID, "pbe_qmom_create: NULL argument");
    *out = nullptr;
    PBE_REQUIRE(order >= 1 && order <= PBE_QMOM_MAX_ORDER, PBE_E_UNSUPPORTED,
                "pbe_qmom_create: quadratureOrder must be in [1,%d] (got %d)", PBE_QMOM_MAX_ORDER,
                order);
    PBE_REQUIRE(spec->coalescence >= PBE_COAL_NONE && spec->coalescence <= PBE_COAL_CT,
                spec->coalescence == PBE_COAL_TABLE ? PBE_E_UNSUPPORTED : PBE_E_UNKNOWN_TYPE,
                "pbe_qmom_create: coalescenceKernel id %d is not available for QMOM (class tables "
                "have no meaning at quadrature nodes)",
                spec->coalescence);
    PBE_REQUIRE(spec->breakup >= PBE_BR_NONE && spec->breakup <= PBE_BR_CT,
                spec->breakup == PBE_BR_TABLE ? PBE_E_UNSUPPORTED : PBE_E_UNKNOWN_TYPE,
                "pbe_qmom_create: breakupKernel id %d is not available for QMOM", spec->breakup);
    PBE_REQUIRE(spec->daughter == PBE_DAUGHTER_UNIFORM_BINARY, PBE_E_UNKNOWN_TYPE,
                "pbe_qmom_create: unknown daughterParticleDistribution id %d", spec->daughter);
    pbe_qmom *q = new pbe_qmom();
    q->ctx = ctx;
    q->order = order;
    q->ks = make_dev_spec(*spec);
    *out = q;
    return PBE_OK;
}

int pbe_qmom_order(pbe_qmom *q, int *order, pbe_ctx **ctx) {
    PBE_REQUIRE(q != nullptr, PBE_E_INVALID, "pbe_qmom_order: NULL handle");
    if (order) *order = q->order;
    if (ctx) *ctx = q->ctx;
    return PBE_OK;
}

int pbe_qmom_destroy(pbe_qmom *q) {
    delete q;
    return PBE_OK;
}

int pbe_qmom_source(pbe_qmom *q, int64_t ncells, int64_t stride, const double *m, const double *eps,
                    const double *rho, const double *nu, double *S, void *cuda_stream) {
    int rc = check_q("pbe_qmom_source", q, ncells, stride, m, eps, rho, nu);
    if (rc) return rc;
    PBE_REQUIRE(ncells == 0 || S, PBE_E_INVALID, "pbe_qmom_source: S is NULL");
    DeviceGuard g(q->ctx->device);
    QmomParams p{};
    p.ncells = ncells;
    p.stride = stride;
    p.m_in = m;
    p.S_out = S;
    p.eps = eps;
    p.rho = rho;
    p.nu = nu;
    p.ks = q->ks;
    return launch_qmom(q, p, (cudaStream_t)cuda_stream);
}

int pbe_qmom_step(pbe_qmom *q, int64_t ncells, int64_t stride, double *m, const double *eps,
                  const double *rho, const double *nu, double dt, double *S, double *d,
                  void *cuda_stream) {
    int rc = check_q("pbe_qmom_step", q, ncells, stride, m, eps, rho, nu);
    if (rc) return rc;
    DeviceGuard g(q->ctx->device);
    QmomParams p{};
    p.ncells = ncells;
    p.stride = stride;
    p.m_in = m;
    p.m_out = m;
    p.S_out = S;
    p.d_out = d;
    p.eps = eps;
    p.rho = rho;
    p.nu = nu;
    p.dt = dt;
    p.do_step = 1;
    p.ks = q->ks;
    return launch_qmom(q, p, (cudaStream_t)cuda_stream);
}

int pbe_wheeler_batch(pbe_ctx *ctx, int order, int64_t ncells, int64_t stride, const double *m,
                      double *x, double *w, void *cuda_stream) {
    PBE_REQUIRE(ctx != nullptr, PBE_E_INVALID, "pbe_wheeler_batch: NULL context");
    PBE_REQUIRE(order >= 1 && order <= PBE_QMOM_MAX_ORDER, PBE_E_UNSUPPORTED,
                "pbe_wheeler_batch: order must be in [1,%d] (got %d)", PBE_QMOM_MAX_ORDER, order);
    PBE_REQUIRE(ncells >= 0 && stride >= ncells, PBE_E_INVALID, "pbe_wheeler_batch: bad ncells/stride");
    if (ncells == 0) return PBE_OK;
    PBE_REQUIRE(m && x && w, PBE_E_INVALID, "pbe_wheeler_batch: NULL pointer");
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    WheelerParams p{ncells, stride, m, x, w};
    const int grid = grid_for(ctx, ncells, 128, 16);
    switch (order) {
    case 1: wheeler_kernel<1><<<grid, 128, 0, st>>>(p); break;
    case 2: wheeler_kernel<2><<<grid, 128, 0, st>>>(p); break;
    case 3: wheeler_kernel<3><<<grid, 128, 0, st>>>(p); break;
    case 4: wheeler_kernel<4><<<grid, 128, 0, st>>>(p); break;
    }
    PBE_CUDA(cudaGetLastError());
    ctx->launches++;
    return PBE_OK;
}

static int qmom_source_host_impl(pbe_qmom *q, int64_t ncells, int64_t stride, const double *m_host,
                                 const double *eps_host, const double *rho_host, const double *nu_host,
                                 double *S_host);
int pbe_qmom_source_host(pbe_qmom *q, int64_t ncells, int64_t stride, const double *m_host,
                         const double *eps_host, const double *rho_host, const double *nu_host,
                         double *S_host) {
    const int rc = qmom_source_host_impl(q, ncells, stride, m_host, eps_host, rho_host, nu_host, S_host);
    // a failure mid-pipeline leaves copies queued that still touch the caller's buffers
    if (rc != PBE_OK && q != nullptr) drain_slot_streams(q->ctx);
    return rc;
}
static int qmom_source_host_impl(pbe_qmom *q, int64_t ncells, int64_t stride, const double *m_host,
                                 const double *eps_host, const double *rho_host, const double *nu_host,
                                 double *S_host) {
    int rc = check_q("pbe_qmom_source_host", q, ncells, stride, m_host, eps_host, rho_host, nu_host);
    if (rc) return rc;
    PBE_REQUIRE(ncells == 0 || S_host, PBE_E_INVALID, "pbe_qmom_source_host: S is NULL");
    if (ncells == 0) return PBE_OK;
    pbe_ctx *ctx = q->ctx;
    DeviceGuard g(ctx->device);
    const int nm = 2 * q->order;
    const int64_t per_cell = (int64_t)(2 * nm + 3) * sizeof(double);
    int64_t C = std::max<int64_t>(1 << 16, (((int64_t)256 << 20) / per_cell) & ~(int64_t)1023);
    C = std::min(C, (ncells + 1023) & ~(int64_t)1023);
    rc = ensure_stage(ctx, (size_t)C * per_cell);
    if (rc) return rc;
    int64_t k = 0;
    for (int64_t c0 = 0; c0 < ncells; c0 += C, ++k) {
        const int64_t nc = std::min(C, ncells - c0);
        const int s = (int)(k % kHostSlots);
        cudaStream_t st = ctx->slot_stream[s];
        double *dm = (double *)ctx->stage[s];
        double *dS = dm + (size_t)nm * C;
        double *de = dS + (size_t)nm * C;
        double *dr = de + C;
        double *dv = dr + C;
        PBE_CUDA(cudaMemcpy2DAsync(dm, (size_t)C * 8, m_host + c0, (size_t)stride * 8, (size_t)nc * 8,
                                   nm, cudaMemcpyHostToDevice, st));
        PBE_CUDA(cudaMemcpyAsync(de, eps_host + c0, (size_t)nc * 8, cudaMemcpyHostToDevice, st));
        PBE_CUDA(cudaMemcpyAsync(dr, rho_host + c0, (size_t)nc * 8, cudaMemcpyHostToDevice, st));
        PBE_CUDA(cudaMemcpyAsync(dv, nu_host + c0, (size_t)nc * 8, cudaMemcpyHostToDevice, st));
        rc = pbe_qmom_source(q, nc, C, dm, de, dr, dv, dS, st);
        if (rc) return rc;
        PBE_CUDA(cudaMemcpy2DAsync(S_host + c0, (size_t)stride * 8, dS, (size_t)C * 8, (size_t)nc * 8,
                                   nm, cudaMemcpyDeviceToHost, st));
    }
    for (int s = 0; s < kHostSlots; ++s) PBE_CUDA(cudaStreamSynchronize(ctx->slot_stream[s]));
    return PBE_OK;
}

// ---- batched scalar functors on the device (host arrays in / out) -----------------------------
int pbe_eval_coalescence(pbe_ctx *ctx, const pbe_kernel_spec *spec, int64_t count, const double *xi1,
                         const double *xi2, const double *eps, const double *rho, const double *nu,
                         double *out) {
    PBE_REQUIRE(ctx && spec, PBE_E_INVALID, "pbe_eval_coalescence: NULL argument");
    PBE_REQUIRE(spec->coalescence >= PBE_COAL_NONE && spec->coalescence <= PBE_COAL_CT,
                PBE_E_UNKNOWN_TYPE, "pbe_eval_coalescence: unknown coalescenceKernel id %d",
                spec->coalescence);
    if (count <= 0) return PBE_OK;
    PBE_REQUIRE(xi1 && xi2 && eps && rho && nu && out, PBE_E_INVALID,
                "pbe_eval_coalescence: NULL array");
    DeviceGuard g(ctx->device);
    HostBatch hb{ctx, {}};
    const size_t B = (size_t)count * sizeof(double);
    void *d1, *d2, *de, *dr, *dn, *dout;
    int rc;
    if ((rc = hb.up(xi1, B, &d1)) || (rc = hb.up(xi2, B, &d2)) || (rc = hb.up(eps, B, &de)) ||
        (rc = hb.up(rho, B, &dr)) || (rc = hb.up(nu, B, &dn)) || (rc = hb.up(nullptr, B, &dout)))
        return rc;
    eval_coal_kernel<<<(unsigned)((count + 127) / 128), 128>>>(
        make_dev_spec(*spec), count, (double *)d1, (double *)d2, (double *)de, (double *)dr,
        (double *)dn, (double *)dout);
    PBE_CUDA(cudaGetLastError());
    ctx->launches++;
    PBE_CUDA(cudaMemcpy(out, dout, B, cudaMemcpyDeviceToHost));
    return PBE_OK;
}

int pbe_eval_breakup(pbe_ctx *ctx, const pbe_kernel_spec *spec, int64_t count, const double *xi,
                     const double *rho, const double *eps, double *out) {
    PBE_REQUIRE(ctx && spec, PBE_E_INVALID, "pbe_eval_breakup: NULL argument");
    PBE_REQUIRE(spec->breakup >= PBE_BR_NONE && spec->breakup <= PBE_BR_CT, PBE_E_UNKNOWN_TYPE,
                "pbe_eval_breakup: unknown breakupKernel id %d", spec->breakup);
    if (count <= 0) return PBE_OK;
    PBE_REQUIRE(xi && rho && eps && out, PBE_E_INVALID, "pbe_eval_breakup: NULL array");
    DeviceGuard g(ctx->device);
    HostBatch hb{ctx, {}};
    const size_t B = (size_t)count * sizeof(double);
    void *dx, *dr, *de, *dout;
    int rc;
    if ((rc = hb.up(xi, B, &dx)) || (rc = hb.up(rho, B, &dr)) || (rc = hb.up(eps, B, &de)) ||
        (rc = hb.up(nullptr, B, &dout)))
        return rc;
    eval_break_kernel<<<(unsigned)((count + 127) / 128), 128>>>(
        make_dev_spec(*spec), count, (double *)dx, (double *)dr, (double *)de, (double *)dout);
    PBE_CUDA(cudaGetLastError());
    ctx->launches++;
    PBE_CUDA(cudaMemcpy(out, dout, B, cudaMemcpyDeviceToHost));
    return PBE_OK;
}

int pbe_eval_daughter_beta(pbe_ctx *ctx, int daughter, int64_t count, const double *xi1,
                           const double *xi2, double *out) {
    PBE_REQUIRE(ctx != nullptr, PBE_E_INVALID, "pbe_eval_daughter_beta: NULL context");
    PBE_REQUIRE(daughter == PBE_DAUGHTER_UNIFORM_BINARY, PBE_E_UNKNOWN_TYPE,
                "pbe_eval_daughter_beta: unknown daughterParticleDistribution id %d", daughter);
    if (count <= 0) return PBE_OK;
    PBE_REQUIRE(xi1 && xi2 && out, PBE_E_INVALID, "pbe_eval_daughter_beta: NULL array");
    DeviceGuard g(ctx->device);
    HostBatch hb{ctx, {}};
    const size_t B = (size_t)count * sizeof(double);
    void *d1, *d2, *dout;
    int rc;
    if ((rc = hb.up(xi1, B, &d1)) || (rc = hb.up(xi2, B, &d2)) || (rc = hb.up(nullptr, B, &dout)))
        return rc;
    eval_beta_kernel<<<(unsigned)((count + 127) / 128), 128>>>(count, (double *)d1, (double *)d2,
                                                               (double *)dout);
    PBE_CUDA(cudaGetLastError());
    ctx->launches++;
    PBE_CUDA(cudaMemcpy(out, dout, B, cudaMemcpyDeviceToHost));
    return PBE_OK;
}

int pbe_eval_daughter_moment(pbe_ctx *ctx, int daughter, int64_t count, const double *xi,
                             const int *k, double *out) {
    PBE_REQUIRE(ctx != nullptr, PBE_E_INVALID, "pbe_eval_daughter_moment: NULL context");
    PBE_REQUIRE(daughter == PBE_DAUGHTER_UNIFORM_BINARY, PBE_E_UNKNOWN_TYPE,
                "pbe_eval_daughter_moment: unknown daughterParticleDistribution id %d", daughter);
    if (count <= 0) return PBE_OK;
    PBE_REQUIRE(xi && k && out, PBE_E_INVALID, "pbe_eval_daughter_moment: NULL array");
    for (int64_t i = 0; i < count; ++i)
        PBE_REQUIRE(k[i] >= 0, PBE_E_INVALID, "pbe_eval_daughter_moment: k[%lld] < 0", (long long)i);
    DeviceGuard g(ctx->device);
    HostBatch hb{ctx, {}};
    const size_t B = (size_t)count * sizeof(double);
    void *dx, *dk, *dout;
    int rc;
    if ((rc = hb.up(xi, B, &dx)) || (rc = hb.up(k, (size_t)count * sizeof(int), &dk)) ||
        (rc = hb.up(nullptr, B, &dout)))
        return rc;
    eval_dmoment_kernel<<<(unsigned)((count + 127) / 128), 128>>>(count, (double *)dx, (int *)dk,
                                                                  (double *)dout);
    PBE_CUDA(cudaGetLastError());
    ctx->launches++;
    PBE_CUDA(cudaMemcpy(out, dout, B, cudaMemcpyDeviceToHost));
    return PBE_OK;
}

int pbe_eval_integrate_exp(pbe_ctx *ctx, int64_t count, const double *rate, const double *lower,
                           double *out) {
    PBE_REQUIRE(ctx != nullptr, PBE_E_INVALID, "pbe_eval_integrate_exp: NULL context");
    if (count <= 0) return PBE_OK;
    PBE_REQUIRE(rate && lower && out, PBE_E_INVALID, "pbe_eval_integrate_exp: NULL array");
    DeviceGuard g(ctx->device);
    HostBatch hb{ctx, {}};
    const size_t B = (size_t)count * sizeof(double);
    void *dr, *dl, *dout;
    int rc;
    if ((rc = hb.up(rate, B, &dr)) || (rc = hb.up(lower, B, &dl)) || (rc = hb.up(nullptr, B, &dout)))
        return rc;
    eval_integrate_exp_kernel<<<(unsigned)((count + 63) / 64), 64>>>(count, (double *)dr,
                                                                     (double *)dl, (double *)dout);
    PBE_CUDA(cudaGetLastError());
    ctx->launches++;
    PBE_CUDA(cudaMemcpy(out, dout, B, cudaMemcpyDeviceToHost));
    return PBE_OK;
}

int pbe_field_avg_max_min(pbe_ctx *ctx, int64_t ncells, const double *f, const double *vcell,
                          double *out4, void *cuda_stream) {
    PBE_REQUIRE(ctx && out4, PBE_E_INVALID, "pbe_field_avg_max_min: NULL argument");
    PBE_REQUIRE(ncells >= 0 && (ncells == 0 || f), PBE_E_INVALID, "pbe_field_avg_max_min: bad field");
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    const int nparts = grid_for(ctx, std::max<int64_t>(ncells, 1), 2048, 2);
    double *part = nullptr;
    PBE_CUDA(cudaMallocAsync(&part, (size_t)nparts * 4 * sizeof(double), st));
    avg_max_min_kernel<<<nparts, 256, 0, st>>>(ncells, f, vcell, part);
    avg_max_min_final_kernel<<<1, 32, 0, st>>>(nparts, part, out4);
    PBE_CUDA(cudaGetLastError());
    PBE_CUDA(cudaFreeAsync(part, st));
    ctx->launches += 2;
    return PBE_OK;
}

} // extern "C"
