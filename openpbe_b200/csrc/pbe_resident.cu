// pbe_resident.cu -- resident-state correct(): the method owns its class / moment fields on the
// device, as the reference's methods own theirs (MOC.H:66, MOC.C:69-77; QMOM.H:62, QMOM.C:72-79),
// and one time step moves only this step's phase fields in (epsilon, rho, nu, alpha) and d out
// (PBEDiameter.H:95-99: method_->correct(); d_ = method_->d()).
//
// Pipeline of one correct() without a mesh (zero-flux update, MOC.C:259-272 with phi = 0): the
// cells are cut into chunks of whole waves of the persistent MOC grid;
//     copy-in stream : H2D eps, rho, nu, alpha of chunk k+1
//     compute stream : pbe_moc_step on chunk k (sources, limiter, update, d, continuity, moments)
//     copy-out stream: D2H d of chunk k-1
//     side stream    : per-class min / max partials of chunk k-1 (HBM-bound, co-resident with the
//                      FP64-bound step kernel of chunk k)
// so that the PCIe traffic (40 B/cell) hides behind the kernel.  With an attached mesh the source
// kernel runs per chunk under the uploads, then the transport update of all fields
// (MOC::solveWithMULES, MOC.C:198-272), d, continuity and the diagnostics.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>

#include "pbe_common.h"
#include "pbe_math.cuh"

using namespace pbe;

namespace {

constexpr int kMaxChunks = 64;

struct Pipeline {
    pbe_ctx *ctx = nullptr;
    cudaStream_t s_in = nullptr, s_comp = nullptr, s_out = nullptr, s_side = nullptr;
    cudaEvent_t ev_in[kMaxChunks] = {}, ev_comp[kMaxChunks] = {};
    cudaEvent_t ev_start = nullptr, ev_side = nullptr, ev_out = nullptr;
    bool ok = false;

    int init(pbe_ctx *c) {
        ctx = c;
        s_in = c->slot_stream[0];
        s_comp = c->slot_stream[1];
        s_out = c->slot_stream[2];
        PBE_CUDA(cudaStreamCreateWithFlags(&s_side, cudaStreamNonBlocking));
        for (int k = 0; k < kMaxChunks; ++k) {
            PBE_CUDA(cudaEventCreateWithFlags(&ev_in[k], cudaEventDisableTiming));
            PBE_CUDA(cudaEventCreateWithFlags(&ev_comp[k], cudaEventDisableTiming));
        }
        PBE_CUDA(cudaEventCreateWithFlags(&ev_start, cudaEventDisableTiming));
        PBE_CUDA(cudaEventCreateWithFlags(&ev_side, cudaEventDisableTiming));
        PBE_CUDA(cudaEventCreateWithFlags(&ev_out, cudaEventDisableTiming));
        ok = true;
        return PBE_OK;
    }
    void destroy() {
        if (s_side) cudaStreamDestroy(s_side);
        for (int k = 0; k < kMaxChunks; ++k) {
            if (ev_in[k]) cudaEventDestroy(ev_in[k]);
            if (ev_comp[k]) cudaEventDestroy(ev_comp[k]);
        }
        if (ev_start) cudaEventDestroy(ev_start);
        if (ev_side) cudaEventDestroy(ev_side);
        if (ev_out) cudaEventDestroy(ev_out);
    }
    // wait for everything this pipeline may have queued (also the error path: queued copies still
    // touch the caller's buffers)
    void drain() {
        drain_slot_streams(ctx);
        if (s_side) cudaStreamSynchronize(s_side);
    }
};

// Cells per chunk: whole `quantum`s (one wave of the persistent grid), about `want` chunks.
int64_t chunk_cells(int64_t ncells, int64_t quantum, int want) {
    if (const char *e = getenv("PBE_RESIDENT_CHUNKS")) want = std::max(1, atoi(e));
    want = std::min(want, kMaxChunks);
    const int64_t quanta = (ncells + quantum - 1) / quantum;
    const int64_t per = std::max<int64_t>(1, (quanta + want - 1) / want);
    return per * quantum;
}

// ---- row statistics: for each of `nrows` rows of f [nrows][stride], over cells [0, ncells):
// {sum f V, sum V, max f, min f} (Utility.C:4-10 printAvgMaxMin; MOC.C:181-183 min(n_k)).
// Stage 1 writes one partial per (row, CTA) at part[(row * parts_total + part0 + blockIdx.x) * 4].
__global__ void __launch_bounds__(256) rows_stats_partial_kernel(int64_t ncells, int64_t stride,
                                                                 const double *__restrict__ f,
                                                                 const double *__restrict__ vcell, int want_sum,
                                                                 double *__restrict__ part, int parts_total,
                                                                 int part0) {
    const int k = blockIdx.y;
    const double *row = f + (size_t)k * stride;
    double sfv = 0.0, sv = 0.0, mx = -INFINITY, mn = INFINITY;
    const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
    for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c < ncells; c += nthreads) {
        const double v = __ldg(row + c);
        if (want_sum) {
            const double V = vcell ? __ldg(vcell + c) : 1.0;
            sfv = fma_(v, V, sfv);
            sv += V;
        }
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
        double *o = part + ((size_t)k * parts_total + part0 + blockIdx.x) * 4;
        o[0] = sfv;
        o[1] = sv;
        o[2] = mx;
        o[3] = mn;
    }
}
// Stage 2: one warp per row folds its partials -> out[4][rows]
__global__ void rows_stats_final_kernel(int nparts, const double *__restrict__ part, double *__restrict__ out) {
    const int k = blockIdx.x;
    double sfv = 0.0, sv = 0.0, mx = -INFINITY, mn = INFINITY;
    for (int b = threadIdx.x; b < nparts; b += 32) {
        const double *p = part + ((size_t)k * nparts + b) * 4;
        sfv += p[0];
        sv += p[1];
        mx = fmax(mx, p[2]);
        mn = fmin(mn, p[3]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sfv += __shfl_xor_sync(0xffffffffu, sfv, o);
        sv += __shfl_xor_sync(0xffffffffu, sv, o);
        mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    }
    if (threadIdx.x == 0) { // structure of arrays [sum f V | sum V | min | max][rows]: each is one NCCL all-reduce
        const size_t nr = gridDim.x;
        out[0 * nr + k] = sfv;
        out[1 * nr + k] = sv;
        out[2 * nr + k] = mn;
        out[3 * nr + k] = mx;
    }
}

int stats_parts(pbe_ctx *ctx, int64_t ncells) {
    return (int)std::max<int64_t>(1, std::min<int64_t>((ncells + 2047) / 2048, (int64_t)ctx->sm_count * 2));
}

// Continuity check and global moments of the class fields after a transport update (the zero-flux
// step kernel fuses them): err = max_c |sum_k xi_k n_k / V - alpha| (MOC.C:254-279), moments[q] +=
// sum_c V_c sum_k xi_k^q n_k.  One thread per cell; block reduction, one atomic per CTA per value.
__global__ void __launch_bounds__(256) moc_continuity_kernel(int N, int64_t ncells, int64_t stride,
                                                             const double *__restrict__ xi_k, // [N] class volumes
                                                             const double *__restrict__ n,
                                                             const double *__restrict__ alpha,
                                                             const double *__restrict__ vcell, int n_moments,
                                                             double *__restrict__ moments,
                                                             double *__restrict__ cont_err_max) {
    double err = 0.0, mom[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const int64_t nthreads = (int64_t)gridDim.x * blockDim.x;
    for (int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; c < ncells; c += nthreads) {
        double cm[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        double vol = 0.0;
        for (int k = 0; k < N; ++k) {
            const double v = __ldg(n + (size_t)k * stride + c);
            const double xi = __ldg(xi_k + k); // (k + 1) xi1 on the reference's grid (MOC.C:79), or the pivot
            vol += xi * v;                   // volumeSum += xi_k N_k, MOC.C:272
            double pw = v;
            for (int q = 0; q < n_moments; ++q) {
                cm[q] += pw;
                pw *= xi;
            }
        }
        const double V = vcell ? __ldg(vcell + c) : 1.0;
        if (alpha && vcell) err = fmax(err, fabs(vol / V - __ldg(alpha + c)));
        for (int q = 0; q < n_moments; ++q) mom[q] = fma_(V, cm[q], mom[q]);
    }
    __shared__ double sh[8][9];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        err = fmax(err, __shfl_xor_sync(0xffffffffu, err, o));
        for (int q = 0; q < 8; ++q) mom[q] += __shfl_xor_sync(0xffffffffu, mom[q], o);
    }
    if ((threadIdx.x & 31) == 0) {
        sh[threadIdx.x >> 5][8] = err;
        for (int q = 0; q < 8; ++q) sh[threadIdx.x >> 5][q] = mom[q];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) {
            err = fmax(err, sh[w][8]);
            for (int q = 0; q < 8; ++q) mom[q] += sh[w][q];
        }
        for (int q = 0; q < n_moments; ++q) atomicAdd(moments + q, mom[q]);
        // err >= 0: the IEEE bit pattern of non-negative doubles orders like the integers
        if (cont_err_max && alpha && vcell)
            atomicMax(reinterpret_cast<unsigned long long *>(cont_err_max),
                      (unsigned long long)__double_as_longlong(err));
    }
}

// d = (6/pi m1/m0)^(1/3) at construction (QMOM.C:81)
__global__ void qmom_d_kernel(int64_t ncells, const double *__restrict__ m0, const double *__restrict__ m1,
                              double *__restrict__ d) {
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c < ncells) d[c] = pow(6.0 / M_PI * m1[c] / m0[c], 1.0 / 3.0);
}

int copy_rows_h2d(double *dst, int64_t dstride, const double *src, int64_t sstride, int64_t ncells, int rows,
                  cudaStream_t st) {
    PBE_CUDA(cudaMemcpy2DAsync(dst, (size_t)dstride * 8, src, (size_t)sstride * 8, (size_t)ncells * 8, rows,
                               cudaMemcpyHostToDevice, st));
    return PBE_OK;
}
int copy_rows_d2h(double *dst, int64_t dstride, const double *src, int64_t sstride, int64_t ncells, int rows,
                  cudaStream_t st) {
    PBE_CUDA(cudaMemcpy2DAsync(dst, (size_t)dstride * 8, src, (size_t)sstride * 8, (size_t)ncells * 8, rows,
                               cudaMemcpyDeviceToHost, st));
    return PBE_OK;
}

} // namespace

// ==================================================================================================
// MOC
// ==================================================================================================
struct pbe_moc_state {
    pbe_moc *moc = nullptr;
    pbe_ctx *ctx = nullptr;
    int N = 0;
    double xi1 = 0.0;
    const double *xi_dev = nullptr; // class volumes, owned by the pbe_moc handle
    int64_t nc = 0;
    int64_t quantum = 1; // cells of one wave of the persistent grid
    double *n = nullptr, *eps = nullptr, *rho = nullptr, *nu = nullptr, *alpha = nullptr, *vcell = nullptr,
           *d = nullptr;
    bool have_vcell = false, have_rho = false, have_nu = false;
    // transport (attached mesh)
    pbe_mesh *mesh = nullptr;
    int scheme = 0;
    int64_t nf = 0, nbnd = 0;
    double *S = nullptr, *nNew = nullptr, *phi = nullptr, *psi_b = nullptr, *psi_max = nullptr;
    // usingMULES false (MOC.C:58, 177-180): solveWithFVM, the implicit solve
    bool implicit = false;
    pbe_fvm_controls fvm{};
    std::vector<pbe_fvm_perf> perf;
    // diagnostics: [N][4] row statistics, then cont_err, then 8 moments
    double *diag = nullptr, *part = nullptr, *diag_host = nullptr;
    int parts_total = 0;
    Pipeline pipe;
};

namespace {

void free_moc_state(pbe_moc_state *s) {
    if (!s) return;
    DeviceGuard g(s->ctx->device);
    s->pipe.drain();
    for (double *p : {s->n, s->eps, s->rho, s->nu, s->alpha, s->vcell, s->d, s->S, s->nNew, s->phi, s->psi_b,
                      s->psi_max, s->diag, s->part})
        if (p) cudaFree(p);
    if (s->diag_host) cudaFreeHost(s->diag_host);
    s->pipe.destroy();
    delete s;
}

// Row statistics of `rows` rows over cells [c0, c0 + cn) into partial slots [slot0, slot0 + parts):
// every chunk launches the same `parts` CTAs per row (CTAs beyond the cells write neutral values),
// so the final fold can run over all slots.
int launch_stats_partial(pbe_ctx *ctx, int rows, int64_t cn, int64_t stride, const double *f, const double *vcell,
                         bool want_sum, double *part, int parts_total, int slot0, int parts, cudaStream_t st) {
    rows_stats_partial_kernel<<<dim3(parts, rows), 256, 0, st>>>(cn, stride, f, vcell, want_sum ? 1 : 0, part,
                                                                parts_total, slot0);
    PBE_CUDA(cudaGetLastError());
    ctx->launches++;
    return PBE_OK;
}

int ensure_parts(double **part, int *have, int rows, int need) {
    if (*have >= need) return PBE_OK;
    if (*part) cudaFree(*part);
    *part = nullptr;
    *have = 0;
    PBE_CUDA(cudaMalloc((void **)part, (size_t)rows * need * 4 * sizeof(double)));
    *have = need;
    return PBE_OK;
}

int moc_correct_impl(pbe_moc_state *s, const double *eps_h, const double *rho_h, const double *nu_h,
                     const double *alpha_h, const double *phi_h, const double *psi_b_h, int64_t stride_b,
                     double dt, double *d_h, pbe_moc_diag *diag) {
    pbe_ctx *ctx = s->ctx;
    Pipeline &P = s->pipe;
    const int N = s->N;
    const int64_t nc = s->nc;
    const bool want_minmax = diag && (diag->class_min || diag->class_max);
    const bool want_cont = alpha_h != nullptr && s->have_vcell;
    const int nmom = (diag && s->have_vcell) ? diag->n_moments : 0;
    double *cont_dev = s->diag + (size_t)N * 4; // [1] continuity error, then [8] moments
    double *mom_dev = cont_dev + 1;

    const int64_t C = chunk_cells(nc, s->quantum, 12);
    const int nchunks = (int)((nc + C - 1) / C);
    PBE_REQUIRE(nchunks <= kMaxChunks, PBE_E_UNSUPPORTED, "pbe_moc_correct_resident: %d chunks", nchunks);
    const int parts = stats_parts(ctx, C);
    // with a mesh the statistics run once over all cells after the transport update
    const int parts_all = s->mesh ? stats_parts(ctx, nc) : parts * nchunks;
    if (want_minmax) {
        int rc = ensure_parts(&s->part, &s->parts_total, N, parts_all);
        if (rc) return rc;
    }

    PBE_CUDA(cudaMemsetAsync(cont_dev, 0, 9 * sizeof(double), P.s_comp));
    PBE_CUDA(cudaEventRecord(P.ev_start, P.s_comp));
    PBE_CUDA(cudaStreamWaitEvent(P.s_in, P.ev_start, 0));
    PBE_CUDA(cudaStreamWaitEvent(P.s_out, P.ev_start, 0));
    PBE_CUDA(cudaStreamWaitEvent(P.s_side, P.ev_start, 0));

    if (s->mesh && s->nf)
        PBE_CUDA(cudaMemcpyAsync(s->phi, phi_h, (size_t)s->nf * 8, cudaMemcpyHostToDevice, P.s_in));
    if (s->mesh && psi_b_h && s->nbnd) {
        int rc = copy_rows_h2d(s->psi_b, s->nbnd, psi_b_h, stride_b, s->nbnd, N, P.s_in);
        if (rc) return rc;
    }

    for (int k = 0; k < nchunks; ++k) {
        const int64_t c0 = (int64_t)k * C, cn = std::min(C, nc - c0);
        // ---- copy-in: this step's phase fields of chunk k ---------------------------------------
        PBE_CUDA(cudaMemcpyAsync(s->eps + c0, eps_h + c0, (size_t)cn * 8, cudaMemcpyHostToDevice, P.s_in));
        if (rho_h)
            PBE_CUDA(cudaMemcpyAsync(s->rho + c0, rho_h + c0, (size_t)cn * 8, cudaMemcpyHostToDevice, P.s_in));
        if (nu_h)
            PBE_CUDA(cudaMemcpyAsync(s->nu + c0, nu_h + c0, (size_t)cn * 8, cudaMemcpyHostToDevice, P.s_in));
        if (alpha_h)
            PBE_CUDA(cudaMemcpyAsync(s->alpha + c0, alpha_h + c0, (size_t)cn * 8, cudaMemcpyHostToDevice, P.s_in));
        PBE_CUDA(cudaEventRecord(P.ev_in[k], P.s_in));
        // ---- compute --------------------------------------------------------------------------------
        PBE_CUDA(cudaStreamWaitEvent(P.s_comp, P.ev_in[k], 0));
        if (s->mesh) { // sources only; the update needs every cell's neighbours
            int rc = pbe_moc_source(s->moc, cn, nc, s->n + c0, s->eps + c0, s->rho + c0, s->nu + c0, dt, s->S + c0,
                                    P.s_comp);
            if (rc) return rc;
            continue;
        }
        pbe_moc_io io{};
        io.d = s->d + c0;
        if (want_cont) {
            io.alpha = s->alpha + c0;
            io.vcell = s->vcell + c0;
            io.cont_err_max = cont_dev;
        }
        if (nmom > 0) {
            io.vcell = s->vcell + c0;
            io.moments = mom_dev;
            io.n_moments = nmom;
        }
        int rc = pbe_moc_step(s->moc, cn, nc, s->n + c0, s->eps + c0, s->rho + c0, s->nu + c0, dt, &io, P.s_comp);
        if (rc) return rc;
        PBE_CUDA(cudaEventRecord(P.ev_comp[k], P.s_comp));
        // ---- copy-out: d of chunk k -----------------------------------------------------------------
        if (d_h) {
            PBE_CUDA(cudaStreamWaitEvent(P.s_out, P.ev_comp[k], 0));
            PBE_CUDA(cudaMemcpyAsync(d_h + c0, s->d + c0, (size_t)cn * 8, cudaMemcpyDeviceToHost, P.s_out));
        }
        // ---- min(n_k) / max(n_k) partials of the updated chunk, beside the next chunk's kernel ------
        if (want_minmax) {
            PBE_CUDA(cudaStreamWaitEvent(P.s_side, P.ev_comp[k], 0));
            rc = launch_stats_partial(ctx, N, cn, nc, s->n + c0, nullptr, false, s->part, parts_all, k * parts, parts,
                                      P.s_side);
            if (rc) return rc;
        }
    }

    if (s->mesh) {
        // MOC::solveWithMULES (MOC.C:198-272): transport + source update of all fields, then d,
        // the continuity check (MOC.C:254-279) and the per-class statistics of the new fields
        PBE_CUDA(cudaStreamWaitEvent(P.s_comp, P.ev_in[nchunks - 1], 0));
        int rc;
        if (s->implicit) { // MOC::solveWithFVM (MOC.C:186-198): fvm::ddt + fvm::div == S, all classes at once
            s->perf.resize((size_t)N);
            rc = pbe_transport_implicit(s->mesh, s->scheme, N, nc, s->n, s->phi, psi_b_h ? s->psi_b : nullptr, s->nbnd,
                                        s->S, dt, &s->fvm, s->nNew, s->perf.data(), P.s_comp);
        } else if (s->scheme == PBE_DIV_UPWIND) { // the limited correction is identically zero
            rc = pbe_transport_explicit(s->mesh, s->scheme, N, nc, s->n, s->phi, psi_b_h ? s->psi_b : nullptr, s->nbnd,
                                        s->S, dt, s->nNew, P.s_comp);
        } else { // MULES::limit against max(alpha V / xi_k) (MOC.C:238-249), limitSum, explicitSolve
            rc = pbe_moc_psi_max(s->moc, nc, s->alpha, s->vcell, s->psi_max, P.s_comp);
            if (rc) return rc;
            rc = pbe_transport_mules(s->mesh, s->scheme, N, nc, s->n, s->phi, psi_b_h ? s->psi_b : nullptr, s->nbnd, s->S,
                                     s->psi_max, 0.0, 3, dt, s->nNew, P.s_comp);
        }
        if (rc) return rc;
        std::swap(s->n, s->nNew);
        rc = pbe_moc_d(s->moc, nc, nc, s->n, s->d, P.s_comp);
        if (rc) return rc;
        PBE_CUDA(cudaEventRecord(P.ev_comp[0], P.s_comp));
        if (d_h) {
            PBE_CUDA(cudaStreamWaitEvent(P.s_out, P.ev_comp[0], 0));
            PBE_CUDA(cudaMemcpyAsync(d_h, s->d, (size_t)nc * 8, cudaMemcpyDeviceToHost, P.s_out));
        }
        if (want_cont || nmom > 0) {
            const int grid = (int)std::min<int64_t>((nc + 255) / 256, (int64_t)ctx->sm_count * 8);
            moc_continuity_kernel<<<grid, 256, 0, P.s_comp>>>(N, nc, nc, s->xi_dev, s->n, want_cont ? s->alpha : nullptr,
                                                              s->have_vcell ? s->vcell : nullptr, nmom, mom_dev,
                                                              cont_dev);
            PBE_CUDA(cudaGetLastError());
            ctx->launches++;
        }
        if (want_minmax) {
            PBE_CUDA(cudaStreamWaitEvent(P.s_side, P.ev_comp[0], 0));
            rc = launch_stats_partial(ctx, N, nc, nc, s->n, nullptr, false, s->part, parts_all, 0, parts_all, P.s_side);
            if (rc) return rc;
        }
    }

    // ---- fold the diagnostics and bring them to the host ------------------------------------------
    if (want_minmax) {
        rows_stats_final_kernel<<<N, 32, 0, P.s_side>>>(parts_all, s->part, s->diag);
        PBE_CUDA(cudaGetLastError());
        ctx->launches++;
    }
    PBE_CUDA(cudaEventRecord(P.ev_side, P.s_side));
    PBE_CUDA(cudaStreamWaitEvent(P.s_comp, P.ev_side, 0));
    if (diag) {
        // with a communicator on the context the printed quantities are the GLOBAL ones, as the reference's
        // gMin / gMax / gSum are (MOC.C:181-183, 238, 276-279): [min N] | [max N, continuity] | [moments 8]
        int rc = pbe_reduce_diagnostics(ctx, 8, mom_dev, want_minmax ? N : 0, s->diag + (size_t)2 * N, want_minmax ? N + 1 : 1,
                                        want_minmax ? s->diag + (size_t)3 * N : cont_dev, P.s_comp);
        if (rc) return rc;
        PBE_CUDA(cudaMemcpyAsync(s->diag_host, s->diag, ((size_t)N * 4 + 9) * sizeof(double), cudaMemcpyDeviceToHost,
                                 P.s_comp));
    }
    PBE_CUDA(cudaStreamSynchronize(P.s_in));
    PBE_CUDA(cudaStreamSynchronize(P.s_comp));
    PBE_CUDA(cudaStreamSynchronize(P.s_out));
    if (rho_h) s->have_rho = true;
    if (nu_h) s->have_nu = true;
    if (diag) {
        for (int k = 0; k < N; ++k) {
            if (diag->class_min) diag->class_min[k] = s->diag_host[(size_t)2 * N + k];
            if (diag->class_max) diag->class_max[k] = s->diag_host[(size_t)3 * N + k];
        }
        diag->cont_err_max = want_cont ? s->diag_host[(size_t)N * 4] : 0.0;
        for (int q = 0; q < 8; ++q) diag->moments[q] = (q < nmom) ? s->diag_host[(size_t)N * 4 + 1 + q] : 0.0;
    }
    return PBE_OK;
}

} // namespace

extern "C" {

int pbe_moc_state_create(pbe_moc *moc, int64_t ncells, const double *vcell_host, pbe_moc_state **out) {
    PBE_REQUIRE(moc && out, PBE_E_INVALID, "pbe_moc_state_create: NULL argument");
    *out = nullptr;
    PBE_REQUIRE(ncells >= 0, PBE_E_INVALID, "pbe_moc_state_create: ncells < 0");
    int N = 0, block = 128, ctas = 1;
    pbe_ctx *ctx = nullptr;
    int rc = pbe_moc_num_classes(moc, &N, &ctx);
    if (rc) return rc;
    pbe_moc_get_kernel(moc, nullptr, &block, &ctas);
    DeviceGuard g(ctx->device);
    pbe_moc_state *s = new pbe_moc_state();
    s->moc = moc;
    s->ctx = ctx;
    s->N = N;
    s->nc = ncells;
    s->quantum = (int64_t)block * ctx->sm_count * ctas;
    pbe_moc_coeffs(moc, &s->xi1, nullptr);
    pbe_moc_grid_device(moc, &s->xi_dev);
    const size_t cells = (size_t)std::max<int64_t>(ncells, 1);
    cudaError_t e = cudaMalloc((void **)&s->n, (size_t)N * cells * 8);
    for (double **p : {&s->eps, &s->rho, &s->nu, &s->alpha, &s->vcell, &s->d})
        if (e == cudaSuccess) e = cudaMalloc((void **)p, cells * 8);
    if (e == cudaSuccess) e = cudaMalloc((void **)&s->diag, ((size_t)N * 4 + 9) * sizeof(double));
    if (e == cudaSuccess) e = cudaHostAlloc((void **)&s->diag_host, ((size_t)N * 4 + 9) * sizeof(double), 0);
    if (e == cudaSuccess) e = cudaMemset(s->n, 0, (size_t)N * cells * 8);
    if (e == cudaSuccess && vcell_host && ncells) {
        e = cudaMemcpy(s->vcell, vcell_host, (size_t)ncells * 8, cudaMemcpyHostToDevice);
        s->have_vcell = true;
    }
    if (e != cudaSuccess) {
        rc = cuda_fail(e, "pbe_moc_state_create", __FILE__, __LINE__);
        free_moc_state(s);
        return rc;
    }
    rc = s->pipe.init(ctx);
    if (rc) {
        free_moc_state(s);
        return rc;
    }
    *out = s;
    return PBE_OK;
}

int pbe_moc_state_destroy(pbe_moc_state *s) {
    free_moc_state(s);
    return PBE_OK;
}

int pbe_moc_state_upload(pbe_moc_state *s, int64_t stride, const double *n_host) {
    PBE_REQUIRE(s && (n_host || s->nc == 0) && stride >= s->nc, PBE_E_INVALID, "pbe_moc_state_upload: bad argument");
    if (s->nc == 0) return PBE_OK;
    DeviceGuard g(s->ctx->device);
    int rc = copy_rows_h2d(s->n, s->nc, n_host, stride, s->nc, s->N, s->pipe.s_comp);
    if (rc) return rc;
    PBE_CUDA(cudaStreamSynchronize(s->pipe.s_comp));
    return PBE_OK;
}

int pbe_moc_state_download(pbe_moc_state *s, int64_t stride, double *n_host) {
    PBE_REQUIRE(s && (n_host || s->nc == 0) && stride >= s->nc, PBE_E_INVALID, "pbe_moc_state_download: bad argument");
    if (s->nc == 0) return PBE_OK;
    DeviceGuard g(s->ctx->device);
    int rc = copy_rows_d2h(n_host, stride, s->n, s->nc, s->nc, s->N, s->pipe.s_comp);
    if (rc) return rc;
    PBE_CUDA(cudaStreamSynchronize(s->pipe.s_comp));
    return PBE_OK;
}

int pbe_moc_state_upload_range(pbe_moc_state *s, int64_t cell0, int64_t count, int64_t stride,
                               const double *n_host) {
    PBE_REQUIRE(s && cell0 >= 0 && count >= 0 && cell0 + count <= s->nc && stride >= count && (n_host || count == 0),
                PBE_E_INVALID, "pbe_moc_state_upload_range: bad argument");
    if (count == 0) return PBE_OK;
    DeviceGuard g(s->ctx->device);
    int rc = copy_rows_h2d(s->n + cell0, s->nc, n_host, stride, count, s->N, s->pipe.s_comp);
    if (rc) return rc;
    PBE_CUDA(cudaStreamSynchronize(s->pipe.s_comp));
    return PBE_OK;
}

int pbe_moc_state_download_range(pbe_moc_state *s, int64_t cell0, int64_t count, int64_t stride, double *n_host) {
    PBE_REQUIRE(s && cell0 >= 0 && count >= 0 && cell0 + count <= s->nc && stride >= count && (n_host || count == 0),
                PBE_E_INVALID, "pbe_moc_state_download_range: bad argument");
    if (count == 0) return PBE_OK;
    DeviceGuard g(s->ctx->device);
    int rc = copy_rows_d2h(n_host, stride, s->n + cell0, s->nc, count, s->N, s->pipe.s_comp);
    if (rc) return rc;
    PBE_CUDA(cudaStreamSynchronize(s->pipe.s_comp));
    return PBE_OK;
}

int pbe_moc_state_d(pbe_moc_state *s, double *d_host) {
    PBE_REQUIRE(s && (d_host || s->nc == 0), PBE_E_INVALID, "pbe_moc_state_d: bad argument");
    if (s->nc == 0) return PBE_OK;
    DeviceGuard g(s->ctx->device);
    int rc = pbe_moc_d(s->moc, s->nc, s->nc, s->n, s->d, s->pipe.s_comp);
    if (rc) return rc;
    PBE_CUDA(cudaMemcpyAsync(d_host, s->d, (size_t)s->nc * 8, cudaMemcpyDeviceToHost, s->pipe.s_comp));
    PBE_CUDA(cudaStreamSynchronize(s->pipe.s_comp));
    return PBE_OK;
}

int pbe_moc_state_device(pbe_moc_state *s, double **n_dev, int64_t *stride, double **d_dev) {
    PBE_REQUIRE(s != nullptr, PBE_E_INVALID, "pbe_moc_state_device: NULL state");
    if (n_dev) *n_dev = s->n;
    if (stride) *stride = s->nc;
    if (d_dev) *d_dev = s->d;
    return PBE_OK;
}

int pbe_moc_state_attach_mesh(pbe_moc_state *s, pbe_mesh *mesh, int scheme) {
    PBE_REQUIRE(s && mesh, PBE_E_INVALID, "pbe_moc_state_attach_mesh: NULL argument");
    int64_t mc = 0, nint = 0, nbnd = 0;
    int rc = pbe_mesh_info(mesh, &mc, &nint, &nbnd, nullptr);
    if (rc) return rc;
    PBE_REQUIRE(mc == s->nc, PBE_E_INVALID, "pbe_moc_state_attach_mesh: the mesh has %lld cells, the state %lld",
                (long long)mc, (long long)s->nc);
    DeviceGuard g(s->ctx->device);
    const size_t field = (size_t)s->N * std::max<int64_t>(s->nc, 1);
    for (double **p : {&s->S, &s->nNew, &s->phi, &s->psi_b, &s->psi_max})
        if (*p) {
            cudaFree(*p);
            *p = nullptr;
        }
    PBE_CUDA(cudaMalloc((void **)&s->psi_max, (size_t)s->N * 8));
    PBE_CUDA(cudaMalloc((void **)&s->S, field * 8));
    PBE_CUDA(cudaMalloc((void **)&s->nNew, field * 8));
    PBE_CUDA(cudaMalloc((void **)&s->phi, (size_t)std::max<int64_t>(nint + nbnd, 1) * 8));
    PBE_CUDA(cudaMalloc((void **)&s->psi_b, (size_t)s->N * std::max<int64_t>(nbnd, 1) * 8));
    s->mesh = mesh;
    s->scheme = scheme;
    s->nf = nint + nbnd;
    s->nbnd = nbnd;
    return PBE_OK;
}

int pbe_moc_state_set_implicit(pbe_moc_state *s, const pbe_fvm_controls *controls) {
    PBE_REQUIRE(s != nullptr, PBE_E_INVALID, "pbe_moc_state_set_implicit: NULL state");
    PBE_REQUIRE(controls == nullptr || s->mesh != nullptr, PBE_E_INVALID,
                "pbe_moc_state_set_implicit: attach a mesh first (pbe_moc_state_attach_mesh)");
    s->implicit = controls != nullptr;
    if (controls) s->fvm = *controls;
    s->perf.clear();
    return PBE_OK;
}

int pbe_moc_state_solver_performance(pbe_moc_state *s, pbe_fvm_perf *perf) {
    PBE_REQUIRE(s && perf, PBE_E_INVALID, "pbe_moc_state_solver_performance: NULL argument");
    PBE_REQUIRE((int)s->perf.size() == s->N, PBE_E_INVALID,
                "pbe_moc_state_solver_performance: no implicit step has run on this state");
    std::copy(s->perf.begin(), s->perf.end(), perf);
    return PBE_OK;
}

int pbe_moc_correct_resident(pbe_moc_state *s, const double *eps_host, const double *rho_host,
                             const double *nu_host, const double *alpha_host, const double *phi_host,
                             const double *psi_b_host, int64_t stride_b, double dt, double *d_host,
                             pbe_moc_diag *diag) {
    PBE_REQUIRE(s != nullptr, PBE_E_INVALID, "pbe_moc_correct_resident: NULL state");
    PBE_REQUIRE(dt > 0.0, PBE_E_INVALID, "pbe_moc_correct_resident: dt must be positive (got %g)", dt);
    PBE_REQUIRE(diag == nullptr || (diag->n_moments >= 0 && diag->n_moments <= 8), PBE_E_INVALID,
                "pbe_moc_correct_resident: n_moments must be in [0,8]");
    if (s->nc == 0) return PBE_OK;
    PBE_REQUIRE(eps_host != nullptr, PBE_E_INVALID, "pbe_moc_correct_resident: eps is NULL");
    PBE_REQUIRE((rho_host || s->have_rho) && (nu_host || s->have_nu), PBE_E_INVALID,
                "pbe_moc_correct_resident: rho / nu may only be NULL after a call that supplied them");
    PBE_REQUIRE(s->mesh == nullptr || s->nf == 0 || phi_host != nullptr, PBE_E_INVALID,
                "pbe_moc_correct_resident: a mesh is attached, phi is NULL");
    PBE_REQUIRE(s->mesh != nullptr || (phi_host == nullptr && psi_b_host == nullptr), PBE_E_INVALID,
                "pbe_moc_correct_resident: phi / psi_b without an attached mesh (pbe_moc_state_attach_mesh)");
    PBE_REQUIRE(psi_b_host == nullptr || stride_b >= s->nbnd, PBE_E_INVALID,
                "pbe_moc_correct_resident: stride_b < boundary faces");
    PBE_REQUIRE(s->mesh == nullptr || s->implicit || s->scheme == PBE_DIV_UPWIND || (alpha_host != nullptr && s->have_vcell),
                PBE_E_INVALID,
                "pbe_moc_correct_resident: the MULES limiter's bound max(alpha V / xi_k) (MOC.C:238) needs alpha "
                "and the cell volumes");
    DeviceGuard g(s->ctx->device);
    const int rc = moc_correct_impl(s, eps_host, rho_host, nu_host, alpha_host, phi_host, psi_b_host, stride_b, dt,
                                    d_host, diag);
    if (rc != PBE_OK) s->pipe.drain(); // queued copies still touch the caller's buffers
    return rc;
}

} // extern "C"

// ==================================================================================================
// QMOM
// ==================================================================================================
struct pbe_qmom_state {
    pbe_qmom *q = nullptr;
    pbe_ctx *ctx = nullptr;
    int nm = 0; // 2 * quadratureOrder
    int64_t nc = 0;
    double *m = nullptr, *S = nullptr, *eps = nullptr, *rho = nullptr, *nu = nullptr, *vcell = nullptr, *d = nullptr;
    bool have_vcell = false, have_rho = false, have_nu = false;
    pbe_mesh *mesh = nullptr;
    int scheme = 0;
    int64_t nf = 0, nbnd = 0;
    double *mNew = nullptr, *phi = nullptr, *psi_b = nullptr, *psi_max = nullptr;
    bool implicit = false; // usingMULES false (QMOM.C:71, 112-115): solveWithFVM
    pbe_fvm_controls fvm{};
    std::vector<pbe_fvm_perf> perf;
    double *diag = nullptr, *part = nullptr, *diag_host = nullptr; // [2 nm + 1][4]
    int parts_total = 0;
    Pipeline pipe;
};

namespace {

void free_qmom_state(pbe_qmom_state *s) {
    if (!s) return;
    DeviceGuard g(s->ctx->device);
    s->pipe.drain();
    for (double *p : {s->m, s->S, s->eps, s->rho, s->nu, s->vcell, s->d, s->mNew, s->phi, s->psi_b, s->psi_max, s->diag,
                      s->part})
        if (p) cudaFree(p);
    if (s->diag_host) cudaFreeHost(s->diag_host);
    s->pipe.destroy();
    delete s;
}

int qmom_correct_impl(pbe_qmom_state *s, const double *eps_h, const double *rho_h, const double *nu_h,
                      const double *phi_h, const double *psi_b_h, int64_t stride_b, double dt, double *d_h,
                      pbe_qmom_diag *diag) {
    pbe_ctx *ctx = s->ctx;
    Pipeline &P = s->pipe;
    const int nm = s->nm;
    const int64_t nc = s->nc;
    const bool want_src = diag && diag->source;
    const bool want_mom = diag && diag->moment;
    const bool want_d = diag && diag->d;
    const bool any_diag = want_src || want_mom || want_d;

    // 1 Mi-cell chunks: 24 MB in, 8 MB out, ~70 us of kernel each -- the bus is the bound here
    const int64_t C = chunk_cells(nc, (int64_t)1 << 17, 12);
    const int nchunks = (int)((nc + C - 1) / C);
    PBE_REQUIRE(nchunks <= kMaxChunks, PBE_E_UNSUPPORTED, "pbe_qmom_correct_resident: %d chunks", nchunks);
    const int parts_all = stats_parts(ctx, nc);
    if (any_diag) {
        int rc = ensure_parts(&s->part, &s->parts_total, 2 * nm + 1, parts_all);
        if (rc) return rc;
    }

    PBE_CUDA(cudaEventRecord(P.ev_start, P.s_comp));
    PBE_CUDA(cudaStreamWaitEvent(P.s_in, P.ev_start, 0));
    PBE_CUDA(cudaStreamWaitEvent(P.s_out, P.ev_start, 0));
    PBE_CUDA(cudaStreamWaitEvent(P.s_side, P.ev_start, 0));
    if (s->mesh && s->nf)
        PBE_CUDA(cudaMemcpyAsync(s->phi, phi_h, (size_t)s->nf * 8, cudaMemcpyHostToDevice, P.s_in));
    if (s->mesh && psi_b_h && s->nbnd) {
        int rc = copy_rows_h2d(s->psi_b, s->nbnd, psi_b_h, stride_b, s->nbnd, nm, P.s_in);
        if (rc) return rc;
    }
    // the scrubbed sources are stored only when somebody reads them (diagnostics, transport)
    double *S_keep = (want_src || s->mesh) ? s->S : nullptr;
    for (int k = 0; k < nchunks; ++k) {
        const int64_t c0 = (int64_t)k * C, cn = std::min(C, nc - c0);
        PBE_CUDA(cudaMemcpyAsync(s->eps + c0, eps_h + c0, (size_t)cn * 8, cudaMemcpyHostToDevice, P.s_in));
        if (rho_h)
            PBE_CUDA(cudaMemcpyAsync(s->rho + c0, rho_h + c0, (size_t)cn * 8, cudaMemcpyHostToDevice, P.s_in));
        if (nu_h)
            PBE_CUDA(cudaMemcpyAsync(s->nu + c0, nu_h + c0, (size_t)cn * 8, cudaMemcpyHostToDevice, P.s_in));
        PBE_CUDA(cudaEventRecord(P.ev_in[k], P.s_in));
        PBE_CUDA(cudaStreamWaitEvent(P.s_comp, P.ev_in[k], 0));
        if (s->mesh) { // momentSourceTerm(k) for every k (QMOM.C:88-96), not yet scrubbed
            int rc = pbe_qmom_source(s->q, cn, nc, s->m + c0, s->eps + c0, s->rho + c0, s->nu + c0, s->S + c0, P.s_comp);
            if (rc) return rc;
            continue;
        }
        // sources, NaN scrub, m += dt S, clamp to SMALL, d (QMOM.C:85-129 with phi = 0), fused
        int rc = pbe_qmom_step(s->q, cn, nc, s->m + c0, s->eps + c0, s->rho + c0, s->nu + c0, dt,
                               S_keep ? S_keep + c0 : nullptr, s->d + c0, P.s_comp);
        if (rc) return rc;
        PBE_CUDA(cudaEventRecord(P.ev_comp[k], P.s_comp));
        if (d_h) {
            PBE_CUDA(cudaStreamWaitEvent(P.s_out, P.ev_comp[k], 0));
            PBE_CUDA(cudaMemcpyAsync(d_h + c0, s->d + c0, (size_t)cn * 8, cudaMemcpyDeviceToHost, P.s_out));
        }
    }
    if (s->mesh) {
        int rc = pbe_field_scrub_nan(ctx, (int64_t)nm * nc, s->S, P.s_comp); // QMOM.C:101-107
        if (rc) return rc;
        if (s->implicit) { // QMOM::solveWithFVM (QMOM.C:201-209)
            s->perf.resize((size_t)nm);
            rc = pbe_transport_implicit(s->mesh, s->scheme, nm, nc, s->m, s->phi, psi_b_h ? s->psi_b : nullptr, s->nbnd,
                                        s->S, dt, &s->fvm, s->mNew, s->perf.data(), P.s_comp);
        } else if (s->scheme == PBE_DIV_UPWIND) {
            rc = pbe_transport_explicit(s->mesh, s->scheme, nm, nc, s->m, s->phi, psi_b_h ? s->psi_b : nullptr, s->nbnd,
                                        s->S, dt, s->mNew, P.s_comp); // QMOM.C:131-197
        } else { // MULES::limit against 1.2 max(m_k) (QMOM.C:171-182), limitSum, explicitSolve
            rc = pbe_qmom_psi_max(ctx, nm, nc, nc, s->m, s->nbnd, s->nbnd, psi_b_h ? s->psi_b : nullptr, s->psi_max,
                                  P.s_comp);
            if (rc) return rc;
            rc = pbe_transport_mules(s->mesh, s->scheme, nm, nc, s->m, s->phi, psi_b_h ? s->psi_b : nullptr, s->nbnd, s->S,
                                     s->psi_max, 0.0, 3, dt, s->mNew, P.s_comp);
        }
        if (rc) return rc;
        std::swap(s->m, s->mNew);
        rc = pbe_qmom_clamp_d(ctx, nm, nc, nc, s->m, s->d, P.s_comp); // QMOM.C:117-125
        if (rc) return rc;
        PBE_CUDA(cudaEventRecord(P.ev_comp[0], P.s_comp));
        if (d_h) {
            PBE_CUDA(cudaStreamWaitEvent(P.s_out, P.ev_comp[0], 0));
            PBE_CUDA(cudaMemcpyAsync(d_h, s->d, (size_t)nc * 8, cudaMemcpyDeviceToHost, P.s_out));
        }
    }
    if (any_diag) { // printAvgMaxMin of the sources, the moments and d (QMOM.C:109, 122, 128)
        const double *V = s->have_vcell ? s->vcell : nullptr;
        int rc = PBE_OK;
        if (want_src) rc = launch_stats_partial(ctx, nm, nc, nc, s->S, V, true, s->part, parts_all, 0, parts_all, P.s_comp);
        if (rc == PBE_OK && want_mom)
            rc = launch_stats_partial(ctx, nm, nc, nc, s->m, V, true, s->part + (size_t)nm * parts_all * 4, parts_all, 0,
                                      parts_all, P.s_comp);
        if (rc == PBE_OK && want_d)
            rc = launch_stats_partial(ctx, 1, nc, nc, s->d, V, true, s->part + (size_t)2 * nm * parts_all * 4, parts_all,
                                      0, parts_all, P.s_comp);
        if (rc) return rc;
        const int nr = 2 * nm + 1;
        rows_stats_final_kernel<<<nr, 32, 0, P.s_comp>>>(parts_all, s->part, s->diag);
        PBE_CUDA(cudaGetLastError());
        ctx->launches++;
        // global printAvgMaxMin (Utility.C:4-10 reduces over the ranks): [sum f V, sum V] | [min] | [max]
        rc = pbe_reduce_diagnostics(ctx, 2 * nr, s->diag, nr, s->diag + 2 * nr, nr, s->diag + 3 * nr, P.s_comp);
        if (rc) return rc;
        PBE_CUDA(cudaMemcpyAsync(s->diag_host, s->diag, (size_t)nr * 4 * sizeof(double), cudaMemcpyDeviceToHost, P.s_comp));
    }
    PBE_CUDA(cudaStreamSynchronize(P.s_in));
    PBE_CUDA(cudaStreamSynchronize(P.s_comp));
    PBE_CUDA(cudaStreamSynchronize(P.s_out));
    if (rho_h) s->have_rho = true;
    if (nu_h) s->have_nu = true;
    {   // device layout [sum f V | sum V | min | max][rows] -> the caller's [row][sum f V, sum V, max, min]
        const int nr = 2 * nm + 1;
        auto row = [&](int r, double *o) {
            o[0] = s->diag_host[r];
            o[1] = s->diag_host[nr + r];
            o[2] = s->diag_host[3 * nr + r];
            o[3] = s->diag_host[2 * nr + r];
        };
        for (int k = 0; k < nm; ++k) {
            if (want_src) row(k, diag->source + 4 * k);
            if (want_mom) row(nm + k, diag->moment + 4 * k);
        }
        if (want_d) row(2 * nm, diag->d);
    }
    return PBE_OK;
}

} // namespace

extern "C" {

int pbe_qmom_state_create(pbe_qmom *q, int64_t ncells, const double *vcell_host, pbe_qmom_state **out) {
    PBE_REQUIRE(q && out, PBE_E_INVALID, "pbe_qmom_state_create: NULL argument");
    *out = nullptr;
    PBE_REQUIRE(ncells >= 0, PBE_E_INVALID, "pbe_qmom_state_create: ncells < 0");
    int order = 0;
    pbe_ctx *ctx = nullptr;
    int rc = pbe_qmom_order(q, &order, &ctx);
    if (rc) return rc;
    DeviceGuard g(ctx->device);
    pbe_qmom_state *s = new pbe_qmom_state();
    s->q = q;
    s->ctx = ctx;
    s->nm = 2 * order;
    s->nc = ncells;
    const size_t cells = (size_t)std::max<int64_t>(ncells, 1);
    cudaError_t e = cudaMalloc((void **)&s->m, (size_t)s->nm * cells * 8);
    if (e == cudaSuccess) e = cudaMalloc((void **)&s->S, (size_t)s->nm * cells * 8);
    for (double **p : {&s->eps, &s->rho, &s->nu, &s->vcell, &s->d})
        if (e == cudaSuccess) e = cudaMalloc((void **)p, cells * 8);
    const size_t dbytes = (size_t)(2 * s->nm + 1) * 4 * sizeof(double);
    if (e == cudaSuccess) e = cudaMalloc((void **)&s->diag, dbytes);
    if (e == cudaSuccess) e = cudaHostAlloc((void **)&s->diag_host, dbytes, 0);
    if (e == cudaSuccess) e = cudaMemset(s->m, 0, (size_t)s->nm * cells * 8);
    if (e == cudaSuccess) e = cudaMemset(s->diag, 0, dbytes);
    if (e == cudaSuccess && vcell_host && ncells) {
        e = cudaMemcpy(s->vcell, vcell_host, (size_t)ncells * 8, cudaMemcpyHostToDevice);
        s->have_vcell = true;
    }
    if (e != cudaSuccess) {
        rc = cuda_fail(e, "pbe_qmom_state_create", __FILE__, __LINE__);
        free_qmom_state(s);
        return rc;
    }
    rc = s->pipe.init(ctx);
    if (rc) {
        free_qmom_state(s);
        return rc;
    }
    *out = s;
    return PBE_OK;
}

int pbe_qmom_state_destroy(pbe_qmom_state *s) {
    free_qmom_state(s);
    return PBE_OK;
}

int pbe_qmom_state_upload(pbe_qmom_state *s, int64_t stride, const double *m_host) {
    PBE_REQUIRE(s && (m_host || s->nc == 0) && stride >= s->nc, PBE_E_INVALID, "pbe_qmom_state_upload: bad argument");
    if (s->nc == 0) return PBE_OK;
    DeviceGuard g(s->ctx->device);
    int rc = copy_rows_h2d(s->m, s->nc, m_host, stride, s->nc, s->nm, s->pipe.s_comp);
    if (rc) return rc;
    PBE_CUDA(cudaStreamSynchronize(s->pipe.s_comp));
    return PBE_OK;
}

int pbe_qmom_state_download(pbe_qmom_state *s, int64_t stride, double *m_host) {
    PBE_REQUIRE(s && (m_host || s->nc == 0) && stride >= s->nc, PBE_E_INVALID,
                "pbe_qmom_state_download: bad argument");
    if (s->nc == 0) return PBE_OK;
    DeviceGuard g(s->ctx->device);
    int rc = copy_rows_d2h(m_host, stride, s->m, s->nc, s->nc, s->nm, s->pipe.s_comp);
    if (rc) return rc;
    PBE_CUDA(cudaStreamSynchronize(s->pipe.s_comp));
    return PBE_OK;
}

int pbe_qmom_state_d(pbe_qmom_state *s, double *d_host) {
    PBE_REQUIRE(s && (d_host || s->nc == 0), PBE_E_INVALID, "pbe_qmom_state_d: bad argument");
    if (s->nc == 0) return PBE_OK;
    PBE_REQUIRE(s->nm >= 2, PBE_E_INVALID, "pbe_qmom_state_d: needs m0 and m1");
    DeviceGuard g(s->ctx->device);
    qmom_d_kernel<<<(unsigned)((s->nc + 255) / 256), 256, 0, s->pipe.s_comp>>>(s->nc, s->m, s->m + s->nc, s->d);
    PBE_CUDA(cudaGetLastError());
    s->ctx->launches++;
    PBE_CUDA(cudaMemcpyAsync(d_host, s->d, (size_t)s->nc * 8, cudaMemcpyDeviceToHost, s->pipe.s_comp));
    PBE_CUDA(cudaStreamSynchronize(s->pipe.s_comp));
    return PBE_OK;
}

int pbe_qmom_state_device(pbe_qmom_state *s, double **m_dev, int64_t *stride, double **d_dev, double **S_dev) {
    PBE_REQUIRE(s != nullptr, PBE_E_INVALID, "pbe_qmom_state_device: NULL state");
    if (m_dev) *m_dev = s->m;
    if (stride) *stride = s->nc;
    if (d_dev) *d_dev = s->d;
    if (S_dev) *S_dev = s->S;
    return PBE_OK;
}

int pbe_qmom_state_attach_mesh(pbe_qmom_state *s, pbe_mesh *mesh, int scheme) {
    PBE_REQUIRE(s && mesh, PBE_E_INVALID, "pbe_qmom_state_attach_mesh: NULL argument");
    int64_t mc = 0, nint = 0, nbnd = 0;
    int rc = pbe_mesh_info(mesh, &mc, &nint, &nbnd, nullptr);
    if (rc) return rc;
    PBE_REQUIRE(mc == s->nc, PBE_E_INVALID, "pbe_qmom_state_attach_mesh: the mesh has %lld cells, the state %lld",
                (long long)mc, (long long)s->nc);
    DeviceGuard g(s->ctx->device);
    for (double **p : {&s->mNew, &s->phi, &s->psi_b, &s->psi_max})
        if (*p) {
            cudaFree(*p);
            *p = nullptr;
        }
    PBE_CUDA(cudaMalloc((void **)&s->psi_max, (size_t)s->nm * 8));
    PBE_CUDA(cudaMalloc((void **)&s->mNew, (size_t)s->nm * std::max<int64_t>(s->nc, 1) * 8));
    PBE_CUDA(cudaMalloc((void **)&s->phi, (size_t)std::max<int64_t>(nint + nbnd, 1) * 8));
    PBE_CUDA(cudaMalloc((void **)&s->psi_b, (size_t)s->nm * std::max<int64_t>(nbnd, 1) * 8));
    s->mesh = mesh;
    s->scheme = scheme;
    s->nf = nint + nbnd;
    s->nbnd = nbnd;
    return PBE_OK;
}

int pbe_qmom_state_set_implicit(pbe_qmom_state *s, const pbe_fvm_controls *controls) {
    PBE_REQUIRE(s != nullptr, PBE_E_INVALID, "pbe_qmom_state_set_implicit: NULL state");
    PBE_REQUIRE(controls == nullptr || s->mesh != nullptr, PBE_E_INVALID,
                "pbe_qmom_state_set_implicit: attach a mesh first (pbe_qmom_state_attach_mesh)");
    s->implicit = controls != nullptr;
    if (controls) s->fvm = *controls;
    s->perf.clear();
    return PBE_OK;
}

int pbe_qmom_state_solver_performance(pbe_qmom_state *s, pbe_fvm_perf *perf) {
    PBE_REQUIRE(s && perf, PBE_E_INVALID, "pbe_qmom_state_solver_performance: NULL argument");
    PBE_REQUIRE((int)s->perf.size() == s->nm, PBE_E_INVALID,
                "pbe_qmom_state_solver_performance: no implicit step has run on this state");
    std::copy(s->perf.begin(), s->perf.end(), perf);
    return PBE_OK;
}

int pbe_qmom_correct_resident(pbe_qmom_state *s, const double *eps_host, const double *rho_host,
                              const double *nu_host, const double *phi_host, const double *psi_b_host,
                              int64_t stride_b, double dt, double *d_host, pbe_qmom_diag *diag) {
    PBE_REQUIRE(s != nullptr, PBE_E_INVALID, "pbe_qmom_correct_resident: NULL state");
    PBE_REQUIRE(dt > 0.0, PBE_E_INVALID, "pbe_qmom_correct_resident: dt must be positive (got %g)", dt);
    if (s->nc == 0) return PBE_OK;
    PBE_REQUIRE(eps_host != nullptr, PBE_E_INVALID, "pbe_qmom_correct_resident: eps is NULL");
    PBE_REQUIRE((rho_host || s->have_rho) && (nu_host || s->have_nu), PBE_E_INVALID,
                "pbe_qmom_correct_resident: rho / nu may only be NULL after a call that supplied them");
    PBE_REQUIRE(s->mesh == nullptr || s->nf == 0 || phi_host != nullptr, PBE_E_INVALID,
                "pbe_qmom_correct_resident: a mesh is attached, phi is NULL");
    PBE_REQUIRE(s->mesh != nullptr || (phi_host == nullptr && psi_b_host == nullptr), PBE_E_INVALID,
                "pbe_qmom_correct_resident: phi / psi_b without an attached mesh (pbe_qmom_state_attach_mesh)");
    PBE_REQUIRE(psi_b_host == nullptr || stride_b >= s->nbnd, PBE_E_INVALID,
                "pbe_qmom_correct_resident: stride_b < boundary faces");
    DeviceGuard g(s->ctx->device);
    const int rc = qmom_correct_impl(s, eps_host, rho_host, nu_host, phi_host, psi_b_host, stride_b, dt, d_host, diag);
    if (rc != PBE_OK) s->pipe.drain();
    return rc;
}

int pbe_host_register(pbe_ctx *ctx, void *p, int64_t bytes) {
    PBE_REQUIRE(ctx && p && bytes > 0, PBE_E_INVALID, "pbe_host_register: bad argument");
    DeviceGuard g(ctx->device);
    PBE_CUDA(cudaHostRegister(p, (size_t)bytes, cudaHostRegisterDefault));
    return PBE_OK;
}
int pbe_host_unregister(pbe_ctx *ctx, void *p) {
    PBE_REQUIRE(ctx && p, PBE_E_INVALID, "pbe_host_unregister: bad argument");
    DeviceGuard g(ctx->device);
    PBE_CUDA(cudaHostUnregister(p));
    return PBE_OK;
}

} // extern "C"
