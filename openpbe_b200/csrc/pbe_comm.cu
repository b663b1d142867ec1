// pbe_comm.cu -- multi-GPU plumbing of the C ABI: one process per GPU, an NCCL communicator over the
// ranks, the tiny global reductions the reference gets from OpenFOAM's implicit MPI all-reduces
// (min(n_k) MOC.C:181-183, max(alpha V / xi_k) :238, max(continuityError) :276-279,
// max(m_k) QMOM.C:171, printAvgMaxMin Utility.C:4-10) and the halo exchange of the transport step.
//
// Cells are independent on the source path, so there is NO data-path collective: every message here
// is <= 3 N doubles (latency-bound) except the halo rows of the transport update.
//
// NCCL is bound at run time (dlopen of libnccl.so.2): the single-GPU library has no link-time
// dependency on it, and a host that already carries an NCCL (PyTorch's, an MPI's) shares that copy.
#include <dlfcn.h>
#include <string.h>

#include <algorithm>

#include "pbe_common.h"

using namespace pbe;

namespace {

// the slice of nccl.h this file needs (ABI-stable since NCCL 2.0; constants from nccl.h 2.27)
typedef struct ncclComm *ncclComm_t;
typedef struct {
    char internal[128];
} ncclUniqueId;
enum { kNcclSuccess = 0, kNcclSum = 0, kNcclMax = 2, kNcclMin = 3, kNcclFloat64 = 8 };

struct Nccl {
    void *so = nullptr;
    int (*GetUniqueId)(ncclUniqueId *) = nullptr;
    int (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*CommCount)(const ncclComm_t, int *) = nullptr;
    int (*CommUserRank)(const ncclComm_t, int *) = nullptr;
    int (*AllReduce)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Send)(const void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Recv)(void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    int (*GetVersion)(int *) = nullptr;
};

Nccl *nccl() {
    static Nccl n;
    static bool tried = false;
    if (tried) return n.so ? &n : nullptr;
    tried = true;
    const char *names[] = {getenv("PBE_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    for (const char *nm : names) {
        if (!nm) continue;
        n.so = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (n.so) break;
    }
    if (!n.so) return nullptr;
#define PBE_NCCL_SYM(f) *(void **)(&n.f) = dlsym(n.so, "nccl" #f)
    PBE_NCCL_SYM(GetUniqueId);
    PBE_NCCL_SYM(CommInitRank);
    PBE_NCCL_SYM(CommDestroy);
    PBE_NCCL_SYM(CommCount);
    PBE_NCCL_SYM(CommUserRank);
    PBE_NCCL_SYM(AllReduce);
    PBE_NCCL_SYM(Send);
    PBE_NCCL_SYM(Recv);
    PBE_NCCL_SYM(GroupStart);
    PBE_NCCL_SYM(GroupEnd);
    PBE_NCCL_SYM(GetErrorString);
    PBE_NCCL_SYM(GetVersion);
#undef PBE_NCCL_SYM
    if (!n.GetUniqueId || !n.CommInitRank || !n.AllReduce || !n.GroupStart || !n.GroupEnd || !n.Send || !n.Recv) {
        dlclose(n.so);
        n.so = nullptr;
        return nullptr;
    }
    return &n;
}

int nccl_fail(int rc, const char *what) {
    Nccl *n = nccl();
    set_error("NCCL error %d (%s) in %s", rc, (n && n->GetErrorString) ? n->GetErrorString(rc) : "?", what);
    return PBE_E_CUDA;
}
#define PBE_NCCL(call)                                       \
    do {                                                     \
        int r_ = (call);                                     \
        if (r_ != kNcclSuccess) return nccl_fail(r_, #call); \
    } while (0)

} // namespace

struct pbe_comm {
    ncclComm_t comm = nullptr;
    int nranks = 1, rank = 0;
    bool owned = false;
    double *scratch = nullptr; // device staging of the host-array reductions
    size_t scratch_doubles = 0;
    cudaStream_t stream = nullptr;
};

extern "C" {

int pbe_comm_unique_id(void *id128) {
    PBE_REQUIRE(id128 != nullptr, PBE_E_INVALID, "pbe_comm_unique_id: NULL");
    Nccl *n = nccl();
    PBE_REQUIRE(n != nullptr, PBE_E_UNSUPPORTED, "pbe_comm_unique_id: libnccl.so.2 not found (set PBE_NCCL_LIB)");
    ncclUniqueId id;
    PBE_NCCL(n->GetUniqueId(&id));
    memcpy(id128, &id, sizeof(id));
    return PBE_OK;
}

static int comm_attach(pbe_ctx *ctx, ncclComm_t c, bool owned) {
    pbe_comm *pc = new pbe_comm();
    pc->comm = c;
    pc->owned = owned;
    Nccl *n = nccl();
    if (n->CommCount) n->CommCount(c, &pc->nranks);
    if (n->CommUserRank) n->CommUserRank(c, &pc->rank);
    cudaError_t e = cudaStreamCreateWithFlags(&pc->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
        delete pc;
        return cuda_fail(e, "pbe_comm: side stream", __FILE__, __LINE__);
    }
    ctx->comm = pc;
    return PBE_OK;
}

int pbe_comm_init(pbe_ctx *ctx, int nranks, int rank, const void *id128) {
    PBE_REQUIRE(ctx && id128 && nranks >= 1 && rank >= 0 && rank < nranks, PBE_E_INVALID, "pbe_comm_init: bad argument");
    PBE_REQUIRE(ctx->comm == nullptr, PBE_E_INVALID, "pbe_comm_init: the context already has a communicator");
    Nccl *n = nccl();
    PBE_REQUIRE(n != nullptr, PBE_E_UNSUPPORTED, "pbe_comm_init: libnccl.so.2 not found (set PBE_NCCL_LIB)");
    DeviceGuard g(ctx->device);
    ncclUniqueId id;
    memcpy(&id, id128, sizeof(id));
    ncclComm_t c = nullptr;
    PBE_NCCL(n->CommInitRank(&c, nranks, id, rank));
    return comm_attach(ctx, c, true);
}

int pbe_comm_init_from(pbe_ctx *ctx, void *nccl_comm) {
    PBE_REQUIRE(ctx && nccl_comm, PBE_E_INVALID, "pbe_comm_init_from: NULL argument");
    PBE_REQUIRE(ctx->comm == nullptr, PBE_E_INVALID, "pbe_comm_init_from: the context already has a communicator");
    PBE_REQUIRE(nccl() != nullptr, PBE_E_UNSUPPORTED, "pbe_comm_init_from: libnccl.so.2 not found (set PBE_NCCL_LIB)");
    DeviceGuard g(ctx->device);
    return comm_attach(ctx, (ncclComm_t)nccl_comm, false);
}

int pbe_comm_destroy(pbe_ctx *ctx) {
    PBE_REQUIRE(ctx != nullptr, PBE_E_INVALID, "pbe_comm_destroy: NULL context");
    pbe_comm *pc = ctx->comm;
    if (!pc) return PBE_OK;
    DeviceGuard g(ctx->device);
    cudaStreamSynchronize(pc->stream);
    if (pc->owned && nccl() && nccl()->CommDestroy) nccl()->CommDestroy(pc->comm);
    if (pc->scratch) cudaFree(pc->scratch);
    cudaStreamDestroy(pc->stream);
    delete pc;
    ctx->comm = nullptr;
    return PBE_OK;
}

int pbe_comm_info(pbe_ctx *ctx, int *nranks, int *rank) {
    PBE_REQUIRE(ctx != nullptr, PBE_E_INVALID, "pbe_comm_info: NULL context");
    if (nranks) *nranks = ctx->comm ? ctx->comm->nranks : 1;
    if (rank) *rank = ctx->comm ? ctx->comm->rank : 0;
    return PBE_OK;
}

// In-place all-reduces of three DEVICE vectors (sum, min, max; any may be empty) as ONE NCCL group on
// `cuda_stream`: what MOC::correct() / QMOM::correct() print every step.  Without a communicator
// (single rank) it is a no-op.
int pbe_reduce_diagnostics(pbe_ctx *ctx, int nsum, double *sums, int nmin, double *mins, int nmax, double *maxs,
                           void *cuda_stream) {
    PBE_REQUIRE(ctx != nullptr && nsum >= 0 && nmin >= 0 && nmax >= 0, PBE_E_INVALID, "pbe_reduce_diagnostics: bad argument");
    pbe_comm *pc = ctx->comm;
    if (!pc || pc->nranks == 1) return PBE_OK;
    Nccl *n = nccl();
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    PBE_NCCL(n->GroupStart());
    if (nsum) PBE_NCCL(n->AllReduce(sums, sums, (size_t)nsum, kNcclFloat64, kNcclSum, pc->comm, st));
    if (nmin) PBE_NCCL(n->AllReduce(mins, mins, (size_t)nmin, kNcclFloat64, kNcclMin, pc->comm, st));
    if (nmax) PBE_NCCL(n->AllReduce(maxs, maxs, (size_t)nmax, kNcclFloat64, kNcclMax, pc->comm, st));
    PBE_NCCL(n->GroupEnd());
    return PBE_OK;
}

// The same for HOST arrays (what a C++ solver holds after correct()): staged through a device scratch
// buffer on the communicator's side stream; synchronous.
int pbe_reduce_diagnostics_host(pbe_ctx *ctx, int nsum, double *sums, int nmin, double *mins, int nmax, double *maxs) {
    PBE_REQUIRE(ctx != nullptr && nsum >= 0 && nmin >= 0 && nmax >= 0, PBE_E_INVALID,
                "pbe_reduce_diagnostics_host: bad argument");
    pbe_comm *pc = ctx->comm;
    if (!pc || pc->nranks == 1) return PBE_OK;
    DeviceGuard g(ctx->device);
    const size_t tot = (size_t)nsum + nmin + nmax;
    if (tot == 0) return PBE_OK;
    if (pc->scratch_doubles < tot) {
        if (pc->scratch) cudaFree(pc->scratch);
        pc->scratch = nullptr;
        pc->scratch_doubles = 0;
        PBE_CUDA(cudaMalloc((void **)&pc->scratch, std::max<size_t>(tot, 1024) * sizeof(double)));
        pc->scratch_doubles = std::max<size_t>(tot, 1024);
    }
    double *ds = pc->scratch, *dm = ds + nsum, *dx = dm + nmin;
    if (nsum) PBE_CUDA(cudaMemcpyAsync(ds, sums, (size_t)nsum * 8, cudaMemcpyHostToDevice, pc->stream));
    if (nmin) PBE_CUDA(cudaMemcpyAsync(dm, mins, (size_t)nmin * 8, cudaMemcpyHostToDevice, pc->stream));
    if (nmax) PBE_CUDA(cudaMemcpyAsync(dx, maxs, (size_t)nmax * 8, cudaMemcpyHostToDevice, pc->stream));
    int rc = pbe_reduce_diagnostics(ctx, nsum, ds, nmin, dm, nmax, dx, pc->stream);
    if (rc) return rc;
    if (nsum) PBE_CUDA(cudaMemcpyAsync(sums, ds, (size_t)nsum * 8, cudaMemcpyDeviceToHost, pc->stream));
    if (nmin) PBE_CUDA(cudaMemcpyAsync(mins, dm, (size_t)nmin * 8, cudaMemcpyDeviceToHost, pc->stream));
    if (nmax) PBE_CUDA(cudaMemcpyAsync(maxs, dx, (size_t)nmax * 8, cudaMemcpyDeviceToHost, pc->stream));
    PBE_CUDA(cudaStreamSynchronize(pc->stream));
    return PBE_OK;
}

// Halo exchange of the transport step on a decomposed mesh (processor patches): for each neighbouring
// rank one send and one receive per field, all in ONE NCCL group on `cuda_stream`.  send [nfields]
// [stride_s] holds the values of this rank's cells along its coupled faces (pbe_gather_cells), recv
// [nfields][stride_r] receives the values across them (psi_halo of pbe_transport_explicit_halo);
// neighbour q exchanges the face range [off[q], off[q] + cnt[q]) on both sides (the two ranks list
// their shared faces in the same order: ascending global face index, decomposePar's convention).
int pbe_halo_exchange(pbe_ctx *ctx, int nneigh, const int *peer, const int64_t *off, const int64_t *cnt, int nfields,
                      const double *send, int64_t stride_s, double *recv, int64_t stride_r, void *cuda_stream) {
    PBE_REQUIRE(ctx != nullptr && nneigh >= 0 && nfields >= 0, PBE_E_INVALID, "pbe_halo_exchange: bad argument");
    if (nneigh == 0 || nfields == 0) return PBE_OK;
    PBE_REQUIRE(peer && off && cnt && send && recv, PBE_E_INVALID, "pbe_halo_exchange: NULL argument");
    pbe_comm *pc = ctx->comm;
    PBE_REQUIRE(pc != nullptr, PBE_E_INVALID, "pbe_halo_exchange: the context has no communicator (pbe_comm_init)");
    Nccl *n = nccl();
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    PBE_NCCL(n->GroupStart());
    for (int q = 0; q < nneigh; ++q)
        for (int k = 0; k < nfields; ++k) {
            PBE_NCCL(n->Send(send + (size_t)k * stride_s + off[q], (size_t)cnt[q], kNcclFloat64, peer[q], pc->comm, st));
            PBE_NCCL(n->Recv(recv + (size_t)k * stride_r + off[q], (size_t)cnt[q], kNcclFloat64, peer[q], pc->comm, st));
        }
    PBE_NCCL(n->GroupEnd());
    return PBE_OK;
}

} // extern "C"
