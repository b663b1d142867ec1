// pbe_common.h -- internals shared by the translation units of libpbe_b200.so.
#pragma once

#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>

#include <atomic>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/pbe_b200.h"

namespace pbe {

void set_error(const char *fmt, ...);
int cuda_fail(cudaError_t e, const char *what, const char *file, int line);

#define PBE_CUDA(call)                                                                  \
    do {                                                                                \
        cudaError_t e_ = (call);                                                        \
        if (e_ != cudaSuccess) return ::pbe::cuda_fail(e_, #call, __FILE__, __LINE__);  \
    } while (0)

#define PBE_REQUIRE(cond, code, ...)      \
    do {                                  \
        if (!(cond)) {                    \
            ::pbe::set_error(__VA_ARGS__); \
            return (code);                \
        }                                 \
    } while (0)

constexpr int kHostSlots = 3; // H2D(k+1) | kernel(k) | D2H(k-1)

} // namespace pbe

struct pbe_comm; // NCCL communicator + side stream (pbe_comm.cu); null on a single-rank context

// The opaque handles of include/pbe_b200.h
struct pbe_ctx {
    int device = 0;
    int sm_count = 0;
    size_t smem_optin = 0;
    size_t smem_per_sm = 0;
    size_t hbm_bytes = 0;
    char name[128] = {0};
    std::atomic<int64_t> launches{0};
    cudaStream_t slot_stream[pbe::kHostSlots] = {nullptr, nullptr, nullptr};
    // staging for the host-buffer entry points (grown on demand, freed with the context)
    void *stage[pbe::kHostSlots] = {nullptr, nullptr, nullptr};
    size_t stage_bytes = 0;
    // cudaFuncAttributeMaxDynamicSharedMemorySize belongs to the kernel instantiation (per device),
    // not to a handle: the largest request any handle of this context has made so far
    std::mutex smem_mu;
    std::map<const void *, size_t> smem_limit;
    pbe_comm *comm = nullptr; // pbe_comm_init / pbe_comm_init_from
};

namespace pbe {
int ensure_stage(pbe_ctx *ctx, size_t bytes_per_slot);
// Raise (never lower) the dynamic shared-memory limit of a kernel before a launch that needs
// `bytes`: handles with different table sizes that share one instantiation stay independent.
cudaError_t raise_dyn_smem(pbe_ctx *ctx, const void *fn, size_t bytes);
// Wait for every host-pipeline stream of the context (error paths of the host-buffer entries:
// asynchronous copies must not outlive the call that was handed the caller's buffers).
void drain_slot_streams(pbe_ctx *ctx);
struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) {
        cudaGetDevice(&prev);
        if (prev != dev) cudaSetDevice(dev);
        else prev = -1;
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};
} // namespace pbe
