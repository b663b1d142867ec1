// Microbenchmark: does a non-FP64 instruction issued next to a DFMA cost FP64 throughput on sm_100a?
// K independent DFMA chains per thread, with J integer (LOP3/IADD) or FP32 (FFMA) instructions
// interleaved per DFMA.  Prints cycles per DFMA warp-instruction per SM sub-partition.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scratch/ubench tools/ubench_fp64_issue.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int J, int KIND>
__global__ void __launch_bounds__(1024) k(double *out, int *iout, float *fout, int iters, double a, int ia, float fa) {
    double x[8];
    int y[4][8];
    float f[4][8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        x[i] = threadIdx.x * 1e-3 + i;
        for (int j = 0; j < 4; ++j) {
            y[j][i] = threadIdx.x + i + j;
            f[j][i] = threadIdx.x * 1e-3f + i + j;
        }
    }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            x[i] = __fma_rn(x[i], a, 1e-9);
#pragma unroll
            for (int j = 0; j < J; ++j) {
                if (KIND == 0) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(y[j][i]) : "r"(ia), "r"(it)); // one LOP3 (3-input xor with the loop counter: not foldable), 8 J independent chains
                if (KIND == 2) asm volatile("shf.l.wrap.b32 %0, %0, %1, %2;" : "+r"(y[j][i]) : "r"(ia), "r"(it)); // one SHF
                if (KIND == 3) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(y[j][i]) : "r"(ia), "r"(it)); // one IMAD
                if (KIND == 1) f[j][i] = __fmaf_rn(f[j][i], fa, 1e-9f);    // one FFMA
            }
        }
    }
    double s = 0;
    int t = 0;
    float g = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        s += x[i];
        for (int j = 0; j < 4; ++j) {
            t += y[j][i];
            g += f[j][i];
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    iout[blockIdx.x * blockDim.x + threadIdx.x] = t;
    fout[blockIdx.x * blockDim.x + threadIdx.x] = g;
}

template <int J, int KIND>
void run(const char *name, int warps_per_smsp) {
    int dev = 0, sms = 0, khz = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
    const int threads = 32 * 4 * warps_per_smsp;
    double *o;
    int *io;
    float *fo;
    cudaMalloc(&o, sms * threads * 8);
    cudaMalloc(&io, sms * threads * 4);
    cudaMalloc(&fo, sms * threads * 4);
    const int iters = 20000;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    k<J, KIND><<<sms, threads>>>(o, io, fo, 100, 1.0000001, 5, 1.0001f);
    if (cudaGetLastError() != cudaSuccess) std::printf("launch failed\n");
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k<J, KIND><<<sms, threads>>>(o, io, fo, iters, 1.0000001, 5, 1.0001f);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double dfma_per_smsp = (double)iters * 8 * warps_per_smsp;
    const double cycles = ms * 1e-3 * khz * 1e3;
    std::printf("%-28s warps/SMSP %d: %.3f ms, %.2f cycles per DFMA warp-instr per SMSP (nominal clock %d MHz)\n", name,
                warps_per_smsp, ms, cycles / dfma_per_smsp, khz / 1000);
    cudaFree(o);
    cudaFree(io);
    cudaFree(fo);
}

int main() {
    for (int w : {2, 8}) {
        run<0, 0>("DFMA only", w);
        run<1, 0>("DFMA + 1 int", w);
        run<2, 0>("DFMA + 2 int", w);
        run<4, 0>("DFMA + 4 int", w);
        run<1, 1>("DFMA + 1 FFMA", w);
        run<2, 1>("DFMA + 2 FFMA", w);
        run<1, 2>("DFMA + 1 SHF", w);
        run<2, 2>("DFMA + 2 SHF", w);
        run<1, 3>("DFMA + 1 IMAD", w);
        run<2, 3>("DFMA + 2 IMAD", w);
    }
    return 0;
}
