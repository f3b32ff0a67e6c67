// Experiment: do DMMA (mma.sync.m8n8k4.f64) and DFMA (fma.rn.f64) share an execution pipe on B200?
// Times register-resident loops: DMMA only, DFMA only, half the warps each, and both interleaved in every warp.
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// mode 0: all warps DMMA; 1: all warps DFMA; 2: even warps DMMA, odd warps DFMA; 3: every warp interleaves both
template <int MODE>
__global__ void __launch_bounds__(512, 1) k(double* out, int iters, double seed)
{
    constexpr int NA = 16;
    double c[NA][2], f[NA * 2];
#pragma unroll
    for (int i = 0; i < NA; i++) { c[i][0] = c[i][1] = 0.0; f[2 * i] = f[2 * i + 1] = 0.0; }
    const double a = seed * (threadIdx.x + 1), b = seed * (threadIdx.x + 2);
    const int warp = threadIdx.x >> 5;
    const bool do_mma = MODE == 0 || MODE == 3 || (MODE == 2 && (warp & 1) == 0);
    const bool do_fma = MODE == 1 || MODE == 3 || (MODE == 2 && (warp & 1) == 1);
    for (int it = 0; it < iters; it++) {
        if (do_mma) {
#pragma unroll
            for (int i = 0; i < NA; i++) dmma884(c[i][0], c[i][1], a, b);
        }
        if (do_fma) {
#pragma unroll
            for (int i = 0; i < 2 * NA; i++) asm volatile("fma.rn.f64 %0, %1, %2, %0;" : "+d"(f[i]) : "d"(a), "d"(b));
        }
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < NA; i++) s += c[i][0] + c[i][1] + f[2 * i] + f[2 * i + 1];
    if (s == 12345.678) out[0] = s;
}

template <int MODE>
static void run(const char* name, int sms)
{
    double* d;
    cudaMalloc(&d, 64);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 4096, threads = 512, warps = threads / 32;
    float best = 1e30f;
    for (int rep = 0; rep < 5; rep++) {
        cudaEventRecord(e0);
        k<MODE><<<sms, threads>>>(d, iters, 1e-3);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (rep) best = ms < best ? ms : best;
    }
    const double mma_warps = MODE == 0 || MODE == 3 ? warps : MODE == 2 ? warps / 2 : 0;
    const double fma_warps = MODE == 1 || MODE == 3 ? warps : MODE == 2 ? warps / 2 : 0;
    const double mma_fl = 512.0 * 16 * iters * mma_warps * sms;          // 8x8x4 x 2 flop per warp-level DMMA
    const double fma_fl = 2.0 * 32 * 32 * iters * fma_warps * sms;       // 32 DFMA per thread per iteration
    printf("%-28s %8.3f ms  DMMA %6.2f TF/s  DFMA %6.2f TF/s  total %6.2f TF/s\n", name, best, mma_fl / best / 1e9, fma_fl / best / 1e9,
           (mma_fl + fma_fl) / best / 1e9);
    cudaFree(d);
}

int main()
{
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    printf("%s, %d SMs\n", p.name, p.multiProcessorCount);
    run<0>("DMMA only (16 warps)", p.multiProcessorCount);
    run<1>("DFMA only (16 warps)", p.multiProcessorCount);
    run<2>("8 warps DMMA + 8 warps DFMA", p.multiProcessorCount);
    run<3>("interleaved in every warp", p.multiProcessorCount);
    return 0;
}
