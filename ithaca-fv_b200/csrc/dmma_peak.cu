// Register-resident DMMA.8x8x4 issue-rate microbenchmark: the measured FP64 tensor-core peak that the
// Gram kernel's roofline fraction is quoted against (no FP64 figure exists in MEASURED_PEAKS.json).
#include "../../include/ithaca_b200.h"
#include "common.cuh"

namespace ib200 {

template <int NACC>
__global__ void __launch_bounds__(512, 1) dmma_peak_kernel(double* out, int iters, double seed)
{
    double c[NACC][2];
#pragma unroll
    for (int i = 0; i < NACC; i++) c[i][0] = c[i][1] = 0.0;
    double a = seed * (threadIdx.x + 1), b = seed * (threadIdx.x + 2);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < NACC; i++) dmma884(c[i][0], c[i][1], a, b);
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < NACC; i++) s += c[i][0] + c[i][1];
    if (s == 12345.678) out[0] = s;  // keep the accumulators live
}

}  // namespace ib200

extern "C" int ithaca_b200_dmma_peak(int device, double* tflops_out)
{
    if (!tflops_out) return IB200_ERR_ARG;
    if (cudaSetDevice(device) != cudaSuccess) return IB200_ERR_CUDA;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return IB200_ERR_CUDA;
    double* d = nullptr;
    if (cudaMalloc(&d, 64) != cudaSuccess) return IB200_ERR_CUDA;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int iters = 4096, nacc = 16, blocks = prop.multiProcessorCount, threads = 512;
    double best = 0.0;
    for (int rep = 0; rep < 6; rep++) {
        cudaEventRecord(e0);
        ib200::dmma_peak_kernel<nacc><<<blocks, threads>>>(d, iters, 1e-3);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(d); return IB200_ERR_CUDA; }
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        const double flops = 512.0 * nacc * (double)iters * (threads / 32) * blocks;
        if (rep > 0) best = std::max(best, flops / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d);
    *tflops_out = best;
    return IB200_OK;
}
