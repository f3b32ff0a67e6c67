// Measurement helper (like dmma_peak.cu; not on the product path): the synthetic snapshot set of SURVEY.md 8(d), generated in
// place on the device from a counter-based generator so that any row range of any snapshot can be (re)produced identically by
// any rank, in any chunking, without ever holding the 480 GB / 800 GB sets of configs C3 / C5 anywhere:
//
//   S(i, j) = sum_{k < R} sigma_k u_k(x_i) coef(j, k)  +  eps * w(i, j)
//     x_i     = (i + 1/2) / NdGlobal                         (i = GLOBAL row = cell * dim + component)
//     u_k(x)  = sin((k + 1) pi x)                            (Chebyshev recurrence in k)
//     sigma_k = 1 / (k + 1)                                  (eigenvalues ~ 1/(k+1)^2: relative gaps ~ 2/k, far above 1e-8)
//     coef    = m x R matrix with orthonormal columns (the caller's, e.g. a QR factor)
//     w(i, j) = sqrt(3) (U1 + U2 + U3 + U4 - 2), U* = the four 32-bit words of Philox4x32-10(key = seed, counter = (i, j))
//               scaled to [0, 1): zero mean, unit variance (Irwin-Hall), independent across (i, j)
//   V(c)    = exp(0.5 n(c)) e^(-1/8) / nCellsGlobal, n(c) = the same unit-variance variate with counter (c, 2^32 - 1): sums to ~1
//
// The kernel is an FP64 FMA-bound rank-R update (rows x m x R FMAs); it runs outside every timed region.
#include "../../include/ithaca_b200.h"
#include "common.cuh"
#include <algorithm>

namespace ib200 {

__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1,
                                              uint32_t (&out)[4])
{
#pragma unroll
    for (int r = 0; r < 10; r++) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        c0 = hi1 ^ c1 ^ k0; c1 = lo1; c2 = hi0 ^ c3 ^ k1; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

__device__ __forceinline__ double unit_variate(unsigned long long row, uint32_t col, unsigned long long seed)
{
    uint32_t w[4];
    philox4x32_10((uint32_t)row, (uint32_t)(row >> 32), col, 0u, (uint32_t)seed, (uint32_t)(seed >> 32), w);
    const double s = ((double)w[0] + (double)w[1] + (double)w[2] + (double)w[3]) * (1.0 / 4294967296.0);
    return 1.7320508075688772 * (s - 2.0);
}

constexpr int SY_ROWS = 128;   // rows per CTA tile (32 lanes x 4 rows)
constexpr int SY_COLS = 32;    // snapshots per inner tile (8 warps x 4)

// grid = row tiles; every CTA builds its U tile once and then walks all snapshot tiles.
__global__ void __launch_bounds__(256)
synth_snapshots_kernel(double* __restrict__ S, long long ld, long long row0, long long nRows, long long NdGlobal, int m,
                       const double* __restrict__ coef, int R, double eps, unsigned long long seed)
{
    extern __shared__ double sm[];
    double* U = sm;                          // [R][SY_ROWS]  sigma_k u_k(x_i)
    double* C = sm + (size_t)R * SY_ROWS;    // [R][SY_COLS]
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const long long tile0 = (long long)blockIdx.x * SY_ROWS;
    if (tid < SY_ROWS) {
        const long long gi = row0 + tile0 + tid;
        const double th = 3.141592653589793 * ((double)gi + 0.5) / (double)NdGlobal;
        double s1, c1;
        sincos(th, &s1, &c1);
        double sp = 0.0, sc = s1;            // sin(0 th), sin(1 th)
        const double two_c = 2.0 * c1;
        for (int k = 0; k < R; k++) {
            U[(size_t)k * SY_ROWS + tid] = sc / (double)(k + 1);
            const double sn = two_c * sc - sp;
            sp = sc; sc = sn;
        }
    }
    for (int j0 = 0; j0 < m; j0 += SY_COLS) {
        __syncthreads();
        for (int idx = tid; idx < R * SY_COLS; idx += 256) {
            const int k = idx / SY_COLS, jj = idx % SY_COLS;
            C[idx] = (j0 + jj < m) ? coef[(size_t)(j0 + jj) + (size_t)k * m] : 0.0;
        }
        __syncthreads();
        double acc[4][4];
#pragma unroll
        for (int a = 0; a < 4; a++)
#pragma unroll
            for (int b = 0; b < 4; b++) acc[a][b] = 0.0;
        for (int k = 0; k < R; k++) {
            const double2 u01 = *reinterpret_cast<const double2*>(U + (size_t)k * SY_ROWS + 4 * lane);
            const double2 u23 = *reinterpret_cast<const double2*>(U + (size_t)k * SY_ROWS + 4 * lane + 2);
            const double2 c01 = *reinterpret_cast<const double2*>(C + (size_t)k * SY_COLS + 4 * warp);
            const double2 c23 = *reinterpret_cast<const double2*>(C + (size_t)k * SY_COLS + 4 * warp + 2);
            const double u[4] = {u01.x, u01.y, u23.x, u23.y}, c[4] = {c01.x, c01.y, c23.x, c23.y};
#pragma unroll
            for (int a = 0; a < 4; a++)
#pragma unroll
                for (int b = 0; b < 4; b++) acc[a][b] = fma(u[a], c[b], acc[a][b]);
        }
#pragma unroll
        for (int b = 0; b < 4; b++) {
            const int j = j0 + 4 * warp + b;
            if (j >= m) continue;
            double* col = S + (size_t)j * ld + tile0 + 4 * lane;
#pragma unroll
            for (int a = 0; a < 4; a++) {
                const long long li = tile0 + 4 * lane + a;
                if (li < nRows) col[a] = acc[a][b] + eps * unit_variate((unsigned long long)(row0 + li), (uint32_t)j, seed);
            }
        }
    }
}

__global__ void synth_volumes_kernel(double* __restrict__ V, long long cell0, long long nCells, long long nCellsGlobal,
                                     unsigned long long seed)
{
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nCells; i += (long long)gridDim.x * blockDim.x)
        V[i] = exp(0.5 * unit_variate((unsigned long long)(cell0 + i), 0xFFFFFFFFu, seed)) * 0.8824969025845955 / (double)nCellsGlobal;
}

}  // namespace ib200

using namespace ib200;

extern "C" {

int ithaca_b200_synth_snapshots(ithaca_b200_t* h, double* S, int64_t ld, int64_t row0, int64_t nRows, int64_t NdGlobal, int m,
                                const double* coef, int R, double eps, uint64_t seed)
{
    if (!h) return IB200_ERR_ARG;
    if (!S || !coef || ld < nRows || nRows <= 0 || row0 < 0 || row0 + nRows > NdGlobal || m <= 0 || R <= 0 || R > 256) {
        set_error(h, "synth_snapshots: bad argument");
        return IB200_ERR_ARG;
    }
    cudaSetDevice(h->device);
    const size_t smem = (size_t)R * (SY_ROWS + SY_COLS) * sizeof(double);
    IB_CUDA(h, cudaFuncSetAttribute(synth_snapshots_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long tiles = (nRows + SY_ROWS - 1) / SY_ROWS;
    synth_snapshots_kernel<<<(unsigned)tiles, 256, smem, h->stream>>>(S, ld, row0, nRows, NdGlobal, m, coef, R, eps, seed);
    IB_CUDA(h, cudaGetLastError());
    return IB200_OK;
}

int ithaca_b200_synth_volumes(ithaca_b200_t* h, double* V, int64_t cell0, int64_t nCells, int64_t nCellsGlobal, uint64_t seed)
{
    if (!h) return IB200_ERR_ARG;
    if (!V || nCells <= 0 || cell0 < 0 || cell0 + nCells > nCellsGlobal) { set_error(h, "synth_volumes: bad argument"); return IB200_ERR_ARG; }
    cudaSetDevice(h->device);
    synth_volumes_kernel<<<(unsigned)std::min<long long>((nCells + 255) / 256, 8LL * h->num_sms), 256, 0, h->stream>>>(
        V, cell0, nCells, nCellsGlobal, seed);
    IB_CUDA(h, cudaGetLastError());
    return IB200_OK;
}

}  // extern "C"
