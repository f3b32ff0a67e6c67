// K2 -- symmetric eigensolve of the m x m Gram matrix by one-sided (Hestenes) Jacobi, FP64, single GPU.
//
// Replaces Spectra::SymEigsSolver<DenseSymMatProd> (ncv = m, LargestAlge) and
// Eigen::SelfAdjointEigenSolver at src/ITHACA_CORE/ITHACAPOD/ITHACAPOD.C:157-180.
//
// Method: W <- G, Q <- I; plane rotations from the right orthogonalise the columns of W while the same
// rotations accumulate in Q, so W = G Q throughout. At convergence W = Q Lambda: lambda_i = q_i . w_i
// (signed, so the reference's negative-eigenvalue branch ITHACAPOD.C:182-186 still sees negative values),
// eigenvectors = columns of Q (orthonormal to working precision for ALL i, including the noise floor).
// Pairs follow the round-robin tournament ordering: m/2 disjoint pairs per round, one warp per pair,
// one grid-wide barrier per round (cooperative launch). Stop when a whole sweep applies no rotation,
// i.e. every pair satisfies |w_i . w_j| <= tol * |w_i| |w_j| with tol = sqrt(m) * 2^-52 (LAPACK dgesvj's
// criterion).
#include "../../include/ithaca_b200.h"
#include "common.cuh"
#include <cooperative_groups.h>
#include <algorithm>
#include <cmath>

namespace cg = cooperative_groups;

namespace ib200 {

constexpr int EIG_THREADS = 256;
constexpr int EIG_MAX_SWEEPS = 40;

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__global__ void eig_init_kernel(const double* __restrict__ G, double* __restrict__ W, double* __restrict__ Q, int m)
{
    const size_t mm = (size_t)m * m;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < mm; idx += (size_t)gridDim.x * blockDim.x) {
        W[idx] = G[idx];
        Q[idx] = ((int)(idx / m) == (int)(idx % m)) ? 1.0 : 0.0;
    }
}

// flags[0..EIG_MAX_SWEEPS): number of rotations applied in each sweep; flags[EIG_MAX_SWEEPS] = sweeps done.
__global__ void __launch_bounds__(EIG_THREADS)
eig_jacobi_kernel(double* W, double* Q, int m, double tol, int* flags)
{
    cg::grid_group grid = cg::this_grid();
    const int lane = threadIdx.x & 31;
    const int gwarp = blockIdx.x * (EIG_THREADS / 32) + (threadIdx.x >> 5);
    const int nwarps = gridDim.x * (EIG_THREADS / 32);
    const int n = (m + 1) & ~1;  // players in the tournament (one dummy when m is odd)
    const int npairs = n / 2;

    int sweep = 0;
    for (; sweep < EIG_MAX_SWEEPS; sweep++) {
        for (int rd = 0; rd < n - 1; rd++) {
            for (int p = gwarp; p < npairs; p += nwarps) {
                int i, j;
                if (p == 0) { i = n - 1; j = rd; }
                else { i = (rd + p) % (n - 1); j = (rd - p + (n - 1)) % (n - 1); }
                if (i >= m || j >= m) continue;  // dummy player
                if (i > j) { int tmp = i; i = j; j = tmp; }
                double* wi = W + (size_t)i * m;
                double* wj = W + (size_t)j * m;
                double a = 0.0, b = 0.0, c = 0.0;
                for (int k = lane; k < m; k += 32) {
                    const double x = wi[k], y = wj[k];
                    a += x * x; b += y * y; c += x * y;
                }
                a = warp_sum(a); b = warp_sum(b); c = warp_sum(c);
                const double ab = a * b;
                if (!(ab > 0.0) || fabs(c) <= tol * sqrt(ab)) continue;
                // rotation that zeroes w_i . w_j
                const double zeta = (b - a) / (2.0 * c);
                const double tt = copysign(1.0, zeta) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                const double cs = 1.0 / sqrt(1.0 + tt * tt);
                const double sn = cs * tt;
                double* qi = Q + (size_t)i * m;
                double* qj = Q + (size_t)j * m;
                for (int k = lane; k < m; k += 32) {
                    const double x = wi[k], y = wj[k];
                    wi[k] = cs * x - sn * y;
                    wj[k] = sn * x + cs * y;
                    const double u = qi[k], v = qj[k];
                    qi[k] = cs * u - sn * v;
                    qj[k] = sn * u + cs * v;
                }
                if (lane == 0) atomicAdd(&flags[sweep], 1);
            }
            grid.sync();
        }
        if (*(volatile int*)&flags[sweep] == 0) { sweep++; break; }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) flags[EIG_MAX_SWEEPS] = sweep;
}

// lambda_i = q_i . w_i, one warp per column.
__global__ void eig_values_kernel(const double* __restrict__ W, const double* __restrict__ Q, int m,
                                  double* __restrict__ lam)
{
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i >= m) return;
    double s = 0.0;
    for (int k = lane; k < m; k += 32) s += Q[(size_t)i * m + k] * W[(size_t)i * m + k];
    s = warp_sum(s);
    if (lane == 0) lam[i] = s;
}

// Descending rank sort (stable) + column permutation with a deterministic sign (largest |component| > 0).
// One block per output column.
__global__ void eig_sort_permute_kernel(const double* __restrict__ lam, const double* __restrict__ Q, int m,
                                        double* __restrict__ lam_sorted, double* __restrict__ Qs)
{
    __shared__ int s_rank;
    __shared__ double s_red[32];
    __shared__ int s_idx[32];
    const int i = blockIdx.x;
    const double li = lam[i];
    int cnt = 0;
    for (int j = threadIdx.x; j < m; j += blockDim.x) {
        const double lj = lam[j];
        if (lj > li || (lj == li && j < i)) cnt++;
    }
    // block reduce cnt
    for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    if ((threadIdx.x & 31) == 0) s_idx[threadIdx.x >> 5] = cnt;
    __syncthreads();
    if (threadIdx.x == 0) {
        int tot = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); w++) tot += s_idx[w];
        s_rank = tot;
    }
    __syncthreads();
    const int rank = s_rank;
    // sign: component of largest magnitude (lowest index on ties)
    double best = -1.0;
    int bidx = 0;
    for (int k = threadIdx.x; k < m; k += blockDim.x) {
        const double v = fabs(Q[(size_t)i * m + k]);
        if (v > best) { best = v; bidx = k; }
    }
    for (int o = 16; o > 0; o >>= 1) {
        const double ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bidx, o);
        if (ob > best || (ob == best && oi < bidx)) { best = ob; bidx = oi; }
    }
    __syncthreads();
    if ((threadIdx.x & 31) == 0) { s_red[threadIdx.x >> 5] = best; s_idx[threadIdx.x >> 5] = bidx; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (int)(blockDim.x >> 5); w++)
            if (s_red[w] > s_red[0] || (s_red[w] == s_red[0] && s_idx[w] < s_idx[0])) { s_red[0] = s_red[w]; s_idx[0] = s_idx[w]; }
    }
    __syncthreads();
    const double sgn = (Q[(size_t)i * m + s_idx[0]] < 0.0) ? -1.0 : 1.0;
    for (int k = threadIdx.x; k < m; k += blockDim.x) Qs[(size_t)rank * m + k] = sgn * Q[(size_t)i * m + k];
    if (threadIdx.x == 0) lam_sorted[rank] = li;
}

int eig_launch(ithaca_b200* h, int* sweeps)
{
    const int m = h->m;
    // dQ doubles as the Jacobi accumulator; the sorted eigenvectors end in dW, then the two are swapped.
    double* lam_tmp = nullptr;
    IB_CUDA(h, cudaMalloc(&lam_tmp, (size_t)m * sizeof(double)));
    IB_CUDA(h, cudaMemsetAsync(h->dFlags, 0, 64 * sizeof(int), h->stream));
    const int ib = 256;
    const int ig = (int)std::min<size_t>(((size_t)m * m + ib - 1) / ib, 8 * (size_t)h->num_sms);
    eig_init_kernel<<<ig, ib, 0, h->stream>>>(h->dG, h->dW, h->dQ, m);

    const int npairs = (m + 1) / 2;
    const int wpb = EIG_THREADS / 32;
    int maxb = 0;
    IB_CUDA(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&maxb, eig_jacobi_kernel, EIG_THREADS, 0));
    int nblocks = std::max(1, std::min((npairs + wpb - 1) / wpb, maxb * h->num_sms));
    double tol = std::sqrt((double)m) * 2.220446049250313e-16;
    double* W = h->dW;
    double* Q = h->dQ;
    int mm = m;
    int* flags = h->dFlags;
    void* args[] = {&W, &Q, &mm, &tol, &flags};
    IB_CUDA(h, cudaLaunchCooperativeKernel((void*)eig_jacobi_kernel, dim3(nblocks), dim3(EIG_THREADS), args, 0, h->stream));
    eig_values_kernel<<<(m + 7) / 8, 256, 0, h->stream>>>(h->dW, h->dQ, m, lam_tmp);
    // sorted eigenvectors are written over W (no longer needed once lambda is known)
    eig_sort_permute_kernel<<<m, 128, 0, h->stream>>>(lam_tmp, h->dQ, m, h->dEvals, h->dW);
    IB_CUDA(h, cudaGetLastError());
    h->launches += 4;
    int hflags[EIG_MAX_SWEEPS + 1];
    IB_CUDA(h, cudaMemcpyAsync(hflags, h->dFlags, sizeof(hflags), cudaMemcpyDeviceToHost, h->stream));
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    cudaFree(lam_tmp);
    std::swap(h->dQ, h->dW);
    const int done = hflags[EIG_MAX_SWEEPS];
    if (sweeps) *sweeps = done;
    if (done >= EIG_MAX_SWEEPS && hflags[EIG_MAX_SWEEPS - 1] != 0) {
        set_error(h, "The Eigenvalue Decomposition did not succeed (Jacobi did not converge)");
        return IB200_ERR_NOCONV;
    }
    return IB200_OK;
}

}  // namespace ib200
