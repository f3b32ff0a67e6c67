// K2 -- symmetric eigensolve of the m x m Gram matrix: Cholesky-preconditioned one-sided (Hestenes) Jacobi, FP64.
//
// Replaces Spectra::SymEigsSolver<DenseSymMatProd> (ncv = m, LargestAlge) and Eigen::SelfAdjointEigenSolver at
// src/ITHACA_CORE/ITHACAPOD/ITHACAPOD.C:157-180.
//
// Method (Veselic-Hari / Drmac): G + sigma I = L L^T (blocked Cholesky on the device), then plane rotations from the
// right orthogonalise the COLUMNS of X = L:  X V = U Sigma  =>  G = U (Sigma^2 - sigma I) U^T.
//   eigenvalues  lambda_i = |x_i|^2 - sigma   (signed, so the reference's negative-shift branch ITHACAPOD.C:182-186
//                still sees negative values), eigenvectors u_i = x_i / |x_i| -- no rotation accumulator is needed.
// Working on the factor instead of on G converges in about half the sweeps (the iteration sees cond(G), not
// cond(G)^2) and halves the work per rotation. sigma = 0 when G is numerically positive definite; when the
// factorisation meets a non-positive pivot (exactly singular or indefinite input) it is repeated with a shift
// (first a few ulps of the largest diagonal entry, then a Gershgorin bound), which leaves the eigenvectors unchanged.
//
// Parallel ordering: block-cyclic. Columns are cut into blocks of b; a tournament over the blocks gives nb-1 rounds
// per sweep; in every round one CTA owns a block pair, keeps its 2b columns in shared memory and rotates all b^2 cross
// pairs (b inner rounds of b disjoint pairs, one warp per pair); intra-block pairs are done once per sweep. One grid
// barrier per block round (cooperative launch) instead of one per column-pair round.
// Stop when a whole sweep applies no rotation: |x_i . x_j| <= tol |x_i| |x_j|, tol = sqrt(m) 2^-52 (LAPACK dgesvj).
#include "../../include/ithaca_b200.h"
#include "common.cuh"
#include <cooperative_groups.h>
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

namespace cg = cooperative_groups;

namespace ib200 {

constexpr int EIG_MAX_SWEEPS = 40;
constexpr int CH_B = 32;  // Cholesky panel width

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// X = lower(scale G) + sigma I, strict upper triangle = 0 (column-major m x m). scale is a power of two (exact).
__global__ void eig_init_kernel(const double* __restrict__ G, double* __restrict__ X, int m, double sigma, double scale)
{
    const size_t mm = (size_t)m * m;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < mm; idx += (size_t)gridDim.x * blockDim.x) {
        const int i = (int)(idx % m), j = (int)(idx / m);
        X[idx] = (i > j) ? scale * G[idx] : (i == j ? scale * G[idx] + sigma : 0.0);
    }
}

// max diagonal entry and Gershgorin bound max_i (sum_{j != i} |G_ij| - G_ii): out[0], out[1]. One block; one warp per
// column (G is symmetric: column i is row i), lanes over the contiguous rows.
__global__ void __launch_bounds__(1024) eig_bounds_kernel(const double* __restrict__ G, int m, double* __restrict__ out)
{
    __shared__ double sd[32], sg[32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    double dmax = -1e300, gmax = -1e300;
    for (int i = w; i < m; i += nw) {
        const double* col = G + (size_t)i * m;
        double s = 0.0;
        for (int j = lane; j < m; j += 32) if (j != i) s += fabs(col[j]);
        s = warp_sum(s);
        const double d = col[i];
        dmax = fmax(dmax, fabs(d));
        gmax = fmax(gmax, s - d);
    }
    if (lane == 0) { sd[w] = dmax; sg[w] = gmax; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < nw; k++) { sd[0] = fmax(sd[0], sd[k]); sg[0] = fmax(sg[0], sg[k]); }
        out[0] = sd[0];
        out[1] = sg[0];
    }
}

// ---- small matrices: init + Cholesky in ONE single-CTA launch ------------------------------------------------------
// X = chol(lower(scale G) + sigma I) (lower factor, strict upper triangle = 0), right-looking, unblocked, on the packed lower
// triangle in shared memory: replaces eig_init_kernel + the 3 m / CH_B launches of the blocked factorisation.
// Measured (scripts/eig_bench.py, whole eigensolve): m = 100 1.265 ms against 1.284 ms blocked; m = 200 2.73 ms against 2.66 ms
// (three block-wide barriers per column cost more than the blocked path's launch gaps), so it serves m <= CHS_MAX = 128.
// flag[0] = 1 on a non-positive (or NaN) pivot (X is then unspecified; the caller retries with a shift).
constexpr int CHS_MAX = 128;
__global__ void __launch_bounds__(1024) chol_small_kernel(const double* __restrict__ G, double* __restrict__ X, int m, double sigma,
                                                          double scale, int* __restrict__ flag)
{
    extern __shared__ double A[];   // column j at off(j) = j m - j (j - 1) / 2, rows j..m-1
    __shared__ int bad;
    const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, w = tid >> 5, nw = nt >> 5;
    auto off = [m](int j) { return j * m - (j * (j - 1)) / 2 - j; };   // A[off(j) + i] = element (i, j), i >= j
    if (tid == 0) bad = 0;
    for (int j = w; j < m; j += nw)
        for (int i = j + lane; i < m; i += 32) A[off(j) + i] = scale * G[(size_t)i + (size_t)j * m] + (i == j ? sigma : 0.0);
    __syncthreads();
    for (int k = 0; k < m; k++) {
        const double d = A[off(k) + k];
        if (!(d > 0.0)) { if (tid == 0) bad = 1; break; }   // uniform: every thread reads the same d
        const double r = sqrt(d), rinv = 1.0 / r;
        __syncthreads();                                     // everyone has read d before it is overwritten
        for (int i = k + tid; i < m; i += nt) A[off(k) + i] = (i == k) ? r : A[off(k) + i] * rinv;
        __syncthreads();
        const double* ck = A + off(k);
        for (int j = k + 1 + w; j < m; j += nw) {
            const double ljk = ck[j];
            double* cj = A + off(j);
            for (int i = j + lane; i < m; i += 32) cj[i] -= ck[i] * ljk;
        }
        __syncthreads();
    }
    __syncthreads();
    if (bad) { if (tid == 0) flag[0] = 1; return; }
    for (int j = w; j < m; j += nw)
        for (int i = lane; i < m; i += 32) X[(size_t)i + (size_t)j * m] = (i >= j) ? A[off(j) + i] : 0.0;
}

// ---- blocked right-looking Cholesky, panel width CH_B ------------------------------------------------------------
// (1) diagonal block: unblocked Cholesky in shared memory; flag[0] = 1 on a non-positive (or NaN) pivot.
__global__ void __launch_bounds__(CH_B * CH_B) chol_diag_kernel(double* __restrict__ X, int m, int k0, int* __restrict__ flag)
{
    __shared__ double a[CH_B][CH_B + 1];
    const int nb = min(CH_B, m - k0);
    const int r = threadIdx.x % CH_B, c = threadIdx.x / CH_B;
    if (r < nb && c < nb) a[r][c] = X[(size_t)(k0 + r) + (size_t)(k0 + c) * m];
    __syncthreads();
    for (int k = 0; k < nb; k++) {
        const double d = a[k][k];
        __syncthreads();
        if (!(d > 0.0)) { if (threadIdx.x == 0) flag[0] = 1; return; }
        const double s = sqrt(d);
        if (c == k && r >= k && r < nb) a[r][k] = (r == k) ? s : a[r][k] / s;
        __syncthreads();
        if (c > k && c < nb && r >= c && r < nb) a[r][c] -= a[r][k] * a[c][k];
        __syncthreads();
    }
    if (r < nb && c < nb) X[(size_t)(k0 + r) + (size_t)(k0 + c) * m] = (r >= c) ? a[r][c] : 0.0;
}

// (2) panel: rows below the diagonal block, X[i, k0:k0+nb] <- X[i, k0:k0+nb] L_kk^{-T}. One thread per row.
__global__ void __launch_bounds__(128) chol_panel_kernel(double* __restrict__ X, int m, int k0)
{
    __shared__ double l[CH_B][CH_B + 1];
    const int nb = min(CH_B, m - k0);
    for (int idx = threadIdx.x; idx < nb * nb; idx += blockDim.x) {
        const int r = idx % nb, c = idx / nb;
        l[r][c] = X[(size_t)(k0 + r) + (size_t)(k0 + c) * m];
    }
    __syncthreads();
    const int i = k0 + nb + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    double x[CH_B];
#pragma unroll
    for (int c = 0; c < CH_B; c++) x[c] = (c < nb) ? X[(size_t)i + (size_t)(k0 + c) * m] : 0.0;
#pragma unroll
    for (int c = 0; c < CH_B; c++) {
        if (c < nb) {
            double s = x[c];
#pragma unroll
            for (int q = 0; q < CH_B; q++) if (q < c) s -= x[q] * l[c][q];
            x[c] = s / l[c][c];
        }
    }
#pragma unroll
    for (int c = 0; c < CH_B; c++) if (c < nb) X[(size_t)i + (size_t)(k0 + c) * m] = x[c];
}

// (3) trailing update: X[i,j] -= sum_c L[i,k0+c] L[j,k0+c] for i >= j >= k0+nb. 32x32 tiles of the lower triangle.
__global__ void __launch_bounds__(256) chol_update_kernel(double* __restrict__ X, int m, int k0)
{
    const int nb = min(CH_B, m - k0);
    const int t0 = k0 + nb;
    if (blockIdx.x > blockIdx.y) return;  // lower triangle of tiles: column tile <= row tile
    const int i0 = t0 + blockIdx.y * 32, j0 = t0 + blockIdx.x * 32;
    __shared__ double li[32][CH_B + 1], lj[32][CH_B + 1];
    for (int idx = threadIdx.x; idx < 32 * CH_B; idx += 256) {
        const int r = idx % 32, c = idx / 32;
        li[r][c] = (i0 + r < m && c < nb) ? X[(size_t)(i0 + r) + (size_t)(k0 + c) * m] : 0.0;
        lj[r][c] = (j0 + r < m && c < nb) ? X[(size_t)(j0 + r) + (size_t)(k0 + c) * m] : 0.0;
    }
    __syncthreads();
    const int r = threadIdx.x % 32;
    for (int cc = threadIdx.x / 32; cc < 32; cc += 8) {
        const int i = i0 + r, j = j0 + cc;
        if (i < m && j < m && i >= j) {
            double s = 0.0;
#pragma unroll
            for (int c = 0; c < CH_B; c++) s += li[r][c] * lj[cc][c];
            X[(size_t)i + (size_t)j * m] -= s;
        }
    }
}

// ---- block-cyclic one-sided Jacobi on the columns of X (m x m, column-major) -----------------------------------
// One warp rotates one column pair held in shared memory. Returns 1 if a rotation was applied.
__device__ __forceinline__ int rotate_pair(double* xi, double* xj, int m, int lane, double tol)
{
    double a = 0.0, b = 0.0, c = 0.0;
    for (int k = lane; k < m; k += 32) {
        const double x = xi[k], y = xj[k];
        a += x * x; b += y * y; c += x * y;
    }
    a = warp_sum(a); b = warp_sum(b); c = warp_sum(c);
    const double ab = a * b;
    if (!(ab > 0.0) || fabs(c) <= tol * sqrt(ab)) return 0;
    const double zeta = (b - a) / (2.0 * c);
    const double tt = copysign(1.0, zeta) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
    const double cs = 1.0 / sqrt(1.0 + tt * tt);
    const double sn = cs * tt;
    for (int k = lane; k < m; k += 32) {
        const double x = xi[k], y = xj[k];
        xi[k] = cs * x - sn * y;
        xj[k] = sn * x + cs * y;
    }
    return 1;
}

// flags[0..EIG_MAX_SWEEPS): rotations applied in each sweep; flags[EIG_MAX_SWEEPS] = sweeps done.
// b columns per block, nb blocks; blockDim = 32*b (one warp per pair of an inner round). smem: 2*b*m doubles.
__global__ void eig_jacobi_block_kernel(double* X, int m, int b, int nb, double tol, int* flags)
{
    cg::grid_group grid = cg::this_grid();
    extern __shared__ double sx[];  // 2b columns of m doubles
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nbp = (nb + 1) & ~1;       // tournament players (one dummy block when nb is odd)
    const int npairs = nbp / 2;
    auto ncols = [&](int blk) { return min(b, m - blk * b); };

    int sweep = 0;
    for (; sweep < EIG_MAX_SWEEPS; sweep++) {
        int rot = 0;
        // ---- phase 0: pairs inside each block (tournament over the block's columns) ----
        for (int blk = blockIdx.x; blk < nb; blk += gridDim.x) {
            const int nc = ncols(blk);
            double* g = X + (size_t)blk * b * m;
            for (int idx = threadIdx.x; idx < nc * m; idx += blockDim.x) sx[idx] = g[idx];
            __syncthreads();
            const int n = (nc + 1) & ~1;
            for (int rd = 0; rd < n - 1; rd++) {
                if (warp < n / 2) {
                    int i, j;
                    if (warp == 0) { i = n - 1; j = rd; }
                    else { i = (rd + warp) % (n - 1); j = (rd - warp + (n - 1)) % (n - 1); }
                    if (i < nc && j < nc) {
                        if (i > j) { const int t = i; i = j; j = t; }
                        rot += rotate_pair(sx + (size_t)i * m, sx + (size_t)j * m, m, lane, tol);
                    }
                }
                __syncthreads();
            }
            for (int idx = threadIdx.x; idx < nc * m; idx += blockDim.x) g[idx] = sx[idx];
            __syncthreads();
        }
        grid.sync();
        // ---- block tournament: nbp-1 rounds, each CTA owns block pairs (I,J) and rotates all cross pairs ----
        for (int rd = 0; rd < nbp - 1; rd++) {
            for (int p = blockIdx.x; p < npairs; p += gridDim.x) {
                int I, J;
                if (p == 0) { I = nbp - 1; J = rd; }
                else { I = (rd + p) % (nbp - 1); J = (rd - p + (nbp - 1)) % (nbp - 1); }
                if (I >= nb || J >= nb) continue;  // dummy block (uniform over the CTA)
                if (I > J) { const int t = I; I = J; J = t; }
                const int ni = ncols(I), nj = ncols(J);
                double* gi = X + (size_t)I * b * m;
                double* gj = X + (size_t)J * b * m;
                double* sj = sx + (size_t)b * m;
                for (int idx = threadIdx.x; idx < ni * m; idx += blockDim.x) sx[idx] = gi[idx];
                for (int idx = threadIdx.x; idx < nj * m; idx += blockDim.x) sj[idx] = gj[idx];
                __syncthreads();
                for (int s = 0; s < b; s++) {
                    const int i = warp, j = (warp + s) % b;
                    if (i < ni && j < nj) rot += rotate_pair(sx + (size_t)i * m, sj + (size_t)j * m, m, lane, tol);
                    __syncthreads();
                }
                for (int idx = threadIdx.x; idx < ni * m; idx += blockDim.x) gi[idx] = sx[idx];
                for (int idx = threadIdx.x; idx < nj * m; idx += blockDim.x) gj[idx] = sj[idx];
                __syncthreads();
            }
            grid.sync();
        }
        if (lane == 0 && rot) atomicAdd(&flags[sweep], rot);
        grid.sync();
        if (*(volatile int*)&flags[sweep] == 0) { sweep++; break; }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) flags[EIG_MAX_SWEEPS] = sweep;
}

// ---- register-resident variant for m <= 32*NE <= 2048 -------------------------------------------------------------
// Same ordering as eig_jacobi_block_kernel, but in a block pair (I, J) warp i keeps column i of block I in REGISTERS for
// all b inner rounds (it meets columns i, i+1, ... of block J in turn), only block J lives in shared memory (padded to
// 32*NE doubles per column, so the hot loop has no bounds checks), and the squared norms travel with the columns and
// are updated by the rotation (|x'|^2 = a - t c, |y'|^2 = b + t c; recomputed from scratch at every block-pair visit),
// so a rotation costs ONE warp reduction (the dot product) instead of three.
// CL = true: the whole grid is ONE thread-block cluster (<= 16 CTAs: m <= 256 with b = 8) and the barrier between block rounds is
// the hardware cluster barrier (barrier.cluster release / acquire, ~380 cycles) instead of the cooperative grid barrier (~2 us of
// global atomics and polling), 27 of them per sweep at m = 200.
template <int NE, bool CL>
__device__ __forceinline__ void eig_grid_sync()
{
    if constexpr (CL) cg::this_cluster().sync();
    else cg::this_grid().sync();
}

template <int NE, bool CL = false>
__global__ void __launch_bounds__(NE > 32 ? 256 : NE > 16 ? 320 : 512, 1) eig_jacobi_reg_kernel(double* X, int m, int b, int nb, double tol, int* flags)
{
    extern __shared__ double sx[];           // b columns of LDC doubles | b squared norms
    constexpr int LDC = 32 * NE;
    double* nrm = sx + (size_t)b * LDC;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nbp = (nb + 1) & ~1;
    const int npairs = nbp / 2;
    auto ncols = [&](int blk) { return min(b, m - blk * b); };
    // block `blk` -> shared memory (zero padded), squared norms by the owning warps
    auto load_block = [&](int blk, int nc) {
        const double* g = X + (size_t)blk * b * m;
        for (int c = warp; c < b; c += (blockDim.x >> 5)) {
            double s2 = 0.0;
#pragma unroll
            for (int e = 0; e < NE; e++) {
                const int k = lane + 32 * e;
                const double v = (c < nc && k < m) ? g[(size_t)c * m + k] : 0.0;
                sx[(size_t)c * LDC + k] = v;
                s2 += v * v;
            }
            s2 = warp_sum(s2);
            if (lane == 0) nrm[c] = s2;
        }
    };
    auto store_block = [&](int blk, int nc) {
        double* g = X + (size_t)blk * b * m;
        for (int c = warp; c < nc; c += (blockDim.x >> 5))
#pragma unroll
            for (int e = 0; e < NE; e++) {
                const int k = lane + 32 * e;
                if (k < m) g[(size_t)c * m + k] = sx[(size_t)c * LDC + k];
            }
    };

    int sweep = 0;
    for (; sweep < EIG_MAX_SWEEPS; sweep++) {
        int rot = 0;
        // ---- phase 0: pairs inside each block (tournament over the block's columns, in shared memory) ----
        for (int blk = blockIdx.x; blk < nb; blk += gridDim.x) {
            const int nc = ncols(blk);
            load_block(blk, nc);
            __syncthreads();
            const int n = (nc + 1) & ~1;
            for (int rd = 0; rd < n - 1; rd++) {
                if (warp < n / 2) {
                    int i, j;
                    if (warp == 0) { i = n - 1; j = rd; }
                    else { i = (rd + warp) % (n - 1); j = (rd - warp + (n - 1)) % (n - 1); }
                    if (i < nc && j < nc) {
                        if (i > j) { const int t = i; i = j; j = t; }
                        rot += rotate_pair(sx + (size_t)i * LDC, sx + (size_t)j * LDC, m, lane, tol);
                    }
                }
                __syncthreads();
            }
            store_block(blk, nc);
            __syncthreads();
        }
        eig_grid_sync<NE, CL>();
        // ---- block tournament ----
        for (int rd = 0; rd < nbp - 1; rd++) {
            for (int p = blockIdx.x; p < npairs; p += gridDim.x) {
                int I, J;
                if (p == 0) { I = nbp - 1; J = rd; }
                else { I = (rd + p) % (nbp - 1); J = (rd - p + (nbp - 1)) % (nbp - 1); }
                if (I >= nb || J >= nb) continue;  // dummy block (uniform over the CTA)
                if (I > J) { const int t = I; I = J; J = t; }
                const int ni = ncols(I), nj = ncols(J);
                // my column of block I -> registers
                double xr[NE];
                double a = 0.0;
                double* gi = X + ((size_t)I * b + warp) * m;
                if (warp < ni) {
#pragma unroll
                    for (int e = 0; e < NE; e++) {
                        const int k = lane + 32 * e;
                        xr[e] = (k < m) ? gi[k] : 0.0;
                        a += xr[e] * xr[e];
                    }
                    a = warp_sum(a);
                } else {
#pragma unroll
                    for (int e = 0; e < NE; e++) xr[e] = 0.0;
                }
                load_block(J, nj);
                __syncthreads();
                for (int s = 0; s < b; s++) {
                    const int j = (warp + s) % b;
                    if (warp < ni && j < nj) {
                        double* y = sx + (size_t)j * LDC + lane;
                        constexpr bool YREG = NE <= 32;   // NE = 64 (m <= 2048): x alone fills the registers, y is re-read
                        double yr[YREG ? NE : 1];
                        double c0 = 0.0, c1 = 0.0;
#pragma unroll
                        for (int e = 0; e < NE; e++) {
                            const double yv = y[32 * e];
                            if (YREG) yr[e] = yv;
                            if (e & 1) c1 += xr[e] * yv; else c0 += xr[e] * yv;
                        }
                        const double c = warp_sum(c0 + c1);
                        const double bn = nrm[j];
                        const double ab = a * bn;
                        if (ab > 0.0 && c * c > tol * tol * ab) {
                            // t = tan(theta) of the smaller rotation angle: 2c / (d + sign(d) sqrt(d^2 + 4c^2)), d = |y|^2 - |x|^2
                            const double d = bn - a;
                            const double tt = 2.0 * c / (d + copysign(sqrt(d * d + 4.0 * c * c), d));
                            const double cs = rsqrt(1.0 + tt * tt);
                            const double sn = cs * tt;
#pragma unroll
                            for (int e = 0; e < NE; e++) {
                                const double xv = xr[e];
                                const double yv = YREG ? yr[e] : y[32 * e];
                                xr[e] = cs * xv - sn * yv;
                                y[32 * e] = sn * xv + cs * yv;
                            }
                            a -= tt * c;
                            __syncwarp();                       // every lane has read nrm[j] (bn) before lane 0 replaces it
                            if (lane == 0) nrm[j] = bn + tt * c;
                            rot++;
                        }
                    }
                    __syncthreads();
                }
                if (warp < ni) {
#pragma unroll
                    for (int e = 0; e < NE; e++) {
                        const int k = lane + 32 * e;
                        if (k < m) gi[k] = xr[e];
                    }
                }
                store_block(J, nj);
                __syncthreads();
            }
            eig_grid_sync<NE, CL>();
        }
        if (lane == 0 && rot) atomicAdd(&flags[sweep], rot);
        eig_grid_sync<NE, CL>();
        if (*(volatile int*)&flags[sweep] == 0) { sweep++; break; }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) flags[EIG_MAX_SWEEPS] = sweep;
}

// lambda_i = (|x_i|^2 - sigma) / scale, u_i = x_i / |x_i| (in place). One warp per column.
__global__ void eig_finalize_kernel(double* __restrict__ X, int m, double sigma, double inv_scale, double* __restrict__ lam)
{
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i >= m) return;
    double* x = X + (size_t)i * m;
    double s = 0.0;
    for (int k = lane; k < m; k += 32) s += x[k] * x[k];
    s = warp_sum(s);
    const double inv = s > 0.0 ? 1.0 / sqrt(s) : 0.0;
    for (int k = lane; k < m; k += 32) x[k] *= inv;
    if (lane == 0) lam[i] = (s - sigma) * inv_scale;
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

static int cholesky(ithaca_b200* h, double* X, int m, int* flag_dev, bool* ok)
{
    IB_CUDA(h, cudaMemsetAsync(flag_dev, 0, sizeof(int), h->stream));
    for (int k0 = 0; k0 < m; k0 += CH_B) {
        const int nb = std::min(CH_B, m - k0), rest = m - k0 - nb;
        chol_diag_kernel<<<1, CH_B * CH_B, 0, h->stream>>>(X, m, k0, flag_dev);
        h->launches += 1;
        if (rest > 0) {
            chol_panel_kernel<<<(rest + 127) / 128, 128, 0, h->stream>>>(X, m, k0);
            const int nt = (rest + 31) / 32;
            chol_update_kernel<<<dim3(nt, nt), 256, 0, h->stream>>>(X, m, k0);
            h->launches += 2;
        }
    }
    IB_CUDA(h, cudaGetLastError());
    int f = 0;
    IB_CUDA(h, cudaMemcpyAsync(&f, flag_dev, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    *ok = (f == 0);
    return IB200_OK;
}

int eig_launch(ithaca_b200* h, int* sweeps)
{
    const int m = h->m;
    const double eps = 2.220446049250313e-16;
    {
        int rc = dev_ensure(h, h->eigTmp, (size_t)(m + 2) * sizeof(double));   // m eigenvalues + 2 bounds, kept across calls
        if (rc) return rc;
    }
    double* lam_tmp = h->eigTmp.as<double>();
    auto fail = [&](int rc) { return rc; };
    int* flags = h->dFlags;       // [0..40] sweep counters, [48] Cholesky flag
    eig_bounds_kernel<<<1, 1024, 0, h->stream>>>(h->dG, m, lam_tmp + m);
    double bounds[2] = {0.0, 0.0};
    if (cudaMemcpyAsync(bounds, lam_tmp + m, 2 * sizeof(double), cudaMemcpyDeviceToHost, h->stream) != cudaSuccess ||
        cudaStreamSynchronize(h->stream) != cudaSuccess) {
        set_error(h, "eig: device failure");
        return fail(IB200_ERR_CUDA);
    }
    h->launches += 1;
    // power-of-two scaling to max |G_ii| in [1, 2): exact, and keeps the squared quantities of the rotation test in range
    double scale = 1.0;
    if (bounds[0] > 0.0 && std::isfinite(bounds[0])) { int ex = 0; std::frexp(bounds[0], &ex); scale = std::ldexp(1.0, 1 - ex); }
    const double dmax = bounds[0] * scale, gersh = std::max(bounds[1], 0.0) * scale;
    const double tiny = 16.0 * m * eps * dmax;
    const double shifts[3] = {0.0, tiny, gersh * (1.0 + 1e-3) + tiny};
    const int ib = 256;
    const int ig = (int)std::min<size_t>(((size_t)m * m + ib - 1) / ib, 8 * (size_t)h->num_sms);
    double sigma = 0.0;
    bool ok = false;
    for (int attempt = 0; attempt < 3 && !ok; attempt++) {
        sigma = shifts[attempt];
        if (attempt > 0 && !(sigma > shifts[attempt - 1])) continue;
        const char* evc = env_knob("ITHACA_B200_CHOL");   // "blocked" keeps the multi-launch factorisation (A/B measurements)
        if (m <= CHS_MAX && !(evc && strcmp(evc, "blocked") == 0)) {
            const size_t csm = (size_t)m * (m + 1) / 2 * sizeof(double);
            IB_CUDA(h, cudaFuncSetAttribute(chol_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)csm));
            IB_CUDA(h, cudaMemsetAsync(flags + 48, 0, sizeof(int), h->stream));
            chol_small_kernel<<<1, 1024, csm, h->stream>>>(h->dG, h->dW, m, sigma, scale, flags + 48);
            h->launches += 1;
            IB_CUDA(h, cudaGetLastError());
            int f = 0;
            IB_CUDA(h, cudaMemcpyAsync(&f, flags + 48, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
            IB_CUDA(h, cudaStreamSynchronize(h->stream));
            ok = (f == 0);
            continue;
        }
        eig_init_kernel<<<ig, ib, 0, h->stream>>>(h->dG, h->dW, m, sigma, scale);
        h->launches += 1;
        int rc = cholesky(h, h->dW, m, flags + 48, &ok);
        if (rc) return fail(rc);
    }
    if (!ok) {
        set_error(h, "The Eigenvalue Decomposition did not succeed (matrix is not finite)");
        return fail(IB200_ERR_NOCONV);
    }

    IB_CUDA(h, cudaMemsetAsync(flags, 0, 48 * sizeof(int), h->stream));
    double tol = std::sqrt((double)m) * eps;
    double* X = h->dW;
    const char* ev = env_knob("ITHACA_B200_EIG");       // "block" selects the shared-memory kernel (A/B measurements)
    const char* evb = env_knob("ITHACA_B200_EIG_B");    // columns per block (tuning)
    const bool use_reg = m <= 2048 && !(ev && strcmp(ev, "block") == 0);
    int b, threads;
    size_t smem;
    void* kern = nullptr;
    if (use_reg) {
        // register-resident kernel: one block of b padded columns (+ b squared norms) in shared memory
        int NE = 2;
        while (32 * NE < m) NE *= 2;
        const int bmax = NE > 32 ? 8 : NE > 16 ? 10 : 16;
        b = std::min(8, m);   // measured (scripts/eig_bench.py): b = 8 beats 4, 12, 16 at m = 200 and m = 1000
        if (evb) b = std::max(2, std::min({atoi(evb), bmax, m}));
        if (b < 2 && m >= 2) b = 2;
        threads = std::max(64, 32 * b);
        smem = (size_t)b * (32 * NE + 1) * sizeof(double);
        switch (NE) {
            case 2: kern = (void*)eig_jacobi_reg_kernel<2>; break;
            case 4: kern = (void*)eig_jacobi_reg_kernel<4>; break;
            case 8: kern = (void*)eig_jacobi_reg_kernel<8>; break;
            case 16: kern = (void*)eig_jacobi_reg_kernel<16>; break;
            case 32: kern = (void*)eig_jacobi_reg_kernel<32>; break;
            default: kern = (void*)eig_jacobi_reg_kernel<64>; break;
        }
    } else {
        // block size: 2b columns of m doubles must fit shared memory; one warp per pair of an inner round
        b = std::min({16, m, (int)((200 * 1024) / (16 * (size_t)m))});
        if (b < 1) { set_error(h, "pod_eig: more than 12800 snapshots are not supported"); return fail(IB200_ERR_ARG); }
        if (evb) b = std::max(1, std::min(atoi(evb), b));
        if (b < 2 && m >= 2) b = std::min(2, m);
        threads = std::max(64, 32 * b);
        smem = (size_t)2 * b * m * sizeof(double);
        kern = (void*)eig_jacobi_block_kernel;
    }
    const int nb = (m + b - 1) / b;
    // one thread-block cluster when the grid has <= 16 CTAs (register kernel, NE <= 8: m <= 256)
    bool clustered = false;
    if (use_reg && m <= 256 && (nb + 1) / 2 <= 16) {
        const char* es = env_knob("ITHACA_B200_EIG_SYNC");   // "grid": keep the cooperative grid barrier (A/B measurements)
        if (!(es && strcmp(es, "grid") == 0)) {
            clustered = true;
            switch (m <= 64 ? 2 : m <= 128 ? 4 : 8) {
                case 2: kern = (void*)eig_jacobi_reg_kernel<2, true>; break;
                case 4: kern = (void*)eig_jacobi_reg_kernel<4, true>; break;
                default: kern = (void*)eig_jacobi_reg_kernel<8, true>; break;
            }
        }
    }
    if (smem > 227 * 1024) { set_error(h, "pod_eig: too many snapshots for the Jacobi block kernel"); return fail(IB200_ERR_ARG); }
    IB_CUDA(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int maxb = 0;
    IB_CUDA(h, cudaOccupancyMaxActiveBlocksPerMultiprocessorWithFlags(&maxb, kern, threads, smem, cudaOccupancyDefault));
    if (maxb < 1) { set_error(h, "pod_eig: Jacobi kernel does not fit one SM"); return fail(IB200_ERR_CUDA); }
    const int want = std::max((nb + 1) / 2, 1);
    const int nblocks = std::max(1, std::min(want, maxb * h->num_sms));
    int mm = m, bb = b, nbb = nb;
    void* args[] = {&X, &mm, &bb, &nbb, &tol, &flags};
    bool launched = false;
    if (clustered) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(nblocks);
        cfg.blockDim = dim3(threads);
        cfg.dynamicSmemBytes = smem;
        cfg.stream = h->stream;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = (unsigned)nblocks;
        at[0].val.clusterDim.y = 1;
        at[0].val.clusterDim.z = 1;
        cfg.attrs = at;
        cfg.numAttrs = 1;
        int ncl = 0;
        if ((nblocks <= 8 || cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess) &&
            cudaOccupancyMaxActiveClusters(&ncl, kern, &cfg) == cudaSuccess && ncl >= 1 &&
            cudaLaunchKernelExC(&cfg, kern, args) == cudaSuccess)
            launched = true;
        else {
            cudaGetLastError();   // the cluster does not fit this device / configuration: cooperative launch of the grid-barrier kernel
            switch (m <= 64 ? 2 : m <= 128 ? 4 : 8) {
                case 2: kern = (void*)eig_jacobi_reg_kernel<2>; break;
                case 4: kern = (void*)eig_jacobi_reg_kernel<4>; break;
                default: kern = (void*)eig_jacobi_reg_kernel<8>; break;
            }
            IB_CUDA(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        }
    }
    if (!launched) IB_CUDA(h, cudaLaunchCooperativeKernel(kern, dim3(nblocks), dim3(threads), args, smem, h->stream));
    eig_finalize_kernel<<<(m + 7) / 8, 256, 0, h->stream>>>(h->dW, m, sigma, 1.0 / scale, lam_tmp);
    eig_sort_permute_kernel<<<m, 128, 0, h->stream>>>(lam_tmp, h->dW, m, h->dEvals, h->dQ);
    IB_CUDA(h, cudaGetLastError());
    h->launches += 3;
    int hflags[EIG_MAX_SWEEPS + 1];
    IB_CUDA(h, cudaMemcpyAsync(hflags, flags, sizeof(hflags), cudaMemcpyDeviceToHost, h->stream));
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    const int done = hflags[EIG_MAX_SWEEPS];
    if (sweeps) *sweeps = done;
    if (done >= EIG_MAX_SWEEPS && hflags[EIG_MAX_SWEEPS - 1] != 0) {
        set_error(h, "The Eigenvalue Decomposition did not succeed (Jacobi did not converge)");
        return IB200_ERR_NOCONV;
    }
    return IB200_OK;
}

}  // namespace ib200
