// K3/K4 -- mode lift Phi = S * Q_r (FP64 DMMA) with the column norms fused into the epilogue, then the
// normalisation pass; plus the small boundary-patch lift.
//
// Replaces src/ITHACA_CORE/ITHACAPOD/ITHACAPOD.C:194 (modesEig = SnapMatrix * eigenVectoreig), :196-243
// (normFact_i = sqrt(phi_i^T diag(V) phi_i) for "L2", sqrt(phi_i^T phi_i) for "Frobenius"; in parallel the
// partial sums are reduced before the sqrt) and :254-264 (division of every column by its factor).
//
// Shape: M = Nd rows (cells x components, 1e6..6e7), N = r (10..100), K = m. Every element of S is used by
// exactly one warp, so the kernel is a streaming pass over S: persistent CTAs, a TMA/mbarrier ring that
// never drains between row tiles, 8 MMA warps each owning 16 rows x all r columns in registers.
// S boxes are {16 rows (128 B) x 16 snapshots} written with the 128-B swizzle; the K permutation
// k = 2t + (ks&1) + 8(ks>>1) makes the A-fragment LDS.64 conflict free, and the Q_r stage is stored
// [k][rpad+2] so that the B-fragment reads (bank 4t+g) are conflict free as well.
#include "../../include/ithaca_b200.h"
#include "common.cuh"
#include <algorithm>

namespace ib200 {

int allreduce_sum(ithaca_b200* h, double* dbuf, size_t count);

constexpr int LIFT_ROWS = 128;   // rows per CTA tile (8 warps x 16)
constexpr int LIFT_KT = 16;      // snapshots per stage
constexpr int LIFT_SBOX = 16 * LIFT_KT * 8;  // bytes of one S box (16 rows x 16 snapshots)
constexpr int LIFT_THREADS = 288;            // 8 MMA warps + 1 producer warp

// Qr[k][n] = Q[k + n*ldq] for k < m, n < r, 0 for the padding columns; pitch = rpad + 2.
__global__ void lift_prep_q_kernel(const double* __restrict__ Q, int m, int ldq, int r, int pitch, double* __restrict__ Qr)
{
    const int total = m * pitch;
    for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
        const int k = idx / pitch, n = idx % pitch;
        Qr[idx] = (n < r) ? Q[(size_t)k + (size_t)n * ldq] : 0.0;
    }
}

template <int NB>
__global__ void __launch_bounds__(LIFT_THREADS, 1)
lift_kernel(const __grid_constant__ CUtensorMap tmS, const __grid_constant__ CUtensorMap tmQ,
            const double* __restrict__ V, int dim, int norm_kind, long long Nd, long long ld /* of Phi */, int m, int r,
            int pitch, int qstage_bytes, int nstages, long long ntiles, double* __restrict__ Phi,
            double* __restrict__ norm_part)
{
    extern __shared__ __align__(1024) unsigned char smem[];
    const int stage_bytes = 8 * LIFT_SBOX + qstage_bytes;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)nstages * stage_bytes);
    uint64_t* empty = full + nstages;
    double* colsum = reinterpret_cast<double*>(empty + nstages);  // [8 warps][NB*8]

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int kiters = (m + LIFT_KT - 1) / LIFT_KT;
    long long my_tiles = 0;
    if ((long long)blockIdx.x < ntiles) my_tiles = (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x;
    const long long total_it = my_tiles * kiters;

    if (tid == 0) {
        for (int s = 0; s < nstages; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], 8); }
        mbar_fence_init();
    }
    for (int i = tid; i < 8 * NB * 8; i += LIFT_THREADS) colsum[i] = 0.0;
    __syncthreads();

    if (warp == 8) {
        // ---------------- producer ----------------
        if (lane == 0) {
            tma_prefetch_desc(&tmS);
            tma_prefetch_desc(&tmQ);
            const uint32_t tx = 8 * LIFT_SBOX + (uint32_t)(LIFT_KT * pitch * 8);
            long long it = 0;
            for (long long tl = 0; tl < my_tiles; tl++) {
                const long long row0 = ((long long)blockIdx.x + tl * gridDim.x) * LIFT_ROWS;
                for (int kk = 0; kk < kiters; kk++, it++) {
                    const int s = (int)(it % nstages);
                    if (it >= nstages) mbar_wait(&empty[s], (uint32_t)((it / nstages) - 1) & 1);
                    unsigned char* dst = smem + (size_t)s * stage_bytes;
                    mbar_arrive_expect_tx(&full[s], tx);
                    for (int b = 0; b < 8; b++)
                        tma_load_2d(dst + b * LIFT_SBOX, &tmS, (int)(row0 + 16 * b), kk * LIFT_KT, &full[s]);
                    tma_load_2d(dst + 8 * LIFT_SBOX, &tmQ, 0, kk * LIFT_KT, &full[s]);
                }
            }
        }
        return;
    }

    // ---------------- MMA warps ----------------
    const int g = lane >> 2, t = lane & 3;
    uint32_t aoff[4][2];
    int krow[4];
#pragma unroll
    for (int ks = 0; ks < 4; ks++) {
        const int k = 2 * t + (ks & 1) + 8 * (ks >> 1);
        krow[ks] = k * pitch + g;
#pragma unroll
        for (int i = 0; i < 2; i++) {
            const int row = 8 * i + g;
            aoff[ks][i] = (uint32_t)(k * 128 + ((((row >> 1) ^ (k & 7)) << 4) | ((row & 1) << 3)));
        }
    }
    double acc[2][NB][2];
#pragma unroll
    for (int i = 0; i < 2; i++)
#pragma unroll
        for (int j = 0; j < NB; j++) acc[i][j][0] = acc[i][j][1] = 0.0;

    double* mycol = colsum + warp * NB * 8;
    long long it = 0;
    int prev = -1;   // stage still to be handed back
    for (long long tl = 0; tl < my_tiles; tl++) {
        for (int kk = 0; kk < kiters; kk++, it++) {
            const int s = (int)(it % nstages);
            mbar_wait(&full[s], (uint32_t)(it / nstages) & 1);
            mbar_release_previous(empty, prev, lane);   // delayed release (common.cuh)
            const unsigned char* pa = smem + (size_t)s * stage_bytes + warp * LIFT_SBOX;
            const double* qs = reinterpret_cast<const double*>(smem + (size_t)s * stage_bytes + 8 * LIFT_SBOX);
#pragma unroll
            for (int ks = 0; ks < 4; ks++) {
                const double a0 = *reinterpret_cast<const double*>(pa + aoff[ks][0]);
                const double a1 = *reinterpret_cast<const double*>(pa + aoff[ks][1]);
#pragma unroll
                for (int j = 0; j < NB; j++) {
                    const double b = qs[krow[ks] + 8 * j];
                    dmma884(acc[0][j][0], acc[0][j][1], a0, b);
                    dmma884(acc[1][j][0], acc[1][j][1], a1, b);
                }
            }
            prev = s;
        }
        // ---- epilogue of this row tile: store unnormalised modes, accumulate weighted column sums ----
        const long long row0 = ((long long)blockIdx.x + tl * gridDim.x) * LIFT_ROWS + warp * 16;
#pragma unroll
        for (int i = 0; i < 2; i++) {
            const long long row = row0 + 8 * i + g;
            const bool ok = row < Nd;
            const double w = ok ? (norm_kind == 0 ? V[row / dim] : 1.0) : 0.0;
#pragma unroll
            for (int j = 0; j < NB; j++) {
                const int c = 8 * j + 2 * t;
                if (ok) {
                    if (c < r) Phi[(size_t)c * ld + row] = acc[i][j][0];
                    if (c + 1 < r) Phi[(size_t)(c + 1) * ld + row] = acc[i][j][1];
                }
                acc[i][j][0] = w * acc[i][j][0] * acc[i][j][0];
                acc[i][j][1] = w * acc[i][j][1] * acc[i][j][1];
            }
        }
#pragma unroll
        for (int j = 0; j < NB; j++) {
            double s0 = acc[0][j][0] + acc[1][j][0], s1 = acc[0][j][1] + acc[1][j][1];
#pragma unroll
            for (int o = 4; o < 32; o <<= 1) {
                s0 += __shfl_xor_sync(0xffffffffu, s0, o);
                s1 += __shfl_xor_sync(0xffffffffu, s1, o);
            }
            if (g == 0) { mycol[8 * j + 2 * t] += s0; mycol[8 * j + 2 * t + 1] += s1; }
            acc[0][j][0] = acc[0][j][1] = acc[1][j][0] = acc[1][j][1] = 0.0;
        }
    }
    // ---- CTA partial: fixed-order sum over the 8 warps (named barrier: the producer warp has exited) ----
    asm volatile("bar.sync 1, 256;" ::: "memory");
    if (tid < NB * 8) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < 8; w++) s += colsum[w * NB * 8 + tid];
        norm_part[(size_t)blockIdx.x * (NB * 8) + tid] = s;
    }
}

// norm2[c] = sum over CTAs (fixed order) of the partial column sums.
__global__ void lift_norm_reduce_kernel(const double* __restrict__ part, int nblocks, int ncol, double* __restrict__ norm2, int accumulate)
{
    const int c = blockIdx.x;
    __shared__ double sh[256];
    double s = 0.0;
    for (int b = threadIdx.x; b < nblocks; b += blockDim.x) s += part[(size_t)b * ncol + c];
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int o = blockDim.x / 2; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) norm2[c] = accumulate ? norm2[c] + sh[0] : sh[0];
}

__global__ void lift_sqrt_kernel(double* __restrict__ nrm, int r)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < r) nrm[i] = sqrt(nrm[i]);
}

// Phi(:,c) /= normfact[c]   (ITHACAPOD.C:256: modesEig.col(i).array() / normFact(i,0))
__global__ void lift_scale_kernel(double* __restrict__ Phi, long long Nd, long long ld, int r,
                                  const double* __restrict__ nrm)
{
    const int c = blockIdx.y;
    const double f = nrm[c];
    double* col = Phi + (size_t)c * ld;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < Nd; i += (long long)gridDim.x * blockDim.x)
        col[i] = col[i] / f;
}

// Boundary rows: out(row, c) = (sum_k Sbc(row,k) Q(k,c)) / normfact[c]   (ITHACAPOD.C:245-264)
__global__ void lift_patch_kernel(const double* __restrict__ Sbc, long long rows, int m, const double* __restrict__ Q,
                                  int r, const double* __restrict__ nrm, double* __restrict__ out)
{
    const long long total = rows * r;
    for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
         idx += (long long)gridDim.x * blockDim.x) {
        const long long row = idx % rows;
        const int c = (int)(idx / rows);
        double s = 0.0;
        for (int k = 0; k < m; k++) s += Sbc[(size_t)k * rows + row] * Q[(size_t)k + (size_t)c * m];
        out[idx] = s / nrm[c];
    }
}

// What one lift launch reads and writes: S = m columns of ldS doubles with Nd valid rows, weights V (one per dim rows),
// Phi columns of ldPhi doubles (the row-streamed path writes a chunk's rows into the rank's full mode matrix).
struct LiftSrc {
    const double* S; long long ldS; long long Nd; const double* V; int dim; long long ldPhi;
};

template <int NB>
static int lift_run(ithaca_b200* h, const LiftSrc& src, int m, double* Phi, int r, int norm_kind, const CUtensorMap& tmS, const CUtensorMap& tmQ,
                    int pitch, int qstage_bytes, long long ntiles, int* grid_out)
{
    const int stage_bytes = 8 * LIFT_SBOX + qstage_bytes;
    const int fixed = 8 * NB * 8 * 8 + 2 * 8 * 8 + 64;
    int nstages = std::min(8, (220 * 1024 - fixed) / (stage_bytes + 16));
    // two CTAs per SM when they fit: more loads in flight for the small-r (HBM-bound) case
    int per_sm = 1;
    if (nstages >= 8) {
        per_sm = 2;
        nstages = std::max(4, std::min(8, (113 * 1024 - fixed) / (stage_bytes + 16)));   // (227 KB - 2 x 1 KB reserved) / 2 CTAs
    }
    if (const char* e = env_knob("ITHACA_B200_LIFT_PERSM")) { if (atoi(e) == 1 && per_sm == 2) { per_sm = 1; nstages = 8; } }   // A/B
    if (const char* e = env_knob("ITHACA_B200_LIFT_STAGES")) nstages = std::max(2, std::min(nstages, atoi(e)));
    const size_t smem = (size_t)nstages * (stage_bytes + 16) + fixed;
    IB_CUDA(h, cudaFuncSetAttribute(lift_kernel<NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int grid = (int)std::min<long long>(ntiles, (long long)per_sm * h->num_sms);
    const size_t need = (size_t)grid * NB * 8 * sizeof(double);
    if (need > h->dNormPart_cap) {
        if (h->dNormPart) cudaFree(h->dNormPart);
        h->dNormPart = nullptr; h->dNormPart_cap = 0;
        IB_CUDA(h, cudaMalloc(&h->dNormPart, need));
        h->dNormPart_cap = need;
    }
    lift_kernel<NB><<<grid, LIFT_THREADS, smem, h->stream>>>(tmS, tmQ, src.V, src.dim, norm_kind, src.Nd, src.ldPhi,
                                                             m, r, pitch, qstage_bytes, nstages, ntiles,
                                                             Phi, h->dNormPart);
    IB_CUDA(h, cudaGetLastError());
    *grid_out = grid;
    return IB200_OK;
}

constexpr int LIFT_MAXC = 104;   // columns per launch (13 n8 fragments); wider lifts run as column chunks

static int lift_pick_nb(int rc)
{
    const int nbs[] = {2, 3, 4, 7, 13};
    for (int c : nbs) if (c * 8 >= rc) return c;
    return 13;
}

// Buffers of a lift of r columns from an m-row Q: the k-major Q stage, the norms, and (phi_ld > 0) r mode columns of phi_ld doubles.
static int lift_ensure(ithaca_b200* h, int m, int r, long long phi_ld)
{
    const int pitch = lift_pick_nb(std::min(r, LIFT_MAXC)) * 8 + 2;
    const size_t qr_need = (size_t)m * pitch * sizeof(double);
    if (qr_need > h->dQr_cap) {
        if (h->dQr) cudaFree(h->dQr);
        h->dQr = nullptr; h->dQr_cap = 0;
        IB_CUDA(h, cudaMalloc(&h->dQr, qr_need));
        h->dQr_cap = qr_need;
    }
    const size_t phi_need = (size_t)phi_ld * r * sizeof(double);
    if (phi_need > h->dPhi_cap) {
        if (h->dPhi) cudaFree(h->dPhi);
        h->dPhi = nullptr; h->dPhi_cap = 0;
        IB_CUDA(h, cudaMalloc(&h->dPhi, phi_need));
        h->dPhi_cap = phi_need;
    }
    if (r > h->norm_cap) {
        if (h->dNorm) cudaFree(h->dNorm);
        h->dNorm = nullptr; h->norm_cap = 0;
        const int cap = std::max(LIFT_MAXC, (r + 63) / 64 * 64);
        IB_CUDA(h, cudaMalloc(&h->dNorm, (size_t)cap * sizeof(double)));
        h->norm_cap = cap;
    }
    return IB200_OK;
}

// Phi = S Q for ONE column chunk (r <= LIFT_MAXC): Q = rows [0, m) and columns [0, r) of a col-major matrix with leading
// dimension ldq. The kernel also leaves the partial squared norms of what it wrote in h->dNormPart (grid x 8 NB).
static int lift_core(ithaca_b200* h, const LiftSrc& src, int m, const double* Q, int ldq, int r, int NB, int norm_kind, double* Phi,
                     int* grid_out)
{
    const int rpad = NB * 8, pitch = rpad + 2;
    lift_prep_q_kernel<<<std::min(1024, (m * pitch + 255) / 256), 256, 0, h->stream>>>(Q, m, ldq, r, pitch, h->dQr);
    CUtensorMap tmS, tmQ;
    int rc = make_tmap_2d(h, &tmS, src.S, (uint64_t)src.Nd, (uint64_t)m, (uint64_t)src.ldS * 8, 16, LIFT_KT, true);
    if (rc) return rc;
    rc = make_tmap_2d(h, &tmQ, h->dQr, (uint64_t)pitch, (uint64_t)m, (uint64_t)pitch * 8, (uint32_t)pitch, LIFT_KT, false);
    if (rc) return rc;
    const int qstage_bytes = (LIFT_KT * pitch * 8 + 1023) / 1024 * 1024;
    const long long ntiles = (src.Nd + LIFT_ROWS - 1) / LIFT_ROWS;
    switch (NB) {
        case 2: rc = lift_run<2>(h, src, m, Phi, r, norm_kind, tmS, tmQ, pitch, qstage_bytes, ntiles, grid_out); break;
        case 3: rc = lift_run<3>(h, src, m, Phi, r, norm_kind, tmS, tmQ, pitch, qstage_bytes, ntiles, grid_out); break;
        case 4: rc = lift_run<4>(h, src, m, Phi, r, norm_kind, tmS, tmQ, pitch, qstage_bytes, ntiles, grid_out); break;
        case 7: rc = lift_run<7>(h, src, m, Phi, r, norm_kind, tmS, tmQ, pitch, qstage_bytes, ntiles, grid_out); break;
        default: rc = lift_run<13>(h, src, m, Phi, r, norm_kind, tmS, tmQ, pitch, qstage_bytes, ntiles, grid_out); break;
    }
    h->launches += 2;
    return rc;
}

// Phi = S Q_r for any r <= m, in column chunks of at most LIFT_MAXC (the reference's default nmodes = 0 means m - 2 or m
// modes, ITHACAPOD.C:111-130). norm2 != null: the r squared column norms (fixed-order sum over the CTAs of each chunk),
// assigned or, for the row-streamed path, added to what is there.
static int lift_cols(ithaca_b200* h, const LiftSrc& src, int m, const double* Q, int ldq, int r, int norm_kind, double* Phi,
                     double* norm2, bool norm_accumulate)
{
    for (int c0 = 0; c0 < r; c0 += LIFT_MAXC) {
        const int rc_ = std::min(LIFT_MAXC, r - c0);
        const int NB = lift_pick_nb(rc_);
        int grid = 0;
        int rc = lift_core(h, src, m, Q + (size_t)c0 * ldq, ldq, rc_, NB, norm_kind, Phi + (size_t)c0 * src.ldPhi, &grid);
        if (rc) return rc;
        if (norm2) {
            lift_norm_reduce_kernel<<<rc_, 256, 0, h->stream>>>(h->dNormPart, grid, NB * 8, norm2 + c0, norm_accumulate ? 1 : 0);
            h->launches += 1;
        }
    }
    IB_CUDA(h, cudaGetLastError());
    return IB200_OK;
}

// squared norms (already in h->dNorm, per rank) -> all-reduce -> sqrt -> Phi(:, c) /= n_c   (Phi: Nd rows, columns of ld doubles)
static int lift_finalize(ithaca_b200* h, int r, long long Nd, long long ld)
{
    int rc = allreduce_sum(h, h->dNorm, (size_t)r);
    if (rc) return rc;
    lift_sqrt_kernel<<<(r + 127) / 128, 128, 0, h->stream>>>(h->dNorm, r);
    cudaEventRecord(h->timers[T_LIFT].b, h->stream);
    h->timers[T_LIFT].used = true;
    cudaEventRecord(h->timers[T_SCALE].a, h->stream);
    dim3 sg((unsigned)std::min<long long>((Nd + 1023) / 1024, 4LL * h->num_sms), (unsigned)r);
    lift_scale_kernel<<<sg, 256, 0, h->stream>>>(h->dPhi, Nd, ld, r, h->dNorm);
    cudaEventRecord(h->timers[T_SCALE].b, h->stream);
    h->timers[T_SCALE].used = true;
    IB_CUDA(h, cudaGetLastError());
    h->launches += 2;
    return IB200_OK;
}

int lift_launch(ithaca_b200* h, int r, int norm_kind)
{
    int rc = lift_ensure(h, h->m, r, h->ld);
    if (rc) return rc;
    h->r = r;
    cudaEventRecord(h->timers[T_LIFT].a, h->stream);
    const LiftSrc src{h->dS, h->ld, h->Nd, h->dV, h->dim, h->ld};
    if ((rc = lift_cols(h, src, h->m, h->dQ, h->m, r, norm_kind, h->dPhi, h->dNorm, false))) return rc;
    return lift_finalize(h, r, h->Nd, h->ld);
}

// Phi(:,c) *= f[c]; nrm[c] = 1 / f[c] (so that lift_patch's division applies the same factor)
__global__ void lift_mul_kernel(double* __restrict__ Phi, long long Nd, long long ld, int r, const double* __restrict__ f)
{
    const int c = blockIdx.y;
    const double v = f[c];
    double* col = Phi + (size_t)c * ld;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < Nd; i += (long long)gridDim.x * blockDim.x)
        col[i] = col[i] * v;
}
__global__ void lift_inv_kernel(const double* __restrict__ f, int r, double* __restrict__ nrm)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < r) nrm[i] = 1.0 / f[i];
}

// modes_i = (S q_i) * factors[i] (getWeightedModes, ITHACAPOD.C:771-777). dF: r factors on the device.
int lift_scaled_launch(ithaca_b200* h, int r, const double* dF)
{
    int rc = lift_ensure(h, h->m, r, h->ld);
    if (rc) return rc;
    h->r = r;
    cudaEventRecord(h->timers[T_LIFT].a, h->stream);
    const LiftSrc src{h->dS, h->ld, h->Nd, h->dV, h->dim, h->ld};
    if ((rc = lift_cols(h, src, h->m, h->dQ, h->m, r, 1, h->dPhi, nullptr, false))) return rc;
    cudaEventRecord(h->timers[T_LIFT].b, h->stream);
    h->timers[T_LIFT].used = true;
    cudaEventRecord(h->timers[T_SCALE].a, h->stream);
    dim3 sg((unsigned)std::min<long long>((h->Nd + 1023) / 1024, 4LL * h->num_sms), (unsigned)r);
    lift_mul_kernel<<<sg, 256, 0, h->stream>>>(h->dPhi, h->Nd, h->ld, r, dF);
    lift_inv_kernel<<<(r + 127) / 128, 128, 0, h->stream>>>(dF, r, h->dNorm);
    cudaEventRecord(h->timers[T_SCALE].b, h->stream);
    h->timers[T_SCALE].used = true;
    IB_CUDA(h, cudaGetLastError());
    h->launches += 2;
    return IB200_OK;
}

// ---- snapshot-blocked lift (memory-efficient POD, ITHACAPOD.C:344-628): Phi (+)= S_block Q[first : first+nb, 0:r] ----
__global__ void axpy_cols_kernel(double* __restrict__ y, const double* __restrict__ x, long long Nd, long long ld, int r)
{
    const int c = blockIdx.y;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < Nd; i += (long long)gridDim.x * blockDim.x)
        y[(size_t)c * ld + i] += x[(size_t)c * ld + i];
}

// part[blockIdx.x][c] = sum over this block's rows of w_k Phi(k,c)^2, w = V[k/dim] (L2) or 1 (Frobenius)
__global__ void __launch_bounds__(256)
phi_norm_part_kernel(const double* __restrict__ Phi, long long Nd, long long ld, int r, const double* __restrict__ V, int dim,
                     int norm_kind, double* __restrict__ part)
{
    __shared__ double sh[256];
    const int c = blockIdx.y;
    const double* col = Phi + (size_t)c * ld;
    double s = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < Nd; i += (long long)gridDim.x * blockDim.x) {
        const double v = col[i];
        s += (norm_kind == 0 ? V[i / dim] : 1.0) * v * v;
    }
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) part[(size_t)blockIdx.x * r + c] = sh[0];
}

int lift_block_launch(ithaca_b200* h, const double* S, int nb, int first, int r, bool accumulate, double* dTmp)
{
    int rc = lift_ensure(h, nb, r, h->ld);
    if (rc) return rc;
    h->r = r;
    const LiftSrc src{S, h->ld, h->Nd, h->dV, h->dim, h->ld};
    if ((rc = lift_cols(h, src, nb, h->dQ + first, h->m, r, 1, accumulate ? dTmp : h->dPhi, nullptr, false))) return rc;
    if (accumulate) {
        dim3 g((unsigned)std::min<long long>((h->Nd + 1023) / 1024, 4LL * h->num_sms), (unsigned)r);
        axpy_cols_kernel<<<g, 256, 0, h->stream>>>(h->dPhi, dTmp, h->Nd, h->ld, r);
        h->launches += 1;
    }
    IB_CUDA(h, cudaGetLastError());
    return IB200_OK;
}

int lift_block_finish(ithaca_b200* h, int r, int norm_kind)
{
    const int nblk = (int)std::min<long long>((h->Nd + 4095) / 4096, 4LL * h->num_sms);
    const size_t need = (size_t)nblk * r * sizeof(double);
    if (need > h->dNormPart_cap) {
        if (h->dNormPart) cudaFree(h->dNormPart);
        h->dNormPart = nullptr; h->dNormPart_cap = 0;
        IB_CUDA(h, cudaMalloc(&h->dNormPart, need));
        h->dNormPart_cap = need;
    }
    cudaEventRecord(h->timers[T_LIFT].a, h->stream);
    phi_norm_part_kernel<<<dim3(nblk, r), 256, 0, h->stream>>>(h->dPhi, h->Nd, h->ld, r, h->dV, h->dim, norm_kind, h->dNormPart);
    lift_norm_reduce_kernel<<<r, 256, 0, h->stream>>>(h->dNormPart, nblk, r, h->dNorm, 0);
    h->launches += 2;
    return lift_finalize(h, r, h->Nd, h->ld);
}

// ---- row-streamed lift (pod_stream_*): the chunk's rows of the rank's mode matrix, squared norms added up over the chunks ----
int lift_stream_add(ithaca_b200* h, const double* S, long long ldS, long long NdChunk, const double* Vchunk, int r, int norm_kind,
                    long long row0, long long ldTotal, bool first_chunk)
{
    int rc = lift_ensure(h, h->m, r, ldTotal);
    if (rc) return rc;
    h->r = r;
    if (first_chunk) IB_CUDA(h, cudaMemsetAsync(h->dNorm, 0, (size_t)r * sizeof(double), h->stream));
    const LiftSrc src{S, ldS, NdChunk, Vchunk, h->dim, ldTotal};
    return lift_cols(h, src, h->m, h->dQ, h->m, r, norm_kind, h->dPhi + row0, h->dNorm, true);
}

int lift_stream_finish(ithaca_b200* h, int r, long long NdTotal, long long ldTotal)
{
    cudaEventRecord(h->timers[T_LIFT].a, h->stream);
    return lift_finalize(h, r, NdTotal, ldTotal);
}

// ---- Modes::reconstruct (Containers/Modes.C:294-305): out_j = Phi[:, 0:n] c_j, up to RC_K columns per launch -------------
// HBM-bound streaming pass: one thread per row, the n x kc coefficient block in shared memory, kc accumulators in
// registers; Phi is read once per launch (coalesced), each output column written once.
constexpr int RC_K = 8;
__global__ void __launch_bounds__(256)
reconstruct_kernel(const double* __restrict__ Phi, long long Nd, long long ld, int n, const double* __restrict__ C, int kc,
                   int clip, double* __restrict__ out)
{
    extern __shared__ double sc[];   // [n][RC_K]
    for (int idx = threadIdx.x; idx < n * RC_K; idx += blockDim.x) {
        const int i = idx / RC_K, j = idx % RC_K;
        sc[idx] = (j < kc) ? C[(size_t)i + (size_t)j * n] : 0.0;
    }
    __syncthreads();
    for (long long row = blockIdx.x * (long long)blockDim.x + threadIdx.x; row < Nd; row += (long long)gridDim.x * blockDim.x) {
        double acc[RC_K];
#pragma unroll
        for (int j = 0; j < RC_K; j++) acc[j] = 0.0;
        for (int i = 0; i < n; i++) {
            const double p = Phi[(size_t)i * ld + row];
#pragma unroll
            for (int j = 0; j < RC_K; j++) acc[j] += p * sc[i * RC_K + j];
        }
#pragma unroll
        for (int j = 0; j < RC_K; j++)
            if (j < kc) out[(size_t)j * ld + row] = (clip && acc[j] < 0.0) ? 0.0 : acc[j];
    }
}

int reconstruct_launch(ithaca_b200* h, int nmodes, const double* dC, int kc, int clip, double* dOut)
{
    const int grid = (int)std::min<long long>((h->Nd + 255) / 256, 16LL * h->num_sms);
    reconstruct_kernel<<<grid, 256, (size_t)nmodes * RC_K * sizeof(double), h->stream>>>(h->dPhi, h->Nd, h->ld, nmodes, dC, kc, clip, dOut);
    IB_CUDA(h, cudaGetLastError());
    h->launches += 1;
    return IB200_OK;
}

int lift_patch_launch(ithaca_b200* h, const double* dSbc, int64_t rows, double* dOut)
{
    const long long total = rows * h->r;
    const int grid = (int)std::min<long long>((total + 255) / 256, 8LL * h->num_sms);
    lift_patch_kernel<<<grid, 256, 0, h->stream>>>(dSbc, rows, h->m, h->dQ, h->r, h->dNorm, dOut);
    IB_CUDA(h, cudaGetLastError());
    h->launches += 1;
    return IB200_OK;
}

}  // namespace ib200
