// K1 -- volume-weighted Gram matrix  G = S^T diag(V) S  (FP64, DMMA.8x8x4 tensor cores).
//
// Replaces ITHACAutilities::getMassMatrix (src/ITHACA_CORE/ITHACAutilities/ITHACAcoeffsMass.C:67-118:
// Eigen GEMM for V.size()<=1e6, rank-1 update loop above) and the V repetition of getMassMatrixFV
// (ITHACAcoeffsMass.C:312-326).
//
// Shape: M = N = m (100..2000) tiny, K = Nd (1e6..6e7) huge -> symmetric (lower-triangle) tile schedule
// + split-K over cell slabs with a deterministic two-stage reduction (no atomics).
//
// Data path: S is snapshot-contiguous ([m][ld] doubles, cells contiguous = OpenFOAM's own field
// storage), so both MMA operands are K-contiguous. A TMA box is {16 cells (128 B) x 8*WF snapshots},
// written with the 128-B swizzle; fragments are read with a K permutation that makes every LDS.64
// bank-conflict free (see kperm below). One producer warp streams boxes + the 16 cell weights of the
// stage through an mbarrier ring; 15 MMA warps each own one (8*WF)x(8*WF) output tile in registers.
#include "../../include/ithaca_b200.h"
#include "common.cuh"
#include <algorithm>
#include <cstring>

namespace ib200 {

constexpr int GRAM_WF = 5;                       // warp tile = WF x WF fragments of 8x8
constexpr int GRAM_BOXROWS = 8 * GRAM_WF;        // snapshots per TMA box
constexpr int GRAM_KT = 16;                      // cells (K) per stage
constexpr int GRAM_BOXB = GRAM_BOXROWS * 128;    // bytes per box
constexpr int GRAM_NBMAX = 10;                   // boxes per job
constexpr int GRAM_MAXTILES = 15;                // MMA warps per CTA
constexpr int GRAM_THREADS = 512;                // 15 MMA warps + 1 producer warp

struct GramJob {
    int nbox;
    int ntiles;
    int boxrow[GRAM_NBMAX];          // first column (snapshot / field) of each box
    int boxmap[GRAM_NBMAX];          // 0 = operand A (tile rows), 1 = operand B (tile columns)
    unsigned char ta[GRAM_MAXTILES]; // box index of the tile's rows (i)
    unsigned char tb[GRAM_MAXTILES]; // box index of the tile's columns (j)
    unsigned char pad[2];
    unsigned int mask[GRAM_MAXTILES]; // bit i*WF+j: fragment (i,j) is on or below the diagonal and inside m
};

__global__ void __launch_bounds__(GRAM_THREADS, 1)
gram_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            const double* __restrict__ V, int dim, long long Nd, int Ma, int Nb,
            const GramJob* __restrict__ jobs, int njobs, long long kchunk, int nstages, int stage_bytes,
            double* __restrict__ ws)
{
    constexpr int WF = GRAM_WF;
    extern __shared__ __align__(1024) unsigned char smem[];
    // [nstages][stage_bytes] panels | [nstages][16] weights | full[nstages] | empty[nstages] | job
    unsigned char* panels = smem;
    double* wsm = reinterpret_cast<double*>(smem + (size_t)nstages * stage_bytes);
    uint64_t* full = reinterpret_cast<uint64_t*>(wsm + nstages * GRAM_KT);
    uint64_t* empty = full + nstages;
    GramJob* job = reinterpret_cast<GramJob*>(empty + nstages);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int jobid = blockIdx.x % njobs;
    const long long ksl = blockIdx.x / njobs;
    const long long k_begin = ksl * kchunk;
    long long k_end = k_begin + kchunk;
    if (k_end > Nd) k_end = Nd;
    const int niter = (k_end > k_begin) ? (int)((k_end - k_begin + GRAM_KT - 1) / GRAM_KT) : 0;

    if (tid < (int)(sizeof(GramJob) / 4))
        reinterpret_cast<int*>(job)[tid] = reinterpret_cast<const int*>(jobs + jobid)[tid];
    if (tid == 0) {
        for (int s = 0; s < nstages; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], GRAM_MAXTILES);
        }
        mbar_fence_init();
    }
    __syncthreads();
    const int ntiles = job->ntiles;

    if (warp == GRAM_MAXTILES) {
        // ---------------- producer warp ----------------
        if (lane == 0) { tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmB); }
        const int nbox = job->nbox;
        for (int it = 0; it < niter; it++) {
            const int s = it % nstages;
            if (it >= nstages) mbar_wait(&empty[s], ((it / nstages) - 1) & 1);
            const long long k0 = k_begin + (long long)it * GRAM_KT;
            if (lane < GRAM_KT) {
                const long long k = k0 + lane;
                wsm[s * GRAM_KT + lane] = (k < Nd) ? (V ? V[k / dim] : 1.0) : 0.0;
            }
            __syncwarp();
            if (lane == 0) {
                mbar_arrive_expect_tx(&full[s], (uint32_t)(nbox * GRAM_BOXB));
                unsigned char* dst = panels + (size_t)s * stage_bytes;
                for (int b = 0; b < nbox; b++)
                    tma_load_2d(dst + b * GRAM_BOXB, job->boxmap[b] ? &tmB : &tmA, (int)k0, job->boxrow[b], &full[s]);
            }
        }
        return;
    }

    // ---------------- MMA warps ----------------
    const bool active = warp < ntiles;
    const int g = lane >> 2, t = lane & 3;
    const int ta = active ? job->ta[warp] : 0, tb = active ? job->tb[warp] : 0;
    const unsigned mask = active ? job->mask[warp] : 0u;

    // K permutation inside a 16-cell stage: at step ks the thread with t = lane&3 takes
    // k = 2*ks + (t&1) + 8*(t>>1). With the 128-B swizzle this makes the 16 lanes of a half-warp hit 16
    // distinct 8-byte banks (chunk (k>>1)^g covers {ks^g} U {(ks+4)^g}, g = 4 consecutive rows).
    uint32_t koff[4];
    int kidx[4];
#pragma unroll
    for (int ks = 0; ks < 4; ks++) {
        const int k = 2 * ks + (t & 1) + 8 * (t >> 1);
        kidx[ks] = k;
        koff[ks] = (uint32_t)(g * 128) + (uint32_t)((((k >> 1) ^ g) << 4) | ((k & 1) << 3));
    }

    double acc[WF][WF][2];
#pragma unroll
    for (int i = 0; i < WF; i++)
#pragma unroll
        for (int j = 0; j < WF; j++) acc[i][j][0] = acc[i][j][1] = 0.0;

    for (int it = 0; it < niter; it++) {
        const int s = it % nstages;
        mbar_wait(&full[s], (it / nstages) & 1);
        if (active) {
            const unsigned char* pa = panels + (size_t)s * stage_bytes + ta * GRAM_BOXB;
            const unsigned char* pb = panels + (size_t)s * stage_bytes + tb * GRAM_BOXB;
            const double* wv = wsm + s * GRAM_KT;
#pragma unroll
            for (int ks = 0; ks < 4; ks++) {
                const double w = wv[kidx[ks]];
                double a[WF], b[WF];
#pragma unroll
                for (int i = 0; i < WF; i++)
                    a[i] = *reinterpret_cast<const double*>(pa + i * 1024 + koff[ks]) * w;
#pragma unroll
                for (int j = 0; j < WF; j++)
                    b[j] = *reinterpret_cast<const double*>(pb + j * 1024 + koff[ks]);
#pragma unroll
                for (int i = 0; i < WF; i++)
#pragma unroll
                    for (int j = 0; j < WF; j++)
                        if (mask & (1u << (i * WF + j))) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
    }

    // ---------------- epilogue: partial tile -> workspace slab (row-major lower triangle) ----------------
    if (active) {
        double* out = ws + (size_t)ksl * (size_t)Ma * (size_t)Nb;
        const int row0 = job->boxrow[ta], col0 = job->boxrow[tb];
#pragma unroll
        for (int i = 0; i < WF; i++) {
            const int gi = row0 + 8 * i + g;
#pragma unroll
            for (int j = 0; j < WF; j++) {
                if (!(mask & (1u << (i * WF + j)))) continue;
                const int gj = col0 + 8 * j + 2 * t;
                if (gi < Ma) {
                    if (gj < Nb) out[(size_t)gi * Nb + gj] = acc[i][j][0];
                    if (gj + 1 < Nb) out[(size_t)gi * Nb + gj + 1] = acc[i][j][1];
                }
            }
        }
    }
}

// Deterministic split-K reduction (+ symmetrisation): out(i,j) = sum_s ws[s][i*Nb+j], written to
// out[i*rs + (j / cinner)*cso + (j % cinner)*csi] (two-level column stride: matrix columns or tensor slices).
__global__ void gram_reduce_kernel(const double* __restrict__ ws, int nsplit, int Ma, int Nb, int symmetric,
                                   double* __restrict__ out, long long rs, int cinner, long long cso, long long csi)
{
    const size_t mm = (size_t)Ma * Nb;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < mm;
         idx += (size_t)gridDim.x * blockDim.x) {
        const int i = (int)(idx / Nb), j = (int)(idx % Nb);
        if (symmetric && j > i) continue;
        double s = 0.0;
        for (int k = 0; k < nsplit; k++) s += ws[(size_t)k * mm + idx];
        out[(size_t)i * rs + (size_t)(j / cinner) * cso + (size_t)(j % cinner) * csi] = s;
        if (symmetric) out[(size_t)j * rs + (size_t)(i / cinner) * cso + (size_t)(i % cinner) * csi] = s;
    }
}

// Host: cover the lower triangle of the (m/8)^2 fragment grid with jobs of <= 15 warp tiles.
static void build_jobs(int m, std::vector<GramJob>& jobs)
{
    constexpr int WF = GRAM_WF;
    const int nfr = (m + 7) / 8;
    const int T = (nfr + WF - 1) / WF;   // warp-tile rows
    const int SG = 5;                    // tile rows per super-tile
    const int nS = (T + SG - 1) / SG;
    auto tile_mask = [&](int ti, int tj) {
        unsigned mk = 0;
        for (int i = 0; i < WF; i++)
            for (int j = 0; j < WF; j++) {
                int fi = ti * WF + i, fj = tj * WF + j;
                if (fi < nfr && fj < nfr && fi >= fj) mk |= 1u << (i * WF + j);
            }
        return mk;
    };
    auto emit = [&](const std::vector<std::pair<int, int>>& tiles) {
        if (tiles.empty()) return;
        GramJob jb;
        std::memset(&jb, 0, sizeof(jb));
        auto box_of = [&](int trow) {
            for (int b = 0; b < jb.nbox; b++)
                if (jb.boxrow[b] == trow * GRAM_BOXROWS) return b;
            jb.boxrow[jb.nbox] = trow * GRAM_BOXROWS;
            return jb.nbox++;
        };
        // balance DMMA counts over the 4 SM sub-partitions (warp w runs on sub-partition w & 3)
        std::vector<std::pair<int, std::pair<int, int>>> order;
        for (auto& tt : tiles) order.push_back({__builtin_popcount(tile_mask(tt.first, tt.second)), tt});
        std::sort(order.begin(), order.end(), [](auto& x, auto& y) { return x.first > y.first; });
        // greedy: heaviest tile first, onto the free warp whose sub-partition carries the least work
        const int nt = (int)order.size();
        int load[4] = {0, 0, 0, 0};
        std::vector<int> warp_tile(nt, -1);
        for (int n = 0; n < nt; n++) {
            int best = -1;
            for (int w = 0; w < nt; w++)
                if (warp_tile[w] < 0 && (best < 0 || load[w & 3] < load[best & 3])) best = w;
            warp_tile[best] = n;
            load[best & 3] += order[n].first;
        }
        jb.ntiles = (int)order.size();
        for (int w = 0; w < jb.ntiles; w++) {
            auto tt = order[warp_tile[w]].second;
            jb.ta[w] = (unsigned char)box_of(tt.first);
            jb.tb[w] = (unsigned char)box_of(tt.second);
            jb.mask[w] = tile_mask(tt.first, tt.second);
        }
        jobs.push_back(jb);
    };
    for (int sI = 0; sI < nS; sI++)
        for (int sJ = 0; sJ <= sI; sJ++) {
            const int r0 = sI * SG, r1 = std::min(T, r0 + SG), c0 = sJ * SG, c1 = std::min(T, c0 + SG);
            if (sI == sJ) {
                std::vector<std::pair<int, int>> tl;
                for (int ti = r0; ti < r1; ti++)
                    for (int tj = c0; tj <= ti; tj++) tl.push_back({ti, tj});
                emit(tl);
            } else {
                for (int part = 0; part < 2; part++) {
                    const int pr0 = part == 0 ? r0 : std::min(r1, r0 + 3);
                    const int pr1 = part == 0 ? std::min(r1, r0 + 3) : r1;
                    std::vector<std::pair<int, int>> tl;
                    for (int ti = pr0; ti < pr1; ti++)
                        for (int tj = c0; tj < c1; tj++) tl.push_back({ti, tj});
                    emit(tl);
                }
            }
        }
}

// Host: rectangular A^T B. The fragment rows (Ma/8) and columns (Nb/8) are cut into balanced tiles of <= WF
// fragments; a job is a group of gr row tiles x gc column tiles (<= 15 warp tiles, <= 10 boxes).
static void build_jobs_rect(int Ma, int Nb, std::vector<GramJob>& jobs)
{
    constexpr int WF = GRAM_WF;
    auto cut = [&](int n, std::vector<std::pair<int, int>>& tiles) {  // (first fragment, #fragments)
        const int nfr = (n + 7) / 8, nt = (nfr + WF - 1) / WF;
        int f = 0;
        for (int t = 0; t < nt; t++) {
            const int len = nfr / nt + (t < nfr % nt ? 1 : 0);
            tiles.push_back({f, len});
            f += len;
        }
    };
    std::vector<std::pair<int, int>> rt, ct;
    cut(Ma, rt);
    cut(Nb, ct);
    const int gr = std::min<int>((int)rt.size(), 3);
    const int gc = std::max(1, std::min<int>({(int)ct.size(), GRAM_MAXTILES / gr, GRAM_NBMAX - gr}));
    for (size_t r0 = 0; r0 < rt.size(); r0 += gr)
        for (size_t c0 = 0; c0 < ct.size(); c0 += gc) {
            GramJob jb;
            std::memset(&jb, 0, sizeof(jb));
            const size_t r1 = std::min(rt.size(), r0 + gr), c1 = std::min(ct.size(), c0 + gc);
            for (size_t r = r0; r < r1; r++) { jb.boxrow[jb.nbox] = 8 * rt[r].first; jb.boxmap[jb.nbox++] = 0; }
            for (size_t c = c0; c < c1; c++) { jb.boxrow[jb.nbox] = 8 * ct[c].first; jb.boxmap[jb.nbox++] = 1; }
            for (size_t r = r0; r < r1; r++)
                for (size_t c = c0; c < c1; c++) {
                    unsigned mk = 0;
                    for (int i = 0; i < rt[r].second; i++)
                        for (int j = 0; j < ct[c].second; j++) mk |= 1u << (i * WF + j);
                    jb.ta[jb.ntiles] = (unsigned char)(r - r0);
                    jb.tb[jb.ntiles] = (unsigned char)((r1 - r0) + (c - c0));
                    jb.mask[jb.ntiles++] = mk;
                }
            jobs.push_back(jb);
        }
}

int atvb_launch(ithaca_b200* h, const double* A, long long lda, int Ma, const double* B, long long ldb, int Nb,
                long long K, const double* V, int dim, int symmetric, double* out, long long rs, int cinner,
                long long cso, long long csi)
{
    std::vector<GramJob> jobs;
    if (symmetric) build_jobs(Ma, jobs);
    else build_jobs_rect(Ma, Nb, jobs);
    const int njobs = (int)jobs.size();
    int nbox_max = 0;
    for (auto& j : jobs) nbox_max = std::max(nbox_max, j.nbox);
    const int stage_bytes = nbox_max * GRAM_BOXB;
    const int smem_cap = 227 * 1024 - 1024;
    int nstages = std::min(8, (int)((smem_cap - (int)sizeof(GramJob) - 64) / (stage_bytes + GRAM_KT * 8 + 16)));
    if (nstages < 2) { set_error(h, "gram: job does not fit shared memory"); return IB200_ERR_ARG; }
    const size_t smem = (size_t)nstages * (stage_bytes + GRAM_KT * 8 + 16) + sizeof(GramJob) + 64;

    // split-K: about two waves of CTAs when there are several jobs, one CTA per SM for a single job;
    // every slab is a multiple of the stage depth and at least 64 stages long.
    int target = (njobs == 1) ? h->num_sms : 2 * h->num_sms;
    int nsplit = std::max(1, target / njobs);
    long long min_chunk = 64LL * GRAM_KT;
    nsplit = (int)std::max(1LL, std::min<long long>(nsplit, (K + min_chunk - 1) / min_chunk));
    long long kchunk = ((K + nsplit - 1) / nsplit + GRAM_KT - 1) / GRAM_KT * GRAM_KT;
    nsplit = (int)((K + kchunk - 1) / kchunk);

    const size_t ws_need = (size_t)nsplit * Ma * Nb * sizeof(double);
    if (ws_need > h->dWs_cap) {
        if (h->dWs) cudaFree(h->dWs);
        h->dWs = nullptr; h->dWs_cap = 0;
        IB_CUDA(h, cudaMalloc(&h->dWs, ws_need));
        h->dWs_cap = ws_need;
    }
    const size_t jb_need = (size_t)njobs * sizeof(GramJob);
    if (jb_need > h->dJobs_cap) {
        if (h->dJobs) cudaFree(h->dJobs);
        h->dJobs = nullptr; h->dJobs_cap = 0;
        IB_CUDA(h, cudaMalloc(&h->dJobs, jb_need));
        h->dJobs_cap = jb_need;
    }
    IB_CUDA(h, cudaMemcpyAsync(h->dJobs, jobs.data(), jb_need, cudaMemcpyHostToDevice, h->stream));

    CUtensorMap tmA, tmB;
    int rc = make_tmap_2d(h, &tmA, A, (uint64_t)K, (uint64_t)Ma, (uint64_t)lda * 8, GRAM_KT, GRAM_BOXROWS, true);
    if (rc) return rc;
    rc = make_tmap_2d(h, &tmB, B, (uint64_t)K, (uint64_t)Nb, (uint64_t)ldb * 8, GRAM_KT, GRAM_BOXROWS, true);
    if (rc) return rc;

    IB_CUDA(h, cudaFuncSetAttribute(gram_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    Timer& tk = h->timers[T_GRAMK];
    IB_CUDA(h, cudaEventRecord(tk.a, h->stream));
    gram_kernel<<<njobs * nsplit, GRAM_THREADS, smem, h->stream>>>(
        tmA, tmB, V, dim, K, Ma, Nb, (const GramJob*)h->dJobs, njobs, kchunk, nstages, stage_bytes, h->dWs);
    IB_CUDA(h, cudaGetLastError());
    IB_CUDA(h, cudaEventRecord(tk.b, h->stream));
    tk.used = true;
    const int rb = 256;
    const int rg = (int)std::min<size_t>(((size_t)Ma * Nb + rb - 1) / rb, 4 * (size_t)h->num_sms);
    gram_reduce_kernel<<<rg, rb, 0, h->stream>>>(h->dWs, nsplit, Ma, Nb, symmetric, out, rs, cinner, cso, csi);
    IB_CUDA(h, cudaGetLastError());
    h->launches += 2;
    return IB200_OK;
}

// POD Gram: G = S^T diag(V) S, symmetric schedule, column-major m x m into h->dG.
int gram_launch(ithaca_b200* h)
{
    return atvb_launch(h, h->dS, h->ld, h->m, h->dS, h->ld, h->m, h->Nd, h->dV, h->dim, 1, h->dG, 1, 0x7fffffff, 0, h->m);
}

}  // namespace ib200
