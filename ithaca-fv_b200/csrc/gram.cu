// K1 / K9 -- weighted contraction  out = A^T diag(V) B  on the FP64 tensor cores (DMMA.8x8x4), with the symmetric
// (SYRK) schedule for the POD Gram matrix  G = S^T diag(V) S.
//
// Replaces ITHACAutilities::getMassMatrix (src/ITHACA_CORE/ITHACAutilities/ITHACAcoeffsMass.C:67-118: Eigen GEMM for
// V.size()<=1e6, rank-1 update loop above), the V repetition of getMassMatrixFV (ITHACAcoeffsMass.C:312-326) and every
// fvc::domainIntegrate(X_i & Y_j) double loop of the steadyNS term builders (steadyNS.C:883-1137).
//
// Shape: M, N (100..2000) tiny, K = cells x components (1e6..6e7) huge -> output-stationary warp tiles + split-K over
// cell slabs with a deterministic reduction (fixed order, no atomics).
//
// Data path: operands are column-contiguous ([cols][ld] doubles, K contiguous = OpenFOAM's own field storage), so both
// MMA operands are K-contiguous. A TMA box is {16 cells (128 B) x 40 columns}, written with the 128-B swizzle;
// fragments are read with a K permutation that makes every LDS.64 bank-conflict free. 16 MMA warps (4 per SM
// sub-partition) each own one <= 40x40 output tile as 5x5 DMMA fragments in registers; lane 0 of warp 0 also drives
// the TMA/mbarrier ring (a dedicated producer warp would cap the MMA warps below the 128 registers they need).
//
// Schedule (host): the output is cut into 5x5-fragment tiles; tiles are packed into jobs of <= 16 tiles sharing <= 10
// boxes (4x4 super-tiles off the diagonal, triangles on it), large tiles are split when warps are spare, and tiles go
// to warps so that the four sub-partitions carry equal DMMA counts. Each job gets a number of K slabs proportional to
// its critical-path cost, so all CTAs of the launch last equally long and fill whole waves of the 148 SMs.
#include "../../include/ithaca_b200.h"
#include "common.cuh"
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <vector>

namespace ib200 {

constexpr int GRAM_WF = 5;                       // warp tile = WF x WF fragments of 8x8
constexpr int GRAM_BOXROWS = 8 * GRAM_WF;        // columns per TMA box
constexpr int GRAM_KT = 16;                      // cells (K) per stage
constexpr int GRAM_BOXB = GRAM_BOXROWS * 128;    // bytes per box
constexpr int GRAM_NBMAX = 10;                   // boxes per job
constexpr int GRAM_WARPS = 16;                   // MMA warps per CTA
constexpr int GRAM_THREADS = GRAM_WARPS * 32;
constexpr int GRAM_TILE = GRAM_BOXROWS * GRAM_BOXROWS;  // doubles per warp tile in the workspace

struct GramJob {
    int nbox;
    int nsplit;                      // K slabs of this job
    long long ws_off;                // first double of this job's workspace: [nsplit][16 tiles][40x40]
    int boxrow[GRAM_NBMAX];          // first column of each box
    int boxmap[GRAM_NBMAX];          // 0 = operand A (tile rows), 1 = operand B (tile columns)
    unsigned char ta[GRAM_WARPS];    // per warp: box index of the tile's rows
    unsigned char tb[GRAM_WARPS];    // per warp: box index of the tile's columns
    unsigned int mask[GRAM_WARPS];   // per warp: bit i*WF+j = fragment (i,j) is computed (0 = idle warp)
};

struct GramCta {
    int job, slot;
    long long k_begin, k_end;        // multiples of GRAM_KT (k_end clipped to K)
};

__global__ void __launch_bounds__(GRAM_THREADS, 1)
gram_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            const double* __restrict__ V, int dim, long long Nd, const GramJob* __restrict__ jobs,
            const GramCta* __restrict__ ctas, int nstages, int stage_bytes, double* __restrict__ ws)
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
    const GramCta cta = ctas[blockIdx.x];
    const long long k_begin = cta.k_begin;
    const int niter = (cta.k_end > k_begin) ? (int)((cta.k_end - k_begin + GRAM_KT - 1) / GRAM_KT) : 0;

    if (tid < (int)(sizeof(GramJob) / 4))
        reinterpret_cast<int*>(job)[tid] = reinterpret_cast<const int*>(jobs + cta.job)[tid];
    if (tid == 0) {
        for (int s = 0; s < nstages; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], GRAM_WARPS);
        }
        mbar_fence_init();
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
    }
    __syncthreads();
    const int nbox = job->nbox;

    // stage loader (warp 0): 16 cell weights by lanes 0..15, then lane 0 arms the barrier and issues the boxes
    auto load_stage = [&](int it, double wv) {
        const int s = it % nstages;
        const long long k0 = k_begin + (long long)it * GRAM_KT;
        if (lane < GRAM_KT) wsm[s * GRAM_KT + lane] = wv;
        __syncwarp();
        if (lane == 0) {
            mbar_arrive_expect_tx(&full[s], (uint32_t)(nbox * GRAM_BOXB));
            unsigned char* dst = panels + (size_t)s * stage_bytes;
            for (int b = 0; b < nbox; b++)
                tma_load_2d(dst + b * GRAM_BOXB, job->boxmap[b] ? &tmB : &tmA, (int)k0, job->boxrow[b], &full[s]);
        }
    };
    auto weight_of = [&](int it) -> double {
        const long long k = k_begin + (long long)it * GRAM_KT + lane;
        return (lane < GRAM_KT && k < Nd) ? (V ? V[k / dim] : 1.0) : 0.0;
    };
    if (warp == 0) {
        const int pre = min(niter, nstages);
        for (int it = 0; it < pre; it++) load_stage(it, weight_of(it));
    }

    const int g = lane >> 2, t = lane & 3;
    const int ta = job->ta[warp], tb = job->tb[warp];
    const unsigned mask = job->mask[warp];

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
        // warp 0: the weights of the stage that will reuse this slot are fetched now and stored after the MMAs
        const bool refill = (warp == 0) && (it + nstages < niter);
        double wnext = 0.0;
        if (refill) wnext = weight_of(it + nstages);
        mbar_wait(&full[s], (it / nstages) & 1);
        if (mask) {
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
        if (refill) {
            mbar_wait(&empty[s], (it / nstages) & 1);  // every warp has released the slot
            load_stage(it + nstages, wnext);
        }
    }

    // ---------------- epilogue: partial tile -> workspace [slot][warp][40][40] ----------------
    if (mask) {
        double* out = ws + job->ws_off + ((size_t)cta.slot * GRAM_WARPS + warp) * GRAM_TILE;
#pragma unroll
        for (int i = 0; i < WF; i++)
#pragma unroll
            for (int j = 0; j < WF; j++) {
                if (!(mask & (1u << (i * WF + j)))) continue;
                *reinterpret_cast<double2*>(out + (8 * i + g) * GRAM_BOXROWS + 8 * j + 2 * t) =
                    make_double2(acc[i][j][0], acc[i][j][1]);
            }
    }
}

// Deterministic split-K reduction (+ symmetrisation), one blockIdx.y per job: out(i,j) = sum over the job's slabs, in
// slab order, written to out[i*rs + (j / cinner)*cso + (j % cinner)*csi] (matrix columns or tensor slices).
__global__ void gram_reduce_kernel(const double* __restrict__ ws, const GramJob* __restrict__ jobs, int Ma, int Nb,
                                   int symmetric, double* __restrict__ out, long long rs, int cinner, long long cso,
                                   long long csi)
{
    const GramJob& jb = jobs[blockIdx.y];
    const int nsplit = jb.nsplit;
    const double* base = ws + jb.ws_off;
    for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < GRAM_WARPS * GRAM_TILE; idx += gridDim.x * blockDim.x) {
        const int w = idx / GRAM_TILE, e = idx % GRAM_TILE, r = e / GRAM_BOXROWS, c = e % GRAM_BOXROWS;
        if (!((jb.mask[w] >> ((r >> 3) * GRAM_WF + (c >> 3))) & 1u)) continue;
        const int i = jb.boxrow[jb.ta[w]] + r, j = jb.boxrow[jb.tb[w]] + c;
        if (i >= Ma || j >= Nb || (symmetric && j > i)) continue;
        double s = 0.0;
        for (int k = 0; k < nsplit; k++) s += base[((size_t)k * GRAM_WARPS + w) * GRAM_TILE + e];
        out[(size_t)i * rs + (size_t)(j / cinner) * cso + (size_t)(j % cinner) * csi] = s;
        if (symmetric) out[(size_t)j * rs + (size_t)(i / cinner) * cso + (size_t)(i % cinner) * csi] = s;
    }
}

// ------------------------------------------------------------------------------------------------
// host: schedule
// ------------------------------------------------------------------------------------------------
namespace {

struct TileB {           // one warp tile while a job is being built
    int rbox, cbox;      // box indices inside the job
    unsigned mask;
    int frags() const { return __builtin_popcount(mask); }
};

struct JobB {
    std::vector<int> boxrow, boxmap;
    std::vector<TileB> tiles;
    int cost = 0;        // critical path: max DMMA count over the four sub-partitions, per k-step
    int box(int row, int map)
    {
        for (size_t b = 0; b < boxrow.size(); b++)
            if (boxrow[b] == row && boxmap[b] == map) return (int)b;
        boxrow.push_back(row);
        boxmap.push_back(map);
        return (int)boxrow.size() - 1;
    }
};

unsigned row_bits(int i) { return 0x1fu << (i * GRAM_WF); }

// Fill spare warps by splitting the heaviest tiles by fragment rows, then spread tiles over the 4 sub-partitions
// (warp w runs on sub-partition w & 3) with the longest-processing-time rule. Warp 0 (which also drives TMA) gets the
// lightest tile of its sub-partition.
void finish_job(JobB& jb, GramJob& out)
{
    std::vector<TileB>& T = jb.tiles;
    while ((int)T.size() < GRAM_WARPS) {
        int best = -1, bestf = 0;
        for (size_t q = 0; q < T.size(); q++) {
            int nrows = 0;
            for (int i = 0; i < GRAM_WF; i++) nrows += (T[q].mask & row_bits(i)) != 0;
            if (nrows >= 2 && T[q].frags() > bestf) { best = (int)q; bestf = T[q].frags(); }
        }
        if (best < 0 || bestf <= 5) break;
        // cut after the fragment row where the running count reaches half of the tile
        unsigned lo = 0;
        int run = 0;
        for (int i = 0; i < GRAM_WF; i++) {
            const unsigned rb = T[best].mask & row_bits(i);
            if (!rb) continue;
            if (2 * run >= bestf) break;
            lo |= rb;
            run += __builtin_popcount(rb);
        }
        TileB hi = T[best];
        hi.mask = T[best].mask & ~lo;
        T[best].mask = lo;
        if (!hi.mask || !lo) { T[best].mask |= hi.mask; break; }
        T.push_back(hi);
    }
    std::sort(T.begin(), T.end(), [](const TileB& x, const TileB& y) { return x.frags() > y.frags(); });
    int load[4] = {0, 0, 0, 0}, cnt[4] = {0, 0, 0, 0};
    std::vector<int> bins[4];
    for (size_t q = 0; q < T.size(); q++) {
        int b = -1;
        for (int s = 0; s < 4; s++)
            if (cnt[s] < GRAM_WARPS / 4 && (b < 0 || load[s] < load[b])) b = s;
        bins[b].push_back((int)q);
        load[b] += T[q].frags();
        cnt[b]++;
    }
    jb.cost = std::max({load[0], load[1], load[2], load[3], 1});
    std::memset(&out, 0, sizeof(out));
    out.nbox = (int)jb.boxrow.size();
    for (int b = 0; b < out.nbox; b++) { out.boxrow[b] = jb.boxrow[b]; out.boxmap[b] = jb.boxmap[b]; }
    for (int s = 0; s < 4; s++) {
        std::vector<int>& v = bins[s];
        if (s == 0) std::reverse(v.begin(), v.end());  // lightest tile of sub-partition 0 first -> warp 0
        for (size_t q = 0; q < v.size(); q++) {
            const int w = s + 4 * (int)q;
            out.ta[w] = (unsigned char)T[v[q]].rbox;
            out.tb[w] = (unsigned char)T[v[q]].cbox;
            out.mask[w] = T[v[q]].mask;
        }
    }
}

// tile mask of fragment rows [fr0, fr0+nr) x columns [fc0, fc0+nc); symmetric: only fragments on/below the diagonal
unsigned tile_mask(int fr0, int nr, int fc0, int nc, bool sym)
{
    unsigned mk = 0;
    for (int i = 0; i < nr; i++)
        for (int j = 0; j < nc; j++)
            if (!sym || fr0 + i >= fc0 + j) mk |= 1u << (i * GRAM_WF + j);
    return mk;
}

// cut n fragments into tiles of <= WF fragments: aligned to WF (symmetric schedule) or balanced (rectangular)
void cut(int nfr, bool aligned, std::vector<std::pair<int, int>>& tiles)
{
    const int nt = (nfr + GRAM_WF - 1) / GRAM_WF;
    int f = 0;
    for (int t = 0; t < nt; t++) {
        const int len = aligned ? std::min(GRAM_WF, nfr - f) : nfr / nt + (t < nfr % nt ? 1 : 0);
        tiles.push_back({f, len});
        f += len;
    }
}

void build_schedule(int Ma, int Nb, bool sym, std::vector<GramJob>& jobs, std::vector<int>& cost)
{
    std::vector<std::pair<int, int>> rt, ct;
    cut((Ma + 7) / 8, sym, rt);
    if (sym) ct = rt;
    else cut((Nb + 7) / 8, false, ct);
    const int SG = 4;
    auto emit = [&](JobB& jb) {
        if (jb.tiles.empty()) return;
        GramJob g;
        finish_job(jb, g);
        jobs.push_back(g);
        cost.push_back(jb.cost);
    };
    for (size_t r0 = 0; r0 < rt.size(); r0 += SG) {
        const int h = (int)std::min<size_t>(SG, rt.size() - r0);
        const size_t cend = sym ? r0 : ct.size();          // rectangular part: tile columns [0, cend)
        const int wc = std::max(1, std::min(GRAM_WARPS / h, GRAM_NBMAX - h));
        for (size_t c0 = 0; c0 < cend; c0 += wc) {
            JobB jb;
            const size_t c1 = std::min(cend, c0 + wc);
            for (int r = 0; r < h; r++)
                for (size_t c = c0; c < c1; c++) {
                    TileB t;
                    t.rbox = jb.box(8 * rt[r0 + r].first, 0);
                    t.cbox = jb.box(8 * ct[c].first, sym ? 0 : 1);
                    t.mask = tile_mask(rt[r0 + r].first, rt[r0 + r].second, ct[c].first, ct[c].second, false);
                    jb.tiles.push_back(t);
                }
            emit(jb);
        }
        if (sym) {  // the triangle of this super row
            JobB jb;
            for (int r = 0; r < h; r++)
                for (int c = 0; c <= r; c++) {
                    TileB t;
                    t.rbox = jb.box(8 * rt[r0 + r].first, 0);
                    t.cbox = jb.box(8 * rt[r0 + c].first, 0);
                    t.mask = tile_mask(rt[r0 + r].first, rt[r0 + r].second, rt[r0 + c].first, rt[r0 + c].second, true);
                    if (t.mask) jb.tiles.push_back(t);
                }
            emit(jb);
        }
    }
}

}  // namespace

int atvb_launch(ithaca_b200* h, const double* A, long long lda, int Ma, const double* B, long long ldb, int Nb,
                long long K, const double* V, int dim, int symmetric, double* out, long long rs, int cinner,
                long long cso, long long csi)
{
    std::vector<GramJob> jobs;
    std::vector<int> cost;
    build_schedule(Ma, Nb, symmetric != 0, jobs, cost);
    const int njobs = (int)jobs.size();
    if (njobs == 0) return IB200_OK;
    int nbox_max = 0;
    for (auto& j : jobs) nbox_max = std::max(nbox_max, j.nbox);
    const int stage_bytes = nbox_max * GRAM_BOXB;
    const int smem_cap = 227 * 1024 - 1024;
    int nstages = std::min(8, (int)((smem_cap - (int)sizeof(GramJob) - 64) / (stage_bytes + GRAM_KT * 8 + 16)));
    if (nstages < 2) { set_error(h, "gram: job does not fit shared memory"); return IB200_ERR_ARG; }
    const size_t smem = (size_t)nstages * (stage_bytes + GRAM_KT * 8 + 16) + sizeof(GramJob) + 64;

    // split-K: every CTA should last equally long. Job j costs cost[j] per k-step on its critical sub-partition;
    // it gets nsplit_j ~ cost[j] * K / tau slabs, with tau chosen for about `waves` full waves of CTAs.
    const long long ksteps = (K + GRAM_KT - 1) / GRAM_KT;  // stages
    double total = 0.0;
    for (int c : cost) total += (double)c;
    const int waves = (njobs >= 8) ? 3 : (njobs > 1 ? 2 : 1);
    const double target_ctas = (double)waves * h->num_sms;
    std::vector<GramCta> ctas;
    long long ws_doubles = 0;
    const long long min_stages = 64;  // a slab is at least 64 stages long (pipeline fill, epilogue)
    for (int j = 0; j < njobs; j++) {
        long long ns = (long long)std::llround(target_ctas * cost[j] / total);
        ns = std::max(1LL, std::min(ns, std::max(1LL, ksteps / min_stages)));
        const long long per = (ksteps + ns - 1) / ns;
        ns = (ksteps + per - 1) / per;
        jobs[j].nsplit = (int)ns;
        jobs[j].ws_off = ws_doubles;
        ws_doubles += ns * GRAM_WARPS * GRAM_TILE;
        for (long long s = 0; s < ns; s++) {
            GramCta c;
            c.job = j;
            c.slot = (int)s;
            c.k_begin = s * per * GRAM_KT;
            c.k_end = std::min<long long>(K, (s + 1) * per * GRAM_KT);
            ctas.push_back(c);
        }
    }
    // longest CTAs first (they are nearly equal; this only orders the rounding leftovers)
    std::stable_sort(ctas.begin(), ctas.end(), [&](const GramCta& x, const GramCta& y) {
        return (double)cost[x.job] * (x.k_end - x.k_begin) > (double)cost[y.job] * (y.k_end - y.k_begin);
    });

    if (getenv("ITHACA_B200_DEBUG")) {
        double useful = 0.0, crit = 0.0, tmax = 0.0, tsum = 0.0;
        for (int j = 0; j < njobs; j++) {
            int fr = 0;
            for (int w = 0; w < GRAM_WARPS; w++) fr += __builtin_popcount(jobs[j].mask[w]);
            useful += fr / 4.0;
            crit += cost[j];
        }
        for (auto& c : ctas) {
            const double t = (double)cost[c.job] * (c.k_end - c.k_begin);
            tmax = std::max(tmax, t);
            tsum += t;
        }
        fprintf(stderr, "[gram] Ma=%d Nb=%d sym=%d jobs=%d ctas=%zu stages=%d sub-partition balance=%.3f cta max/mean=%.3f waves=%.2f\n",
                Ma, Nb, symmetric, njobs, ctas.size(), nstages, useful / crit, tmax / (tsum / ctas.size()),
                (double)ctas.size() / h->num_sms);
    }
    const size_t ws_need = (size_t)ws_doubles * sizeof(double);
    if (ws_need > h->dWs_cap) {
        if (h->dWs) cudaFree(h->dWs);
        h->dWs = nullptr; h->dWs_cap = 0;
        IB_CUDA(h, cudaMalloc(&h->dWs, ws_need));
        h->dWs_cap = ws_need;
    }
    const size_t jb_bytes = (size_t)njobs * sizeof(GramJob), ct_bytes = ctas.size() * sizeof(GramCta);
    const size_t jb_pad = (jb_bytes + 255) / 256 * 256;
    if (jb_pad + ct_bytes > h->dJobs_cap) {
        if (h->dJobs) cudaFree(h->dJobs);
        h->dJobs = nullptr; h->dJobs_cap = 0;
        IB_CUDA(h, cudaMalloc(&h->dJobs, jb_pad + ct_bytes));
        h->dJobs_cap = jb_pad + ct_bytes;
    }
    // the previous launch on this stream may still read dJobs: the copies below are stream-ordered after it
    // (pageable sources are staged before cudaMemcpyAsync returns, so the host vectors may go out of scope)
    IB_CUDA(h, cudaMemcpyAsync(h->dJobs, jobs.data(), jb_bytes, cudaMemcpyHostToDevice, h->stream));
    IB_CUDA(h, cudaMemcpyAsync((char*)h->dJobs + jb_pad, ctas.data(), ct_bytes, cudaMemcpyHostToDevice, h->stream));

    CUtensorMap tmA, tmB;
    int rc = make_tmap_2d(h, &tmA, A, (uint64_t)K, (uint64_t)Ma, (uint64_t)lda * 8, GRAM_KT, GRAM_BOXROWS, true);
    if (rc) return rc;
    rc = make_tmap_2d(h, &tmB, B, (uint64_t)K, (uint64_t)Nb, (uint64_t)ldb * 8, GRAM_KT, GRAM_BOXROWS, true);
    if (rc) return rc;

    IB_CUDA(h, cudaFuncSetAttribute(gram_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    Timer& tk = h->timers[T_GRAMK];
    IB_CUDA(h, cudaEventRecord(tk.a, h->stream));
    gram_kernel<<<(unsigned)ctas.size(), GRAM_THREADS, smem, h->stream>>>(
        tmA, tmB, V, dim, K, (const GramJob*)h->dJobs, (const GramCta*)((char*)h->dJobs + jb_pad), nstages, stage_bytes,
        h->dWs);
    IB_CUDA(h, cudaGetLastError());
    IB_CUDA(h, cudaEventRecord(tk.b, h->stream));
    tk.used = true;
    gram_reduce_kernel<<<dim3(GRAM_WARPS * GRAM_TILE / 256 / 4, njobs), 256, 0, h->stream>>>(
        h->dWs, (const GramJob*)h->dJobs, Ma, Nb, symmetric, out, rs, cinner, cso, csi);
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
