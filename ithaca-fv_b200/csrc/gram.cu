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
// fragments are read with a K permutation that makes every LDS.64 bank-conflict free. The 16 cell weights of a stage
// arrive with a bulk copy from a dim-times repeated copy of V (built once per volume vector). 15 MMA warps (4,4,4,3 per
// SM sub-partition) each own one <= 5x5-fragment output tile in registers; the tile shape is a template parameter
// (rectangular NR x NC, or rows [I0, I0+NR) of a lower-triangular diagonal tile), so the inner loop carries no masks.
// One lane of a 16th warp keeps the TMA/mbarrier ring full.
//
// Schedule (host): the output is cut into 5x5-fragment tiles; tiles are packed into jobs of <= 15 tiles sharing <= 10
// boxes (4x3 / 3x4 super-tiles off the diagonal, triangles on it, the whole triangle for m <= 200), heavy tiles are
// split by fragment rows when warps are spare, and tiles go to warps so that the four sub-partitions carry equal DMMA
// counts. The K range of all jobs is then laid end to end, weighted by each job's measured stage time, and cut into
// one equal share per SM (stream-K): a CTA runs one to a few (job, K slab) segments back to back through one ring.
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
constexpr int GRAM_MMA_WARPS = 15;               // MMA warps per CTA (warp w runs on SM sub-partition w & 3)
constexpr int GRAM_WARPS = GRAM_MMA_WARPS;       // warp tiles per job
constexpr int GRAM_THREADS = (GRAM_MMA_WARPS + 1) * 32;  // + 1 producer warp (on sub-partition 3)
constexpr int GRAM_TILE = GRAM_BOXROWS * GRAM_BOXROWS;  // doubles per warp tile in the workspace
constexpr int GRAM_PAT_IDLE = 255;
constexpr int GRAM_PAT_LOWER = 25;               // pattern ids >= 25: lower-triangular row ranges

struct GramJob {
    int nbox;
    int nsplit;                      // K slabs of this job
    long long ws_off;                // first double of this job's workspace: [nsplit][16 tiles][40x40]
    int boxrow[GRAM_NBMAX];          // first column of each box
    int boxmap[GRAM_NBMAX];          // 0 = operand A (tile rows), 1 = operand B (tile columns)
    unsigned char ta[GRAM_WARPS];    // per warp: box index of the tile's rows
    unsigned char tb[GRAM_WARPS];    // per warp: box index of the tile's columns
    unsigned char pat[GRAM_WARPS];   // per warp: tile shape. (NR-1)*5+(NC-1): NR x NC rectangle starting at fragment
                                     // row i0; 25 + lower_id(I0, NR): rows [I0, I0+NR) of a lower triangle; 255 idle
    unsigned char i0[GRAM_WARPS];    // per warp: first fragment row inside the box
    unsigned int mask[GRAM_WARPS];   // per warp: bit i*WF+j = fragment (i,j) is computed (for the reduction)
};

struct GramSeg {
    int job, slot;
    long long k_begin, k_end;        // cells; k_begin is a multiple of GRAM_KT, k_end is clipped to K
};

__host__ __device__ constexpr int gram_lower_id(int i0, int nr)
{
    // (I0, NR) with I0 + NR <= 5, enumerated I0-major: I0 = 0 has 5 entries, I0 = 1 has 4, ...
    return (i0 == 0 ? 0 : i0 == 1 ? 5 : i0 == 2 ? 9 : i0 == 3 ? 12 : 14) + (nr - 1);
}

// DMMA without `volatile`: the scheduler may interleave it with the shared-memory loads of the next k-step
__device__ __forceinline__ void gram_dmma(double& c0, double& c1, double a, double b)
{
    asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// ring position of a consumer warp
struct GramRing {
    int s, ph;
    int prev;   // stage this warp still has to hand back, -1: none (delayed release, common.cuh)
};

// One (job, K slab) segment of one MMA warp: NR x NC fragments (LI0 < 0; the runtime row offset is folded into a_off)
// or the rows [LI0, LI0+NR) of a lower-triangular tile (local row i has columns 0..LI0+i).
template <int NR, int NC, int LI0>
__device__ __forceinline__ void gram_run(GramRing& r, int niter, int nstages, int stage_bytes, const unsigned char* panels,
                                         const double* wsm, uint64_t* full, uint64_t* empty, uint32_t a_off, uint32_t b_off,
                                         int row0, double* __restrict__ out)
{
    constexpr int NCM = (LI0 < 0) ? NC : (LI0 + NR);
    const int lane = threadIdx.x & 31;
    const int g = lane >> 2, t = lane & 3;
    // K permutation inside a 16-cell stage: at step ks the thread with t = lane&3 takes k = 2*ks + (t&1) + 8*(t>>1).
    // With the 128-B swizzle the 16 lanes of a half-warp then hit 16 distinct 8-byte banks. The 16-B chunk of k in
    // row g is ((k>>1) ^ g) = ks ^ (4*(t>>1) ^ g), so the byte offset of step ks is kb ^ (ks << 4).
    const uint32_t kb = (uint32_t)(g * 128) + (uint32_t)((((4 * (t >> 1)) ^ g) << 4) | ((t & 1) << 3));
    const int kw = (t & 1) + 8 * (t >> 1);

    double acc[NR][NCM][2];
#pragma unroll
    for (int i = 0; i < NR; i++)
#pragma unroll
        for (int j = 0; j < NCM; j++) acc[i][j][0] = acc[i][j][1] = 0.0;

    int s = r.s, ph = r.ph, prev = r.prev;
    for (int it = 0; it < niter; it++) {
        mbar_wait(full + s, (uint32_t)ph);
        mbar_release_previous(empty, prev, lane);   // delayed release (common.cuh)
        const unsigned char* st = panels + (size_t)s * stage_bytes;
        const unsigned char* pa = st + a_off;
        const unsigned char* pb = st + b_off;
        const double* wv = wsm + s * GRAM_KT + kw;
#pragma unroll
        for (int ks = 0; ks < 4; ks++) {
            const uint32_t off = kb ^ (uint32_t)(ks << 4);
            const double w = wv[2 * ks];
            double b[NCM];
#pragma unroll
            for (int j = 0; j < NCM; j++) b[j] = *reinterpret_cast<const double*>(pb + j * 1024 + off);
#pragma unroll
            for (int i = 0; i < NR; i++) {
                const double a = *reinterpret_cast<const double*>(pa + i * 1024 + off) * w;
#pragma unroll
                for (int j = 0; j < NCM; j++)
                    if (LI0 < 0 || j <= LI0 + i) gram_dmma(acc[i][j][0], acc[i][j][1], a, b[j]);
            }
        }
        prev = s;
        if (++s == nstages) { s = 0; ph ^= 1; }
    }
    r.s = s;
    r.ph = ph;
    r.prev = prev;
    // partial tile -> workspace [40][40]
#pragma unroll
    for (int i = 0; i < NR; i++)
#pragma unroll
        for (int j = 0; j < NCM; j++)
            if (LI0 < 0 || j <= LI0 + i)
                *reinterpret_cast<double2*>(out + (8 * (row0 + i) + g) * GRAM_BOXROWS + 8 * j + 2 * t) =
                    make_double2(acc[i][j][0], acc[i][j][1]);
}

__global__ void __launch_bounds__(GRAM_THREADS, 1)
gram_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            const double* __restrict__ wexp, const GramJob* __restrict__ jobs, const GramSeg* __restrict__ segs,
            const int* __restrict__ seg_off, int maxseg, int nstages, int stage_bytes, double* __restrict__ ws,
            long long* __restrict__ dbg)
{
    extern __shared__ __align__(1024) unsigned char smem[];
    // [nstages][stage_bytes] panels | [nstages][16] weights | full[nstages] | empty[nstages] | segs[maxseg] | jobs[maxseg]
    unsigned char* const panels = smem;
    double* const wsm = reinterpret_cast<double*>(smem + (size_t)nstages * stage_bytes);
    uint64_t* const full = reinterpret_cast<uint64_t*>(wsm + nstages * GRAM_KT);
    uint64_t* const empty = full + nstages;
    GramSeg* const sseg = reinterpret_cast<GramSeg*>(empty + nstages);
    GramJob* const sjob = reinterpret_cast<GramJob*>(sseg + maxseg);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int s0 = seg_off[blockIdx.x], nseg = seg_off[blockIdx.x + 1] - s0;
    if (nseg == 0) return;

    for (int q = tid; q < nseg * (int)(sizeof(GramSeg) / 4); q += GRAM_THREADS)
        reinterpret_cast<int*>(sseg)[q] = reinterpret_cast<const int*>(segs + s0)[q];
    for (int sgi = 0; sgi < nseg; sgi++) {
        const int j = segs[s0 + sgi].job;
        for (int q = tid; q < (int)(sizeof(GramJob) / 4); q += GRAM_THREADS)
            reinterpret_cast<int*>(sjob + sgi)[q] = reinterpret_cast<const int*>(jobs + j)[q];
    }
    if (tid == 0) {
        for (int s = 0; s < nstages; s++) {
            mbar_init(full + s, 1);
            mbar_init(empty + s, GRAM_MMA_WARPS);
        }
        mbar_fence_init();
    }
    __syncthreads();

    if (warp == GRAM_MMA_WARPS) {
        // ---------------- producer warp: one lane walks the segments and keeps the ring full ----------------
        if (lane == 0) {
            tma_prefetch_desc(&tmA);
            tma_prefetch_desc(&tmB);
            int s = 0, ph = 0, git = 0;
            for (int sgi = 0; sgi < nseg; sgi++) {
                const GramJob* jb = sjob + sgi;
                const int nbox = jb->nbox;
                const long long k_end = sseg[sgi].k_end;
                for (long long k0 = sseg[sgi].k_begin; k0 < k_end; k0 += GRAM_KT, git++) {
                    if (git >= nstages) mbar_wait(empty + s, (uint32_t)(ph ^ 1));
                    uint64_t* bar = full + s;
                    mbar_arrive_expect_tx(bar, (uint32_t)(nbox * GRAM_BOXB + GRAM_KT * 8));
                    bulk_load_1d(wsm + s * GRAM_KT, wexp + k0, GRAM_KT * 8, bar);
                    unsigned char* dst = panels + (size_t)s * stage_bytes;
                    for (int b = 0; b < nbox; b++)
                        tma_load_2d(dst + b * GRAM_BOXB, jb->boxmap[b] ? &tmB : &tmA, (int)k0, jb->boxrow[b], bar);
                    if (++s == nstages) { s = 0; ph ^= 1; }
                }
            }
        }
        return;
    }

    // ---------------- MMA warps ----------------
    GramRing ring;
    ring.s = 0;
    ring.ph = 0;
    ring.prev = -1;
    for (int sgi = 0; sgi < nseg; sgi++) {
        const GramJob* jb = sjob + sgi;
        const int niter = (int)((sseg[sgi].k_end - sseg[sgi].k_begin + GRAM_KT - 1) / GRAM_KT);
        const int pat = jb->pat[warp], i0 = jb->i0[warp];
        long long t0 = 0;
        if (dbg) t0 = clock64();
        const uint32_t a_off = (uint32_t)(jb->ta[warp] * GRAM_BOXB + i0 * 1024);
        const uint32_t b_off = (uint32_t)(jb->tb[warp] * GRAM_BOXB);
        double* out = ws + jb->ws_off + ((size_t)sseg[sgi].slot * GRAM_MMA_WARPS + warp) * GRAM_TILE;
#define GRAM_ARGS ring, niter, nstages, stage_bytes, panels, wsm, full, empty, a_off, b_off
#define GRAM_RECT(NR, NC) case ((NR - 1) * 5 + (NC - 1)): gram_run<NR, NC, -1>(GRAM_ARGS, i0, out); break;
#define GRAM_RECT_ROW(NR) GRAM_RECT(NR, 1) GRAM_RECT(NR, 2) GRAM_RECT(NR, 3) GRAM_RECT(NR, 4) GRAM_RECT(NR, 5)
#define GRAM_LOW(I0, NR) case (GRAM_PAT_LOWER + gram_lower_id(I0, NR)): gram_run<NR, 5, I0>(GRAM_ARGS, I0, out); break;
        switch (pat) {
            GRAM_RECT_ROW(1) GRAM_RECT_ROW(2) GRAM_RECT_ROW(3) GRAM_RECT_ROW(4) GRAM_RECT_ROW(5)
            GRAM_LOW(0, 1) GRAM_LOW(0, 2) GRAM_LOW(0, 3) GRAM_LOW(0, 4) GRAM_LOW(0, 5)
            GRAM_LOW(1, 1) GRAM_LOW(1, 2) GRAM_LOW(1, 3) GRAM_LOW(1, 4)
            GRAM_LOW(2, 1) GRAM_LOW(2, 2) GRAM_LOW(2, 3)
            GRAM_LOW(3, 1) GRAM_LOW(3, 2)
            GRAM_LOW(4, 1)
        default:  // idle warp: keep the ring moving (it reads nothing from the stages: immediate release)
            mbar_release_previous(empty, ring.prev, lane);
            ring.prev = -1;
            for (int it = 0; it < niter; it++) {
                mbar_wait(full + ring.s, (uint32_t)ring.ph);
                __syncwarp();
                if (lane == 0) mbar_arrive(empty + ring.s);
                if (++ring.s == nstages) { ring.s = 0; ring.ph ^= 1; }
            }
            break;
        }
#undef GRAM_ARGS
#undef GRAM_RECT
#undef GRAM_RECT_ROW
#undef GRAM_LOW
        if (dbg && lane == 0) dbg[((size_t)(s0 + sgi) * GRAM_MMA_WARPS + warp)] = clock64() - t0;   // cycles of this warp in this segment
    }
}

// dim-times repeated weights, zero padded to a multiple of 16: wexp[k] = V[k / dim] (1 when V is null) for k < K
__global__ void gram_wexp_kernel(const double* __restrict__ V, int dim, long long K, long long Kpad, double* __restrict__ wexp)
{
    for (long long k = blockIdx.x * (long long)blockDim.x + threadIdx.x; k < Kpad; k += (long long)gridDim.x * blockDim.x)
        wexp[k] = (k < K) ? (V ? V[k / dim] : 1.0) : 0.0;
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
        // fixed summation order (reproducible run to run): four interleaved partial sums over the slabs keep four loads in
        // flight per thread instead of one dependent add chain
        const size_t sstr = (size_t)GRAM_WARPS * GRAM_TILE;
        const double* p = base + (size_t)w * GRAM_TILE + e;
        double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
        int k = 0;
        for (; k + 4 <= nsplit; k += 4) {
            s0 += p[(size_t)k * sstr]; s1 += p[(size_t)(k + 1) * sstr]; s2 += p[(size_t)(k + 2) * sstr]; s3 += p[(size_t)(k + 3) * sstr];
        }
        for (; k < nsplit; k++) s0 += p[(size_t)k * sstr];
        const double s = (s0 + s1) + (s2 + s3);
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
    int i0, nr, nc;      // fragment rows [i0, i0+nr); nc columns (rectangular)
    bool lower;          // rows of a lower-triangular diagonal tile: row i has columns 0..i
    int frags() const
    {
        if (!lower) return nr * nc;
        int n = 0;
        for (int i = i0; i < i0 + nr; i++) n += i + 1;
        return n;
    }
    unsigned mask() const
    {
        unsigned mk = 0;
        for (int i = i0; i < i0 + nr; i++)
            for (int j = 0; j < (lower ? i + 1 : nc); j++) mk |= 1u << (i * GRAM_WF + j);
        return mk;
    }
    int pat() const { return lower ? GRAM_PAT_LOWER + gram_lower_id(i0, nr) : (nr - 1) * 5 + (nc - 1); }
};

struct JobB {
    std::vector<int> boxrow, boxmap;
    std::vector<TileB> tiles;
    double cost = 0;     // estimated clocks per 16-cell stage (gram_stage_clocks)
    int box(int row, int map)
    {
        for (size_t b = 0; b < boxrow.size(); b++)
            if (boxrow[b] == row && boxmap[b] == map) return (int)b;
        boxrow.push_back(row);
        boxmap.push_back(map);
        return (int)boxrow.size() - 1;
    }
};

// Spread the tiles of a job over the 4 SM sub-partitions (MMA warp w runs on sub-partition w & 3; sub-partitions 0..2 host
// 4 MMA warps, sub-partition 3 hosts 3 and the producer warp): longest-processing-time rule, then single moves / swaps
// while they lower the larger of two loads. Returns the critical-path cost (largest DMMA count per k-step).
int pack_tiles(const std::vector<TileB>& T, std::vector<int> (&bins)[4])
{
    std::vector<int> order(T.size());
    for (size_t q = 0; q < T.size(); q++) order[q] = (int)q;
    std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return T[x].frags() > T[y].frags(); });
    const int cap[4] = {4, 4, 4, GRAM_MMA_WARPS - 12};
    int load[4] = {0, 0, 0, 0};
    for (int s = 0; s < 4; s++) bins[s].clear();
    for (int q : order) {
        int b = -1;
        for (int s = 3; s >= 0; s--)   // ties go to the 3-warp sub-partition first: it must take big tiles
            if ((int)bins[s].size() < cap[s] && (b < 0 || load[s] < load[b])) b = s;
        bins[b].push_back(q);
        load[b] += T[q].frags();
    }
    for (bool improved = true; improved;) {
        improved = false;
        for (int x = 0; x < 4 && !improved; x++)
            for (int y = 0; y < 4 && !improved; y++) {
                if (x == y || load[x] <= load[y]) continue;
                const int cur = load[x];
                for (size_t u = 0; u < bins[x].size() && !improved; u++) {
                    const int fu = T[bins[x][u]].frags();
                    if ((int)bins[y].size() < cap[y] && std::max(load[x] - fu, load[y] + fu) < cur) {
                        bins[y].push_back(bins[x][u]);
                        bins[x].erase(bins[x].begin() + u);
                        load[x] -= fu; load[y] += fu;
                        improved = true;
                        break;
                    }
                    for (size_t v = 0; v < bins[y].size(); v++) {
                        const int fv = T[bins[y][v]].frags();
                        if (fv < fu && std::max(load[x] - fu + fv, load[y] + fu - fv) < cur) {
                            std::swap(bins[x][u], bins[y][v]);
                            load[x] += fv - fu; load[y] += fu - fv;
                            improved = true;
                            break;
                        }
                    }
                }
            }
    }
    return std::max({load[0], load[1], load[2], load[3], 1});
}

// Clocks of one 16-cell stage of a job, fitted to per-segment cycle counts measured on B200 (ITHACA_B200_GRAM_TRACE):
// 510 fixed + 63.7 per DMMA fragment on the busiest sub-partition (= 4 k-steps x 15.9 clk, the DMMA issue rate), but
// never less than 104 clocks per fragment of the largest warp tile (one warp alone issues a DMMA every 26 clk).
double gram_stage_clocks(int crit, int maxwarp) { return std::max(510.0 + 63.7 * crit, 104.0 * maxwarp); }

// Split the heaviest tiles by fragment rows while warps are spare, keep the split whose packing has the lowest
// estimated stage time, and write the job descriptor.
void finish_job(JobB& jb, GramJob& out)
{
    std::vector<TileB> T = jb.tiles, bestT;
    std::vector<int> bins[4], bestbins[4];
    double bestcost = 1e300;
    for (;;) {
        int maxwarp = 1;
        for (const TileB& t : T) maxwarp = std::max(maxwarp, t.frags());
        const double c = gram_stage_clocks(pack_tiles(T, bins), maxwarp);
        if (c < bestcost) {
            bestcost = c;
            bestT = T;
            for (int s = 0; s < 4; s++) bestbins[s] = bins[s];
        }
        if ((int)T.size() >= GRAM_WARPS) break;
        int best = -1, bestf = 0;
        for (size_t q = 0; q < T.size(); q++)
            if (T[q].nr >= 2 && T[q].frags() > bestf) { best = (int)q; bestf = T[q].frags(); }
        if (best < 0 || bestf <= 5) break;
        // cut after the fragment row where the running count reaches half of the tile
        TileB lo = T[best], hi = T[best];
        int run = 0, cutrows = 0;
        for (int i = 0; i < T[best].nr; i++) {
            if (2 * run >= bestf) break;
            run += T[best].lower ? (T[best].i0 + i + 1) : T[best].nc;
            cutrows++;
        }
        if (cutrows <= 0 || cutrows >= T[best].nr) cutrows = T[best].nr / 2;
        lo.nr = cutrows;
        hi.i0 = T[best].i0 + cutrows;
        hi.nr = T[best].nr - cutrows;
        T[best] = lo;
        T.push_back(hi);
    }
    jb.cost = bestcost;
    std::memset(&out, 0, sizeof(out));
    std::memset(out.pat, GRAM_PAT_IDLE, sizeof(out.pat));
    out.nbox = (int)jb.boxrow.size();
    for (int b = 0; b < out.nbox; b++) { out.boxrow[b] = jb.boxrow[b]; out.boxmap[b] = jb.boxmap[b]; }
    for (int s = 0; s < 4; s++)
        for (size_t q = 0; q < bestbins[s].size(); q++) {
            const int w = s + 4 * (int)q;
            const TileB& t = bestT[bestbins[s][q]];
            out.ta[w] = (unsigned char)t.rbox;
            out.tb[w] = (unsigned char)t.cbox;
            out.pat[w] = (unsigned char)t.pat();
            out.i0[w] = (unsigned char)t.i0;
            out.mask[w] = t.mask();
        }
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

void build_schedule(int Ma, int Nb, bool sym, std::vector<GramJob>& jobs, std::vector<double>& cost)
{
    std::vector<std::pair<int, int>> rt, ct;
    cut((Ma + 7) / 8, sym, rt);
    if (sym) ct = rt;
    else cut((Nb + 7) / 8, false, ct);
    auto emit = [&](JobB& jb) {
        if (jb.tiles.empty()) return;
        GramJob g;
        finish_job(jb, g);
        jobs.push_back(g);
        cost.push_back(jb.cost);
    };
    // the whole lower triangle in one job when it fits 16 warps and 10 boxes (m <= 200: every SM then runs the same
    // perfectly balanced job on its own K slab)
    if (sym && (int)rt.size() <= 5) {
        JobB jb;
        for (size_t r = 0; r < rt.size(); r++)
            for (size_t c = 0; c <= r; c++) {
                TileB t;
                t.rbox = jb.box(8 * rt[r].first, 0);
                t.cbox = jb.box(8 * rt[c].first, 0);
                t.i0 = 0; t.nr = rt[r].second; t.nc = rt[c].second; t.lower = (c == r);
                jb.tiles.push_back(t);
            }
        if ((int)jb.tiles.size() <= GRAM_WARPS) { emit(jb); return; }
    }
    // super rows of 4 or 3 tile rows (never a 1- or 2-row remainder once there are >= 6 tile rows): 4 x 3 and 3 x 4
    // rectangles both fill 12 warps = 3 tiles on each sub-partition, the best fit for 15 MMA warps
    std::vector<int> groups;
    {
        const int T = (int)rt.size();
        int n3 = (T >= 6) ? ((4 - T % 4) % 4) : 0;      // T = 4a + 3b with the fewest 3-groups
        if (3 * n3 > T) n3 = 0;
        int left = T - 3 * n3;
        while (left >= 4 && (left % 4 == 0 || left > 4)) { groups.push_back(4); left -= 4; }
        if (left > 0) groups.push_back(left);
        for (int q = 0; q < n3; q++) groups.push_back(3);
    }
    size_t r0 = 0;
    for (size_t gi = 0; gi < groups.size(); r0 += groups[gi], gi++) {
        const int h = groups[gi];
        const size_t cend = sym ? r0 : ct.size();          // rectangular part: tile columns [0, cend)
        const int wc = (h == 4) ? 3 : (h == 3) ? 4 : std::max(1, std::min(GRAM_WARPS / h, GRAM_NBMAX - h));
        for (size_t c0 = 0; c0 < cend; c0 += wc) {
            JobB jb;
            const size_t c1 = std::min(cend, c0 + wc);
            for (int r = 0; r < h; r++)
                for (size_t c = c0; c < c1; c++) {
                    TileB t;
                    t.rbox = jb.box(8 * rt[r0 + r].first, 0);
                    t.cbox = jb.box(8 * ct[c].first, sym ? 0 : 1);
                    t.i0 = 0; t.nr = rt[r0 + r].second; t.nc = ct[c].second; t.lower = false;
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
                    t.i0 = 0; t.nr = rt[r0 + r].second; t.nc = rt[r0 + c].second; t.lower = (c == r);
                    jb.tiles.push_back(t);
                }
            emit(jb);
        }
    }
}

// stream-K: CTA boundaries as global stage positions j*ksteps + x (job j, stage x), cut at equal shares of the
// cost-weighted work and snapped away from slab ends shorter than minlen; every CTA then gets the (job, K slab)
// segments between two boundaries.
void gram_partition(const std::vector<double>& cost, long long ksteps, long long K, long long nctas, long long minlen,
                    std::vector<GramSeg>& segs, std::vector<int>& seg_off, std::vector<int>& nsplit)
{
    const int njobs = (int)cost.size();
    segs.clear();
    seg_off.assign(1, 0);
    nsplit.assign(njobs, 0);
    std::vector<double> prefix(njobs + 1, 0.0);
    for (int j = 0; j < njobs; j++) prefix[j + 1] = prefix[j] + (double)cost[j] * (double)ksteps;
    const double total = prefix[njobs];
    std::vector<long long> bound(nctas + 1, 0);
    bound[nctas] = (long long)njobs * ksteps;
    int j = 0;
    for (long long c = 1; c < nctas; c++) {
        const double goal = total * (double)c / (double)nctas;
        while (j + 1 < njobs && prefix[j + 1] <= goal) j++;
        long long x = (long long)std::llround((goal - prefix[j]) / cost[j]);
        x = std::max(0LL, std::min(x, ksteps));
        if (x < minlen) x = 0;
        else if (ksteps - x < minlen) x = ksteps;
        bound[c] = std::max(bound[c - 1], (long long)j * ksteps + x);
    }
    for (long long c = 0; c < nctas; c++) {
        long long pos = bound[c];
        while (pos < bound[c + 1]) {
            const int jj = (int)(pos / ksteps);
            const long long x0 = pos - (long long)jj * ksteps;
            const long long x1 = std::min(ksteps, x0 + (bound[c + 1] - pos));
            GramSeg sg;
            sg.job = jj;
            sg.slot = nsplit[jj]++;
            sg.k_begin = x0 * GRAM_KT;
            sg.k_end = std::min<long long>(K, x1 * GRAM_KT);
            segs.push_back(sg);
            pos += x1 - x0;
        }
        seg_off.push_back((int)segs.size());
    }
}

// Type-balanced static partition (round 2). The stage-time model above is off by -9 .. +7.5 % for the minority job shapes
// (measured per segment with ITHACA_B200_GRAM_TRACE at m = 1000: CTA cycles max / mean = 1.08 under gram_partition, i.e. the
// slowest CTA set the kernel time 8 % above the mean). Instead of trusting a model ACROSS shapes, jobs are grouped by their exact
// composition (boxes + fragments per warp: jobs of one type cost the same per stage whatever that cost is), the concatenated
// stage range of every type is cut into equal slices, and CTA c gets slice c of EVERY type: all CTAs run the same mix, so the
// balance no longer depends on the model. Still static (fixed slab order in the reduction: bit-reproducible run to run).
// A type too small for nctas slices of minlen stages gets fewer slices, spread over the CTAs.
void gram_partition_typed(const std::vector<int>& type, int ntypes, long long ksteps, long long K, long long nctas, long long minlen,
                          std::vector<GramSeg>& segs, std::vector<int>& seg_off, std::vector<int>& nsplit)
{
    const int njobs = (int)type.size();
    std::vector<std::vector<GramSeg>> per(nctas);
    for (int t = 0; t < ntypes; t++) {
        std::vector<int> js;
        for (int j = 0; j < njobs; j++) if (type[j] == t) js.push_back(j);
        if (js.empty()) continue;
        const long long L = (long long)js.size() * ksteps;
        const long long nt = std::max<long long>(1, std::min<long long>(nctas, L / std::max<long long>(minlen, 1)));
        std::vector<long long> b(nt + 1, 0);
        b[nt] = L;
        for (long long i = 1; i < nt; i++) {
            long long pos = (long long)std::llround((double)L * (double)i / (double)nt);
            const long long jj = std::min<long long>(pos / ksteps, (long long)js.size() - 1);
            long long x = pos - jj * ksteps;
            if (x < minlen) x = 0;
            else if (ksteps - x < minlen) x = ksteps;
            b[i] = std::max(b[i - 1], jj * ksteps + x);
        }
        for (long long i = 0; i < nt; i++) {
            const long long cta = ((i * nctas) / nt + 37LL * t) % nctas;   // small types do not all land on the same CTAs
            long long pos = b[i];
            while (pos < b[i + 1]) {
                const long long jj = pos / ksteps;
                const long long x0 = pos - jj * ksteps;
                const long long x1 = std::min(ksteps, x0 + (b[i + 1] - pos));
                GramSeg sg;
                sg.job = js[(size_t)jj];
                sg.slot = 0;
                sg.k_begin = x0 * GRAM_KT;
                sg.k_end = std::min<long long>(K, x1 * GRAM_KT);
                per[(size_t)cta].push_back(sg);
                pos += x1 - x0;
            }
        }
    }
    segs.clear();
    seg_off.assign(1, 0);
    nsplit.assign(njobs, 0);
    for (long long c = 0; c < nctas; c++) {
        for (GramSeg sg : per[(size_t)c]) {
            sg.slot = nsplit[sg.job]++;
            segs.push_back(sg);
        }
        seg_off.push_back((int)segs.size());
    }
}

}  // namespace

// cached dim-times repeated weights (two slots: the projection uses V with dim 3 and dim 1 alternately)
static int get_wexp(ithaca_b200* h, const double* V, int dim, long long K, const double** out)
{
    const long long Kpad = (K + GRAM_KT - 1) / GRAM_KT * GRAM_KT;
    for (int q = 0; q < 2; q++)
        if (h->wexp[q].p && h->wexp_V[q] == V && h->wexp_dim[q] == dim && h->wexp_K[q] == K && h->wexp_gen[q] == h->vol_gen) {
            *out = h->wexp[q].as<double>();
            return IB200_OK;
        }
    const int q = (h->wexp_next++) & 1;
    int rc = dev_ensure(h, h->wexp[q], (size_t)Kpad * sizeof(double));
    if (rc) return rc;
    gram_wexp_kernel<<<(unsigned)std::min<long long>((Kpad + 255) / 256, 8LL * h->num_sms), 256, 0, h->stream>>>(
        V, dim, K, Kpad, h->wexp[q].as<double>());
    IB_CUDA(h, cudaGetLastError());
    h->launches += 1;
    h->wexp_V[q] = V; h->wexp_dim[q] = dim; h->wexp_K[q] = K; h->wexp_gen[q] = h->vol_gen;
    *out = h->wexp[q].as<double>();
    return IB200_OK;
}

int atvb_launch(ithaca_b200* h, const double* A, long long lda, int Ma, const double* B, long long ldb, int Nb,
                long long K, const double* V, int dim, int symmetric, double* out, long long rs, int cinner,
                long long cso, long long csi)
{
    std::vector<GramJob> jobs;
    std::vector<double> cost;
    build_schedule(Ma, Nb, symmetric != 0, jobs, cost);
    const int njobs = (int)jobs.size();
    if (njobs == 0 || K <= 0) return IB200_OK;
    int nbox_max = 0;
    for (auto& j : jobs) nbox_max = std::max(nbox_max, j.nbox);
    const int stage_bytes = nbox_max * GRAM_BOXB;

    // stream-K: lay the jobs' K ranges end to end, weighted by cost, and give every CTA an equal share
    const long long ksteps = (K + GRAM_KT - 1) / GRAM_KT;  // stages per job
    double total = 0.0;
    for (double c : cost) total += c * (double)ksteps;
    const long long minlen = std::min<long long>(64, ksteps);     // shortest slab worth its pipeline fill and epilogue
    long long nctas = std::min<long long>(h->num_sms, std::max<long long>(1, (long long)njobs * ksteps / 128));
    if (const char* e = env_knob("ITHACA_B200_GRAM_CTAS")) nctas = std::max(1, atoi(e));
    std::vector<GramSeg> segs;
    std::vector<int> seg_off, nsplit;
    // job types = exact compositions (boxes, per-warp pattern / fragment count): see gram_partition_typed
    std::vector<int> type(njobs, 0);
    int ntypes = 0;
    {
        std::vector<std::vector<int>> keys;
        for (int j = 0; j < njobs; j++) {
            std::vector<int> key(1, jobs[j].nbox);
            for (int w = 0; w < GRAM_WARPS; w++) { key.push_back(jobs[j].pat[w]); key.push_back(__builtin_popcount(jobs[j].mask[w])); }
            size_t t = 0;
            while (t < keys.size() && keys[t] != key) t++;
            if (t == keys.size()) keys.push_back(key);
            type[j] = (int)t;
        }
        ntypes = (int)keys.size();
    }
    const char* part = env_knob("ITHACA_B200_GRAM_PART");   // "model": the cost-model stream-K of round 1 (A/B measurements)
    if (part && strcmp(part, "model") == 0) gram_partition(cost, ksteps, K, nctas, minlen, segs, seg_off, nsplit);
    else gram_partition_typed(type, ntypes, ksteps, K, nctas, minlen, segs, seg_off, nsplit);
    int maxseg = 1;
    for (size_t c = 0; c + 1 < seg_off.size(); c++) maxseg = std::max(maxseg, seg_off[c + 1] - seg_off[c]);
    long long ws_doubles = 0;
    for (int j = 0; j < njobs; j++) {
        jobs[j].nsplit = nsplit[j];
        jobs[j].ws_off = ws_doubles;
        ws_doubles += (long long)nsplit[j] * GRAM_WARPS * GRAM_TILE;
    }

    const int fixed = maxseg * (int)(sizeof(GramSeg) + sizeof(GramJob)) + 64;
    const int smem_cap = 227 * 1024 - 1024;
    // Ring depth: a stage of the big shapes lasts ~2.8 us (1500 DMMAs), far longer than a TMA round trip, so depth buys nothing
    // there -- and shared memory beyond ~160 KB costs the L1 the job tables, weights and workspace stores go through. Measured
    // (scripts/gram_sweep.py, 1M cells x 3 x m = 1000): 6 stages (216 KB) 96.2 ms, 5 / 4 / 3 stages 92.6 ms; m = 2000: 190.9 ->
    // 181.7 ms; m = 200 at C2: 2.91 -> 2.73 ms. Small stages (m = 50: 10 KB, 0.26 ms at >= 4 stages, 0.36 ms at 3) keep up to 8.
    const int per_stage = stage_bytes + GRAM_KT * 8 + 16;
    int nstages = std::min((smem_cap - fixed) / per_stage, std::max(4, std::min(8, 160 * 1024 / per_stage)));
    if (const char* e = env_knob("ITHACA_B200_GRAM_STAGES")) nstages = std::max(2, std::min(nstages, atoi(e)));
    if (nstages < 2) { set_error(h, "gram: job does not fit shared memory"); return IB200_ERR_ARG; }
    const size_t smem = (size_t)nstages * (stage_bytes + GRAM_KT * 8 + 16) + fixed;

    if (env_knob("ITHACA_B200_DEBUG")) {
        double useful = 0.0, crit = 0.0, tmax = 0.0, tsum = 0.0;
        for (int j = 0; j < njobs; j++) {
            int ld[4] = {0, 0, 0, 0};
            for (int w = 0; w < GRAM_WARPS; w++) ld[w & 3] += __builtin_popcount(jobs[j].mask[w]);
            useful += (ld[0] + ld[1] + ld[2] + ld[3]) / 4.0;
            crit += std::max({ld[0], ld[1], ld[2], ld[3]});
        }
        for (size_t c = 0; c + 1 < seg_off.size(); c++) {
            double t = 0.0;
            for (int q = seg_off[c]; q < seg_off[c + 1]; q++)
                t += (double)cost[segs[q].job] * (double)((segs[q].k_end - segs[q].k_begin + GRAM_KT - 1) / GRAM_KT);
            tmax = std::max(tmax, t);
            tsum += t;
        }
        fprintf(stderr, "[gram] Ma=%d Nb=%d sym=%d K=%lld jobs=%d ctas=%lld segs=%zu maxseg=%d stages=%d sub-partition balance=%.3f "
                        "cta max/mean=%.3f\n",
                Ma, Nb, symmetric, K, njobs, nctas, segs.size(), maxseg, nstages, useful / crit, tmax / (tsum / nctas));
    }
    const size_t ws_need = (size_t)ws_doubles * sizeof(double);
    if (ws_need > h->dWs_cap) {
        if (h->dWs) cudaFree(h->dWs);
        h->dWs = nullptr; h->dWs_cap = 0;
        IB_CUDA(h, cudaMalloc(&h->dWs, ws_need));
        h->dWs_cap = ws_need;
    }
    const size_t jb_bytes = (size_t)njobs * sizeof(GramJob), sg_bytes = segs.size() * sizeof(GramSeg),
                 so_bytes = seg_off.size() * sizeof(int);
    const size_t jb_pad = (jb_bytes + 255) / 256 * 256, sg_pad = (sg_bytes + 255) / 256 * 256;
    if (jb_pad + sg_pad + so_bytes > h->dJobs_cap) {
        if (h->dJobs) cudaFree(h->dJobs);
        h->dJobs = nullptr; h->dJobs_cap = 0;
        IB_CUDA(h, cudaMalloc(&h->dJobs, jb_pad + sg_pad + so_bytes));
        h->dJobs_cap = jb_pad + sg_pad + so_bytes;
    }
    // the previous launch on this stream may still read dJobs: the copies below are stream-ordered after it
    // (pageable sources are staged before cudaMemcpyAsync returns, so the host vectors may go out of scope)
    char* dj = (char*)h->dJobs;
    IB_CUDA(h, cudaMemcpyAsync(dj, jobs.data(), jb_bytes, cudaMemcpyHostToDevice, h->stream));
    IB_CUDA(h, cudaMemcpyAsync(dj + jb_pad, segs.data(), sg_bytes, cudaMemcpyHostToDevice, h->stream));
    IB_CUDA(h, cudaMemcpyAsync(dj + jb_pad + sg_pad, seg_off.data(), so_bytes, cudaMemcpyHostToDevice, h->stream));

    const double* wexp = nullptr;
    int rc = get_wexp(h, V, dim, K, &wexp);
    if (rc) return rc;

    CUtensorMap tmA, tmB;
    rc = make_tmap_2d(h, &tmA, A, (uint64_t)K, (uint64_t)Ma, (uint64_t)lda * 8, GRAM_KT, GRAM_BOXROWS, true);
    if (rc) return rc;
    rc = make_tmap_2d(h, &tmB, B, (uint64_t)K, (uint64_t)Nb, (uint64_t)ldb * 8, GRAM_KT, GRAM_BOXROWS, true);
    if (rc) return rc;

    long long* dbg = nullptr;   // ITHACA_B200_GRAM_TRACE=file: per (segment, warp) cycle counts, for calibrating the cost model
    const char* trace = env_knob("ITHACA_B200_GRAM_TRACE");
    if (trace) IB_CUDA(h, cudaMalloc(&dbg, segs.size() * GRAM_MMA_WARPS * sizeof(long long)));
    IB_CUDA(h, cudaFuncSetAttribute(gram_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    Timer& tk = h->timers[T_GRAMK];
    IB_CUDA(h, cudaEventRecord(tk.a, h->stream));
    gram_kernel<<<(unsigned)nctas, GRAM_THREADS, smem, h->stream>>>(
        tmA, tmB, wexp, (const GramJob*)dj, (const GramSeg*)(dj + jb_pad), (const int*)(dj + jb_pad + sg_pad), maxseg, nstages,
        stage_bytes, h->dWs, dbg);
    IB_CUDA(h, cudaGetLastError());
    IB_CUDA(h, cudaEventRecord(tk.b, h->stream));
    tk.used = true;
    gram_reduce_kernel<<<dim3((GRAM_WARPS * GRAM_TILE + 255) / 256, njobs), 256, 0, h->stream>>>(
        h->dWs, (const GramJob*)dj, Ma, Nb, symmetric, out, rs, cinner, cso, csi);
    IB_CUDA(h, cudaGetLastError());
    h->launches += 2;
    if (trace) {
        std::vector<long long> cyc(segs.size() * GRAM_MMA_WARPS);
        IB_CUDA(h, cudaMemcpyAsync(cyc.data(), dbg, cyc.size() * sizeof(long long), cudaMemcpyDeviceToHost, h->stream));
        IB_CUDA(h, cudaStreamSynchronize(h->stream));
        cudaFree(dbg);
        if (FILE* f = fopen(trace, "w")) {
            fprintf(f, "cta seg job cost nbox stages warp_cycles_max frags_per_warp...\n");
            for (size_t c = 0; c + 1 < seg_off.size(); c++)
                for (int q = seg_off[c]; q < seg_off[c + 1]; q++) {
                    long long mx = 0;
                    for (int w = 0; w < GRAM_MMA_WARPS; w++) mx = std::max(mx, cyc[(size_t)q * GRAM_MMA_WARPS + w]);
                    const GramJob& jb = jobs[segs[q].job];
                    fprintf(f, "%zu %d %d %.0f %d %lld %lld", c, q, segs[q].job, cost[segs[q].job], jb.nbox,
                            (segs[q].k_end - segs[q].k_begin + GRAM_KT - 1) / GRAM_KT, mx);
                    for (int w = 0; w < GRAM_MMA_WARPS; w++) fprintf(f, " %d", __builtin_popcount(jb.mask[w]));
                    fprintf(f, "\n");
                }
            fclose(f);
        }
    }
    return IB200_OK;
}

// POD Gram: G = S^T diag(V) S, symmetric schedule, column-major m x m into h->dG.
int gram_launch(ithaca_b200* h)
{
    return atvb_launch(h, h->dS, h->ld, h->m, h->dS, h->ld, h->m, h->Nd, h->dV, h->dim, 1, h->dG, 1, 0x7fffffff, 0, h->m);
}

}  // namespace ib200
