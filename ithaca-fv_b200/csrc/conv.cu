// K6f/K9f -- fused convective tensor  C(i,j,k) = int U_i . div(interpolate(U_j) & Sf, U_k)   (FP64, DMMA).
//
// Replaces steadyNS::convective_term_tens / _cache (src/ITHACA_FOMPROBLEMS/steadyNS/steadyNS.C:997-1083), whose loop
// nest evaluates N^3 (resp. N^2) full-mesh div fields and N^3 domainIntegrate reductions.
//
// Reformulation (exact in real arithmetic, summation order differs from the reference at the 1e-16 level):
//   sum_c V_c U_i(c) . (1/V_c) sum_{f in c} +-phi_j(f) U_k,f(f)  =  sum_f a_j(f) * ( b_k(f) . d_i(f) )
//   a_j(f) = (w U_j[o] + (1-w) U_j[n]) . Sf        the face flux of mode j          (boundary: Ub_j . Sf, 0 on `empty`)
//   b_k(f) =  w U_k[o] + (1-w) U_k[n]              the face value of mode k         (boundary: Ub_k)
//   d_i(f) =  U_i[o] - U_i[n]                      the scatter of surfaceIntegrate  (boundary: U_i[o])
// so the face->cell scatter disappears: every face contributes a rank-one term, and C is ONE contraction over faces
//   C[(k,i), j] = sum_f E[f,(k,i)] * a[f,j],   E[f,(k,i)] = b_k(f) . d_i(f)  (generated in registers, never stored).
// Two kernels: conv_face_prep_kernel gathers both cells of every face once for all N modes (HBM-bound, coalesced over
// the mode index thanks to a cell-major copy of the basis) and writes a, b, d per face; conv_contract_kernel streams
// those per-face records through a TMA/bulk-copy mbarrier ring and contracts them on the FP64 tensor cores.
#include "../../include/ithaca_b200.h"
#include "common.cuh"
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace ib200 {

constexpr int CV_KTMAX = 32;   // faces per pipeline stage: template parameter KT = 16 or 32 (4 / 8 DMMA k-steps)
constexpr int CV_ROWS = 240;   // output rows (k,i) per CTA
constexpr int CV_KC = 7;       // distinct k per CTA when the test fields are the basis (NI = N): the b box is 4*CV_KC doubles per
                               // face (pitch 28 = 12 mod 16). With fewer test fields (G: NI = Np) a chunk takes up to CV_KCMAX k's
constexpr int CV_KCMAX = 25;   // (odd: 4*kc = 4 or 12 mod 16 keeps the b reads of the four k-lanes on distinct banks)
constexpr int CV_WARPS = 15;   // MMA warps, two row fragments (16 rows) each; + 1 producer warp = 512 threads
constexpr int CV_THREADS = (CV_WARPS + 1) * 32;
constexpr int CV_JF = 7;       // column fragments (8 j each) per CTA

// ---- cell-major copy of the basis: out[c][comp][j] (pitch NP per component) from N field-major columns ----------
// in: column j at in + j*ld, `rows` doubles each (3 per cell / boundary face). A (N x 96) tile is transposed in smem.
__global__ void __launch_bounds__(256)
conv_transpose_kernel(const double* __restrict__ in, long long ld, long long rows, int N, int NP, double* __restrict__ out)
{
    __shared__ double tile[96][33];
    const long long r0 = (long long)blockIdx.x * 96;
    for (int j0 = 0; j0 < NP; j0 += 32) {
        for (int idx = threadIdx.x; idx < 32 * 96; idx += 256) {
            const int jj = idx / 96, r = idx % 96, j = j0 + jj;
            tile[r][jj] = (j < N && r0 + r < rows) ? in[(size_t)j * ld + r0 + r] : 0.0;
        }
        __syncthreads();
        for (int idx = threadIdx.x; idx < 96 * 32; idx += 256) {
            const int r = idx / 32, jj = idx % 32, j = j0 + jj;
            if (j < NP && r0 + r < rows) out[(size_t)(r0 + r) * NP + j] = tile[r][jj];
        }
        __syncthreads();
    }
}

// ---- per-face records for all N modes ---------------------------------------------------------------------------
// Aj[f][NP], Bk[f][N][4] (component padded to 4: 32 B per mode, TMA-box friendly), Dd[f][3][NP]. One block = 4 faces,
// lanes over the mode index (coalesced 8*NP-byte rows of the cell-major basis).
template <bool UPW>   // UPW: "Gauss upwind" face value, two records per internal face (ConvMesh::upwind)
__global__ void __launch_bounds__(256)
conv_face_prep_kernel(ConvMesh m, const double* __restrict__ Ucm, const double* __restrict__ Ubcm, const double* __restrict__ Dcm,
                      int N, int NP, double* __restrict__ Aj, double* __restrict__ Bk, double* __restrict__ Dd)
{
    // Dcm: cell-major test fields of the d_i records (null = the basis itself: the convective tensor; grad(P_i) padded
    // to N fields: the divMomentum tensor)
    const long long f = (long long)blockIdx.x * 4 + threadIdx.y;
    const long long nF = m.nInt + m.nBnd;
    if (f >= nF) return;
    const double s0 = m.Sf[3 * f], s1 = m.Sf[3 * f + 1], s2 = m.Sf[3 * f + 2];
    const int o = m.owner[f];
    const double* uo = Ucm + (size_t)o * 3 * NP;
    const double mo = (m.cmask && !m.cmask[o]) ? 0.0 : 1.0;   // halo cells of a shard take no part in the integral
    if (f < m.nInt) {
        const int nn = m.neigh[f];
        const double* un = Ucm + (size_t)nn * 3 * NP;
        const double mn = (m.cmask && !m.cmask[nn]) ? 0.0 : 1.0;
        const double w = m.w[f], w1 = 1.0 - w;
        for (int j = threadIdx.x; j < NP; j += 64) {
            const double o0 = uo[j], o1 = uo[NP + j], o2 = uo[2 * NP + j];
            const double n0 = un[j], n1 = un[NP + j], n2 = un[2 * NP + j];
            const double b0 = w * o0 + w1 * n0, b1 = w * o1 + w1 * n1, b2 = w * o2 + w1 * n2;
            double a = 0.0;
            a += b0 * s0; a += b1 * s1; a += b2 * s2;   // the flux is linearInterpolate(U_j) & Sf in every scheme ...
            if (m.flux) a = j < N ? m.flux[(size_t)j * m.ldF + f] : 0.0;   // ... unless the caller supplied L_PHImodes[j] ("consistent")
            if (!UPW) {
                Aj[(size_t)f * NP + j] = a;
                double* d = Dd + (size_t)f * 3 * NP;
                if (Dcm) {
                    const double* qo = Dcm + (size_t)o * 3 * NP;
                    const double* qn = Dcm + (size_t)nn * 3 * NP;
                    d[j] = mo * qo[j] - mn * qn[j]; d[NP + j] = mo * qo[NP + j] - mn * qn[NP + j];
                    d[2 * NP + j] = mo * qo[2 * NP + j] - mn * qn[2 * NP + j];
                } else {
                    d[j] = mo * o0 - mn * n0; d[NP + j] = mo * o1 - mn * n1; d[2 * NP + j] = mo * o2 - mn * n2;
                }
                if (j < N) {
                    double4* b = reinterpret_cast<double4*>(Bk + ((size_t)f * N + j) * 4);
                    *b = make_double4(b0, b1, b2, 0.0);
                }
            } else {
                // upwind: phi_j >= 0 carries the owner's U_k, phi_j < 0 the neighbour's -> two records of the same face
                double d0, d1, d2;
                if (Dcm) {
                    const double* qo = Dcm + (size_t)o * 3 * NP;
                    const double* qn = Dcm + (size_t)nn * 3 * NP;
                    d0 = mo * qo[j] - mn * qn[j]; d1 = mo * qo[NP + j] - mn * qn[NP + j];
                    d2 = mo * qo[2 * NP + j] - mn * qn[2 * NP + j];
                } else {
                    d0 = mo * o0 - mn * n0; d1 = mo * o1 - mn * n1; d2 = mo * o2 - mn * n2;
                }
                const long long f2 = nF + f;
                Aj[(size_t)f * NP + j] = a >= 0.0 ? a : 0.0;
                Aj[(size_t)f2 * NP + j] = a >= 0.0 ? 0.0 : a;
                double* d = Dd + (size_t)f * 3 * NP;
                double* dd = Dd + (size_t)f2 * 3 * NP;
                d[j] = d0; d[NP + j] = d1; d[2 * NP + j] = d2;
                dd[j] = d0; dd[NP + j] = d1; dd[2 * NP + j] = d2;
                if (j < N) {
                    *reinterpret_cast<double4*>(Bk + ((size_t)f * N + j) * 4) = make_double4(o0, o1, o2, 0.0);
                    *reinterpret_cast<double4*>(Bk + ((size_t)f2 * N + j) * 4) = make_double4(n0, n1, n2, 0.0);
                }
            }
        }
    } else {
        const long long bf = f - m.nInt;
        const double* ub = Ubcm + (size_t)bf * 3 * NP;
        const bool on = m.bmask[bf] != 0;
        for (int j = threadIdx.x; j < NP; j += 64) {
            const double b0 = ub[j], b1 = ub[NP + j], b2 = ub[2 * NP + j];
            double a = 0.0;
            if (on) { a += b0 * s0; a += b1 * s1; a += b2 * s2; }
            if (m.flux) a = (on && j < N) ? m.flux[(size_t)j * m.ldF + f] : 0.0;
            Aj[(size_t)f * NP + j] = a;
            double* d = Dd + (size_t)f * 3 * NP;
            const double* qo = Dcm ? Dcm + (size_t)o * 3 * NP : uo;
            d[j] = mo * qo[j]; d[NP + j] = mo * qo[NP + j]; d[2 * NP + j] = mo * qo[2 * NP + j];
            if (j < N) {
                double4* b = reinterpret_cast<double4*>(Bk + ((size_t)f * N + j) * 4);
                *b = make_double4(b0, b1, b2, 0.0);
            }
        }
    }
}

// ---- contraction over faces -------------------------------------------------------------------------------------
struct ConvChunk { int r_lo, r_hi, k_lo, pad; };  // output rows [r_lo, r_hi) of the flattened (k,i) index, first k
// one CTA = (row chunk, column block, K slab [t_begin, t_end) in stages of KT faces). Chunks with fewer rows get fewer slabs
// (round 2): the slab counts follow the chunks' cost so that all SMs finish together.
struct ConvCta { int chunk, jblk, slab, pad; long long t_begin, t_end; };

// grid: one CTA per ConvCta descriptor. Stage layout in smem: [A: KT x NP][B: KT x bp][D: KT x 3NP].
// DMMA without `volatile` (see gram.cu)
__device__ __forceinline__ void conv_dmma(double& c0, double& c1, double a, double b)
{
    asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// one pipeline stage of one MMA warp with NR active row fragments: E = b_k . d_i in registers, 7 DMMAs per fragment row
template <int KT, int NR>
__device__ __forceinline__ void conv_stage(double (&acc)[2][CV_JF][2], const double* As, const double* Bs, const double* Ds, int NP,
                                           int bp, int j0g, int t, const int (&boff)[2], const int (&doff)[2])
{
#pragma unroll
    for (int ks = 0; ks < KT / 4; ks++) {
        const int fa = 4 * ks + t;
        const double* ap = As + fa * NP + j0g;
        double bj[CV_JF];
#pragma unroll
        for (int jf = 0; jf < CV_JF; jf++) bj[jf] = ap[8 * jf];
#pragma unroll
        for (int r = 0; r < NR; r++) {
            const double* bq = Bs + fa * bp + boff[r];
            const double* dq = Ds + fa * 3 * NP + doff[r];
            const double e = bq[0] * dq[0] + bq[1] * dq[NP] + bq[2] * dq[2 * NP];
#pragma unroll
            for (int jf = 0; jf < CV_JF; jf++) conv_dmma(acc[r][jf][0], acc[r][jf][1], e, bj[jf]);
        }
    }
}

template <int KT>
__global__ void __launch_bounds__(CV_THREADS, 1)
conv_contract_kernel(const __grid_constant__ CUtensorMap tmB, const double* __restrict__ Aj, const double* __restrict__ Dd,
                     const ConvChunk* __restrict__ chunks, const ConvCta* __restrict__ ctas, int N, int NI, int NP, int bp, int kc,
                     int Jpad, int nstages, double* __restrict__ ws)
{
    extern __shared__ __align__(1024) unsigned char smem[];
    const int a_bytes = KT * NP * 8, b_bytes = KT * bp * 8, d_bytes = KT * 3 * NP * 8;
    const int stage_bytes = a_bytes + b_bytes + d_bytes;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)nstages * stage_bytes);
    uint64_t* empty = full + nstages;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const ConvCta me = ctas[blockIdx.x];
    const int jblk = me.jblk;
    const long long slab = me.slab;
    const ConvChunk ch = chunks[me.chunk];
    const long long t_begin = me.t_begin;
    const int niter = (int)max(0LL, me.t_end - t_begin);

    if (tid == 0) {
        for (int s = 0; s < nstages; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], CV_WARPS); }
        mbar_fence_init();
    }
    __syncthreads();

    if (warp == CV_WARPS) {
        // ---------------- producer: two bulk copies + one TMA box per stage ----------------
        if (lane == 0) {
            tma_prefetch_desc(&tmB);
            for (int it = 0; it < niter; it++) {
                const int s = it % nstages;
                if (it >= nstages) mbar_wait(&empty[s], ((it / nstages) - 1) & 1);
                const long long f0 = (t_begin + it) * KT;
                unsigned char* dst = smem + (size_t)s * stage_bytes;
                mbar_arrive_expect_tx(&full[s], (uint32_t)stage_bytes);
                bulk_load_1d(dst, Aj + (size_t)f0 * NP, (uint32_t)a_bytes, &full[s]);
                tma_load_2d(dst + a_bytes, &tmB, 4 * ch.k_lo, (int)f0, &full[s]);
                bulk_load_1d(dst + a_bytes + b_bytes, Dd + (size_t)f0 * 3 * NP, (uint32_t)d_bytes, &full[s]);
            }
        }
        return;
    }

    // ---------------- MMA warps ----------------
    const int g = lane >> 2, t = lane & 3;
    int boff[2], doff[2];
    bool act[2];
#pragma unroll
    for (int r = 0; r < 2; r++) {
        const int row0 = ch.r_lo + 8 * (2 * warp + r);
        act[r] = row0 < ch.r_hi;                       // warp-uniform
        const int row = min(row0 + g, N * NI - 1);   // rows are (k, i) pairs, i < NI test fields
        const int k = row / NI, i = row - k * NI;
        boff[r] = 4 * min(max(k - ch.k_lo, 0), kc - 1);
        doff[r] = i;
    }
    const int j0 = jblk * (8 * CV_JF);
    const int nact = (act[0] ? 1 : 0) + (act[1] ? 1 : 0);   // warp-uniform; act[1] implies act[0]
    double acc[2][CV_JF][2];
#pragma unroll
    for (int r = 0; r < 2; r++)
#pragma unroll
        for (int jf = 0; jf < CV_JF; jf++) acc[r][jf][0] = acc[r][jf][1] = 0.0;

    for (int it = 0; it < niter; it++) {
        const int s = it % nstages;
        mbar_wait(&full[s], (it / nstages) & 1);
        mbar_release_previous(empty, it > 0 ? (it - 1) % nstages : -1, lane);   // delayed release (common.cuh)
        const double* As = reinterpret_cast<const double*>(smem + (size_t)s * stage_bytes);
        const double* Bs = As + KT * NP;
        const double* Ds = Bs + KT * bp;
        if (nact == 2) conv_stage<KT, 2>(acc, As, Bs, Ds, NP, bp, j0 + g, t, boff, doff);
        else if (nact == 1) conv_stage<KT, 1>(acc, As, Bs, Ds, NP, bp, j0 + g, t, boff, doff);
    }

    // ---------------- epilogue: partial rows -> workspace slab [row][Jpad] ----------------
    double* out = ws + (size_t)slab * (size_t)N * NI * Jpad;
#pragma unroll
    for (int r = 0; r < 2; r++) {
        const int row = ch.r_lo + 8 * (2 * warp + r) + g;
        if (!act[r] || row >= ch.r_hi) continue;
#pragma unroll
        for (int jf = 0; jf < CV_JF; jf++) {
            const int col = j0 + 8 * jf + 2 * t;
            if (col < Jpad) out[(size_t)row * Jpad + col] = acc[r][jf][0];
            if (col + 1 < Jpad) out[(size_t)row * Jpad + col + 1] = acc[r][jf][1];
        }
    }
}

// fixed-order split-K reduction and scatter into the Eigen tensor layout: C(i,j,k) at i + NI*(j + N*k), row = k*NI + i
__global__ void conv_reduce_kernel(const double* __restrict__ ws, int nslabs, int N, int NI, int Jpad, double* __restrict__ C)
{
    const long long total = (long long)N * NI * N;
    const size_t slab = (size_t)N * NI * Jpad;
    for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
        const int j = (int)(idx % N);
        const long long row = idx / N;
        const int k = (int)(row / NI), i = (int)(row % NI);
        double s = 0.0;
        for (int q = 0; q < nslabs; q++) s += ws[q * slab + (size_t)row * Jpad + j];
        C[(size_t)i + (size_t)NI * (j + (size_t)N * k)] = s;
    }
}

// ---- host -------------------------------------------------------------------------------------------------------
int conv_fused_launch(ithaca_b200* h, ConvBufs* cb, const ConvMesh& m, long long nCells, const double* U, long long ldU,
                      const double* Ub, long long ldUb, int N, double* dC, double* ms_prep, double* ms_contract, const double* Dfm,
                      long long ldD, int Ni)
{
    const int NP = (N + 3) / 8 * 8 + 4;  // smallest NP >= N with NP = 4 (mod 8): conflict-free smem pitches
    const long long nFaces = m.nInt + m.nBnd;
    const long long nF = nFaces + (m.upwind ? m.nInt : 0);   // rank-one records: upwind doubles the internal faces
    const int NI = Dfm ? Ni : N;   // test fields of the d_i records: rows of the contraction are (k, i), i < NI
    // distinct k per row chunk: enough to fill the CTA's CV_ROWS rows (G has NI = Np = 10 test fields: 7 k's were 70 of 240 rows)
    int kc = std::min(N, std::max(CV_KC, std::min(CV_KCMAX, (CV_ROWS + NI - 1) / NI)));
    if (kc > CV_KC && (kc & 1) == 0) kc--;
    const int bp = std::min(4 * kc, 4 * N);
    int CV_KT = 32;   // faces per stage; 16 when the wider b box would leave fewer than 3 stages (or ITHACA_B200_CONV_KT=16)
    if (((226 * 1024 - 256) / (32 * 8 * (4 * NP + bp) + 16)) < 3) CV_KT = 16;
    if (const char* e = env_knob("ITHACA_B200_CONV_KT")) CV_KT = (atoi(e) == 16) ? 16 : 32;
    const long long ntiles = (nF + CV_KT - 1) / CV_KT, nFpad = ntiles * CV_KT;
    const int Jpad = (N + 7) / 8 * 8;
    const int njblk = (Jpad + 8 * CV_JF - 1) / (8 * CV_JF);
    int rc;
    if ((rc = dev_ensure(h, cb->Ucm, (size_t)nCells * 3 * NP * 8))) return rc;
    if ((rc = dev_ensure(h, cb->Ubcm, (size_t)std::max<long long>(m.nBnd, 1) * 3 * NP * 8))) return rc;
    if (Dfm && (rc = dev_ensure(h, cb->Dcm, (size_t)nCells * 3 * NP * 8))) return rc;
    if ((rc = dev_ensure(h, cb->Aj, (size_t)nFpad * NP * 8 + 1024))) return rc;
    if ((rc = dev_ensure(h, cb->Bk, (size_t)nFpad * N * 32))) return rc;
    if ((rc = dev_ensure(h, cb->Dd, (size_t)nFpad * 3 * NP * 8))) return rc;

    // output row chunks: <= CV_ROWS rows and <= kc distinct k each
    std::vector<ConvChunk> chunks;
    for (int r = 0; r < N * NI;) {
        const int k_lo = r / NI;
        const int r_hi = std::min({r + CV_ROWS, (k_lo + kc) * NI, N * NI});
        chunks.push_back({r, r_hi, k_lo, 0});
        r = r_hi;
    }
    const int nchunks = (int)chunks.size();
    // K slabs per chunk: a CTA's time per stage follows the DMMAs of its busiest sub-partition (warp w runs on sub-partition w & 3,
    // two 8-row fragments per warp) plus a fixed part (~2 fragments' worth: barrier, A / B / D loads, E generation); one slab to
    // every chunk, then the remaining SMs one by one to the chunk with the largest cost per slab
    std::vector<double> ccost(nchunks);
    for (int c = 0; c < nchunks; c++) {
        const int frags = (chunks[c].r_hi - chunks[c].r_lo + 7) / 8;
        int per[4] = {0, 0, 0, 0};
        for (int f = 0; f < frags; f++) per[(f / 2) & 3]++;
        ccost[c] = 2.0 + std::max({per[0], per[1], per[2], per[3]});
    }
    std::vector<int> nsl(nchunks, 1);
    const bool uniform = [] { const char* e = env_knob("ITHACA_B200_CONV_SLABS"); return e && strcmp(e, "uniform") == 0; }();
    if (uniform) {
        const int u = (int)std::max<long long>(1, std::min<long long>(h->num_sms / (nchunks * njblk), ntiles));
        for (int c = 0; c < nchunks; c++) nsl[c] = u;
    } else {
        for (int left = h->num_sms / njblk - nchunks; left > 0; left--) {
            int best = -1;
            for (int c = 0; c < nchunks; c++)
                if (nsl[c] < ntiles && (best < 0 || ccost[c] / nsl[c] > ccost[best] / nsl[best])) best = c;
            if (best < 0) break;
            nsl[best]++;
        }
    }
    std::vector<ConvCta> ctas;
    int nslabs = 1;
    for (int c = 0; c < nchunks; c++) {
        const long long tps = (ntiles + nsl[c] - 1) / nsl[c];
        const int ns = (int)((ntiles + tps - 1) / tps);
        nslabs = std::max(nslabs, ns);
        for (int jb = 0; jb < njblk; jb++)
            for (int q = 0; q < ns; q++) ctas.push_back({c, jb, q, 0, q * tps, std::min(ntiles, (q + 1) * tps)});
    }
    if ((rc = dev_ensure(h, cb->chunks, chunks.size() * sizeof(ConvChunk) + 256 + ctas.size() * sizeof(ConvCta)))) return rc;
    const size_t cta_off = (chunks.size() * sizeof(ConvChunk) + 255) / 256 * 256;
    IB_CUDA(h, cudaMemcpyAsync(cb->chunks.p, chunks.data(), chunks.size() * sizeof(ConvChunk), cudaMemcpyHostToDevice, h->stream));
    IB_CUDA(h, cudaMemcpyAsync((char*)cb->chunks.p + cta_off, ctas.data(), ctas.size() * sizeof(ConvCta), cudaMemcpyHostToDevice, h->stream));
    const size_t ws_bytes = (size_t)nslabs * N * NI * Jpad * 8;
    if ((rc = dev_ensure(h, cb->ws, ws_bytes))) return rc;
    // chunks with fewer slabs leave their rows of the higher slabs untouched: the reduction sums all nslabs slabs
    IB_CUDA(h, cudaMemsetAsync(cb->ws.p, 0, ws_bytes, h->stream));

    cudaEvent_t e0, e1, e2;
    cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventCreate(&e2);
    cudaEventRecord(e0, h->stream);
    // zero the padding faces of the last tile (a = 0: they contribute nothing)
    if (nFpad > nF) {
        IB_CUDA(h, cudaMemsetAsync(cb->Aj.as<double>() + (size_t)nF * NP, 0, (size_t)(nFpad - nF) * NP * 8, h->stream));
        IB_CUDA(h, cudaMemsetAsync(cb->Bk.as<double>() + (size_t)nF * N * 4, 0, (size_t)(nFpad - nF) * N * 32, h->stream));
        IB_CUDA(h, cudaMemsetAsync(cb->Dd.as<double>() + (size_t)nF * 3 * NP, 0, (size_t)(nFpad - nF) * 3 * NP * 8, h->stream));
    }
    conv_transpose_kernel<<<(unsigned)((3 * nCells + 95) / 96), 256, 0, h->stream>>>(U, ldU, 3 * nCells, N, NP, cb->Ucm.as<double>());
    if (m.nBnd)
        conv_transpose_kernel<<<(unsigned)((3 * m.nBnd + 95) / 96), 256, 0, h->stream>>>(Ub, ldUb, 3 * m.nBnd, N, NP, cb->Ubcm.as<double>());
    if (Dfm)   // Ni test fields, zero-padded to N by the transpose
        conv_transpose_kernel<<<(unsigned)((3 * nCells + 95) / 96), 256, 0, h->stream>>>(Dfm, ldD, 3 * nCells, Ni, NP, cb->Dcm.as<double>());
    if (m.upwind)
        conv_face_prep_kernel<true><<<(unsigned)((nFaces + 3) / 4), dim3(64, 4), 0, h->stream>>>(
            m, cb->Ucm.as<double>(), cb->Ubcm.as<double>(), Dfm ? cb->Dcm.as<double>() : nullptr, N, NP, cb->Aj.as<double>(),
            cb->Bk.as<double>(), cb->Dd.as<double>());
    else
        conv_face_prep_kernel<false><<<(unsigned)((nFaces + 3) / 4), dim3(64, 4), 0, h->stream>>>(
            m, cb->Ucm.as<double>(), cb->Ubcm.as<double>(), Dfm ? cb->Dcm.as<double>() : nullptr, N, NP, cb->Aj.as<double>(),
            cb->Bk.as<double>(), cb->Dd.as<double>());
    IB_CUDA(h, cudaGetLastError());
    cudaEventRecord(e1, h->stream);

    CUtensorMap tmB;
    rc = make_tmap_2d(h, &tmB, cb->Bk.p, (uint64_t)4 * N, (uint64_t)nFpad, (uint64_t)N * 32, (uint32_t)bp, CV_KT, false);
    if (rc) return rc;
    const int stage_bytes = CV_KT * 8 * (4 * NP + bp);
    int nstages = std::min(8, (226 * 1024 - 256) / (stage_bytes + 16));
    if (const char* e = env_knob("ITHACA_B200_CONV_STAGES")) nstages = std::max(2, std::min(nstages, atoi(e)));   // A/B
    if (nstages < 2) { set_error(h, "project(C): too many modes for one pipeline stage"); return IB200_ERR_ARG; }
    const size_t smem = (size_t)nstages * (stage_bytes + 16) + 64;
    auto kern = (CV_KT == 16) ? conv_contract_kernel<16> : conv_contract_kernel<32>;
    IB_CUDA(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)ctas.size(), CV_THREADS, smem, h->stream>>>(
        tmB, cb->Aj.as<double>(), cb->Dd.as<double>(), cb->chunks.as<ConvChunk>(),
        reinterpret_cast<const ConvCta*>((const char*)cb->chunks.p + cta_off), N, NI, NP, bp, kc, Jpad, nstages, cb->ws.as<double>());
    IB_CUDA(h, cudaGetLastError());
    conv_reduce_kernel<<<std::min(4 * h->num_sms, (N * NI * N + 255) / 256), 256, 0, h->stream>>>(cb->ws.as<double>(), nslabs, N, NI, Jpad, dC);
    IB_CUDA(h, cudaGetLastError());
    cudaEventRecord(e2, h->stream);
    IB_CUDA(h, cudaEventSynchronize(e2));
    float a = 0.f, b = 0.f;
    cudaEventElapsedTime(&a, e0, e1);
    cudaEventElapsedTime(&b, e1, e2);
    if (ms_prep) *ms_prep += a;
    if (ms_contract) *ms_contract += b;
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2);
    h->launches += 4 + (m.nBnd ? 1 : 0);
    return IB200_OK;
}

}  // namespace ib200
