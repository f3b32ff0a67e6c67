// K5-K9 -- Galerkin projection of the steadyNS / unsteadyNS operators on the POD basis.
//
// Replaces the term builders of src/ITHACA_FOMPROBLEMS/steadyNS/steadyNS.C:
//   mass_term :1085-1108 (M), diffusive_term :883-907 (B), pressure_gradient_term :935-960 (K),
//   divergence_term :1112-1137 (P), convective_term_tens :997-1036 / _cache :1038-1083 (C)
// and the OpenFOAM fvc operators they call (Gauss linear grad/div/interpolate, Gauss linear corrected or
// uncorrected laplacian), over OpenFOAM's own owner/neighbour face addressing.
//
// Structure (the reference evaluates one fvc operator and one domainIntegrate -> one MPI reduce per ENTRY;
// here each distinct field is evaluated once and each operator is one contraction):
//   K5 face_flux_kernel       phi_j(f) = (w U_j[own] + (1-w) U_j[nei]) . Sf      gather, face-parallel
//   K6 div_phi_kernel         div(phi_j, U_k)(c)                                 gather over the cell's faces
//   K7 grad_vec / laplacian   grad(U_j) tensor, then laplacian(1, U_j)           two gathers
//   K8 grad_scalar / surface_integrate   grad(P_j), div(U_j)                     gather
//   K9 atvb (gram.cu, DMMA)   X^T diag(V) Y for every operator                   one contraction per operator
// Scatter is replaced by a cell -> face CSR (ascending face order, built on the device) so every cell sums its
// own faces in OpenFOAM's order: no atomics on doubles, results are reproducible run to run.
#include "../../include/ithaca_b200.h"
#include "common.cuh"
#include <algorithm>
#include <cstring>
#include <vector>

#ifndef LAP_UNROLL_N
#define LAP_UNROLL_N 2
#endif

namespace ib200 {

constexpr int LAP_UNROLL = LAP_UNROLL_N;   // faces per trip of the K7 face loops: the loads of that many faces are in flight together


// renumber.cu
void mesh_renumber(long long nCells, long long nInt, const int* owner, const int* neigh, std::vector<int>& old2new,
                   std::vector<int>& fnew2old);
void mesh_renumber_blocked(long long nCells, long long nInt, const int* owner, const int* neigh, int block, std::vector<int>& old2new,
                           std::vector<int>& fnew2old);

struct MeshDev {
    long long nCells, nInt, nBnd;
    const int *owner, *neigh;
    const double *Sf, *magSf, *w, *dc, *dcB, *corr, *V;
    const unsigned char* bmask;
    const int *off, *ent, *other;
    // scheme selection beyond "Gauss linear" (ithaca_b200_mesh_set_schemes); limK <= 0 / snPsi < 0: off
    const int *entL, *otherL;     // the CSR entries of each cell in the CALLER's face order (the limiter is order dependent)
    const double *Cc, *Cf;        // cell / face centres, internal order
    double limK, snPsi;
    int conv;                     // convection interpolation of fvc::div(phi, U): 0 linear, 1 upwind, 2 linearUpwind grad(U)
};

struct ProjState {
    long long nCells = 0, nInt = 0, nBnd = 0;
    bool have_mesh = false, corrected = false;
    DevBuf owner, neigh, Sf, magSf, w, dc, dcB, corr, V, bmask, off, ent, other, cnt, blocksum;   // renumbered mesh (renumber.cu)
    DevBuf owner0, neigh0, off0, ent0, other0;   // the caller's addressing (mesh_addressing reports this one)
    bool have_csr0 = false;
    DevBuf perm;    // old2new cell label (int, nCells)
    DevBuf Vw, cmask;   // shard: integration weights (V on owned cells, 0 on halo cells) and owned mask, internal order
    long long nOwned = 0;   // == nCells unless mesh_set_owned marked halo cells
    DevBuf fperm;   // original index of the internal face at each new position (int, nInt)
    DevBuf entL, otherL, Cc, Cf;   // cellMDLimited gradient limiter (mesh_set_schemes)
    double limK = 0.0, snPsi = -1.0;
    int convScheme = 0;
    bool have_centres = false, have_Tall = false;
    DevBuf Tall;   // grad(U_k) of all basis fields (linearUpwind), component-major per field like T
    DevBuf stage;   // one field in the caller's order (upload / fvc_eval staging)
    int nU = 0, nP = 0;
    long long ldU = 0, ldUb = 0, ldP = 0, ldPb = 0, ldF = 0;
    DevBuf U, Ub, P, Pb;
    DevBuf flux;   // nU columns of ldF doubles (all faces)
    bool have_flux = false;
    DevBuf fluxUser;          // caller-supplied convecting fluxes (ithaca_b200_flux_upload: L_PHImodes of fluxMethod "consistent")
    bool flux_user = false;   // the convective operators take fluxUser instead of linearInterpolate(U_j) & Sf
    DevBuf work;   // operator fields, columns of ldU (vector) or ldP (scalar) doubles
    DevBuf T;      // gradient tensors of a batch of fields, 9*nCells doubles each
    DevBuf gradP;      // grad(P_i), Np columns of ldU doubles (D and G operators)
    DevBuf bndA, bndB; // boundary-face fields of the BC3 / BC4 operators
    DevBuf Tcm, Lcm;   // cell-major gradient tensors [c][9][NP] and laplacians [c][3][NP] of all fields
    DevBuf out;    // device copy of the projected operator
    ConvBufs conv; // fused convective tensor (conv.cu)
    double ms[8] = {0};  // last timings: flux, div_phi, lap, grad, div, contraction
};

int dev_ensure(ithaca_b200* h, DevBuf& b, size_t bytes)
{
    if (bytes <= b.cap) return IB200_OK;
    if (b.p) cudaFree(b.p);
    b.p = nullptr;
    b.cap = 0;
    IB_CUDA(h, cudaMalloc(&b.p, bytes));
    b.cap = bytes;
    return IB200_OK;
}

static int ensure(ithaca_b200* h, DevBuf& b, size_t bytes) { return dev_ensure(h, b, bytes); }

static int upload(ithaca_b200* h, DevBuf& b, const void* src, size_t bytes)
{
    int rc = ensure(h, b, std::max<size_t>(bytes, 16));
    if (rc) return rc;
    if (bytes) IB_CUDA(h, cudaMemcpyAsync(b.p, src, bytes, cudaMemcpyDefault, h->stream));
    return IB200_OK;
}

void proj_destroy(ithaca_b200* h)
{
    ProjState* s = h->proj;
    if (!s) return;
    for (DevBuf* b : {&s->owner, &s->neigh, &s->Sf, &s->magSf, &s->w, &s->dc, &s->dcB, &s->corr, &s->V, &s->bmask,
                      &s->off, &s->ent, &s->other, &s->cnt, &s->blocksum, &s->U, &s->Ub, &s->P, &s->Pb, &s->flux, &s->fluxUser,
                      &s->owner0, &s->neigh0, &s->off0, &s->ent0, &s->other0, &s->entL, &s->otherL, &s->Cc, &s->Cf, &s->Tall, &s->perm, &s->fperm, &s->stage, &s->Vw, &s->cmask, &s->Tcm, &s->Lcm, &s->gradP, &s->bndA, &s->bndB,
                      &s->work, &s->T, &s->out, &s->conv.Ucm, &s->conv.Ubcm, &s->conv.Dcm, &s->conv.Aj, &s->conv.Bk, &s->conv.Dd,
                      &s->conv.ws, &s->conv.chunks})
        if (b->p) cudaFree(b->p);
    delete s;
    h->proj = nullptr;
}

static MeshDev mesh_dev(const ProjState* s)
{
    MeshDev m;
    m.nCells = s->nCells; m.nInt = s->nInt; m.nBnd = s->nBnd;
    m.owner = s->owner.as<int>(); m.neigh = s->neigh.as<int>();
    m.Sf = s->Sf.as<double>(); m.magSf = s->magSf.as<double>(); m.w = s->w.as<double>();
    m.dc = s->dc.as<double>(); m.dcB = s->dcB.as<double>();
    m.corr = s->corrected ? s->corr.as<double>() : nullptr;
    m.V = s->V.as<double>(); m.bmask = s->bmask.as<unsigned char>();
    m.off = s->off.as<int>(); m.ent = s->ent.as<int>(); m.other = s->other.as<int>();
    m.entL = s->entL.as<int>(); m.otherL = s->otherL.as<int>(); m.Cc = s->Cc.as<double>(); m.Cf = s->Cf.as<double>();
    m.limK = s->limK; m.snPsi = s->corrected ? s->snPsi : -1.0;
    m.conv = s->convScheme;
    return m;
}

// ------------------------------------------------------------------------------------------------
// cell -> face CSR, built on the device: count, exclusive scan, fill, per-cell sort by face index.
// entry = +(f+1): the cell owns face f; -(f+1): it is the neighbour. other = cell across the face, or -(b+1)
// for boundary face b. Ascending face order == primitiveMesh::cells() order == the oracle's fvo_cell_faces.
// ------------------------------------------------------------------------------------------------
__global__ void csr_count_kernel(const int* __restrict__ owner, const int* __restrict__ neigh, long long nInt,
                                 long long nF, int* __restrict__ cnt)
{
    for (long long f = blockIdx.x * (long long)blockDim.x + threadIdx.x; f < nF; f += (long long)gridDim.x * blockDim.x) {
        atomicAdd(&cnt[owner[f]], 1);
        if (f < nInt) atomicAdd(&cnt[neigh[f]], 1);
    }
}

constexpr int SCAN_B = 1024;
// per-block exclusive scan of cnt -> off (block-local), block totals -> blocksum
__global__ void __launch_bounds__(SCAN_B) scan_block_kernel(const int* __restrict__ cnt, long long n, int* __restrict__ off,
                                                            int* __restrict__ blocksum)
{
    __shared__ int sh[SCAN_B];
    const long long i = blockIdx.x * (long long)SCAN_B + threadIdx.x;
    const int v = (i < n) ? cnt[i] : 0;
    sh[threadIdx.x] = v;
    __syncthreads();
    for (int o = 1; o < SCAN_B; o <<= 1) {
        const int t = (threadIdx.x >= o) ? sh[threadIdx.x - o] : 0;
        __syncthreads();
        sh[threadIdx.x] += t;
        __syncthreads();
    }
    if (i < n) off[i] = sh[threadIdx.x] - v;
    if (threadIdx.x == SCAN_B - 1) blocksum[blockIdx.x] = sh[threadIdx.x];
}
// single block: exclusive scan of blocksum in place (nb can exceed the block size: serial chunks)
__global__ void __launch_bounds__(SCAN_B) scan_sums_kernel(int* __restrict__ blocksum, int nb)
{
    __shared__ int sh[SCAN_B];
    __shared__ int carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < nb; base += SCAN_B) {
        const int i = base + threadIdx.x;
        const int v = (i < nb) ? blocksum[i] : 0;
        sh[threadIdx.x] = v;
        __syncthreads();
        for (int o = 1; o < SCAN_B; o <<= 1) {
            const int t = (threadIdx.x >= o) ? sh[threadIdx.x - o] : 0;
            __syncthreads();
            sh[threadIdx.x] += t;
            __syncthreads();
        }
        if (i < nb) blocksum[i] = carry + sh[threadIdx.x] - v;
        __syncthreads();
        if (threadIdx.x == SCAN_B - 1) carry += sh[threadIdx.x];
        __syncthreads();
    }
}
__global__ void scan_add_kernel(int* __restrict__ off, const int* __restrict__ blocksum, long long n, int* __restrict__ cursor,
                                int total_slot)
{
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i < n) {
        const int v = off[i] + blocksum[i / SCAN_B];
        off[i] = v;
        cursor[i] = v;
    }
    (void)total_slot;
}
__global__ void csr_fill_kernel(const int* __restrict__ owner, const int* __restrict__ neigh, long long nInt, long long nF,
                                int* __restrict__ cursor, int* __restrict__ ent, int* __restrict__ other)
{
    for (long long f = blockIdx.x * (long long)blockDim.x + threadIdx.x; f < nF; f += (long long)gridDim.x * blockDim.x) {
        const int o = owner[f];
        int p = atomicAdd(&cursor[o], 1);
        ent[p] = (int)(f + 1);
        other[p] = (f < nInt) ? neigh[f] : -(int)(f - nInt + 1);
        if (f < nInt) {
            const int n = neigh[f];
            p = atomicAdd(&cursor[n], 1);
            ent[p] = -(int)(f + 1);
            other[p] = o;
        }
    }
}
// insertion sort of each cell's segment by face index (segments are short: the faces of one cell)
__global__ void csr_sort_kernel(const int* __restrict__ off, long long nCells, int* __restrict__ ent, int* __restrict__ other)
{
    const long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (c >= nCells) return;
    const int a = off[c], b = off[c + 1];
    for (int i = a + 1; i < b; i++) {
        const int e = ent[i], o = other[i], key = abs(e);
        int j = i - 1;
        while (j >= a && abs(ent[j]) > key) {
            ent[j + 1] = ent[j];
            other[j + 1] = other[j];
            j--;
        }
        ent[j + 1] = e;
        other[j + 1] = o;
    }
}

// ------------------------------------------------------------------------------------------------
// renumbered copy of the mesh (see renumber.cu): cells old -> perm[old], internal faces new -> fperm[new] (orientation
// kept, so no sign or weight changes), boundary faces keep their order.
// ------------------------------------------------------------------------------------------------
struct MeshSrc {
    const int *owner, *neigh;
    const double *Sf, *magSf, *w, *dc, *corr;
};
__global__ void permute_faces_kernel(MeshSrc in, const int* __restrict__ perm, const int* __restrict__ fperm, long long nInt,
                                     long long nF, int* __restrict__ owner, int* __restrict__ neigh, double* __restrict__ Sf,
                                     double* __restrict__ magSf, double* __restrict__ w, double* __restrict__ dc,
                                     double* __restrict__ corr)
{
    for (long long f = blockIdx.x * (long long)blockDim.x + threadIdx.x; f < nF; f += (long long)gridDim.x * blockDim.x) {
        const long long g = (f < nInt) ? fperm[f] : f;
        owner[f] = perm[in.owner[g]];
        Sf[3 * f] = in.Sf[3 * g]; Sf[3 * f + 1] = in.Sf[3 * g + 1]; Sf[3 * f + 2] = in.Sf[3 * g + 2];
        magSf[f] = in.magSf[g];
        if (f < nInt) {
            neigh[f] = perm[in.neigh[g]];
            w[f] = in.w[g];
            dc[f] = in.dc[g];
            if (in.corr) { corr[3 * f] = in.corr[3 * g]; corr[3 * f + 1] = in.corr[3 * g + 1]; corr[3 * f + 2] = in.corr[3 * g + 2]; }
        }
    }
}
// out[dim*perm[c] + d] = in[dim*c + d]   (caller's order -> internal order)
__global__ void permute_cells_kernel(const double* __restrict__ in, const int* __restrict__ perm, long long n, int dim,
                                     double* __restrict__ out)
{
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n * dim; i += (long long)gridDim.x * blockDim.x) {
        const long long c = i / dim;
        out[(long long)perm[c] * dim + (i - c * dim)] = in[i];
    }
}
// out[dim*c + d] = in[dim*perm[c] + d]   (internal order -> caller's order)
__global__ void unpermute_cells_kernel(const double* __restrict__ in, const int* __restrict__ perm, long long n, int dim,
                                       double* __restrict__ out)
{
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n * dim; i += (long long)gridDim.x * blockDim.x) {
        const long long c = i / dim;
        out[i] = in[(long long)perm[c] * dim + (i - c * dim)];
    }
}
// face field back to the caller's face order
__global__ void unpermute_faces_kernel(const double* __restrict__ in, const int* __restrict__ fperm, long long nInt, long long nF,
                                       double* __restrict__ out)
{
    for (long long f = blockIdx.x * (long long)blockDim.x + threadIdx.x; f < nF; f += (long long)gridDim.x * blockDim.x)
        out[(f < nInt) ? fperm[f] : f] = in[f];
}

// each cell's CSR segment re-sorted by the CALLER's face index (internal face f sat at fperm[f] before the renumbering)
__global__ void csr_sort_caller_order_kernel(const int* __restrict__ off, long long nCells, const int* __restrict__ fperm, long long nInt,
                                             int* __restrict__ ent, int* __restrict__ other)
{
    const long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (c >= nCells) return;
    const int a = off[c], b = off[c + 1];
    auto key = [&](int e) { const int f = abs(e) - 1; return f < nInt ? fperm[f] : f; };
    for (int i = a + 1; i < b; i++) {
        const int e = ent[i], o = other[i], k = key(e);
        int j = i - 1;
        while (j >= a && key(ent[j]) > k) {
            ent[j + 1] = ent[j];
            other[j + 1] = other[j];
            j--;
        }
        ent[j + 1] = e;
        other[j + 1] = o;
    }
}
// face centres into the internal face order
__global__ void permute_face_vectors_kernel(const double* __restrict__ in, const int* __restrict__ fperm, long long nInt, long long nF,
                                            double* __restrict__ out)
{
    for (long long f = blockIdx.x * (long long)blockDim.x + threadIdx.x; f < nF; f += (long long)gridDim.x * blockDim.x) {
        const long long g = (f < nInt) ? fperm[f] : f;
        out[3 * f] = in[3 * g]; out[3 * f + 1] = in[3 * g + 1]; out[3 * f + 2] = in[3 * g + 2];
    }
}

// ------------------------------------------------------------------------------------------------
// gradSchemes "cellMDLimited Gauss linear k" (tutorials/CFD/04unsteadyNS/system/fvSchemes; OpenFOAM
// limitedGradSchemes/cellMDLimitedGrad: calcGrad + limitFaceCmpt). One scalar component of one cell: bounds = max/min
// of the cell's value and its face neighbours' (boundary faces: the patch value) minus the cell value, widened by
// (1/k - 1)(max - min) for k < 1; then the gradient is clipped face by face along dcf = Cf - C, in the CALLER's face
// order (OpenFOAM's loop: internal faces ascending, then boundary faces) -- the clipping is order dependent, so the
// renumbered CSR order of the summation kernels cannot be used here.
// ------------------------------------------------------------------------------------------------
template <class NbVal, class BndVal>
__device__ __forceinline__ void md_limit(const MeshDev& m, long long c, double vc, NbVal nbval, BndVal bndval, double& gx, double& gy,
                                         double& gz)
{
    const int e0 = m.off[c], e1 = m.off[c + 1];
    double mx = vc, mn = vc;
    for (int e = e0; e < e1; e++) {
        const int ot = m.otherL[e];
        double v;
        if (ot >= 0) v = nbval(ot);
        else {
            const long long b = -(long long)ot - 1;
            if (!m.bmask[b]) continue;
            v = bndval(b);
        }
        mx = fmax(mx, v);
        mn = fmin(mn, v);
    }
    mx -= vc;
    mn -= vc;
    if (m.limK < 1.0) {
        const double dd = (1.0 / m.limK - 1.0) * (mx - mn);
        mx += dd;
        mn -= dd;
    }
    const double cx = m.Cc[3 * c], cy = m.Cc[3 * c + 1], cz = m.Cc[3 * c + 2];
    for (int e = e0; e < e1; e++) {
        const int ot = m.otherL[e];
        if (ot < 0 && !m.bmask[-(long long)ot - 1]) continue;
        const long long f = abs(m.entL[e]) - 1;
        const double dx = m.Cf[3 * f] - cx, dy = m.Cf[3 * f + 1] - cy, dz = m.Cf[3 * f + 2] - cz;
        const double ex = dx * gx + dy * gy + dz * gz;
        double sft;
        if (ex > mx) sft = mx - ex;
        else if (ex < mn) sft = mn - ex;
        else continue;
        const double msq = dx * dx + dy * dy + dz * dz;
        gx = gx + dx * sft / msq;
        gy = gy + dy * sft / msq;
        gz = gz + dz * sft / msq;
    }
}

// ------------------------------------------------------------------------------------------------
// K5: face flux of nf fields. phi[j*ldF + f] = (w U_j[o] + (1-w) U_j[n]) . Sf(f); boundary: Ub_j(b) . Sf,
// 0 on `empty` patches. Face geometry is loaded once per face and reused for FB fields.
// ------------------------------------------------------------------------------------------------
constexpr int PROJ_T = 256;
constexpr int FB = 4;  // fields per thread in the gather kernels

__global__ void __launch_bounds__(PROJ_T)
face_flux_kernel(MeshDev m, const double* __restrict__ U, long long ldU, const double* __restrict__ Ub, long long ldUb,
                 int nf, double* __restrict__ phi, long long ldF)
{
    const long long f = blockIdx.x * (long long)PROJ_T + threadIdx.x;
    const long long nF = m.nInt + m.nBnd;
    if (f >= nF) return;
    const int j0 = blockIdx.y * FB;
    const double s0 = m.Sf[3 * f], s1 = m.Sf[3 * f + 1], s2 = m.Sf[3 * f + 2];
    if (f < m.nInt) {
        const int o = m.owner[f], n = m.neigh[f];
        const double w = m.w[f], w1 = 1.0 - w;
#pragma unroll
        for (int jj = 0; jj < FB; jj++) {
            const int j = j0 + jj;
            if (j >= nf) break;
            const double* u = U + (size_t)j * ldU;
            double s = 0.0;
            s += (w * u[3 * (size_t)o] + w1 * u[3 * (size_t)n]) * s0;
            s += (w * u[3 * (size_t)o + 1] + w1 * u[3 * (size_t)n + 1]) * s1;
            s += (w * u[3 * (size_t)o + 2] + w1 * u[3 * (size_t)n + 2]) * s2;
            phi[(size_t)j * ldF + f] = s;
        }
    } else {
        const long long b = f - m.nInt;
        const bool on = m.bmask[b] != 0;
#pragma unroll
        for (int jj = 0; jj < FB; jj++) {
            const int j = j0 + jj;
            if (j >= nf) break;
            const double* ub = Ub + (size_t)j * ldUb + 3 * b;
            double s = 0.0;
            if (on) { s += ub[0] * s0; s += ub[1] * s1; s += ub[2] * s2; }
            phi[(size_t)j * ldF + f] = s;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// K6: out_k = fvc::div(phi, U_k) = surfaceIntegrate(phi * linearInterpolate(U_k)), FB fields k per thread,
// one thread per cell walking the cell's CSR segment (OpenFOAM's summation order, then / V).
// ------------------------------------------------------------------------------------------------
// CONV selects the interpolation of U_k onto the internal faces (divSchemes "Gauss <scheme>"; boundary faces always take
// the patch value): 0 linear; 1 upwind (owner value when phi >= 0, upwind::weights = pos0(faceFlux)); 2 linearUpwind
// grad(U): upwind value + (Cf - C_upwind) . grad(U_k)[upwind] (linearUpwind::correction), grad from Tall.
template <int CONV>
__device__ __forceinline__ double conv_face_value(const MeshDev& m, long long f, bool own, double ph, double w, double w1, double vc, double vn,
                                                  long long c, long long ot, const double* __restrict__ tk, long long ldc, int d)
{
    if (CONV == 0) return own ? (w * vc + w1 * vn) : (w * vn + w1 * vc);
    const bool up_is_c = (ph >= 0.0) == own;
    double v = up_is_c ? vc : vn;
    if (CONV == 2) {
        const long long u = up_is_c ? c : ot;
        const double dx = m.Cf[3 * f] - m.Cc[3 * u], dy = m.Cf[3 * f + 1] - m.Cc[3 * u + 1], dz = m.Cf[3 * f + 2] - m.Cc[3 * u + 2];
        v += dx * tk[(size_t)d * ldc + u] + dy * tk[(size_t)(3 + d) * ldc + u] + dz * tk[(size_t)(6 + d) * ldc + u];
    }
    return v;
}

template <int CONV>
__global__ void __launch_bounds__(PROJ_T)
div_phi_kernel(MeshDev m, const double* __restrict__ phi, const double* __restrict__ U, long long ldU,
               const double* __restrict__ Ub, long long ldUb, int nk, double* __restrict__ out, long long ldo,
               const double* __restrict__ Tall, long long ldT)
{
    const long long ldc = ldT / 9;
    const long long c = blockIdx.x * (long long)PROJ_T + threadIdx.x;
    if (c >= m.nCells) return;
    const int k0 = blockIdx.y * FB;
    double uc[FB][3], acc[FB][3];
#pragma unroll
    for (int kk = 0; kk < FB; kk++) {
        const int k = min(k0 + kk, nk - 1);
#pragma unroll
        for (int d = 0; d < 3; d++) { uc[kk][d] = U[(size_t)k * ldU + 3 * c + d]; acc[kk][d] = 0.0; }
    }
    const int e0 = m.off[c], e1 = m.off[c + 1];
    for (int e = e0; e < e1; e++) {
        const int en = m.ent[e], ot = m.other[e];
        const long long f = abs(en) - 1;
        const double ph = phi[f];
        if (ot >= 0) {
            const double w = m.w[f], w1 = 1.0 - w;
            const bool own = en > 0;
#pragma unroll
            for (int kk = 0; kk < FB; kk++) {
                const int k = min(k0 + kk, nk - 1);
                const double* un = U + (size_t)k * ldU + 3 * (size_t)ot;
                const double* tk = CONV == 2 ? Tall + (size_t)k * ldT : nullptr;
#pragma unroll
                for (int d = 0; d < 3; d++) {
                    const double x = un[d];
                    const double uf = conv_face_value<CONV>(m, f, own, ph, w, w1, uc[kk][d], x, c, ot, tk, ldc, d);
                    const double v = ph * uf;
                    acc[kk][d] += own ? v : -v;
                }
            }
        } else {
            const long long b = -(long long)ot - 1;
            if (!m.bmask[b]) continue;
#pragma unroll
            for (int kk = 0; kk < FB; kk++) {
                const int k = min(k0 + kk, nk - 1);
                const double* ub = Ub + (size_t)k * ldUb + 3 * b;
#pragma unroll
                for (int d = 0; d < 3; d++) acc[kk][d] += ph * ub[d];
            }
        }
    }
    const double V = m.V[c];
#pragma unroll
    for (int kk = 0; kk < FB; kk++) {
        const int k = k0 + kk;
        if (k >= nk) break;
#pragma unroll
        for (int d = 0; d < 3; d++) out[(size_t)k * ldo + 3 * c + d] = acc[kk][d] / V;
    }
}

// K8a: out_j = surfaceIntegrate(phi_j) = fvc::div(U_j) when phi_j is the face flux of U_j. Scalar output.
__global__ void __launch_bounds__(PROJ_T)
surface_integrate_kernel(MeshDev m, const double* __restrict__ phi, long long ldF, int nf, double* __restrict__ out, long long ldo)
{
    const long long c = blockIdx.x * (long long)PROJ_T + threadIdx.x;
    if (c >= m.nCells) return;
    const int j0 = blockIdx.y * FB;
    double acc[FB];
#pragma unroll
    for (int jj = 0; jj < FB; jj++) acc[jj] = 0.0;
    const int e0 = m.off[c], e1 = m.off[c + 1];
    for (int e = e0; e < e1; e++) {
        const int en = m.ent[e], ot = m.other[e];
        const long long f = abs(en) - 1;
        if (ot < 0 && !m.bmask[-(long long)ot - 1]) continue;
#pragma unroll
        for (int jj = 0; jj < FB; jj++) {
            const int j = min(j0 + jj, nf - 1);
            const double ph = phi[(size_t)j * ldF + f];
            acc[jj] += (en > 0) ? ph : -ph;
        }
    }
    const double V = m.V[c];
#pragma unroll
    for (int jj = 0; jj < FB; jj++)
        if (j0 + jj < nf) out[(size_t)(j0 + jj) * ldo + c] = acc[jj] / V;
}

// K8b: out_j = fvc::grad(P_j), Gauss linear: sum_f +-Sf * (w p[o] + (1-w) p[n]) / V; boundary Sf * pb.
__global__ void __launch_bounds__(PROJ_T)
grad_scalar_kernel(MeshDev m, const double* __restrict__ P, long long ldP, const double* __restrict__ Pb, long long ldPb,
                   int np, double* __restrict__ out, long long ldo)
{
    const long long c = blockIdx.x * (long long)PROJ_T + threadIdx.x;
    if (c >= m.nCells) return;
    const int j0 = blockIdx.y * FB;
    double pc[FB], acc[FB][3];
#pragma unroll
    for (int jj = 0; jj < FB; jj++) {
        pc[jj] = P[(size_t)min(j0 + jj, np - 1) * ldP + c];
        acc[jj][0] = acc[jj][1] = acc[jj][2] = 0.0;
    }
    const int e0 = m.off[c], e1 = m.off[c + 1];
    for (int e = e0; e < e1; e++) {
        const int en = m.ent[e], ot = m.other[e];
        const long long f = abs(en) - 1;
        const double s[3] = {m.Sf[3 * f], m.Sf[3 * f + 1], m.Sf[3 * f + 2]};
        if (ot >= 0) {
            const double w = m.w[f], w1 = 1.0 - w;
            const bool own = en > 0;
#pragma unroll
            for (int jj = 0; jj < FB; jj++) {
                const double x = P[(size_t)min(j0 + jj, np - 1) * ldP + ot];
                const double pf = own ? (w * pc[jj] + w1 * x) : (w * x + w1 * pc[jj]);
#pragma unroll
                for (int d = 0; d < 3; d++) {
                    const double v = s[d] * pf;
                    acc[jj][d] += own ? v : -v;
                }
            }
        } else {
            const long long b = -(long long)ot - 1;
            if (!m.bmask[b]) continue;
#pragma unroll
            for (int jj = 0; jj < FB; jj++) {
                const double pb = Pb[(size_t)min(j0 + jj, np - 1) * ldPb + b];
#pragma unroll
                for (int d = 0; d < 3; d++) acc[jj][d] += s[d] * pb;
            }
        }
    }
    const double V = m.V[c];
#pragma unroll
    for (int jj = 0; jj < FB; jj++)
        if (j0 + jj < np) {
            double g0 = acc[jj][0] / V, g1 = acc[jj][1] / V, g2 = acc[jj][2] / V;
            if (m.limK > 0.0) {
                const double* p = P + (size_t)(j0 + jj) * ldP;
                const double* pb = Pb + (size_t)(j0 + jj) * ldPb;
                md_limit(m, c, pc[jj], [&](int o) { return p[o]; }, [&](long long b) { return pb[b]; }, g0, g1, g2);
            }
            double* o = out + (size_t)(j0 + jj) * ldo + 3 * c;
            o[0] = g0; o[1] = g1; o[2] = g2;
        }
}

// Adjoint of `fvc::div` with owner-extrapolated boundary values: int P_i div(A) = sum_c V_c A[c] . t_i[c] for every vector
// field A whose boundary value is its owner cell's (the fvc:: results of the flux-method operators BP / Cf), with
//   t_i[c] = ( sum_{f in c, internal} wc(f) Sf (P_i[own(f)] - P_i[nei(f)]) + sum_{b in c} Sf_b P_i[c] ) / V_c,
// wc(f) = w_f when c owns f, 1 - w_f otherwise. Same cell walk as grad_scalar_kernel.
__global__ void __launch_bounds__(PROJ_T)
div_adjoint_kernel(MeshDev m, const double* __restrict__ P, long long ldP, int np, double* __restrict__ out, long long ldo)
{
    const long long c = blockIdx.x * (long long)PROJ_T + threadIdx.x;
    if (c >= m.nCells) return;
    const int j0 = blockIdx.y * FB;
    double pc[FB], acc[FB][3];
#pragma unroll
    for (int jj = 0; jj < FB; jj++) {
        pc[jj] = P[(size_t)min(j0 + jj, np - 1) * ldP + c];
        acc[jj][0] = acc[jj][1] = acc[jj][2] = 0.0;
    }
    const int e0 = m.off[c], e1 = m.off[c + 1];
    for (int e = e0; e < e1; e++) {
        const int en = m.ent[e], ot = m.other[e];
        const long long f = abs(en) - 1;
        const double s[3] = {m.Sf[3 * f], m.Sf[3 * f + 1], m.Sf[3 * f + 2]};
        if (ot >= 0) {
            const double w = m.w[f];
            const bool own = en > 0;
            const double wc = own ? w : 1.0 - w;
#pragma unroll
            for (int jj = 0; jj < FB; jj++) {
                const double x = P[(size_t)min(j0 + jj, np - 1) * ldP + ot];
                const double dp = wc * (own ? pc[jj] - x : x - pc[jj]);
#pragma unroll
                for (int d = 0; d < 3; d++) acc[jj][d] += s[d] * dp;
            }
        } else {
            const long long b = -(long long)ot - 1;
            if (!m.bmask[b]) continue;
#pragma unroll
            for (int jj = 0; jj < FB; jj++)
#pragma unroll
                for (int d = 0; d < 3; d++) acc[jj][d] += s[d] * pc[jj];
        }
    }
    const double V = m.V[c];
#pragma unroll
    for (int jj = 0; jj < FB; jj++)
        if (j0 + jj < np)
#pragma unroll
            for (int d = 0; d < 3; d++) out[(size_t)(j0 + jj) * ldo + 3 * c + d] = acc[jj][d] / V;
}

// K7a: T_j = fvc::grad(U_j), Gauss linear: T[c][i][d] = sum_f +-Sf_i * Uf_d / V. One field per block.
// Occupancy experiments: -DGRADV_MINB=n / -DLAP_MINB=n force n CTAs per SM. The default leaves the register count to ptxas
// (62 / 80 registers, 4 / 3 CTAs per SM): NOT the same as a minimum of 1, which lets ptxas take 78 / 108 registers and costs
// 2 ms of the 15.4 ms laplacian stage at C4.
#ifdef GRADV_MINB
#define GRADV_BOUNDS __launch_bounds__(PROJ_T, GRADV_MINB)
#else
#define GRADV_BOUNDS __launch_bounds__(PROJ_T)
#endif
#ifdef LAP_MINB
#define LAP_BOUNDS __launch_bounds__(PROJ_T, LAP_MINB)
#else
#define LAP_BOUNDS __launch_bounds__(PROJ_T)
#endif
template <bool SCH>   // SCH: a scheme of mesh_set_schemes is active (kept out of the default instantiation)
__global__ void GRADV_BOUNDS
grad_vec_kernel(MeshDev m, const double* __restrict__ U, long long ldU, const double* __restrict__ Ub, long long ldUb,
                int j0, int nb, double* __restrict__ T, long long ldT)
{
    // blockIdx.x = cell block * nb + field: the nb fields of one cell block run together, so the mesh arrays of the
    // block are fetched from DRAM once and re-read from L2 by the other fields
    const int jl = blockIdx.x % nb;
    const long long c = (blockIdx.x / nb) * (long long)PROJ_T + threadIdx.x;
    if (c >= m.nCells) return;
    const int j = j0 + jl;
    const double* u = U + (size_t)j * ldU;
    const double uc[3] = {u[3 * c], u[3 * c + 1], u[3 * c + 2]};
    double acc[3][3];
#pragma unroll
    for (int i = 0; i < 3; i++) acc[i][0] = acc[i][1] = acc[i][2] = 0.0;
    const int e0 = m.off[c], e1 = m.off[c + 1];
#pragma unroll LAP_UNROLL
    for (int e = e0; e < e1; e++) {
        const int en = m.ent[e], ot = m.other[e];
        const long long f = abs(en) - 1;
        const double s[3] = {m.Sf[3 * f], m.Sf[3 * f + 1], m.Sf[3 * f + 2]};
        if (ot >= 0) {
            const double w = m.w[f], w1 = 1.0 - w;
            const bool own = en > 0;
#pragma unroll
            for (int d = 0; d < 3; d++) {
                const double x = u[3 * (size_t)ot + d];
                const double uf = own ? (w * uc[d] + w1 * x) : (w * x + w1 * uc[d]);
#pragma unroll
                for (int i = 0; i < 3; i++) {
                    const double v = s[i] * uf;
                    acc[i][d] += own ? v : -v;
                }
            }
        } else {
            const long long b = -(long long)ot - 1;
            if (!m.bmask[b]) continue;
            const double* ub = Ub + (size_t)j * ldUb + 3 * b;
#pragma unroll
            for (int d = 0; d < 3; d++)
#pragma unroll
                for (int i = 0; i < 3; i++) acc[i][d] += s[i] * ub[d];
        }
    }
    const double V = m.V[c];
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int d = 0; d < 3; d++) acc[i][d] /= V;
    if (SCH && m.limK > 0.0) {
        const double* ub = Ub + (size_t)j * ldUb;
#pragma unroll
        for (int d = 0; d < 3; d++)   // component d: its gradient is column d of the tensor
            md_limit(m, c, uc[d], [&](int o) { return u[3 * (size_t)o + d]; }, [&](long long b) { return ub[3 * b + d]; }, acc[0][d],
                     acc[1][d], acc[2][d]);
    }
    // component-major within a field, T[j][3i+d][c] (ldc = ldT/9 cells per component): consecutive cells of a warp
    // read and write consecutive doubles (the 72-byte cell-major records cost 3-4x the L1 sector accesses)
    const long long ldc = ldT / 9;
    double* t = T + (size_t)jl * ldT + c;
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int d = 0; d < 3; d++) t[(3 * i + d) * ldc] = acc[i][d];
}

// K7b: out_j = fvc::laplacian(1, U_j) = surfaceIntegrate(|Sf| snGrad(U_j)); snGrad = dc (U[n]-U[o])
// (+ corr . (w T[o] + (1-w) T[n]) when corrected); boundary |Sf| dcB (Ub - U[c]). Grid as in grad_vec_kernel.
template <bool SCH>
__global__ void LAP_BOUNDS
laplacian_kernel(MeshDev m, const double* __restrict__ U, long long ldU, const double* __restrict__ Ub, long long ldUb,
                 int j0, int nb, const double* __restrict__ T, long long ldT, double* __restrict__ out, long long ldo)
{
    const int jl = blockIdx.x % nb;
    const long long c = (blockIdx.x / nb) * (long long)PROJ_T + threadIdx.x;
    if (c >= m.nCells) return;
    const int j = j0 + jl;
    const double* u = U + (size_t)j * ldU;
    const double* t = T ? T + (size_t)jl * ldT : nullptr;
    const double uc[3] = {u[3 * c], u[3 * c + 1], u[3 * c + 2]};
    const long long ldc = ldT / 9;
    double tc[9];
    if (t)
#pragma unroll
        for (int k = 0; k < 9; k++) tc[k] = t[k * ldc + c];
    double acc[3] = {0.0, 0.0, 0.0};
    const int e0 = m.off[c], e1 = m.off[c + 1];
#pragma unroll LAP_UNROLL
    for (int e = e0; e < e1; e++) {
        const int en = m.ent[e], ot = m.other[e];
        const long long f = abs(en) - 1;
        const double ms = m.magSf[f];
        if (ot >= 0) {
            const bool own = en > 0;
            const double dc = m.dc[f];
            double cr[3] = {0.0, 0.0, 0.0};
            if (t) {
                const double w = m.w[f], w1 = 1.0 - w;
                const double* tn = t + (size_t)ot;
#pragma unroll
                for (int i = 0; i < 3; i++) {
                    const double k = m.corr[3 * f + i];
#pragma unroll
                    for (int d = 0; d < 3; d++) {
                        const double x = tn[(3 * i + d) * ldc];
                        cr[d] += k * (own ? (w * tc[3 * i + d] + w1 * x) : (w * x + w1 * tc[3 * i + d]));
                    }
                }
            }
            if (SCH && t && m.snPsi >= 0.0) {
                // snGrad "limited psi" (limitedSnGrad::correction): limiter = min(psi |snGrad_unc| / ((1-psi) |corr| + SMALL), 1);
                // both magnitudes are the same seen from either side of the face
                double snu[3];
#pragma unroll
                for (int d = 0; d < 3; d++) {
                    const double x = u[3 * (size_t)ot + d];
                    snu[d] = own ? dc * (x - uc[d]) : dc * (uc[d] - x);
                }
                const double a = sqrt(snu[0] * snu[0] + snu[1] * snu[1] + snu[2] * snu[2]);
                const double cm = sqrt(cr[0] * cr[0] + cr[1] * cr[1] + cr[2] * cr[2]);
                const double lim = fmin(m.snPsi * a / ((1.0 - m.snPsi) * cm + 1e-15), 1.0);
#pragma unroll
                for (int d = 0; d < 3; d++) {
                    const double v = ms * (snu[d] + lim * cr[d]);
                    acc[d] += own ? v : -v;
                }
            } else {
#pragma unroll
                for (int d = 0; d < 3; d++) {
                    const double x = u[3 * (size_t)ot + d];
                    double sn = own ? dc * (x - uc[d]) : dc * (uc[d] - x);
                    if (t) sn += cr[d];
                    const double v = ms * sn;
                    acc[d] += own ? v : -v;
                }
            }
        } else {
            const long long b = -(long long)ot - 1;
            if (!m.bmask[b]) continue;
            const double* ub = Ub + (size_t)j * ldUb + 3 * b;
            const double dcb = m.dcB[b];
#pragma unroll
            for (int d = 0; d < 3; d++) acc[d] += ms * (dcb * (ub[d] - uc[d]));
        }
    }
    const double V = m.V[c];
#pragma unroll
    for (int d = 0; d < 3; d++) out[(size_t)j * ldo + 3 * c + d] = acc[d] / V;
}

// ------------------------------------------------------------------------------------------------
// K10: boundary-face kernels of the PPE operators (steadyNS::pressure_BC3 / pressure_BC4, steadyNS.C:1439-1513).
// One thread per (boundary face, field). The boundary value of an fvc::grad result on face b with owner cell P is
// grad_P + n (snGrad_b - n . grad_P), snGrad_b = deltaCoeffs_b (value_b - value_P) (gaussGrad::correctBoundaryConditions);
// grad_P is recomputed here from the cell's CSR segment (same order as grad_vec_kernel / grad_scalar_kernel), so no
// N x 9 x nCells gradient array is materialised for the few cells that touch the boundary.
// Faces of masked patches and (in a shard) faces whose owner is a halo cell produce zeros.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool bnd_face_active(const MeshDev& m, const unsigned char* cmask, long long b, int c)
{
    return m.bmask[b] != 0 && (!cmask || cmask[c] != 0);
}

// Fb[j][b] = (n . Ub_j) |Sf|   (BC4 = Pb^T Fb)
__global__ void __launch_bounds__(PROJ_T)
bnd_normal_flux_kernel(MeshDev m, const unsigned char* __restrict__ cmask, const double* __restrict__ Ub, long long ldUb, int N,
                       double* __restrict__ out, long long ldo)
{
    const long long b = blockIdx.x * (long long)PROJ_T + threadIdx.x;
    if (b >= m.nBnd) return;
    const int j = blockIdx.y;
    const long long f = m.nInt + b;
    double v = 0.0;
    if (bnd_face_active(m, cmask, b, m.owner[f])) {
        const double ms = m.magSf[f];
        const double* u = Ub + (size_t)j * ldUb + 3 * b;
        v = ((m.Sf[3 * f] / ms) * u[0] + (m.Sf[3 * f + 1] / ms) * u[1] + (m.Sf[3 * f + 2] / ms) * u[2]) * ms;
    }
    out[(size_t)j * ldo + b] = v;
}

// Lb[j][b] = laplacian(U_j)[owner(b)] . Sf_b   (BC1 = Pb^T Lb; lap: N columns of ldl doubles)
__global__ void __launch_bounds__(PROJ_T)
bnd_owner_dot_sf_kernel(MeshDev m, const unsigned char* __restrict__ cmask, const double* __restrict__ lap, long long ldl, int N,
                        double* __restrict__ out, long long ldo)
{
    const long long b = blockIdx.x * (long long)PROJ_T + threadIdx.x;
    if (b >= m.nBnd) return;
    const int j = blockIdx.y;
    const long long f = m.nInt + b;
    const int c = m.owner[f];
    double v = 0.0;
    if (bnd_face_active(m, cmask, b, c)) {
        const double* l = lap + (size_t)j * ldl + 3 * (size_t)c;
        v = l[0] * m.Sf[3 * f] + l[1] * m.Sf[3 * f + 1] + l[2] * m.Sf[3 * f + 2];
    }
    out[(size_t)j * ldo + b] = v;
}

// Db[k][b] = fvc::div(phi, U_k)[owner(b)] . Sf_b for the boundary faces b (BC2 = Pb^T Db per flux field): the div(phi_j, U_k)
// value of the owner cell is recomputed from the cell's CSR segment with div_phi_kernel's own arithmetic (same order, same
// rounding), so only the N_b boundary-owner cells are visited instead of materialising N^2 whole div fields.
template <int CONV>
__global__ void __launch_bounds__(PROJ_T)
bnd_div_phi_dot_sf_kernel(MeshDev m, const unsigned char* __restrict__ cmask, const double* __restrict__ phi,
                          const double* __restrict__ U, long long ldU, const double* __restrict__ Ub, long long ldUb, int nk,
                          double* __restrict__ out, long long ldo, const double* __restrict__ Tall, long long ldT)
{
    const long long ldc = ldT / 9;
    const long long bf = blockIdx.x * (long long)PROJ_T + threadIdx.x;
    if (bf >= m.nBnd) return;
    const int k0 = blockIdx.y * FB;
    const long long fb = m.nInt + bf;
    const int c = m.owner[fb];
    if (!bnd_face_active(m, cmask, bf, c)) {
#pragma unroll
        for (int kk = 0; kk < FB; kk++)
            if (k0 + kk < nk) out[(size_t)(k0 + kk) * ldo + bf] = 0.0;
        return;
    }
    double uc[FB][3], acc[FB][3];
#pragma unroll
    for (int kk = 0; kk < FB; kk++) {
        const int k = min(k0 + kk, nk - 1);
#pragma unroll
        for (int d = 0; d < 3; d++) { uc[kk][d] = U[(size_t)k * ldU + 3 * (size_t)c + d]; acc[kk][d] = 0.0; }
    }
    const int e0 = m.off[c], e1 = m.off[c + 1];
    for (int e = e0; e < e1; e++) {
        const int en = m.ent[e], ot = m.other[e];
        const long long f = abs(en) - 1;
        const double ph = phi[f];
        if (ot >= 0) {
            const double w = m.w[f], w1 = 1.0 - w;
            const bool own = en > 0;
#pragma unroll
            for (int kk = 0; kk < FB; kk++) {
                const int k = min(k0 + kk, nk - 1);
                const double* un = U + (size_t)k * ldU + 3 * (size_t)ot;
                const double* tk = CONV == 2 ? Tall + (size_t)k * ldT : nullptr;
#pragma unroll
                for (int d = 0; d < 3; d++) {
                    const double x = un[d];
                    const double uf = conv_face_value<CONV>(m, f, own, ph, w, w1, uc[kk][d], x, c, ot, tk, ldc, d);
                    const double v = ph * uf;
                    acc[kk][d] += own ? v : -v;
                }
            }
        } else {
            const long long b = -(long long)ot - 1;
            if (!m.bmask[b]) continue;
#pragma unroll
            for (int kk = 0; kk < FB; kk++) {
                const int k = min(k0 + kk, nk - 1);
                const double* ub = Ub + (size_t)k * ldUb + 3 * b;
#pragma unroll
                for (int d = 0; d < 3; d++) acc[kk][d] += ph * ub[d];
            }
        }
    }
    const double V = m.V[c];
    const double sx = m.Sf[3 * fb], sy = m.Sf[3 * fb + 1], sz = m.Sf[3 * fb + 2];
#pragma unroll
    for (int kk = 0; kk < FB; kk++) {
        const int k = k0 + kk;
        if (k >= nk) break;
        out[(size_t)k * ldo + bf] = (acc[kk][0] / V) * sx + (acc[kk][1] / V) * sy + (acc[kk][2] / V) * sz;
    }
}

// Xp[j][b] = component `comp` of U_j on patch face start+b (j < N), Xp[N][b] = 1   (bcVelocityVec / bcVelocityMat)
__global__ void __launch_bounds__(PROJ_T)
bnd_patch_component_kernel(MeshDev m, const unsigned char* __restrict__ cmask, const double* __restrict__ Ub, long long ldUb, int N,
                           long long start, long long size, int comp, double* __restrict__ out, long long ldo)
{
    const long long b = blockIdx.x * (long long)PROJ_T + threadIdx.x;
    if (b >= size) return;
    const int j = blockIdx.y;
    const long long bf = start + b;
    const bool on = !cmask || cmask[m.owner[m.nInt + bf]] != 0;   // shard: faces of halo cells belong to another rank
    out[(size_t)j * ldo + b] = !on ? 0.0 : (j < N ? Ub[(size_t)j * ldUb + 3 * bf + comp] : 1.0);
}

// Cb[j][3b+d] = curl(U_j) on boundary face b = 2 *skew(grad_b): (g_yz - g_zy, g_zx - g_xz, g_xy - g_yx), g_id = d_i U_d
__global__ void __launch_bounds__(PROJ_T)
bnd_curl_kernel(MeshDev m, const unsigned char* __restrict__ cmask, const double* __restrict__ U, long long ldU,
                const double* __restrict__ Ub, long long ldUb, int N, double* __restrict__ out, long long ldo)
{
    const long long b0 = blockIdx.x * (long long)PROJ_T + threadIdx.x;
    if (b0 >= m.nBnd) return;
    const int j = blockIdx.y;
    const long long f0 = m.nInt + b0;
    const int c = m.owner[f0];
    double* o = out + (size_t)j * ldo + 3 * b0;
    if (!bnd_face_active(m, cmask, b0, c)) { o[0] = o[1] = o[2] = 0.0; return; }
    const double* u = U + (size_t)j * ldU;
    const double uc[3] = {u[3 * (size_t)c], u[3 * (size_t)c + 1], u[3 * (size_t)c + 2]};
    double g[3][3];
#pragma unroll
    for (int i = 0; i < 3; i++) g[i][0] = g[i][1] = g[i][2] = 0.0;
    const int e0 = m.off[c], e1 = m.off[c + 1];
    for (int e = e0; e < e1; e++) {
        const int en = m.ent[e], ot = m.other[e];
        const long long f = abs(en) - 1;
        const double s[3] = {m.Sf[3 * f], m.Sf[3 * f + 1], m.Sf[3 * f + 2]};
        if (ot >= 0) {
            const double w = m.w[f], w1 = 1.0 - w;
            const bool own = en > 0;
#pragma unroll
            for (int d = 0; d < 3; d++) {
                const double x = u[3 * (size_t)ot + d];
                const double uf = own ? (w * uc[d] + w1 * x) : (w * x + w1 * uc[d]);
#pragma unroll
                for (int i = 0; i < 3; i++) {
                    const double v = s[i] * uf;
                    g[i][d] += own ? v : -v;
                }
            }
        } else {
            const long long b = -(long long)ot - 1;
            if (!m.bmask[b]) continue;
            const double* ub = Ub + (size_t)j * ldUb + 3 * b;
#pragma unroll
            for (int d = 0; d < 3; d++)
#pragma unroll
                for (int i = 0; i < 3; i++) g[i][d] += s[i] * ub[d];
        }
    }
    const double V = m.V[c];
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int d = 0; d < 3; d++) g[i][d] /= V;
    if (m.limK > 0.0) {
        const double* ubj = Ub + (size_t)j * ldUb;
#pragma unroll
        for (int d = 0; d < 3; d++)
            md_limit(m, c, uc[d], [&](int o) { return u[3 * (size_t)o + d]; }, [&](long long b) { return ubj[3 * b + d]; }, g[0][d],
                     g[1][d], g[2][d]);
    }
    const double ms = m.magSf[f0];
    const double n[3] = {m.Sf[3 * f0] / ms, m.Sf[3 * f0 + 1] / ms, m.Sf[3 * f0 + 2] / ms};
    const double* ub = Ub + (size_t)j * ldUb + 3 * b0;
    const double dcb = m.dcB[b0];
#pragma unroll
    for (int d = 0; d < 3; d++) {
        const double sn = dcb * (ub[d] - uc[d]);
        const double ng = n[0] * g[0][d] + n[1] * g[1][d] + n[2] * g[2][d];
        const double corr = sn - ng;
#pragma unroll
        for (int i = 0; i < 3; i++) g[i][d] = g[i][d] + n[i] * corr;
    }
    o[0] = g[1][2] - g[2][1];
    o[1] = g[2][0] - g[0][2];
    o[2] = g[0][1] - g[1][0];
}

// Gb[i][3b+d] = (n x grad(P_i)_b) |Sf| on boundary face b
__global__ void __launch_bounds__(PROJ_T)
bnd_ncross_gradp_kernel(MeshDev m, const unsigned char* __restrict__ cmask, const double* __restrict__ P, long long ldP,
                        const double* __restrict__ Pb, long long ldPb, int Np, double* __restrict__ out, long long ldo)
{
    const long long b0 = blockIdx.x * (long long)PROJ_T + threadIdx.x;
    if (b0 >= m.nBnd) return;
    const int j = blockIdx.y;
    const long long f0 = m.nInt + b0;
    const int c = m.owner[f0];
    double* o = out + (size_t)j * ldo + 3 * b0;
    if (!bnd_face_active(m, cmask, b0, c)) { o[0] = o[1] = o[2] = 0.0; return; }
    const double* p = P + (size_t)j * ldP;
    const double pc = p[c];
    double g[3] = {0.0, 0.0, 0.0};
    const int e0 = m.off[c], e1 = m.off[c + 1];
    for (int e = e0; e < e1; e++) {
        const int en = m.ent[e], ot = m.other[e];
        const long long f = abs(en) - 1;
        const double s[3] = {m.Sf[3 * f], m.Sf[3 * f + 1], m.Sf[3 * f + 2]};
        if (ot >= 0) {
            const double w = m.w[f], w1 = 1.0 - w;
            const bool own = en > 0;
            const double x = p[ot];
            const double pf = own ? (w * pc + w1 * x) : (w * x + w1 * pc);
#pragma unroll
            for (int d = 0; d < 3; d++) {
                const double v = s[d] * pf;
                g[d] += own ? v : -v;
            }
        } else {
            const long long b = -(long long)ot - 1;
            if (!m.bmask[b]) continue;
            const double pb = Pb[(size_t)j * ldPb + b];
#pragma unroll
            for (int d = 0; d < 3; d++) g[d] += s[d] * pb;
        }
    }
    const double V = m.V[c];
#pragma unroll
    for (int d = 0; d < 3; d++) g[d] /= V;
    if (m.limK > 0.0) {
        const double* pbj = Pb + (size_t)j * ldPb;
        md_limit(m, c, pc, [&](int o) { return p[o]; }, [&](long long b) { return pbj[b]; }, g[0], g[1], g[2]);
    }
    const double ms = m.magSf[f0];
    const double n[3] = {m.Sf[3 * f0] / ms, m.Sf[3 * f0 + 1] / ms, m.Sf[3 * f0 + 2] / ms};
    const double sn = m.dcB[b0] * (Pb[(size_t)j * ldPb + b0] - pc);
    const double ng = n[0] * g[0] + n[1] * g[1] + n[2] * g[2];
#pragma unroll
    for (int d = 0; d < 3; d++) g[d] = g[d] + n[d] * (sn - ng);
    o[0] = (n[1] * g[2] - n[2] * g[1]) * ms;
    o[1] = (n[2] * g[0] - n[0] * g[2]) * ms;
    o[2] = (n[0] * g[1] - n[1] * g[0]) * ms;
}

// ------------------------------------------------------------------------------------------------
// K7 cell-major: all N laplacians in one pass over the mesh. The basis is held cell-major, Ucm[c][d][NP] (fields
// innermost, the layout of conv.cu), one thread = (cell, field): the lanes of a warp read 8*NP-byte rows (coalesced)
// and share every mesh scalar (one broadcast load), so the addressing and geometry are fetched once per cell for all
// fields instead of once per field. Same arithmetic, in the same order, as grad_vec_kernel / laplacian_kernel.
// ------------------------------------------------------------------------------------------------
constexpr int CM_CELLS = 4;   // cells per block (threadIdx.y); threadIdx.x = 64 lanes over the fields

// out[r*NP + j] = in[j*ld + r] (j < N, else 0): field-major columns -> cell-major rows. Block (32, 8), 32x32 tiles.
__global__ void __launch_bounds__(256)
tr_fm2cm_kernel(const double* __restrict__ in, long long ld, long long rows, int N, int NP, double* __restrict__ out)
{
    __shared__ double tile[32][33];
    const long long r0 = (long long)blockIdx.x * 32;
    for (int j0 = 0; j0 < NP; j0 += 32) {
        for (int jj = threadIdx.y; jj < 32; jj += 8) {
            const int j = j0 + jj;
            const long long r = r0 + threadIdx.x;
            tile[jj][threadIdx.x] = (j < N && r < rows) ? in[(size_t)j * ld + r] : 0.0;
        }
        __syncthreads();
        for (int rr = threadIdx.y; rr < 32; rr += 8) {
            const int j = j0 + threadIdx.x;
            const long long r = r0 + rr;
            if (j < NP && r < rows) out[(size_t)r * NP + j] = tile[threadIdx.x][rr];
        }
        __syncthreads();
    }
}
// out[j*ldo + r] = in[r*NP + j] for j < N: cell-major rows -> field-major columns.
__global__ void __launch_bounds__(256)
tr_cm2fm_kernel(const double* __restrict__ in, long long rows, int NP, int N, double* __restrict__ out, long long ldo)
{
    __shared__ double tile[32][33];
    const long long r0 = (long long)blockIdx.x * 32;
    for (int j0 = 0; j0 < N; j0 += 32) {
        for (int rr = threadIdx.y; rr < 32; rr += 8) {
            const int j = j0 + threadIdx.x;
            const long long r = r0 + rr;
            tile[rr][threadIdx.x] = (j < NP && r < rows) ? in[(size_t)r * NP + j] : 0.0;
        }
        __syncthreads();
        for (int jj = threadIdx.y; jj < 32; jj += 8) {
            const int j = j0 + jj;
            const long long r = r0 + threadIdx.x;
            if (j < N && r < rows) out[(size_t)j * ldo + r] = tile[threadIdx.x][jj];
        }
        __syncthreads();
    }
}

// Tcm[c][3i+d][NP] = fvc::grad(U_j)[c][i][d] for every field j (lane)
__global__ void __launch_bounds__(64 * CM_CELLS)
grad_cm_kernel(MeshDev m, const double* __restrict__ Ucm, const double* __restrict__ Ubcm, int NP, double* __restrict__ Tcm)
{
    const long long c = (long long)blockIdx.x * CM_CELLS + threadIdx.y;
    if (c >= m.nCells) return;
    const int e0 = m.off[c], e1 = m.off[c + 1];
    const double V = m.V[c];
    for (int j = threadIdx.x; j < NP; j += 64) {
        const double* uo = Ucm + (size_t)c * 3 * NP + j;
        const double uc[3] = {uo[0], uo[NP], uo[2 * NP]};
        double acc[3][3];
#pragma unroll
        for (int i = 0; i < 3; i++) acc[i][0] = acc[i][1] = acc[i][2] = 0.0;
        for (int e = e0; e < e1; e++) {
            const int en = m.ent[e], ot = m.other[e];
            const long long f = abs(en) - 1;
            const double s[3] = {m.Sf[3 * f], m.Sf[3 * f + 1], m.Sf[3 * f + 2]};
            if (ot >= 0) {
                const double w = m.w[f], w1 = 1.0 - w;
                const bool own = en > 0;
                const double* un = Ucm + (size_t)ot * 3 * NP + j;
#pragma unroll
                for (int d = 0; d < 3; d++) {
                    const double x = un[d * NP];
                    const double uf = own ? (w * uc[d] + w1 * x) : (w * x + w1 * uc[d]);
#pragma unroll
                    for (int i = 0; i < 3; i++) {
                        const double v = s[i] * uf;
                        acc[i][d] += own ? v : -v;
                    }
                }
            } else {
                const long long b = -(long long)ot - 1;
                if (!m.bmask[b]) continue;
                const double* ub = Ubcm + (size_t)b * 3 * NP + j;
#pragma unroll
                for (int d = 0; d < 3; d++) {
                    const double x = ub[d * NP];
#pragma unroll
                    for (int i = 0; i < 3; i++) acc[i][d] += s[i] * x;
                }
            }
        }
        double* t = Tcm + (size_t)c * 9 * NP + j;
#pragma unroll
        for (int i = 0; i < 3; i++)
#pragma unroll
            for (int d = 0; d < 3; d++) t[(3 * i + d) * NP] = acc[i][d] / V;
    }
}

// Lcm[c][d][NP] = fvc::laplacian(1, U_j)[c][d] for every field j (lane); Tcm = nullptr: uncorrected
__global__ void __launch_bounds__(64 * CM_CELLS)
laplacian_cm_kernel(MeshDev m, const double* __restrict__ Ucm, const double* __restrict__ Ubcm, int NP,
                    const double* __restrict__ Tcm, double* __restrict__ Lcm)
{
    const long long c = (long long)blockIdx.x * CM_CELLS + threadIdx.y;
    if (c >= m.nCells) return;
    const int e0 = m.off[c], e1 = m.off[c + 1];
    const double V = m.V[c];
    for (int j = threadIdx.x; j < NP; j += 64) {
        const double* uo = Ucm + (size_t)c * 3 * NP + j;
        const double uc[3] = {uo[0], uo[NP], uo[2 * NP]};
        double tc[9];
        if (Tcm) {
            const double* t = Tcm + (size_t)c * 9 * NP + j;
#pragma unroll
            for (int k = 0; k < 9; k++) tc[k] = t[k * NP];
        }
        double acc[3] = {0.0, 0.0, 0.0};
        for (int e = e0; e < e1; e++) {
            const int en = m.ent[e], ot = m.other[e];
            const long long f = abs(en) - 1;
            const double ms = m.magSf[f];
            if (ot >= 0) {
                const bool own = en > 0;
                const double dc = m.dc[f];
                double cr[3] = {0.0, 0.0, 0.0};
                if (Tcm) {
                    const double w = m.w[f], w1 = 1.0 - w;
                    const double* tn = Tcm + (size_t)ot * 9 * NP + j;
#pragma unroll
                    for (int i = 0; i < 3; i++) {
                        const double k = m.corr[3 * f + i];
#pragma unroll
                        for (int d = 0; d < 3; d++) {
                            const double x = tn[(3 * i + d) * NP];
                            cr[d] += k * (own ? (w * tc[3 * i + d] + w1 * x) : (w * x + w1 * tc[3 * i + d]));
                        }
                    }
                }
                const double* un = Ucm + (size_t)ot * 3 * NP + j;
#pragma unroll
                for (int d = 0; d < 3; d++) {
                    const double x = un[d * NP];
                    double sn = own ? dc * (x - uc[d]) : dc * (uc[d] - x);
                    if (Tcm) sn += cr[d];
                    const double v = ms * sn;
                    acc[d] += own ? v : -v;
                }
            } else {
                const long long b = -(long long)ot - 1;
                if (!m.bmask[b]) continue;
                const double* ub = Ubcm + (size_t)b * 3 * NP + j;
                const double dcb = m.dcB[b];
#pragma unroll
                for (int d = 0; d < 3; d++) acc[d] += ms * (dcb * (ub[d * NP] - uc[d]));
            }
        }
        double* o = Lcm + (size_t)c * 3 * NP + j;
#pragma unroll
        for (int d = 0; d < 3; d++) o[d * NP] = acc[d] / V;
    }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
// shard bookkeeping: caller-order cell c is owned iff c < nOwned. Internal order: cmask[perm[c]], Vw = V on owned cells else 0.
__global__ void owned_mask_kernel(const int* __restrict__ perm, long long nCells, long long nOwned, const double* __restrict__ V,
                                  double* __restrict__ Vw, unsigned char* __restrict__ cmask)
{
    for (long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x; c < nCells; c += (long long)gridDim.x * blockDim.x) {
        const int n = perm[c];
        const bool own = c < nOwned;
        cmask[n] = own ? 1 : 0;
        Vw[n] = own ? V[n] : 0.0;
    }
}

__global__ void negate_kernel(double* __restrict__ x, long long n)
{
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i < n) x[i] = -x[i];
}

static inline long long pad16(long long n) { return (n + 15) / 16 * 16; }
static inline unsigned gx(long long n) { return (unsigned)((n + PROJ_T - 1) / PROJ_T); }
static inline unsigned gy(int n) { return (unsigned)((n + FB - 1) / FB); }

struct StageTimer {
    cudaEvent_t a, b;
    cudaStream_t s;
    double* acc;
    StageTimer(cudaStream_t st, double* where) : s(st), acc(where)
    {
        cudaEventCreate(&a);
        cudaEventCreate(&b);
        cudaEventRecord(a, s);
    }
    void stop()
    {
        cudaEventRecord(b, s);
        cudaEventSynchronize(b);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, a, b);
        *acc += ms;
        cudaEventDestroy(a);
        cudaEventDestroy(b);
    }
};

static int build_csr(ithaca_b200* h, ProjState* s, const int* owner, const int* neigh, DevBuf& off, DevBuf& ent, DevBuf& other)
{
    const long long nC = s->nCells, nF = s->nInt + s->nBnd, nnz = 2 * s->nInt + s->nBnd, n1 = nC + 1;
    int rc;
    if ((rc = ensure(h, s->cnt, n1 * sizeof(int)))) return rc;
    if ((rc = ensure(h, off, n1 * sizeof(int)))) return rc;
    if ((rc = ensure(h, ent, std::max<long long>(nnz, 1) * sizeof(int)))) return rc;
    if ((rc = ensure(h, other, std::max<long long>(nnz, 1) * sizeof(int)))) return rc;
    const int nb = (int)((n1 + SCAN_B - 1) / SCAN_B);
    if ((rc = ensure(h, s->blocksum, (size_t)nb * sizeof(int)))) return rc;
    IB_CUDA(h, cudaMemsetAsync(s->cnt.p, 0, n1 * sizeof(int), h->stream));
    const int fg = (int)std::min<long long>((nF + 255) / 256, 64LL * h->num_sms);
    csr_count_kernel<<<std::max(fg, 1), 256, 0, h->stream>>>(owner, neigh, s->nInt, nF, s->cnt.as<int>());
    scan_block_kernel<<<nb, SCAN_B, 0, h->stream>>>(s->cnt.as<int>(), n1, off.as<int>(), s->blocksum.as<int>());
    scan_sums_kernel<<<1, SCAN_B, 0, h->stream>>>(s->blocksum.as<int>(), nb);
    scan_add_kernel<<<(unsigned)((n1 + 255) / 256), 256, 0, h->stream>>>(off.as<int>(), s->blocksum.as<int>(), n1, s->cnt.as<int>(), 0);
    csr_fill_kernel<<<std::max(fg, 1), 256, 0, h->stream>>>(owner, neigh, s->nInt, nF, s->cnt.as<int>(), ent.as<int>(), other.as<int>());
    csr_sort_kernel<<<(unsigned)((nC + 255) / 256), 256, 0, h->stream>>>(off.as<int>(), nC, ent.as<int>(), other.as<int>());
    IB_CUDA(h, cudaGetLastError());
    h->launches += 6;
    return IB200_OK;
}

static int need_flux(ithaca_b200* h, ProjState* s)
{
    if (s->have_flux) return IB200_OK;
    int rc = ensure(h, s->flux, (size_t)s->nU * s->ldF * sizeof(double));
    if (rc) return rc;
    StageTimer t(h->stream, &s->ms[0]);
    face_flux_kernel<<<dim3(gx(s->nInt + s->nBnd), gy(s->nU)), PROJ_T, 0, h->stream>>>(
        mesh_dev(s), s->U.as<double>(), s->ldU, s->Ub.as<double>(), s->ldUb, s->nU, s->flux.as<double>(), s->ldF);
    t.stop();
    IB_CUDA(h, cudaGetLastError());
    h->launches += 1;
    s->have_flux = true;
    return IB200_OK;
}

// the convecting flux of mode j in fvc::div(phi_j, U_k): the caller's (fluxMethod "consistent", steadyNS.C:1010-1015) or the computed one
static inline const double* conv_flux(const ProjState* s) { return s->flux_user ? s->fluxUser.as<double>() : s->flux.as<double>(); }

// grad(U_k) of ALL basis fields for the linearUpwind correction (limited when the cellMDLimited gradient is selected, as the
// scheme's "grad(U)" argument names the case's gradient scheme)
static int need_gradU_all(ithaca_b200* h, ProjState* s)
{
    if (s->convScheme != 2 || s->have_Tall) return IB200_OK;
    const long long ldT = 9 * pad16(s->nCells);
    int rc = ensure(h, s->Tall, (size_t)s->nU * ldT * sizeof(double));
    if (rc) return rc;
    const MeshDev m = mesh_dev(s);
    if (s->limK > 0.0)
        grad_vec_kernel<true><<<gx(s->nCells) * s->nU, PROJ_T, 0, h->stream>>>(m, s->U.as<double>(), s->ldU, s->Ub.as<double>(), s->ldUb, 0, s->nU,
                                                                                s->Tall.as<double>(), ldT);
    else
        grad_vec_kernel<false><<<gx(s->nCells) * s->nU, PROJ_T, 0, h->stream>>>(m, s->U.as<double>(), s->ldU, s->Ub.as<double>(), s->ldUb, 0, s->nU,
                                                                                 s->Tall.as<double>(), ldT);
    h->launches += 1;
    IB_CUDA(h, cudaGetLastError());
    s->have_Tall = true;
    return IB200_OK;
}
// one launch of K (div_phi_kernel or its boundary-owner twin) with the convection scheme of the mesh
#define LAUNCH_DIV_PHI(K, GRID, ...)                                                                                              \
    do {                                                                                                                          \
        int rc_ldp_ = need_gradU_all(h, s);                                                                                       \
        if (rc_ldp_) return rc_ldp_;                                                                                              \
        const long long ldT_ldp_ = 9 * pad16(s->nCells);                                                                          \
        if (s->convScheme == 2) K<2><<<GRID, PROJ_T, 0, h->stream>>>(__VA_ARGS__, s->Tall.as<double>(), ldT_ldp_);                      \
        else if (s->convScheme == 1) K<1><<<GRID, PROJ_T, 0, h->stream>>>(__VA_ARGS__, nullptr, 0);                                     \
        else K<0><<<GRID, PROJ_T, 0, h->stream>>>(__VA_ARGS__, nullptr, 0);                                                       \
    } while (0)

// laplacian(1, U_j) for j in [0, nU) into `out` (columns of ldU doubles)
static int run_laplacians(ithaca_b200* h, ProjState* s, double* out)
{
    const MeshDev m = mesh_dev(s);
    // ITHACA_B200_LAP=cm selects the cell-major pass (all fields per cell, bitwise the same results). Measured at C4
    // (2M cells, N = 50): 23.8 ms against 16.2 ms for the field-major kernels below -- both are bound by the latency of
    // the per-face dependent loads, and the field-major blocks (64 consecutive cells of one field) reuse neighbour
    // rows in L1 -- so field-major stays the default.
    {
        const int N = s->nU, NP = (N + 3) / 8 * 8 + 4;
        const size_t per = (size_t)NP * 8;
        const size_t need = per * ((size_t)s->nCells * (3 + 3 + (s->corrected ? 9 : 0)) + (size_t)std::max<long long>(s->nBnd, 1) * 3);
        size_t have = s->conv.Ucm.cap + s->conv.Ubcm.cap + s->Tcm.cap + s->Lcm.cap, fr = 0, tot = 0;
        cudaMemGetInfo(&fr, &tot);
        const char* ev = env_knob("ITHACA_B200_LAP");
        if (ev && strcmp(ev, "cm") == 0 && need <= have + fr / 2 && !(s->limK > 0.0) && s->snPsi < 0.0) {
            int rc;
            if ((rc = ensure(h, s->conv.Ucm, (size_t)s->nCells * 3 * per)) ||
                (rc = ensure(h, s->conv.Ubcm, (size_t)std::max<long long>(s->nBnd, 1) * 3 * per)) ||
                (s->corrected && (rc = ensure(h, s->Tcm, (size_t)s->nCells * 9 * per))) ||
                (rc = ensure(h, s->Lcm, (size_t)s->nCells * 3 * per)))
                return rc;
            StageTimer t(h->stream, &s->ms[2]);
            const long long rows = 3 * s->nCells, rowsB = 3 * s->nBnd;
            tr_fm2cm_kernel<<<(unsigned)((rows + 31) / 32), dim3(32, 8), 0, h->stream>>>(s->U.as<double>(), s->ldU, rows, N, NP,
                                                                                        s->conv.Ucm.as<double>());
            if (rowsB)
                tr_fm2cm_kernel<<<(unsigned)((rowsB + 31) / 32), dim3(32, 8), 0, h->stream>>>(s->Ub.as<double>(), s->ldUb, rowsB, N, NP,
                                                                                             s->conv.Ubcm.as<double>());
            const unsigned gc = (unsigned)((s->nCells + CM_CELLS - 1) / CM_CELLS);
            if (s->corrected)
                grad_cm_kernel<<<gc, dim3(64, CM_CELLS), 0, h->stream>>>(m, s->conv.Ucm.as<double>(), s->conv.Ubcm.as<double>(), NP,
                                                                         s->Tcm.as<double>());
            laplacian_cm_kernel<<<gc, dim3(64, CM_CELLS), 0, h->stream>>>(m, s->conv.Ucm.as<double>(), s->conv.Ubcm.as<double>(), NP,
                                                                          s->corrected ? s->Tcm.as<double>() : nullptr, s->Lcm.as<double>());
            tr_cm2fm_kernel<<<(unsigned)((rows + 31) / 32), dim3(32, 8), 0, h->stream>>>(s->Lcm.as<double>(), rows, NP, N, out, s->ldU);
            t.stop();
            h->launches += 3 + (rowsB ? 1 : 0) + (s->corrected ? 1 : 0);
            IB_CUDA(h, cudaGetLastError());
            return IB200_OK;
        }
    }
    int batch = 8;   // fields per launch (T holds batch gradient tensors); ITHACA_B200_LAP_BATCH overrides (tuning)
    if (const char* e = env_knob("ITHACA_B200_LAP_BATCH")) batch = std::max(1, std::min(atoi(e), 64));
    const long long ldT = 9 * pad16(s->nCells);
    if (s->corrected) {
        int rc = ensure(h, s->T, (size_t)batch * ldT * sizeof(double));
        if (rc) return rc;
    }
    const bool sch = s->limK > 0.0 || (s->corrected && s->snPsi >= 0.0);
    StageTimer t(h->stream, &s->ms[2]);
    for (int j0 = 0; j0 < s->nU; j0 += batch) {
        const int nb = std::min(batch, s->nU - j0);
        if (s->corrected) {
            if (sch)
                grad_vec_kernel<true><<<gx(s->nCells) * nb, PROJ_T, 0, h->stream>>>(m, s->U.as<double>(), s->ldU, s->Ub.as<double>(),
                                                                                     s->ldUb, j0, nb, s->T.as<double>(), ldT);
            else
                grad_vec_kernel<false><<<gx(s->nCells) * nb, PROJ_T, 0, h->stream>>>(m, s->U.as<double>(), s->ldU, s->Ub.as<double>(),
                                                                                      s->ldUb, j0, nb, s->T.as<double>(), ldT);
            h->launches += 1;
        }
        if (sch)
            laplacian_kernel<true><<<gx(s->nCells) * nb, PROJ_T, 0, h->stream>>>(m, s->U.as<double>(), s->ldU, s->Ub.as<double>(), s->ldUb, j0, nb,
                                                                                  s->corrected ? s->T.as<double>() : nullptr, ldT, out, s->ldU);
        else
            laplacian_kernel<false><<<gx(s->nCells) * nb, PROJ_T, 0, h->stream>>>(m, s->U.as<double>(), s->ldU, s->Ub.as<double>(), s->ldUb, j0, nb,
                                                                                   s->corrected ? s->T.as<double>() : nullptr, ldT, out, s->ldU);
        h->launches += 1;
    }
    t.stop();
    IB_CUDA(h, cudaGetLastError());
    return IB200_OK;
}

}  // namespace ib200

using namespace ib200;

#define PROJ_REQUIRE(cond, msg)                 \
    do {                                        \
        if (!(cond)) {                          \
            set_error(h, msg);                  \
            return IB200_ERR_ARG;               \
        }                                       \
    } while (0)

extern "C" {

int ithaca_b200_mesh_upload(ithaca_b200_t* h, int64_t nCells, int64_t nIntFaces, int64_t nBndFaces, const int32_t* owner,
                            const int32_t* neighbour, const double* Sf, const double* magSf, const double* weights,
                            const double* nonOrthDeltaCoeffs, const double* bndDeltaCoeffs, const double* nonOrthCorrVec,
                            const double* V, const unsigned char* bndMask)
{
    if (!h) return IB200_ERR_ARG;
    PROJ_REQUIRE(nCells > 0 && nIntFaces >= 0 && nBndFaces >= 0 && owner && (neighbour || nIntFaces == 0) && Sf && magSf &&
                     (weights || nIntFaces == 0) && (nonOrthDeltaCoeffs || nIntFaces == 0) && (bndDeltaCoeffs || nBndFaces == 0) && V,
                 "mesh_upload: bad argument");
    PROJ_REQUIRE(2 * nIntFaces + nBndFaces < 0x7fffffffLL && nCells < 0x7fffffffLL, "mesh_upload: mesh too large for 32-bit labels");
    cudaSetDevice(h->device);
    if (!h->proj) h->proj = new ProjState();
    ProjState* s = h->proj;
    s->have_mesh = false;
    s->have_flux = false; s->flux_user = false;
    h->vol_gen++;
    s->nCells = nCells; s->nInt = nIntFaces; s->nBnd = nBndFaces;
    const long long nF = nIntFaces + nBndFaces;
    int rc;
    // the caller's arrays: addressing is kept (mesh_addressing), geometry only until the renumbered copy is made
    DevBuf Sf0, magSf0, w0, dc0, corr0, V0;
    auto drop = [&]() { for (DevBuf* b : {&Sf0, &magSf0, &w0, &dc0, &corr0, &V0}) if (b->p) { cudaFree(b->p); b->p = nullptr; } };
    s->corrected = nonOrthCorrVec != nullptr;
    s->have_csr0 = false;
    if ((rc = upload(h, s->owner0, owner, nF * sizeof(int))) || (rc = upload(h, s->neigh0, neighbour, nIntFaces * sizeof(int))) ||
        (rc = upload(h, Sf0, Sf, 3 * nF * sizeof(double))) || (rc = upload(h, magSf0, magSf, nF * sizeof(double))) ||
        (rc = upload(h, w0, weights, nIntFaces * sizeof(double))) || (rc = upload(h, dc0, nonOrthDeltaCoeffs, nIntFaces * sizeof(double))) ||
        (s->corrected && (rc = upload(h, corr0, nonOrthCorrVec, 3 * nIntFaces * sizeof(double)))) ||
        (rc = upload(h, V0, V, nCells * sizeof(double))) || (rc = upload(h, s->dcB, bndDeltaCoeffs, nBndFaces * sizeof(double)))) {
        drop();
        return rc;
    }
    if ((rc = ensure(h, s->bmask, std::max<long long>(nBndFaces, 16)))) { drop(); return rc; }
    if (bndMask) IB_CUDA(h, cudaMemcpyAsync(s->bmask.p, bndMask, nBndFaces, cudaMemcpyDefault, h->stream));
    else IB_CUDA(h, cudaMemsetAsync(s->bmask.p, 1, std::max<long long>(nBndFaces, 16), h->stream));

    // locality renumbering on the host (renumber.cu); ITHACA_B200_NO_RENUMBER=1 keeps the caller's labels
    std::vector<int> old2new, fnew2old;
    {
        std::vector<int> ho((size_t)nIntFaces), hn((size_t)nIntFaces);
        IB_CUDA(h, cudaMemcpyAsync(ho.data(), s->owner0.p, nIntFaces * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
        IB_CUDA(h, cudaMemcpyAsync(hn.data(), s->neigh0.p, nIntFaces * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
        IB_CUDA(h, cudaStreamSynchronize(h->stream));
        bool bad = false;
        for (long long f = 0; f < nIntFaces && !bad; f++)
            bad = ho[f] < 0 || ho[f] >= nCells || hn[f] < 0 || hn[f] >= nCells;
        if (bad) { drop(); set_error(h, "mesh_upload: owner/neighbour label out of range"); return IB200_ERR_ARG; }
        const char* off_env = env_knob("ITHACA_B200_NO_RENUMBER");
        if (off_env && *off_env == '1') {
            old2new.resize((size_t)nCells);
            fnew2old.resize((size_t)nIntFaces);
            for (long long c = 0; c < nCells; c++) old2new[c] = (int)c;
            for (long long f = 0; f < nIntFaces; f++) fnew2old[f] = (int)f;
        } else {
            // ITHACA_B200_RENUMBER_BLOCK: cells per compact cluster of the two-level ordering (renumber.cu); 0 = plain
            // Cuthill-McKee, the default: measured at C4, clusters of 64..4096 cells make every gather stage SLOWER
            // (laplacian 15.7 -> 17.8..25.2 ms): inside a breadth-first ball the neighbours of consecutive cells are no
            // longer consecutive, and coalescing matters more than L1 reuse
            int blk = 0;
            if (const char* e = env_knob("ITHACA_B200_RENUMBER_BLOCK")) blk = std::max(0, atoi(e));
            if (blk > 1) mesh_renumber_blocked(nCells, nIntFaces, ho.data(), hn.data(), blk, old2new, fnew2old);
            else mesh_renumber(nCells, nIntFaces, ho.data(), hn.data(), old2new, fnew2old);
        }
    }
    if ((rc = upload(h, s->perm, old2new.data(), nCells * sizeof(int))) ||
        (rc = upload(h, s->fperm, fnew2old.data(), nIntFaces * sizeof(int))) ||
        (rc = ensure(h, s->owner, nF * sizeof(int))) || (rc = ensure(h, s->neigh, std::max<long long>(nIntFaces, 4) * sizeof(int))) ||
        (rc = ensure(h, s->Sf, 3 * nF * sizeof(double))) || (rc = ensure(h, s->magSf, nF * sizeof(double))) ||
        (rc = ensure(h, s->w, std::max<long long>(nIntFaces, 2) * sizeof(double))) ||
        (rc = ensure(h, s->dc, std::max<long long>(nIntFaces, 2) * sizeof(double))) ||
        (s->corrected && (rc = ensure(h, s->corr, 3 * std::max<long long>(nIntFaces, 1) * sizeof(double)))) ||
        (rc = ensure(h, s->V, nCells * sizeof(double))) ||
        (rc = ensure(h, s->stage, (size_t)std::max<long long>(3 * nCells, nF) * sizeof(double)))) {
        drop();
        return rc;
    }
    {
        MeshSrc src;
        src.owner = s->owner0.as<int>(); src.neigh = s->neigh0.as<int>();
        src.Sf = Sf0.as<double>(); src.magSf = magSf0.as<double>(); src.w = w0.as<double>(); src.dc = dc0.as<double>();
        src.corr = s->corrected ? corr0.as<double>() : nullptr;
        const int fg = (int)std::min<long long>((nF + 255) / 256, 64LL * h->num_sms);
        permute_faces_kernel<<<std::max(fg, 1), 256, 0, h->stream>>>(src, s->perm.as<int>(), s->fperm.as<int>(), nIntFaces, nF,
                                                                     s->owner.as<int>(), s->neigh.as<int>(), s->Sf.as<double>(),
                                                                     s->magSf.as<double>(), s->w.as<double>(), s->dc.as<double>(),
                                                                     s->corrected ? s->corr.as<double>() : nullptr);
        const int cg = (int)std::min<long long>((nCells + 255) / 256, 64LL * h->num_sms);
        permute_cells_kernel<<<std::max(cg, 1), 256, 0, h->stream>>>(V0.as<double>(), s->perm.as<int>(), nCells, 1, s->V.as<double>());
        h->launches += 2;
    }
    rc = build_csr(h, s, s->owner.as<int>(), s->neigh.as<int>(), s->off, s->ent, s->other);
    cudaError_t e = cudaStreamSynchronize(h->stream);
    drop();
    if (rc) return rc;
    IB_CUDA(h, e);
    IB_CUDA(h, cudaGetLastError());
    s->have_mesh = true;
    s->nOwned = nCells;
    s->nU = s->nP = 0;
    s->limK = 0.0;     // a new mesh starts with the "Gauss linear" / "corrected" scheme set
    s->snPsi = -1.0;
    s->convScheme = 0;
    s->have_centres = s->have_Tall = false;
    return IB200_OK;
}

int ithaca_b200_mesh_set_convection(ithaca_b200_t* h, int scheme)
{
    if (!h) return IB200_ERR_ARG;
    ProjState* s = h->proj;
    PROJ_REQUIRE(s && s->have_mesh, "mesh_set_convection: no mesh uploaded");
    PROJ_REQUIRE(scheme >= IB200_CONV_LINEAR && scheme <= IB200_CONV_LINEAR_UPWIND, "mesh_set_convection: unknown interpolation scheme");
    PROJ_REQUIRE(scheme != IB200_CONV_LINEAR_UPWIND || s->have_centres,
                 "mesh_set_convection: linearUpwind needs mesh.C() and mesh.Cf() (pass them to mesh_set_schemes first)");
    s->convScheme = scheme;
    s->have_Tall = false;
    return IB200_OK;
}

int ithaca_b200_mesh_set_schemes(ithaca_b200_t* h, double grad_limit_k, double sngrad_psi, const double* C, const double* Cf)
{
    if (!h) return IB200_ERR_ARG;
    ProjState* s = h->proj;
    PROJ_REQUIRE(s && s->have_mesh, "mesh_set_schemes: no mesh uploaded");
    PROJ_REQUIRE(!(grad_limit_k > 1.0), "mesh_set_schemes: cellMDLimited coefficient k = " + std::to_string(grad_limit_k) + " should be >= 0 and <= 1");
    PROJ_REQUIRE(!(sngrad_psi > 1.0), "mesh_set_schemes: limited snGrad coefficient = " + std::to_string(sngrad_psi) + " should be >= 0 and <= 1");
    PROJ_REQUIRE(!(sngrad_psi >= 0.0) || s->corrected, "mesh_set_schemes: limited snGrad needs the nonOrthCorrectionVectors (mesh_upload got NULL)");
    // OpenFOAM returns the unlimited gradient when k < SMALL (cellMDLimitedGrads.C)
    const bool lim = grad_limit_k >= 1e-15;
    PROJ_REQUIRE(!lim || (C && Cf), "mesh_set_schemes: the cellMDLimited gradient needs mesh.C() and mesh.Cf()");
    cudaSetDevice(h->device);
    s->have_Tall = false;
    if (C && Cf) {
        const long long nC = s->nCells, nF = s->nInt + s->nBnd, nnz = 2 * s->nInt + s->nBnd;
        int rc;
        DevBuf C0, Cf0;
        auto drop = [&]() { for (DevBuf* b : {&C0, &Cf0}) if (b->p) { cudaFree(b->p); b->p = nullptr; } };
        if ((rc = upload(h, C0, C, 3 * nC * sizeof(double))) || (rc = upload(h, Cf0, Cf, 3 * nF * sizeof(double))) ||
            (rc = ensure(h, s->Cc, 3 * nC * sizeof(double))) || (rc = ensure(h, s->Cf, 3 * std::max<long long>(nF, 1) * sizeof(double))) ||
            (rc = ensure(h, s->entL, std::max<long long>(nnz, 1) * sizeof(int))) ||
            (rc = ensure(h, s->otherL, std::max<long long>(nnz, 1) * sizeof(int)))) {
            drop();
            return rc;
        }
        const int cg = (int)std::min<long long>((nC + 255) / 256, 64LL * h->num_sms);
        const int fg = (int)std::min<long long>((nF + 255) / 256, 64LL * h->num_sms);
        permute_cells_kernel<<<std::max(cg, 1), 256, 0, h->stream>>>(C0.as<double>(), s->perm.as<int>(), nC, 3, s->Cc.as<double>());
        permute_face_vectors_kernel<<<std::max(fg, 1), 256, 0, h->stream>>>(Cf0.as<double>(), s->fperm.as<int>(), s->nInt, nF, s->Cf.as<double>());
        cudaMemcpyAsync(s->entL.p, s->ent.p, nnz * sizeof(int), cudaMemcpyDeviceToDevice, h->stream);
        cudaMemcpyAsync(s->otherL.p, s->other.p, nnz * sizeof(int), cudaMemcpyDeviceToDevice, h->stream);
        csr_sort_caller_order_kernel<<<(unsigned)((nC + 255) / 256), 256, 0, h->stream>>>(s->off.as<int>(), nC, s->fperm.as<int>(), s->nInt,
                                                                                         s->entL.as<int>(), s->otherL.as<int>());
        h->launches += 3;
        cudaError_t e = cudaStreamSynchronize(h->stream);
        drop();
        IB_CUDA(h, e);
        IB_CUDA(h, cudaGetLastError());
        s->have_centres = true;
    }
    s->limK = lim ? grad_limit_k : 0.0;
    s->snPsi = sngrad_psi >= 0.0 ? sngrad_psi : -1.0;
    return IB200_OK;
}

int ithaca_b200_mesh_set_owned(ithaca_b200_t* h, int64_t nOwned)
{
    if (!h) return IB200_ERR_ARG;
    ProjState* s = h->proj;
    PROJ_REQUIRE(s && s->have_mesh, "mesh_set_owned: no mesh uploaded");
    PROJ_REQUIRE(nOwned >= 0 && nOwned <= s->nCells, "mesh_set_owned: nOwned out of range");
    cudaSetDevice(h->device);
    s->nOwned = nOwned;
    h->vol_gen++;
    if (nOwned == s->nCells) return IB200_OK;
    int rc;
    if ((rc = ensure(h, s->Vw, s->nCells * sizeof(double))) || (rc = ensure(h, s->cmask, std::max<long long>(s->nCells, 16)))) return rc;
    const int g = (int)std::min<long long>((s->nCells + 255) / 256, 64LL * h->num_sms);
    owned_mask_kernel<<<std::max(g, 1), 256, 0, h->stream>>>(s->perm.as<int>(), s->nCells, nOwned, s->V.as<double>(), s->Vw.as<double>(),
                                                             s->cmask.as<unsigned char>());
    h->launches += 1;
    IB_CUDA(h, cudaGetLastError());
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    return IB200_OK;
}

int ithaca_b200_mesh_addressing(ithaca_b200_t* h, int32_t* off, int32_t* ent, int32_t* other, uint64_t* digest)
{
    if (!h) return IB200_ERR_ARG;
    ProjState* s = h->proj;
    PROJ_REQUIRE(s && s->have_mesh, "mesh_addressing: call mesh_upload first");
    cudaSetDevice(h->device);
    const long long n1 = s->nCells + 1, nnz = 2 * s->nInt + s->nBnd;
    if (!s->have_csr0) {   // addressing of the caller's own arrays (the kernels run on the renumbered copy)
        int rc = build_csr(h, s, s->owner0.as<int>(), s->neigh0.as<int>(), s->off0, s->ent0, s->other0);
        if (rc) return rc;
        s->have_csr0 = true;
    }
    std::vector<int> o(n1), e(nnz), t(nnz);
    IB_CUDA(h, cudaMemcpyAsync(o.data(), s->off0.p, n1 * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    IB_CUDA(h, cudaMemcpyAsync(e.data(), s->ent0.p, nnz * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    IB_CUDA(h, cudaMemcpyAsync(t.data(), s->other0.p, nnz * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    if (off) memcpy(off, o.data(), n1 * sizeof(int));
    if (ent) memcpy(ent, e.data(), nnz * sizeof(int));
    if (other) memcpy(other, t.data(), nnz * sizeof(int));
    if (digest) {  // FNV-1a over the three arrays
        uint64_t d = 1469598103934665603ULL;
        auto mix = [&](const std::vector<int>& v) {
            const unsigned char* p = reinterpret_cast<const unsigned char*>(v.data());
            for (size_t i = 0; i < v.size() * sizeof(int); i++) { d ^= p[i]; d *= 1099511628211ULL; }
        };
        mix(o); mix(e); mix(t);
        *digest = d;
    }
    return IB200_OK;
}

int ithaca_b200_basis_upload(ithaca_b200_t* h, int which, const double* const* int_ptrs, const double* const* bnd_ptrs, int n)
{
    if (!h) return IB200_ERR_ARG;
    ProjState* s = h->proj;
    PROJ_REQUIRE(s && s->have_mesh, "basis_upload: call mesh_upload first");
    PROJ_REQUIRE((which == 0 || which == 1) && n > 0 && int_ptrs && (bnd_ptrs || s->nBnd == 0), "basis_upload: bad argument");
    cudaSetDevice(h->device);
    const int dim = which == 0 ? 3 : 1;
    const long long ld = pad16(dim * s->nCells), ldb = pad16(std::max<long long>(dim * s->nBnd, 1));
    DevBuf& I = which == 0 ? s->U : s->P;
    DevBuf& Bd = which == 0 ? s->Ub : s->Pb;
    int rc;
    if ((rc = ensure(h, I, (size_t)n * ld * sizeof(double)))) return rc;
    if ((rc = ensure(h, Bd, (size_t)n * ldb * sizeof(double)))) return rc;
    for (int j = 0; j < n; j++) {
        PROJ_REQUIRE(int_ptrs[j] && (s->nBnd == 0 || bnd_ptrs[j]), "basis_upload: null field pointer");
        IB_CUDA(h, cudaMemcpyAsync(s->stage.p, int_ptrs[j], (size_t)dim * s->nCells * sizeof(double), cudaMemcpyDefault, h->stream));
        permute_cells_kernel<<<(unsigned)std::min<long long>((dim * s->nCells + 255) / 256, 64LL * h->num_sms), 256, 0, h->stream>>>(
            s->stage.as<double>(), s->perm.as<int>(), s->nCells, dim, I.as<double>() + (size_t)j * ld);
        h->launches += 1;
        if (s->nBnd)
            IB_CUDA(h, cudaMemcpyAsync(Bd.as<double>() + (size_t)j * ldb, bnd_ptrs[j], (size_t)dim * s->nBnd * sizeof(double),
                                       cudaMemcpyDefault, h->stream));
    }
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    if (which == 0) { s->nU = n; s->ldU = ld; s->ldUb = ldb; s->ldF = pad16(s->nInt + s->nBnd); s->have_flux = false; s->flux_user = false; s->have_Tall = false; }
    else { s->nP = n; s->ldP = ld; s->ldPb = ldb; }
    return IB200_OK;
}

// phi in the caller's face order -> the kernels' order (internal face at new position f sat at fperm[f]); 0 on `empty` faces
__global__ void permute_flux_in_kernel(const double* __restrict__ in, const int* __restrict__ fperm, const unsigned char* __restrict__ bmask,
                                       long long nInt, long long nF, double* __restrict__ out)
{
    for (long long f = blockIdx.x * (long long)blockDim.x + threadIdx.x; f < nF; f += (long long)gridDim.x * blockDim.x)
        out[f] = (f < nInt) ? in[fperm[f]] : (bmask[f - nInt] ? in[f] : 0.0);
}

int ithaca_b200_flux_upload(ithaca_b200_t* h, const double* const* phi_ptrs, int n)
{
    if (!h) return IB200_ERR_ARG;
    ProjState* s = h->proj;
    PROJ_REQUIRE(s && s->have_mesh && s->nU > 0, "flux_upload: call mesh_upload and basis_upload(U) first");
    cudaSetDevice(h->device);
    if (!phi_ptrs) { s->flux_user = false; return IB200_OK; }   // back to linearInterpolate(U_j) & Sf
    PROJ_REQUIRE(n == s->nU, "flux_upload: one flux field per velocity basis field (L_PHImodes has the size of L_U_SUPmodes)");
    const long long nF = s->nInt + s->nBnd;
    int rc = ensure(h, s->fluxUser, (size_t)s->nU * s->ldF * sizeof(double));
    if (rc) return rc;
    for (int j = 0; j < n; j++) {
        PROJ_REQUIRE(phi_ptrs[j], "flux_upload: null field pointer");
        IB_CUDA(h, cudaMemcpyAsync(s->stage.p, phi_ptrs[j], (size_t)nF * sizeof(double), cudaMemcpyDefault, h->stream));
        permute_flux_in_kernel<<<(unsigned)std::min<long long>((nF + 255) / 256, 64LL * h->num_sms), 256, 0, h->stream>>>(
            s->stage.as<double>(), s->fperm.as<int>(), mesh_dev(s).bmask, s->nInt, nF, s->fluxUser.as<double>() + (size_t)j * s->ldF);
        h->launches += 1;
    }
    IB_CUDA(h, cudaGetLastError());
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    s->flux_user = true;
    return IB200_OK;
}

}  // extern "C"

extern "C" {

// Materialised operator field of basis function j (and k), for the parity tests of K5-K8.
int ithaca_b200_fvc_eval(ithaca_b200_t* h, int op, int j, int k, double* out)
{
    if (!h) return IB200_ERR_ARG;
    ProjState* s = h->proj;
    PROJ_REQUIRE(s && s->have_mesh && out, "fvc_eval: call mesh_upload and basis_upload first");
    cudaSetDevice(h->device);
    const MeshDev m = mesh_dev(s);
    int rc;
    size_t nout = 0;
    auto ug = [&](size_t n) { return (unsigned)std::min<size_t>((n + 255) / 256, 64 * (size_t)h->num_sms); };
    switch (op) {
    case IB200_FVC_FLUX: {
        PROJ_REQUIRE(j >= 0 && j < s->nU, "fvc_eval: field index out of range");
        if ((rc = need_flux(h, s))) return rc;
        nout = (size_t)(s->nInt + s->nBnd);
        unpermute_faces_kernel<<<ug(nout), 256, 0, h->stream>>>(s->flux.as<double>() + (size_t)j * s->ldF, s->fperm.as<int>(), s->nInt,
                                                                 s->nInt + s->nBnd, s->stage.as<double>());
        break;
    }
    case IB200_FVC_DIV_PHI: {
        PROJ_REQUIRE(j >= 0 && j < s->nU && k >= 0 && k < s->nU, "fvc_eval: field index out of range");
        if ((rc = need_flux(h, s))) return rc;
        if ((rc = ensure(h, s->work, (size_t)s->nU * s->ldU * sizeof(double)))) return rc;
        LAUNCH_DIV_PHI(div_phi_kernel, dim3(gx(s->nCells), gy(s->nU)), 
            m, conv_flux(s) + (size_t)j * s->ldF, s->U.as<double>(), s->ldU, s->Ub.as<double>(), s->ldUb, s->nU,
            s->work.as<double>(), s->ldU);
        h->launches += 1;
        nout = (size_t)3 * s->nCells;
        unpermute_cells_kernel<<<ug(nout), 256, 0, h->stream>>>(s->work.as<double>() + (size_t)k * s->ldU, s->perm.as<int>(), s->nCells, 3,
                                                                 s->stage.as<double>());
        break;
    }
    case IB200_FVC_LAPLACIAN: {
        PROJ_REQUIRE(j >= 0 && j < s->nU, "fvc_eval: field index out of range");
        if ((rc = ensure(h, s->work, (size_t)s->nU * s->ldU * sizeof(double)))) return rc;
        if ((rc = run_laplacians(h, s, s->work.as<double>()))) return rc;
        nout = (size_t)3 * s->nCells;
        unpermute_cells_kernel<<<ug(nout), 256, 0, h->stream>>>(s->work.as<double>() + (size_t)j * s->ldU, s->perm.as<int>(), s->nCells, 3,
                                                                 s->stage.as<double>());
        break;
    }
    case IB200_FVC_GRAD_P: {
        PROJ_REQUIRE(j >= 0 && j < s->nP, "fvc_eval: field index out of range");
        const long long ldo = pad16(3 * s->nCells);
        if ((rc = ensure(h, s->work, (size_t)s->nP * ldo * sizeof(double)))) return rc;
        grad_scalar_kernel<<<dim3(gx(s->nCells), gy(s->nP)), PROJ_T, 0, h->stream>>>(m, s->P.as<double>(), s->ldP, s->Pb.as<double>(),
                                                                                      s->ldPb, s->nP, s->work.as<double>(), ldo);
        h->launches += 1;
        nout = (size_t)3 * s->nCells;
        unpermute_cells_kernel<<<ug(nout), 256, 0, h->stream>>>(s->work.as<double>() + (size_t)j * ldo, s->perm.as<int>(), s->nCells, 3,
                                                                 s->stage.as<double>());
        break;
    }
    case IB200_FVC_DIV_U: {
        PROJ_REQUIRE(j >= 0 && j < s->nU, "fvc_eval: field index out of range");
        if ((rc = need_flux(h, s))) return rc;
        const long long ldo = pad16(s->nCells);
        if ((rc = ensure(h, s->work, (size_t)s->nU * ldo * sizeof(double)))) return rc;
        surface_integrate_kernel<<<dim3(gx(s->nCells), gy(s->nU)), PROJ_T, 0, h->stream>>>(m, s->flux.as<double>(), s->ldF, s->nU,
                                                                                            s->work.as<double>(), ldo);
        h->launches += 1;
        nout = (size_t)s->nCells;
        unpermute_cells_kernel<<<ug(nout), 256, 0, h->stream>>>(s->work.as<double>() + (size_t)j * ldo, s->perm.as<int>(), s->nCells, 1,
                                                                 s->stage.as<double>());
        break;
    }
    default:
        PROJ_REQUIRE(false, "fvc_eval: unknown operator");
    }
    IB_CUDA(h, cudaGetLastError());
    h->launches += 1;
    // the field was gathered back to the caller's cell / face order into the staging buffer
    IB_CUDA(h, cudaMemcpyAsync(out, s->stage.p, nout * sizeof(double), cudaMemcpyDefault, h->stream));
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    return IB200_OK;
}

// One projected operator. Output layouts are the reference's Eigen objects (column-major; C with i fastest).
int ithaca_b200_project(ithaca_b200_t* h, int op, double* out)
{
    if (!h) return IB200_ERR_ARG;
    ProjState* s = h->proj;
    PROJ_REQUIRE(s && s->have_mesh && s->nU > 0, "project: call mesh_upload and basis_upload(U) first");
    PROJ_REQUIRE(out, "project: null output");
    PROJ_REQUIRE((op != IB200_OP_K && op != IB200_OP_P && op != IB200_OP_D && op != IB200_OP_G && op != IB200_OP_BC3 &&
                  op != IB200_OP_BC4 && op != IB200_OP_BC1 && op != IB200_OP_BC2 && op != IB200_OP_BP && op != IB200_OP_CF) || s->nP > 0,
                 "project: K, P, D, G, BC1-BC4, BP and Cf need basis_upload(P)");
    cudaSetDevice(h->device);
    const MeshDev m = mesh_dev(s);
    const int N = s->nU, Np = s->nP, BIG = 0x7fffffff;
    const long long K3 = 3 * s->nCells;
    const double* U = s->U.as<double>();
    const bool shard = s->nOwned < s->nCells;
    const double* V = shard ? s->Vw.as<double>() : s->V.as<double>();   // integration weights (0 on halo cells)
    size_t nout = 0;
    int rc;
    for (double& x : s->ms) x = 0.0;
    switch (op) {
    case IB200_OP_M: {  // mass_term, steadyNS.C:1085-1108
        nout = (size_t)N * N;
        if ((rc = ensure(h, s->out, nout * sizeof(double)))) return rc;
        StageTimer t(h->stream, &s->ms[5]);
        rc = atvb_launch(h, U, s->ldU, N, U, s->ldU, N, K3, V, 3, 1, s->out.as<double>(), 1, BIG, 0, N);
        t.stop();
        if (rc) return rc;
        break;
    }
    case IB200_OP_B: {  // diffusive_term, steadyNS.C:883-907
        nout = (size_t)N * N;
        if ((rc = ensure(h, s->out, nout * sizeof(double)))) return rc;
        if ((rc = ensure(h, s->work, (size_t)N * s->ldU * sizeof(double)))) return rc;
        if ((rc = run_laplacians(h, s, s->work.as<double>()))) return rc;
        StageTimer t(h->stream, &s->ms[5]);
        rc = atvb_launch(h, U, s->ldU, N, s->work.as<double>(), s->ldU, N, K3, V, 3, 0, s->out.as<double>(), 1, BIG, 0, N);
        t.stop();
        if (rc) return rc;
        break;
    }
    case IB200_OP_K: {  // pressure_gradient_term, steadyNS.C:935-960: N x Np
        nout = (size_t)N * Np;
        if ((rc = ensure(h, s->out, nout * sizeof(double)))) return rc;
        if ((rc = ensure(h, s->work, (size_t)Np * s->ldU * sizeof(double)))) return rc;
        {
            StageTimer t(h->stream, &s->ms[3]);
            grad_scalar_kernel<<<dim3(gx(s->nCells), gy(Np)), PROJ_T, 0, h->stream>>>(m, s->P.as<double>(), s->ldP, s->Pb.as<double>(),
                                                                                       s->ldPb, Np, s->work.as<double>(), s->ldU);
            t.stop();
            h->launches += 1;
        }
        StageTimer t(h->stream, &s->ms[5]);
        rc = atvb_launch(h, U, s->ldU, N, s->work.as<double>(), s->ldU, Np, K3, V, 3, 0, s->out.as<double>(), 1, BIG, 0, N);
        t.stop();
        if (rc) return rc;
        break;
    }
    case IB200_OP_P: {  // divergence_term, steadyNS.C:1112-1137: Np x N
        nout = (size_t)Np * N;
        if ((rc = ensure(h, s->out, nout * sizeof(double)))) return rc;
        if ((rc = need_flux(h, s))) return rc;
        if ((rc = ensure(h, s->work, (size_t)N * s->ldP * sizeof(double)))) return rc;
        {
            StageTimer t(h->stream, &s->ms[4]);
            surface_integrate_kernel<<<dim3(gx(s->nCells), gy(N)), PROJ_T, 0, h->stream>>>(m, s->flux.as<double>(), s->ldF, N,
                                                                                            s->work.as<double>(), s->ldP);
            t.stop();
            h->launches += 1;
        }
        StageTimer t(h->stream, &s->ms[5]);
        rc = atvb_launch(h, s->P.as<double>(), s->ldP, Np, s->work.as<double>(), s->ldP, N, s->nCells, V, 1, 0, s->out.as<double>(), 1,
                         BIG, 0, Np);
        t.stop();
        if (rc) return rc;
        break;
    }
    case IB200_OP_C:    // convective_term_tens (== _cache), steadyNS.C:997-1083, fused face-based contraction (conv.cu)
        if (s->convScheme == 2) goto materialised_C;   // linearUpwind: the face value needs grad(U_k) of the upwind cell -> div fields
    {
        nout = (size_t)N * N * N;
        if ((rc = ensure(h, s->out, nout * sizeof(double)))) return rc;
        ConvMesh cm;
        cm.nInt = s->nInt; cm.nBnd = s->nBnd; cm.owner = m.owner; cm.neigh = m.neigh; cm.Sf = m.Sf; cm.w = m.w; cm.bmask = m.bmask;
        cm.cmask = shard ? s->cmask.as<unsigned char>() : nullptr;
        cm.upwind = s->convScheme == 1 ? 1 : 0;
        if (s->flux_user) { cm.flux = s->fluxUser.as<double>(); cm.ldF = s->ldF; }
        rc = conv_fused_launch(h, &s->conv, cm, s->nCells, U, s->ldU, s->Ub.as<double>(), s->ldUb, N, s->out.as<double>(), &s->ms[6],
                               &s->ms[5]);
        if (rc) return rc;
        break;
    }
    materialised_C:
    case IB200_OP_C_MATERIALISED: {  // same tensor through materialised div(phi_j, U_k) fields + cell-based contraction
        nout = (size_t)N * N * N;
        if ((rc = ensure(h, s->out, nout * sizeof(double)))) return rc;
        if ((rc = need_flux(h, s))) return rc;
        const size_t percol = (size_t)s->ldU * sizeof(double);
        int jb = (int)std::max<size_t>(1, std::min<size_t>({(size_t)N, 8, (size_t)12e9 / (percol * N)}));
        if ((rc = ensure(h, s->work, (size_t)jb * N * percol))) return rc;
        for (int j0 = 0; j0 < N; j0 += jb) {
            const int nj = std::min(jb, N - j0);
            {
                StageTimer t(h->stream, &s->ms[1]);
                for (int jj = 0; jj < nj; jj++) {
                    LAUNCH_DIV_PHI(div_phi_kernel, dim3(gx(s->nCells), gy(N)), 
                        m, conv_flux(s) + (size_t)(j0 + jj) * s->ldF, U, s->ldU, s->Ub.as<double>(), s->ldUb, N,
                        s->work.as<double>() + (size_t)jj * N * s->ldU, s->ldU);
                    h->launches += 1;
                }
                t.stop();
            }
            StageTimer t(h->stream, &s->ms[5]);
            // column jj*N + k of the work matrix = div(phi_{j0+jj}, U_k)  ->  C(i, j0+jj, k)
            rc = atvb_launch(h, U, s->ldU, N, s->work.as<double>(), s->ldU, nj * N, K3, V, 3, 0,
                             s->out.as<double>() + (size_t)N * j0, 1, N, N, (long long)N * N);
            t.stop();
            if (rc) return rc;
        }
        break;
    }
    case IB200_OP_D: {  // laplacian_pressure, steadyNS.C:1295-1317: Np x Np
        nout = (size_t)Np * Np;
        if ((rc = ensure(h, s->out, nout * sizeof(double)))) return rc;
        if ((rc = ensure(h, s->gradP, (size_t)Np * s->ldU * sizeof(double)))) return rc;
        {
            StageTimer t(h->stream, &s->ms[3]);
            grad_scalar_kernel<<<dim3(gx(s->nCells), gy(Np)), PROJ_T, 0, h->stream>>>(m, s->P.as<double>(), s->ldP, s->Pb.as<double>(),
                                                                                       s->ldPb, Np, s->gradP.as<double>(), s->ldU);
            t.stop();
            h->launches += 1;
        }
        StageTimer t(h->stream, &s->ms[5]);
        rc = atvb_launch(h, s->gradP.as<double>(), s->ldU, Np, s->gradP.as<double>(), s->ldU, Np, K3, V, 3, 1, s->out.as<double>(), 1,
                         BIG, 0, Np);
        t.stop();
        if (rc) return rc;
        break;
    }
    case IB200_OP_G: {  // divMomentum (== divMomentum_cache), steadyNS.C:1176-1254: Np x N x N, i fastest
        nout = (size_t)Np * N * N;
        if ((rc = ensure(h, s->out, nout * sizeof(double)))) return rc;
        if ((rc = ensure(h, s->gradP, (size_t)Np * s->ldU * sizeof(double)))) return rc;
        {
            StageTimer t(h->stream, &s->ms[3]);
            grad_scalar_kernel<<<dim3(gx(s->nCells), gy(Np)), PROJ_T, 0, h->stream>>>(m, s->P.as<double>(), s->ldP, s->Pb.as<double>(),
                                                                                       s->ldPb, Np, s->gradP.as<double>(), s->ldU);
            t.stop();
            h->launches += 1;
        }
        // fused face-based form (conv.cu), as for C: G(i,j,k) = sum_f flux_j(f) (Uf_k(f) . (g_i[own] - g_i[nei])), g_i = grad(P_i);
        // ITHACA_B200_G=mat keeps the materialised div(phi_j, U_k) fields (cross-check)
        const char* gev = env_knob("ITHACA_B200_G");
        if (Np <= N && !(gev && strcmp(gev, "mat") == 0) && s->convScheme != 2) {
            ConvMesh cm;
            cm.nInt = s->nInt; cm.nBnd = s->nBnd; cm.owner = m.owner; cm.neigh = m.neigh; cm.Sf = m.Sf; cm.w = m.w; cm.bmask = m.bmask;
            cm.cmask = shard ? s->cmask.as<unsigned char>() : nullptr;
        cm.upwind = s->convScheme == 1 ? 1 : 0;
        if (s->flux_user) { cm.flux = s->fluxUser.as<double>(); cm.ldF = s->ldF; }
            rc = conv_fused_launch(h, &s->conv, cm, s->nCells, U, s->ldU, s->Ub.as<double>(), s->ldUb, N, s->out.as<double>(), &s->ms[6],
                                   &s->ms[5], s->gradP.as<double>(), s->ldU, Np);   // writes G(i,j,k) at i + Np*(j + N*k)
            if (rc) return rc;
            break;
        }
        if ((rc = need_flux(h, s))) return rc;
        const size_t percol = (size_t)s->ldU * sizeof(double);
        int jb = (int)std::max<size_t>(1, std::min<size_t>({(size_t)N, 8, (size_t)12e9 / (percol * N)}));
        if ((rc = ensure(h, s->work, (size_t)jb * N * percol))) return rc;
        for (int j0 = 0; j0 < N; j0 += jb) {
            const int nj = std::min(jb, N - j0);
            {
                StageTimer t(h->stream, &s->ms[1]);
                for (int jj = 0; jj < nj; jj++) {
                    LAUNCH_DIV_PHI(div_phi_kernel, dim3(gx(s->nCells), gy(N)), 
                        m, conv_flux(s) + (size_t)(j0 + jj) * s->ldF, U, s->ldU, s->Ub.as<double>(), s->ldUb, N,
                        s->work.as<double>() + (size_t)jj * N * s->ldU, s->ldU);
                    h->launches += 1;
                }
                t.stop();
            }
            StageTimer t(h->stream, &s->ms[5]);
            // column jj*N + k of the work matrix = div(phi_{j0+jj}, U_k)  ->  G(i, j0+jj, k) at i + Np*(j + N*k)
            rc = atvb_launch(h, s->gradP.as<double>(), s->ldU, Np, s->work.as<double>(), s->ldU, nj * N, K3, V, 3, 0,
                             s->out.as<double>() + (size_t)Np * j0, 1, N, Np, (long long)Np * N);
            t.stop();
            if (rc) return rc;
        }
        break;
    }
    case IB200_OP_BC4: {  // pressure_BC4, steadyNS.C:1477-1513: Np x N, boundary-face sum of P_i (n . U_j) |Sf|
        nout = (size_t)Np * N;
        if ((rc = ensure(h, s->out, nout * sizeof(double)))) return rc;
        IB_CUDA(h, cudaMemsetAsync(s->out.p, 0, nout * sizeof(double), h->stream));
        if (s->nBnd > 0) {
            const long long ldB = pad16(s->nBnd);
            if ((rc = ensure(h, s->bndA, (size_t)N * ldB * sizeof(double)))) return rc;
            {
                StageTimer t(h->stream, &s->ms[0]);
                bnd_normal_flux_kernel<<<dim3(gx(s->nBnd), N), PROJ_T, 0, h->stream>>>(m, shard ? s->cmask.as<unsigned char>() : nullptr,
                                                                                        s->Ub.as<double>(), s->ldUb, N, s->bndA.as<double>(), ldB);
                t.stop();
                h->launches += 1;
            }
            StageTimer t(h->stream, &s->ms[5]);
            rc = atvb_launch(h, s->Pb.as<double>(), s->ldPb, Np, s->bndA.as<double>(), ldB, N, s->nBnd, nullptr, 1, 0, s->out.as<double>(), 1,
                             BIG, 0, Np);
            t.stop();
            if (rc) return rc;
        }
        break;
    }
    case IB200_OP_BC1: {  // pressure_BC1, steadyNS.C:1319-1354: Np x N, boundary-face sum of (lap(U_j)[owner] . Sf) P_i
        nout = (size_t)Np * N;
        if ((rc = ensure(h, s->out, nout * sizeof(double)))) return rc;
        IB_CUDA(h, cudaMemsetAsync(s->out.p, 0, nout * sizeof(double), h->stream));
        if (s->nBnd > 0) {
            const long long ldB = pad16(s->nBnd);
            if ((rc = ensure(h, s->work, (size_t)N * s->ldU * sizeof(double))) || (rc = ensure(h, s->bndA, (size_t)N * ldB * sizeof(double))))
                return rc;
            if ((rc = run_laplacians(h, s, s->work.as<double>()))) return rc;
            bnd_owner_dot_sf_kernel<<<dim3(gx(s->nBnd), N), PROJ_T, 0, h->stream>>>(m, shard ? s->cmask.as<unsigned char>() : nullptr,
                                                                                     s->work.as<double>(), s->ldU, N, s->bndA.as<double>(), ldB);
            h->launches += 1;
            StageTimer t(h->stream, &s->ms[5]);
            rc = atvb_launch(h, s->Pb.as<double>(), s->ldPb, Np, s->bndA.as<double>(), ldB, N, s->nBnd, nullptr, 1, 0, s->out.as<double>(), 1,
                             BIG, 0, Np);
            t.stop();
            if (rc) return rc;
        }
        break;
    }
    case IB200_OP_BC3: {  // pressure_BC3, steadyNS.C:1439-1475: Np x N, boundary-face sum of curl(U_j) . (n x grad P_i) |Sf|
        nout = (size_t)Np * N;
        if ((rc = ensure(h, s->out, nout * sizeof(double)))) return rc;
        IB_CUDA(h, cudaMemsetAsync(s->out.p, 0, nout * sizeof(double), h->stream));
        if (s->nBnd > 0) {
            const long long ld3 = pad16(3 * s->nBnd);
            const unsigned char* cm = shard ? s->cmask.as<unsigned char>() : nullptr;
            if ((rc = ensure(h, s->bndA, (size_t)N * ld3 * sizeof(double))) || (rc = ensure(h, s->bndB, (size_t)Np * ld3 * sizeof(double))))
                return rc;
            {
                StageTimer t(h->stream, &s->ms[3]);
                bnd_curl_kernel<<<dim3(gx(s->nBnd), N), PROJ_T, 0, h->stream>>>(m, cm, U, s->ldU, s->Ub.as<double>(), s->ldUb, N,
                                                                                 s->bndA.as<double>(), ld3);
                bnd_ncross_gradp_kernel<<<dim3(gx(s->nBnd), Np), PROJ_T, 0, h->stream>>>(m, cm, s->P.as<double>(), s->ldP, s->Pb.as<double>(),
                                                                                          s->ldPb, Np, s->bndB.as<double>(), ld3);
                t.stop();
                h->launches += 2;
            }
            StageTimer t(h->stream, &s->ms[5]);
            rc = atvb_launch(h, s->bndB.as<double>(), ld3, Np, s->bndA.as<double>(), ld3, N, 3 * s->nBnd, nullptr, 1, 0, s->out.as<double>(), 1,
                             BIG, 0, Np);
            t.stop();
            if (rc) return rc;
        }
        break;
    }
    case IB200_OP_BC2: {  // pressureBC2 (== pressure_BC2), steadyNS.C:1356-1437: Np x N x N, i fastest; boundary-face sum of
                          // (div(interpolate(U_j) & Sf, U_k)[owner] . Sf) P_i
        nout = (size_t)Np * N * N;
        if ((rc = ensure(h, s->out, nout * sizeof(double)))) return rc;
        IB_CUDA(h, cudaMemsetAsync(s->out.p, 0, nout * sizeof(double), h->stream));
        if (s->nBnd > 0) {
            if ((rc = need_flux(h, s))) return rc;
            const long long ldB = pad16(s->nBnd);
            const int jb = (int)std::max<size_t>(1, std::min<size_t>((size_t)N, (size_t)2e9 / ((size_t)ldB * N * sizeof(double))));
            if ((rc = ensure(h, s->bndA, (size_t)jb * N * ldB * sizeof(double)))) return rc;
            for (int j0 = 0; j0 < N; j0 += jb) {
                const int nj = std::min(jb, N - j0);
                {
                    StageTimer t(h->stream, &s->ms[1]);
                    for (int jj = 0; jj < nj; jj++) {
                        LAUNCH_DIV_PHI(bnd_div_phi_dot_sf_kernel, dim3(gx(s->nBnd), gy(N)), 
                            m, shard ? s->cmask.as<unsigned char>() : nullptr, conv_flux(s) + (size_t)(j0 + jj) * s->ldF, U, s->ldU,
                            s->Ub.as<double>(), s->ldUb, N, s->bndA.as<double>() + (size_t)jj * N * ldB, ldB);
                        h->launches += 1;
                    }
                    t.stop();
                }
                StageTimer t(h->stream, &s->ms[5]);
                // column jj*N + k of bndA = div(phi_{j0+jj}, U_k)[owner] . Sf  ->  BC2(i, j0+jj, k) at i + Np*(j + N*k)
                rc = atvb_launch(h, s->Pb.as<double>(), s->ldPb, Np, s->bndA.as<double>(), ldB, nj * N, s->nBnd, nullptr, 1, 0,
                                 s->out.as<double>() + (size_t)Np * j0, 1, N, Np, (long long)Np * N);
                t.stop();
                if (rc) return rc;
            }
        }
        break;
    }
    case IB200_OP_BP:    // diffusive_term_flux_method, steadyNS.C:1584-1610: Np x N, int P_i div(laplacian(U_j))
    case IB200_OP_CF: {  // convective_term_flux_tens, steadyNS.C:1704-1746: Np x N x N, int P_i div(div(flux(U_j), U_k)), i fastest
        // both through the adjoint of the outer div (div_adjoint_kernel): int P_i div(A) = int t_i . A, so BP is one
        // contraction against the laplacian fields and Cf is the fused face-based contraction of G with t_i as test fields
        PROJ_REQUIRE(op == IB200_OP_BP || s->convScheme != 2, "project(Cf): not built for the linearUpwind convection scheme (fused contraction only)");
        nout = op == IB200_OP_BP ? (size_t)Np * N : (size_t)Np * N * N;
        if ((rc = ensure(h, s->out, nout * sizeof(double)))) return rc;
        if ((rc = ensure(h, s->gradP, (size_t)Np * s->ldU * sizeof(double)))) return rc;
        {
            StageTimer t(h->stream, &s->ms[3]);
            div_adjoint_kernel<<<dim3(gx(s->nCells), gy(Np)), PROJ_T, 0, h->stream>>>(m, s->P.as<double>(), s->ldP, Np,
                                                                                       s->gradP.as<double>(), s->ldU);
            t.stop();
            h->launches += 1;
        }
        if (op == IB200_OP_BP) {
            if ((rc = ensure(h, s->work, (size_t)N * s->ldU * sizeof(double)))) return rc;
            if ((rc = run_laplacians(h, s, s->work.as<double>()))) return rc;
            StageTimer t(h->stream, &s->ms[5]);
            rc = atvb_launch(h, s->gradP.as<double>(), s->ldU, Np, s->work.as<double>(), s->ldU, N, K3, V, 3, 0, s->out.as<double>(), 1,
                             BIG, 0, Np);
            t.stop();
            if (rc) return rc;
            break;
        }
        if (Np <= N) {
            ConvMesh cm;
            cm.nInt = s->nInt; cm.nBnd = s->nBnd; cm.owner = m.owner; cm.neigh = m.neigh; cm.Sf = m.Sf; cm.w = m.w; cm.bmask = m.bmask;
            cm.cmask = shard ? s->cmask.as<unsigned char>() : nullptr;
        cm.upwind = s->convScheme == 1 ? 1 : 0;
        if (s->flux_user) { cm.flux = s->fluxUser.as<double>(); cm.ldF = s->ldF; }
            rc = conv_fused_launch(h, &s->conv, cm, s->nCells, U, s->ldU, s->Ub.as<double>(), s->ldUb, N, s->out.as<double>(), &s->ms[6],
                                   &s->ms[5], s->gradP.as<double>(), s->ldU, Np);   // writes Cf(i,j,k) at i + Np*(j + N*k)
            if (rc) return rc;
            break;
        }
        if ((rc = need_flux(h, s))) return rc;
        const size_t percol = (size_t)s->ldU * sizeof(double);
        int jb = (int)std::max<size_t>(1, std::min<size_t>({(size_t)N, 8, (size_t)12e9 / (percol * N)}));
        if ((rc = ensure(h, s->work, (size_t)jb * N * percol))) return rc;
        for (int j0 = 0; j0 < N; j0 += jb) {
            const int nj = std::min(jb, N - j0);
            for (int jj = 0; jj < nj; jj++) {
                LAUNCH_DIV_PHI(div_phi_kernel, dim3(gx(s->nCells), gy(N)), 
                    m, conv_flux(s) + (size_t)(j0 + jj) * s->ldF, U, s->ldU, s->Ub.as<double>(), s->ldUb, N,
                    s->work.as<double>() + (size_t)jj * N * s->ldU, s->ldU);
                h->launches += 1;
            }
            StageTimer t(h->stream, &s->ms[5]);
            rc = atvb_launch(h, s->gradP.as<double>(), s->ldU, Np, s->work.as<double>(), s->ldU, nj * N, K3, V, 3, 0,
                             s->out.as<double>() + (size_t)Np * j0, 1, N, Np, (long long)Np * N);
            t.stop();
            if (rc) return rc;
        }
        break;
    }
    default:
        PROJ_REQUIRE(false, "project: unknown operator");
    }
    IB_CUDA(h, cudaGetLastError());
    // cell-sharded run: one all-reduce of the whole operator (the reference: one gSum per entry inside domainIntegrate)
    if (h->nccl_comm && h->nranks > 1) {
        StageTimer t(h->stream, &s->ms[7]);
        rc = allreduce_sum(h, s->out.as<double>(), nout);
        t.stop();
        if (rc) return rc;
    }
    IB_CUDA(h, cudaMemcpyAsync(out, s->out.p, nout * sizeof(double), cudaMemcpyDefault, h->stream));
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    return IB200_OK;
}

int ithaca_b200_project_bc_velocity(ithaca_b200_t* h, int64_t patch_start, int64_t patch_size, int comp, double* vec_out, double* mat_out)
{
    if (!h) return IB200_ERR_ARG;
    ProjState* s = h->proj;
    PROJ_REQUIRE(s && s->have_mesh && s->nU > 0, "project_bc_velocity: call mesh_upload and basis_upload(U) first");
    PROJ_REQUIRE(patch_start >= 0 && patch_size > 0 && patch_start + patch_size <= s->nBnd && comp >= 0 && comp < 3 && (vec_out || mat_out),
                 "project_bc_velocity: bad argument");
    cudaSetDevice(h->device);
    const MeshDev m = mesh_dev(s);
    const int N = s->nU, N1 = N + 1;
    const bool shard = s->nOwned < s->nCells;
    const long long ldp = pad16(patch_size);
    int rc;
    if ((rc = ensure(h, s->bndA, (size_t)N1 * ldp * sizeof(double))) || (rc = ensure(h, s->out, (size_t)N1 * N1 * sizeof(double)))) return rc;
    bnd_patch_component_kernel<<<dim3(gx(patch_size), N1), PROJ_T, 0, h->stream>>>(m, shard ? s->cmask.as<unsigned char>() : nullptr,
                                                                                  s->Ub.as<double>(), s->ldUb, N, patch_start, patch_size, comp,
                                                                                  s->bndA.as<double>(), ldp);
    h->launches += 1;
    // Gram of [X, 1]: the N x N block is bcVelMat, the last column bcVelVec
    rc = atvb_launch(h, s->bndA.as<double>(), ldp, N1, s->bndA.as<double>(), ldp, N1, patch_size, nullptr, 1, 1, s->out.as<double>(), 1,
                     0x7fffffff, 0, N1);
    if (rc) return rc;
    if (h->nccl_comm && h->nranks > 1 && (rc = allreduce_sum(h, s->out.as<double>(), (size_t)N1 * N1))) return rc;
    std::vector<double> hM((size_t)N1 * N1);
    IB_CUDA(h, cudaMemcpyAsync(hM.data(), s->out.p, hM.size() * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    for (int j = 0; j < N; j++) {
        if (vec_out) vec_out[j] = hM[(size_t)j + (size_t)N * N1];
        if (mat_out)
            for (int i = 0; i < N; i++) mat_out[(size_t)i + (size_t)j * N] = hM[(size_t)i + (size_t)j * N1];
    }
    return IB200_OK;
}

int ithaca_b200_project_background(ithaca_b200_t* h, int kind, const double* vls_int, const double* vls_bnd, double* out)
{
    if (!h) return IB200_ERR_ARG;
    ProjState* s = h->proj;
    PROJ_REQUIRE(s && s->have_mesh && s->nU > 0, "project_background: call mesh_upload and basis_upload(U) first");
    PROJ_REQUIRE((kind == 0 || kind == 1) && vls_int && (vls_bnd || s->nBnd == 0) && out, "project_background: bad argument");
    PROJ_REQUIRE(kind == 0 || s->nP > 0, "project_background: L_D needs basis_upload(P)");
    cudaSetDevice(h->device);
    const MeshDev m = mesh_dev(s);
    const int N = s->nU, Np = s->nP, BIG = 0x7fffffff;
    const long long K3 = 3 * s->nCells;
    const bool shard = s->nOwned < s->nCells;
    const double* V = shard ? s->Vw.as<double>() : s->V.as<double>();
    const int rows = kind == 0 ? N : Np;
    const size_t nout = (size_t)rows * N;
    for (double& x : s->ms) x = 0.0;
    int rc;
    // buffers: vls (internal order) + its boundary values + its face flux in bndB; div fields in work
    const long long ldv = s->ldU, ldvb = s->ldUb, ldF = s->ldF;
    if ((rc = ensure(h, s->bndB, (size_t)(ldv + ldvb + ldF) * sizeof(double))) || (rc = ensure(h, s->work, (size_t)N * s->ldU * sizeof(double))) ||
        (rc = ensure(h, s->out, nout * sizeof(double))))
        return rc;
    double* vl = s->bndB.as<double>();
    double* vb = vl + ldv;
    double* phi = vb + ldvb;
    IB_CUDA(h, cudaMemcpyAsync(s->stage.p, vls_int, (size_t)3 * s->nCells * sizeof(double), cudaMemcpyDefault, h->stream));
    permute_cells_kernel<<<(unsigned)std::min<long long>((3 * s->nCells + 255) / 256, 64LL * h->num_sms), 256, 0, h->stream>>>(
        s->stage.as<double>(), s->perm.as<int>(), s->nCells, 3, vl);
    if (s->nBnd) IB_CUDA(h, cudaMemcpyAsync(vb, vls_bnd, (size_t)3 * s->nBnd * sizeof(double), cudaMemcpyDefault, h->stream));
    {
        StageTimer t(h->stream, &s->ms[0]);
        face_flux_kernel<<<dim3(gx(s->nInt + s->nBnd), gy(1)), PROJ_T, 0, h->stream>>>(m, vl, ldv, vb, ldvb, 1, phi, ldF);
        t.stop();
    }
    {
        StageTimer t(h->stream, &s->ms[1]);
        LAUNCH_DIV_PHI(div_phi_kernel, dim3(gx(s->nCells), gy(N)), m, phi, s->U.as<double>(), s->ldU, s->Ub.as<double>(), s->ldUb, N,
                                                                             s->work.as<double>(), s->ldU);
        t.stop();
    }
    h->launches += 3;
    const double* A = s->U.as<double>();
    if (kind == 1) {
        if ((rc = ensure(h, s->gradP, (size_t)Np * s->ldU * sizeof(double)))) return rc;
        StageTimer t(h->stream, &s->ms[3]);
        grad_scalar_kernel<<<dim3(gx(s->nCells), gy(Np)), PROJ_T, 0, h->stream>>>(m, s->P.as<double>(), s->ldP, s->Pb.as<double>(), s->ldPb, Np,
                                                                                   s->gradP.as<double>(), s->ldU);
        t.stop();
        h->launches += 1;
        A = s->gradP.as<double>();
    }
    {
        StageTimer t(h->stream, &s->ms[5]);
        rc = atvb_launch(h, A, s->ldU, rows, s->work.as<double>(), s->ldU, N, K3, V, 3, 0, s->out.as<double>(), 1, BIG, 0, rows);
        t.stop();
    }
    if (rc) return rc;
    negate_kernel<<<(unsigned)((nout + 255) / 256), 256, 0, h->stream>>>(s->out.as<double>(), (long long)nout);
    h->launches += 1;
    IB_CUDA(h, cudaGetLastError());
    if (h->nccl_comm && h->nranks > 1 && (rc = allreduce_sum(h, s->out.as<double>(), nout))) return rc;
    IB_CUDA(h, cudaMemcpyAsync(out, s->out.p, nout * sizeof(double), cudaMemcpyDefault, h->stream));
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    return IB200_OK;
}

// Device times (ms) of the last ithaca_b200_project call: out[0]=face flux out[1]=div(phi,U) out[2]=laplacian (+grad tensor)
// out[3]=grad(p) out[4]=div(U) out[5]=contraction (DMMA) out[6]=face records of the fused C (transpose + gather).
int ithaca_b200_proj_timings(ithaca_b200_t* h, double* out, int n)
{
    if (!h || !out || !h->proj) return IB200_ERR_ARG;
    for (int i = 0; i < n; i++) out[i] = (i < 8) ? h->proj->ms[i] : 0.0;
    return IB200_OK;
}

}  // extern "C"
