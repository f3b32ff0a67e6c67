// Shared device helpers for libithaca_b200 (sm_100a only): mbarrier / TMA PTX wrappers, the FP64
// tensor-core instruction, and the handle layout. Internal; the public boundary is include/ithaca_b200.h.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>

namespace ib200 {

// ---- FP64 tensor core: DMMA.8x8x4 is the only native f64 MMA shape on sm_100a (larger PTX shapes
// are lowered to it, checked with cuobjdump). Fragment layout (lane = 4*g + t):
//   A (8x4, row-major): a = A[g][t]     B (4x8, col-major): b = B[t][g]
//   C/D (8x8): c0 = C[g][2t], c1 = C[g][2t+1]
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

__device__ __forceinline__ uint32_t smem_u32(const void* p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init()
{
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// DELAYED RELEASE -- the rule for every consumer warp of a TMA ring that reads a stage with ordinary shared-memory loads
// (LDS) feeding DMMAs (gram.cu, lift.cu, conv.cu): stage s is handed back to the producer only after the wait for stage
// s + 1, never at the end of its own iteration.
// A `__syncwarp(); if (lane == 0) mbar_arrive(&empty[s]);` at the end of the iteration is a write-after-read race: nothing
// orders the arrive against the DMMAs, so ptxas schedules it right behind the last LDS *issue* -- ahead of the DMMAs that consume
// the data and without a scoreboard wait -- and the arrive then overtakes loads that are still in flight: the producer's next
// TMA box lands under them. Measured with scripts/exp/lift_stress.py (m = 256, r = 24, two CTAs per SM): one warp's A or B
// fragment of one stage wrong in 8 of 40 passes over 1.2 M rows with the arrive at 3 LDS in flight, 0 of 300 since.
// Behind the NEXT wait (a spin loop: no instruction is moved across it) every DMMA of the released stage has issued, and a
// DMMA cannot issue before the LDS results it reads have returned -- whatever the instruction scheduler does inside the loop
// body. Cost: the ring holds one stage less in flight. tests/test_sass_ring_release.py checks the order in the SASS of the build.
__device__ __forceinline__ void mbar_release_previous(uint64_t* empty, int prev, int lane)
{
    if (prev >= 0) {
        __syncwarp();
        if (lane == 0) mbar_arrive(empty + prev);
    }
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile(
        "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
        : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    while (!mbar_try_wait(bar, parity)) { }
}

// TMA: 2-D tiled bulk tensor load global -> shared, completion counted in bytes on an mbarrier.
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* tmap, int c0, int c1,
                                            uint64_t* bar)
{
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
        ::"r"(smem_u32(smem_dst)), "l"(tmap), "r"(c0), "r"(c1), "r"(smem_u32(bar)) : "memory");
}
// 1-D bulk copy global -> shared (bytes multiple of 16, both addresses 16-B aligned), completion on an mbarrier.
__device__ __forceinline__ void bulk_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* tmap)
{
    asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}

// Byte offset of element (row, k) inside a TMA box whose rows are 16 doubles (128 B) written with
// CU_TENSOR_MAP_SWIZZLE_128B into 1024-B aligned shared memory: the 16-B chunk index is XOR-ed with
// (row & 7).
__device__ __forceinline__ uint32_t swz128(uint32_t row, uint32_t k)
{
    return row * 128u + ((((k >> 1) ^ (row & 7u)) << 4) | ((k & 1u) << 3));
}

}  // namespace ib200

// ------------------------------------------------------------------------------------------------
// Host side
// ------------------------------------------------------------------------------------------------
#define IB_CUDA(h, call)                                                                          \
    do {                                                                                          \
        cudaError_t e__ = (call);                                                                 \
        if (e__ != cudaSuccess) {                                                                 \
            ib200::set_error(h, std::string(#call) + ": " + cudaGetErrorString(e__));             \
            return IB200_ERR_CUDA;                                                                \
        }                                                                                         \
    } while (0)

struct ithaca_b200;

namespace ib200 {
struct ProjState;  // proj.cu
void proj_destroy(ithaca_b200* h);
// growable device buffer (proj.cu / conv.cu)
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};
int dev_ensure(ithaca_b200* h, DevBuf& b, size_t bytes);
// conv.cu: fused convective tensor
struct ConvMesh {
    long long nInt, nBnd;
    const int *owner, *neigh;
    const double *Sf, *w;
    const unsigned char* bmask;
    const unsigned char* cmask;   // per cell: 1 owned, 0 halo (null = all owned); scales U_i in d_i = U_i[own] - U_i[nei]
    const double* flux = nullptr; // caller-supplied face fluxes (fluxMethod "consistent": L_PHImodes), field-major flux[j*ldF + f] in
    long long ldF = 0;            // the kernels' face order; null: a_j = linearInterpolate(U_j) & Sf
    int upwind = 0;               // 1: "Gauss upwind" face value. Every internal face becomes TWO rank-one records,
                                  // (max(a,0), U_k[own], d) at f and (min(a,0), U_k[nei], d) at nInt + nBnd + f
};
struct ConvBufs {
    DevBuf Ucm, Ubcm, Dcm, Aj, Bk, Dd, ws, chunks;
};
int conv_fused_launch(ithaca_b200* h, ConvBufs* cb, const ConvMesh& m, long long nCells, const double* U, long long ldU,
                      const double* Ub, long long ldUb, int N, double* dC, double* ms_prep, double* ms_contract,
                      const double* Dfm = nullptr, long long ldD = 0, int Ni = 0);   // Dfm: Ni test fields for d_i (null: the basis)
void set_error(ithaca_b200* h, const std::string& msg);
// getenv for the ITHACA_B200_* measurement knobs: an active override is logged to stderr once per name, so that a drop-in
// library never changes kernel selection or summation order silently.
const char* env_knob(const char* name);

// cuTensorMapEncodeTiled through cudaGetDriverEntryPoint (no link-time libcuda dependency).
// 2-D FP64 tensor: inner dimension `inner` (contiguous), outer dimension `outer` with `pitch_bytes`.
int make_tmap_2d(ithaca_b200* h, CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer,
                 uint64_t pitch_bytes, uint32_t box_inner, uint32_t box_outer, bool swizzle128);

struct Timer {
    cudaEvent_t a = nullptr, b = nullptr;
    bool used = false;
};
enum { T_UPLOAD = 0, T_GRAM, T_ALLREDUCE, T_EIG, T_LIFT, T_SCALE, T_DOWNLOAD, T_GRAMK, T_COUNT };

}  // namespace ib200

// ---- launchers implemented in the .cu files ----
namespace ib200 {
// gram.cu: ws/partials live in the handle. Writes the full symmetric m x m Gram (col-major) to dG.
int gram_launch(ithaca_b200* h);
// eig.cu: one-sided Jacobi of h->dG. Fills h->dEvals (m, descending) and h->dQ (m x m col-major, sorted).
int eig_launch(ithaca_b200* h, int* sweeps);
// lift.cu
int lift_launch(ithaca_b200* h, int r, int norm_kind);
int lift_patch_launch(ithaca_b200* h, const double* dSbc, int64_t rows, double* dOut);
int lift_scaled_launch(ithaca_b200* h, int r, const double* dF);   // modes_i = (S q_i) * dF[i]
// snapshot-blocked lift: Phi (+)= S Q[first : first+nb, 0:r] (dTmp: r columns of ld doubles when accumulating); finish = norms + scale
int lift_block_launch(ithaca_b200* h, const double* S, int nb, int first, int r, bool accumulate, double* dTmp);
int lift_block_finish(ithaca_b200* h, int r, int norm_kind);
// row-streamed lift: rows [row0, row0 + NdChunk) of the rank's mode matrix (columns of ldTotal doubles) = S_chunk Q_r, squared norms
// added to h->dNorm (zeroed at the first chunk); finish = all-reduce, sqrt, scale of all NdTotal rows
int lift_stream_add(ithaca_b200* h, const double* S, long long ldS, long long NdChunk, const double* Vchunk, int r, int norm_kind,
                    long long row0, long long ldTotal, bool first_chunk);
int lift_stream_finish(ithaca_b200* h, int r, long long NdTotal, long long ldTotal);
// out[:, j] = Phi[:, 0:nmodes] * C[:, j] for j < kc <= 8 (C: nmodes x kc col-major on the device); out columns of ld doubles
int reconstruct_launch(ithaca_b200* h, int nmodes, const double* dC, int kc, int clip, double* dOut);
// gram.cu: general weighted contraction  out(i,j) = sum_k A(k,i) * V[k/dim] * B(k,j)  (V == nullptr: unweighted).
// A: Ma columns of lda doubles, B: Nb columns of ldb doubles, both with K valid rows (K-contiguous columns).
// symmetric != 0: B == A, only the lower triangle is computed and mirrored. out(i,j) is written to
// out[i*rs + (j / cinner)*cso + (j % cinner)*csi] (doubles), so the caller can scatter straight into a matrix
// (cinner = INT_MAX, csi = column stride) or into slices of a tensor.
int allreduce_sum(ithaca_b200* h, double* dbuf, size_t count);   // capi.cu: ncclAllReduce(sum, f64) on the handle stream, no-op for 1 rank
int atvb_launch(ithaca_b200* h, const double* A, long long lda, int Ma, const double* B, long long ldb, int Nb,
                long long K, const double* V, int dim, int symmetric, double* out, long long rs, int cinner,
                long long cso, long long csi);
}  // namespace ib200

typedef int (*ib200_allreduce_fn)(void* comm, const void* send, void* recv, size_t count, int dtype,
                                  int op, void* comm2, cudaStream_t stream);

struct ithaca_b200 {
    int device = 0;
    int num_sms = 148;
    cudaStream_t stream = nullptr;
    std::string err;
    int64_t launches = 0;
    ib200::Timer timers[ib200::T_COUNT];

    // communicator (NCCL loaded with dlopen; null = single rank)
    void* nccl_comm = nullptr;
    int rank = 0, nranks = 1;

    // POD state
    int64_t nCells = 0;  // cells of this rank
    int dim = 0, m = 0;
    int64_t Nd = 0;      // nCells*dim
    int64_t ld = 0;      // leading dimension (doubles) of dS, multiple of 16
    double* dV = nullptr;
    int64_t dV_cap = 0;
    double* dS = nullptr;  // snapshot matrix, m columns of ld doubles
    size_t dS_cap = 0;
    double* dG = nullptr;  // m x m Gram (col-major, symmetric, all-reduced)
    double* dWs = nullptr; // split-K workspace
    size_t dWs_cap = 0;
    double* dQ = nullptr;      // m x m eigenvectors (col-major, sorted descending)
    double* dW = nullptr;      // Jacobi work matrix
    double* dEvals = nullptr;  // m
    int mat_cap = 0;           // capacity (m) of dG/dQ/dW/dEvals
    bool have_gram = false, have_eig = false;
    bool second_list = false;   // dS holds the second field list of pod_upload_second (liftable, but not the Gram's matrix)
    // lift state
    int r = 0;
    double* dQr = nullptr;  // [m][8 NB + 2] k-major copy of one column chunk of Q_r for the lift kernel
    size_t dQr_cap = 0;
    double* dPhi = nullptr;  // modes, r columns of ld doubles
    size_t dPhi_cap = 0;
    double* dNormPart = nullptr;
    size_t dNormPart_cap = 0;
    double* dNorm = nullptr;  // >= r doubles: squared norms then norm factors
    int norm_cap = 0;
    bool have_lift = false;
    bool have_modes = false;
    // snapshot-blocked (memory-efficient) POD: two resident snapshot blocks, slot 0 = dS
    double* dS2 = nullptr;
    size_t dS2_cap = 0;
    int blk_n[2] = {0, 0}, blk_first[2] = {0, 0};
    bool block_mode = false;
    int blk_lifted = 0;   // blocks accumulated into dPhi so far
    double* dPhiTmp = nullptr;
    size_t dPhiTmp_cap = 0;   // dPhi holds h->r modes (from pod_lift or pod_set_modes)
    void* dJobs = nullptr;
    size_t dJobs_cap = 0;
    // dim-times repeated weight vectors of the contraction kernel (gram.cu), keyed on (V, dim, K, vol_gen)
    ib200::DevBuf wexp[2];
    const double* wexp_V[2] = {nullptr, nullptr};
    int wexp_dim[2] = {0, 0};
    long long wexp_K[2] = {0, 0};
    unsigned long long wexp_gen[2] = {0, 0};
    unsigned long long vol_gen = 1;  // bumped whenever a volume vector on the device may have changed
    int wexp_next = 0;
    int* dFlags = nullptr;
    // row-streamed POD (pod_stream_*): the rank's shard passes through two chunk slots (slot 0 = dS, slot 1 = dS2)
    struct StreamSlot {
        double* V = nullptr;            // the chunk's cell volumes
        int64_t V_cap = 0;
        int64_t cell0 = 0, ncells = 0;  // which cells of the shard the slot holds (ncells = 0: empty)
        int64_t ld = 0;
        cudaEvent_t ready = nullptr;    // recorded on the copy stream after the chunk's last H2D copy
        cudaEvent_t freed = nullptr;    // recorded on the compute stream after the slot's last consumer
        bool ready_pending = false, freed_pending = false;
    };
    StreamSlot slot[2];
    bool stream_mode = false;
    int64_t strm_cells = 0, strm_Nd = 0, strm_ld = 0;   // the whole shard of this rank
    int strm_gram_chunks = 0, strm_lift_chunks = 0;
    cudaStream_t copy_stream = nullptr;
    double* dGacc = nullptr;   // m x m partial Gram summed over the chunks
    int gacc_cap = 0;
    ib200::DevBuf eigTmp; // eigensolve: m unsorted eigenvalues + 2 bounds
    ib200::DevBuf volT;   // pod_gram_volumes: the snapshot matrix scaled row-wise by sqrt|V_j|

    // ---- projection state (proj.cu) ----
    ib200::ProjState* proj = nullptr;
};
