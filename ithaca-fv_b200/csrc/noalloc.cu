// Allocation-free variant of the norm-factor computation (no counterpart in the reference: added
// for callers that cannot call cudaMalloc -- CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// What it computes is ITHACAPOD.C:196-243's normFact for caller-held columns:  n_c = sqrt(sum_k w_k Phi(k,c)^2),
// w_k = V[k / dim] ("L2") or 1 ("Frobenius") -- the arithmetic of phi_norm_part_kernel + lift_norm_reduce_kernel in lift.cu,
// whose per-block partial sums live in a cudaMalloc'ed buffer of the handle. Here the partial sums live in a file-scope
// __device__ array of fixed size instead, so the entry point calls no cudaMalloc / cudaFree / allocating routine at all.
//
// The scratch array: NA_BLOCKS x NA_MAX_COLS doubles = 592 x 256 x 8 B = 1 212 416 bytes, held by EVERY GPU that loads the
// library until the process exits. It is shared by all calls and all handles of a device and keeps the last call's contents:
// every (block, column) slot a call reads is written by that call first (blocks without rows write 0), so nothing has to be
// cleared; but two calls that overlap on the GPU (different streams, or two host threads) RACE on it -- serialise them.
//
// In host code `g_na_scratch` is only a host-side stand-in for the symbol, not a device address: the address the kernels get
// comes from cudaGetSymbolAddress, taken once per device with that device current (below).
#include "../../include/ithaca_b200.h"
#include "common.cuh"
#include <algorithm>
#include <mutex>

namespace ib200 {

constexpr int NA_MAX_COLS = IB200_NOALLOC_MAX_COLS;   // 256
constexpr int NA_BLOCKS = 592;                        // 4 x 148 SMs: row slabs per column
constexpr int NA_MAX_DEVICES = 64;

static __device__ __align__(256) double g_na_scratch[(size_t)NA_BLOCKS * NA_MAX_COLS];

// part[b * ncols + c] = sum over the rows of slab b of w_k Phi(k,c)^2 (fixed order inside the block; 0 for an empty slab)
__global__ void __launch_bounds__(256)
na_norm_part_kernel(const double* __restrict__ Phi, long long Nd, long long ld, int ncols, const double* __restrict__ V, int dim,
                    int norm_kind, double* __restrict__ part)
{
    __shared__ double sh[256];
    const int c = blockIdx.y;
    const double* col = Phi + (size_t)c * ld;
    double s = 0.0;
    for (long long i = blockIdx.x * 256LL + threadIdx.x; i < Nd; i += (long long)gridDim.x * 256) {
        const double v = col[i];
        s += (norm_kind == 0 ? V[i / dim] : 1.0) * v * v;
    }
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) part[(size_t)blockIdx.x * ncols + c] = sh[0];
}

// out[c] = sqrt(sum over the nblocks slabs, fixed order). One block per column.
__global__ void __launch_bounds__(256)
na_norm_reduce_kernel(const double* __restrict__ part, int nblocks, int ncols, double* __restrict__ out)
{
    __shared__ double sh[256];
    const int c = blockIdx.x;
    double s = 0.0;
    for (int b = threadIdx.x; b < nblocks; b += 256) s += part[(size_t)b * ncols + c];
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[c] = sqrt(sh[0]);
}

// device address of the scratch array, looked up once per device with that device current
static int na_scratch_ptr(ithaca_b200* h, double** out)
{
    static std::mutex mu;
    static double* ptrs[NA_MAX_DEVICES] = {nullptr};
    if (h->device < 0 || h->device >= NA_MAX_DEVICES) { set_error(h, "col_norms_noalloc: device index out of range"); return IB200_ERR_ARG; }
    std::lock_guard<std::mutex> lk(mu);
    if (!ptrs[h->device]) {
        IB_CUDA(h, cudaSetDevice(h->device));
        double* p = nullptr;
        IB_CUDA(h, cudaGetSymbolAddress((void**)&p, g_na_scratch));
        if (!p) { set_error(h, "col_norms_noalloc: cudaGetSymbolAddress returned a null address"); return IB200_ERR_CUDA; }
        ptrs[h->device] = p;
    }
    *out = ptrs[h->device];
    return IB200_OK;
}

}  // namespace ib200

using namespace ib200;

extern "C" {

int ithaca_b200_col_norms_noalloc(ithaca_b200_t* h, const double* Phi, int64_t nRows, int64_t ld, int nCols, const double* V,
                                  int dim, int norm_kind, double* norms_out, void* stream)
{
    if (!h) return IB200_ERR_ARG;
    if (nCols > NA_MAX_COLS) {
        set_error(h, "col_norms_noalloc: " + std::to_string(nCols) + " columns exceed the fixed scratch of IB200_NOALLOC_MAX_COLS = " +
                         std::to_string(NA_MAX_COLS) + " columns (use pod_lift / pod_modes_mass, which allocate)");
        return IB200_ERR_ARG;
    }
    if (!Phi || !norms_out || nRows <= 0 || ld < nRows || nCols <= 0 || dim <= 0 || (norm_kind != 0 && norm_kind != 1)) {
        set_error(h, norm_kind != 0 && norm_kind != 1 ? "col_norms_noalloc: The PODnorm can be only L2 or Frobenius"
                                                       : "col_norms_noalloc: bad argument");
        return IB200_ERR_ARG;
    }
    if (norm_kind == 0 && (!V || nRows % dim != 0)) { set_error(h, "col_norms_noalloc: the L2 norm needs nRows / dim cell volumes"); return IB200_ERR_ARG; }
    IB_CUDA(h, cudaSetDevice(h->device));
    double* scratch = nullptr;
    int rc = na_scratch_ptr(h, &scratch);
    if (rc) return rc;
    cudaStream_t st = stream ? (cudaStream_t)stream : h->stream;
    const int nblk = (int)std::min<long long>((nRows + 4095) / 4096, NA_BLOCKS);
    na_norm_part_kernel<<<dim3(nblk, nCols), 256, 0, st>>>(Phi, nRows, ld, nCols, V, dim, norm_kind, scratch);
    na_norm_reduce_kernel<<<nCols, 256, 0, st>>>(scratch, nblk, nCols, norms_out);
    IB_CUDA(h, cudaGetLastError());
    h->launches += 2;
    if (!stream) IB_CUDA(h, cudaStreamSynchronize(st));   // blocking like the other entry points; with a caller stream: enqueue only
    return IB200_OK;
}

}  // extern "C"
