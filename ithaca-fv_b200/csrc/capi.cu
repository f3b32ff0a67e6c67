// C ABI of libithaca_b200.so (include/ithaca_b200.h): handle, device-resident snapshot matrix,
// NCCL communicator (loaded with dlopen, so the library links against the CUDA runtime only), and the
// call sequence gram -> eig -> lift that replaces the dense part of ITHACAPOD::getModes
// (src/ITHACA_CORE/ITHACAPOD/ITHACAPOD.C:132-264).
#include "../../include/ithaca_b200.h"
#include "common.cuh"
#include <dlfcn.h>
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <set>

static thread_local std::string g_create_error;

namespace ib200 {

void set_error(ithaca_b200* h, const std::string& msg)
{
    if (h) h->err = msg;
    else g_create_error = msg;
}

const char* env_knob(const char* name)
{
    const char* v = getenv(name);
    if (!v || !*v) return nullptr;
    static std::mutex mu;
    static std::set<std::string> seen;
    std::lock_guard<std::mutex> lk(mu);
    if (seen.insert(name).second)
        fprintf(stderr, "[libithaca_b200] override active: %s=%s (measurement knob: changes kernel selection / schedule, not the "
                        "defaults the parity tests cover)\n", name, v);
    return v;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;

int make_tmap_2d(ithaca_b200* h, CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer,
                 uint64_t pitch_bytes, uint32_t box_inner, uint32_t box_outer, bool swizzle128)
{
    if (!g_encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
        if (e != cudaSuccess || !fn) {
            set_error(h, "cuTensorMapEncodeTiled entry point not available");
            return IB200_ERR_CUDA;
        }
        g_encode = (EncodeTiledFn)fn;
    }
    cuuint64_t dims[2] = {inner, outer};
    cuuint64_t strides[1] = {pitch_bytes};
    cuuint32_t box[2] = {box_inner, box_outer};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, const_cast<void*>(base), dims, strides, box,
                          estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        char buf[160];
        snprintf(buf, sizeof(buf), "cuTensorMapEncodeTiled failed (%d): inner=%llu outer=%llu pitch=%llu box=%ux%u",
                 (int)r, (unsigned long long)inner, (unsigned long long)outer, (unsigned long long)pitch_bytes,
                 box_inner, box_outer);
        set_error(h, buf);
        return IB200_ERR_CUDA;
    }
    return IB200_OK;
}

// ---- NCCL through dlopen: only the five entry points the path needs ----
struct NcclApi {
    void* lib = nullptr;
    int (*GetUniqueId)(void*) = nullptr;
    int (*CommInitRank)(void**, int, /* ncclUniqueId by value: 128 bytes */ struct Id128, int) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*CommDestroy)(void*) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};
struct Id128 { char b[128]; };
static NcclApi g_nccl;

static int load_nccl(ithaca_b200* h)
{
    if (g_nccl.lib) return IB200_OK;
    const char* names[] = {getenv("ITHACA_B200_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    void* lib = nullptr;
    for (const char* n : names) {
        if (!n || !*n) continue;
        lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (lib) break;
    }
    if (!lib) {
        set_error(h, "cannot dlopen libnccl.so.2 (set ITHACA_B200_NCCL_LIB)");
        return IB200_ERR_NCCL;
    }
    g_nccl.GetUniqueId = (decltype(g_nccl.GetUniqueId))dlsym(lib, "ncclGetUniqueId");
    g_nccl.CommInitRank = (decltype(g_nccl.CommInitRank))dlsym(lib, "ncclCommInitRank");
    g_nccl.AllReduce = (decltype(g_nccl.AllReduce))dlsym(lib, "ncclAllReduce");
    g_nccl.CommDestroy = (decltype(g_nccl.CommDestroy))dlsym(lib, "ncclCommDestroy");
    g_nccl.GetErrorString = (decltype(g_nccl.GetErrorString))dlsym(lib, "ncclGetErrorString");
    if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.AllReduce || !g_nccl.CommDestroy) {
        set_error(h, "libnccl is missing required symbols");
        return IB200_ERR_NCCL;
    }
    g_nccl.lib = lib;
    return IB200_OK;
}

static int nccl_fail(ithaca_b200* h, const char* what, int rc)
{
    std::string msg = std::string(what) + " failed: ";
    msg += g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "nccl error";
    set_error(h, msg);
    return IB200_ERR_NCCL;
}

// In-place sum all-reduce of `count` doubles on the handle's stream (no-op for a single rank).
int allreduce_sum(ithaca_b200* h, double* dbuf, size_t count)
{
    if (h->nranks <= 1 || !h->nccl_comm) return IB200_OK;
    const int ncclFloat64 = 8, ncclSum = 0;
    int rc = g_nccl.AllReduce(dbuf, dbuf, count, ncclFloat64, ncclSum, h->nccl_comm, h->stream);
    if (rc) return nccl_fail(h, "ncclAllReduce", rc);
    return IB200_OK;
}

static int ensure(ithaca_b200* h, double** p, size_t* cap, size_t bytes)
{
    if (bytes <= *cap) return IB200_OK;
    if (*p) cudaFree(*p);
    *p = nullptr;
    *cap = 0;
    IB_CUDA(h, cudaMalloc(p, bytes));
    *cap = bytes;
    return IB200_OK;
}

static int ensure_mats(ithaca_b200* h, int m)
{
    if (m <= h->mat_cap) return IB200_OK;
    for (double** p : {&h->dG, &h->dQ, &h->dW, &h->dEvals}) {
        if (*p) cudaFree(*p);
        *p = nullptr;
    }
    h->mat_cap = 0;
    const size_t mm = (size_t)m * m * sizeof(double);
    IB_CUDA(h, cudaMalloc(&h->dG, mm));
    IB_CUDA(h, cudaMalloc(&h->dQ, mm));
    IB_CUDA(h, cudaMalloc(&h->dW, mm));
    IB_CUDA(h, cudaMalloc(&h->dEvals, (size_t)m * sizeof(double)));
    h->mat_cap = m;
    return IB200_OK;
}

static void tstart(ithaca_b200* h, int which) { cudaEventRecord(h->timers[which].a, h->stream); }
static void tstop(ithaca_b200* h, int which)
{
    cudaEventRecord(h->timers[which].b, h->stream);
    h->timers[which].used = true;
}

}  // namespace ib200

using namespace ib200;

extern "C" {

int ithaca_b200_create(ithaca_b200_t** out, int device)
{
    if (!out) return IB200_ERR_ARG;
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        set_error(nullptr, std::string("no CUDA device: ") + cudaGetErrorString(e) +
                               " (libithaca_b200 has no CPU fallback)");
        return IB200_ERR_CUDA;
    }
    if (device < 0 || device >= ndev) {
        set_error(nullptr, "device index out of range");
        return IB200_ERR_ARG;
    }
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, device);
    if (e != cudaSuccess) {
        set_error(nullptr, cudaGetErrorString(e));
        return IB200_ERR_CUDA;
    }
    if (prop.major != 10) {
        set_error(nullptr, "libithaca_b200 is built for sm_100a only; device is sm_" +
                               std::to_string(prop.major) + std::to_string(prop.minor));
        return IB200_ERR_CUDA;
    }
    ithaca_b200* h = new ithaca_b200();
    h->device = device;
    h->num_sms = prop.multiProcessorCount;
    if (cudaSetDevice(device) != cudaSuccess ||
        cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess) {
        set_error(nullptr, "cannot create stream");
        delete h;
        return IB200_ERR_CUDA;
    }
    for (int i = 0; i < T_COUNT; i++) {
        cudaEventCreate(&h->timers[i].a);
        cudaEventCreate(&h->timers[i].b);
    }
    cudaMalloc(&h->dFlags, 64 * sizeof(int));
    *out = h;
    return IB200_OK;
}

int ithaca_b200_destroy(ithaca_b200_t* h)
{
    if (!h) return IB200_OK;
    cudaSetDevice(h->device);
    cudaStreamSynchronize(h->stream);
    proj_destroy(h);
    if (h->nccl_comm && g_nccl.CommDestroy) g_nccl.CommDestroy(h->nccl_comm);
    for (void* p : {(void*)h->dV, (void*)h->dS, (void*)h->dG, (void*)h->dWs, (void*)h->dQ, (void*)h->dW,
                    (void*)h->dEvals, (void*)h->dQr, (void*)h->dPhi, (void*)h->dNormPart, (void*)h->dNorm,
                    h->dJobs, (void*)h->dFlags, h->wexp[0].p, h->wexp[1].p, (void*)h->dS2, (void*)h->dPhiTmp, h->volT.p, h->eigTmp.p})
        if (p) cudaFree(p);
    for (int i = 0; i < T_COUNT; i++) {
        cudaEventDestroy(h->timers[i].a);
        cudaEventDestroy(h->timers[i].b);
    }
    for (auto& sl : h->slot) {
        if (sl.V) cudaFree(sl.V);
        if (sl.ready) cudaEventDestroy(sl.ready);
        if (sl.freed) cudaEventDestroy(sl.freed);
    }
    if (h->dGacc) cudaFree(h->dGacc);
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    cudaStreamDestroy(h->stream);
    delete h;
    return IB200_OK;
}

const char* ithaca_b200_last_error(const ithaca_b200_t* h) { return h ? h->err.c_str() : g_create_error.c_str(); }

void* ithaca_b200_stream(ithaca_b200_t* h) { return h ? (void*)h->stream : nullptr; }

int64_t ithaca_b200_launch_count(ithaca_b200_t* h, int reset)
{
    if (!h) return 0;
    int64_t n = h->launches;
    if (reset) h->launches = 0;
    return n;
}

int ithaca_b200_timings(ithaca_b200_t* h, double* out, int n)
{
    if (!h || !out) return IB200_ERR_ARG;
    cudaSetDevice(h->device);
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    for (int i = 0; i < n; i++) {
        out[i] = 0.0;
        if (i < T_COUNT && h->timers[i].used) {
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, h->timers[i].a, h->timers[i].b) == cudaSuccess) out[i] = ms;
        }
    }
    return IB200_OK;
}

int ithaca_b200_nccl_unique_id(void* id128)
{
    if (!id128) return IB200_ERR_ARG;
    int rc = load_nccl(nullptr);
    if (rc) return rc;
    rc = g_nccl.GetUniqueId(id128);
    if (rc) return nccl_fail(nullptr, "ncclGetUniqueId", rc);
    return IB200_OK;
}

int ithaca_b200_comm_init(ithaca_b200_t* h, const void* id128, int rank, int nranks)
{
    if (!h || nranks < 1 || rank < 0 || rank >= nranks) return IB200_ERR_ARG;
    h->rank = rank;
    h->nranks = nranks;
    if (nranks == 1) return IB200_OK;
    if (!id128) { set_error(h, "comm_init: id128 is null"); return IB200_ERR_ARG; }
    int rc = load_nccl(h);
    if (rc) return rc;
    cudaSetDevice(h->device);
    Id128 id;
    memcpy(&id, id128, sizeof(id));
    rc = g_nccl.CommInitRank(&h->nccl_comm, nranks, id, rank);
    if (rc) return nccl_fail(h, "ncclCommInitRank", rc);
    return IB200_OK;
}

int ithaca_b200_cell_range(int64_t nCells, int rank, int nranks, int64_t* begin, int64_t* end)
{
    if (nCells < 0 || nranks < 1 || rank < 0 || rank >= nranks || !begin || !end) return IB200_ERR_ARG;
    *begin = (int64_t)(((__int128)rank * nCells) / nranks);
    *end = (int64_t)(((__int128)(rank + 1) * nCells) / nranks);
    return IB200_OK;
}

int ithaca_b200_pod_set_volumes(ithaca_b200_t* h, const double* V, int64_t nCells)
{
    if (!h || !V || nCells <= 0) { set_error(h, "pod_set_volumes: bad argument"); return IB200_ERR_ARG; }
    cudaSetDevice(h->device);
    if (nCells > h->dV_cap) {
        if (h->dV) cudaFree(h->dV);
        h->dV = nullptr; h->dV_cap = 0;
        IB_CUDA(h, cudaMalloc(&h->dV, (size_t)nCells * sizeof(double)));
        h->dV_cap = nCells;
    }
    IB_CUDA(h, cudaMemcpyAsync(h->dV, V, (size_t)nCells * sizeof(double), cudaMemcpyDefault, h->stream));
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    h->nCells = nCells;
    h->vol_gen++;
    h->have_gram = h->have_eig = h->have_lift = false;
    return IB200_OK;
}

int ithaca_b200_pod_upload(ithaca_b200_t* h, const double* const* snap_ptrs, int m, int dim, int64_t nCells)
{
    if (!h || !snap_ptrs || m <= 0 || dim <= 0 || nCells <= 0) { set_error(h, "pod_upload: bad argument"); return IB200_ERR_ARG; }
    if (h->nCells != nCells) { set_error(h, "pod_upload: nCells differs from pod_set_volumes"); return IB200_ERR_ARG; }
    cudaSetDevice(h->device);
    const int64_t Nd = nCells * dim;
    const int64_t ld = (Nd + 15) / 16 * 16;
    int rc = ensure(h, &h->dS, &h->dS_cap, (size_t)ld * m * sizeof(double));
    if (rc) return rc;
    tstart(h, T_UPLOAD);
    for (int j = 0; j < m; j++) {
        if (!snap_ptrs[j]) { set_error(h, "pod_upload: null snapshot pointer"); return IB200_ERR_ARG; }
        IB_CUDA(h, cudaMemcpyAsync(h->dS + (size_t)j * ld, snap_ptrs[j], (size_t)Nd * sizeof(double),
                                   cudaMemcpyDefault, h->stream));
    }
    tstop(h, T_UPLOAD);
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    h->m = m; h->dim = dim; h->Nd = Nd; h->ld = ld;
    h->block_mode = false;
    h->stream_mode = false;
    h->second_list = false;
    h->slot[0].ncells = h->slot[1].ncells = 0;
    h->have_gram = h->have_eig = h->have_lift = h->have_modes = false;
    return IB200_OK;
}

// getModes(snapshots, modes, fields2, ...) (ITHACAPOD.C:1940-2086): the correlation matrix and the eigenvectors come from fields2,
// the modes are combinations of ANOTHER field list of the same length. Replaces the resident snapshot matrix, keeps the eigensolve.
int ithaca_b200_pod_upload_second(ithaca_b200_t* h, const double* const* snap_ptrs, int m, int64_t nValues)
{
    if (!h) return IB200_ERR_ARG;
    if (!snap_ptrs || m <= 0 || nValues <= 0) { set_error(h, "pod_upload_second: bad argument"); return IB200_ERR_ARG; }
    if (!h->have_eig || h->block_mode || h->stream_mode || m != h->m) {
        set_error(h, "pod_upload_second: needs pod_upload / pod_gram / pod_eig of a field list with the same number of snapshots");
        return IB200_ERR_ARG;
    }
    cudaSetDevice(h->device);
    const int64_t ld = (nValues + 15) / 16 * 16;
    int rc = ensure(h, &h->dS, &h->dS_cap, (size_t)ld * m * sizeof(double));
    if (rc) return rc;
    tstart(h, T_UPLOAD);
    for (int j = 0; j < m; j++) {
        if (!snap_ptrs[j]) { set_error(h, "pod_upload_second: null snapshot pointer"); return IB200_ERR_ARG; }
        IB_CUDA(h, cudaMemcpyAsync(h->dS + (size_t)j * ld, snap_ptrs[j], (size_t)nValues * sizeof(double), cudaMemcpyDefault, h->stream));
    }
    tstop(h, T_UPLOAD);
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    h->dim = 1; h->Nd = nValues; h->ld = ld;
    h->second_list = true;         // pod_gram is refused until the next pod_upload: the volumes belong to the first list
    h->have_gram = false;          // the resident matrix no longer belongs to the Gram; the eigenvectors stay
    h->have_lift = h->have_modes = false;
    return IB200_OK;
}

int ithaca_b200_pod_gram(ithaca_b200_t* h, double* G_out)
{
    if (!h) return IB200_ERR_ARG;
    if (!h->dS || h->m <= 0 || !h->dV || h->stream_mode || h->second_list) { set_error(h, "pod_gram: upload snapshots and volumes first"); return IB200_ERR_ARG; }
    cudaSetDevice(h->device);
    int rc = ensure_mats(h, h->m);
    if (rc) return rc;
    tstart(h, T_GRAM);
    rc = gram_launch(h);
    if (rc) return rc;
    tstop(h, T_GRAM);
    tstart(h, T_ALLREDUCE);
    rc = allreduce_sum(h, h->dG, (size_t)h->m * h->m);
    if (rc) return rc;
    tstop(h, T_ALLREDUCE);
    if (G_out)
        IB_CUDA(h, cudaMemcpyAsync(G_out, h->dG, (size_t)h->m * h->m * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    h->have_gram = true;
    h->have_eig = h->have_lift = false;
    return IB200_OK;
}

// T[:, j] = S[:, j] * sqrt|Vs[:, j]| in place over the uploaded volume columns (pad rows -> 0)
__global__ void __launch_bounds__(256)
scale_by_sqrt_abs_kernel(const double* __restrict__ S, double* __restrict__ T, long long Nd, long long ld, int m)
{
    const long long n = ld * m;
    for (long long i = blockIdx.x * 256LL + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
        const long long k = i % ld;
        T[i] = k < Nd ? S[i] * sqrt(fabs(T[i])) : 0.0;
    }
}

int ithaca_b200_pod_gram_volumes(ithaca_b200_t* h, const double* const* vol_ptrs, double* G_out)
{
    if (!h) return IB200_ERR_ARG;
    if (!h->dS || h->m <= 0 || h->block_mode || h->stream_mode || h->second_list) { set_error(h, "pod_gram_volumes: upload snapshots first"); return IB200_ERR_ARG; }
    if (!vol_ptrs) { set_error(h, "pod_gram_volumes: null volume list"); return IB200_ERR_ARG; }
    cudaSetDevice(h->device);
    int rc = ensure_mats(h, h->m);
    if (rc) return rc;
    if ((rc = dev_ensure(h, h->volT, (size_t)h->ld * h->m * sizeof(double)))) return rc;
    double* T = h->volT.as<double>();
    tstart(h, T_UPLOAD);
    for (int j = 0; j < h->m; j++) {
        if (!vol_ptrs[j]) { set_error(h, "pod_gram_volumes: null volume pointer"); return IB200_ERR_ARG; }
        IB_CUDA(h, cudaMemcpyAsync(T + (size_t)j * h->ld, vol_ptrs[j], (size_t)h->Nd * sizeof(double), cudaMemcpyDefault, h->stream));
    }
    tstop(h, T_UPLOAD);
    tstart(h, T_GRAM);
    scale_by_sqrt_abs_kernel<<<8 * h->num_sms, 256, 0, h->stream>>>(h->dS, T, h->Nd, h->ld, h->m);
    IB_CUDA(h, cudaGetLastError());
    h->launches += 1;
    rc = atvb_launch(h, T, h->ld, h->m, T, h->ld, h->m, h->Nd, nullptr, 1, 1, h->dG, 1, 0x7fffffff, 0, h->m);
    if (rc) return rc;
    tstop(h, T_GRAM);
    tstart(h, T_ALLREDUCE);
    rc = allreduce_sum(h, h->dG, (size_t)h->m * h->m);
    if (rc) return rc;
    tstop(h, T_ALLREDUCE);
    if (G_out)
        IB_CUDA(h, cudaMemcpyAsync(G_out, h->dG, (size_t)h->m * h->m * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    h->have_gram = true;
    h->have_eig = h->have_lift = false;
    return IB200_OK;
}

int ithaca_b200_pod_set_gram(ithaca_b200_t* h, const double* G, int m)
{
    if (!h || !G || m <= 0) return IB200_ERR_ARG;
    cudaSetDevice(h->device);
    int rc = ensure_mats(h, m);
    if (rc) return rc;
    if (h->m != m) { h->m = m; h->Nd = 0; }  // detached from any snapshot matrix
    IB_CUDA(h, cudaMemcpyAsync(h->dG, G, (size_t)m * m * sizeof(double), cudaMemcpyDefault, h->stream));
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    h->have_gram = true;
    h->have_eig = h->have_lift = false;
    return IB200_OK;
}

int ithaca_b200_pod_eig(ithaca_b200_t* h, int r, double* evals_out, double* evecs_out, int* sweeps_out)
{
    if (!h) return IB200_ERR_ARG;
    if (!h->have_gram) { set_error(h, "pod_eig: no Gram matrix (call pod_gram first)"); return IB200_ERR_ARG; }
    if (r < 0 || r > h->m) { set_error(h, "pod_eig: The number of requested modes cannot be bigger than the number of Snapshots"); return IB200_ERR_ARG; }
    cudaSetDevice(h->device);
    int sweeps = 0;
    tstart(h, T_EIG);
    int rc = eig_launch(h, &sweeps);
    if (rc) return rc;
    tstop(h, T_EIG);
    if (sweeps_out) *sweeps_out = sweeps;
    if (evals_out)
        IB_CUDA(h, cudaMemcpyAsync(evals_out, h->dEvals, (size_t)h->m * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    if (evecs_out && r > 0)
        IB_CUDA(h, cudaMemcpyAsync(evecs_out, h->dQ, (size_t)h->m * r * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    h->have_eig = true;
    h->have_lift = false;
    return IB200_OK;
}

int ithaca_b200_pod_postprocess_eigenvalues(const double* evals_all, int m, int nmodes, int solver,
                                            double* evals_norm_out, double* cum_out, int* n_out)
{
    if (!evals_all || m <= 0 || nmodes <= 0 || nmodes > m || !evals_norm_out || !n_out) return IB200_ERR_ARG;
    // ITHACAPOD.C:169 (spectra keeps nmodes values) / :179 (eigen keeps all m)
    const int n = (solver == 0) ? nmodes : m;
    double mn = evals_all[0];
    for (int i = 0; i < n; i++) mn = std::min(mn, evals_all[i]);
    // ITHACAPOD.C:182-186
    const double shift = (mn < 0) ? 2 * std::fabs(mn) : 0.0;
    double sum = 0.0;
    for (int i = 0; i < n; i++) { evals_norm_out[i] = evals_all[i] + shift; sum += evals_norm_out[i]; }
    // ITHACAPOD.C:281-287
    double c = 0.0;
    for (int i = 0; i < n; i++) {
        evals_norm_out[i] /= sum;
        c += evals_norm_out[i];
        if (cum_out) cum_out[i] = c;
    }
    *n_out = n;
    return IB200_OK;
}

int ithaca_b200_pod_lift(ithaca_b200_t* h, int r, int norm_kind, double* const* mode_ptrs, double* normfact_out)
{
    if (!h) return IB200_ERR_ARG;
    if (!h->have_eig || h->Nd <= 0 || h->block_mode || h->stream_mode) { set_error(h, "pod_lift: call pod_upload, pod_gram and pod_eig first"); return IB200_ERR_ARG; }
    if (r <= 0 || r > h->m) { set_error(h, "pod_lift: The number of requested modes cannot be bigger than the number of Snapshots"); return IB200_ERR_ARG; }
    if (norm_kind != 0 && norm_kind != 1) { set_error(h, "pod_lift: The PODnorm can be only L2 or Frobenius"); return IB200_ERR_ARG; }
    cudaSetDevice(h->device);
    int rc = lift_launch(h, r, norm_kind);
    if (rc) return rc;
    tstart(h, T_DOWNLOAD);
    if (mode_ptrs)
        for (int i = 0; i < r; i++) {
            if (!mode_ptrs[i]) { set_error(h, "pod_lift: null mode pointer"); return IB200_ERR_ARG; }
            IB_CUDA(h, cudaMemcpyAsync(mode_ptrs[i], h->dPhi + (size_t)i * h->ld, (size_t)h->Nd * sizeof(double),
                                       cudaMemcpyDefault, h->stream));
        }
    if (normfact_out)
        IB_CUDA(h, cudaMemcpyAsync(normfact_out, h->dNorm, (size_t)r * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    tstop(h, T_DOWNLOAD);
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    h->have_lift = true;
    h->have_modes = true;
    return IB200_OK;
}

int ithaca_b200_pod_lift_scaled(ithaca_b200_t* h, int r, const double* factors, double* const* mode_ptrs)
{
    if (!h) return IB200_ERR_ARG;
    if (!h->have_eig || h->Nd <= 0 || h->block_mode || h->stream_mode) { set_error(h, "pod_lift_scaled: call pod_upload, pod_gram and pod_eig first"); return IB200_ERR_ARG; }
    if (r <= 0 || r > h->m || !factors) { set_error(h, "pod_lift_scaled: bad argument"); return IB200_ERR_ARG; }
    cudaSetDevice(h->device);
    double* dF = nullptr;
    IB_CUDA(h, cudaMalloc(&dF, (size_t)r * sizeof(double)));
    cudaError_t e = cudaMemcpyAsync(dF, factors, (size_t)r * sizeof(double), cudaMemcpyDefault, h->stream);
    int rc = e == cudaSuccess ? lift_scaled_launch(h, r, dF) : IB200_ERR_CUDA;
    if (!rc && mode_ptrs)
        for (int i = 0; i < r && e == cudaSuccess; i++)
            e = mode_ptrs[i] ? cudaMemcpyAsync(mode_ptrs[i], h->dPhi + (size_t)i * h->ld, (size_t)h->Nd * sizeof(double), cudaMemcpyDefault, h->stream)
                             : cudaErrorInvalidValue;
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    cudaFree(dF);
    if (rc) return rc;
    IB_CUDA(h, e);
    h->have_lift = true;
    h->have_modes = true;
    return IB200_OK;
}

int ithaca_b200_pod_lift_patch(ithaca_b200_t* h, const double* Sbc, int64_t rows, double* out)
{
    if (!h || !Sbc || !out || rows <= 0) return IB200_ERR_ARG;
    if (!h->have_lift) { set_error(h, "pod_lift_patch: call pod_lift first"); return IB200_ERR_ARG; }
    cudaSetDevice(h->device);
    double *dIn = nullptr, *dOut = nullptr;
    IB_CUDA(h, cudaMalloc(&dIn, (size_t)rows * h->m * sizeof(double)));
    if (cudaMalloc(&dOut, (size_t)rows * h->r * sizeof(double)) != cudaSuccess) {
        cudaFree(dIn);
        set_error(h, "pod_lift_patch: out of device memory");
        return IB200_ERR_CUDA;
    }
    int rc = IB200_OK;
    do {
        if (cudaMemcpyAsync(dIn, Sbc, (size_t)rows * h->m * sizeof(double), cudaMemcpyDefault, h->stream) != cudaSuccess) { rc = IB200_ERR_CUDA; break; }
        rc = lift_patch_launch(h, dIn, rows, dOut);
        if (rc) break;
        if (cudaMemcpyAsync(out, dOut, (size_t)rows * h->r * sizeof(double), cudaMemcpyDefault, h->stream) != cudaSuccess) { rc = IB200_ERR_CUDA; break; }
        if (cudaStreamSynchronize(h->stream) != cudaSuccess) { rc = IB200_ERR_CUDA; break; }
    } while (0);
    if (rc == IB200_ERR_CUDA && h->err.empty()) set_error(h, cudaGetErrorString(cudaGetLastError()));
    cudaFree(dIn);
    cudaFree(dOut);
    return rc;
}

// ---- snapshot projection / coefficients / reconstruction on the resident modes ----------------

int ithaca_b200_pod_set_modes(ithaca_b200_t* h, const double* const* mode_ptrs, int r)
{
    if (!h) return IB200_ERR_ARG;
    if (!mode_ptrs || r <= 0) { set_error(h, "pod_set_modes: bad argument"); return IB200_ERR_ARG; }
    if (h->Nd <= 0) { set_error(h, "pod_set_modes: call pod_upload first (it fixes the field size)"); return IB200_ERR_ARG; }
    cudaSetDevice(h->device);
    const size_t need = (size_t)h->ld * r * sizeof(double);
    if (need > h->dPhi_cap) {
        if (h->dPhi) cudaFree(h->dPhi);
        h->dPhi = nullptr; h->dPhi_cap = 0;
        IB_CUDA(h, cudaMalloc(&h->dPhi, need));
        h->dPhi_cap = need;
    }
    for (int i = 0; i < r; i++) {
        if (!mode_ptrs[i]) { set_error(h, "pod_set_modes: null mode pointer"); return IB200_ERR_ARG; }
        IB_CUDA(h, cudaMemcpyAsync(h->dPhi + (size_t)i * h->ld, mode_ptrs[i], (size_t)h->Nd * sizeof(double), cudaMemcpyDefault, h->stream));
    }
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    h->r = r;
    h->have_modes = true;
    h->have_lift = false;   // lift_patch needs the eigenvectors and norm factors of a pod_lift
    return IB200_OK;
}

// Phi_n^T diag(w) X on the device into a fresh buffer (n x cols, col-major), all-reduced. X = dS (cols = m) or dPhi (cols = n).
static int modes_contract(ithaca_b200* h, int n, bool weighted, bool with_modes, double** dOut)
{
    const int cols = with_modes ? n : h->m;
    IB_CUDA(h, cudaMalloc(dOut, (size_t)n * cols * sizeof(double)));
    int rc = atvb_launch(h, h->dPhi, h->ld, n, with_modes ? h->dPhi : h->dS, h->ld, cols, h->Nd, weighted ? h->dV : nullptr, h->dim,
                         with_modes ? 1 : 0, *dOut, 1, 0x7fffffff, 0, n);
    if (!rc) rc = allreduce_sum(h, *dOut, (size_t)n * cols);
    if (rc) { cudaFree(*dOut); *dOut = nullptr; }
    return rc;
}

static int modes_args(ithaca_b200* h, const char* who, int& nmodes, const void* out)
{
    if (!h->have_modes || h->Nd <= 0) { set_error(h, std::string(who) + ": no resident modes (call pod_lift or pod_set_modes)"); return IB200_ERR_ARG; }
    if (h->second_list) { set_error(h, std::string(who) + ": the resident fields are the second list of pod_upload_second (no volumes of their own)"); return IB200_ERR_ARG; }
    if (!out) { set_error(h, std::string(who) + ": null output"); return IB200_ERR_ARG; }
    if (nmodes == 0) nmodes = h->r;
    if (nmodes < 0 || nmodes > h->r) {
        set_error(h, std::string(who) + ": The Number of requested modes is larger then the available quantity.");
        return IB200_ERR_ARG;
    }
    cudaSetDevice(h->device);
    return IB200_OK;
}

int ithaca_b200_pod_project(ithaca_b200_t* h, int nmodes, int proj_type, double* out)
{
    if (!h) return IB200_ERR_ARG;
    if (proj_type != 0 && proj_type != 1) { set_error(h, "pod_project: Projection type can be G for Galerking or F for Frobenius"); return IB200_ERR_ARG; }
    int rc = modes_args(h, "pod_project", nmodes, out);
    if (rc) return rc;
    double* d = nullptr;
    if ((rc = modes_contract(h, nmodes, proj_type == 0, false, &d))) return rc;
    cudaError_t e = cudaMemcpyAsync(out, d, (size_t)nmodes * h->m * sizeof(double), cudaMemcpyDefault, h->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    cudaFree(d);
    IB_CUDA(h, e);
    return IB200_OK;
}

int ithaca_b200_pod_modes_mass(ithaca_b200_t* h, int nmodes, double* out)
{
    if (!h) return IB200_ERR_ARG;
    int rc = modes_args(h, "pod_modes_mass", nmodes, out);
    if (rc) return rc;
    double* d = nullptr;
    if ((rc = modes_contract(h, nmodes, true, true, &d))) return rc;
    cudaError_t e = cudaMemcpyAsync(out, d, (size_t)nmodes * nmodes * sizeof(double), cudaMemcpyDefault, h->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    cudaFree(d);
    IB_CUDA(h, e);
    return IB200_OK;
}

// x = A^-1 B by LU with complete pivoting (what Eigen's fullPivLu().solve does), A: n x n, B: n x k, col-major, in place in B.
// Returns false when a pivot is exactly zero.
static bool full_piv_lu_solve(std::vector<double>& A, int n, double* B, int k)
{
    std::vector<int> colperm(n);
    for (int i = 0; i < n; i++) colperm[i] = i;
    for (int s = 0; s < n; s++) {
        int pr = s, pc = s;
        double best = -1.0;
        for (int c = s; c < n; c++)
            for (int r = s; r < n; r++)
                if (std::fabs(A[r + (size_t)c * n]) > best) { best = std::fabs(A[r + (size_t)c * n]); pr = r; pc = c; }
        if (!(best > 0.0)) return false;
        if (pr != s) {
            for (int c = 0; c < n; c++) std::swap(A[s + (size_t)c * n], A[pr + (size_t)c * n]);
            for (int j = 0; j < k; j++) std::swap(B[s + (size_t)j * n], B[pr + (size_t)j * n]);
        }
        if (pc != s) {
            for (int r = 0; r < n; r++) std::swap(A[r + (size_t)s * n], A[r + (size_t)pc * n]);
            std::swap(colperm[s], colperm[pc]);
        }
        const double piv = A[s + (size_t)s * n];
        for (int r = s + 1; r < n; r++) {
            const double l = A[r + (size_t)s * n] / piv;
            if (l == 0.0) continue;
            for (int c = s + 1; c < n; c++) A[r + (size_t)c * n] -= l * A[s + (size_t)c * n];
            for (int j = 0; j < k; j++) B[r + (size_t)j * n] -= l * B[s + (size_t)j * n];
        }
    }
    std::vector<double> y(n);
    for (int j = 0; j < k; j++) {
        double* b = B + (size_t)j * n;
        for (int r = n - 1; r >= 0; r--) {
            double v = b[r];
            for (int c = r + 1; c < n; c++) v -= A[r + (size_t)c * n] * y[c];
            y[r] = v / A[r + (size_t)r * n];
        }
        for (int c = 0; c < n; c++) b[colperm[c]] = y[c];
    }
    return true;
}

int ithaca_b200_pod_get_coeffs(ithaca_b200_t* h, int nmodes, double* out)
{
    if (!h) return IB200_ERR_ARG;
    int rc = modes_args(h, "pod_get_coeffs", nmodes, out);
    if (rc) return rc;
    std::vector<double> M((size_t)nmodes * nmodes);
    if ((rc = ithaca_b200_pod_modes_mass(h, nmodes, M.data()))) return rc;
    if ((rc = ithaca_b200_pod_project(h, nmodes, 0, out))) return rc;
    if (!full_piv_lu_solve(M, nmodes, out, h->m)) { set_error(h, "pod_get_coeffs: the mass matrix of the modes is singular"); return IB200_ERR_ARG; }
    return IB200_OK;
}

int ithaca_b200_pod_reconstruct(ithaca_b200_t* h, int nmodes, const double* coeff, int k, int clip_negative, double* const* out_ptrs)
{
    if (!h) return IB200_ERR_ARG;
    int rc = modes_args(h, "pod_reconstruct", nmodes, out_ptrs);
    if (rc) return rc;
    if (!coeff || k <= 0) { set_error(h, "pod_reconstruct: bad argument"); return IB200_ERR_ARG; }
    const int KB = 8;   // columns per pass (RC_K in lift.cu)
    double *dC = nullptr, *dOut = nullptr;
    IB_CUDA(h, cudaMalloc(&dC, (size_t)nmodes * k * sizeof(double)));
    if (cudaMalloc(&dOut, (size_t)h->ld * KB * sizeof(double)) != cudaSuccess) {
        cudaFree(dC);
        set_error(h, "pod_reconstruct: out of device memory");
        return IB200_ERR_CUDA;
    }
    cudaError_t e = cudaMemcpyAsync(dC, coeff, (size_t)nmodes * k * sizeof(double), cudaMemcpyDefault, h->stream);
    for (int j0 = 0; j0 < k && e == cudaSuccess && rc == IB200_OK; j0 += KB) {
        const int kc = std::min(KB, k - j0);
        rc = reconstruct_launch(h, nmodes, dC + (size_t)j0 * nmodes, kc, clip_negative, dOut);
        for (int j = 0; j < kc && e == cudaSuccess && rc == IB200_OK; j++) {
            if (!out_ptrs[j0 + j]) { set_error(h, "pod_reconstruct: null output pointer"); rc = IB200_ERR_ARG; break; }
            e = cudaMemcpyAsync(out_ptrs[j0 + j], dOut + (size_t)j * h->ld, (size_t)h->Nd * sizeof(double), cudaMemcpyDefault, h->stream);
        }
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    cudaFree(dC);
    cudaFree(dOut);
    if (rc) return rc;
    IB_CUDA(h, e);
    return IB200_OK;
}

// ---- snapshot-blocked POD (getModesMemoryEfficient, ITHACAPOD.C:344-628) ----------------------------------------

// G[J, I] = G[I, J]^T for an off-diagonal block pair (rows i0.., cols j0.. of the m x m col-major matrix)
__global__ void mirror_block_kernel(double* __restrict__ G, int m, int i0, int ni, int j0, int nj)
{
    const int total = ni * nj;
    for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
        const int i = idx % ni, j = idx / ni;
        G[(size_t)(j0 + j) + (size_t)(i0 + i) * m] = G[(size_t)(i0 + i) + (size_t)(j0 + j) * m];
    }
}

int ithaca_b200_pod_block_begin(ithaca_b200_t* h, int m_total, int dim, int64_t nCells)
{
    if (!h) return IB200_ERR_ARG;
    if (m_total <= 0 || dim <= 0 || nCells <= 0) { set_error(h, "pod_block_begin: bad argument"); return IB200_ERR_ARG; }
    if (h->nCells != nCells) { set_error(h, "pod_block_begin: nCells differs from pod_set_volumes"); return IB200_ERR_ARG; }
    cudaSetDevice(h->device);
    int rc = ensure_mats(h, m_total);
    if (rc) return rc;
    h->m = m_total; h->dim = dim; h->Nd = nCells * dim; h->ld = (h->Nd + 15) / 16 * 16;
    h->stream_mode = false;
    h->second_list = false;
    h->slot[0].ncells = h->slot[1].ncells = 0;
    IB_CUDA(h, cudaMemsetAsync(h->dG, 0, (size_t)m_total * m_total * sizeof(double), h->stream));
    h->blk_n[0] = h->blk_n[1] = 0;
    h->blk_lifted = 0;
    h->block_mode = true;
    h->have_gram = h->have_eig = h->have_lift = h->have_modes = false;
    return IB200_OK;
}

int ithaca_b200_pod_block_upload(ithaca_b200_t* h, int slot, const double* const* snap_ptrs, int nb, int first)
{
    if (!h) return IB200_ERR_ARG;
    if (!h->block_mode) { set_error(h, "pod_block_upload: call pod_block_begin first"); return IB200_ERR_ARG; }
    if ((slot != 0 && slot != 1) || !snap_ptrs || nb <= 0 || first < 0 || first + nb > h->m) {
        set_error(h, "pod_block_upload: bad argument");
        return IB200_ERR_ARG;
    }
    cudaSetDevice(h->device);
    double** buf = slot == 0 ? &h->dS : &h->dS2;
    size_t* cap = slot == 0 ? &h->dS_cap : &h->dS2_cap;
    int rc = ensure(h, buf, cap, (size_t)h->ld * nb * sizeof(double));
    if (rc) return rc;
    tstart(h, T_UPLOAD);
    for (int j = 0; j < nb; j++) {
        if (!snap_ptrs[j]) { set_error(h, "pod_block_upload: null snapshot pointer"); return IB200_ERR_ARG; }
        IB_CUDA(h, cudaMemcpyAsync(*buf + (size_t)j * h->ld, snap_ptrs[j], (size_t)h->Nd * sizeof(double), cudaMemcpyDefault, h->stream));
    }
    tstop(h, T_UPLOAD);
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    h->blk_n[slot] = nb; h->blk_first[slot] = first;
    return IB200_OK;
}

int ithaca_b200_pod_block_gram(ithaca_b200_t* h, int slotI, int slotJ, int norm_kind)
{
    if (!h) return IB200_ERR_ARG;
    if (!h->block_mode || (slotI != 0 && slotI != 1) || (slotJ != 0 && slotJ != 1) || h->blk_n[slotI] <= 0 || h->blk_n[slotJ] <= 0) {
        set_error(h, "pod_block_gram: upload both blocks first");
        return IB200_ERR_ARG;
    }
    if (norm_kind != 0 && norm_kind != 1) { set_error(h, "pod_block_gram: The PODnorm can be only L2 or Frobenius"); return IB200_ERR_ARG; }
    cudaSetDevice(h->device);
    const double* SI = slotI == 0 ? h->dS : h->dS2;
    const double* SJ = slotJ == 0 ? h->dS : h->dS2;
    const int ni = h->blk_n[slotI], nj = h->blk_n[slotJ], i0 = h->blk_first[slotI], j0 = h->blk_first[slotJ];
    tstart(h, T_GRAM);
    int rc = atvb_launch(h, SI, h->ld, ni, SJ, h->ld, nj, h->Nd, norm_kind == 0 ? h->dV : nullptr, h->dim, slotI == slotJ ? 1 : 0,
                         h->dG + (size_t)i0 + (size_t)j0 * h->m, 1, 0x7fffffff, 0, h->m);
    if (rc) return rc;
    if (slotI != slotJ) {
        mirror_block_kernel<<<std::min(1024, (ni * nj + 255) / 256), 256, 0, h->stream>>>(h->dG, h->m, i0, ni, j0, nj);
        h->launches += 1;
    }
    tstop(h, T_GRAM);
    IB_CUDA(h, cudaGetLastError());
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    return IB200_OK;
}

int ithaca_b200_pod_block_gram_finish(ithaca_b200_t* h, double* G_out)
{
    if (!h) return IB200_ERR_ARG;
    if (!h->block_mode) { set_error(h, "pod_block_gram_finish: call pod_block_begin first"); return IB200_ERR_ARG; }
    cudaSetDevice(h->device);
    tstart(h, T_ALLREDUCE);
    int rc = allreduce_sum(h, h->dG, (size_t)h->m * h->m);
    tstop(h, T_ALLREDUCE);
    if (rc) return rc;
    if (G_out) IB_CUDA(h, cudaMemcpyAsync(G_out, h->dG, (size_t)h->m * h->m * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    h->have_gram = true;
    h->have_eig = false;
    h->blk_lifted = 0;
    return IB200_OK;
}

// Checkpoint support of the blocked correlation matrix (PODTemplate::buildCovMatrix saves the matrix after every block
// pair, PODTemplate.C:487-505, and restarts from the file, :265-341).
int ithaca_b200_pod_block_gram_get(ithaca_b200_t* h, double* G_out)
{
    if (!h) return IB200_ERR_ARG;
    if (!h->block_mode || !G_out) { set_error(h, "pod_block_gram_get: call pod_block_begin first"); return IB200_ERR_ARG; }
    cudaSetDevice(h->device);
    const size_t n = (size_t)h->m * h->m;
    double* src = h->dG;
    if (h->nranks > 1 && h->nccl_comm) {
        // the resident matrix holds this rank's partial sums until pod_block_gram_finish: reduce a copy
        int rc = ensure(h, &h->dPhiTmp, &h->dPhiTmp_cap, n * sizeof(double));
        if (rc) return rc;
        IB_CUDA(h, cudaMemcpyAsync(h->dPhiTmp, h->dG, n * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
        if ((rc = allreduce_sum(h, h->dPhiTmp, n))) return rc;
        src = h->dPhiTmp;
    }
    IB_CUDA(h, cudaMemcpyAsync(G_out, src, n * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    return IB200_OK;
}

int ithaca_b200_pod_block_gram_set(ithaca_b200_t* h, const double* G)
{
    if (!h) return IB200_ERR_ARG;
    if (!h->block_mode || !G) { set_error(h, "pod_block_gram_set: call pod_block_begin first"); return IB200_ERR_ARG; }
    cudaSetDevice(h->device);
    const size_t n = (size_t)h->m * h->m;
    // G is the all-reduced matrix: rank 0 carries it, the other ranks restart their partial sums from zero
    if (h->rank == 0) IB_CUDA(h, cudaMemcpyAsync(h->dG, G, n * sizeof(double), cudaMemcpyDefault, h->stream));
    else IB_CUDA(h, cudaMemsetAsync(h->dG, 0, n * sizeof(double), h->stream));
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    h->have_gram = h->have_eig = false;
    return IB200_OK;
}

int ithaca_b200_pod_block_lift(ithaca_b200_t* h, int slot, int r)
{
    if (!h) return IB200_ERR_ARG;
    if (!h->block_mode || !h->have_eig || (slot != 0 && slot != 1) || h->blk_n[slot] <= 0) {
        set_error(h, "pod_block_lift: needs pod_block_gram_finish, pod_eig and an uploaded block");
        return IB200_ERR_ARG;
    }
    if (r <= 0 || r > h->m) { set_error(h, "pod_block_lift: The number of requested modes cannot be bigger than the number of Snapshots"); return IB200_ERR_ARG; }
    cudaSetDevice(h->device);
    const bool acc = h->blk_lifted > 0;
    if (acc) {
        int rc = ensure(h, &h->dPhiTmp, &h->dPhiTmp_cap, (size_t)h->ld * r * sizeof(double));
        if (rc) return rc;
    }
    int rc = lift_block_launch(h, slot == 0 ? h->dS : h->dS2, h->blk_n[slot], h->blk_first[slot], r, acc, h->dPhiTmp);
    if (rc) return rc;
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    h->blk_lifted++;
    return IB200_OK;
}

int ithaca_b200_pod_block_lift_finish(ithaca_b200_t* h, int r, int norm_kind, double* const* mode_ptrs, double* normfact_out)
{
    if (!h) return IB200_ERR_ARG;
    if (!h->block_mode || h->blk_lifted <= 0 || r != h->r) { set_error(h, "pod_block_lift_finish: call pod_block_lift for every block first (same r)"); return IB200_ERR_ARG; }
    if (norm_kind != 0 && norm_kind != 1) { set_error(h, "pod_block_lift_finish: The PODnorm can be only L2 or Frobenius"); return IB200_ERR_ARG; }
    cudaSetDevice(h->device);
    int rc = lift_block_finish(h, r, norm_kind);
    if (rc) return rc;
    tstart(h, T_DOWNLOAD);
    if (mode_ptrs)
        for (int i = 0; i < r; i++) {
            if (!mode_ptrs[i]) { set_error(h, "pod_block_lift_finish: null mode pointer"); return IB200_ERR_ARG; }
            IB_CUDA(h, cudaMemcpyAsync(mode_ptrs[i], h->dPhi + (size_t)i * h->ld, (size_t)h->Nd * sizeof(double), cudaMemcpyDefault, h->stream));
        }
    if (normfact_out) IB_CUDA(h, cudaMemcpyAsync(normfact_out, h->dNorm, (size_t)r * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    tstop(h, T_DOWNLOAD);
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    h->have_modes = true;
    h->have_lift = true;   // pod_lift_patch: eigenvectors, norm factors and r are those of this lift (ITHACAPOD.C:573-578)
    h->blk_lifted = 0;
    return IB200_OK;
}

// ---- row-streamed POD: a rank's cell shard larger than HBM passes through two chunk slots -----------------------------
// The reference's own decomposition over cells (ITHACAcoeffsMass.C:85-115: every rank forms the Gram of its rows, the partial
// matrices are summed) applied with more cell chunks than GPUs: G = sum over chunks of S_c^T V_c S_c, then modes rows = S_c Q_r.

__global__ void stream_acc_kernel(double* __restrict__ acc, const double* __restrict__ x, size_t n, int assign)
{
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        acc[i] = assign ? x[i] : acc[i] + x[i];
}

static int stream_slot_prepare(ithaca_b200* h, int slot, int64_t cell0, int64_t nc, bool host_waits)
{
    ithaca_b200::StreamSlot& sl = h->slot[slot];
    const int64_t Nd = nc * h->dim, ld = (Nd + 15) / 16 * 16;
    // the previous occupant's consumers (gram_add / lift_add on the compute stream) must be done before the slot is rewritten
    if (sl.freed_pending) {
        if (host_waits) IB_CUDA(h, cudaEventSynchronize(sl.freed));
        else IB_CUDA(h, cudaStreamWaitEvent(h->copy_stream, sl.freed, 0));
    }
    double** buf = slot == 0 ? &h->dS : &h->dS2;
    size_t* cap = slot == 0 ? &h->dS_cap : &h->dS2_cap;
    if ((size_t)ld * h->m * sizeof(double) > *cap || nc > sl.V_cap) {
        // growing a slot frees memory the device may still read: drain both streams first (only at the first chunks)
        IB_CUDA(h, cudaStreamSynchronize(h->stream));
        IB_CUDA(h, cudaStreamSynchronize(h->copy_stream));
        int rc = ensure(h, buf, cap, (size_t)ld * h->m * sizeof(double));
        if (rc) return rc;
        if (nc > sl.V_cap) {
            if (sl.V) cudaFree(sl.V);
            sl.V = nullptr; sl.V_cap = 0;
            IB_CUDA(h, cudaMalloc(&sl.V, (size_t)nc * sizeof(double)));
            sl.V_cap = nc;
        }
    }
    sl.cell0 = cell0; sl.ncells = nc; sl.ld = ld;
    sl.ready_pending = false;
    h->vol_gen++;   // the weights of this slot changed: the contraction rebuilds its repeated copy
    return IB200_OK;
}

static int stream_args(ithaca_b200* h, const char* who, int slot, int64_t cell0, int64_t nc)
{
    if (!h->stream_mode) { set_error(h, std::string(who) + ": call pod_stream_begin first"); return IB200_ERR_ARG; }
    if ((slot != 0 && slot != 1) || cell0 < 0 || nc <= 0 || cell0 + nc > h->strm_cells) { set_error(h, std::string(who) + ": bad argument"); return IB200_ERR_ARG; }
    cudaSetDevice(h->device);
    return IB200_OK;
}

int ithaca_b200_pod_stream_begin(ithaca_b200_t* h, int m, int dim, int64_t nCellsTotal)
{
    if (!h) return IB200_ERR_ARG;
    if (m <= 0 || dim <= 0 || nCellsTotal <= 0) { set_error(h, "pod_stream_begin: bad argument"); return IB200_ERR_ARG; }
    cudaSetDevice(h->device);
    int rc = ensure_mats(h, m);
    if (rc) return rc;
    if (!h->copy_stream) IB_CUDA(h, cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    for (auto& sl : h->slot) {
        if (!sl.ready) IB_CUDA(h, cudaEventCreateWithFlags(&sl.ready, cudaEventDisableTiming));
        if (!sl.freed) IB_CUDA(h, cudaEventCreateWithFlags(&sl.freed, cudaEventDisableTiming));
        sl.ncells = 0; sl.ready_pending = sl.freed_pending = false;
    }
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    IB_CUDA(h, cudaStreamSynchronize(h->copy_stream));
    if (m > h->gacc_cap) {
        if (h->dGacc) cudaFree(h->dGacc);
        h->dGacc = nullptr; h->gacc_cap = 0;
        IB_CUDA(h, cudaMalloc(&h->dGacc, (size_t)m * m * sizeof(double)));
        h->gacc_cap = m;
    }
    h->m = m; h->dim = dim;
    h->strm_cells = nCellsTotal; h->strm_Nd = nCellsTotal * dim; h->strm_ld = (h->strm_Nd + 15) / 16 * 16;
    h->Nd = 0; h->ld = 0; h->nCells = 0;          // no resident snapshot matrix in this mode
    h->strm_gram_chunks = h->strm_lift_chunks = 0;
    h->stream_mode = true; h->block_mode = false; h->second_list = false;
    h->have_gram = h->have_eig = h->have_lift = h->have_modes = false;
    return IB200_OK;
}

int ithaca_b200_pod_stream_slot(ithaca_b200_t* h, int slot, int64_t cell0, int64_t nCellsChunk, double** S_out, int64_t* ld_out,
                                double** V_out)
{
    if (!h) return IB200_ERR_ARG;
    int rc = stream_args(h, "pod_stream_slot", slot, cell0, nCellsChunk);
    if (rc) return rc;
    if ((rc = stream_slot_prepare(h, slot, cell0, nCellsChunk, true))) return rc;
    if (S_out) *S_out = slot == 0 ? h->dS : h->dS2;
    if (ld_out) *ld_out = h->slot[slot].ld;
    if (V_out) *V_out = h->slot[slot].V;
    return IB200_OK;
}

int ithaca_b200_pod_stream_upload(ithaca_b200_t* h, int slot, int64_t cell0, int64_t nCellsChunk, const double* V_chunk,
                                  const double* const* snap_ptrs)
{
    if (!h) return IB200_ERR_ARG;
    int rc = stream_args(h, "pod_stream_upload", slot, cell0, nCellsChunk);
    if (rc) return rc;
    if (!V_chunk || !snap_ptrs) { set_error(h, "pod_stream_upload: null pointer"); return IB200_ERR_ARG; }
    if ((rc = stream_slot_prepare(h, slot, cell0, nCellsChunk, false))) return rc;
    ithaca_b200::StreamSlot& sl = h->slot[slot];
    double* S = slot == 0 ? h->dS : h->dS2;
    const size_t bytes = (size_t)nCellsChunk * h->dim * sizeof(double);
    IB_CUDA(h, cudaMemcpyAsync(sl.V, V_chunk, (size_t)nCellsChunk * sizeof(double), cudaMemcpyDefault, h->copy_stream));
    for (int j = 0; j < h->m; j++) {
        if (!snap_ptrs[j]) { set_error(h, "pod_stream_upload: null snapshot pointer"); return IB200_ERR_ARG; }
        IB_CUDA(h, cudaMemcpyAsync(S + (size_t)j * sl.ld, snap_ptrs[j], bytes, cudaMemcpyDefault, h->copy_stream));
    }
    IB_CUDA(h, cudaEventRecord(sl.ready, h->copy_stream));
    sl.ready_pending = true;
    return IB200_OK;
}

// the compute stream picks up a slot: wait for its copies, remember that a consumer is in flight
static int stream_consume_begin(ithaca_b200* h, int slot)
{
    ithaca_b200::StreamSlot& sl = h->slot[slot];
    if (sl.ready_pending) IB_CUDA(h, cudaStreamWaitEvent(h->stream, sl.ready, 0));
    return IB200_OK;
}
static int stream_consume_end(ithaca_b200* h, int slot)
{
    ithaca_b200::StreamSlot& sl = h->slot[slot];
    IB_CUDA(h, cudaEventRecord(sl.freed, h->stream));
    sl.freed_pending = true;
    return IB200_OK;
}

int ithaca_b200_pod_stream_gram_add(ithaca_b200_t* h, int slot)
{
    if (!h) return IB200_ERR_ARG;
    if (!h->stream_mode || (slot != 0 && slot != 1) || h->slot[slot].ncells <= 0) { set_error(h, "pod_stream_gram_add: fill the slot first"); return IB200_ERR_ARG; }
    cudaSetDevice(h->device);
    const ithaca_b200::StreamSlot& sl = h->slot[slot];
    const double* S = slot == 0 ? h->dS : h->dS2;
    int rc = stream_consume_begin(h, slot);
    if (rc) return rc;
    tstart(h, T_GRAM);
    rc = atvb_launch(h, S, sl.ld, h->m, S, sl.ld, h->m, sl.ncells * h->dim, sl.V, h->dim, 1, h->dG, 1, 0x7fffffff, 0, h->m);
    if (rc) return rc;
    const size_t n = (size_t)h->m * h->m;
    stream_acc_kernel<<<(unsigned)std::min<size_t>((n + 255) / 256, 1024), 256, 0, h->stream>>>(h->dGacc, h->dG, n, h->strm_gram_chunks == 0);
    h->launches += 1;
    tstop(h, T_GRAM);
    IB_CUDA(h, cudaGetLastError());
    h->strm_gram_chunks++;
    h->have_gram = h->have_eig = false;
    return stream_consume_end(h, slot);
}

int ithaca_b200_pod_stream_gram_finish(ithaca_b200_t* h, double* G_out)
{
    if (!h) return IB200_ERR_ARG;
    if (!h->stream_mode || h->strm_gram_chunks <= 0) { set_error(h, "pod_stream_gram_finish: no chunk was added"); return IB200_ERR_ARG; }
    cudaSetDevice(h->device);
    const size_t n = (size_t)h->m * h->m;
    tstart(h, T_ALLREDUCE);
    IB_CUDA(h, cudaMemcpyAsync(h->dG, h->dGacc, n * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
    int rc = allreduce_sum(h, h->dG, n);
    tstop(h, T_ALLREDUCE);
    if (rc) return rc;
    if (G_out) IB_CUDA(h, cudaMemcpyAsync(G_out, h->dG, n * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    h->have_gram = true;
    h->have_eig = h->have_lift = false;
    h->strm_gram_chunks = 0;
    h->strm_lift_chunks = 0;
    return IB200_OK;
}

int ithaca_b200_pod_stream_lift_add(ithaca_b200_t* h, int slot, int r, int norm_kind)
{
    if (!h) return IB200_ERR_ARG;
    if (!h->stream_mode || !h->have_eig || (slot != 0 && slot != 1) || h->slot[slot].ncells <= 0) {
        set_error(h, "pod_stream_lift_add: needs pod_stream_gram_finish, pod_eig and a filled slot");
        return IB200_ERR_ARG;
    }
    if (r <= 0 || r > h->m) { set_error(h, "pod_stream_lift_add: The number of requested modes cannot be bigger than the number of Snapshots"); return IB200_ERR_ARG; }
    if (norm_kind != 0 && norm_kind != 1) { set_error(h, "pod_stream_lift_add: The PODnorm can be only L2 or Frobenius"); return IB200_ERR_ARG; }
    if (h->strm_lift_chunks > 0 && r != h->r) { set_error(h, "pod_stream_lift_add: r differs from the previous chunk"); return IB200_ERR_ARG; }
    cudaSetDevice(h->device);
    const ithaca_b200::StreamSlot& sl = h->slot[slot];
    int rc = stream_consume_begin(h, slot);
    if (rc) return rc;
    rc = lift_stream_add(h, slot == 0 ? h->dS : h->dS2, sl.ld, sl.ncells * h->dim, sl.V, r, norm_kind, sl.cell0 * h->dim, h->strm_ld,
                         h->strm_lift_chunks == 0);
    if (rc) return rc;
    h->strm_lift_chunks++;
    return stream_consume_end(h, slot);
}

int ithaca_b200_pod_stream_lift_finish(ithaca_b200_t* h, int r, double* const* mode_ptrs, double* normfact_out)
{
    if (!h) return IB200_ERR_ARG;
    if (!h->stream_mode || h->strm_lift_chunks <= 0 || r != h->r) { set_error(h, "pod_stream_lift_finish: call pod_stream_lift_add for every chunk first (same r)"); return IB200_ERR_ARG; }
    cudaSetDevice(h->device);
    int rc = lift_stream_finish(h, r, h->strm_Nd, h->strm_ld);
    if (rc) return rc;
    tstart(h, T_DOWNLOAD);
    if (mode_ptrs)
        for (int i = 0; i < r; i++) {
            if (!mode_ptrs[i]) { set_error(h, "pod_stream_lift_finish: null mode pointer"); return IB200_ERR_ARG; }
            IB_CUDA(h, cudaMemcpyAsync(mode_ptrs[i], h->dPhi + (size_t)i * h->strm_ld, (size_t)h->strm_Nd * sizeof(double), cudaMemcpyDefault, h->stream));
        }
    if (normfact_out) IB_CUDA(h, cudaMemcpyAsync(normfact_out, h->dNorm, (size_t)r * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    tstop(h, T_DOWNLOAD);
    IB_CUDA(h, cudaStreamSynchronize(h->stream));
    h->have_lift = true;    // pod_lift_patch works on the eigenvectors and norm factors of this lift
    h->strm_lift_chunks = 0;
    return IB200_OK;
}

}  // extern "C"
