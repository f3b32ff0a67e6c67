/*
 * ithaca_b200.h -- C ABI of libithaca_b200.so: the B200 (sm_100a) implementation of ITHACA-FV's
 * reduced-basis hot path (method-of-snapshots POD + Galerkin operator projection).
 *
 * This header is the drop-in boundary. Each entry point names the reference interface it replaces
 * (paths relative to the ITHACA-FV source tree). Plain pointers and sizes only; no C++ types, no
 * exceptions cross the boundary. All functions return 0 on success and a non-zero status otherwise;
 * the message is available from ithaca_b200_last_error(). There is no CPU fallback: every compute
 * call fails with IB200_ERR_CUDA when no sm_100 device is usable.
 *
 * Conventions
 *   - all matrices are FP64, column-major (Eigen's default), sizes as 64-bit where they scale with
 *     the mesh;
 *   - a "snapshot" is one field's internal values exactly as OpenFOAM stores them: nCells*dim
 *     doubles, components interleaved per cell (x0,y0,z0,x1,...) -- src/ITHACA_CORE/Foam2Eigen/
 *     Foam2Eigen.C:49-56 (field2Eigen maps the raw storage);
 *   - "host or device pointer": the library copies with cudaMemcpyDefault, so either is accepted;
 *   - calls on one handle are blocking and not re-entrant; one handle per process and GPU.
 */
#ifndef ITHACA_B200_H
#define ITHACA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ithaca_b200 ithaca_b200_t;

enum {
    IB200_OK = 0,
    IB200_ERR_ARG = 1,    /* bad argument / call order (the reference would M_Assert -> abort) */
    IB200_ERR_CUDA = 2,   /* CUDA / driver failure, or no usable sm_100 device */
    IB200_ERR_NCCL = 3,   /* communicator failure */
    IB200_ERR_NOCONV = 4  /* eigensolver did not converge ("The Eigenvalue Decomposition did not succeed",
                             src/ITHACA_CORE/ITHACAPOD/ITHACAPOD.C:167,175) */
};

/* ---- lifetime ----------------------------------------------------------------------------- */
int ithaca_b200_create(ithaca_b200_t** h, int device);
int ithaca_b200_destroy(ithaca_b200_t* h);
/* h may be NULL: returns the last error of a failed create in this thread. */
const char* ithaca_b200_last_error(const ithaca_b200_t* h);
/* cudaStream_t every kernel of this handle is launched on (for event timing by the caller). */
void* ithaca_b200_stream(ithaca_b200_t* h);
/* Number of kernels launched by this handle since create / since the last reset. */
int64_t ithaca_b200_launch_count(ithaca_b200_t* h, int reset);
/* Per-stage device times of the last POD call sequence, milliseconds, measured with CUDA events:
 * out[0]=upload(H2D) out[1]=gram out[2]=allreduce out[3]=eig out[4]=lift(+norm) out[5]=scale
 * out[6]=download(D2H) out[7]=gram kernel only. n = number of doubles available in out. */
int ithaca_b200_timings(ithaca_b200_t* h, double* out, int n);

/* ---- multi-GPU: cells are sharded over ranks, one rank per GPU ----------------------------- *
 * Replaces OpenFOAM's Pstream reduce of the Gram matrix and of the mode norms
 * (src/ITHACA_CORE/ITHACAutilities/ITHACAcoeffsMass.C:112-115, src/ITHACA_CORE/ITHACAPOD/
 * ITHACAPOD.C:146-149 and :235-243) by ONE NCCL all-reduce each. Rank 0 obtains a unique id,
 * the caller distributes its 128 bytes to all ranks (the OpenFOAM shim does that with Pstream,
 * bench.py with torch.distributed), every rank calls comm_init. */
int ithaca_b200_nccl_unique_id(void* id128);
int ithaca_b200_comm_init(ithaca_b200_t* h, const void* id128, int rank, int nranks);
/* Contiguous cell range of `rank` out of `nranks`: [floor(rank*nCells/nranks), floor((rank+1)*nCells/nranks)).
 * The same rule as the reference's parity statement in SURVEY.md section 8(e). Pure host arithmetic. */
int ithaca_b200_cell_range(int64_t nCells, int rank, int nranks, int64_t* begin, int64_t* end);

/* ---- POD: ITHACAPOD::getModes (src/ITHACA_CORE/ITHACAPOD/ITHACAPOD.C:92-323) --------------- */

/* Cell volumes of this rank's cells: mesh.V(). The library repeats each entry dim times on the fly;
 * replaces ITHACAutilities::getMassMatrixFV (ITHACAcoeffsMass.C:312-326). V: host or device. */
int ithaca_b200_pod_set_volumes(ithaca_b200_t* h, const double* V, int64_t nCells);

/* Stream m snapshots into the device-resident snapshot matrix (one async copy per snapshot, no host
 * staging copy). snap_ptrs[j] points at nCells*dim contiguous doubles = &field.ref()[0][0].
 * Replaces Foam2Eigen::PtrList2Eigen (Foam2Eigen.C:664-692), which the reference calls twice per
 * getModes (ITHACAPOD.C:132 and ITHACAcoeffsMass.C:85). Pointers: host or device. */
int ithaca_b200_pod_upload(ithaca_b200_t* h, const double* const* snap_ptrs, int m, int dim,
                           int64_t nCells);

/* G = S^T diag(V) S, all-reduced over ranks when a communicator is set. Replaces
 * ITHACAutilities::getMassMatrix, 1-list form (ITHACAcoeffsMass.C:67-118), including its reduce.
 * The Gram is volume-weighted under both the "L2" and the "Frobenius" POD norm, as in the reference
 * (ITHACAcoeffsMass.C:87-105 ignores consider_volumes for vol fields). G_out: m*m doubles, host, or NULL. */
int ithaca_b200_pod_gram(ithaca_b200_t* h, double* G_out);

/* Symmetric eigensolve of the Gram by one-sided Jacobi; replaces Spectra::SymEigsSolver /
 * Eigen::SelfAdjointEigenSolver at ITHACAPOD.C:157-180. evals_out: ALL m eigenvalues, descending
 * (algebraic), un-shifted, un-normalised: the caller applies the reference's solver-dependent
 * post-processing (ITHACAPOD.C:169,179,182-186,281-287) -- see ithaca_b200_pod_postprocess_eigenvalues.
 * evecs_out: m*r doubles col-major, the r leading eigenvectors (may be NULL). sweeps_out may be NULL. */
int ithaca_b200_pod_eig(ithaca_b200_t* h, int r, double* evals_out, double* evecs_out, int* sweeps_out);

/* Install a caller-supplied m*m Gram (host) instead of computing it; used by tests of the eigensolver. */
int ithaca_b200_pod_set_gram(ithaca_b200_t* h, const double* G, int m);

/* Host arithmetic of ITHACAPOD.C:169/179 (which values are kept), :182-186 (negative shift) and
 * :281-287 (normalisation, cumulative sum). solver: 0 = "spectra" (keeps nmodes values),
 * 1 = "eigen" (keeps all m). evals_all: m values from pod_eig. Outputs hold *n_out entries (<= m). */
int ithaca_b200_pod_postprocess_eigenvalues(const double* evals_all, int m, int nmodes, int solver,
                                            double* evals_norm_out, double* cum_out, int* n_out);

/* modes = S * Q_r, then each column divided by its norm: sqrt(phi^T diag(V) phi) (norm_kind 0 = "L2")
 * or sqrt(phi^T phi) (1 = "Frobenius"); partial sums all-reduced over ranks before the sqrt.
 * Replaces ITHACAPOD.C:194-243 and the internal-field part of :254-264. mode_ptrs[i]: nCells*dim
 * doubles each (host or device), or mode_ptrs == NULL to keep the modes on the device only.
 * normfact_out: r doubles (host) or NULL. */
int ithaca_b200_pod_lift(ithaca_b200_t* h, int r, int norm_kind, double* const* mode_ptrs,
                         double* normfact_out);

/* Boundary rows: out = (Sbc * Q_r) with column i divided by normfact_i. Replaces ITHACAPOD.C:245-264
 * (modesEigBC). Sbc: rows*m col-major (what Foam2Eigen::PtrList2EigenBC returns for one patch,
 * Foam2Eigen.C:272-411); out: rows*r col-major. Host pointers. Must follow pod_lift. */
int ithaca_b200_pod_lift_patch(ithaca_b200_t* h, const double* Sbc, int64_t rows, double* out);

/* ---- measurement helpers -------------------------------------------------------------------- */
/* Register-resident DMMA.8x8x4 (mma.sync.m8n8k4.f64) issue-rate microbenchmark on `device`:
 * the measured FP64 tensor peak used as the Gram kernel's roofline denominator. */
int ithaca_b200_dmma_peak(int device, double* tflops_out);

#ifdef __cplusplus
}
#endif
#endif /* ITHACA_B200_H */
