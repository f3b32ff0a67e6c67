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

/* Correlation matrix of the moving-mesh getModes overload, one volume field PER SNAPSHOT
 * (getModes(snapshots, modes, Volumes, ...), ITHACAPOD.C:1392-1541, loop :1425-1445):
 *   G(i,j) = S_i^T diag(sqrt|V_i V_j|) S_j  =  (S_i sqrt|V_i|)^T (S_j sqrt|V_j|),
 * formed as ONE unweighted symmetric contraction of the row-scaled snapshot matrix (the reference: m(m+1)/2 weighted dot
 * products, each building an N_d temporary), all-reduced over ranks. vol_ptrs[j]: as many doubles as snapshot j (the
 * reference multiplies V.col(j) and SnapMatrix.col(j) coefficient-wise: scalar fields in practice), host or device.
 * Follow with pod_eig and pod_lift_scaled(factors = lambda^-1/2) (:1481-1484). G_out: m*m doubles, host, or NULL. */
int ithaca_b200_pod_gram_volumes(ithaca_b200_t* h, const double* const* vol_ptrs, double* G_out);

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

/* Lift with caller-given column factors instead of the norm normalisation: modes_i = (S q_i) * factors[i].
 * ITHACAPOD::getWeightedModes (ITHACAPOD.C:715-845) is this with factors[i] = lambda_i^(-1/2) * lambda_i (:771-777,
 * eigenvalues of pod_eig, spectra convention); pod_lift_patch afterwards applies the same factors to the boundary rows
 * (:780-785). factors: r doubles, host. mode_ptrs as in pod_lift. */
int ithaca_b200_pod_lift_scaled(ithaca_b200_t* h, int r, const double* factors, double* const* mode_ptrs);

/* getModes(snapshots, modes, fields2, ...) (ITHACAPOD.C:1940-2086): correlation matrix, eigensolve from `fields2` (pod_upload,
 * pod_gram, pod_eig as usual), the modes are combinations of ANOTHER list of m fields (surfaceScalarField fluxes or
 * volScalarFields in the reference's instantiations): modes_i = lambda_i^(-1/2) sum_j Q(j,i) snapshots[j] (:2016-2033).
 * This call replaces the resident snapshot matrix by that second list (nValues doubles per field, any field type, host or
 * device) and KEEPS the eigenvectors; follow with pod_lift_scaled(factors = |lambda_i|^(-1/2)) and, for the boundary values,
 * pod_lift_patch. pod_gram afterwards is refused until the next pod_upload. */
int ithaca_b200_pod_upload_second(ithaca_b200_t* h, const double* const* snap_ptrs, int m, int64_t nValues);

/* Boundary rows: out = (Sbc * Q_r) with column i divided by normfact_i. Replaces ITHACAPOD.C:245-264
 * (modesEigBC). Sbc: rows*m col-major (what Foam2Eigen::PtrList2EigenBC returns for one patch,
 * Foam2Eigen.C:272-411); out: rows*r col-major. Host pointers. Must follow pod_lift. */
int ithaca_b200_pod_lift_patch(ithaca_b200_t* h, const double* Sbc, int64_t rows, double* out);

/* ---- the callers either side of getModes: snapshot projection and reconstruction ----------
 * All of them work on the snapshot matrix of pod_upload and on the RESIDENT MODES: the result of pod_lift, or a mode
 * list uploaded with pod_set_modes. nmodes = 0 means all resident modes (the reference's "Nmodes == 0" convention). */

/* Install r modes (nCells*dim doubles each, host or device) as the resident modes; requires pod_upload (sizes). */
int ithaca_b200_pod_set_modes(ithaca_b200_t* h, const double* const* mode_ptrs, int r);

/* Modes::project over a field list (src/ITHACA_CORE/Containers/Modes.C:228-290):
 *   proj_type 0 = "G": Phi_n^T diag(V) S;  1 = "F": Phi_n^T S (the list overload replaces vol by ones, :246-249, :271-275).
 * out: nmodes x m col-major (host). All-reduced over ranks. */
int ithaca_b200_pod_project(ithaca_b200_t* h, int nmodes, int proj_type, double* out);

/* ITHACAutilities::getMassMatrix over the mode list (ITHACAcoeffsMass.C:67-118): Phi_n^T diag(V) Phi_n, nmodes x nmodes. */
int ithaca_b200_pod_modes_mass(ithaca_b200_t* h, int nmodes, double* out);

/* ITHACAutilities::getCoeffs over the snapshot list (ITHACAcoeffsMass.C:348-388 per snapshot, :424-443):
 * coeff(:, i) = M.fullPivLu().solve(Phi_n^T diag(V) s_i), M = the mass matrix above. out: nmodes x m col-major. */
int ithaca_b200_pod_get_coeffs(ithaca_b200_t* h, int nmodes, double* out);

/* Modes::reconstruct, internal field (Modes.C:294-305): out_j = Phi[:, 0:nmodes] * coeff(:, j) for j < k.
 * coeff: nmodes x k col-major (host). clip_negative != 0 applies the reference's "nut" rule (negative entries -> 0).
 * out_ptrs[j]: nCells*dim doubles (host or device). Boundary patches: ithaca_b200_pod_lift_patch-style rows are the
 * caller's (EigenModes[i+1] are small host matrices in the reference too). */
int ithaca_b200_pod_reconstruct(ithaca_b200_t* h, int nmodes, const double* coeff, int k, int clip_negative,
                                double* const* out_ptrs);

/* ---- snapshot-blocked ("memory efficient") POD: ITHACAPOD::getModesMemoryEfficient (ITHACAPOD.C:344-628) -------
 * For snapshot sets larger than HBM. The reference reads the snapshots pair by pair from disk and
 * fills the correlation matrix one inner product at a time (:414-451), then accumulates the modes snapshot by snapshot;
 * here the caller streams BLOCKS of snapshots (as many as fit: two blocks are resident at a time), G is filled one
 * block pair per contraction launch on the device, and the modes are accumulated block by block.
 *   block_begin(m_total, dim, nCells)            after pod_set_volumes; zeroes the device correlation matrix
 *   block_upload(slot, ptrs, nb, first)          slot 0 or 1 <- snapshots [first, first+nb) (host or device pointers)
 *   block_gram(slotI, slotJ, norm_kind)          G[I, J] (and G[J, I]) = S_I^T W S_J, W = diag(V) for norm_kind 0 ("L2",
 *                                                computeInnerProduct) or I for 1 ("Frobenius", computeFrobeniusInnerProduct:
 *                                                unlike getModes, this path does NOT weigh the Frobenius correlation)
 *   block_gram_finish(G_out)                     all-reduce over ranks, install as the Gram of pod_eig; G_out may be NULL
 *   pod_eig(...)                                 as for the resident path
 *   block_lift(slot, r)                          Phi (+)= S_slot Q[first : first+nb, 0:r]   (first call after pod_eig starts a new sum)
 *   block_lift_finish(r, norm_kind, mode_ptrs, normfact_out)   norms, scaling, download: as pod_lift */
int ithaca_b200_pod_block_begin(ithaca_b200_t* h, int m_total, int dim, int64_t nCells);
int ithaca_b200_pod_block_upload(ithaca_b200_t* h, int slot, const double* const* snap_ptrs, int nb, int first);
int ithaca_b200_pod_block_gram(ithaca_b200_t* h, int slotI, int slotJ, int norm_kind);
int ithaca_b200_pod_block_gram_finish(ithaca_b200_t* h, double* G_out);
/* Checkpoint / resume of the blocked correlation matrix (PODTemplate::buildCovMatrix saves it after every block pair and
 * restarts from the newest complete file, src/ITHACA_CORE/ITHACAPOD/PODTemplate.C:419-577, findTempFile :265-341):
 *   block_gram_get(G_out)   the matrix accumulated so far, summed over ranks, WITHOUT finishing (m x m col-major)
 *   block_gram_set(G)       after block_begin: continue from a saved matrix (every rank passes the same G) */
int ithaca_b200_pod_block_gram_get(ithaca_b200_t* h, double* G_out);
int ithaca_b200_pod_block_gram_set(ithaca_b200_t* h, const double* G);
int ithaca_b200_pod_block_lift(ithaca_b200_t* h, int slot, int r);
int ithaca_b200_pod_block_lift_finish(ithaca_b200_t* h, int r, int norm_kind, double* const* mode_ptrs, double* normfact_out);

/* ---- row-streamed POD: a rank's cell shard larger than HBM (C3 on 1 or 2 GPUs: 480 GB / 240 GB per rank) ------------
 * The reference's own parallel decomposition is over CELLS: every rank forms the Gram of its rows and the partial matrices are
 * summed (ITHACAcoeffsMass.C:85-115, ITHACAPOD.C:146-149); the lift and the norms are row-local (ITHACAPOD.C:194-243). The same
 * identity with more cell chunks than GPUs: the shard is cut into chunks of consecutive cells that pass through two device
 * slots; each snapshot row crosses PCIe twice (once per pass) instead of once per block pair as in the snapshot-blocked path
 * above, and the m-snapshot contraction keeps its full-size schedule. The r modes of the whole shard stay resident (r << m).
 *   stream_begin(m, dim, nCellsShard)
 *   pass 1, per chunk:  stream_upload(slot, cell0, n, V_chunk, ptrs)  or  stream_slot(...) + an in-place producer
 *                       stream_gram_add(slot)                 G_partial += S_c^T diag(V_c) S_c
 *   stream_gram_finish(G_out)                                 all-reduce over ranks; installs the Gram of pod_eig
 *   pod_eig(...)
 *   pass 2, per chunk:  fill a slot again (a chunk still resident from pass 1 needs no refill), stream_lift_add(slot, r, norm)
 *   stream_lift_finish(r, mode_ptrs, normfact_out)            norms all-reduced, modes scaled and downloaded: as pod_lift
 * stream_upload, gram_add and lift_add only ENQUEUE (copies on a copy stream, kernels on ithaca_b200_stream): with the two
 * slots the upload of chunk k+1 overlaps the contraction of chunk k. The *_finish calls synchronise and report errors.
 *   stream_upload: V_chunk = n volumes, snap_ptrs[j] -> the chunk's n*dim values of snapshot j (host or device; host memory
 *                  must stay valid until the next *_finish; pinned memory makes the copies asynchronous).
 *   stream_slot:   returns the slot's device addresses for a producer that writes the chunk in place (S_out: m columns of
 *                  *ld_out doubles, V_out: n doubles). Blocks the host until the slot's previous consumers are done. The
 *                  producer's writes must be complete, or issued on ithaca_b200_stream(h), before gram_add / lift_add. */
int ithaca_b200_pod_stream_begin(ithaca_b200_t* h, int m, int dim, int64_t nCellsShard);
int ithaca_b200_pod_stream_upload(ithaca_b200_t* h, int slot, int64_t cell0, int64_t nCellsChunk, const double* V_chunk,
                                  const double* const* snap_ptrs);
int ithaca_b200_pod_stream_slot(ithaca_b200_t* h, int slot, int64_t cell0, int64_t nCellsChunk, double** S_out, int64_t* ld_out,
                                double** V_out);
int ithaca_b200_pod_stream_gram_add(ithaca_b200_t* h, int slot);
int ithaca_b200_pod_stream_gram_finish(ithaca_b200_t* h, double* G_out);
int ithaca_b200_pod_stream_lift_add(ithaca_b200_t* h, int slot, int r, int norm_kind);
int ithaca_b200_pod_stream_lift_finish(ithaca_b200_t* h, int r, double* const* mode_ptrs, double* normfact_out);

/* ---- Galerkin projection: steadyNS::projectSUP term builders ---------------------------------------
 * (src/ITHACA_FOMPROBLEMS/steadyNS/steadyNS.C:556-707 calls :883-1137; inherited by unsteadyNS,
 *  src/ITHACA_FOMPROBLEMS/unsteadyNS/unsteadyNS.H:61). Scheme set: Gauss linear (grad, div, interpolate),
 *  Gauss linear corrected | uncorrected (laplacian); ithaca_b200_mesh_set_schemes adds the cellMDLimited gradient
 *  and the limited snGrad. */

/* Mesh addressing and geometry exactly as OpenFOAM holds them (32-bit labels):
 *   owner[nInt+nBnd], neighbour[nInt]      mesh.owner(), mesh.neighbour() (faces in OpenFOAM order: internal, then patches)
 *   Sf[3*(nInt+nBnd)], magSf[nInt+nBnd]    mesh.Sf(), mesh.magSf() (internal field then boundary patches in order)
 *   weights[nInt]                          mesh.weights() (linear interpolation)
 *   nonOrthDeltaCoeffs[nInt]               mesh.nonOrthDeltaCoeffs()
 *   bndDeltaCoeffs[nBnd]                   patch.deltaCoeffs() of every boundary face
 *   nonOrthCorrVec[3*nInt] or NULL         mesh.nonOrthCorrectionVectors(); NULL selects the "uncorrected" laplacian
 *   V[nCells]                              mesh.V()
 *   bndMask[nBnd] or NULL                  0 for faces of `empty` patches (they hold no values and never contribute)
 * The cell->face addressing is built on the device. Host or device pointers. */
int ithaca_b200_mesh_upload(ithaca_b200_t* h, int64_t nCells, int64_t nIntFaces, int64_t nBndFaces,
                            const int32_t* owner, const int32_t* neighbour, const double* Sf, const double* magSf,
                            const double* weights, const double* nonOrthDeltaCoeffs, const double* bndDeltaCoeffs,
                            const double* nonOrthCorrVec, const double* V, const unsigned char* bndMask);

/* Scheme selection beyond "Gauss linear" for the mesh just uploaded: the set that
 * tutorials/CFD/04unsteadyNS/system/fvSchemes ships ("gradSchemes default cellMDLimited Gauss linear 0.5",
 * "laplacianSchemes default Gauss linear limited 1.0", "snGradSchemes default limited 1").
 *   grad_limit_k  in (0,1]: every fvc::grad on the path -- grad(P_j) of K, D, G, BC3 and the divergent background term,
 *                 grad(U_j) inside the non-orthogonal correction of fvc::laplacian (B, BC1, BP) and inside curl(U_j) of
 *                 BC3 -- is the cellMDLimited gradient (OpenFOAM limitedGradSchemes/cellMDLimitedGrad): the Gauss linear
 *                 cell gradient clipped face by face, in OpenFOAM's face order, so that the value extrapolated to each
 *                 face centre stays within the min/max of the cell's neighbourhood (widened by 1/k - 1). 0 (or < SMALL,
 *                 as in OpenFOAM): plain Gauss linear. k > 1 is refused (OpenFOAM: FatalIOError).
 *   sngrad_psi    in [0,1]: the laplacian's snGrad is "limited psi" (limitedSnGrad): the non-orthogonal correction is scaled
 *                 per face by min(psi |snGrad_uncorrected| / ((1 - psi) |correction| + SMALL), 1). Negative: "corrected"
 *                 (or "uncorrected" when mesh_upload got nonOrthCorrVec = NULL). Needs nonOrthCorrVec.
 *   C[3*nCells], Cf[3*(nInt+nBnd)]   mesh.C() internal field and mesh.Cf() (internal then boundary), in the order of the
 *                 uploaded mesh (a shard's local order after mesh_upload_shard); needed when grad_limit_k > 0 and for
 *                 ithaca_b200_mesh_set_convection(LINEAR_UPWIND); stored whenever both are given.
 * mesh_upload / mesh_upload_shard reset both to off. Host or device pointers. */
int ithaca_b200_mesh_set_schemes(ithaca_b200_t* h, double grad_limit_k, double sngrad_psi, const double* C, const double* Cf);

/* Interpolation scheme of the convection operators fvc::div(phi, U_k) ("divSchemes  div(phi,U)  Gauss <scheme>"; the flux
 * phi itself is always linearInterpolate(U_j) & Sf, steadyNS.C:1019-1021). Boundary faces take the patch value in every scheme.
 *   IB200_CONV_LINEAR         "Gauss linear" (the shipped tutorials; default after mesh_upload)
 *   IB200_CONV_UPWIND         "Gauss upwind": the owner's value when phi >= 0, else the neighbour's (upwind::weights = pos0(phi))
 *   IB200_CONV_LINEAR_UPWIND  "Gauss linearUpwind grad(U)": upwind value + (Cf - C_upwind) . grad(U_k)[upwind]
 *                             (linearUpwind::correction; grad(U_k) by the gradient scheme of mesh_set_schemes, which must
 *                             have been given C and Cf)
 * upwind keeps the fused face-based contraction of C, G and Cf: the face value depends on the sign of phi_j only, so every
 * internal face becomes two rank-one records (max(phi_j,0) with the owner's U_k, min(phi_j,0) with the neighbour's).
 * linearUpwind forms C and G through materialised div(phi_j, U_k) fields and refuses Cf. BC2 and the background terms follow
 * the scheme in both. */
enum { IB200_CONV_LINEAR = 0, IB200_CONV_UPWIND = 1, IB200_CONV_LINEAR_UPWIND = 2 };
int ithaca_b200_mesh_set_convection(ithaca_b200_t* h, int scheme);

/* The device-built cell->face CSR, for the bit-exact addressing check: off[nCells+1]; ent[2*nInt+nBnd] = +(f+1) when
 * the cell owns face f, -(f+1) when it is its neighbour, ascending face index per cell (primitiveMesh::cells() order);
 * other[...] = the cell across the face, -(b+1) for boundary face b. digest = FNV-1a 64 of the three arrays.
 * Any output may be NULL. */
int ithaca_b200_mesh_addressing(ithaca_b200_t* h, int32_t* off, int32_t* ent, int32_t* other, uint64_t* digest);

/* Basis fields. which = 0: the n vector fields of L_U_SUPmodes (lift, U modes, supremizer modes in that order,
 * steadyNS.C:561-585), int_ptrs[j] -> 3*nCells doubles (&field.ref()[0][0]), bnd_ptrs[j] -> 3*nBnd doubles (boundary
 * patch values in face order). which = 1: the n scalar fields of Pmodes (nCells / nBnd doubles). */
int ithaca_b200_basis_upload(ithaca_b200_t* h, int which, const double* const* int_ptrs,
                             const double* const* bnd_ptrs, int n);

/* fluxMethod == "consistent" (steadyNS.C:1010-1015, :1056, :1722): the convecting flux of mode j is the STORED face-flux field
 * L_PHImodes[j] (the POD modes of phi, steadyNS.C:733-740) instead of linearInterpolate(L_U_SUPmodes[j]) & Sf.
 * phi_ptrs[j] -> nInt+nBnd doubles in the caller's face order (internal field, then the boundary patches in order), n = the
 * number of velocity basis fields; call after basis_upload(which = 0) (which resets to the interpolated flux, like
 * phi_ptrs == NULL). Affects every fvc::div(phi_j, U_k) on the path: C (fused and materialised), G, BC2, Cf and
 * fvc_eval(DIV_PHI); fvc::div(U_j) of the divergence term keeps the interpolated flux. Host or device pointers. */
int ithaca_b200_flux_upload(ithaca_b200_t* h, const double* const* phi_ptrs, int n);

enum { IB200_OP_M = 0, IB200_OP_B = 1, IB200_OP_K = 2, IB200_OP_P = 3, IB200_OP_C = 4, IB200_OP_C_MATERIALISED = 5,
       IB200_OP_D = 6, IB200_OP_G = 7, IB200_OP_BC3 = 8, IB200_OP_BC4 = 9, IB200_OP_BC1 = 10, IB200_OP_BC2 = 11,
       IB200_OP_BP = 12, IB200_OP_CF = 13 };
/* One projected operator, laid out as the reference's Eigen object (column-major):
 *   IB200_OP_M  mass_term              steadyNS.C:1085-1108   N x N        M(i,j) = int U_i . U_j
 *   IB200_OP_B  diffusive_term         steadyNS.C:883-907     N x N        B(i,j) = int U_i . laplacian(1, U_j)
 *   IB200_OP_K  pressure_gradient_term steadyNS.C:935-960     N x Np       K(i,j) = int U_i . grad(P_j)
 *   IB200_OP_P  divergence_term        steadyNS.C:1112-1137   Np x N       P(i,j) = int P_i div(U_j)
 *   IB200_OP_C  convective_term_tens   steadyNS.C:997-1036    N x N x N    C(i,j,k) = int U_i . div(interpolate(U_j) & Sf, U_k),
 *               (values identical to convective_term_tens_cache :1038-1083), stored with i fastest. Computed as one
 *               face-based contraction sum_f flux_j(f) * (Uf_k(f) . (U_i[own] - U_i[nei])) without materialising div fields.
 *   IB200_OP_C_MATERIALISED  the same tensor through N^2 materialised fvc::div(flux_j, U_k) fields (the reference's
 *               convective_term_tens_cache structure); kept as a cross-check and for the HBM-roofline measurement.
 * The operators steadyNS::projectPPE adds (steadyNS.C:354-553):
 *   IB200_OP_D    laplacian_pressure   steadyNS.C:1295-1317   Np x Np      D(i,j) = int grad(P_i) . grad(P_j)
 *   IB200_OP_G    divMomentum          steadyNS.C:1176-1254   Np x N x N   G(i,j,k) = int grad(P_i) . div(interpolate(U_j) & Sf, U_k),
 *                 i fastest (Eigen::Tensor), through materialised div fields (the reference's divMomentum_cache structure)
 *   IB200_OP_BC3  pressure_BC3         steadyNS.C:1439-1475   Np x N       sum over boundary faces of
 *                 (interpolate(curl U_j) & (n ^ interpolate(grad P_i))) |Sf|
 *   IB200_OP_BC4  pressure_BC4         steadyNS.C:1477-1513   Np x N       sum over boundary faces of P_i (n & U_j) |Sf|
 *   IB200_OP_BC1  pressure_BC1         steadyNS.C:1319-1354   Np x N       sum over boundary faces of
 *                 (interpolate(laplacian(U_j)) & Sf) P_i: the boundary value of an fvc::laplacian result is the owner cell's
 *                 value (extrapolatedCalculated patch of fvc::surfaceIntegrate)
 *   IB200_OP_BC2  pressureBC2          steadyNS.C:1396-1437   Np x N x N   BC2(i,j,k) = sum over boundary faces of
 *                 (interpolate(div(interpolate(U_j) & Sf, U_k)) & Sf) P_i, i fastest (Eigen::Tensor; slice i = the List entry
 *                 BC2_matrix[i] of pressure_BC2 :1356-1394); the boundary value of the fvc::div result is the owner cell's
 * The fvc-only operators of steadyNS::discretizeThenProject (steadyNS.C:709-881, "inconsistent" flux method):
 *   IB200_OP_BP   diffusive_term_flux_method  steadyNS.C:1584-1610  Np x N      BP(i,j) = int P_i div(laplacian(1, U_j))
 *   IB200_OP_CF   convective_term_flux_tens   steadyNS.C:1704-1746  Np x N x N  Cf(i,j,k) = int P_i div(div(flux(U_j), U_k)),
 *                 i fastest. The inner fvc:: result is assigned to a volVectorField before the outer div, so its boundary
 *                 value is the owner cell's (extrapolatedCalculated patch copied by GeometricField::operator=). Computed
 *                 through the adjoint of the outer div: int P_i div(A) = int t_i . A with one precomputed vector field t_i
 *                 per pressure mode, so BP is one contraction and Cf is the fused face-based contraction of G.
 * BC1, BC3 and BC4 are returned for ALL N velocity fields; the reference keeps the leading lift + NUmodes columns
 * (P_BC2size = NUmodes + liftfield.size()): take leftCols. Faces of masked (`empty`) patches do not contribute. */
int ithaca_b200_project(ithaca_b200_t* h, int op, double* out);

/* Background ("large scale") convection with an external advecting field vls (host or device: 3*nCells internal values
 * and 3*nBnd boundary values, the caller's order):
 *   kind 0  convective_background            steadyNS.C:1256-1273   N x N    L(i,j)   = - int U_i . div(interpolate(vls) & Sf, U_j)
 *   kind 1  divergent_convective_background  steadyNS.C:1275-1293   Np x N   L_D(i,j) = - int grad(P_i) . div(interpolate(vls) & Sf, U_j)
 * for ALL N velocity fields (the reference keeps the leading NUmodes + liftfield.size() rows / columns). */
int ithaca_b200_project_background(ithaca_b200_t* h, int kind, const double* vls_int, const double* vls_bnd, double* out);

/* Penalty-method boundary terms of one (patch, component) pair (steadyNS::bcVelocityVec / bcVelocityMat,
 * steadyNS.C:1515-1582; inletIndex(k,0) = patch, inletIndex(k,1) = component):
 *   vec_out[i]    = sum over the patch faces of U_i.boundaryField()[patch].component(comp)              (N)
 *   mat_out(i,j)  = sum over the patch faces of U_i...component(comp) * U_j...component(comp)           (N x N)
 * for ALL N velocity fields (the reference fills BCsize = NUmodes + NSUPmodes entries indexed L_U_SUPmodes[0 .. BCsize): as
 * coded that is the LEADING block, lift fields included; host/steadyNS_b200.H takes exactly that block).
 * The patch is the boundary-face range [patch_start, patch_start + patch_size) (patch.start() - nInternalFaces). */
int ithaca_b200_project_bc_velocity(ithaca_b200_t* h, int64_t patch_start, int64_t patch_size, int comp, double* vec_out,
                                    double* mat_out);

enum { IB200_FVC_FLUX = 0, IB200_FVC_DIV_PHI = 1, IB200_FVC_LAPLACIAN = 2, IB200_FVC_GRAD_P = 3, IB200_FVC_DIV_U = 4 };
/* One materialised fvc operator field (used by the parity tests of the gather kernels):
 *   FLUX      linearInterpolate(U_j) & Sf                 nInt+nBnd doubles
 *   DIV_PHI   fvc::div(flux_j, U_k)                       3*nCells
 *   LAPLACIAN fvc::laplacian(1, U_j)                      3*nCells
 *   GRAD_P    fvc::grad(P_j)                              3*nCells
 *   DIV_U     fvc::div(U_j)                               nCells */
int ithaca_b200_fvc_eval(ithaca_b200_t* h, int op, int j, int k, double* out);

/* ---- cell-sharded projection ---------------------------------------------------------
 * Rank g of G owns the contiguous cell range of ithaca_b200_cell_range (the rule POD uses) and works on a SHARD MESH:
 * its owned cells followed by `halo_layers` layers of face-neighbour cells (1 layer is enough for Gauss linear
 * div/grad/interpolate, 2 for the "corrected" laplacian, whose face gradient needs grad in the first layer).
 * The addressing follows OpenFOAM's decomposition files as ITHACAparallel reads them
 * (src/ITHACA_CORE/ITHACAparallel/ITHACAparallel.C:43-74, used at :133-165):
 *   cellProcAddressing[nLocalCells]  global label of each shard cell: owned cells first (ascending), then halo cells
 *                                    (ascending global label)
 *   faceProcAddressing[nLocalFaces]  +-(global face + 1); the sign is negative when the shard face is FLIPPED with
 *                                    respect to the global face (the shard cell is the global neighbour), exactly the
 *                                    processor-face convention of decomposePar. Shard faces in this order:
 *                                    [internal: both cells in the shard][boundary faces of shard cells][cut faces:
 *                                    faces between the outermost halo layer and cells outside the shard]; ascending
 *                                    global index inside each group. Cut faces are boundary faces of the shard mesh
 *                                    with mask 0 (they only touch outer-halo cells, whose results are never used).
 * Ownership, `method`:
 *   IB200_PART_LABEL_RANGE     owned = the contiguous LABEL range of ithaca_b200_cell_range (SURVEY 8(e)'s rule; what a
 *                              decomposePar "simple"-like split of an already ordered mesh gives);
 *   IB200_PART_LOCALITY_RANGE  owned = a contiguous range of the library's locality ordering (the Cuthill-McKee
 *                              renumbering every mesh_upload applies internally, ithaca_b200_mesh_renumbering): same
 *                              range rule, applied to ranks in that ordering, so that shards are spatially compact
 *                              whatever labels the mesher produced (the C4 benchmark mesh is randomly relabelled, and a
 *                              label range of it touches every other cell: its halo is the whole mesh).
 * Host code only (no GPU needed). sizes[5] = {nLocalCells, nOwned, nLocalInt, nLocalBnd (true boundary), nCut}. */
enum { IB200_PART_LABEL_RANGE = 0, IB200_PART_LOCALITY_RANGE = 1 };
typedef struct ithaca_b200_partition ithaca_b200_partition_t;
int ithaca_b200_partition_create(ithaca_b200_partition_t** p, int64_t nCells, int64_t nIntFaces, int64_t nBndFaces,
                                 const int32_t* owner, const int32_t* neighbour, int rank, int nranks, int halo_layers,
                                 int method);
/* The locality ordering itself: old2new[c] = position of cell c (breadth-first Cuthill-McKee from a pseudo-peripheral
 * cell, neighbours by increasing degree then label; one search per connected component in label order). */
int ithaca_b200_mesh_renumbering(int64_t nCells, int64_t nIntFaces, const int32_t* owner, const int32_t* neighbour,
                                 int32_t* old2new);
int ithaca_b200_partition_destroy(ithaca_b200_partition_t* p);
int ithaca_b200_partition_sizes(const ithaca_b200_partition_t* p, int64_t* sizes5);
int ithaca_b200_partition_addressing(const ithaca_b200_partition_t* p, int32_t* cellProcAddressing,
                                     int32_t* faceProcAddressing);
/* Shard copy of a global field. kind 0: cell field (dim doubles per cell) -> nLocalCells*dim; kind 1: boundary-face
 * field (dim per global boundary face) -> (nLocalBnd+nCut)*dim, zeros on cut faces. */
int ithaca_b200_partition_gather(const ithaca_b200_partition_t* p, int kind, int dim, const double* global_field,
                                 double* shard_field);
/* ithaca_b200_mesh_upload of this rank's shard, from the GLOBAL mesh arrays (host pointers, same meaning as in
 * ithaca_b200_mesh_upload). Marks the halo cells: they get integration weight 0 in every projected operator, and
 * ithaca_b200_project all-reduces the operator over the communicator of ithaca_b200_comm_init (one
 * ncclAllReduce per operator instead of the reference's one gSum per entry, fvc::domainIntegrate at steadyNS.C:896-897 et al.).
 * After this call ithaca_b200_basis_upload expects SHARD fields (ithaca_b200_partition_gather). */
int ithaca_b200_mesh_upload_shard(ithaca_b200_t* h, const ithaca_b200_partition_t* p, const int32_t* owner,
                                  const int32_t* neighbour, const double* Sf, const double* magSf, const double* weights,
                                  const double* nonOrthDeltaCoeffs, const double* bndDeltaCoeffs,
                                  const double* nonOrthCorrVec, const double* V, const unsigned char* bndMask);
/* Lower-level form of the above for callers that already hold a decomposed mesh with halo cells: the first nOwned
 * cells (caller's order) of the uploaded mesh are owned, the rest are halo. nOwned = nCells restores the default. */
int ithaca_b200_mesh_set_owned(ithaca_b200_t* h, int64_t nOwned);

/* Device times (ms) of the last ithaca_b200_project call: out[0]=face flux, [1]=div(phi,U), [2]=laplacian (+grad tensor),
 * [3]=grad(p), [4]=div(U), [5]=contraction, [6]=per-face records of the fused C. */
int ithaca_b200_proj_timings(ithaca_b200_t* h, double* out, int n);

/* ---- on-disk formats of the reduced operators (host code, no GPU needed) -------------
 * The reference caches every projected operator under ./ITHACAoutput/Matrices/ and reads it back instead of
 * recomputing (steadyNS.C:384-467); these write the ABI's output buffers in exactly those formats so that an
 * unmodified ITHACA-FV (and its python tooling) reads what this library computed. Errors: last_error(NULL).
 *   save/read_dense_matrix  ITHACAstream::SaveDenseMatrix / ReadDenseMatrix (ITHACAstream.H:722-762): rows, cols as
 *                           Eigen::Index (int64), then rows*cols column-major doubles. read: A may be NULL (sizes only).
 *   save/read_dense_tensor  SaveDenseTensor / ReadDenseTensor (ITHACAstream.H:765-809): Eigen::DSizes<Index,3>
 *                           (3 x int64), then the data with the FIRST index fastest.
 *   save_npy                cnpy::save of an Eigen matrix / rank-3 tensor (cnpy.C:524-568, header cnpy.H:454-488):
 *                           NPY 1.0, '<f8', C order; `data` is the Eigen object's storage (first index fastest),
 *                           ndim = 1, 2 or 3.
 *   export_matrix/tensor    ITHACAstream::exportMatrix / exportTensor (ITHACAstream.C:54-147, :263-358): the text
 *                           files <folder>/<name>_mat.py | .m | .txt ("python": setprecision(10) np.array literal,
 *                           "matlab", "eigen": precision 20, one file <name><i>_mat.txt per slice T(i,:,:) for tensors). */
enum { IB200_EXPORT_PYTHON = 0, IB200_EXPORT_MATLAB = 1, IB200_EXPORT_EIGEN = 2 };
int ithaca_b200_save_dense_matrix(const char* path, const double* A, int64_t rows, int64_t cols);
int ithaca_b200_read_dense_matrix(const char* path, int64_t* rows, int64_t* cols, double* A, int64_t capacity);
int ithaca_b200_save_dense_tensor(const char* path, const double* T, int64_t d0, int64_t d1, int64_t d2);
int ithaca_b200_read_dense_tensor(const char* path, int64_t* dims3, double* T, int64_t capacity);
int ithaca_b200_save_npy(const char* path, const double* data, int ndim, const int64_t* dims);
int ithaca_b200_export_matrix(const char* folder, const char* name, int type, const double* A, int64_t rows, int64_t cols);
int ithaca_b200_export_tensor(const char* folder, const char* name, int type, const double* T, int64_t d0, int64_t d1,
                              int64_t d2);

/* ---- allocation-free entry point ------------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid allocation). It is a variant of the norm-factor
 * computation of pod_lift / pod_block_lift_finish (n_c = sqrt(sum_k w_k Phi(k,c)^2), w_k = V[k / dim] for norm_kind 0 "L2",
 * 1 for norm_kind 1 "Frobenius") whose per-block partial sums live in a file-scope static __device__ array instead of a
 * cudaMalloc'ed buffer. The call performs NO cudaMalloc, cudaFree or allocating library routine.
 *   Phi        DEVICE pointer: nCols columns of ld doubles, nRows valid rows each (ld >= nRows)
 *   V          DEVICE pointer: nRows / dim cell volumes (may be NULL for norm_kind 1)
 *   norms_out  DEVICE pointer: nCols doubles
 *   stream     NULL: the handle's stream, and the call blocks until the result is there (like every other entry point);
 *              otherwise a cudaStream_t of the caller's on which the two kernels are only enqueued.
 * Limit: nCols <= IB200_NOALLOC_MAX_COLS (256); more columns are refused with IB200_ERR_ARG and a message. The scratch array
 * (592 x 256 doubles = 1.2 MB) is held by every GPU that loads the library until the process exits; all calls on a device share
 * it and it keeps the last call's contents (every slot read by a call is first written by that call, nothing needs clearing).
 * Calls that overlap on the GPU -- two streams, two host threads -- RACE on it: serialise them. */
#define IB200_NOALLOC_MAX_COLS 256
int ithaca_b200_col_norms_noalloc(ithaca_b200_t* h, const double* Phi, int64_t nRows, int64_t ld, int nCols, const double* V,
                                  int dim, int norm_kind, double* norms_out, void* stream);

/* ---- measurement helpers -------------------------------------------------------------------- */
/* Register-resident DMMA.8x8x4 (mma.sync.m8n8k4.f64) issue-rate microbenchmark on `device`:
 * the measured FP64 tensor peak used as the Gram kernel's roofline denominator. */
int ithaca_b200_dmma_peak(int device, double* tflops_out);

/* The synthetic snapshot set of the benchmark configurations, written in place on the device (kernels on
 * ithaca_b200_stream(h)) from a counter-based generator, so that any row range is reproduced identically by any rank in any
 * chunking and the 480 GB / 800 GB sets of C3 / C5 never have to exist anywhere:
 *   S(i, j) = sum_{k<R} u_k(x_i) coef(j, k) / (k + 1) + eps * w(i, j),  x_i = (i + 1/2) / NdGlobal,  u_k(x) = sin((k + 1) pi x),
 *   w = zero-mean unit-variance variate from Philox4x32-10(key = seed, counter = (global row i, snapshot j)).
 * S: device, rows [row0, row0 + nRows) of the m snapshots, column j at S + j*ld. coef: device, m x R col-major (R <= 256).
 *   V(c) = exp(0.5 w(c)) e^(-1/8) / nCellsGlobal  (sums to ~1 over the global mesh). */
int ithaca_b200_synth_snapshots(ithaca_b200_t* h, double* S, int64_t ld, int64_t row0, int64_t nRows, int64_t NdGlobal, int m,
                                const double* coef, int R, double eps, uint64_t seed);
int ithaca_b200_synth_volumes(ithaca_b200_t* h, double* V, int64_t cell0, int64_t nCells, int64_t nCellsGlobal, uint64_t seed);

#ifdef __cplusplus
}
#endif
#endif /* ITHACA_B200_H */
