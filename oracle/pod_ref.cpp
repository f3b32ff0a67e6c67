// TEST INFRASTRUCTURE ONLY -- never linked, imported or executed by the product path.
//
// pod_ref: the reference's OWN dense arithmetic for ITHACAPOD::getModes, compiled against the
// reference's OWN vendored Eigen (src/thirdparty/Eigen) and Spectra (src/thirdparty/spectra)
// headers where they lie under /root/reference (see oracle/Makefile; output goes to oracle/_ref/).
// Only the OpenFOAM PtrList<GeometricField> -> raw pointer marshalling is restated, because
// OpenFOAM itself is not installed in this image. Every expression below cites the reference
// line it follows (paths relative to /root/reference/src/ITHACA_CORE).
//
//   Gram        : ITHACAutilities/ITHACAcoeffsMass.C:85-105   (both branches, threshold on V.size())
//   V repetition: ITHACAutilities/ITHACAcoeffsMass.C:312-326, EigenFunctions/EigenFunctions.C:125-143
//   eigensolve  : ITHACAPOD/ITHACAPOD.C:157-180               (Spectra / Eigen)
//   neg. shift  : ITHACAPOD/ITHACAPOD.C:182-186
//   lift        : ITHACAPOD/ITHACAPOD.C:194
//   norm factors: ITHACAPOD/ITHACAPOD.C:196-243
//   BC lift     : ITHACAPOD/ITHACAPOD.C:245-264
//   eigenvalue normalisation / cumulative sum: ITHACAPOD/ITHACAPOD.C:281-287
//   getCoeffs   : ITHACAutilities/ITHACAcoeffsMass.C:348-388, :424-443
//   Modes::project / reconstruct: Containers/Modes.C:228-290, :294-305
//   getModesSVD : ITHACAPOD/ITHACAPOD.C:858-941               (JacobiSVD of diag(sqrt V) S)
//   getModes with per-snapshot Volumes: ITHACAPOD/ITHACAPOD.C:1392-1541
//
// Used by: tests/ (checker), tests/golden/make_golden.py (fixture generator),
// bench.py cpu_baseline / --impl reference (timed CPU baseline). Nothing else.

#include <Eigen/Eigen>
#include <Spectra/SymEigsSolver.h>
#include <Spectra/MatOp/DenseSymMatProd.h>
#include <cmath>
#include <cstring>

typedef long long i64;
typedef Eigen::Map<const Eigen::MatrixXd> CMap;

// getMassMatrixFV (ITHACAcoeffsMass.C:312-326) + repeatElements (EigenFunctions.C:125-143)
static Eigen::VectorXd mass_fv(const double* Vc, i64 Nc, int dim)
{
    Eigen::VectorXd out(Nc * dim);
    for (i64 i = 0; i < Nc; i++)
        for (int r = 0; r < dim; r++)
            out(i * dim + r) = Vc[i];
    return out;
}

// branch: -1 = the reference's own rule (V.size() > 1e6 -> loop), 0 = force GEMM, 1 = force loop
static Eigen::MatrixXd gram(const CMap& F, const Eigen::VectorXd& V, int branch)
{
    const int Msize = (int)F.cols();
    Eigen::MatrixXd M = Eigen::MatrixXd::Zero(Msize, Msize);
    bool loop = (branch < 0) ? (V.size() > 1e6) : (branch == 1);
    if (loop)
    {
        // ITHACAcoeffsMass.C:95-101
        for (i64 i = 0; i < V.size(); i++)
        {
            M += V(i) * F.transpose().topRows(Msize).col(i) * F.leftCols(Msize).row(i);
        }
    }
    else
    {
        // ITHACAcoeffsMass.C:104
        M = F.transpose().topRows(Msize) * V.asDiagonal() * F.leftCols(Msize);
    }
    return M;
}

extern "C" {

// G_out: m x m col-major.  S: Nd x m col-major (Nd = dim*Nc), AoS-interleaved components.
int podref_gram(const double* S, i64 Nc, int dim, int m, const double* Vc, int branch,
                double* G_out)
{
    CMap F(S, Nc * dim, m);
    Eigen::VectorXd V = mass_fv(Vc, Nc, dim);
    Eigen::MatrixXd M = gram(F, V, branch);
    std::memcpy(G_out, M.data(), sizeof(double) * m * m);
    return 0;
}

// Eigensolve + post-processing only, on a given Gram matrix (ITHACAPOD.C:157-186).
// solver: 0 = spectra, 1 = eigen. evals_out must hold m doubles; *n_evals = nmodes (spectra) or m (eigen).
// evecs_out: m x nmodes col-major. Returns 0 on success, 1 if the solver reports failure.
int podref_eig(const double* G, int m, int nmodes, int solver, double* evals_out, int* n_evals,
               double* evecs_out)
{
    Eigen::MatrixXd _corMatrix = CMap(G, m, m);
    Eigen::VectorXd eigenValueseig;
    Eigen::MatrixXd eigenVectoreig;
    if (solver == 0)
    {
        int ncv = m;
        Spectra::DenseSymMatProd<double> op(_corMatrix);
        Spectra::SymEigsSolver<Spectra::DenseSymMatProd<double>> es(op, nmodes, ncv);
        es.init();
        es.compute(Spectra::SortRule::LargestAlge);
        if (es.info() != Spectra::CompInfo::Successful) return 1;
        eigenVectoreig = es.eigenvectors().real();
        eigenValueseig = es.eigenvalues().real();
    }
    else
    {
        Eigen::SelfAdjointEigenSolver<Eigen::MatrixXd> esEg;
        esEg.compute(_corMatrix);
        if (esEg.info() != Eigen::Success) return 1;
        eigenVectoreig = esEg.eigenvectors().real().rowwise().reverse().leftCols(nmodes);
        eigenValueseig = esEg.eigenvalues().real().array().reverse();
    }
    if (eigenValueseig.array().minCoeff() < 0)
    {
        eigenValueseig = eigenValueseig.array() + 2 * std::abs(eigenValueseig.array().minCoeff());
    }
    *n_evals = (int)eigenValueseig.size();
    std::memcpy(evals_out, eigenValueseig.data(), sizeof(double) * eigenValueseig.size());
    std::memcpy(evecs_out, eigenVectoreig.data(), sizeof(double) * m * nmodes);
    return 0;
}

// The whole dense part of getModes (ITHACAPOD.C:132-287), serial (Pstream::parRun() == false).
//   norm: 0 = "L2", 1 = "Frobenius".  (Quirk: the Gram is V-weighted under both, ITHACAcoeffsMass.C:87-105.)
//   Sbc : all boundary patches stacked row-wise, nbc x m col-major (may be null when nbc == 0); the
//         reference processes each patch separately (ITHACAPOD.C:245-264) with per-row arithmetic
//         identical to the stacked form.
// Outputs (any may be null): G m*m; evals (post-shift, un-normalised) m; evals_norm, cum (n_evals
// entries); evecs m*nmodes; normfact nmodes; modes Nd*nmodes col-major; modes_bc nbc*nmodes.
int podref_getmodes(const double* S, i64 Nc, int dim, int m, const double* Vc, int nmodes,
                    int solver, int norm, int branch, const double* Sbc, i64 nbc,
                    double* G_out, double* evals_out, int* n_evals, double* evals_norm_out,
                    double* cum_out, double* evecs_out, double* normfact_out, double* modes_out,
                    double* modes_bc_out)
{
    const i64 Nd = Nc * dim;
    CMap SnapMatrix(S, Nd, m);
    Eigen::VectorXd V = mass_fv(Vc, Nc, dim);
    Eigen::MatrixXd _corMatrix = gram(SnapMatrix, V, branch);
    if (G_out) std::memcpy(G_out, _corMatrix.data(), sizeof(double) * m * m);

    Eigen::VectorXd ev(m);
    Eigen::MatrixXd eigenVectoreig(m, nmodes);
    int ne = 0;
    int rc = podref_eig(_corMatrix.data(), m, nmodes, solver, ev.data(), &ne, eigenVectoreig.data());
    if (rc) return rc;
    Eigen::VectorXd eigenValueseig = ev.head(ne);
    if (n_evals) *n_evals = ne;
    if (evals_out) std::memcpy(evals_out, eigenValueseig.data(), sizeof(double) * ne);
    if (evecs_out) std::memcpy(evecs_out, eigenVectoreig.data(), sizeof(double) * m * nmodes);

    // ITHACAPOD.C:194
    Eigen::MatrixXd modesEig = (SnapMatrix * eigenVectoreig);
    // ITHACAPOD.C:196-233
    Eigen::MatrixXd normFact(nmodes, 1);
    for (int i = 0; i < nmodes; i++)
    {
        if (norm == 0)
        {
            normFact(i, 0) = std::sqrt((modesEig.col(i).transpose() * V.asDiagonal() *
                                        modesEig.col(i))(0, 0));
        }
        else
        {
            normFact(i, 0) = std::sqrt((modesEig.col(i).transpose() * modesEig.col(i))(0, 0));
        }
    }
    // ITHACAPOD.C:245-264
    Eigen::MatrixXd modesEigBC;
    if (nbc > 0)
    {
        CMap SnapMatrixBC(Sbc, nbc, m);
        modesEigBC = SnapMatrixBC * eigenVectoreig;
    }
    for (int i = 0; i < nmodes; i++)
    {
        modesEig.col(i) = modesEig.col(i).array() / normFact(i, 0);
        if (nbc > 0) modesEigBC.col(i) = modesEigBC.col(i).array() / normFact(i, 0);
    }
    if (normfact_out) std::memcpy(normfact_out, normFact.data(), sizeof(double) * nmodes);
    if (modes_out) std::memcpy(modes_out, modesEig.data(), sizeof(double) * Nd * nmodes);
    if (modes_bc_out && nbc > 0)
        std::memcpy(modes_bc_out, modesEigBC.data(), sizeof(double) * nbc * nmodes);

    // ITHACAPOD.C:281-287
    eigenValueseig = eigenValueseig / eigenValueseig.sum();
    Eigen::VectorXd cumEigenValues(eigenValueseig);
    for (int j = 1; j < cumEigenValues.size(); ++j)
    {
        cumEigenValues(j) += cumEigenValues(j - 1);
    }
    if (evals_norm_out) std::memcpy(evals_norm_out, eigenValueseig.data(), sizeof(double) * ne);
    if (cum_out) std::memcpy(cum_out, cumEigenValues.data(), sizeof(double) * ne);
    return 0;
}

// Lift + normalise only, for a given eigenvector block (ITHACAPOD.C:194-264); used to time the
// CPU lift separately and to check the CUDA lift against the reference with identical Q.
int podref_lift(const double* S, i64 Nc, int dim, int m, const double* Vc, const double* Q, int nmodes,
                int norm, double* normfact_out, double* modes_out)
{
    const i64 Nd = Nc * dim;
    CMap SnapMatrix(S, Nd, m);
    CMap eigenVectoreig(Q, m, nmodes);
    Eigen::VectorXd V = mass_fv(Vc, Nc, dim);
    Eigen::MatrixXd modesEig = (SnapMatrix * eigenVectoreig);
    for (int i = 0; i < nmodes; i++)
    {
        double nf = (norm == 0)
            ? std::sqrt((modesEig.col(i).transpose() * V.asDiagonal() * modesEig.col(i))(0, 0))
            : std::sqrt((modesEig.col(i).transpose() * modesEig.col(i))(0, 0));
        modesEig.col(i) = modesEig.col(i).array() / nf;
        if (normfact_out) normfact_out[i] = nf;
    }
    if (modes_out) std::memcpy(modes_out, modesEig.data(), sizeof(double) * Nd * nmodes);
    return 0;
}

// getCoeffs over a snapshot list (ITHACAcoeffsMass.C:348-388 per snapshot, :424-443 over the list):
// b = F^T.topRows(Msize) V snap; a = getMassMatrix(modes, Nmodes).fullPivLu().solve(b). The mass matrix is the
// 1-list getMassMatrix (ITHACAcoeffsMass.C:67-118, GEMM branch). coeff_out: Msize x nsnap col-major.
int podref_getcoeffs(const double* S, i64 Nc, int dim, int nsnap, const double* Vc, const double* modes, int nmodes_avail,
                     int Nmodes, double* coeff_out)
{
    const i64 Nd = Nc * dim;
    const int Msize = (Nmodes == 0) ? nmodes_avail : Nmodes;
    if (nmodes_avail < Msize) return 1;  // "The Number of requested modes is larger then the available quantity."
    CMap F(modes, Nd, nmodes_avail);
    Eigen::VectorXd V = mass_fv(Vc, Nc, dim);
    Eigen::MatrixXd M_matrix = F.transpose().topRows(Msize) * V.asDiagonal() * F.leftCols(Msize);
    Eigen::Map<Eigen::MatrixXd> coeff(coeff_out, Msize, nsnap);
    for (int i = 0; i < nsnap; i++)
    {
        CMap snapEigen(S + (size_t)i * Nd, Nd, 1);
        Eigen::VectorXd b = F.transpose().topRows(Msize) * V.asDiagonal() * snapEigen;
        Eigen::VectorXd a = M_matrix.fullPivLu().solve(b);
        coeff.col(i) = a;
    }
    return 0;
}

// Modes::project over a field list (Containers/Modes.C:228-290): projType "G" (0) and "F" (1; this overload
// REPLACES vol by ones, :246-249 and :271-275, and does not solve with a mass matrix). proj_out: nmodes x nsnap.
int podref_project(const double* S, i64 Nc, int dim, int nsnap, const double* Vc, const double* modes, int nmodes_avail,
                   int numberOfModes, int projType, double* proj_out)
{
    const i64 Nd = Nc * dim;
    const int nm = (numberOfModes == 0) ? nmodes_avail : numberOfModes;
    if (nm > nmodes_avail) return 1;
    CMap fieldEig(S, Nd, nsnap);
    CMap EigenModes0(modes, Nd, nmodes_avail);
    Eigen::VectorXd vol = mass_fv(Vc, Nc, dim);
    if (projType == 1) vol = Eigen::VectorXd::Ones(vol.size());
    Eigen::MatrixXd projField = (EigenModes0.leftCols(nm)).transpose() * vol.asDiagonal() * fieldEig;
    std::memcpy(proj_out, projField.data(), sizeof(double) * nm * nsnap);
    return 0;
}

// Modes::reconstruct, internal field (Containers/Modes.C:294-305): InField = EigenModes[0].leftCols(Nmodes) * Coeff,
// negative entries clipped when the field is called "nut" (clip != 0). Coeff: Nmodes x k col-major.
int podref_reconstruct(const double* modes, i64 Nd, int nmodes_avail, const double* Coeff, int Nmodes, int k, int clip,
                       double* out)
{
    if (Nmodes > nmodes_avail) return 1;
    CMap EigenModes0(modes, Nd, nmodes_avail);
    CMap C(Coeff, Nmodes, k);
    for (int j = 0; j < k; j++)
    {
        Eigen::VectorXd InField = EigenModes0.leftCols(Nmodes) * C.col(j);
        if (clip) InField = (InField.array() < 0).select(0, InField);
        std::memcpy(out + (size_t)j * Nd, InField.data(), sizeof(double) * Nd);
    }
    return 0;
}

// getModesSVD (ITHACAPOD.C:872-905): thin JacobiSVD of diag(sqrt(V)) * S; modes = diag(1/sqrt(V)) * U.leftCols(nmodes);
// "eigenvalues" = ALL singular values / their sum, then the running sum.
// sv_norm_out, cum_out: min(Nd, m) doubles; modes_out: Nd x nmodes col-major.
int podref_getmodes_svd(const double* S, i64 Nc, int dim, int m, const double* Vc, int nmodes, double* sv_norm_out,
                        double* cum_out, double* modes_out)
{
    const i64 Nd = Nc * dim;
    Eigen::MatrixXd SnapMatrix = CMap(S, Nd, m);
    Eigen::VectorXd V = mass_fv(Vc, Nc, dim);
    Eigen::VectorXd V3dSqrt = V.array().sqrt();
    Eigen::VectorXd V3dInv = V3dSqrt.array().cwiseInverse();
    auto VMsqr = V3dSqrt.asDiagonal();
    auto VMsqrInv = V3dInv.asDiagonal();
    Eigen::MatrixXd SnapMatrix2 = VMsqr * SnapMatrix;
    Eigen::JacobiSVD<Eigen::MatrixXd> svd(SnapMatrix2, Eigen::ComputeThinU | Eigen::ComputeThinV);
    Eigen::VectorXd eigenValueseig;
    Eigen::MatrixXd eigenVectoreig;
    eigenValueseig = svd.singularValues().real();
    eigenVectoreig = svd.matrixU().real();
    Eigen::MatrixXd modesEig = VMsqrInv * eigenVectoreig;
    for (int i = 0; i < nmodes; i++)
        std::memcpy(modes_out + (size_t)i * Nd, modesEig.col(i).data(), sizeof(double) * Nd);
    eigenValueseig = eigenValueseig / eigenValueseig.sum();
    Eigen::VectorXd cumEigenValues(eigenValueseig);
    for (int j = 1; j < cumEigenValues.size(); ++j)
    {
        cumEigenValues(j) += cumEigenValues(j - 1);
    }
    std::memcpy(sv_norm_out, eigenValueseig.data(), sizeof(double) * eigenValueseig.size());
    std::memcpy(cum_out, cumEigenValues.data(), sizeof(double) * cumEigenValues.size());
    return 0;
}

// getModes with one volume field per snapshot (moving meshes; ITHACAPOD.C:1392-1541):
//   _corMatrix(i,j) = S_i^T diag(sqrt|V_i V_j|) S_j (:1425-1435), mirrored; eigensolve as getModes WITHOUT the negative
//   shift; modesEig = (S Q) diag(lambda^-1/2) (:1481-1484); eigenvalues / sum of the KEPT ones (eigen solver: head(nmodes),
//   :1476). Vs: Nd x m col-major (the reference multiplies V.col(i) and SnapMatrix.col(i) coefficient-wise, so the
//   volume fields must have as many rows as the snapshots: scalar fields in practice).
int podref_getmodes_volumes(const double* S, i64 Nd, int m, const double* Vs, int nmodes, int solver, double* G_out,
                            double* evals_norm_out, double* cum_out, double* evecs_out, double* modes_out)
{
    CMap SnapMatrix(S, Nd, m);
    CMap V(Vs, Nd, m);
    Eigen::MatrixXd _corMatrix(m, m);
    for (int i = 0; i < m; i++)
    {
        for (int j = 0; j <= i; j++)
        {
            Eigen::VectorXd Mij = (V.col(i).array() * V.col(j).array());
            Mij = Mij.array().abs().sqrt();
            _corMatrix(i, j) = SnapMatrix.col(i).transpose() * Mij.asDiagonal() * SnapMatrix.col(j);
        }
    }
    for (int i = 1; i < m; i++)
    {
        for (int j = 0; j < i; j++)
        {
            _corMatrix(j, i) = _corMatrix(i, j);
        }
    }
    Eigen::VectorXd eigenValueseig;
    Eigen::MatrixXd eigenVectoreig;
    int ncv = m;
    Spectra::DenseSymMatProd<double> op(_corMatrix);
    Eigen::SelfAdjointEigenSolver<Eigen::MatrixXd> esEg;
    if (solver == 0)
    {
        Spectra::SymEigsSolver<Spectra::DenseSymMatProd<double>> es(op, nmodes, ncv);
        es.init();
        es.compute(Spectra::SortRule::LargestAlge);
        if (es.info() != Spectra::CompInfo::Successful) return 1;
        eigenVectoreig = es.eigenvectors().real();
        eigenValueseig = es.eigenvalues().real();
    }
    else
    {
        esEg.compute(_corMatrix);
        if (esEg.info() != Eigen::Success) return 1;
        eigenVectoreig = esEg.eigenvectors().real().rowwise().reverse().leftCols(nmodes);
        eigenValueseig = esEg.eigenvalues().real().reverse().head(nmodes);
    }
    Eigen::VectorXd eigenValueseigLam = eigenValueseig.real().array().cwiseInverse().sqrt();
    Eigen::MatrixXd modesEig = (SnapMatrix * eigenVectoreig) * eigenValueseigLam.asDiagonal();
    eigenValueseig = eigenValueseig / eigenValueseig.sum();
    Eigen::VectorXd cumEigenValues(eigenValueseig);
    for (int j = 1; j < cumEigenValues.size(); ++j)
    {
        cumEigenValues(j) += cumEigenValues(j - 1);
    }
    if (G_out) std::memcpy(G_out, _corMatrix.data(), sizeof(double) * m * m);
    std::memcpy(evals_norm_out, eigenValueseig.data(), sizeof(double) * nmodes);
    std::memcpy(cum_out, cumEigenValues.data(), sizeof(double) * nmodes);
    if (evecs_out) std::memcpy(evecs_out, eigenVectoreig.data(), sizeof(double) * m * nmodes);
    std::memcpy(modes_out, modesEig.data(), sizeof(double) * Nd * nmodes);
    return 0;
}

const char* podref_version() { return "eigen+spectra vendored by ITHACA-FV @8e7d6c09"; }

} // extern "C"
