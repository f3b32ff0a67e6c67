/* TEST INFRASTRUCTURE ONLY -- never linked, imported or executed by the product path.
 *
 * fvc_oracle: plain-C restatement of (a) the OpenFOAM finite-volume-calculus operators that ITHACA-FV's
 * Galerkin projection calls and (b) the reference's own term-builder loop nests.
 *
 * Parity status: UNPINNED at the OpenFOAM boundary. The term builders (b) follow the reference line by
 * line (/root/reference/src/ITHACA_FOMPROBLEMS/steadyNS/steadyNS.C, cited at each function); the operators
 * (a) live in OpenFOAM, a third-party dependency absent from /root/reference (pinned by the reference to
 * v2412, README.md:19) and not installable here, so they restate OpenFOAM's published algorithms for the
 * scheme set "Gauss linear" (grad, div, interpolate) and "Gauss linear corrected|uncorrected" (laplacian):
 *   fvc::surfaceIntegrate  finiteVolume/fvc/fvcSurfaceIntegrate.C : += owner, -= neighbour over internal faces,
 *                          += faceCells over boundary patches, then /= V
 *   fvc::domainIntegrate   finiteVolume/fvc/fvcVolumeIntegrate.C  : sum_c V_c * vf_c
 *   linearInterpolate      interpolation/surfaceInterpolation/schemes/linear/linear.H : w*vf[own] + (1-w)*vf[nei];
 *                          boundary face = patch value
 *   fvc::div(flux, vf)     convectionSchemes/gaussConvectionScheme : surfaceIntegrate(flux * interpolate(vf))
 *   fvc::div(vf)           fvc/fvcDiv.C + divSchemes/gaussDivScheme : surfaceIntegrate(Sf & interpolate(vf))
 *   fvc::grad(vf)          gradSchemes/gaussGrad/gaussGrad.C : surfaceIntegrate-like sum of Sf * vf_f (boundary-value
 *                          correction of grad only alters boundary values, which domainIntegrate never reads)
 *   fvc::laplacian(1, vf)  laplacianSchemes/gaussLaplacianScheme : surfaceIntegrate(|Sf| * snGrad(vf)),
 *                          snGradSchemes/correctedSnGrad : nonOrthDeltaCoeffs*(vf[nei]-vf[own])
 *                          + nonOrthCorrectionVectors & linearInterpolate(grad(vf.component)); boundary faces:
 *                          fvPatchField::snGrad = deltaCoeffs*(patchValue - patchInternalValue)
 * Faces of `empty` patches hold no values in OpenFOAM and never contribute (bmask == 0).
 * Known answers that check (a): tests/test_fvc_oracle.py.
 *
 * Layouts: vector fields AoS (x0,y0,z0,x1,...) as OpenFOAM stores them; boundary values as one flat array over
 * boundary faces (face nInt+b <-> entry b); matrices column-major (Eigen); the C tensor column-major with the test
 * index i fastest (Eigen::Tensor<double,3>, steadyNS.C:1002-1003).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    long long nCells, nInt, nBnd;
    const int* owner;      /* nInt + nBnd */
    const int* neighbour;  /* nInt */
    const double* Sf;      /* 3*(nInt+nBnd) */
    const double* magSf;   /* nInt+nBnd */
    const double* w;       /* nInt: linear interpolation weights */
    const double* dc;      /* nInt: nonOrthDeltaCoeffs */
    const double* dcB;     /* nBnd: boundary deltaCoeffs */
    const double* corr;    /* 3*nInt nonOrthCorrectionVectors, or NULL for "uncorrected" */
    const double* V;       /* nCells */
    const unsigned char* bmask; /* nBnd: 0 = face of an `empty` patch */
    /* scheme selection beyond "Gauss linear" (tutorials/CFD/04unsteadyNS/system/fvSchemes): */
    const double* Cc;      /* 3*nCells cell centres mesh.C()  (only read when gradK > 0) */
    const double* Cf;      /* 3*(nInt+nBnd) face centres mesh.Cf() */
    double gradK;          /* > 0: gradSchemes "cellMDLimited Gauss linear <gradK>"; <= 0: "Gauss linear" */
    double snPsi;          /* >= 0: snGrad "limited <snPsi>" (needs corr); < 0: "corrected" / "uncorrected" */
    int convScheme;        /* divSchemes "Gauss <scheme>" of fvc::div(phi, U): 0 linear, 1 upwind, 2 linearUpwind grad(U) */
} fvo_mesh;

void fvo_grad_vec(const fvo_mesh* m, const double* U, const double* Ub, double* out);

/* cellMDLimited gradient limiter (OpenFOAM finiteVolume/gradSchemes/limitedGradSchemes/cellMDLimitedGrad:
 * cellMDLimitedGrads.C calcGrad + cellMDLimitedGrad.H limitFaceCmpt), applied in place to the Gauss linear cell
 * gradients g[c][i*dim + d] = d_i vf_d of a dim-component field vf (boundary values vb on non-empty faces):
 *   max/min of vf over the cell and its face neighbours (boundary faces: the patch value), minus the cell value,
 *   widened by (1/k - 1)(max - min) when k < 1; then, face by face IN OPENFOAM'S ORDER (internal faces ascending, each
 *   limiting its owner then its neighbour; then the boundary faces), every component's gradient is clipped along
 *   dcf = Cf - C so that the extrapolated increment dcf . g stays inside [min, max]. The vector version limits each
 *   component's gradient (a column of the tensor) with that component's own bounds. */
static void limit_face_cmpt(double* g, int stride, double maxDelta, double minDelta, const double* dcf)
{
    const double ex = dcf[0] * g[0] + dcf[1] * g[stride] + dcf[2] * g[2 * stride];
    double s;
    if (ex > maxDelta) s = maxDelta - ex;
    else if (ex < minDelta) s = minDelta - ex;
    else return;
    const double msq = dcf[0] * dcf[0] + dcf[1] * dcf[1] + dcf[2] * dcf[2];
    for (int i = 0; i < 3; i++) g[i * stride] = g[i * stride] + dcf[i] * s / msq;
}

static void fvo_limit_grad(const fvo_mesh* m, int dim, const double* vf, const double* vb, double* g)
{
    const double k = m->gradK;
    if (!(k > 0.0)) return;
    const long long nC = m->nCells;
    double* mx = (double*)malloc(sizeof(double) * dim * nC);
    double* mn = (double*)malloc(sizeof(double) * dim * nC);
    memcpy(mx, vf, sizeof(double) * dim * nC);
    memcpy(mn, vf, sizeof(double) * dim * nC);
    for (long long f = 0; f < m->nInt; f++) {
        const int o = m->owner[f], n = m->neighbour[f];
        for (int d = 0; d < dim; d++) {
            const double vo = vf[dim * o + d], vn = vf[dim * n + d];
            mx[dim * o + d] = fmax(mx[dim * o + d], vn); mn[dim * o + d] = fmin(mn[dim * o + d], vn);
            mx[dim * n + d] = fmax(mx[dim * n + d], vo); mn[dim * n + d] = fmin(mn[dim * n + d], vo);
        }
    }
    for (long long b = 0; b < m->nBnd; b++) {
        if (!m->bmask[b]) continue;
        const int o = m->owner[m->nInt + b];
        for (int d = 0; d < dim; d++) {
            mx[dim * o + d] = fmax(mx[dim * o + d], vb[dim * b + d]);
            mn[dim * o + d] = fmin(mn[dim * o + d], vb[dim * b + d]);
        }
    }
    for (long long i = 0; i < dim * nC; i++) { mx[i] -= vf[i]; mn[i] -= vf[i]; }
    if (k < 1.0)
        for (long long i = 0; i < dim * nC; i++) {
            const double dd = (1.0 / k - 1.0) * (mx[i] - mn[i]);
            mx[i] += dd;
            mn[i] -= dd;
        }
    for (long long f = 0; f < m->nInt + m->nBnd; f++) {
        if (f >= m->nInt && !m->bmask[f - m->nInt]) continue;
        const int o = m->owner[f];
        double dcf[3];
        for (int i = 0; i < 3; i++) dcf[i] = m->Cf[3 * f + i] - m->Cc[3 * o + i];
        for (int d = 0; d < dim; d++) limit_face_cmpt(g + 3 * dim * (long long)o + d, dim, mx[dim * o + d], mn[dim * o + d], dcf);
        if (f < m->nInt) {
            const int n = m->neighbour[f];
            for (int i = 0; i < 3; i++) dcf[i] = m->Cf[3 * f + i] - m->Cc[3 * n + i];
            for (int d = 0; d < dim; d++) limit_face_cmpt(g + 3 * dim * (long long)n + d, dim, mx[dim * n + d], mn[dim * n + d], dcf);
        }
    }
    free(mx);
    free(mn);
}

/* ---- operators ------------------------------------------------------------------------------------- */

/* phi_f = linearInterpolate(U) & Sf   (steadyNS.C:1019-1021). phi: nInt+nBnd. */
void fvo_flux(const fvo_mesh* m, const double* U, const double* Ub, double* phi)
{
    for (long long f = 0; f < m->nInt; f++) {
        const int o = m->owner[f], n = m->neighbour[f];
        const double w = m->w[f];
        double s = 0.0;
        for (int c = 0; c < 3; c++) s += (w * U[3 * o + c] + (1.0 - w) * U[3 * n + c]) * m->Sf[3 * f + c];
        phi[f] = s;
    }
    for (long long b = 0; b < m->nBnd; b++) {
        const long long f = m->nInt + b;
        double s = 0.0;
        if (m->bmask[b])
            for (int c = 0; c < 3; c++) s += Ub[3 * b + c] * m->Sf[3 * f + c];
        phi[f] = s;
    }
}

/* out = fvc::div(phi, U) with Gauss linear: surfaceIntegrate(phi * linearInterpolate(U)). out: 3*nCells. */
void fvo_div_phi_vec(const fvo_mesh* m, const double* phi, const double* U, const double* Ub, double* out)
{
    /* face value of U by the convection scheme (OpenFOAM interpolation/surfaceInterpolation):
     *   linear        schemes/linear/linear.H            w U[own] + (1 - w) U[nei]
     *   upwind        limitedSchemes/upwind/upwind.H     weights = pos0(faceFlux): U[own] when phi >= 0, else U[nei]
     *   linearUpwind  schemes/linearUpwind/linearUpwind.C  upwind + correction, correction = (Cf - C[up]) & grad(U)[up]
     *                 (faceFlux > 0 ? owner : neighbour; at phi == 0 the product phi * U_f vanishes anyway), grad(U) by the
     *                 gradient scheme the dictionary entry names ("linearUpwind grad(U)") */
    double* T = NULL;
    if (m->convScheme == 2) {
        T = (double*)malloc(sizeof(double) * 9 * m->nCells);
        fvo_grad_vec(m, U, Ub, T);
    }
    memset(out, 0, sizeof(double) * 3 * m->nCells);
    for (long long f = 0; f < m->nInt; f++) {
        const int o = m->owner[f], n = m->neighbour[f];
        const double w = m->w[f];
        const int up = phi[f] >= 0.0 ? o : n;
        for (int c = 0; c < 3; c++) {
            double uf;
            if (m->convScheme == 0) uf = w * U[3 * o + c] + (1.0 - w) * U[3 * n + c];
            else {
                uf = U[3 * up + c];
                if (T)
                    for (int i = 0; i < 3; i++) uf += (m->Cf[3 * f + i] - m->Cc[3 * up + i]) * T[9 * up + 3 * i + c];
            }
            const double v = phi[f] * uf;
            out[3 * o + c] += v;
            out[3 * n + c] -= v;
        }
    }
    free(T);
    for (long long b = 0; b < m->nBnd; b++) {
        if (!m->bmask[b]) continue;
        const int o = m->owner[m->nInt + b];
        for (int c = 0; c < 3; c++) out[3 * o + c] += phi[m->nInt + b] * Ub[3 * b + c];
    }
    for (long long i = 0; i < m->nCells; i++)
        for (int c = 0; c < 3; c++) out[3 * i + c] /= m->V[i];
}

/* out = fvc::div(U): surfaceIntegrate(Sf & linearInterpolate(U)). out: nCells. */
void fvo_div_vec(const fvo_mesh* m, const double* U, const double* Ub, double* out)
{
    double* phi = (double*)malloc(sizeof(double) * (m->nInt + m->nBnd));
    fvo_flux(m, U, Ub, phi);
    memset(out, 0, sizeof(double) * m->nCells);
    for (long long f = 0; f < m->nInt; f++) {
        out[m->owner[f]] += phi[f];
        out[m->neighbour[f]] -= phi[f];
    }
    for (long long b = 0; b < m->nBnd; b++)
        if (m->bmask[b]) out[m->owner[m->nInt + b]] += phi[m->nInt + b];
    for (long long i = 0; i < m->nCells; i++) out[i] /= m->V[i];
    free(phi);
}

/* out = fvc::grad(p), Gauss linear. out: 3*nCells. */
void fvo_grad_scalar(const fvo_mesh* m, const double* p, const double* pb, double* out)
{
    memset(out, 0, sizeof(double) * 3 * m->nCells);
    for (long long f = 0; f < m->nInt; f++) {
        const int o = m->owner[f], n = m->neighbour[f];
        const double pf = m->w[f] * p[o] + (1.0 - m->w[f]) * p[n];
        for (int c = 0; c < 3; c++) {
            const double v = m->Sf[3 * f + c] * pf;
            out[3 * o + c] += v;
            out[3 * n + c] -= v;
        }
    }
    for (long long b = 0; b < m->nBnd; b++) {
        if (!m->bmask[b]) continue;
        const int o = m->owner[m->nInt + b];
        for (int c = 0; c < 3; c++) out[3 * o + c] += m->Sf[3 * (m->nInt + b) + c] * pb[b];
    }
    for (long long i = 0; i < m->nCells; i++)
        for (int c = 0; c < 3; c++) out[3 * i + c] /= m->V[i];
    fvo_limit_grad(m, 1, p, pb, out);
}

/* T = fvc::grad(U), Gauss linear: T[cell][i][j] = sum_f Sf_i U_f,j / V (outer product Sf * U_f). out: 9*nCells. */
void fvo_grad_vec(const fvo_mesh* m, const double* U, const double* Ub, double* out)
{
    memset(out, 0, sizeof(double) * 9 * m->nCells);
    for (long long f = 0; f < m->nInt; f++) {
        const int o = m->owner[f], n = m->neighbour[f];
        const double w = m->w[f];
        for (int j = 0; j < 3; j++) {
            const double uf = w * U[3 * o + j] + (1.0 - w) * U[3 * n + j];
            for (int i = 0; i < 3; i++) {
                const double v = m->Sf[3 * f + i] * uf;
                out[9 * o + 3 * i + j] += v;
                out[9 * n + 3 * i + j] -= v;
            }
        }
    }
    for (long long b = 0; b < m->nBnd; b++) {
        if (!m->bmask[b]) continue;
        const int o = m->owner[m->nInt + b];
        for (int j = 0; j < 3; j++)
            for (int i = 0; i < 3; i++) out[9 * o + 3 * i + j] += m->Sf[3 * (m->nInt + b) + i] * Ub[3 * b + j];
    }
    for (long long c = 0; c < m->nCells; c++)
        for (int k = 0; k < 9; k++) out[9 * c + k] /= m->V[c];
    fvo_limit_grad(m, 3, U, Ub, out);
}

/* out = fvc::laplacian(dimensionedScalar(1), U), Gauss linear corrected (m->corr != NULL) or uncorrected.
 * gamma = 1 so gammaMagSf = |Sf| (gaussLaplacianScheme::fvcLaplacian). out: 3*nCells. */
void fvo_laplacian_vec(const fvo_mesh* m, const double* U, const double* Ub, double* out)
{
    double* T = NULL;
    if (m->corr) {
        T = (double*)malloc(sizeof(double) * 9 * m->nCells);
        fvo_grad_vec(m, U, Ub, T);
    }
    memset(out, 0, sizeof(double) * 3 * m->nCells);
    for (long long f = 0; f < m->nInt; f++) {
        const int o = m->owner[f], n = m->neighbour[f];
        const double w = m->w[f];
        double snu[3], cr[3] = {0.0, 0.0, 0.0}, lim = 1.0;
        for (int j = 0; j < 3; j++) {
            snu[j] = m->dc[f] * (U[3 * n + j] - U[3 * o + j]);
            if (T) /* nonOrthCorrectionVectors & linearInterpolate(grad(U)) (correctedSnGrad<vector>::correction) */
                for (int i = 0; i < 3; i++)
                    cr[j] += m->corr[3 * f + i] * (w * T[9 * o + 3 * i + j] + (1.0 - w) * T[9 * n + 3 * i + j]);
        }
        if (T && m->snPsi >= 0.0) {
            /* snGradSchemes/limitedSnGrad/limitedSnGrad.C correction(): limiter * corr,
             * limiter = min(psi |snGrad_uncorrected| / ((1 - psi) |corr| + SMALL), 1), |.| the vector magnitude */
            const double a = sqrt(snu[0] * snu[0] + snu[1] * snu[1] + snu[2] * snu[2]);
            const double c = sqrt(cr[0] * cr[0] + cr[1] * cr[1] + cr[2] * cr[2]);
            lim = fmin(m->snPsi * a / ((1.0 - m->snPsi) * c + 1e-15), 1.0);
        }
        for (int j = 0; j < 3; j++) {
            double sn = snu[j];
            if (T) sn += (m->snPsi >= 0.0) ? lim * cr[j] : cr[j];
            const double v = m->magSf[f] * sn;
            out[3 * o + j] += v;
            out[3 * n + j] -= v;
        }
    }
    for (long long b = 0; b < m->nBnd; b++) {
        if (!m->bmask[b]) continue;
        const long long f = m->nInt + b;
        const int o = m->owner[f];
        for (int j = 0; j < 3; j++) out[3 * o + j] += m->magSf[f] * (m->dcB[b] * (Ub[3 * b + j] - U[3 * o + j]));
    }
    for (long long c = 0; c < m->nCells; c++)
        for (int j = 0; j < 3; j++) out[3 * c + j] /= m->V[c];
    free(T);
}

/* fvc::domainIntegrate(a & b).value() for dim-component fields: sum_c V_c * (a_c . b_c) */
double fvo_domain_integrate_dot(const fvo_mesh* m, const double* a, const double* b, int dim)
{
    double s = 0.0;
    for (long long c = 0; c < m->nCells; c++) {
        double d = 0.0;
        for (int k = 0; k < dim; k++) d += a[dim * c + k] * b[dim * c + k];
        s += m->V[c] * d;
    }
    return s;
}

/* ---- term builders: the reference's loop nests ----------------------------------------------------------
 * U[i], Ub[i]: the N fields of L_U_SUPmodes (lift, U modes, supremizer modes; steadyNS.C:561-585), internal and
 * boundary values; P[j], Pb[j]: the Np pressure modes. */

/* steadyNS::mass_term, steadyNS.C:1085-1108.  M: N x N col-major. */
void fvo_mass_term(const fvo_mesh* m, int N, const double* const* U, double* M)
{
    for (int i = 0; i < N; i++)
        for (int j = 0; j < N; j++) M[i + (size_t)j * N] = fvo_domain_integrate_dot(m, U[i], U[j], 3);
}

/* steadyNS::diffusive_term, steadyNS.C:883-907: B(i,j) = domainIntegrate(U_i & laplacian(1, U_j)).
 * The laplacian of U_j is evaluated once per j instead of once per (i,j): same values, same order inside. */
void fvo_diffusive_term_range(const fvo_mesh* m, int N, const double* const* U, const double* const* Ub, double* B, int j0, int j1)
{
    double* L = (double*)malloc(sizeof(double) * 3 * m->nCells);
    for (int j = j0; j < j1; j++) {
        fvo_laplacian_vec(m, U[j], Ub[j], L);
        for (int i = 0; i < N; i++) B[i + (size_t)j * N] = fvo_domain_integrate_dot(m, U[i], L, 3);
    }
    free(L);
}

/* steadyNS::pressure_gradient_term, steadyNS.C:935-960: K(i,j) = domainIntegrate(U_i & grad(P_j)). K: N x Np. */
void fvo_pressure_gradient_term_range(const fvo_mesh* m, int N, int Np, const double* const* U, const double* const* P,
                                      const double* const* Pb, double* K, int j0, int j1)
{
    double* G = (double*)malloc(sizeof(double) * 3 * m->nCells);
    for (int j = j0; j < j1; j++) {
        fvo_grad_scalar(m, P[j], Pb[j], G);
        for (int i = 0; i < N; i++) K[i + (size_t)j * N] = fvo_domain_integrate_dot(m, U[i], G, 3);
    }
    free(G);
}

/* steadyNS::divergence_term, steadyNS.C:1112-1137: P(i,j) = domainIntegrate(P_i * div(U_j)). Pm: Np x N. */
void fvo_divergence_term_range(const fvo_mesh* m, int N, int Np, const double* const* U, const double* const* Ub,
                               const double* const* P, double* Pm, int j0, int j1)
{
    double* D = (double*)malloc(sizeof(double) * m->nCells);
    for (int j = j0; j < j1; j++) {
        fvo_div_vec(m, U[j], Ub[j], D);
        for (int i = 0; i < Np; i++) Pm[i + (size_t)j * Np] = fvo_domain_integrate_dot(m, P[i], D, 1);
    }
    free(D);
}

/* steadyNS::convective_term_tens, steadyNS.C:997-1036 (fluxMethod != "consistent"): the reference's own loop
 * nest, flux and div re-evaluated for every (i,j,k).  C(i,j,k) at C[i + N*(j + N*k)]. */
void fvo_convective_term_tens_range(const fvo_mesh* m, int N, const double* const* U, const double* const* Ub, double* C, int i0, int i1)
{
    double* phi = (double*)malloc(sizeof(double) * (m->nInt + m->nBnd));
    double* D = (double*)malloc(sizeof(double) * 3 * m->nCells);
    for (int i = i0; i < i1; i++)
        for (int j = 0; j < N; j++)
            for (int k = 0; k < N; k++) {
                fvo_flux(m, U[j], Ub[j], phi);
                fvo_div_phi_vec(m, phi, U[k], Ub[k], D);
                C[i + (size_t)N * (j + (size_t)N * k)] = fvo_domain_integrate_dot(m, U[i], D, 3);
            }
    free(phi);
    free(D);
}

/* steadyNS::convective_term_tens_cache, steadyNS.C:1038-1083: flux hoisted per j, div per (j,k). */
void fvo_convective_term_tens_cache_range(const fvo_mesh* m, int N, const double* const* U, const double* const* Ub, double* C, int j0, int j1)
{
    double* phi = (double*)malloc(sizeof(double) * (m->nInt + m->nBnd));
    double* D = (double*)malloc(sizeof(double) * 3 * m->nCells);
    for (int j = j0; j < j1; j++) {
        fvo_flux(m, U[j], Ub[j], phi);
        for (int k = 0; k < N; k++) {
            fvo_div_phi_vec(m, phi, U[k], Ub[k], D);
            for (int i = 0; i < N; i++) C[i + (size_t)N * (j + (size_t)N * k)] = fvo_domain_integrate_dot(m, U[i], D, 3);
        }
    }
    free(phi);
    free(D);
}

/* The whole operator = the slab [0, N) of its outer loop; the *_range forms let the CPU-baseline harness (bench_proj.py) give
 * every host thread one slab (each entry is computed by the same serial arithmetic either way). */
void fvo_diffusive_term(const fvo_mesh* m, int N, const double* const* U, const double* const* Ub, double* B)
{
    fvo_diffusive_term_range(m, N, U, Ub, B, 0, N);
}
void fvo_pressure_gradient_term(const fvo_mesh* m, int N, int Np, const double* const* U, const double* const* P,
                                const double* const* Pb, double* K)
{
    fvo_pressure_gradient_term_range(m, N, Np, U, P, Pb, K, 0, Np);
}
void fvo_divergence_term(const fvo_mesh* m, int N, int Np, const double* const* U, const double* const* Ub,
                         const double* const* P, double* Pm)
{
    fvo_divergence_term_range(m, N, Np, U, Ub, P, Pm, 0, N);
}
void fvo_convective_term_tens(const fvo_mesh* m, int N, const double* const* U, const double* const* Ub, double* C)
{
    fvo_convective_term_tens_range(m, N, U, Ub, C, 0, N);
}
void fvo_convective_term_tens_cache(const fvo_mesh* m, int N, const double* const* U, const double* const* Ub, double* C)
{
    fvo_convective_term_tens_cache_range(m, N, U, Ub, C, 0, N);
}

/* ---- PPE operators (steadyNS::projectPPE, steadyNS.C:354-553) ------------------------------------------------
 * Boundary values of fvc results, restated from OpenFOAM v2412 (recalled, unpinned like the operators above):
 *   fvc::grad(vf) on a non-coupled patch face b with cell P and unit normal n (gaussGrad::correctBoundaryConditions):
 *       grad_b = grad_P + n * (snGrad_b - (n & grad_P)),   snGrad_b = deltaCoeffs_b * (vf_b - vf_P)
 *       (fvPatchField::snGrad; for a zeroGradient patch vf_b == vf_P and it vanishes)
 *   fvc::curl(vf) = 2 * (*skew(fvc::grad(vf)))  (fvcCurl.C), boundary field = the same expression of grad_b;
 *       with grad_id = d_i vf_d: curl = (g_yz - g_zy, g_zx - g_xz, g_xy - g_yx)
 *   fvc::interpolate(vol field) on a boundary face = the field's boundary value. */

/* steadyNS::laplacian_pressure, steadyNS.C:1295-1317: D(i,j) = domainIntegrate(grad(P_i) & grad(P_j)). D: Np x Np. */
void fvo_laplacian_pressure(const fvo_mesh* m, int Np, const double* const* P, const double* const* Pb, double* D)
{
    double** G = (double**)malloc(sizeof(double*) * Np);
    for (int i = 0; i < Np; i++) {
        G[i] = (double*)malloc(sizeof(double) * 3 * m->nCells);
        fvo_grad_scalar(m, P[i], Pb[i], G[i]);
    }
    for (int i = 0; i < Np; i++)
        for (int j = 0; j < Np; j++) D[i + (size_t)j * Np] = fvo_domain_integrate_dot(m, G[i], G[j], 3);
    for (int i = 0; i < Np; i++) free(G[i]);
    free(G);
}

/* steadyNS::divMomentum_cache, steadyNS.C:1207-1254 (values identical to divMomentum :1176-1205):
 * G(i,j,k) = domainIntegrate(grad(P_i) & div(interpolate(U_j) & Sf, U_k)) at G[i + Np*(j + N*k)]. */
void fvo_div_momentum(const fvo_mesh* m, int N, int Np, const double* const* U, const double* const* Ub,
                      const double* const* P, const double* const* Pb, double* G)
{
    double** gp = (double**)malloc(sizeof(double*) * Np);
    for (int i = 0; i < Np; i++) {
        gp[i] = (double*)malloc(sizeof(double) * 3 * m->nCells);
        fvo_grad_scalar(m, P[i], Pb[i], gp[i]);
    }
    double* phi = (double*)malloc(sizeof(double) * (m->nInt + m->nBnd));
    double* D = (double*)malloc(sizeof(double) * 3 * m->nCells);
    for (int j = 0; j < N; j++) {
        fvo_flux(m, U[j], Ub[j], phi);
        for (int k = 0; k < N; k++) {
            fvo_div_phi_vec(m, phi, U[k], Ub[k], D);
            for (int i = 0; i < Np; i++) G[i + (size_t)Np * (j + (size_t)N * k)] = fvo_domain_integrate_dot(m, gp[i], D, 3);
        }
    }
    for (int i = 0; i < Np; i++) free(gp[i]);
    free(gp);
    free(phi);
    free(D);
}

/* steadyNS::pressure_BC4, steadyNS.C:1477-1513: BC4(i,j) = sum over boundary faces of
 * interpolate(P_i) * (n & interpolate(U_j)) * magSf, n = Sf/magSf. BC4: Np x N. */
void fvo_pressure_bc4(const fvo_mesh* m, int N, int Np, const double* const* Ub, const double* const* Pb, double* BC4)
{
    for (int i = 0; i < Np; i++)
        for (int j = 0; j < N; j++) {
            double s = 0.0;
            for (long long b = 0; b < m->nBnd; b++) {
                if (!m->bmask[b]) continue;
                const long long f = m->nInt + b;
                const double ms = m->magSf[f];
                const double n[3] = {m->Sf[3 * f] / ms, m->Sf[3 * f + 1] / ms, m->Sf[3 * f + 2] / ms};
                const double* u = Ub[j] + 3 * b;
                const double bc4 = n[0] * u[0] + n[1] * u[1] + n[2] * u[2];
                s += (Pb[i][b] * bc4) * ms;
            }
            BC4[i + (size_t)j * Np] = s;
        }
}

/* boundary value of fvc::grad of a field with `dim` components at boundary face b (see the note above).
 * gP: the cell gradient of the owner, g[i*dim + d] = d_i vf_d; vb, vP: boundary and owner-cell values. */
static void bnd_grad(const fvo_mesh* m, long long b, int dim, const double* gP, const double* vb, const double* vP, double* g)
{
    const long long f = m->nInt + b;
    const double ms = m->magSf[f];
    const double n[3] = {m->Sf[3 * f] / ms, m->Sf[3 * f + 1] / ms, m->Sf[3 * f + 2] / ms};
    for (int d = 0; d < dim; d++) {
        const double sn = m->dcB[b] * (vb[d] - vP[d]);
        const double ng = n[0] * gP[0 * dim + d] + n[1] * gP[1 * dim + d] + n[2] * gP[2 * dim + d];
        for (int i = 0; i < 3; i++) g[i * dim + d] = gP[i * dim + d] + n[i] * (sn - ng);
    }
}

/* steadyNS::pressure_BC3, steadyNS.C:1439-1475: BC3(i,j) = sum over boundary faces of
 * (interpolate(curl(U_j)) & (n ^ interpolate(grad(P_i)))) * magSf. BC3: Np x N. */
void fvo_pressure_bc3(const fvo_mesh* m, int N, int Np, const double* const* U, const double* const* Ub,
                      const double* const* P, const double* const* Pb, double* BC3)
{
    const long long nB = m->nBnd;
    double* T = (double*)malloc(sizeof(double) * 9 * m->nCells);
    double* gp = (double*)malloc(sizeof(double) * 3 * m->nCells);
    double* curl = (double*)malloc(sizeof(double) * 3 * nB * (size_t)N);
    double* ngp = (double*)malloc(sizeof(double) * 3 * nB * (size_t)Np);
    for (int j = 0; j < N; j++) {
        fvo_grad_vec(m, U[j], Ub[j], T);
        for (long long b = 0; b < nB; b++) {
            const long long c = m->owner[m->nInt + b];
            double g[9];
            bnd_grad(m, b, 3, T + 9 * c, Ub[j] + 3 * b, U[j] + 3 * c, g);
            double* cu = curl + 3 * (b + nB * (size_t)j);
            cu[0] = g[1 * 3 + 2] - g[2 * 3 + 1];
            cu[1] = g[2 * 3 + 0] - g[0 * 3 + 2];
            cu[2] = g[0 * 3 + 1] - g[1 * 3 + 0];
        }
    }
    for (int i = 0; i < Np; i++) {
        fvo_grad_scalar(m, P[i], Pb[i], gp);
        for (long long b = 0; b < nB; b++) {
            const long long f = m->nInt + b, c = m->owner[f];
            const double ms = m->magSf[f];
            const double n[3] = {m->Sf[3 * f] / ms, m->Sf[3 * f + 1] / ms, m->Sf[3 * f + 2] / ms};
            double g[3];
            bnd_grad(m, b, 1, gp + 3 * c, Pb[i] + b, P[i] + c, g);
            double* o = ngp + 3 * (b + nB * (size_t)i);
            o[0] = n[1] * g[2] - n[2] * g[1];
            o[1] = n[2] * g[0] - n[0] * g[2];
            o[2] = n[0] * g[1] - n[1] * g[0];
        }
    }
    for (int i = 0; i < Np; i++)
        for (int j = 0; j < N; j++) {
            double s = 0.0;
            for (long long b = 0; b < nB; b++) {
                if (!m->bmask[b]) continue;
                const double* a = curl + 3 * (b + nB * (size_t)j);
                const double* q = ngp + 3 * (b + nB * (size_t)i);
                s += (a[0] * q[0] + a[1] * q[1] + a[2] * q[2]) * m->magSf[m->nInt + b];
            }
            BC3[i + (size_t)j * Np] = s;
        }
    free(T); free(gp); free(curl); free(ngp);
}

/* steadyNS::pressure_BC1, steadyNS.C:1319-1354: BC1(i,j) = sum over boundary faces of
 * (interpolate(laplacian(U_j)) & Sf) * interpolate(P_i). The boundary value of the fvc::laplacian result is the owner
 * cell's value (fvc::surfaceIntegrate builds its result with extrapolatedCalculated patches). BC1: Np x N. */
void fvo_pressure_bc1(const fvo_mesh* m, int N, int Np, const double* const* U, const double* const* Ub, const double* const* Pb,
                      double* BC1)
{
    double* L = (double*)malloc(sizeof(double) * 3 * m->nCells);
    for (int j = 0; j < N; j++) {
        fvo_laplacian_vec(m, U[j], Ub[j], L);
        for (int i = 0; i < Np; i++) {
            double s = 0.0;
            for (long long b = 0; b < m->nBnd; b++) {
                if (!m->bmask[b]) continue;
                const long long f = m->nInt + b, c = m->owner[f];
                s += (L[3 * c] * m->Sf[3 * f] + L[3 * c + 1] * m->Sf[3 * f + 1] + L[3 * c + 2] * m->Sf[3 * f + 2]) * Pb[i][b];
            }
            BC1[i + (size_t)j * Np] = s;
        }
    }
    free(L);
}

/* steadyNS::pressureBC2 / pressure_BC2, steadyNS.C:1356-1437: BC2(i,j,k) = sum over boundary faces of
 * interpolate(div(interpolate(U_j) & Sf, U_k)) & (Sf * interpolate(P_i))  (C++ precedence: `*` binds before `&`). The
 * boundary value of the fvc::div result is the owner cell's value (extrapolatedCalculated patch of surfaceIntegrate).
 * Np x N x N with i fastest (Eigen::Tensor storage; slice i is BC2_matrix[i] of the List form). */
void fvo_pressure_bc2(const fvo_mesh* m, int N, int Np, const double* const* U, const double* const* Ub, const double* const* Pb,
                      double* BC2)
{
    double* phi = (double*)malloc(sizeof(double) * (m->nInt + m->nBnd));
    double* D = (double*)malloc(sizeof(double) * 3 * m->nCells);
    for (int i = 0; i < Np; i++)
        for (int j = 0; j < N; j++)
            for (int k = 0; k < N; k++) {
                fvo_flux(m, U[j], Ub[j], phi);
                fvo_div_phi_vec(m, phi, U[k], Ub[k], D);
                double s = 0.0;
                for (long long b = 0; b < m->nBnd; b++) {
                    if (!m->bmask[b]) continue;
                    const long long f = m->nInt + b, c = m->owner[f];
                    const double p = Pb[i][b];
                    s += D[3 * c] * (m->Sf[3 * f] * p) + D[3 * c + 1] * (m->Sf[3 * f + 1] * p) + D[3 * c + 2] * (m->Sf[3 * f + 2] * p);
                }
                BC2[i + (size_t)Np * (j + (size_t)N * k)] = s;
            }
    free(phi); free(D);
}

/* fvc::div(A) of a vector field A that came out of another fvc:: operator and was assigned to a volVectorField
 * (`L_U_SUPmodesaux = dt_dummy * fvc::laplacian(...)`, steadyNS.C:1598-1601, :1727-1736): GeometricField::operator=
 * copies the result's extrapolatedCalculated patch values, i.e. the boundary value is the owner cell's. out: nCells. */
static void div_vec_extrapolated(const fvo_mesh* m, const double* A, double* Ab, double* out)
{
    for (long long b = 0; b < m->nBnd; b++) {
        const long long c = m->owner[m->nInt + b];
        for (int d = 0; d < 3; d++) Ab[3 * b + d] = A[3 * c + d];
    }
    fvo_div_vec(m, A, Ab, out);
}

/* steadyNS::diffusive_term_flux_method, steadyNS.C:1584-1610: BP(i,j) = domainIntegrate(P_i * div(dt_dummy *
 * laplacian(nu_dummy, U_j))), dt_dummy = nu_dummy = 1. BP: Np x N. */
void fvo_diffusive_term_flux_method(const fvo_mesh* m, int N, int Np, const double* const* U, const double* const* Ub,
                                    const double* const* P, double* BP)
{
    double* L = (double*)malloc(sizeof(double) * 3 * m->nCells);
    double* Lb = (double*)malloc(sizeof(double) * 3 * (m->nBnd + 1));
    double* D = (double*)malloc(sizeof(double) * m->nCells);
    for (int i = 0; i < Np; i++)
        for (int j = 0; j < N; j++) {
            fvo_laplacian_vec(m, U[j], Ub[j], L);
            div_vec_extrapolated(m, L, Lb, D);
            BP[i + (size_t)j * Np] = fvo_domain_integrate_dot(m, P[i], D, 1);
        }
    free(L); free(Lb); free(D);
}

/* steadyNS::convective_term_flux_tens, steadyNS.C:1704-1746 (fluxMethod != "consistent"): Cf(i,j,k) =
 * domainIntegrate(P_i * div(dt_dummy * div(flux(U_j), U_k))), fvc::flux(U) = interpolate(U) & Sf. Np x N x N, i fastest. */
void fvo_convective_term_flux_tens(const fvo_mesh* m, int N, int Np, const double* const* U, const double* const* Ub,
                                   const double* const* P, double* Cf)
{
    double* phi = (double*)malloc(sizeof(double) * (m->nInt + m->nBnd));
    double* A = (double*)malloc(sizeof(double) * 3 * m->nCells);
    double* Ab = (double*)malloc(sizeof(double) * 3 * (m->nBnd + 1));
    double* D = (double*)malloc(sizeof(double) * m->nCells);
    for (int j = 0; j < N; j++) {
        fvo_flux(m, U[j], Ub[j], phi);
        for (int k = 0; k < N; k++) {
            fvo_div_phi_vec(m, phi, U[k], Ub[k], A);
            div_vec_extrapolated(m, A, Ab, D);
            for (int i = 0; i < Np; i++) Cf[i + (size_t)Np * (j + (size_t)N * k)] = fvo_domain_integrate_dot(m, P[i], D, 1);
        }
    }
    free(phi); free(A); free(Ab); free(D);
}

/* steadyNS::convective_background (kind 0, steadyNS.C:1256-1273): L(i,j) = -domainIntegrate(U_i & div(interpolate(vls) & Sf, U_j)),
 * N x N; steadyNS::divergent_convective_background (kind 1, :1275-1293): L_D(i,j) = -domainIntegrate(grad(P_i) & div(...)), Np x N. */
void fvo_convective_background(const fvo_mesh* m, int kind, int N, int Np, const double* const* U, const double* const* Ub,
                               const double* const* P, const double* const* Pb, const double* vls, const double* vlsb, double* out)
{
    const int rows = kind == 0 ? N : Np;
    double* phi = (double*)malloc(sizeof(double) * (m->nInt + m->nBnd));
    double* D = (double*)malloc(sizeof(double) * 3 * m->nCells);
    double* g = (double*)malloc(sizeof(double) * 3 * m->nCells);
    fvo_flux(m, vls, vlsb, phi);
    for (int j = 0; j < N; j++) {
        fvo_div_phi_vec(m, phi, U[j], Ub[j], D);
        for (int i = 0; i < rows; i++) {
            const double* a = U ? U[i] : 0;
            if (kind == 1) { fvo_grad_scalar(m, P[i], Pb[i], g); a = g; }
            out[i + (size_t)j * rows] = -fvo_domain_integrate_dot(m, a, D, 3);
        }
    }
    free(phi); free(D); free(g);
}

/* ---- addressing: cell -> face CSR in ascending face order (= primitiveMesh::cells() order) -------------
 * entry = +(f+1) when the cell owns face f (Sf points outwards), -(f+1) when it is the neighbour.
 * off: nCells+1, ent/other: 2*nInt+nBnd. other = the cell across the face, or -(b+1) for boundary face b. */
void fvo_cell_faces(const fvo_mesh* m, int* off, int* ent, int* other)
{
    const long long nF = m->nInt + m->nBnd;
    int* cnt = (int*)calloc(m->nCells + 1, sizeof(int));
    for (long long f = 0; f < nF; f++) cnt[m->owner[f] + 1]++;
    for (long long f = 0; f < m->nInt; f++) cnt[m->neighbour[f] + 1]++;
    off[0] = 0;
    for (long long c = 0; c < m->nCells; c++) off[c + 1] = off[c] + cnt[c + 1];
    for (long long c = 0; c < m->nCells; c++) cnt[c] = off[c];
    for (long long f = 0; f < nF; f++) {
        const int o = m->owner[f];
        ent[cnt[o]] = (int)(f + 1);
        other[cnt[o]++] = (f < m->nInt) ? m->neighbour[f] : -(int)(f - m->nInt + 1);
        if (f < m->nInt) {
            const int n = m->neighbour[f];
            ent[cnt[n]] = -(int)(f + 1);
            other[cnt[n]++] = o;
        }
    }
    free(cnt);
}
