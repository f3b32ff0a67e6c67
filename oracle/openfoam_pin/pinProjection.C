/*---------------------------------------------------------------------------*\
  pinProjection -- dumps the golden files that would PIN oracle/fvc_oracle.c at the OpenFOAM boundary.
  CANNOT be built in this image (no OpenFOAM, no network): a maintainer with OpenFOAM v2412 + ITHACA-FV runs it once on
  tutorials/CFD/04unsteadyNS (mesh + fvSchemes as shipped) and commits the output under tests/golden/openfoam_pin/.
  It fills N vector / Np scalar fields with the SAME seeded synthetic basis as oracle/fvc.py::synthetic_basis (values read from
  basis.bin written by make_basis.py, so no RNG has to be reproduced), evaluates the reference's own expressions
  (steadyNS.C:883-1137) with real fvc:: calls and writes B, K, P, M, C as raw column-major doubles.
\*---------------------------------------------------------------------------*/
#include "fvCFD.H"
#include <fstream>
#include <vector>

static void dump(const word& name, const std::vector<double>& a)
{
    std::ofstream f((name + ".bin").c_str(), std::ios::binary);
    f.write(reinterpret_cast<const char*>(a.data()), a.size() * sizeof(double));
}

int main(int argc, char* argv[])
{
    #include "setRootCase.H"
    #include "createTime.H"
    #include "createMesh.H"
    std::ifstream in("basis.bin", std::ios::binary);   // header: N, Np (int64), then per field: internal values, boundary values
    int64_t N = 0, Np = 0;
    in.read(reinterpret_cast<char*>(&N), 8);
    in.read(reinterpret_cast<char*>(&Np), 8);
    PtrList<volVectorField> U(N);
    PtrList<volScalarField> P(Np);
    auto fill = [&](auto& fld, int nc)
    {
        std::vector<double> v(nc * mesh.nCells());
        in.read(reinterpret_cast<char*>(v.data()), v.size() * 8);
        forAll(fld, c) for (int d = 0; d < nc; d++) setComponent(fld.ref()[c], d) = v[nc * c + d];
        forAll(fld.boundaryField(), p)
        {
            auto& pf = fld.boundaryFieldRef()[p];
            std::vector<double> b(nc * pf.size());
            in.read(reinterpret_cast<char*>(b.data()), b.size() * 8);   // empty patches have size 0 and read nothing
            forAll(pf, i) for (int d = 0; d < nc; d++) setComponent(pf[i], d) = b[nc * i + d];
        }
    };
    for (label j = 0; j < N; j++)
    {
        U.set(j, new volVectorField(IOobject("U" + name(j), runTime.timeName(), mesh), mesh, dimensionedVector(dimless, Zero), "calculated"));
        fill(U[j], 3);
    }
    for (label j = 0; j < Np; j++)
    {
        P.set(j, new volScalarField(IOobject("P" + name(j), runTime.timeName(), mesh), mesh, dimensionedScalar(dimless, 0), "calculated"));
        fill(P[j], 1);
    }
    std::vector<double> B(N * N), M(N * N), K(N * Np), Pm(Np * N), C(N * N * N);
    for (label i = 0; i < N; i++) for (label j = 0; j < N; j++)
    {
        B[i + N * j] = fvc::domainIntegrate(U[i] & fvc::laplacian(dimensionedScalar("1", dimless, 1), U[j])).value();   // steadyNS.C:895-897
        M[i + N * j] = fvc::domainIntegrate(U[i] & U[j]).value();                                                        // :1096-1097
        for (label k = 0; k < N; k++)
            C[i + N * (j + N * k)] = fvc::domainIntegrate(U[i] & fvc::div(linearInterpolate(U[j]) & mesh.Sf(), U[k])).value();   // :1019-1021
    }
    for (label i = 0; i < N; i++) for (label j = 0; j < Np; j++)
    {
        K[i + N * j] = fvc::domainIntegrate(U[i] & fvc::grad(P[j])).value();     // :947-948
        Pm[j + Np * i] = fvc::domainIntegrate(P[j] * fvc::div(U[i])).value();    // :1124-1125
    }
    dump("B", B); dump("M", M); dump("K", K); dump("P", Pm); dump("C", C);
    Info << "pinProjection: wrote B M K P C for N = " << N << ", Np = " << Np << endl;
    return 0;
}
