// Test driver of host/steadyNS_b200.H (the C++ shim of steadyNS::projectSUP / projectPPE above the C ABI).
// usage: test_projectsup <input.bin> <matrices_dir/> ; reads mesh + basis views from the file, runs the projection
// exactly as an ITHACA-FV tutorial would (example.projectSUP("./Matrices", NU, NP, NSUP)) and leaves the operators in
// the reference's files under <matrices_dir>. Prints "computed <n> read <n>".
#include "../../host/steadyNS_b200.H"

#include <cstdint>
#include <cstdio>
#include <fstream>

template<class T>
static std::vector<T> rd(std::ifstream& f, size_t n)
{
    std::vector<T> v(n);
    if (n) f.read(reinterpret_cast<char*>(v.data()), sizeof(T) * n);
    return v;
}

int main(int argc, char** argv)
{
    if (argc < 3) { std::fprintf(stderr, "usage: %s input.bin matrices_dir/\n", argv[0]); return 2; }
    std::ifstream f(argv[1], std::ios::binary);
    if (!f) { std::fprintf(stderr, "cannot open %s\n", argv[1]); return 2; }
    auto hd = rd<int64_t>(f, 16);
    const int64_t nC = hd[0], nI = hd[1], nB = hd[2], nLift = hd[3], nU = hd[4], nSup = hd[5], nP = hd[6], NU = hd[7], NP = hd[8],
                  NSUP = hd[9], corrected = hd[10], ppe = hd[11], penalty = hd[12], nPatches = hd[13], nInlet = hd[14], exports = hd[15];
    auto sch = rd<double>(f, 2);
    const int64_t nF = nI + nB;
    auto owner = rd<int32_t>(f, nF);
    auto nei = rd<int32_t>(f, nI);
    auto Sf = rd<double>(f, 3 * nF);
    auto magSf = rd<double>(f, nF);
    auto w = rd<double>(f, nI);
    auto dc = rd<double>(f, nI);
    auto dcB = rd<double>(f, nB);
    auto corr = rd<double>(f, corrected ? 3 * nI : 0);
    auto V = rd<double>(f, nC);
    auto bmask = rd<unsigned char>(f, nB);
    auto C = rd<double>(f, 3 * nC);
    auto Cf = rd<double>(f, 3 * nF);
    auto pstart = rd<int64_t>(f, nPatches);
    auto psize = rd<int64_t>(f, nPatches);
    auto inlet = rd<int64_t>(f, 2 * nInlet);
    std::vector<std::vector<double>> store;
    auto field = [&](int dim) {
        steadyNS_B200::FieldView v;
        store.push_back(rd<double>(f, dim * nC));
        v.internal = store.back().data();
        store.push_back(rd<double>(f, dim * nB));
        v.boundary = store.back().data();
        return v;
    };
    store.reserve(2 * (nLift + nU + nSup + nP) + 4);

    steadyNS_B200::MeshView mesh;
    mesh.nCells = nC; mesh.nInternalFaces = nI; mesh.nBoundaryFaces = nB;
    mesh.owner = owner.data(); mesh.neighbour = nei.data(); mesh.Sf = Sf.data(); mesh.magSf = magSf.data();
    mesh.weights = w.data(); mesh.nonOrthDeltaCoeffs = dc.data(); mesh.boundaryDeltaCoeffs = dcB.data();
    mesh.nonOrthCorrectionVectors = corrected ? corr.data() : nullptr;
    mesh.V = V.data(); mesh.boundaryMask = bmask.data(); mesh.C = C.data(); mesh.Cf = Cf.data();
    steadyNS_B200::Schemes schemes;
    schemes.gradLimitK = sch[0]; schemes.snGradPsi = sch[1];
    steadyNS_B200::Parameters para;
    para.matricesDir = argv[2];
    para.bcMethod = penalty ? "penalty" : "lift";
    para.exportPython = exports & 1; para.exportMatlab = exports & 2; para.exportTxt = exports & 4; para.exportNpy = exports & 8;

    steadyNS_B200::steadyNS example(mesh, schemes, para);
    for (int64_t k = 0; k < nLift; k++) example.liftfield.push_back(field(3));
    for (int64_t k = 0; k < nU; k++) example.Umodes.push_back(field(3));
    for (int64_t k = 0; k < nSup; k++) example.supmodes.push_back(field(3));
    for (int64_t k = 0; k < nP; k++) example.Pmodes.push_back(field(1));
    for (int64_t p = 0; p < nPatches; p++) { example.patchStart.push_back(pstart[p]); example.patchSize.push_back(psize[p]); }
    for (int64_t k = 0; k < nInlet; k++) example.inletIndex.push_back({(long)inlet[2 * k], (long)inlet[2 * k + 1]});

    if (ppe) example.projectPPE("./Matrices", NU, NP, NSUP);
    else example.projectSUP("./Matrices", NU, NP, NSUP);
    std::printf("computed %d read %d\n", example.computed, example.readFromFile);
    return 0;
}
