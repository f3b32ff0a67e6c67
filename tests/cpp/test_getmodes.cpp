// Test driver for host/ITHACAPOD_b200.H without OpenFOAM: a minimal field type + adapter, inputs read from a binary
// file written by tests/test_cpp_shim.py, modes written back for comparison with the oracle.
//   usage: test_getmodes <input.bin> <output.bin> <outDir>
#include "../../host/ITHACAPOD_b200.H"
#include <unistd.h>

#include <cstdint>
#include <cstring>
#include <sys/stat.h>

struct MiniMesh
{
    long long nCells = 0;
    int dim = 1;
    std::vector<long long> patchSize;
    std::vector<int> patchLifted;
    std::vector<double> V;
};

struct MiniField
{
    const MiniMesh* mesh;
    std::string name;
    std::vector<double> internal;
    std::vector<std::vector<double>> patches;
};

namespace ITHACAPOD_B200
{
template<>
struct FieldAdapter<MiniField>
{
    static int g_dim;
    static int nComponents() { return g_dim; }
    static long long nCells(const MiniField& f) { return f.mesh->nCells; }
    static const double* internalPtr(const MiniField& f) { return f.internal.data(); }
    static double* internalPtr(MiniField& f) { return f.internal.data(); }
    static std::string name(const MiniField& f) { return f.name; }
    static int nPatches(const MiniField& f) { return (int)f.mesh->patchSize.size(); }
    static bool patchIsLifted(const MiniField& f, int p) { return f.mesh->patchLifted[p] != 0; }
    static long long patchSize(const MiniField& f, int p) { return f.mesh->patchSize[p]; }
    static const double* patchPtr(const MiniField& f, int p) { return f.patches[p].data(); }
    static void assignPatch(MiniField& f, int p, const double* v, bool) { std::copy(v, v + f.patches[p].size(), f.patches[p].begin()); }
    static const double* cellVolumes(const MiniField& f) { return f.mesh->V.data(); }
    static MiniField* clone(const MiniField& f) { return new MiniField(f); }
};
int FieldAdapter<MiniField>::g_dim = 1;
}  // namespace ITHACAPOD_B200

template<class T>
static void rd(FILE* f, T* p, size_t n)
{
    if (fread(p, sizeof(T), n, f) != n)
    {
        std::fprintf(stderr, "short read\n");
        std::exit(2);
    }
}

int main(int argc, char** argv)
{
    if (argc < 4)
    {
        std::fprintf(stderr, "usage: %s in.bin out.bin outDir [rank nranks idfile]\n", argv[0]);
        return 2;
    }
    FILE* fi = std::fopen(argv[1], "rb");
    if (!fi) return 2;
    int64_t hdr[8];
    rd(fi, hdr, 8);  // nCells, dim, m, nmodes, npatch, solver, norm, podex
    MiniMesh mesh;
    mesh.nCells = hdr[0];
    mesh.dim = (int)hdr[1];
    const int m = (int)hdr[2], nmodes = (int)hdr[3], np = (int)hdr[4];
    ITHACAPOD_B200::FieldAdapter<MiniField>::g_dim = mesh.dim;
    mesh.patchSize.resize(np);
    mesh.patchLifted.resize(np);
    std::vector<int64_t> tmp(np);
    rd(fi, tmp.data(), np);
    for (int p = 0; p < np; p++) mesh.patchSize[p] = tmp[p];
    rd(fi, tmp.data(), np);
    for (int p = 0; p < np; p++) mesh.patchLifted[p] = (int)tmp[p];
    mesh.V.resize(mesh.nCells);
    rd(fi, mesh.V.data(), mesh.nCells);
    std::vector<MiniField*> snaps(m), modes;
    for (int j = 0; j < m; j++)
    {
        MiniField* f = new MiniField();
        f->mesh = &mesh;
        f->name = "U";
        f->internal.resize(mesh.nCells * mesh.dim);
        rd(fi, f->internal.data(), f->internal.size());
        f->patches.resize(np);
        for (int p = 0; p < np; p++)
        {
            f->patches[p].resize(mesh.patchSize[p] * mesh.dim);
            rd(fi, f->patches[p].data(), f->patches[p].size());
        }
        snaps[j] = f;
    }
    std::fclose(fi);

    ITHACAPOD_B200::Parameters para;
    para.eigensolver = hdr[5] == 0 ? "spectra" : "eigen";
    para.podNorm = hdr[6] == 0 ? "L2" : "Frobenius";
    para.outDir = argv[3];
    para.precision = 16;
    para.outytpe = std::ios_base::scientific;
    if (argc >= 7)
    {
        // Pstream::parRun() stand-in: this process is one rank with its own cell shard and GPU; rank 0's NCCL id travels
        // through a file (in ITHACA-FV: Pstream::broadcast)
        para.parallel.rank = std::atoi(argv[4]);
        para.parallel.nranks = std::atoi(argv[5]);
        para.device = para.parallel.rank;
        const std::string idfile = argv[6];
        const int rank = para.parallel.rank;
        para.parallel.broadcastId = [idfile, rank](void* id128)
        {
            if (rank == 0)
            {
                FILE* f = std::fopen((idfile + ".tmp").c_str(), "wb");
                std::fwrite(id128, 1, 128, f);
                std::fclose(f);
                std::rename((idfile + ".tmp").c_str(), idfile.c_str());
                return;
            }
            for (int tries = 0; tries < 6000; tries++)
            {
                FILE* f = std::fopen(idfile.c_str(), "rb");
                if (f)
                {
                    const size_t n = std::fread(id128, 1, 128, f);
                    std::fclose(f);
                    if (n == 128) return;
                }
                usleep(10000);
            }
            std::fprintf(stderr, "id transport timed out\n");
            std::abort();
        };
    }
    mkdir(argv[3], 0777);
    mkdir((std::string(argv[3]) + "/POD").c_str(), 0777);
    int exported = 0, readCalls = 0;
    auto exportFields = [&](std::vector<MiniField*>& mds, const std::string&, const std::string&) { exported = (int)mds.size(); };
    auto readFields = [&](std::vector<MiniField*>&, const std::string&, const std::string&) { readCalls++; };
    ITHACAPOD_B200::Result r = ITHACAPOD_B200::getModes<MiniField>(snaps, modes, "U", hdr[7] != 0, false, false, nmodes, true, para,
                                                                   exportFields, readFields);
    FILE* fo = std::fopen(argv[2], "wb");
    int64_t oh[4] = {(int64_t)modes.size(), (int64_t)r.eigenvalues.size(), exported, readCalls};
    std::fwrite(oh, sizeof(int64_t), 4, fo);
    std::fwrite(r.eigenvalues.data(), sizeof(double), r.eigenvalues.size(), fo);
    std::fwrite(r.cumEigenvalues.data(), sizeof(double), r.cumEigenvalues.size(), fo);
    std::fwrite(r.normFactors.data(), sizeof(double), r.normFactors.size(), fo);
    for (MiniField* f : modes)
    {
        std::fwrite(f->internal.data(), sizeof(double), f->internal.size(), fo);
        for (auto& p : f->patches) std::fwrite(p.data(), sizeof(double), p.size(), fo);
    }
    std::fclose(fo);
    return 0;
}
