// Cell-range partitioner with halo layers for the sharded Galerkin projection (host code).
//
// Ownership rule: rank g of G owns cells [floor(g N/G), floor((g+1) N/G)) -- ithaca_b200_cell_range, the rule the POD
// path shards by. The shard mesh = owned cells + `halo` layers of face neighbours; its addressing is reported in the
// form OpenFOAM's decomposePar writes and ITHACAparallel reads (src/ITHACA_CORE/ITHACAparallel/ITHACAparallel.C:43-74):
// cellProcAddressing = global cell label, faceProcAddressing = +-(global face + 1), negative when the shard face is
// flipped (the shard-side cell is the global neighbour). The same rule is restated in oracle/partition.py and the two
// are compared integer for integer (tests/test_partition.py).
#include "../../include/ithaca_b200.h"
#include "common.cuh"
#include <algorithm>
#include <vector>

namespace ib200 {
void mesh_renumber(long long nCells, long long nInt, const int* owner, const int* neigh, std::vector<int>& old2new,
                   std::vector<int>& fnew2old);   // renumber.cu
}

struct ithaca_b200_partition {
    int64_t nCells = 0, nInt = 0, nBnd = 0;          // global mesh
    int64_t nOwned = 0, nLocInt = 0, nLocBnd = 0, nCut = 0;
    std::vector<int32_t> cpa;                        // shard cell -> global cell
    std::vector<int32_t> fpa;                        // shard face -> +-(global face + 1)
    std::vector<int32_t> g2l;                        // global cell -> shard cell or -1
};

extern "C" {

int ithaca_b200_partition_create(ithaca_b200_partition_t** out, int64_t nCells, int64_t nInt, int64_t nBnd, const int32_t* owner,
                                 const int32_t* neighbour, int rank, int nranks, int halo, int method)
{
    if (!out) return IB200_ERR_ARG;
    *out = nullptr;
    if (nCells <= 0 || nInt < 0 || nBnd < 0 || !owner || (!neighbour && nInt > 0) || rank < 0 || nranks < 1 || rank >= nranks ||
        halo < 0 || (method != IB200_PART_LABEL_RANGE && method != IB200_PART_LOCALITY_RANGE) || nCells >= 0x7fffffffLL || nInt + nBnd >= 0x7fffffffLL)
        return IB200_ERR_ARG;
    for (int64_t f = 0; f < nInt + nBnd; f++)
        if (owner[f] < 0 || owner[f] >= nCells || (f < nInt && (neighbour[f] < 0 || neighbour[f] >= nCells))) return IB200_ERR_ARG;
    int64_t c0 = 0, c1 = 0;
    ithaca_b200_cell_range(nCells, rank, nranks, &c0, &c1);
    ithaca_b200_partition* p = new ithaca_b200_partition();
    p->nCells = nCells; p->nInt = nInt; p->nBnd = nBnd;
    p->nOwned = c1 - c0;
    // level[c]: 0 owned, L = halo layer L, -1 outside
    std::vector<signed char> level((size_t)nCells, -1);
    if (method == IB200_PART_LOCALITY_RANGE) {
        std::vector<int> old2new, fnew2old;
        ib200::mesh_renumber(nCells, nInt, owner, neighbour, old2new, fnew2old);
        for (int64_t c = 0; c < nCells; c++)
            if (old2new[c] >= c0 && old2new[c] < c1) level[c] = 0;
    } else {
        for (int64_t c = c0; c < c1; c++) level[c] = 0;
    }
    const int nl = std::min(halo, 120);
    for (int L = 1; L <= nl; L++) {
        bool any = false;
        for (int64_t f = 0; f < nInt; f++) {
            const int o = owner[f], n = neighbour[f];
            if (level[o] == L - 1 && level[n] < 0) { level[n] = (signed char)L; any = true; }
            else if (level[n] == L - 1 && level[o] < 0) { level[o] = (signed char)L; any = true; }
        }
        if (!any) break;
    }
    p->g2l.assign((size_t)nCells, -1);
    p->cpa.reserve((size_t)(c1 - c0));
    for (int64_t c = 0; c < nCells; c++)
        if (level[c] == 0) { p->g2l[c] = (int32_t)p->cpa.size(); p->cpa.push_back((int32_t)c); }
    for (int64_t c = 0; c < nCells; c++)
        if (level[c] > 0) { p->g2l[c] = (int32_t)p->cpa.size(); p->cpa.push_back((int32_t)c); }
    std::vector<int32_t> cut;
    for (int64_t f = 0; f < nInt; f++) {
        const bool lo = level[owner[f]] >= 0, ln = level[neighbour[f]] >= 0;
        if (lo && ln) p->fpa.push_back((int32_t)(f + 1));
        else if (lo) cut.push_back((int32_t)(f + 1));
        else if (ln) cut.push_back((int32_t)-(f + 1));
    }
    p->nLocInt = (int64_t)p->fpa.size();
    for (int64_t f = nInt; f < nInt + nBnd; f++)
        if (level[owner[f]] >= 0) p->fpa.push_back((int32_t)(f + 1));
    p->nLocBnd = (int64_t)p->fpa.size() - p->nLocInt;
    p->nCut = (int64_t)cut.size();
    p->fpa.insert(p->fpa.end(), cut.begin(), cut.end());
    *out = p;
    return IB200_OK;
}

int ithaca_b200_mesh_renumbering(int64_t nCells, int64_t nInt, const int32_t* owner, const int32_t* neighbour, int32_t* old2new)
{
    if (nCells <= 0 || nInt < 0 || !old2new || (nInt > 0 && (!owner || !neighbour)) || nCells >= 0x7fffffffLL) return IB200_ERR_ARG;
    for (int64_t f = 0; f < nInt; f++)
        if (owner[f] < 0 || owner[f] >= nCells || neighbour[f] < 0 || neighbour[f] >= nCells) return IB200_ERR_ARG;
    std::vector<int> o2n, f2o;
    ib200::mesh_renumber(nCells, nInt, owner, neighbour, o2n, f2o);
    std::copy(o2n.begin(), o2n.end(), old2new);
    return IB200_OK;
}

int ithaca_b200_partition_destroy(ithaca_b200_partition_t* p)
{
    delete p;
    return IB200_OK;
}

int ithaca_b200_partition_sizes(const ithaca_b200_partition_t* p, int64_t* s)
{
    if (!p || !s) return IB200_ERR_ARG;
    s[0] = (int64_t)p->cpa.size(); s[1] = p->nOwned; s[2] = p->nLocInt; s[3] = p->nLocBnd; s[4] = p->nCut;
    return IB200_OK;
}

int ithaca_b200_partition_addressing(const ithaca_b200_partition_t* p, int32_t* cpa, int32_t* fpa)
{
    if (!p) return IB200_ERR_ARG;
    if (cpa) std::copy(p->cpa.begin(), p->cpa.end(), cpa);
    if (fpa) std::copy(p->fpa.begin(), p->fpa.end(), fpa);
    return IB200_OK;
}

int ithaca_b200_partition_gather(const ithaca_b200_partition_t* p, int kind, int dim, const double* g, double* l)
{
    if (!p || !g || !l || dim < 1 || (kind != 0 && kind != 1)) return IB200_ERR_ARG;
    if (kind == 0) {
        for (size_t c = 0; c < p->cpa.size(); c++)
            for (int d = 0; d < dim; d++) l[c * dim + d] = g[(size_t)p->cpa[c] * dim + d];
    } else {
        const size_t nb = (size_t)(p->nLocBnd + p->nCut);
        for (size_t b = 0; b < nb; b++) {
            const bool real = (int64_t)b < p->nLocBnd;
            const size_t gb = real ? (size_t)(p->fpa[p->nLocInt + b] - 1 - p->nInt) : 0;
            for (int d = 0; d < dim; d++) l[b * dim + d] = real ? g[gb * dim + d] : 0.0;
        }
    }
    return IB200_OK;
}

int ithaca_b200_mesh_upload_shard(ithaca_b200_t* h, const ithaca_b200_partition_t* p, const int32_t* owner, const int32_t* neighbour,
                                  const double* Sf, const double* magSf, const double* weights, const double* dc, const double* dcB,
                                  const double* corr, const double* V, const unsigned char* bndMask)
{
    if (!h) return IB200_ERR_ARG;
    if (!p || !owner || (!neighbour && p->nInt > 0) || !Sf || !magSf || (!weights && p->nInt > 0) || (!dc && p->nInt > 0) ||
        (!dcB && p->nBnd > 0) || !V) {
        ib200::set_error(h, "mesh_upload_shard: bad argument");
        return IB200_ERR_ARG;
    }
    const size_t nC = p->cpa.size(), nI = (size_t)p->nLocInt, nB = (size_t)(p->nLocBnd + p->nCut), nF = nI + nB;
    std::vector<int32_t> lo(nF), ln(std::max<size_t>(nI, 1));
    std::vector<double> lSf(3 * nF), lmag(nF), lw(std::max<size_t>(nI, 1)), ldc(std::max<size_t>(nI, 1)), ldcB(std::max<size_t>(nB, 1)),
        lcorr(corr ? 3 * std::max<size_t>(nI, 1) : 0), lV(nC);
    std::vector<unsigned char> lmask(std::max<size_t>(nB, 1), 1);
    for (size_t f = 0; f < nF; f++) {
        const int32_t a = p->fpa[f];
        const size_t g = (size_t)(a > 0 ? a : -a) - 1;
        const double sgn = a > 0 ? 1.0 : -1.0;
        lSf[3 * f] = sgn * Sf[3 * g]; lSf[3 * f + 1] = sgn * Sf[3 * g + 1]; lSf[3 * f + 2] = sgn * Sf[3 * g + 2];
        lmag[f] = magSf[g];
        if (f < nI) {
            lo[f] = p->g2l[owner[g]];
            ln[f] = p->g2l[neighbour[g]];
            lw[f] = weights[g];
            ldc[f] = dc[g];
            if (corr) { lcorr[3 * f] = corr[3 * g]; lcorr[3 * f + 1] = corr[3 * g + 1]; lcorr[3 * f + 2] = corr[3 * g + 2]; }
        } else {
            const size_t b = f - nI;
            if ((int64_t)b < p->nLocBnd) {
                lo[f] = p->g2l[owner[g]];
                ldcB[b] = dcB[g - (size_t)p->nInt];
                lmask[b] = bndMask ? bndMask[g - (size_t)p->nInt] : 1;
            } else {  // cut face: the shard-side cell owns it; holds no values
                lo[f] = p->g2l[a > 0 ? owner[g] : neighbour[g]];
                ldcB[b] = dc[g];
                lmask[b] = 0;
            }
        }
    }
    for (size_t c = 0; c < nC; c++) lV[c] = V[p->cpa[c]];
    int rc = ithaca_b200_mesh_upload(h, (int64_t)nC, (int64_t)nI, (int64_t)nB, lo.data(), ln.data(), lSf.data(), lmag.data(), lw.data(),
                                     ldc.data(), ldcB.data(), corr ? lcorr.data() : nullptr, lV.data(), lmask.data());
    if (rc) return rc;
    return ithaca_b200_mesh_set_owned(h, p->nOwned);
}

}  // extern "C"
