// Internal locality renumbering of the mesh for the projection gather kernels (host code, run once per mesh_upload).
//
// OpenFOAM meshes arrive with whatever cell labels the mesher produced (the C4 benchmark mesh is
// randomly relabelled on purpose). Every projected operator is a SUM over cells or faces, so its value does not depend
// on the labelling (only the rounding order does, at the 1e-16 level against the 1e-10 gate); the gather kernels of
// proj.cu / conv.cu, however, read the two cells of every face, and with random labels each of those reads is its own
// DRAM sector. The library therefore works on a relabelled copy: cells in Cuthill-McKee (breadth-first, low degree
// first) order, internal faces sorted by their (lower, higher) new cell label -- the same idea as OpenFOAM's
// renumberMesh, applied behind the ABI. The caller's labelling is kept for everything that crosses the boundary:
// ithaca_b200_mesh_addressing reports the cell->face addressing of the ORIGINAL arrays, ithaca_b200_fvc_eval returns
// fields in the original cell / face order, and ithaca_b200_basis_upload takes fields in the original order.
#include "common.cuh"
#include <algorithm>
#include <numeric>
#include <vector>

namespace ib200 {

// internal faces sorted by (lower new label, higher new label): counting sort on the lower label, then the (few)
// faces of one cell by the higher label
static void sort_faces(long long nCells, long long nInt, const int* owner, const int* neigh, const std::vector<int>& old2new,
                       std::vector<int>& fnew2old)
{
    const size_t nC = (size_t)nCells, nF = (size_t)nInt;
    std::vector<int> lo(nF), hi(nF), cnt(nC + 1, 0);
    for (size_t f = 0; f < nF; f++) {
        const int a = old2new[owner[f]], b = old2new[neigh[f]];
        lo[f] = std::min(a, b);
        hi[f] = std::max(a, b);
        cnt[lo[f] + 1]++;
    }
    for (size_t c = 0; c < nC; c++) cnt[c + 1] += cnt[c];
    fnew2old.assign(nF, 0);
    {
        std::vector<int> pos(cnt.begin(), cnt.end() - 1);
        for (size_t f = 0; f < nF; f++) fnew2old[pos[lo[f]]++] = (int)f;
    }
    for (size_t c = 0; c < nC; c++)
        std::sort(fnew2old.begin() + cnt[c], fnew2old.begin() + cnt[c + 1],
                  [&](int x, int y) { return hi[x] != hi[y] ? hi[x] < hi[y] : x < y; });
}

// old2new[c] = new label of cell c; fnew2old[f'] = original index of the internal face at new position f'.
void mesh_renumber(long long nCells, long long nInt, const int* owner, const int* neigh, std::vector<int>& old2new,
                   std::vector<int>& fnew2old)
{
    const size_t nC = (size_t)nCells, nF = (size_t)nInt;
    // cell -> cell adjacency (CSR) over internal faces
    std::vector<int> off(nC + 1, 0);
    for (size_t f = 0; f < nF; f++) { off[owner[f] + 1]++; off[neigh[f] + 1]++; }
    for (size_t c = 0; c < nC; c++) off[c + 1] += off[c];
    std::vector<int> adj(2 * nF), cur(off.begin(), off.end() - 1);
    for (size_t f = 0; f < nF; f++) {
        adj[cur[owner[f]]++] = neigh[f];
        adj[cur[neigh[f]]++] = owner[f];
    }
    auto degree = [&](int c) { return off[c + 1] - off[c]; };

    // breadth-first search from `start` over unvisited cells, neighbours in order of increasing degree; appends to order
    std::vector<int> order;
    order.reserve(nC);
    std::vector<unsigned char> seen(nC, 0);
    std::vector<int> nb;
    auto bfs = [&](int start, std::vector<unsigned char>& mark, std::vector<int>& out) {
        const size_t head0 = out.size();
        out.push_back(start);
        mark[start] = 1;
        for (size_t head = head0; head < out.size(); head++) {
            const int c = out[head];
            nb.clear();
            for (int e = off[c]; e < off[c + 1]; e++)
                if (!mark[adj[e]]) { mark[adj[e]] = 1; nb.push_back(adj[e]); }
            std::sort(nb.begin(), nb.end(), [&](int x, int y) { return degree(x) != degree(y) ? degree(x) < degree(y) : x < y; });
            out.insert(out.end(), nb.begin(), nb.end());
        }
    };
    std::vector<unsigned char> probe_mark;
    std::vector<int> probe;
    for (size_t c0 = 0; c0 < nC; c0++) {
        if (seen[c0]) continue;
        // pseudo-peripheral start: the last cell reached by a search from the first unvisited cell of this component
        probe_mark = seen;
        probe.clear();
        bfs((int)c0, probe_mark, probe);
        bfs(probe.back(), seen, order);
    }
    old2new.assign(nC, 0);
    for (size_t k = 0; k < nC; k++) old2new[order[k]] = (int)k;

    sort_faces(nCells, nInt, owner, neigh, old2new, fnew2old);
}

// Two-level ordering for the gather kernels: the Cuthill-McKee sequence is cut into compact clusters of <= `block` cells
// (a breadth-first ball grown from the first still unassigned cell of the sequence, over unassigned cells only), and the
// clusters are numbered in the order their seeds appear. A thread block of the cell kernels then covers one cluster, most
// of whose face neighbours lie INSIDE it (L1 hits), instead of a 1-D run of a breadth-first level whose neighbours in the
// levels before and after are an L2 round trip away.
void mesh_renumber_blocked(long long nCells, long long nInt, const int* owner, const int* neigh, int block, std::vector<int>& old2new,
                           std::vector<int>& fnew2old)
{
    const size_t nC = (size_t)nCells, nF = (size_t)nInt;
    std::vector<int> cm, ftmp;
    mesh_renumber(nCells, nInt, owner, neigh, cm, ftmp);
    std::vector<int> seq(nC);
    for (size_t c = 0; c < nC; c++) seq[cm[c]] = (int)c;
    std::vector<int> off(nC + 1, 0);
    for (size_t f = 0; f < nF; f++) { off[owner[f] + 1]++; off[neigh[f] + 1]++; }
    for (size_t c = 0; c < nC; c++) off[c + 1] += off[c];
    std::vector<int> adj(2 * nF), cur(off.begin(), off.end() - 1);
    for (size_t f = 0; f < nF; f++) {
        adj[cur[owner[f]]++] = neigh[f];
        adj[cur[neigh[f]]++] = owner[f];
    }
    old2new.assign(nC, -1);
    std::vector<int> ball;
    ball.reserve((size_t)block);
    int next = 0;
    for (size_t k = 0; k < nC; k++) {
        const int seed = seq[k];
        if (old2new[seed] >= 0) continue;
        ball.clear();
        ball.push_back(seed);
        old2new[seed] = next++;
        for (size_t head = 0; head < ball.size() && (int)ball.size() < block; head++) {
            const int c = ball[head];
            for (int e = off[c]; e < off[c + 1] && (int)ball.size() < block; e++) {
                const int x = adj[e];
                if (old2new[x] < 0) { old2new[x] = next++; ball.push_back(x); }
            }
        }
    }
    sort_faces(nCells, nInt, owner, neigh, old2new, fnew2old);
}

}  // namespace ib200
