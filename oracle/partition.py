"""TEST INFRASTRUCTURE (oracle): numpy restatement of the cell-range partition with halo layers.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module; the product path is
ithaca_b200_partition_* in libithaca_b200.so.

Semantics follow the decomposition files ITHACAparallel reads
(/root/reference/src/ITHACA_CORE/ITHACAparallel/ITHACAparallel.C:43-74, used at :133-165 and :201-240):
  cellProcAddressing[local cell] = global cell label (0-based);
  faceProcAddressing[local face] = +-(global face + 1), negative when the local face is flipped with respect to the
  global one (decomposePar's processor-face convention: the local cell owns the face, so a face whose global owner
  lives elsewhere is turned round).
Ownership is the contiguous-range rule of SURVEY 8(e): rank g of G owns cells [floor(g N/G), floor((g+1) N/G)).
"""
import numpy as np


def cell_range(nCells, rank, nranks):
    return (rank * nCells) // nranks, ((rank + 1) * nCells) // nranks


def cuthill_mckee(nCells, owner, neighbour, nInt):
    """Restatement of the product's locality ordering (ithaca-fv_b200/csrc/renumber.cu: mesh_renumber), pure Python
    (small meshes only). Returns old2new. Rule: for every connected component, in order of its lowest unvisited label,
    a breadth-first search from that label finds a pseudo-peripheral cell (the LAST cell it reaches); the numbering is
    the breadth-first order from there, the newly reached neighbours of a cell appended by (degree, label)."""
    own = np.asarray(owner, dtype=np.int64)[:nInt]
    nei = np.asarray(neighbour, dtype=np.int64)[:nInt]
    adj = [[] for _ in range(nCells)]
    for o, n in zip(own.tolist(), nei.tolist()):   # adjacency in face order (only its content matters: lists are sorted)
        adj[o].append(n)
        adj[n].append(o)
    deg = [len(a) for a in adj]

    def bfs(start, mark, out):
        out.append(start)
        mark[start] = True
        head = len(out) - 1
        while head < len(out):
            c = out[head]
            head += 1
            nb = []
            for x in adj[c]:
                if not mark[x]:
                    mark[x] = True
                    nb.append(x)
            nb.sort(key=lambda x: (deg[x], x))
            out.extend(nb)

    seen = [False] * nCells
    order = []
    for c0 in range(nCells):
        if seen[c0]:
            continue
        probe_mark, probe = list(seen), []
        bfs(c0, probe_mark, probe)
        bfs(probe[-1], seen, order)
    old2new = np.zeros(nCells, dtype=np.int32)
    old2new[np.asarray(order, dtype=np.int64)] = np.arange(nCells, dtype=np.int32)
    return old2new


def partition(nCells, owner, neighbour, nInt, rank, nranks, halo_layers=2, old2new=None):
    """Returns dict(cpa, fpa, nOwned, nLocalInt, nLocalBnd, nCut). owner: nInt+nBnd labels, neighbour: nInt labels.
    old2new = None: owned cells are the label range (IB200_PART_LABEL_RANGE); else the cells whose position in that
    ordering falls in the range (IB200_PART_LOCALITY_RANGE)."""
    owner = np.asarray(owner, dtype=np.int64)
    nei = np.asarray(neighbour, dtype=np.int64)[:nInt]
    own_i = owner[:nInt]
    c0, c1 = cell_range(nCells, rank, nranks)
    level = np.full(nCells, -1, dtype=np.int64)
    if old2new is None:
        level[c0:c1] = 0
    else:
        o2n = np.asarray(old2new, dtype=np.int64)
        level[(o2n >= c0) & (o2n < c1)] = 0
    for L in range(1, halo_layers + 1):
        # face neighbours of layer L-1 that are still outside (simultaneous update = breadth-first layer)
        grow_n = (level[own_i] == L - 1) & (level[nei] < 0)
        grow_o = (level[nei] == L - 1) & (level[own_i] < 0)
        new = np.unique(np.concatenate([nei[grow_n], own_i[grow_o]]))
        if new.size == 0:
            break
        level[new] = L
    halo = np.flatnonzero(level > 0)
    cpa = np.concatenate([np.flatnonzero(level == 0), halo]).astype(np.int32)
    inside = level >= 0
    lo, ln = inside[own_i], inside[nei]
    f_int = np.flatnonzero(lo & ln)
    f_bnd = nInt + np.flatnonzero(inside[owner[nInt:]])
    cut_idx = np.flatnonzero(lo ^ ln)
    cut = np.where(lo[cut_idx], cut_idx + 1, -(cut_idx + 1))
    fpa = np.concatenate([f_int + 1, f_bnd + 1, cut]).astype(np.int32)
    return dict(cpa=cpa, fpa=fpa, nOwned=c1 - c0, nLocalInt=int(f_int.size), nLocalBnd=int(f_bnd.size), nCut=int(cut.size))


class ShardMesh:
    """The shard as a mesh object with the attributes oracle/fvc.Oracle and the Projector binding use."""

    def __init__(self, mesh, part):
        cpa, fpa = part["cpa"].astype(np.int64), part["fpa"].astype(np.int64)
        nI, nB, nCut = part["nLocalInt"], part["nLocalBnd"], part["nCut"]
        g = np.abs(fpa) - 1
        sgn = np.sign(fpa).astype(np.float64)
        g2l = np.full(mesh.nCells, -1, dtype=np.int64)
        g2l[cpa] = np.arange(cpa.size)
        owner = np.asarray(mesh.owner, dtype=np.int64)
        nei = np.asarray(mesh.neighbour, dtype=np.int64)
        gi, gb, gc = g[:nI], g[nI:nI + nB], g[nI + nB:]
        cut_cell = np.where(fpa[nI + nB:] > 0, owner[gc], nei[np.minimum(gc, max(mesh.nInt - 1, 0))]) if nCut else np.zeros(0, np.int64)
        self.nCells, self.nInt, self.nBnd = int(cpa.size), nI, nB + nCut
        self.owner = np.concatenate([g2l[owner[gi]], g2l[owner[gb]], g2l[cut_cell]]).astype(np.int32)
        self.neighbour = g2l[nei[gi]].astype(np.int32)
        Sf = np.asarray(mesh.Sf, dtype=np.float64).reshape(-1, 3)
        self.Sf = (Sf[g] * sgn[:, None]).copy()
        self.magSf = np.asarray(mesh.magSf, dtype=np.float64)[g].copy()
        self.w = np.asarray(mesh.w, dtype=np.float64)[gi].copy()
        self.dc = np.asarray(mesh.dc, dtype=np.float64)[gi].copy()
        self.corr = np.asarray(mesh.corr, dtype=np.float64).reshape(-1, 3)[gi].copy()
        self.dcB = np.concatenate([np.asarray(mesh.dcB, dtype=np.float64)[gb - mesh.nInt],
                                   np.asarray(mesh.dc, dtype=np.float64)[gc] if nCut else np.zeros(0)])
        self.bmask = np.concatenate([np.asarray(mesh.bmask, dtype=np.uint8)[gb - mesh.nInt], np.zeros(nCut, dtype=np.uint8)])
        self.V = np.asarray(mesh.V, dtype=np.float64)[cpa].copy()
        self.nOwned = int(part["nOwned"])
        if hasattr(mesh, "Cc"):   # centres for the cellMDLimited gradient (mesh_set_schemes), shard order
            self.Cc = np.asarray(mesh.Cc, dtype=np.float64)[cpa].copy()
            self.Cf = np.asarray(mesh.Cf, dtype=np.float64)[g].copy()
        self._cpa, self._gb, self._nCut, self._nInt_g = cpa, gb, nCut, mesh.nInt

    def cell_field(self, f, dim):
        return np.asarray(f, dtype=np.float64).reshape(-1, dim)[self._cpa].reshape(-1).copy()

    def bnd_field(self, f, dim):
        f = np.asarray(f, dtype=np.float64).reshape(-1, dim)
        return np.concatenate([f[self._gb - self._nInt_g], np.zeros((self._nCut, dim))]).reshape(-1).copy()
