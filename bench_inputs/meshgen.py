"""BENCHMARK / TEST INPUT GENERATOR -- finite-volume mesh arrays without OpenFOAM.

Neither the product nor the oracle's arithmetic lives here: this module only FABRICATES INPUTS (the arrays OpenFOAM itself hands
to ithaca_b200_mesh_upload in production: owner / neighbour / Sf / weights / deltaCoeffs / V ...) for the benchmark's C4 mesh
 and for the parity tests. bench.py / bench_proj.py import it; oracle/polymesh.py (the ASCII polyMesh
reader used by the tests) builds on the same FvMesh.

The geometry formulas restate OpenFOAM v2412's published algorithms (OpenFOAM is absent from /root/reference):
  face centres / area vectors : primitiveMeshTools::faceCentresAndAreas (triangle fan about the vertex average)
  cell centres / volumes      : primitiveMeshTools::cellCentresAndVols (pyramids about the face-centre average)
  linear weights              : surfaceInterpolation::makeWeights      w = |Sf.(Cn-Cf)| / (|Sf.(Cf-Co)| + |Sf.(Cn-Cf)|)
  nonOrthDeltaCoeffs          : surfaceInterpolation::makeNonOrthDeltaCoeffs   1/max(nf.d, 0.05|d|), d = Cn-Co
  nonOrthCorrectionVectors    : surfaceInterpolation::makeNonOrthCorrectionVectors   nf - d*nonOrthDeltaCoeffs
  boundary deltaCoeffs        : fvPatch::delta() = nf*(nf.(Cf-Cc)) for non-coupled patches, coefficient 1/|delta|
Known answers used to check them: tests/test_fvc_oracle.py (uniform cavity V = 1e-8, closed cells sum Sf = 0,
Gauss theorem, exact gradients of linear fields on orthogonal meshes).
"""
import numpy as np


def faces_to_padded(faces):
    """list of vertex lists -> (F, maxv) int array padded with -1, and vertex counts."""
    nv = np.array([len(f) for f in faces])
    A = -np.ones((len(faces), nv.max()), dtype=np.int64)
    for k in np.unique(nv):
        idx = np.where(nv == k)[0]
        A[idx, :k] = np.array([faces[i] for i in idx])
    return A, nv


# ------------------------------------------------------------------------------------------------
# geometry
# ------------------------------------------------------------------------------------------------
def face_centres_areas(points, F, nv):
    nF = F.shape[0]
    Cf = np.zeros((nF, 3))
    Sf = np.zeros((nF, 3))
    for k in np.unique(nv):
        idx = np.where(nv == k)[0]
        P = points[F[idx, :k]]  # (n, k, 3)
        if k == 3:
            Cf[idx] = P.sum(1) / 3.0
            Sf[idx] = 0.5 * np.cross(P[:, 1] - P[:, 0], P[:, 2] - P[:, 0])
            continue
        fc = P.sum(1) / k
        sumN = np.zeros((len(idx), 3))
        sumA = np.zeros(len(idx))
        sumAc = np.zeros((len(idx), 3))
        for pi in range(k):
            this, nxt = P[:, pi], P[:, (pi + 1) % k]
            c = this + nxt + fc
            n = np.cross(nxt - this, fc - this)
            a = np.linalg.norm(n, axis=1)
            sumN += n
            sumA += a
            sumAc += a[:, None] * c
        Cf[idx] = (1.0 / 3.0) * sumAc / sumA[:, None]
        Sf[idx] = 0.5 * sumN
    return Cf, Sf


def cell_centres_volumes(Cf, Sf, owner, neighbour, nCells):
    nI = len(neighbour)
    cEst = np.zeros((nCells, 3))
    nf = np.zeros(nCells)
    np.add.at(cEst, owner, Cf)
    np.add.at(nf, owner, 1)
    np.add.at(cEst, neighbour, Cf[:nI])
    np.add.at(nf, neighbour, 1)
    cEst /= nf[:, None]
    Cc = np.zeros((nCells, 3))
    V = np.zeros(nCells)
    pyr = np.einsum("ij,ij->i", Sf, Cf - cEst[owner])
    pc = 0.75 * Cf + 0.25 * cEst[owner]
    np.add.at(Cc, owner, pyr[:, None] * pc)
    np.add.at(V, owner, pyr)
    pyr = np.einsum("ij,ij->i", Sf[:nI], cEst[neighbour] - Cf[:nI])
    pc = 0.75 * Cf[:nI] + 0.25 * cEst[neighbour]
    np.add.at(Cc, neighbour, pyr[:, None] * pc)
    np.add.at(V, neighbour, pyr)
    Cc /= V[:, None]
    V *= 1.0 / 3.0
    return Cc, V


class FvMesh:
    """The arrays ithaca_b200_mesh_upload takes, computed from a polyMesh (int32 labels as in the shipped
    OpenFOAM builds, tutorials/CFD/01POD/log.compare_POD_par:9 'label=32')."""

    def __init__(self, points, F, nv, owner, neighbour, patches):
        self.nCells = int(max(owner.max(), neighbour.max() if len(neighbour) else 0)) + 1
        self.nInt = nI = len(neighbour)
        self.nBnd = len(owner) - nI
        self.owner = owner.astype(np.int32)
        self.neighbour = neighbour.astype(np.int32)
        self.patches = patches  # (name, type, nFaces, startFace)
        Cf, Sf = face_centres_areas(points, F, nv)
        Cc, V = cell_centres_volumes(Cf, Sf, owner, neighbour, self.nCells)
        self.Cf, self.Sf, self.Cc, self.V = Cf, np.ascontiguousarray(Sf), Cc, V
        self.magSf = np.linalg.norm(Sf, axis=1)
        o, n = owner[:nI], neighbour
        SfdOwn = np.abs(np.einsum("ij,ij->i", Sf[:nI], Cf[:nI] - Cc[o]))
        SfdNei = np.abs(np.einsum("ij,ij->i", Sf[:nI], Cc[n] - Cf[:nI]))
        self.w = SfdNei / (SfdOwn + SfdNei)
        d = Cc[n] - Cc[o]
        nf = Sf[:nI] / self.magSf[:nI, None]
        self.dc = 1.0 / np.maximum(np.einsum("ij,ij->i", nf, d), 0.05 * np.linalg.norm(d, axis=1))
        self.corr = np.ascontiguousarray(nf - d * self.dc[:, None])
        nfb = Sf[nI:] / self.magSf[nI:, None]
        db = nfb * np.einsum("ij,ij->i", nfb, Cf[nI:] - Cc[owner[nI:]])[:, None]
        self.dcB = 1.0 / np.linalg.norm(db, axis=1)
        # faces of `empty` patches carry no field values in OpenFOAM: they never contribute
        self.bmask = np.ones(self.nBnd, dtype=np.uint8)
        for (_, typ, nF, start) in patches:
            if typ == "empty":
                self.bmask[start - nI:start - nI + nF] = 0

    def patch_slices(self):
        return [(name, typ, slice(start - self.nInt, start - self.nInt + nF)) for (name, typ, nF, start) in self.patches]


# ------------------------------------------------------------------------------------------------
# synthetic hexahedral meshes: structured (i,j,k) blocks mapped to a box or
# to a cylinder O-grid, vertices perturbed (non-orthogonality), cells randomly relabelled, faces re-sorted
# to OpenFOAM's upper-triangular order => genuinely unstructured owner/neighbour addressing.
# ------------------------------------------------------------------------------------------------
def synthetic_hex(ni, nj, nk, kind="box", seed=0, perturb=0.15, relabel=True, periodic_j=None):
    """kind 'box': unit box, patches xmin,xmax,ymin,ymax,zmin,zmax.
    kind 'cylinder': O-grid, i = radial (graded), j = angular (periodic), k = axial; patches cylinder, farfield,
    front, back. Returns FvMesh."""
    rng = np.random.default_rng(seed)
    per = (kind == "cylinder") if periodic_j is None else periodic_j
    npj = nj if per else nj + 1
    I, J, K = np.meshgrid(np.arange(ni + 1), np.arange(npj), np.arange(nk + 1), indexing="ij")
    u, v, w = I / ni, J / nj, K / nk
    jit = perturb * (rng.random((3,) + I.shape) - 0.5)
    inner = (I > 0) & (I < ni)  # keep the boundary surfaces smooth in i
    u = u + np.where(inner, jit[0] / ni, 0.0)
    v = v + (jit[1] / nj if per else np.where((J > 0) & (J < nj), jit[1] / nj, 0.0))
    w = w + np.where((K > 0) & (K < nk), jit[2] / nk, 0.0)
    if kind == "box":
        X, Y, Z = u, v, w
    else:
        r = 0.5 * (20.0 ** u)  # geometric radial grading, cylinder radius 0.5, farfield radius 10
        th = 2 * np.pi * v
        X, Y, Z = r * np.cos(th), r * np.sin(th), 2.0 * w
    points = np.stack([X.ravel(), Y.ravel(), Z.ravel()], axis=1)

    def pid(i, j, k):
        return (i * npj + (j % npj if per else j)) * (nk + 1) + k

    def cid(i, j, k):
        return (i * nj + j) * nk + k

    quads, ca, cb, pat = [], [], [], []

    def add(i0, j0, k0, quad, a, b, patch):
        quads.append(np.stack([q.ravel() for q in quad], axis=1))
        ca.append(a.ravel())
        cb.append(b.ravel() if b is not None else np.full(a.size, -1))
        pat.append(np.full(a.size, patch))

    # i-faces (normal +i): vertices (i; j,k),(j+1,k),(j+1,k+1),(j,k+1)
    i, j, k = np.meshgrid(np.arange(ni + 1), np.arange(nj), np.arange(nk), indexing="ij")
    q = [pid(i, j, k), pid(i, j + 1, k), pid(i, j + 1, k + 1), pid(i, j, k + 1)]
    lo, hi = i == 0, i == ni
    mid = ~lo & ~hi
    add(0, 0, 0, [x[mid] for x in q], cid(i - 1, j, k)[mid], cid(i, j, k)[mid], -1)
    add(0, 0, 0, [x[lo] for x in q[::-1]], cid(i, j, k)[lo], None, 0)
    add(0, 0, 0, [x[hi] for x in q], cid(i - 1, j, k)[hi], None, 1)
    # j-faces (normal +j): vertices (j; k,i) order (i,k),(i,k+1),(i+1,k+1),(i+1,k)
    i, j, k = np.meshgrid(np.arange(ni), np.arange(nj if per else nj + 1), np.arange(nk), indexing="ij")
    q = [pid(i, j, k), pid(i, j, k + 1), pid(i + 1, j, k + 1), pid(i + 1, j, k)]
    if per:
        add(0, 0, 0, q, cid(i, (j - 1) % nj, k), cid(i, j, k), -1)
    else:
        lo, hi = j == 0, j == nj
        mid = ~lo & ~hi
        add(0, 0, 0, [x[mid] for x in q], cid(i, j - 1, k)[mid], cid(i, j, k)[mid], -1)
        add(0, 0, 0, [x[lo] for x in q[::-1]], cid(i, j, k)[lo], None, 2)
        add(0, 0, 0, [x[hi] for x in q], cid(i, j - 1, k)[hi], None, 3)
    # k-faces (normal +k): (i,j),(i+1,j),(i+1,j+1),(i,j+1)
    i, j, k = np.meshgrid(np.arange(ni), np.arange(nj), np.arange(nk + 1), indexing="ij")
    q = [pid(i, j, k), pid(i + 1, j, k), pid(i + 1, j + 1, k), pid(i, j + 1, k)]
    lo, hi = k == 0, k == nk
    mid = ~lo & ~hi
    add(0, 0, 0, [x[mid] for x in q], cid(i, j, k - 1)[mid], cid(i, j, k)[mid], -1)
    add(0, 0, 0, [x[lo] for x in q[::-1]], cid(i, j, k)[lo], None, 4)
    add(0, 0, 0, [x[hi] for x in q], cid(i, j, k - 1)[hi], None, 5)

    Q = np.concatenate(quads)
    a = np.concatenate(ca)
    b = np.concatenate(cb)
    p = np.concatenate(pat)
    nC = ni * nj * nk
    perm = rng.permutation(nC) if relabel else np.arange(nC)
    a = perm[a]
    inter = b >= 0
    b = np.where(inter, perm[np.where(inter, b, 0)], -1)
    swap = inter & (a > b)
    a2 = np.where(swap, b, a)
    b2 = np.where(swap, a, b)
    Q[swap] = Q[swap][:, ::-1]
    ii = np.where(inter)[0]
    ii = ii[np.lexsort((b2[ii], a2[ii]))]          # upper-triangular order: by owner, then neighbour
    bi = np.where(~inter)[0]
    bi = bi[np.lexsort((a2[bi], p[bi]))]           # boundary: by patch, then owner
    order = np.concatenate([ii, bi])
    Q, own, nei = Q[order], a2[order], b2[ii]
    if kind == "box":
        names = ["xmin", "xmax", "ymin", "ymax", "zmin", "zmax"]
    else:
        names = ["cylinder", "farfield", "jmin", "jmax", "front", "back"]
    patches, start = [], len(ii)
    pb = p[bi]
    for pi, nm in enumerate(names):
        n = int((pb == pi).sum())
        if n:
            patches.append((nm, "wall" if nm == "cylinder" else "patch", n, start))
            start += n
    nv = np.full(len(Q), 4, dtype=np.int64)
    return FvMesh(points, Q.astype(np.int64), nv, own.astype(np.int64), nei.astype(np.int64), patches)
