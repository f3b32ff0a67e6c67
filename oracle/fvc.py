"""TEST INFRASTRUCTURE ONLY -- ctypes binding of oracle/libfvc_oracle.so (oracle/fvc_oracle.c) and the seeded
synthetic basis used by the projection tests and benchmark. Parity status: see fvc_oracle.c (UNPINNED at the
OpenFOAM boundary; the term-builder loop nests follow steadyNS.C:883-1137)."""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(_HERE, "libfvc_oracle.so")
_lib = None
P = ctypes.c_void_p


class _Mesh(ctypes.Structure):
    _fields_ = [("nCells", ctypes.c_longlong), ("nInt", ctypes.c_longlong), ("nBnd", ctypes.c_longlong),
                ("owner", P), ("neighbour", P), ("Sf", P), ("magSf", P), ("w", P), ("dc", P), ("dcB", P),
                ("corr", P), ("V", P), ("bmask", P), ("Cc", P), ("Cf", P), ("gradK", ctypes.c_double),
                ("snPsi", ctypes.c_double), ("convScheme", ctypes.c_int)]


def lib():
    global _lib
    if _lib is None:
        L = ctypes.CDLL(LIB)
        L.fvo_domain_integrate_dot.restype = ctypes.c_double
        _lib = L
    return _lib


def _c(a, dt=np.float64):
    return np.ascontiguousarray(a, dtype=dt)


class Oracle:
    """Holds the C mesh struct (and keeps the numpy arrays it points into alive)."""

    def __init__(self, mesh, corrected=True, grad_limit_k=0.0, sngrad_psi=-1.0, convection="linear"):
        """grad_limit_k > 0: gradSchemes 'cellMDLimited Gauss linear k'; sngrad_psi >= 0: snGrad 'limited psi'
        (tutorials/CFD/04unsteadyNS/system/fvSchemes); defaults = the 'Gauss linear' / 'corrected' set."""
        self.mesh = mesh
        self.keep = dict(owner=_c(mesh.owner, np.int32), neighbour=_c(mesh.neighbour, np.int32), Sf=_c(mesh.Sf),
                         magSf=_c(mesh.magSf), w=_c(mesh.w), dc=_c(mesh.dc), dcB=_c(mesh.dcB), corr=_c(mesh.corr),
                         V=_c(mesh.V), bmask=_c(mesh.bmask, np.uint8), Cc=_c(getattr(mesh, "Cc", np.zeros(3))), Cf=_c(getattr(mesh, "Cf", np.zeros(3))))
        k = self.keep
        assert not (grad_limit_k > 0 or convection == "linearUpwind") or (k["Cc"].size == 3 * mesh.nCells and k["Cf"].size == 3 * (mesh.nInt + mesh.nBnd))
        self.m = _Mesh(mesh.nCells, mesh.nInt, mesh.nBnd, k["owner"].ctypes.data, k["neighbour"].ctypes.data,
                       k["Sf"].ctypes.data, k["magSf"].ctypes.data, k["w"].ctypes.data, k["dc"].ctypes.data,
                       k["dcB"].ctypes.data, k["corr"].ctypes.data if corrected else None, k["V"].ctypes.data,
                       k["bmask"].ctypes.data, k["Cc"].ctypes.data, k["Cf"].ctypes.data, float(grad_limit_k),
                       float(sngrad_psi), {"linear": 0, "upwind": 1, "linearUpwind": 2}[convection])
        self.L = lib()

    def _ref(self):
        return ctypes.byref(self.m)

    @staticmethod
    def _ptrs(fields):
        fs = [_c(f) for f in fields]
        return fs, (P * len(fs))(*[f.ctypes.data for f in fs])

    # ---- operators ----
    def flux(self, U, Ub):
        U, Ub = _c(U), _c(Ub)
        out = np.zeros(self.mesh.nInt + self.mesh.nBnd)
        self.L.fvo_flux(self._ref(), P(U.ctypes.data), P(Ub.ctypes.data), P(out.ctypes.data))
        return out

    def div_phi(self, phi, U, Ub):
        phi, U, Ub = _c(phi), _c(U), _c(Ub)
        out = np.zeros(3 * self.mesh.nCells)
        self.L.fvo_div_phi_vec(self._ref(), P(phi.ctypes.data), P(U.ctypes.data), P(Ub.ctypes.data), P(out.ctypes.data))
        return out

    def div(self, U, Ub):
        U, Ub = _c(U), _c(Ub)
        out = np.zeros(self.mesh.nCells)
        self.L.fvo_div_vec(self._ref(), P(U.ctypes.data), P(Ub.ctypes.data), P(out.ctypes.data))
        return out

    def grad(self, p, pb):
        p, pb = _c(p), _c(pb)
        out = np.zeros(3 * self.mesh.nCells)
        self.L.fvo_grad_scalar(self._ref(), P(p.ctypes.data), P(pb.ctypes.data), P(out.ctypes.data))
        return out

    def grad_vec(self, U, Ub):
        U, Ub = _c(U), _c(Ub)
        out = np.zeros(9 * self.mesh.nCells)
        self.L.fvo_grad_vec(self._ref(), P(U.ctypes.data), P(Ub.ctypes.data), P(out.ctypes.data))
        return out

    def laplacian(self, U, Ub):
        U, Ub = _c(U), _c(Ub)
        out = np.zeros(3 * self.mesh.nCells)
        self.L.fvo_laplacian_vec(self._ref(), P(U.ctypes.data), P(Ub.ctypes.data), P(out.ctypes.data))
        return out

    # ---- term builders ----
    # threads > 1 (CPU-baseline timing only): the operator's outer loop is cut into slabs, one per host thread, through the
    # *_range entry points (ctypes releases the GIL); every entry is computed by the same serial arithmetic either way.
    @staticmethod
    def _slabs(fn, args, n, threads):
        import threading
        threads = max(1, min(threads, n))
        if threads == 1:
            fn(*args, 0, n)
            return
        cuts = [n * t // threads for t in range(threads + 1)]
        th = [threading.Thread(target=fn, args=(*args, cuts[t], cuts[t + 1])) for t in range(threads)]
        [x.start() for x in th]
        [x.join() for x in th]

    def mass_term(self, U):
        fs, pu = self._ptrs(U)
        N = len(fs)
        M = np.zeros((N, N), order="F")
        self.L.fvo_mass_term(self._ref(), N, pu, P(M.ctypes.data))
        return M

    def diffusive_term(self, U, Ub, threads=1):
        fs, pu = self._ptrs(U)
        fb, pb = self._ptrs(Ub)
        N = len(fs)
        B = np.zeros((N, N), order="F")
        self._slabs(self.L.fvo_diffusive_term_range, (self._ref(), N, pu, pb, P(B.ctypes.data)), N, threads)
        return B

    def pressure_gradient_term(self, U, Pm, Pb, threads=1):
        fs, pu = self._ptrs(U)
        fp, pp = self._ptrs(Pm)
        fb, pb = self._ptrs(Pb)
        K = np.zeros((len(fs), len(fp)), order="F")
        self._slabs(self.L.fvo_pressure_gradient_term_range, (self._ref(), len(fs), len(fp), pu, pp, pb, P(K.ctypes.data)), len(fp), threads)
        return K

    def divergence_term(self, U, Ub, Pm, threads=1):
        fs, pu = self._ptrs(U)
        fb, pb = self._ptrs(Ub)
        fp, pp = self._ptrs(Pm)
        Pmat = np.zeros((len(fp), len(fs)), order="F")
        self._slabs(self.L.fvo_divergence_term_range, (self._ref(), len(fs), len(fp), pu, pb, pp, P(Pmat.ctypes.data)), len(fs), threads)
        return Pmat

    def convective_term_tens(self, U, Ub, cache=True, threads=1):
        fs, pu = self._ptrs(U)
        fb, pb = self._ptrs(Ub)
        N = len(fs)
        C = np.zeros((N, N, N), order="F")
        fn = self.L.fvo_convective_term_tens_cache_range if cache else self.L.fvo_convective_term_tens_range
        self._slabs(fn, (self._ref(), N, pu, pb, P(C.ctypes.data)), N, threads)
        return C

    def convective_term_tens_consistent(self, U, Ub, Phi):
        """fluxMethod == "consistent" (steadyNS.C:1010-1015): C(i,j,k) = domainIntegrate(U_i & fvc::div(L_PHImodes[j], U_k)),
        the reference's loop nest with the stored flux fields Phi[j] (nInt+nBnd values, 0 on `empty` faces)."""
        N = len(U)
        C = np.zeros((N, N, N), order="F")
        V3 = np.repeat(_c(self.mesh.V), 3)
        for j in range(N):
            for k in range(N):
                D = self.div_phi(_c(Phi[j]), U[k], Ub[k])
                for i in range(N):
                    C[i, j, k] = np.dot(V3 * _c(U[i]), D)
        return C

    # ---- PPE operators (steadyNS::projectPPE) ----
    def laplacian_pressure(self, Pm, Pb):
        fp, pp = self._ptrs(Pm)
        fb, pb = self._ptrs(Pb)
        D = np.zeros((len(fp), len(fp)), order="F")
        self.L.fvo_laplacian_pressure(self._ref(), len(fp), pp, pb, P(D.ctypes.data))
        return D

    def div_momentum(self, U, Ub, Pm, Pb):
        fs, pu = self._ptrs(U)
        fb, pub = self._ptrs(Ub)
        fp, pp = self._ptrs(Pm)
        fpb, ppb = self._ptrs(Pb)
        G = np.zeros((len(fp), len(fs), len(fs)), order="F")
        self.L.fvo_div_momentum(self._ref(), len(fs), len(fp), pu, pub, pp, ppb, P(G.ctypes.data))
        return G

    def pressure_bc4(self, Ub, Pb):
        fb, pub = self._ptrs(Ub)
        fpb, ppb = self._ptrs(Pb)
        out = np.zeros((len(fpb), len(fb)), order="F")
        self.L.fvo_pressure_bc4(self._ref(), len(fb), len(fpb), pub, ppb, P(out.ctypes.data))
        return out

    def pressure_bc3(self, U, Ub, Pm, Pb):
        fs, pu = self._ptrs(U)
        fb, pub = self._ptrs(Ub)
        fp, pp = self._ptrs(Pm)
        fpb, ppb = self._ptrs(Pb)
        out = np.zeros((len(fp), len(fs)), order="F")
        self.L.fvo_pressure_bc3(self._ref(), len(fs), len(fp), pu, pub, pp, ppb, P(out.ctypes.data))
        return out

    def pressure_bc1(self, U, Ub, Pb):
        fs, pu = self._ptrs(U)
        fb, pub = self._ptrs(Ub)
        fpb, ppb = self._ptrs(Pb)
        out = np.zeros((len(fpb), len(fs)), order="F")
        self.L.fvo_pressure_bc1(self._ref(), len(fs), len(fpb), pu, pub, ppb, P(out.ctypes.data))
        return out

    def pressure_bc2(self, U, Ub, Pb):
        """pressureBC2 (steadyNS.C:1396-1437): Np x N x N, i fastest."""
        fs, pu = self._ptrs(U)
        fb, pub = self._ptrs(Ub)
        fpb, ppb = self._ptrs(Pb)
        out = np.zeros((len(fpb), len(fs), len(fs)), order="F")
        self.L.fvo_pressure_bc2(self._ref(), len(fs), len(fpb), pu, pub, ppb, P(out.ctypes.data))
        return out

    def diffusive_term_flux_method(self, U, Ub, Pm):
        """BP (steadyNS.C:1584-1610): Np x N."""
        fs, pu = self._ptrs(U)
        fb, pub = self._ptrs(Ub)
        fp, pp = self._ptrs(Pm)
        out = np.zeros((len(fp), len(fs)), order="F")
        self.L.fvo_diffusive_term_flux_method(self._ref(), len(fs), len(fp), pu, pub, pp, P(out.ctypes.data))
        return out

    def convective_term_flux_tens(self, U, Ub, Pm):
        """Cf (steadyNS.C:1704-1746, inconsistent flux method): Np x N x N, i fastest."""
        fs, pu = self._ptrs(U)
        fb, pub = self._ptrs(Ub)
        fp, pp = self._ptrs(Pm)
        out = np.zeros((len(fp), len(fs), len(fs)), order="F")
        self.L.fvo_convective_term_flux_tens(self._ref(), len(fs), len(fp), pu, pub, pp, P(out.ctypes.data))
        return out

    def convective_background(self, U, Ub, vls, vlsb, Pm=None, Pb=None):
        """Pm None: L (N x N, steadyNS.C:1256-1273); else L_D (Np x N, :1275-1293)."""
        fs, pu = self._ptrs(U)
        fb, pub = self._ptrs(Ub)
        vls, vlsb = _c(vls), _c(vlsb)
        kind = 0 if Pm is None else 1
        if kind:
            fp, pp = self._ptrs(Pm)
            fpb, ppb = self._ptrs(Pb)
            rows, npm = len(fp), len(fp)
        else:
            pp = ppb = None
            rows, npm = len(fs), 0
        out = np.zeros((rows, len(fs)), order="F")
        self.L.fvo_convective_background(self._ref(), kind, len(fs), npm, pu, pub, pp, ppb, P(vls.ctypes.data), P(vlsb.ctypes.data),
                                         P(out.ctypes.data))
        return out

    def cell_faces(self):
        nnz = 2 * self.mesh.nInt + self.mesh.nBnd
        off = np.zeros(self.mesh.nCells + 1, dtype=np.int32)
        ent = np.zeros(nnz, dtype=np.int32)
        other = np.zeros(nnz, dtype=np.int32)
        self.L.fvo_cell_faces(self._ref(), P(off.ctypes.data), P(ent.ctypes.data), P(other.ctypes.data))
        return off, ent, other


def synthetic_basis(mesh, N, Np, seed=0, noise=1e-3):
    """N vector fields + Np scalar fields: smooth functions of the cell / boundary-face centres plus noise;
    boundary values = the same functions evaluated at the face centres (SURVEY.md section 8(d), C4 inputs)."""
    rng = np.random.default_rng(seed)
    nI = mesh.nInt
    Cc, Cb = mesh.Cc, mesh.Cf[nI:]
    L = np.abs(np.concatenate([Cc, Cb])).max() + 1e-30

    def vec(k, X, nz):
        x, y, z = X[:, 0] / L, X[:, 1] / L, X[:, 2] / L
        out = np.stack([np.sin((k + 1) * x + 0.3) * np.cos((k + 2) * y) + 0.2 * z,
                        np.cos((k + 1) * y - 0.1) * np.sin((k + 3) * x) - 0.1 * z * x,
                        0.3 * np.sin((k + 1) * (x + y + z))], axis=1)
        return np.ascontiguousarray((out + nz).ravel())

    U, Ub, Pm, Pb = [], [], [], []
    for k in range(N):
        U.append(vec(k, Cc, noise * rng.standard_normal((mesh.nCells, 3))))
        Ub.append(vec(k, Cb, noise * rng.standard_normal((mesh.nBnd, 3))))
    for k in range(Np):
        f = lambda X: np.cos((k + 1) * X[:, 0] / L) * np.sin((k + 2) * X[:, 1] / L + 0.5) + 0.1 * X[:, 2] / L
        Pm.append(f(Cc) + noise * rng.standard_normal(mesh.nCells))
        Pb.append(f(Cb) + noise * rng.standard_normal(mesh.nBnd))
    return U, Ub, Pm, Pb
