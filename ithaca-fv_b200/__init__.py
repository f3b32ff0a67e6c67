"""ithaca-fv_b200 -- ctypes binding of libithaca_b200.so (include/ithaca_b200.h) for the test and
benchmark harnesses, plus a host-side mirror of ITHACAPOD::getModes' argument handling.

The product is the CUDA library; this module only marshals pointers. There is NO CPU fallback: if the
library is missing, or no sm_100 device is present, every compute call raises.

Reference semantics mirrored here (paths relative to /root/reference/src/ITHACA_CORE):
  ITHACAPOD/ITHACAPOD.C:92-323   getModes (nmodes rules, solver-dependent eigenvalue handling, norms)
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libithaca_b200.so")

# every symbol include/ithaca_b200.h declares (checked by tests/test_abi.py)
SYMBOLS = [
    "ithaca_b200_create", "ithaca_b200_destroy", "ithaca_b200_last_error", "ithaca_b200_stream",
    "ithaca_b200_launch_count", "ithaca_b200_timings", "ithaca_b200_nccl_unique_id",
    "ithaca_b200_comm_init", "ithaca_b200_cell_range", "ithaca_b200_pod_set_volumes",
    "ithaca_b200_pod_upload", "ithaca_b200_pod_upload_second", "ithaca_b200_pod_gram", "ithaca_b200_pod_eig", "ithaca_b200_pod_set_gram",
    "ithaca_b200_pod_postprocess_eigenvalues", "ithaca_b200_pod_lift", "ithaca_b200_pod_lift_patch",
    "ithaca_b200_dmma_peak",
    "ithaca_b200_pod_set_modes", "ithaca_b200_pod_project", "ithaca_b200_pod_modes_mass", "ithaca_b200_pod_get_coeffs",
    "ithaca_b200_pod_reconstruct", "ithaca_b200_pod_lift_scaled", "ithaca_b200_pod_gram_volumes",
    "ithaca_b200_pod_block_begin", "ithaca_b200_pod_block_upload", "ithaca_b200_pod_block_gram",
    "ithaca_b200_pod_block_gram_finish", "ithaca_b200_pod_block_gram_get", "ithaca_b200_pod_block_gram_set", "ithaca_b200_pod_block_lift", "ithaca_b200_pod_block_lift_finish",
    "ithaca_b200_pod_stream_begin", "ithaca_b200_pod_stream_upload", "ithaca_b200_pod_stream_slot", "ithaca_b200_pod_stream_gram_add",
    "ithaca_b200_pod_stream_gram_finish", "ithaca_b200_pod_stream_lift_add", "ithaca_b200_pod_stream_lift_finish",
    "ithaca_b200_synth_snapshots", "ithaca_b200_synth_volumes", "ithaca_b200_col_norms_noalloc",
    "ithaca_b200_mesh_upload", "ithaca_b200_mesh_addressing", "ithaca_b200_basis_upload", "ithaca_b200_flux_upload", "ithaca_b200_project",
    "ithaca_b200_fvc_eval", "ithaca_b200_proj_timings", "ithaca_b200_project_background", "ithaca_b200_project_bc_velocity",
    "ithaca_b200_partition_create", "ithaca_b200_partition_destroy", "ithaca_b200_partition_sizes",
    "ithaca_b200_partition_addressing", "ithaca_b200_partition_gather", "ithaca_b200_mesh_upload_shard",
    "ithaca_b200_mesh_set_owned", "ithaca_b200_mesh_renumbering",
    "ithaca_b200_mesh_set_schemes", "ithaca_b200_mesh_set_convection",
    "ithaca_b200_save_dense_matrix", "ithaca_b200_read_dense_matrix", "ithaca_b200_save_dense_tensor",
    "ithaca_b200_read_dense_tensor", "ithaca_b200_save_npy", "ithaca_b200_export_matrix", "ithaca_b200_export_tensor",
]

_lib = None


class IthacaB200Error(RuntimeError):
    pass


def lib():
    """Load libithaca_b200.so; raises (never falls back) when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise IthacaB200Error(
                "libithaca_b200.so is not built (run `python -c 'import __graft_entry__ as g; g.build()'`); "
                "there is no CPU fallback")
        L = ctypes.CDLL(LIB_PATH)
        P, I, L64, D = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_double
        L.ithaca_b200_create.argtypes = [ctypes.POINTER(P), I]
        L.ithaca_b200_destroy.argtypes = [P]
        L.ithaca_b200_last_error.argtypes = [P]
        L.ithaca_b200_last_error.restype = ctypes.c_char_p
        L.ithaca_b200_stream.argtypes = [P]
        L.ithaca_b200_stream.restype = P
        L.ithaca_b200_launch_count.argtypes = [P, I]
        L.ithaca_b200_launch_count.restype = L64
        L.ithaca_b200_timings.argtypes = [P, P, I]
        L.ithaca_b200_nccl_unique_id.argtypes = [P]
        L.ithaca_b200_comm_init.argtypes = [P, P, I, I]
        L.ithaca_b200_cell_range.argtypes = [L64, I, I, ctypes.POINTER(L64), ctypes.POINTER(L64)]
        L.ithaca_b200_pod_set_volumes.argtypes = [P, P, L64]
        L.ithaca_b200_pod_upload.argtypes = [P, P, I, I, L64]
        L.ithaca_b200_pod_upload_second.argtypes = [P, P, I, L64]
        L.ithaca_b200_pod_gram.argtypes = [P, P]
        L.ithaca_b200_pod_gram_volumes.argtypes = [P, P, P]
        L.ithaca_b200_pod_eig.argtypes = [P, I, P, P, ctypes.POINTER(I)]
        L.ithaca_b200_pod_set_gram.argtypes = [P, P, I]
        L.ithaca_b200_pod_postprocess_eigenvalues.argtypes = [P, I, I, I, P, P, ctypes.POINTER(I)]
        L.ithaca_b200_pod_lift.argtypes = [P, I, I, P, P]
        L.ithaca_b200_pod_lift_patch.argtypes = [P, P, L64, P]
        L.ithaca_b200_dmma_peak.argtypes = [I, ctypes.POINTER(D)]
        L.ithaca_b200_pod_set_modes.argtypes = [P, P, I]
        L.ithaca_b200_pod_lift_scaled.argtypes = [P, I, P, P]
        L.ithaca_b200_pod_block_begin.argtypes = [P, I, I, L64]
        L.ithaca_b200_pod_block_upload.argtypes = [P, I, P, I, I]
        L.ithaca_b200_pod_block_gram.argtypes = [P, I, I, I]
        L.ithaca_b200_pod_block_gram_finish.argtypes = [P, P]
        L.ithaca_b200_pod_block_gram_get.argtypes = [P, P]
        L.ithaca_b200_pod_block_gram_set.argtypes = [P, P]
        L.ithaca_b200_pod_block_lift.argtypes = [P, I, I]
        L.ithaca_b200_pod_block_lift_finish.argtypes = [P, I, I, P, P]
        L.ithaca_b200_pod_stream_begin.argtypes = [P, I, I, L64]
        L.ithaca_b200_pod_stream_upload.argtypes = [P, I, L64, L64, P, P]
        L.ithaca_b200_pod_stream_slot.argtypes = [P, I, L64, L64, ctypes.POINTER(P), ctypes.POINTER(L64), ctypes.POINTER(P)]
        L.ithaca_b200_pod_stream_gram_add.argtypes = [P, I]
        L.ithaca_b200_pod_stream_gram_finish.argtypes = [P, P]
        L.ithaca_b200_pod_stream_lift_add.argtypes = [P, I, I, I]
        L.ithaca_b200_pod_stream_lift_finish.argtypes = [P, I, P, P]
        L.ithaca_b200_synth_snapshots.argtypes = [P, P, L64, L64, L64, L64, I, P, I, D, ctypes.c_uint64]
        L.ithaca_b200_synth_volumes.argtypes = [P, P, L64, L64, L64, ctypes.c_uint64]
        L.ithaca_b200_col_norms_noalloc.argtypes = [P, P, L64, L64, I, P, I, I, P, P]
        L.ithaca_b200_pod_project.argtypes = [P, I, I, P]
        L.ithaca_b200_pod_modes_mass.argtypes = [P, I, P]
        L.ithaca_b200_pod_get_coeffs.argtypes = [P, I, P]
        L.ithaca_b200_pod_reconstruct.argtypes = [P, I, P, I, I, P]
        L.ithaca_b200_mesh_upload.argtypes = [P, L64, L64, L64] + [P] * 10
        L.ithaca_b200_mesh_addressing.argtypes = [P, P, P, P, ctypes.POINTER(ctypes.c_uint64)]
        L.ithaca_b200_basis_upload.argtypes = [P, I, P, P, I]
        L.ithaca_b200_project.argtypes = [P, I, P]
        L.ithaca_b200_flux_upload.argtypes = [P, P, I]
        L.ithaca_b200_fvc_eval.argtypes = [P, I, I, I, P]
        L.ithaca_b200_proj_timings.argtypes = [P, P, I]
        L.ithaca_b200_project_background.argtypes = [P, I, P, P, P]
        L.ithaca_b200_project_bc_velocity.argtypes = [P, L64, L64, I, P, P]
        L.ithaca_b200_partition_create.argtypes = [ctypes.POINTER(P), L64, L64, L64, P, P, I, I, I, I]
        L.ithaca_b200_mesh_renumbering.argtypes = [L64, L64, P, P, P]
        L.ithaca_b200_partition_destroy.argtypes = [P]
        L.ithaca_b200_partition_sizes.argtypes = [P, P]
        L.ithaca_b200_partition_addressing.argtypes = [P, P, P]
        L.ithaca_b200_partition_gather.argtypes = [P, I, I, P, P]
        L.ithaca_b200_mesh_upload_shard.argtypes = [P, P] + [P] * 10
        L.ithaca_b200_mesh_set_owned.argtypes = [P, L64]
        L.ithaca_b200_mesh_set_schemes.argtypes = [P, D, D, P, P]
        L.ithaca_b200_mesh_set_convection.argtypes = [P, I]
        C = ctypes.c_char_p
        L.ithaca_b200_save_dense_matrix.argtypes = [C, P, L64, L64]
        L.ithaca_b200_read_dense_matrix.argtypes = [C, ctypes.POINTER(L64), ctypes.POINTER(L64), P, L64]
        L.ithaca_b200_save_dense_tensor.argtypes = [C, P, L64, L64, L64]
        L.ithaca_b200_read_dense_tensor.argtypes = [C, P, P, L64]
        L.ithaca_b200_save_npy.argtypes = [C, P, I, P]
        L.ithaca_b200_export_matrix.argtypes = [C, C, I, P, L64, L64]
        L.ithaca_b200_export_tensor.argtypes = [C, C, I, P, L64, L64, L64]
        _lib = L
    return _lib


def cell_range(nCells, rank, nranks):
    b, e = ctypes.c_int64(), ctypes.c_int64()
    rc = lib().ithaca_b200_cell_range(nCells, rank, nranks, ctypes.byref(b), ctypes.byref(e))
    if rc:
        raise IthacaB200Error("cell_range: bad argument")
    return b.value, e.value


def postprocess_eigenvalues(evals_all, nmodes, solver):
    """Host arithmetic of ITHACAPOD.C:169/179, :182-186, :281-287 (through the C ABI)."""
    ev = np.ascontiguousarray(evals_all, dtype=np.float64)
    m = len(ev)
    out = np.zeros(m)
    cum = np.zeros(m)
    n = ctypes.c_int(0)
    rc = lib().ithaca_b200_pod_postprocess_eigenvalues(ev.ctypes.data, m, nmodes, 0 if solver == "spectra" else 1,
                                                       out.ctypes.data, cum.ctypes.data, ctypes.byref(n))
    if rc:
        raise IthacaB200Error("postprocess_eigenvalues: bad argument")
    return out[:n.value].copy(), cum[:n.value].copy()


def dmma_peak(device=0):
    v = ctypes.c_double(0)
    rc = lib().ithaca_b200_dmma_peak(device, ctypes.byref(v))
    if rc:
        raise IthacaB200Error("dmma_peak failed (no sm_100 device?)")
    return v.value


def _io_chk(rc):
    if rc:
        raise IthacaB200Error(lib().ithaca_b200_last_error(None).decode())


def _eigen_storage(A):
    """The bytes an Eigen::MatrixXd / Eigen::Tensor<double,3> of this shape holds: first index fastest."""
    return np.asfortranarray(np.asarray(A, dtype=np.float64))


def SaveDenseMatrix(A, folder, name):
    """ITHACAstream::SaveDenseMatrix (ITHACAstream.H:722-737): folder + name, folder created."""
    A = _eigen_storage(np.atleast_2d(A))
    os.makedirs(folder, exist_ok=True)
    _io_chk(lib().ithaca_b200_save_dense_matrix((folder + name).encode(), A.ctypes.data, A.shape[0], A.shape[1]))


def ReadDenseMatrix(folder, name):
    """ITHACAstream::ReadDenseMatrix (ITHACAstream.H:739-762)."""
    path = (folder + name).encode()
    r, c = ctypes.c_int64(), ctypes.c_int64()
    _io_chk(lib().ithaca_b200_read_dense_matrix(path, ctypes.byref(r), ctypes.byref(c), None, 0))
    A = np.zeros((r.value, c.value), order="F")
    _io_chk(lib().ithaca_b200_read_dense_matrix(path, ctypes.byref(r), ctypes.byref(c), A.ctypes.data, A.size))
    return A


def SaveDenseTensor(T, folder, name):
    """ITHACAstream::SaveDenseTensor (ITHACAstream.H:765-784); like the reference it does not create the folder."""
    T = _eigen_storage(T)
    assert T.ndim == 3
    _io_chk(lib().ithaca_b200_save_dense_tensor((folder + name).encode(), T.ctypes.data, *T.shape))


def ReadDenseTensor(folder, name):
    """ITHACAstream::ReadDenseTensor (ITHACAstream.H:786-809)."""
    path = (folder + name).encode()
    d = np.zeros(3, dtype=np.int64)
    _io_chk(lib().ithaca_b200_read_dense_tensor(path, d.ctypes.data, None, 0))
    T = np.zeros(tuple(int(x) for x in d), order="F")
    _io_chk(lib().ithaca_b200_read_dense_tensor(path, d.ctypes.data, T.ctypes.data, T.size))
    return T


def save_npy(A, fname):
    """cnpy::save of an Eigen vector / matrix / rank-3 tensor (cnpy.C:524-568)."""
    A = _eigen_storage(A)
    dims = np.array(A.shape, dtype=np.int64)
    _io_chk(lib().ithaca_b200_save_npy(fname.encode(), A.ctypes.data, A.ndim, dims.ctypes.data))


_EXPORT = {"python": 0, "matlab": 1, "eigen": 2}


def exportMatrix(A, name, type, folder):
    """ITHACAstream::exportMatrix (ITHACAstream.C:54-147)."""
    if type not in _EXPORT:
        raise IthacaB200Error('The extension "%s" was not implemented. Check the list of possible extensions.' % type)
    A = _eigen_storage(np.atleast_2d(A))
    _io_chk(lib().ithaca_b200_export_matrix(folder.encode(), name.encode(), _EXPORT[type], A.ctypes.data, *A.shape))


def exportTensor(T, name, type, folder):
    """ITHACAstream::exportTensor (ITHACAstream.C:263-358)."""
    if type not in _EXPORT:
        raise IthacaB200Error('The extension "%s" was not implemented. Check the list of possible extensions.' % type)
    T = _eigen_storage(T)
    assert T.ndim == 3
    _io_chk(lib().ithaca_b200_export_tensor(folder.encode(), name.encode(), _EXPORT[type], T.ctypes.data, *T.shape))


def _ptr(x):
    """Address of a numpy array / torch tensor / raw int."""
    if x is None:
        return None
    if isinstance(x, int):
        return x
    if isinstance(x, np.ndarray):
        return x.ctypes.data
    return x.data_ptr()  # torch tensor (host pinned or device)


class Pod:
    """One handle = one GPU = one rank's cell shard. Mirrors the call sequence of getModes."""

    TIMER_NAMES = ["upload", "gram", "allreduce", "eig", "lift", "scale", "download", "gram_kernel"]

    def __init__(self, device=0):
        self.L = lib()
        self.h = ctypes.c_void_p()
        rc = self.L.ithaca_b200_create(ctypes.byref(self.h), device)
        if rc:
            raise IthacaB200Error(self.L.ithaca_b200_last_error(None).decode())
        self.m = self.dim = self.nCells = 0

    def close(self):
        if self.h:
            self.L.ithaca_b200_destroy(self.h)
            self.h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _chk(self, rc):
        if rc:
            raise IthacaB200Error("status %d: %s" % (rc, self.L.ithaca_b200_last_error(self.h).decode()))

    @property
    def stream(self):
        return self.L.ithaca_b200_stream(self.h)

    def launch_count(self, reset=False):
        return self.L.ithaca_b200_launch_count(self.h, 1 if reset else 0)

    def timings(self):
        out = np.zeros(8)
        self._chk(self.L.ithaca_b200_timings(self.h, out.ctypes.data, 8))
        return dict(zip(self.TIMER_NAMES, out.tolist()))

    def comm_init(self, id128, rank, nranks):
        buf = (ctypes.c_char * 128).from_buffer_copy(bytes(id128)) if id128 is not None else None
        self._chk(self.L.ithaca_b200_comm_init(self.h, buf, rank, nranks))

    def set_volumes(self, V):
        n = V.shape[0]
        self._chk(self.L.ithaca_b200_pod_set_volumes(self.h, _ptr(V), n))
        self.nCells = n

    def upload(self, snaps, dim, nCells=None):
        """snaps: list of m arrays/tensors (each nCells*dim contiguous doubles), or a Fortran-ordered
        (Nd, m) numpy array whose columns are the snapshots."""
        if isinstance(snaps, np.ndarray) and snaps.ndim == 2:
            assert snaps.flags.f_contiguous and snaps.dtype == np.float64
            ptrs = [snaps.ctypes.data + j * snaps.strides[1] for j in range(snaps.shape[1])]
            Nd = snaps.shape[0]
            self._keep = snaps
        else:
            ptrs = [_ptr(s) for s in snaps]
            Nd = int(snaps[0].numel()) if hasattr(snaps[0], "numel") else int(snaps[0].size)
            self._keep = snaps
        m = len(ptrs)
        nCells = Nd // dim if nCells is None else nCells
        arr = (ctypes.c_void_p * m)(*ptrs)
        self._chk(self.L.ithaca_b200_pod_upload(self.h, arr, m, dim, nCells))
        self.m, self.dim = m, dim

    def upload_second(self, snaps):
        """The second field list of getModes(snapshots, modes, fields2, ...): replaces the resident matrix, keeps the eigensolve."""
        if isinstance(snaps, np.ndarray) and snaps.ndim == 2:
            assert snaps.flags.f_contiguous and snaps.dtype == np.float64
            ptrs = [snaps.ctypes.data + j * snaps.strides[1] for j in range(snaps.shape[1])]
            n = snaps.shape[0]
        else:
            ptrs = [_ptr(s) for s in snaps]
            n = int(snaps[0].numel()) if hasattr(snaps[0], "numel") else int(snaps[0].size)
        self._keep = snaps
        arr = (ctypes.c_void_p * len(ptrs))(*ptrs)
        self._chk(self.L.ithaca_b200_pod_upload_second(self.h, arr, len(ptrs), n))
        self.nCells, self.dim = n, 1

    def gram(self, want=True):
        G = np.zeros((self.m, self.m), order="F") if want else None
        self._chk(self.L.ithaca_b200_pod_gram(self.h, _ptr(G)))
        return G

    def gram_volumes(self, vols, want=True):
        """Correlation matrix with one volume field per snapshot (ITHACAPOD.C:1425-1445). vols: Fortran (Nd, m) array or a
        list of m buffers."""
        if isinstance(vols, np.ndarray) and vols.ndim == 2:
            assert vols.flags.f_contiguous and vols.dtype == np.float64 and vols.shape[1] == self.m
            ptrs = [vols.ctypes.data + j * vols.strides[1] for j in range(vols.shape[1])]
        else:
            assert len(vols) == self.m
            ptrs = [_ptr(v) for v in vols]
        arr = (ctypes.c_void_p * self.m)(*ptrs)
        G = np.zeros((self.m, self.m), order="F") if want else None
        self._chk(self.L.ithaca_b200_pod_gram_volumes(self.h, arr, _ptr(G)))
        return G

    def set_gram(self, G):
        G = np.asfortranarray(G, dtype=np.float64)
        self._chk(self.L.ithaca_b200_pod_set_gram(self.h, G.ctypes.data, G.shape[0]))
        self.m = G.shape[0]

    def eig(self, r):
        ev = np.zeros(self.m)
        Q = np.zeros((self.m, r), order="F")
        sw = ctypes.c_int(0)
        self._chk(self.L.ithaca_b200_pod_eig(self.h, r, ev.ctypes.data, Q.ctypes.data if r else None, ctypes.byref(sw)))
        self.sweeps = sw.value
        return ev, Q

    def lift(self, r, norm="L2", out=None, want=True):
        """out: optional list of r host/device buffers; want=False keeps the modes on the device."""
        if norm not in ("L2", "Frobenius"):
            raise IthacaB200Error("The PODnorm can be only L2 or Frobenius")
        nf = np.zeros(r)
        Phi = None
        arr = None
        if out is not None:
            arr = (ctypes.c_void_p * r)(*[_ptr(o) for o in out])
        elif want:
            Phi = np.zeros((self.nCells * self.dim, r), order="F")
            arr = (ctypes.c_void_p * r)(*[Phi.ctypes.data + i * Phi.strides[1] for i in range(r)])
        self._chk(self.L.ithaca_b200_pod_lift(self.h, r, 0 if norm == "L2" else 1, arr, nf.ctypes.data))
        self.r = r
        return Phi, nf

    # ---- callers either side of getModes: Modes::project / getCoeffs / Modes::reconstruct on the resident modes ----
    def set_modes(self, modes):
        """modes: list of r arrays/tensors (nCells*dim doubles each) or a Fortran-ordered (Nd, r) array."""
        if isinstance(modes, np.ndarray) and modes.ndim == 2:
            assert modes.flags.f_contiguous and modes.dtype == np.float64
            ptrs = [modes.ctypes.data + j * modes.strides[1] for j in range(modes.shape[1])]
        else:
            ptrs = [_ptr(x) for x in modes]
        self._keep_modes = modes
        arr = (ctypes.c_void_p * len(ptrs))(*ptrs)
        self._chk(self.L.ithaca_b200_pod_set_modes(self.h, arr, len(ptrs)))
        self.r = len(ptrs)

    def project(self, nmodes, proj_type="G", r=None):
        """Modes::project over the uploaded snapshots (Modes.C:228-290): (nmodes or all) x m."""
        if proj_type not in ("G", "F"):
            raise IthacaB200Error("Projection type can be G for Galerking or F for Frobenius")
        n = nmodes if nmodes else (r if r is not None else getattr(self, "r", 0))
        out = np.zeros((n, self.m), order="F")
        self._chk(self.L.ithaca_b200_pod_project(self.h, nmodes, 0 if proj_type == "G" else 1, out.ctypes.data))
        return out

    def modes_mass(self, nmodes, r=None):
        n = nmodes if nmodes else (r if r is not None else getattr(self, "r", 0))
        out = np.zeros((n, n), order="F")
        self._chk(self.L.ithaca_b200_pod_modes_mass(self.h, nmodes, out.ctypes.data))
        return out

    def get_coeffs(self, nmodes, r=None):
        """ITHACAutilities::getCoeffs over the uploaded snapshots (ITHACAcoeffsMass.C:348-443)."""
        n = nmodes if nmodes else (r if r is not None else getattr(self, "r", 0))
        out = np.zeros((n, self.m), order="F")
        self._chk(self.L.ithaca_b200_pod_get_coeffs(self.h, nmodes, out.ctypes.data))
        return out

    def reconstruct(self, coeff, clip_negative=False, out=None):
        """Modes::reconstruct (Modes.C:294-305): Phi[:, :Nmodes] @ coeff, coeff (Nmodes, k)."""
        coeff = np.asfortranarray(coeff, dtype=np.float64)
        if coeff.ndim == 1:
            coeff = np.asfortranarray(coeff[:, None])
        n, k = coeff.shape
        R = None
        if out is None:
            R = np.zeros((self.nCells * self.dim, k), order="F")
            ptrs = [R.ctypes.data + j * R.strides[1] for j in range(k)]
        else:
            ptrs = [_ptr(o) for o in out]
        arr = (ctypes.c_void_p * k)(*ptrs)
        self._chk(self.L.ithaca_b200_pod_reconstruct(self.h, n, coeff.ctypes.data, k, 1 if clip_negative else 0, arr))
        return R

    def lift_scaled(self, factors, want=True):
        """modes_i = (S q_i) * factors[i] (getWeightedModes, ITHACAPOD.C:771-777)."""
        f = np.ascontiguousarray(factors, dtype=np.float64)
        r = f.shape[0]
        Phi = arr = None
        if want:
            Phi = np.zeros((self.nCells * self.dim, r), order="F")
            arr = (ctypes.c_void_p * r)(*[Phi.ctypes.data + i * Phi.strides[1] for i in range(r)])
        self._chk(self.L.ithaca_b200_pod_lift_scaled(self.h, r, f.ctypes.data, arr))
        self.r = r
        return Phi

    # ---- row-streamed POD (a shard larger than HBM passes through two chunk slots) ----
    def stream_begin(self, m, dim, nCellsShard):
        self._chk(self.L.ithaca_b200_pod_stream_begin(self.h, m, dim, nCellsShard))
        self.m, self.dim, self.nCells = m, dim, nCellsShard

    def stream_upload(self, slot, cell0, ncells, V_chunk, snap_ptrs):
        """snap_ptrs: m addresses of the chunk's first value in each snapshot (host pinned, host pageable or device)."""
        arr = (ctypes.c_void_p * self.m)(*snap_ptrs)
        self._chk(self.L.ithaca_b200_pod_stream_upload(self.h, slot, cell0, ncells, _ptr(V_chunk), arr))

    def stream_slot(self, slot, cell0, ncells):
        """Device addresses (S, ld, V) of a slot for an in-place producer."""
        S, V, ld = ctypes.c_void_p(), ctypes.c_void_p(), ctypes.c_int64()
        self._chk(self.L.ithaca_b200_pod_stream_slot(self.h, slot, cell0, ncells, ctypes.byref(S), ctypes.byref(ld), ctypes.byref(V)))
        return S.value, ld.value, V.value

    def stream_gram_add(self, slot):
        self._chk(self.L.ithaca_b200_pod_stream_gram_add(self.h, slot))

    def stream_gram_finish(self, want=True):
        G = np.zeros((self.m, self.m), order="F") if want else None
        self._chk(self.L.ithaca_b200_pod_stream_gram_finish(self.h, _ptr(G)))
        return G

    def stream_lift_add(self, slot, r, norm="L2"):
        self._chk(self.L.ithaca_b200_pod_stream_lift_add(self.h, slot, r, 0 if norm == "L2" else 1))

    def stream_lift_finish(self, r, out=None, want=True):
        nf = np.zeros(r)
        Phi = arr = None
        if out is not None:
            arr = (ctypes.c_void_p * r)(*[_ptr(o) for o in out])
        elif want:
            Phi = np.zeros((self.nCells * self.dim, r), order="F")
            arr = (ctypes.c_void_p * r)(*[Phi.ctypes.data + i * Phi.strides[1] for i in range(r)])
        self._chk(self.L.ithaca_b200_pod_stream_lift_finish(self.h, r, arr, nf.ctypes.data))
        self.r = r
        return Phi, nf

    # ---- synthetic benchmark data written in place on the device (measurement helper) ----
    def synth_snapshots(self, S_ptr, ld, row0, nrows, Nd_global, m, coef_ptr, R, eps, seed):
        self._chk(self.L.ithaca_b200_synth_snapshots(self.h, S_ptr, ld, row0, nrows, Nd_global, m, coef_ptr, R, eps, seed))

    def synth_volumes(self, V_ptr, cell0, ncells, ncells_global, seed):
        self._chk(self.L.ithaca_b200_synth_volumes(self.h, V_ptr, cell0, ncells, ncells_global, seed))

    def col_norms_noalloc(self, Phi_ptr, nrows, ld, ncols, V_ptr, dim, norm="L2", out_ptr=None, stream=None):
        """Norm factors of caller-held DEVICE columns through the allocation-free entry point (all pointers are device addresses)."""
        self._chk(self.L.ithaca_b200_col_norms_noalloc(self.h, Phi_ptr, nrows, ld, ncols, V_ptr, dim, 0 if norm == "L2" else 1,
                                                       out_ptr, stream))

    def lift_patch(self, Sbc, r):
        Sbc = np.asfortranarray(Sbc, dtype=np.float64)
        out = np.zeros((Sbc.shape[0], r), order="F")
        self._chk(self.L.ithaca_b200_pod_lift_patch(self.h, Sbc.ctypes.data, Sbc.shape[0], out.ctypes.data))
        return out


def mesh_renumbering(mesh):
    """old2new of the library's locality (Cuthill-McKee) ordering; host code."""
    own = np.ascontiguousarray(mesh.owner, dtype=np.int32)
    nei = np.ascontiguousarray(mesh.neighbour, dtype=np.int32)
    out = np.zeros(int(mesh.nCells), dtype=np.int32)
    if lib().ithaca_b200_mesh_renumbering(int(mesh.nCells), int(mesh.nInt), _ptr(own), _ptr(nei), _ptr(out)):
        raise IthacaB200Error("mesh_renumbering: bad argument")
    return out


class Partition:
    """One rank's shard of a mesh: contiguous owned cell range + halo layers (host code of the library, no GPU needed).
    Addressing in the form of decomposePar's cellProcAddressing / faceProcAddressing
    (src/ITHACA_CORE/ITHACAparallel/ITHACAparallel.C:43-74)."""

    METHODS = {"label": 0, "locality": 1}

    def __init__(self, mesh, rank, nranks, halo_layers=2, method="label"):
        self.L = lib()
        self.p = ctypes.c_void_p()
        self._own = np.ascontiguousarray(mesh.owner, dtype=np.int32)
        self._nei = np.ascontiguousarray(mesh.neighbour, dtype=np.int32)
        rc = self.L.ithaca_b200_partition_create(ctypes.byref(self.p), int(mesh.nCells), int(mesh.nInt), int(mesh.nBnd),
                                                 _ptr(self._own), _ptr(self._nei), rank, nranks, halo_layers,
                                                 self.METHODS[method])
        if rc:
            raise IthacaB200Error("partition_create: bad argument (status %d)" % rc)
        sz = np.zeros(5, dtype=np.int64)
        self.L.ithaca_b200_partition_sizes(self.p, sz.ctypes.data)
        self.nLocalCells, self.nOwned, self.nLocalInt, self.nLocalBnd, self.nCut = [int(x) for x in sz]
        self.rank, self.nranks = rank, nranks

    def addressing(self):
        cpa = np.zeros(self.nLocalCells, dtype=np.int32)
        fpa = np.zeros(self.nLocalInt + self.nLocalBnd + self.nCut, dtype=np.int32)
        self.L.ithaca_b200_partition_addressing(self.p, cpa.ctypes.data, fpa.ctypes.data)
        return cpa, fpa

    def gather(self, field, kind, dim):
        """Shard copy of a global cell field (kind 'cell') or boundary-face field (kind 'bnd')."""
        g = np.ascontiguousarray(field, dtype=np.float64)
        n = self.nLocalCells if kind == "cell" else self.nLocalBnd + self.nCut
        out = np.zeros(n * dim)
        rc = self.L.ithaca_b200_partition_gather(self.p, 0 if kind == "cell" else 1, dim, g.ctypes.data, out.ctypes.data)
        if rc:
            raise IthacaB200Error("partition_gather: bad argument")
        return out

    def close(self):
        if self.p:
            self.L.ithaca_b200_partition_destroy(self.p)
            self.p = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


OPS = {"M": 0, "B": 1, "K": 2, "P": 3, "C": 4, "Cmat": 5, "D": 6, "G": 7, "BC3": 8, "BC4": 9, "BC1": 10, "BC2": 11, "BP": 12, "Cf": 13}
FVC = {"flux": 0, "div_phi": 1, "laplacian": 2, "grad_p": 3, "div_u": 4}


class Projector:
    """Galerkin projection on a Pod handle's GPU: mirrors the steadyNS term builders
    (src/ITHACA_FOMPROBLEMS/steadyNS/steadyNS.C:883-1137). `mesh` is any object with the attributes of
    ithaca_b200_mesh_upload's arguments (nCells, nInt, nBnd, owner, neighbour, Sf, magSf, w, dc, dcB, corr, V, bmask)."""

    TIMER_NAMES = ["flux", "div_phi", "laplacian", "grad_p", "div_u", "contraction", "face_records"]

    def __init__(self, pod):
        self.pod, self.L, self.h = pod, pod.L, pod.h
        self.N = self.Np = 0

    def set_mesh(self, mesh, corrected=True):
        c = lambda a, dt: np.ascontiguousarray(a, dtype=dt)
        own = np.concatenate([c(mesh.owner, np.int32)])
        k = dict(owner=own, nei=c(mesh.neighbour, np.int32), Sf=c(mesh.Sf, np.float64), magSf=c(mesh.magSf, np.float64),
                 w=c(mesh.w, np.float64), dc=c(mesh.dc, np.float64), dcB=c(mesh.dcB, np.float64),
                 corr=c(mesh.corr, np.float64) if corrected else None, V=c(mesh.V, np.float64),
                 bmask=c(mesh.bmask, np.uint8))
        self.nCells, self.nInt, self.nBnd = int(mesh.nCells), int(mesh.nInt), int(mesh.nBnd)
        self.pod._chk(self.L.ithaca_b200_mesh_upload(
            self.h, self.nCells, self.nInt, self.nBnd, _ptr(k["owner"]), _ptr(k["nei"]), _ptr(k["Sf"]), _ptr(k["magSf"]),
            _ptr(k["w"]), _ptr(k["dc"]), _ptr(k["dcB"]), _ptr(k["corr"]), _ptr(k["V"]), _ptr(k["bmask"])))

    def set_schemes(self, grad_limit_k=0.0, sngrad_psi=-1.0, C=None, Cf=None):
        """fvSchemes beyond 'Gauss linear': gradSchemes 'cellMDLimited Gauss linear k', snGrad 'limited psi'
        (tutorials/CFD/04unsteadyNS/system/fvSchemes). C, Cf: mesh.C(), mesh.Cf() in the uploaded mesh's order."""
        c = lambda a: None if a is None else np.ascontiguousarray(a, dtype=np.float64)
        C, Cf = c(C), c(Cf)
        self.pod._chk(self.L.ithaca_b200_mesh_set_schemes(self.h, float(grad_limit_k), float(sngrad_psi), _ptr(C), _ptr(Cf)))

    def set_convection(self, scheme="linear"):
        """divSchemes 'Gauss linear | upwind | linearUpwind grad(U)' of fvc::div(phi, U_k)."""
        if scheme not in ("linear", "upwind", "linearUpwind"):
            raise IthacaB200Error("unknown convection scheme " + str(scheme))
        self.pod._chk(self.L.ithaca_b200_mesh_set_convection(self.h, {"linear": 0, "upwind": 1, "linearUpwind": 2}[scheme]))

    def set_mesh_shard(self, mesh, part, corrected=True):
        """Upload `part`'s shard of the GLOBAL mesh; set_basis then takes shard fields (part.gather) and project()
        all-reduces over the Pod's communicator."""
        c = lambda a, dt: np.ascontiguousarray(a, dtype=dt)
        k = [c(mesh.owner, np.int32), c(mesh.neighbour, np.int32), c(mesh.Sf, np.float64), c(mesh.magSf, np.float64),
             c(mesh.w, np.float64), c(mesh.dc, np.float64), c(mesh.dcB, np.float64),
             c(mesh.corr, np.float64) if corrected else None, c(mesh.V, np.float64), c(mesh.bmask, np.uint8)]
        self.nCells, self.nInt, self.nBnd = part.nLocalCells, part.nLocalInt, part.nLocalBnd + part.nCut
        self.pod._chk(self.L.ithaca_b200_mesh_upload_shard(self.h, part.p, *[_ptr(a) for a in k]))

    def addressing(self):
        nnz = 2 * self.nInt + self.nBnd
        off = np.zeros(self.nCells + 1, dtype=np.int32)
        ent = np.zeros(nnz, dtype=np.int32)
        other = np.zeros(nnz, dtype=np.int32)
        d = ctypes.c_uint64(0)
        self.pod._chk(self.L.ithaca_b200_mesh_addressing(self.h, off.ctypes.data, ent.ctypes.data, other.ctypes.data, ctypes.byref(d)))
        return off, ent, other, d.value

    def set_basis(self, which, fields, bfields):
        keep = [np.ascontiguousarray(f, dtype=np.float64) if isinstance(f, np.ndarray) else f for f in fields]
        keepb = [np.ascontiguousarray(f, dtype=np.float64) if isinstance(f, np.ndarray) else f for f in bfields]
        n = len(keep)
        a = (ctypes.c_void_p * n)(*[_ptr(f) for f in keep])
        b = (ctypes.c_void_p * n)(*[_ptr(f) for f in keepb])
        self.pod._chk(self.L.ithaca_b200_basis_upload(self.h, 0 if which == "U" else 1, a, b, n))
        if which == "U":
            self.N = n
        else:
            self.Np = n

    def set_flux(self, phis):
        """fluxMethod 'consistent': the stored face-flux modes L_PHImodes (steadyNS.C:1010-1015) convect instead of
        linearInterpolate(U_j) & Sf. phis: N arrays of nInt+nBnd doubles (caller's face order), or None to reset."""
        if phis is None:
            self.pod._chk(self.L.ithaca_b200_flux_upload(self.h, None, 0))
            return
        keep = [np.ascontiguousarray(f, dtype=np.float64) if isinstance(f, np.ndarray) else f for f in phis]
        a = (ctypes.c_void_p * len(keep))(*[_ptr(f) for f in keep])
        self.pod._chk(self.L.ithaca_b200_flux_upload(self.h, a, len(keep)))

    def project(self, op):
        N, Np = self.N, self.Np
        shape = {"M": (N, N), "B": (N, N), "K": (N, Np), "P": (Np, N), "C": (N, N, N), "Cmat": (N, N, N), "D": (Np, Np),
                 "G": (Np, N, N), "BC3": (Np, N), "BC4": (Np, N), "BC1": (Np, N), "BC2": (Np, N, N), "BP": (Np, N), "Cf": (Np, N, N)}[op]
        out = np.zeros(shape, order="F")
        self.pod._chk(self.L.ithaca_b200_project(self.h, OPS[op], out.ctypes.data))
        return out

    def project_background(self, vls, vlsb, divergent=False):
        """convective_background (L, N x N) or divergent_convective_background (L_D, Np x N), steadyNS.C:1256-1293."""
        vls = np.ascontiguousarray(vls, dtype=np.float64) if isinstance(vls, np.ndarray) else vls
        vlsb = np.ascontiguousarray(vlsb, dtype=np.float64) if isinstance(vlsb, np.ndarray) else vlsb
        out = np.zeros((self.Np if divergent else self.N, self.N), order="F")
        self.pod._chk(self.L.ithaca_b200_project_background(self.h, 1 if divergent else 0, _ptr(vls), _ptr(vlsb), out.ctypes.data))
        return out

    def bc_velocity(self, patch_start, patch_size, comp):
        """bcVelocityVec / bcVelocityMat of one (patch, component) pair (steadyNS.C:1515-1582): (vec[N], mat[N, N])."""
        vec = np.zeros(self.N)
        mat = np.zeros((self.N, self.N), order="F")
        self.pod._chk(self.L.ithaca_b200_project_bc_velocity(self.h, patch_start, patch_size, comp, vec.ctypes.data, mat.ctypes.data))
        return vec, mat

    def fvc(self, op, j, k=0):
        n = {"flux": self.nInt + self.nBnd, "div_phi": 3 * self.nCells, "laplacian": 3 * self.nCells,
             "grad_p": 3 * self.nCells, "div_u": self.nCells}[op]
        out = np.zeros(n)
        self.pod._chk(self.L.ithaca_b200_fvc_eval(self.h, FVC[op], j, k, out.ctypes.data))
        return out

    def timings(self):
        out = np.zeros(7)
        self.pod._chk(self.L.ithaca_b200_proj_timings(self.h, out.ctypes.data, 7))
        return dict(zip(self.TIMER_NAMES, out.tolist()))


def getWeightedModes(pod, snapshots, V, dim, nmodes=0, solver="spectra", Sbc=None):
    """ITHACAPOD::getWeightedModes (ITHACAPOD.C:715-845): same Gram and eigensolve, modes_i = (S q_i) lambda_i^(-1/2) lambda_i.
    nmodes rule of this variant: at most m - 2 whatever the solver (:723-729)."""
    m = snapshots.shape[1] if isinstance(snapshots, np.ndarray) else len(snapshots)
    if nmodes == 0:
        nmodes = m - 2
    if nmodes > m - 2:
        raise IthacaB200Error("The number of requested modes cannot be bigger than the number of Snapshots - 2")
    pod.set_volumes(V)
    pod.upload(snapshots, dim)
    G = pod.gram()
    ev_all, Q = pod.eig(nmodes)
    ev = ev_all[:nmodes] if solver == "spectra" else ev_all
    lam = 1.0 / np.sqrt(ev[:nmodes])
    factors = lam * ev[:nmodes]
    Phi = pod.lift_scaled(factors)
    Pbc = pod.lift_patch(Sbc, nmodes) if Sbc is not None else None
    evn = ev / ev.sum()
    return dict(G=G, evals=ev, evals_norm=evn, cum=np.cumsum(evn), evecs=Q, factors=factors, modes=Phi, modes_bc=Pbc)


def corMatrix(pod, snapshots, V, dim):
    """ITHACAPOD::corMatrix (ITHACAPOD.C:955-1008): matrix(i, j) = fvc::domainIntegrate(snapshots[i] (&) snapshots[j]),
    the volume-weighted correlation matrix, lower triangle mirrored: one Gram launch."""
    pod.set_volumes(V)
    pod.upload(snapshots, dim)
    return pod.gram()


def getNestedSnapshotMatrix(pod, snapshots, V, dim, Npar, NnestedOut, solver="spectra"):
    """ITHACAPOD::getNestedSnapshotMatrix (ITHACAPOD.C:41-82): the snapshot set holds Npar parameter groups of
    Nt = m / Npar consecutive time snapshots; every group is compressed by getWeightedModes to NnestedOut modes and the
    groups' modes are appended in order (the nested-POD snapshot matrix). snapshots: (Nd, m) array."""
    m = snapshots.shape[1]
    Nt = m // Npar
    out = []
    for i in range(Npar):
        grp = np.asfortranarray(snapshots[:, Nt * i:Nt * (i + 1)])
        out.append(getWeightedModes(pod, grp, V, dim, NnestedOut, solver)["modes"])
    return np.asfortranarray(np.concatenate(out, axis=1))


def getMeanMemoryEfficient(pod, block_source, m, V, dim, block=64):
    """ITHACAPOD::getMeanMemoryEfficient (ITHACAPOD.C:660-697): mean of a snapshot set streamed block by block
    (block_source(first, nb) -> nb buffers); each block contributes S_block (1/m) through the reconstruct kernel."""
    L, h = pod.L, pod.h
    nCells = V.shape[0]
    pod.set_volumes(V)
    pod._chk(L.ithaca_b200_pod_block_begin(h, m, dim, nCells))
    pod.m, pod.dim = m, dim
    mean = np.zeros(nCells * dim)
    for first in range(0, m, block):
        nb = min(block, m - first)
        pod.set_modes(block_source(first, nb))
        mean += pod.reconstruct(np.full((nb, 1), 1.0 / m))[:, 0]
    return mean


def getModesSVD(pod, snapshots, V, dim, nmodes):
    """ITHACAPOD::getModesSVD (ITHACAPOD.C:858-941) without forming the N_d x m SVD: the left singular vectors of
    diag(sqrt V) S are diag(sqrt V) S q_i / sigma_i with (sigma_i^2, q_i) the eigenpairs of the Gram matrix, so
    modes_i = diag(1/sqrt V) u_i = (S q_i) / sigma_i: one Gram, one eigensolve, one scaled lift. The exported "eigenvalues"
    are ALL m singular values over their sum (:905). Singular values below sqrt(eps) sigma_0 carry the squaring's loss."""
    m = snapshots.shape[1] if isinstance(snapshots, np.ndarray) else len(snapshots)
    if nmodes <= 0 or nmodes > m:
        raise IthacaB200Error("The number of requested modes cannot be bigger than the number of Snapshots")
    pod.set_volumes(V)
    pod.upload(snapshots, dim)
    G = pod.gram()
    ev_all, Q = pod.eig(nmodes)
    sv = np.sqrt(np.maximum(ev_all, 0.0))
    Phi = pod.lift_scaled(1.0 / sv[:nmodes])
    evn = sv / sv.sum()
    return dict(G=G, sv=sv, evals_norm=evn, cum=np.cumsum(evn), evecs=Q, modes=Phi)


def getModesVolumes(pod, snapshots, volumes, nmodes=0, solver="spectra", Sbc=None):
    """getModes(snapshots, modes, Volumes, ...) for moving meshes (ITHACAPOD.C:1392-1541): one volume field per snapshot,
    G(i,j) = S_i^T diag(sqrt|V_i V_j|) S_j, no negative shift, modes = (S Q) diag(lambda^-1/2), eigenvalues of the kept
    modes over their sum. Scalar fields (the reference multiplies volume and snapshot columns coefficient-wise)."""
    m = snapshots.shape[1] if isinstance(snapshots, np.ndarray) else len(snapshots)
    nmodes = check_nmodes(m, nmodes, solver)
    v0 = volumes[:, 0] if isinstance(volumes, np.ndarray) and volumes.ndim == 2 else volumes[0]
    pod.set_volumes(np.ascontiguousarray(v0))
    pod.upload(snapshots, 1)
    G = pod.gram_volumes(volumes)
    ev_all, Q = pod.eig(nmodes)
    ev = ev_all[:nmodes]
    factors = 1.0 / np.sqrt(ev)
    Phi = pod.lift_scaled(factors)
    Pbc = pod.lift_patch(Sbc, nmodes) if Sbc is not None else None   # same factors applied to the boundary rows
    evn = ev / ev.sum()
    return dict(G=G, evals=ev, evals_norm=evn, cum=np.cumsum(evn), evecs=Q, modes=Phi, modes_bc=Pbc)


def getModesFields2(pod, snapshots, fields2, V, dim2, nmodes=0, solver="spectra", Sbc=None):
    """getModes(snapshots, modes, fields2, ...) (ITHACAPOD.C:1940-2086): Gram and eigenvectors of `fields2` (dim2 components,
    volumes V), modes_i = |lambda_i|^(-1/2) sum_j Q(j,i) snapshots[j] for a second list of the same length (any field type:
    (n, m) Fortran array or m buffers; Sbc: its boundary rows). Both solvers keep nmodes eigenvalues here (:2004, :2011), there
    is no negative shift, and they are normalised by their own sum (:2036)."""
    m = fields2.shape[1] if isinstance(fields2, np.ndarray) else len(fields2)
    nmodes = check_nmodes(m, nmodes, solver)
    pod.set_volumes(V)
    pod.upload(fields2, dim2)
    G = pod.gram()
    ev_all, Q = pod.eig(nmodes)
    ev = ev_all[:nmodes]
    factors = 1.0 / np.sqrt(np.abs(ev))
    pod.upload_second(snapshots)
    Phi = pod.lift_scaled(factors)
    Pbc = pod.lift_patch(Sbc, nmodes) if Sbc is not None else None
    evn = ev / ev.sum()
    return dict(G=G, evals=ev, evals_norm=evn, cum=np.cumsum(evn), evecs=Q, factors=factors, modes=Phi, modes_bc=Pbc)


def DEIMmodes(pod, snapshots, V, dim, nmodes=0, solver="spectra", norm="L2", Sbc=None):
    """ITHACAPOD::DEIMmodes (ITHACAPOD.C:1739-1937): the dense arithmetic is getModes' own (same correlation matrix,
    eigensolve, negative shift, norm factors and boundary lift, :1774-1885); only the export folder differs
    (./ITHACAoutput/DEIM/<FunctionName>), which stays with the caller."""
    return getModes(pod, snapshots, V, dim, nmodes, solver, norm, Sbc)


def getModesMemoryEfficient(pod, block_source, m, V, dim, nmodes=0, solver="spectra", norm="L2", block=64, want_modes=True, Sbc=None):
    """Snapshot-blocked POD (ITHACAPOD::getModesMemoryEfficient, ITHACAPOD.C:344-628) through the C ABI.
    block_source(first, nb) -> list of nb snapshot buffers (numpy arrays or torch tensors, nCells*dim doubles each):
    the caller streams them from disk / host / a generator; two blocks are resident on the device at a time."""
    if norm not in ("L2", "Frobenius"):
        raise IthacaB200Error("The PODnorm can be only L2 or Frobenius")
    nmodes = check_nmodes(m, nmodes, solver)
    nk = 0 if norm == "L2" else 1
    L, h = pod.L, pod.h
    nCells = V.shape[0]
    pod.set_volumes(V)
    pod._chk(L.ithaca_b200_pod_block_begin(h, m, dim, nCells))
    pod.m, pod.dim = m, dim
    starts = list(range(0, m, block))

    def up(slot, first):
        nb = min(block, m - first)
        bufs = block_source(first, nb)
        arr = (ctypes.c_void_p * nb)(*[_ptr(b) for b in bufs])
        pod._chk(L.ithaca_b200_pod_block_upload(h, slot, arr, nb, first))
        return bufs

    for bi, fi in enumerate(starts):          # block pairs (I, J >= I): block I stays in slot 0, J streams through slot 1
        keep = up(0, fi)
        pod._chk(L.ithaca_b200_pod_block_gram(h, 0, 0, nk))
        for fj in starts[bi + 1:]:
            keep2 = up(1, fj)
            pod._chk(L.ithaca_b200_pod_block_gram(h, 0, 1, nk))
    G = np.zeros((m, m), order="F")
    pod._chk(L.ithaca_b200_pod_block_gram_finish(h, G.ctypes.data))
    ev_all, Q = pod.eig(nmodes)
    for fi in starts:
        keep = up(0, fi)
        pod._chk(L.ithaca_b200_pod_block_lift(h, 0, nmodes))
    nf = np.zeros(nmodes)
    Phi = arr = None
    if want_modes:
        Phi = np.zeros((nCells * dim, nmodes), order="F")
        arr = (ctypes.c_void_p * nmodes)(*[Phi.ctypes.data + i * Phi.strides[1] for i in range(nmodes)])
    pod._chk(L.ithaca_b200_pod_block_lift_finish(h, nmodes, nk, arr, nf.ctypes.data))
    pod.r = nmodes
    Pbc = pod.lift_patch(Sbc, nmodes) if Sbc is not None else None   # modesEigBC = SnapMatrixBC * eigvec / normFactor (ITHACAPOD.C:573-578)
    evn, cum = postprocess_eigenvalues(ev_all, nmodes, solver)
    return dict(G=G, evals_all=ev_all, evals_norm=evn, cum=cum, evecs=Q, normfact=nf, modes=Phi, modes_bc=Pbc)


def getModesStreamed(pod, snapshots, V, dim, nmodes=0, solver="spectra", norm="L2", chunk_cells=None, Sbc=None, want_modes=True):
    """getModes for a cell shard larger than HBM, through the row-streamed calls (pod_stream_*): the shard is cut into chunks of
    `chunk_cells` consecutive cells; pass 1 sums the chunk Grams (the reference's per-rank Gram + reduce, ITHACAcoeffsMass.C:85-115,
    with more chunks than GPUs), pass 2 lifts chunk by chunk (ITHACAPOD.C:194-243 is row-local). snapshots: Fortran (Nd, m) host
    array or a list of m host buffers (numpy / pinned torch); the upload of chunk k+1 overlaps the work on chunk k. The chunks
    still resident after pass 1 are lifted first, without a second upload."""
    if norm not in ("L2", "Frobenius"):
        raise IthacaB200Error("The PODnorm can be only L2 or Frobenius")
    if isinstance(snapshots, np.ndarray) and snapshots.ndim == 2:
        assert snapshots.flags.f_contiguous and snapshots.dtype == np.float64
        base = [snapshots.ctypes.data + j * snapshots.strides[1] for j in range(snapshots.shape[1])]
    else:
        base = [_ptr(x) for x in snapshots]
    m = len(base)
    nmodes = check_nmodes(m, nmodes, solver)
    Vp = _ptr(V)
    nCells = int(V.shape[0])
    if not chunk_cells:
        chunk_cells = nCells
    chunks = [(c0, min(chunk_cells, nCells - c0)) for c0 in range(0, nCells, chunk_cells)]
    pod.stream_begin(m, dim, nCells)

    def up(slot, k):
        c0, nc = chunks[k]
        pod.stream_upload(slot, c0, nc, Vp + 8 * c0, [b + 8 * c0 * dim for b in base])

    up(0, 0)
    for k in range(len(chunks)):
        if k + 1 < len(chunks):
            up((k + 1) & 1, k + 1)
        pod.stream_gram_add(k & 1)
    G = pod.stream_gram_finish()
    ev_all, Q = pod.eig(nmodes)
    n = len(chunks)
    slot_of = {n - 1: (n - 1) & 1}              # the last two chunks of pass 1 are still in the slots: lift them first
    if n > 1:
        slot_of[n - 2] = (n - 2) & 1
    order = sorted(slot_of, reverse=True) + list(range(n - 3, -1, -1))
    pending = order[len(slot_of):]
    for k in order:
        pod.stream_lift_add(slot_of[k], nmodes, norm)
        if pending:                              # the slot just consumed takes the next chunk while the other one is lifted
            kn = pending.pop(0)
            up(slot_of[k], kn)
            slot_of[kn] = slot_of[k]
    Phi, nf = pod.stream_lift_finish(nmodes, want=want_modes)
    Pbc = pod.lift_patch(Sbc, nmodes) if Sbc is not None else None
    evn, cum = postprocess_eigenvalues(ev_all, nmodes, solver)
    return dict(G=G, evals_all=ev_all, evals_norm=evn, cum=cum, evecs=Q, normfact=nf, modes=Phi, modes_bc=Pbc)


COV_UNREAL = 999999.999999   # PODTemplate::buildCovMatrix marks entries not computed yet with this value (PODTemplate.C:422)


def buildCovMatrix(pod, block_source, m, V, dim, nBlocks, folder, name, norm="L2", keep_temp=3, on_checkpoint=None):
    """PODTemplate::buildCovMatrix (src/ITHACA_CORE/ITHACAPOD/PODTemplate.C:419-616) on the device: the m x m correlation
    matrix of a snapshot set streamed in nBlocks blocks of q = m // nBlocks snapshots (+ one remainder block), with the
    reference's checkpoint / resume protocol:
      * after every square block pair (i, j < i) the matrix is written to  folder/name_temp_<i>_<j>.npy  (cnpy layout,
        entries not computed yet = 999999.999999, lower triangle only; nameTempCovMatrix :341-352), the file `keep_temp`
        steps back is removed;
      * a run that finds such files restarts from the OLDEST one (findTempFile :265-341) at the next block pair;
      * the finished matrix (lower part mirrored, :537) is written to  folder/name.npy  and read back by later runs.
    block_source(first, nb) -> nb snapshot buffers. on_checkpoint(i, j) is called after each temp file is written.
    The matrix stays installed on the device as the Gram of pod.eig (block_gram_finish). Returns the m x m matrix."""
    import glob
    import re
    if norm not in ("L2", "Frobenius"):
        raise IthacaB200Error("The PODnorm can be only L2 or Frobenius")
    nk = 0 if norm == "L2" else 1
    L, h = pod.L, pod.h
    nCells = V.shape[0]
    os.makedirs(folder, exist_ok=True)
    final = os.path.join(folder, name + ".npy")
    pod.set_volumes(V)
    pod._chk(L.ithaca_b200_pod_block_begin(h, m, dim, nCells))
    pod.m, pod.dim = m, dim
    if os.path.exists(final):      # "Reading the covariance matrix" branch (:553-558)
        G = np.asfortranarray(np.load(final))
        if G.shape != (m, m):
            raise IthacaB200Error("buildCovMatrix: %s holds a %s matrix, expected %d x %d" % (final, G.shape, m, m))
        pod._chk(L.ithaca_b200_pod_block_gram_set(h, G.ctypes.data))
        pod._chk(L.ithaca_b200_pod_block_gram_finish(h, None))
        return G
    q, r = m // nBlocks, m % nBlocks
    if q == 0:
        raise IthacaB200Error("buildCovMatrix: more blocks than snapshots")
    blocks = [(b * q, q) for b in range(nBlocks)] + ([(nBlocks * q, r)] if r else [])
    done = np.zeros((m, m), dtype=bool)          # lower-triangle entries computed so far
    # restart point
    temps = []
    for f in glob.glob(os.path.join(folder, name + "_temp_*_*.npy")):
        mt = re.match(re.escape(name) + r"_temp_(\d+)_(\d+)\.npy$", os.path.basename(f))
        if mt:
            temps.append((int(mt.group(1)), int(mt.group(2)), f))
    etI, etJ = 0, 0
    if temps:
        ti, tj, tf = min(temps)
        C = np.load(tf)
        if C.shape != (m, m):
            raise IthacaB200Error("buildCovMatrix: temp file %s has the wrong size" % tf)
        done = np.tril(C != COV_UNREAL)
        Gl = np.where(done, C, 0.0)
        G0 = np.asfortranarray(Gl + np.tril(Gl, -1).T)      # the device keeps both halves of every computed block
        pod._chk(L.ithaca_b200_pod_block_gram_set(h, G0.ctypes.data))
        etI, etJ = (ti + 1, 0) if tj == ti - 1 else (ti, tj + 1)    # PODTemplate.C:452-466

    def up(slot, b):
        first, nb = blocks[b]
        bufs = block_source(first, nb)
        arr = (ctypes.c_void_p * nb)(*[_ptr(x) for x in bufs])
        pod._chk(L.ithaca_b200_pod_block_upload(h, slot, arr, nb, first))
        return bufs

    def mark(bi, bj):
        (i0, ni), (j0, nj) = blocks[bi], blocks[bj]
        done[i0:i0 + ni, j0:j0 + nj] = True

    saved = [(ti, tj) for ti, tj, _ in sorted(temps)]
    G = np.zeros((m, m), order="F")
    for bi in range(etI, len(blocks)):
        # the remainder block restarts its row from j = 0 as the reference does (:509-529)
        j_start = etJ if (bi == etI and bi < nBlocks) else 0
        keep = up(0, bi)
        pod._chk(L.ithaca_b200_pod_block_gram(h, 0, 0, nk))      # addCovMatrixTriCoeff
        mark(bi, bi)
        for bj in range(j_start, bi):
            keep2 = up(1, bj)
            pod._chk(L.ithaca_b200_pod_block_gram(h, 0, 1, nk))  # addCovMatrixSquareCoeff
            mark(bi, bj)
            pod._chk(L.ithaca_b200_pod_block_gram_get(h, G.ctypes.data))
            save_npy(np.where(np.tril(done), G, COV_UNREAL), os.path.join(folder, "%s_temp_%d_%d.npy" % (name, bi, bj)))
            saved.append((bi, bj))
            while len(saved) > keep_temp + 1:                     # deletePreviousTempCovMatrix_N
                oi, oj = saved.pop(0)
                old = os.path.join(folder, "%s_temp_%d_%d.npy" % (name, oi, oj))
                if os.path.exists(old):
                    os.remove(old)
            if on_checkpoint is not None:
                on_checkpoint(bi, bj)
    pod._chk(L.ithaca_b200_pod_block_gram_finish(h, G.ctypes.data))
    if not np.tril(done).sum() == m * (m + 1) // 2:
        raise IthacaB200Error("!!! OUPS !!! Unreal value [%s] found in triangular part of %s" % (COV_UNREAL, name))
    G = np.asfortranarray(np.tril(G) + np.tril(G, -1).T)          # covMatrix.selfadjointView<Eigen::Lower>()
    save_npy(G, final)
    for oi, oj in saved:
        old = os.path.join(folder, "%s_temp_%d_%d.npy" % (name, oi, oj))
        if os.path.exists(old):
            os.remove(old)
    return G


def check_nmodes(m, nmodes, solver):
    """nmodes rules of getModes (ITHACAPOD.C:109-130), same messages as the reference's M_Assert."""
    if solver == "spectra":
        if nmodes == 0:
            nmodes = m - 2
        if nmodes > m - 2:
            raise IthacaB200Error(
                "The number of requested modes cannot be bigger than the number of Snapshots - 2")
    elif solver == "eigen":
        if nmodes == 0:
            nmodes = m
        if nmodes > m:
            raise IthacaB200Error("The number of requested modes cannot be bigger than the number of Snapshots")
    else:
        raise IthacaB200Error("unknown EigenSolver '%s' (spectra|eigen)" % solver)
    return nmodes


def getModes(pod, snapshots, V, dim, nmodes=0, solver="spectra", norm="L2", Sbc=None):
    """Host mirror of the dense part of ITHACAPOD::getModes (ITHACAPOD.C:109-287) on a Pod handle.
    snapshots: Fortran (Nd, m) array or list of per-snapshot buffers; V: cell volumes of this rank.
    Returns dict(G, evals_norm, cum, evecs, normfact, modes, modes_bc) laid out like the reference's
    Eigen objects."""
    if norm not in ("L2", "Frobenius"):
        raise IthacaB200Error("The PODnorm can be only L2 or Frobenius")
    m = snapshots.shape[1] if isinstance(snapshots, np.ndarray) else len(snapshots)
    nmodes = check_nmodes(m, nmodes, solver)
    pod.set_volumes(V)
    pod.upload(snapshots, dim)
    G = pod.gram()
    ev_all, Q = pod.eig(nmodes)
    evn, cum = postprocess_eigenvalues(ev_all, nmodes, solver)
    Phi, nf = pod.lift(nmodes, norm)
    Pbc = pod.lift_patch(Sbc, nmodes) if Sbc is not None else None
    return dict(G=G, evals_all=ev_all, evals_norm=evn, cum=cum, evecs=Q, normfact=nf, modes=Phi, modes_bc=Pbc)
