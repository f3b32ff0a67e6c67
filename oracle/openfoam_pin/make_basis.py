"""Writes basis.bin for pinProjection (run in this repo, any machine): the seeded synthetic basis of the parity tests
(oracle/fvc.py::synthetic_basis, seed 11, N = 7, Np = 3 = tests/test_proj_gpu.py's tutorial-04 case) on the tutorial-04 mesh,
boundary values split per patch in OpenFOAM's patch order (empty patches write nothing)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import fvc, polymesh as pm  # noqa: E402

mesh = pm.load_compact(os.path.join(ROOT, "tests", "golden", "mesh_04unsteadyNS.npz"))
N, Np = 7, 3
U, Ub, P, Pb = fvc.synthetic_basis(mesh, N, Np, seed=11)
with open("basis.bin", "wb") as f:
    f.write(np.array([N, Np], dtype=np.int64).tobytes())
    for fields, bnds, nc in ((U, Ub, 3), (P, Pb, 1)):
        for a, b in zip(fields, bnds):
            f.write(np.ascontiguousarray(a, dtype=np.float64).tobytes())
            b = np.asarray(b, dtype=np.float64).reshape(-1, nc)
            for name, typ, sl in mesh.patch_slices():
                if typ != "empty":
                    f.write(np.ascontiguousarray(b[sl]).tobytes())
print("wrote basis.bin: N = %d, Np = %d, %d cells" % (N, Np, mesh.nCells))
