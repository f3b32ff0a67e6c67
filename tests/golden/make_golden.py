"""Generates tests/golden/pod_*.npz from the reference's own expressions compiled on its vendored
Eigen/Spectra (oracle/_ref/libpod_ref.so, built from oracle/pod_ref.cpp against /root/reference).
Run in the build container only:  python tests/golden/make_golden.py
Inputs are regenerated from the seed at test time (oracle/synth.py); outputs are stored."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pod_oracle as po, synth  # noqa: E402

CASES = [
    # name, Nc, dim, m, r, seed, solver, norm, nbc
    ("scalar_spectra_L2", 2500, 1, 20, 6, 101, "spectra", "L2", 40),
    ("vector_eigen_L2", 1700, 3, 16, 5, 102, "eigen", "L2", 30),
    ("vector_spectra_frob", 1301, 3, 18, 7, 103, "spectra", "Frobenius", 0),
    ("scalar_eigen_frob_odd", 999, 1, 13, 13, 104, "eigen", "Frobenius", 12),
]


def inputs(Nc, dim, m, r, seed, nbc):
    S = synth.snapshots(Nc, dim, m, 2 * r, seed)
    V = synth.volumes(Nc, seed)
    Sbc = None
    if nbc:
        Sbc = np.asfortranarray(np.random.default_rng(seed + 99).standard_normal((nbc * dim, m)))
    return S, V, Sbc


if __name__ == "__main__":
    here = os.path.dirname(os.path.abspath(__file__))
    for name, Nc, dim, m, r, seed, solver, norm, nbc in CASES:
        S, V, Sbc = inputs(Nc, dim, m, r, seed, nbc)
        out = po.ref_getmodes(S, V, dim, r, solver, norm, Sbc)
        # modes are stored at a stride to keep the fixture small; G/evals/evecs/normfact in full
        stride = max(1, (Nc * dim) // 400)
        np.savez_compressed(
            os.path.join(here, "pod_%s.npz" % name), G=out["G"], evals=out["evals"], evals_norm=out["evals_norm"],
            cum=out["cum"], evecs=out["evecs"], normfact=out["normfact"], modes_sub=out["modes"][::stride],
            stride=stride, modes_bc=(out["modes_bc"] if nbc else np.zeros((0, r))),
            meta=np.array([Nc, dim, m, r, seed, nbc]), solver=solver, norm=norm)
        print("wrote", name, out["evals_norm"][:3])
