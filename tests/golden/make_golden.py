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


# the benchmark shapes (C3: m = 1000 vector snapshots, 50 modes; C5: m = 2000 scalar, 100 modes) on few cells: the 25- and
# 50-tile-row Gram schedules, eig_jacobi_reg_kernel<32/64>, lift_kernel<7/13>. Outputs stored at a stride (G: every gs-th
# row and column), the reference takes 8 s / 22 s per case, the test regenerates the inputs from the seed.
SHAPE_CASES = [
    # name, Nc, dim, m, r, seed, solver, gs
    ("shape_c3", 6007, 3, 1000, 50, 77, "spectra", 7),
    ("shape_c5", 6007, 1, 2000, 100, 78, "spectra", 13),
]


def variant_inputs():
    """Inputs of the getModesSVD / per-snapshot-volume getModes fixture (pod_variants.npz)."""
    Nc, dim, m, r = 1203, 3, 14, 4
    S = synth.snapshots(Nc, dim, m, 2 * r, 105)
    V = synth.volumes(Nc, 105)
    Ss = synth.snapshots(Nc, 1, m, 2 * r, 106)
    rng = np.random.default_rng(106)
    Vs = np.asfortranarray(V[:, None] * (1.0 + 0.2 * rng.uniform(-1, 1, (Nc, m))))
    Vs[::53, 2] *= -1.0
    return Nc, dim, m, r, S, V, Ss, Vs


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

    for name, Nc, dim, m, r, seed, solver, gs in SHAPE_CASES:
        S, V, _ = inputs(Nc, dim, m, r, seed, 0)
        out = po.ref_getmodes(S, V, dim, r, solver, "L2")
        stride = max(1, (Nc * dim) // 200)
        np.savez_compressed(os.path.join(here, "pod_%s.npz" % name), G_sub=out["G"][::gs, ::gs], G_trace=np.trace(out["G"]),
                            evals_norm=out["evals_norm"], cum=out["cum"], normfact=out["normfact"],
                            modes_sub=out["modes"][::stride], stride=stride, gs=gs)
        print("wrote", name, out["evals_norm"][:3])

    Nc, dim, m, r, S, V, Ss, Vs = variant_inputs()
    a = po.ref_getmodes_svd(S, V, dim, r)
    b = po.ref_getmodes_volumes(Ss, Vs, r, "eigen")
    np.savez_compressed(os.path.join(here, "pod_variants.npz"), svd_evals_norm=a["evals_norm"], svd_cum=a["cum"],
                        svd_modes_sub=a["modes"][::9], vol_G=b["G"], vol_evals_norm=b["evals_norm"], vol_cum=b["cum"],
                        vol_modes_sub=b["modes"][::3])
    print("wrote variants", a["evals_norm"][:3], b["evals_norm"][:3])
