"""CPU tests: the numpy restatement (oracle/pod_oracle.py) against the committed golden vectors produced
by the reference's own Eigen/Spectra expressions, and against oracle/_ref directly when it is present."""
import glob
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "golden"))
from make_golden import CASES, SHAPE_CASES, inputs  # noqa: E402
from oracle import pod_oracle as po, synth  # noqa: E402

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def load(name):
    return np.load(os.path.join(GOLD, "pod_%s.npz" % name), allow_pickle=False)


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_numpy_oracle_matches_golden(case):
    name, Nc, dim, m, r, seed, solver, norm, nbc = case
    g = load(name)
    S, V, Sbc = inputs(Nc, dim, m, r, seed, nbc)
    out = po.getmodes(S, V, dim, r, solver, norm, Sbc)
    assert np.abs(out["G"] - g["G"]).max() <= 1e-13 * np.abs(g["G"]).max()
    assert len(out["evals_norm"]) == len(g["evals_norm"])  # quirk Q4: m values (eigen) or nmodes (spectra)
    # eigenvalues: 1e-10 relative on the kept modes; absolute (relative to lambda_1) on the rest
    assert np.all(np.abs(out["evals_norm"][:r] - g["evals_norm"][:r]) <= 1e-10 * g["evals_norm"][:r])
    assert np.abs(out["evals_norm"] - g["evals_norm"]).max() <= 1e-12 * g["evals_norm"][0]
    assert np.abs(out["cum"] - g["cum"]).max() <= 1e-12
    assert np.all(np.abs(out["normfact"] - g["normfact"]) <= 1e-9 * g["normfact"])
    st = int(g["stride"])
    A = po.sign_align(out["modes"], np.zeros_like(out["modes"]) + 0)  # no-op, keeps shape
    sub = out["modes"][::st]
    sub = po.sign_align(sub, g["modes_sub"])
    rel = np.abs(sub - g["modes_sub"]).max(axis=0) / np.abs(g["modes_sub"]).max(axis=0)
    assert rel.max() <= 1e-8, rel
    if nbc:
        s = np.sign(np.einsum("ij,ij->j", out["modes"][::st], g["modes_sub"]))
        rel = np.abs(out["modes_bc"] * s - g["modes_bc"]).max(axis=0) / np.abs(g["modes_bc"]).max(axis=0)
        assert rel.max() <= 1e-8


@pytest.mark.skipif(not po.ref_available(), reason="oracle/_ref not built (reference tree absent)")
def test_numpy_oracle_matches_reference_build_both_gram_branches():
    Nc, dim, m = 4000, 3, 15
    S = synth.snapshots(Nc, dim, m, 10, 5)
    V = synth.volumes(Nc, 5)
    G = po.gram(S, V, dim)
    for branch in (0, 1):  # GEMM branch and the V.size()>1e6 rank-1 loop (ITHACAcoeffsMass.C:95-104)
        Gr = po.ref_gram(S, V, dim, branch)
        assert np.abs(G - Gr).max() <= 1e-13 * np.abs(Gr).max()


@pytest.mark.skipif(not po.ref_available(), reason="oracle/_ref not built (reference tree absent)")
def test_golden_is_reproducible_from_reference_build():
    name, Nc, dim, m, r, seed, solver, norm, nbc = CASES[1]
    g = load(name)
    S, V, Sbc = inputs(Nc, dim, m, r, seed, nbc)
    out = po.ref_getmodes(S, V, dim, r, solver, norm, Sbc)
    assert np.array_equal(out["G"], g["G"])
    assert np.array_equal(out["evals_norm"], g["evals_norm"])


def test_nmodes_rules():
    assert po.check_nmodes(20, 0, "spectra") == 18
    assert po.check_nmodes(20, 0, "eigen") == 20
    with pytest.raises(ValueError):
        po.check_nmodes(20, 19, "spectra")
    with pytest.raises(ValueError):
        po.check_nmodes(20, 21, "eigen")


def test_negative_shift_rule():
    # ITHACAPOD.C:182-186: lambda += 2|min| when the smallest kept value is negative
    G = np.diag([3.0, 1.0, -0.5])
    ev, _ = po.eig(G, 2, "eigen")
    assert np.allclose(ev, [4.0, 2.0, 0.5])
    ev, _ = po.eig(G, 2, "spectra")  # only the kept values are inspected: no shift
    assert np.allclose(ev, [3.0, 1.0])


@pytest.mark.skipif(not po.ref_available(), reason="oracle/_ref not built")
def test_project_getcoeffs_reconstruct_restatement_matches_reference_build():
    """The numpy restatement of getCoeffs / Modes::project / Modes::reconstruct equals the reference's own Eigen
    expressions (oracle/pod_ref.cpp: ITHACAcoeffsMass.C:348-443, Modes.C:228-305) compiled on the vendored Eigen."""
    Nc, dim, m, r = 501, 3, 12, 5
    S = synth.snapshots(Nc, dim, m, 2 * r, 3)
    V = synth.volumes(Nc, 3)
    Phi = po.getmodes(S, V, dim, r, "eigen", "L2")["modes"]
    rng = np.random.default_rng(2)
    W = Phi @ (np.eye(r) + 0.3 * rng.standard_normal((r, r)))   # non-orthonormal modes: the fullPivLu solve matters
    for nm in (0, 3):
        a, b = po.get_coeffs(S, V, dim, W, nm), po.ref_get_coeffs(S, V, dim, W, nm)
        assert np.abs(a - b).max() <= 1e-12 * np.abs(b).max()
        for t in ("G", "F"):
            a, b = po.project(S, V, dim, W, nm, t), po.ref_project(S, V, dim, W, nm, t)
            assert np.abs(a - b).max() <= 1e-13 * np.abs(b).max()
    c = rng.standard_normal((4, 3))
    for clip in (False, True):
        a, b = po.reconstruct(W, c, clip), po.ref_reconstruct(W, c, clip)
        assert np.abs(a - b).max() <= 1e-14 * np.abs(b).max()


def test_svd_and_per_snapshot_volume_variants_match_golden_and_reference_build():
    """getModesSVD (ITHACAPOD.C:858-941) and getModes with per-snapshot Volumes (:1392-1541): numpy restatement against
    the committed outputs of the reference's own expressions (JacobiSVD / the weighted double loop on the vendored Eigen),
    and against oracle/_ref directly when present."""
    from make_golden import variant_inputs
    Nc, dim, m, r, S, V, Ss, Vs = variant_inputs()
    g = np.load(os.path.join(GOLD, "pod_variants.npz"))
    a = po.getmodes_svd(S, V, dim, r)
    assert np.abs(a["evals_norm"] - g["svd_evals_norm"]).max() <= 1e-13
    assert np.abs(a["cum"] - g["svd_cum"]).max() <= 1e-12
    sub = a["modes"][::9]
    sgn = np.sign(np.einsum("ij,ij->j", sub, g["svd_modes_sub"]))
    assert np.abs(sub * sgn - g["svd_modes_sub"]).max() <= 1e-9 * np.abs(g["svd_modes_sub"]).max()
    b = po.getmodes_volumes(Ss, Vs, r, "eigen")
    assert np.abs(b["G"] - g["vol_G"]).max() <= 1e-13 * np.abs(g["vol_G"]).max()
    assert np.abs(b["evals_norm"] - g["vol_evals_norm"]).max() <= 1e-12
    sub = b["modes"][::3]
    sgn = np.sign(np.einsum("ij,ij->j", sub, g["vol_modes_sub"]))
    assert np.abs(sub * sgn - g["vol_modes_sub"]).max() <= 1e-9 * np.abs(g["vol_modes_sub"]).max()
    if po.ref_available():
        c = po.ref_getmodes_volumes(Ss, Vs, r, "spectra")
        d = po.getmodes_volumes(Ss, Vs, r, "spectra")
        assert np.abs(c["evals_norm"] - d["evals_norm"]).max() <= 1e-12
        e = po.ref_getmodes_svd(S, V, dim, r)
        assert np.abs(e["evals_norm"] - g["svd_evals_norm"]).max() <= 1e-14


@pytest.mark.parametrize("case", SHAPE_CASES, ids=[c[0] for c in SHAPE_CASES])
def test_numpy_oracle_matches_reference_at_benchmark_shapes(case):
    """m = 1000 vector (C3) and m = 2000 scalar (C5) on few cells: the restatement against the reference's stored results."""
    name, Nc, dim, m, r, seed, solver, gs = case
    g = load(name)
    S, V, _ = inputs(Nc, dim, m, r, seed, 0)
    out = po.getmodes(S, V, dim, r, solver, "L2")
    assert np.abs(out["G"][::gs, ::gs] - g["G_sub"]).max() <= 1e-13 * np.abs(g["G_sub"]).max()
    assert np.all(np.abs(out["evals_norm"][:r] - g["evals_norm"][:r]) <= 1e-10 * g["evals_norm"][:r])
    assert np.all(np.abs(out["normfact"] - g["normfact"]) <= 1e-9 * g["normfact"])
    sub = po.sign_align(out["modes"][::int(g["stride"])], g["modes_sub"])
    assert (np.abs(sub - g["modes_sub"]).max(axis=0) / np.abs(g["modes_sub"]).max(axis=0)).max() <= 1e-8


def test_getmodes_fields2_definition():
    """getModes(snapshots, modes, fields2, ...) restatement (ITHACAPOD.C:1940-2086): with snapshots == fields2 the modes
    (S Q) lambda^-1/2 are V-orthonormal; a second list only changes what is combined, not the eigenpairs."""
    Nc, m, r = 700, 12, 4
    S2 = synth.snapshots(Nc, 3, m, 6, 1)
    V = synth.volumes(Nc, 1)
    o = po.getmodes_fields2(S2, S2, V, 3, r, "eigen")
    W = np.repeat(V, 3)
    assert np.abs(o["modes"].T @ (W[:, None] * o["modes"]) - np.eye(r)).max() <= 1e-12
    S = np.random.default_rng(0).standard_normal((333, m))
    o2 = po.getmodes_fields2(S, S2, V, 3, r, "spectra")
    assert np.array_equal(o2["evals"], o["evals"]) and len(o2["evals_norm"]) == r
    assert np.abs(o2["modes"] - (S @ o["evecs"]) / np.sqrt(o["evals"])).max() <= 1e-12 * np.abs(o2["modes"]).max()
    assert abs(o2["evals_norm"].sum() - 1.0) <= 1e-14
