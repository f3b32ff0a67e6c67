"""GPU parity tests: the CUDA path, called through the C ABI, against the oracle on the same seeded inputs,
against the committed golden vectors, and through size-independent properties at larger sizes.
Tolerances are BASELINE.json's: eigenvalues 1e-10 relative, sign-aligned modes 1e-8 max relative; the Gram
itself is held to 1e-13 relative (max norm)."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "golden"))
from make_golden import CASES, inputs  # noqa: E402
from oracle import pod_oracle as po, synth  # noqa: E402

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def rel_max(A, B):
    return np.abs(A - B).max() / np.abs(B).max()


def modes_rel(A, B):
    A = po.sign_align(A, B)
    return (np.abs(A - B).max(axis=0) / np.abs(B).max(axis=0)).max()


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_getmodes_matches_golden(pkg, pod, case):
    name, Nc, dim, m, r, seed, solver, norm, nbc = case
    g = np.load(os.path.join(GOLD, "pod_%s.npz" % name))
    S, V, Sbc = inputs(Nc, dim, m, r, seed, nbc)
    out = pkg.getModes(pod, S, V, dim, r, solver, norm, Sbc)
    assert rel_max(out["G"], g["G"]) <= 1e-13
    assert len(out["evals_norm"]) == len(g["evals_norm"])
    assert np.all(np.abs(out["evals_norm"][:r] - g["evals_norm"][:r]) <= 1e-10 * g["evals_norm"][:r])
    assert np.abs(out["evals_norm"] - g["evals_norm"]).max() <= 1e-12 * g["evals_norm"][0]
    assert np.abs(out["cum"] - g["cum"]).max() <= 1e-12
    assert np.all(np.abs(out["normfact"] - g["normfact"]) <= 1e-9 * g["normfact"])
    st = int(g["stride"])
    assert modes_rel(out["modes"][::st], g["modes_sub"]) <= 1e-8
    if nbc:
        s = np.sign(np.einsum("ij,ij->j", out["modes"][::st], g["modes_sub"]))
        assert rel_max(out["modes_bc"] * s, g["modes_bc"]) <= 1e-8


@pytest.mark.parametrize("Nc,dim,m", [(1, 1, 3), (5, 3, 4), (17, 1, 8), (1000, 1, 40), (1000, 3, 41), (777, 3, 100),
                                       (3001, 1, 200), (2111, 3, 233), (12000, 1, 411)])
def test_gram_matches_oracle_ragged(pkg, pod, Nc, dim, m):
    """Every tile-edge case: m not a multiple of 8/40/200, Nd not a multiple of 16, single cell."""
    rng = np.random.default_rng(Nc * 31 + m)
    S = np.asfortranarray(rng.standard_normal((Nc * dim, m)))
    V = rng.random(Nc) + 0.1
    pod.set_volumes(V)
    pod.upload(S, dim)
    G = pod.gram()
    Gr = po.gram(S, V, dim)
    assert rel_max(G, Gr) <= 1e-13
    assert np.array_equal(G, G.T)  # symmetric by construction (quirk-free: the reference's is up to rounding)


def test_gram_vs_reference_build(pkg, pod):
    if not po.ref_available():
        pytest.skip("oracle/_ref not present")
    Nc, dim, m = 20000, 3, 64
    S = synth.snapshots(Nc, dim, m, 20, 3)
    V = synth.volumes(Nc, 3)
    pod.set_volumes(V)
    pod.upload(S, dim)
    G = pod.gram()
    assert rel_max(G, po.ref_gram(S, V, dim, 0)) <= 1e-13
    assert rel_max(G, po.ref_gram(S, V, dim, 1)) <= 1e-13


@pytest.mark.parametrize("m", [2, 3, 16, 57, 200, 333, 600, 1100, 2100])
def test_eig_matches_numpy(pkg, pod, m):
    """Every kernel variant: register-resident NE = 2..64 (m <= 2048) and the shared-memory block kernel (m = 2100)."""
    rng = np.random.default_rng(m)
    A = rng.standard_normal((m + 5, m))
    G = A.T @ A
    pod.set_gram(G)
    ev, Q = pod.eig(min(m, 104) if m > 400 else m)
    m_full, m = m, Q.shape[1]
    if m < m_full:
        w = np.linalg.eigvalsh(G)[::-1]
        assert np.abs(ev - w).max() <= 1e-12 * w[0]
        assert np.abs(Q.T @ Q - np.eye(m)).max() <= 1e-12
        assert np.abs(G @ Q - Q * ev[:m]).max() <= 1e-12 * w[0]
        return
    w = np.linalg.eigvalsh(G)[::-1]
    assert np.abs(ev - w).max() <= 1e-12 * w[0]
    assert np.abs(Q.T @ Q - np.eye(m)).max() <= 1e-12
    assert np.abs(G @ Q - Q * ev).max() <= 1e-12 * w[0]


def test_eig_indefinite_and_degenerate(pkg, pod):
    """Negative eigenvalues must come out signed (the reference then shifts, ITHACAPOD.C:182-186);
    repeated eigenvalues and an exactly singular matrix must converge."""
    rng = np.random.default_rng(9)
    Qo, _ = np.linalg.qr(rng.standard_normal((30, 30)))
    lam = np.array([5.0, 5.0, 3.0, 1.0, 1.0, 1.0] + [0.0] * 20 + [-1e-3, -0.5, -0.5, -2.0])
    G = (Qo * lam) @ Qo.T
    G = 0.5 * (G + G.T)
    pod.set_gram(G)
    ev, Q = pod.eig(30)
    assert np.abs(ev - np.sort(lam)[::-1]).max() <= 1e-13 * 5
    assert np.abs(Q.T @ Q - np.eye(30)).max() <= 1e-13
    evn, cum = pkg.postprocess_eigenvalues(ev, 30, "eigen")
    ref_ev, _ = po.eig(G, 30, "eigen")
    assert np.allclose(evn, ref_ev / ref_ev.sum(), rtol=0, atol=1e-13)


def test_lift_matches_reference_with_identical_eigenvectors(pkg, pod):
    """Lift + norms checked in isolation: feed the oracle the GPU's own eigenvectors."""
    Nc, dim, m, r = 5003, 3, 48, 20
    S = synth.snapshots(Nc, dim, m, 2 * r, 21)
    V = synth.volumes(Nc, 21)
    for norm in ("L2", "Frobenius"):
        out = pkg.getModes(pod, S, V, dim, r, "eigen", norm)
        Phi, nf, _ = po.lift(S, V, dim, out["evecs"], norm)
        assert np.all(np.abs(out["normfact"] - nf) <= 1e-12 * nf)
        assert (np.abs(out["modes"] - Phi).max(axis=0) / np.abs(Phi).max(axis=0)).max() <= 1e-11


@pytest.mark.parametrize("r", [1, 10, 24, 25, 50, 100])
def test_lift_all_column_block_sizes(pkg, pod, r):
    Nc, dim, m = 2000, 1, 120
    rng = np.random.default_rng(r)
    S = np.asfortranarray(rng.standard_normal((Nc, m)))
    V = rng.random(Nc) + 0.5
    out = pkg.getModes(pod, S, V, dim, r, "eigen", "L2")
    Phi, nf, _ = po.lift(S, V, dim, out["evecs"], "L2")
    assert np.all(np.abs(out["normfact"] - nf) <= 1e-12 * nf)
    assert (np.abs(out["modes"] - Phi).max(axis=0) / np.abs(Phi).max(axis=0)).max() <= 1e-11
    # modes are V-orthonormal (property of the method)
    Mm = out["modes"].T @ (out["modes"] * V[:, None])
    assert np.abs(Mm - np.eye(r)).max() <= 1e-9


def test_full_size_properties_c2_shape(pkg, pod):
    """BASELINE config 2 shape (2M cells x 200 scalar snapshots) through properties that need no oracle pass:
    a rank-2 snapshot set has the closed-form Gram  G = aa^T F11 + (ab^T + ba^T) F12 + bb^T F22."""
    Nc, m, r = 2_000_000, 200, 2
    rng = np.random.default_rng(77)
    x = np.linspace(0, 1, Nc)
    f, g2 = np.sin(3 * x) + 0.2, np.cos(5 * x)
    a, b = rng.standard_normal(m), rng.standard_normal(m)
    V = (1.0 + 0.5 * np.sin(11 * x)) / Nc
    S = np.asfortranarray(np.outer(f, a) + np.outer(g2, b))
    pod.set_volumes(V)
    pod.upload(S, 1)
    G = pod.gram()
    F11, F12, F22 = (V * f * f).sum(), (V * f * g2).sum(), (V * g2 * g2).sum()
    Gx = np.outer(a, a) * F11 + (np.outer(a, b) + np.outer(b, a)) * F12 + np.outer(b, b) * F22
    assert rel_max(G, Gx) <= 1e-12
    ev, Q = pod.eig(r)
    w = np.linalg.eigvalsh(Gx)[::-1]
    assert np.all(np.abs(ev[:2] - w[:2]) <= 1e-10 * w[:2])
    assert np.abs(ev[2:]).max() <= 1e-12 * w[0]
    Phi, nf = pod.lift(r, "L2")
    assert np.abs(Phi.T @ (Phi * V[:, None]) - np.eye(2)).max() <= 1e-10
    # modes span {f, g}: residual of projecting f on the modes vanishes
    c = Phi.T @ (V * f)
    assert np.abs(Phi @ c - f).max() <= 1e-9 * np.abs(f).max()


def test_error_paths(pkg):
    p = pkg.Pod(0)
    with pytest.raises(pkg.IthacaB200Error):
        p.gram()  # nothing uploaded
    V = np.ones(10)
    p.set_volumes(V)
    with pytest.raises(pkg.IthacaB200Error):
        p.upload(np.asfortranarray(np.ones((12, 3))), 1)  # nCells mismatch
    p.upload(np.asfortranarray(np.random.default_rng(0).standard_normal((10, 3))), 1)
    p.gram()
    with pytest.raises(pkg.IthacaB200Error):
        p.eig(4)  # more modes than snapshots
    p.eig(3)
    with pytest.raises(pkg.IthacaB200Error):
        p.lift(2, "H1")
    p.close()


@pytest.mark.parametrize("dim,solver", [(3, "spectra"), (3, "eigen"), (1, "spectra"), (1, "eigen")])
def test_c1_tutorial_scale_matches_oracle(pkg, pod, dim, solver):
    """BASELINE config C1: the 01POD cavity, 10 000 cells of 1e-8 m^3 each
    (tutorials/CFD/01POD/system/blockMeshDict: uniform 100x100x1 block of a 0.1 x 0.1 x 0.01 m box), 100 U / p
    snapshots, 10 modes, both EigenSolver settings, L2 norm."""
    Nc, m, r = 10_000, 100, 10
    S = synth.snapshots(Nc, dim, m, 2 * r, 20251104 + 1)
    V = np.full(Nc, 1e-8)
    out = pkg.getModes(pod, S, V, dim, r, solver, "L2")
    ref = po.getmodes(S, V, dim, r, solver, "L2")
    assert rel_max(out["G"], ref["G"]) <= 1e-13
    assert len(out["evals_norm"]) == len(ref["evals_norm"]) == (r if solver == "spectra" else m)
    assert np.all(np.abs(out["evals_norm"][:r] - ref["evals_norm"][:r]) <= 1e-10 * ref["evals_norm"][:r])
    assert np.abs(out["cum"] - ref["cum"]).max() <= 1e-12
    assert modes_rel(out["modes"], ref["modes"]) <= 1e-8


@pytest.mark.parametrize("Nc,dim,m,r", [(4099, 3, 37, 6), (10_000, 1, 100, 10), (777, 3, 9, 9)])
def test_project_getcoeffs_reconstruct_match_oracle(pkg, pod, Nc, dim, m, r):
    """Modes::project (G / F), getMassMatrix over modes, getCoeffs, Modes::reconstruct on the
    device-resident snapshots and modes, against the numpy restatement (pinned to the compiled reference expressions by
    tests/test_oracle_pod.py)."""
    S = synth.snapshots(Nc, dim, m, 2 * r, 77)
    V = synth.volumes(Nc, 77)
    ref = po.getmodes(S, V, dim, r, "eigen", "L2")
    out = pkg.getModes(pod, S, V, dim, r, "eigen", "L2")          # leaves snapshots + lifted modes resident
    Phi = out["modes"]
    for nm in (0, max(1, r // 2)):
        for t in ("G", "F"):
            assert rel_max(pod.project(nm, t), po.project(S, V, dim, Phi, nm, t)) <= 1e-12
        assert rel_max(pod.modes_mass(nm), po.modes_mass(Phi, V, dim, nm)) <= 1e-12
        assert rel_max(pod.get_coeffs(nm), po.get_coeffs(S, V, dim, Phi, nm)) <= 1e-10
    # externally supplied modes (not orthonormal: the mass matrix solve matters)
    rng = np.random.default_rng(5)
    W = np.asfortranarray(Phi @ (np.eye(r) + 0.3 * rng.standard_normal((r, r))))
    pod.set_modes(W)
    assert rel_max(pod.get_coeffs(0), po.get_coeffs(S, V, dim, W, 0)) <= 1e-10
    coeff = rng.standard_normal((r, 11))
    assert rel_max(pod.reconstruct(coeff), po.reconstruct(W, coeff)) <= 1e-13
    clipped = pod.reconstruct(coeff[:2, :3], clip_negative=True)
    assert np.array_equal(clipped, po.reconstruct(W, coeff[:2, :3], clip=True)) or rel_max(clipped, po.reconstruct(W, coeff[:2, :3], clip=True)) <= 1e-13
    assert clipped.min() >= 0.0
    # projecting the snapshots and reconstructing from all m = rank modes returns them (round trip)
    if r == m:
        pod.set_modes(np.asfortranarray(Phi))
        back = pod.reconstruct(pod.get_coeffs(0))
        assert rel_max(back, S) <= 1e-8
    with pytest.raises(pkg.IthacaB200Error):
        pod.get_coeffs(r + 1)


@pytest.mark.parametrize("Nc,dim,m,r,block,solver,norm", [(4099, 3, 37, 6, 16, "eigen", "L2"), (3001, 1, 50, 8, 7, "spectra", "Frobenius"),
                                                          (777, 3, 20, 5, 64, "eigen", "L2")])
def test_memory_efficient_blocked_pod_matches_oracle(pkg, pod, Nc, dim, m, r, block, solver, norm):
    """getModesMemoryEfficient (ITHACAPOD.C:344-628) with the snapshots streamed in blocks (two
    resident at a time); ragged last block, one-block case, both norms (Frobenius: unweighted correlation here)."""
    S = synth.snapshots(Nc, dim, m, 2 * r, 91)
    V = synth.volumes(Nc, 91)
    calls = []

    def source(first, nb):
        calls.append((first, nb))
        return [np.ascontiguousarray(S[:, j]) for j in range(first, first + nb)]

    out = pkg.getModesMemoryEfficient(pod, source, m, V, dim, r, solver, norm, block=block)
    ref = po.getmodes_memory_efficient(S, V, dim, r, solver, norm)
    assert rel_max(out["G"], ref["G"]) <= 1e-13
    assert np.array_equal(out["G"], out["G"].T)
    assert len(out["evals_norm"]) == len(ref["evals_norm"])
    assert np.all(np.abs(out["evals_norm"][:r] - ref["evals_norm"][:r]) <= 1e-10 * ref["evals_norm"][:r])
    assert np.all(np.abs(out["normfact"] - ref["normfact"]) <= 1e-9 * ref["normfact"])
    assert modes_rel(out["modes"], ref["modes"]) <= 1e-8
    assert max(nb for _, nb in calls) <= block
    if norm == "L2":   # same numbers as the resident path
        res = pkg.getModes(pod, S, V, dim, r, solver, norm)
        assert rel_max(out["G"], res["G"]) <= 1e-13 and modes_rel(out["modes"], res["modes"]) <= 1e-8
    # the resident-mode follow-ups work on the blocked result too
    pod.upload(S, dim)
    pod.set_modes(np.asfortranarray(out["modes"]))
    assert rel_max(pod.project(0, "G"), po.project(S, V, dim, out["modes"], 0, "G")) <= 1e-12


@pytest.mark.parametrize("solver", ["spectra", "eigen"])
def test_weighted_modes_match_oracle(pkg, pod, solver):
    """getWeightedModes (ITHACAPOD.C:715-845): modes_i = (S q_i) lambda_i^(-1/2) lambda_i, boundary rows alike."""
    Nc, dim, m, r = 3001, 3, 30, 6
    S = synth.snapshots(Nc, dim, m, 2 * r, 55)
    V = synth.volumes(Nc, 55)
    Sbc = np.asfortranarray(np.random.default_rng(3).standard_normal((21, m)))
    out = pkg.getWeightedModes(pod, S, V, dim, r, solver, Sbc)
    ref = po.getweightedmodes(S, V, dim, r, solver, Sbc)
    assert len(out["evals_norm"]) == len(ref["evals_norm"])
    assert np.all(np.abs(out["evals_norm"][:r] - ref["evals_norm"][:r]) <= 1e-10 * ref["evals_norm"][:r])
    assert modes_rel(out["modes"], ref["modes"]) <= 1e-8
    s = np.sign(np.einsum("ij,ij->j", out["modes"], ref["modes"]))
    assert rel_max(out["modes_bc"] * s, ref["modes_bc"]) <= 1e-8
    with pytest.raises(pkg.IthacaB200Error):
        pkg.getWeightedModes(pod, S, V, dim, m - 1, solver)


def test_svd_modes_match_oracle(pkg, pod):
    """getModesSVD (ITHACAPOD.C:858-941) through Gram + eigensolve + scaled lift against a direct thin SVD of
    diag(sqrt V) S (numpy) and, when built, the reference's own JacobiSVD expression."""
    Nc, dim, m, r = 2503, 3, 24, 6
    S = synth.snapshots(Nc, dim, m, 2 * r, 77)
    V = synth.volumes(Nc, 77)
    out = pkg.getModesSVD(pod, S, V, dim, r)
    refs = [po.getmodes_svd(S, V, dim, r)] + ([po.ref_getmodes_svd(S, V, dim, r)] if po.ref_available() else [])
    for ref in refs:
        assert len(out["evals_norm"]) == len(ref["evals_norm"]) == m
        # leading singular values; the noise-floor ones (sigma ~ 1e-6 sigma_0) lose half the digits to the squaring
        assert np.all(np.abs(out["evals_norm"][:r] - ref["evals_norm"][:r]) <= 1e-10 * ref["evals_norm"][:r])
        assert np.abs(out["evals_norm"] - ref["evals_norm"]).max() <= 1e-9 * ref["evals_norm"][0]
        assert modes_rel(out["modes"], ref["modes"]) <= 1e-8
    with pytest.raises(pkg.IthacaB200Error):
        pkg.getModesSVD(pod, S, V, dim, m + 1)


@pytest.mark.parametrize("solver", ["spectra", "eigen"])
def test_per_snapshot_volume_modes_match_oracle(pkg, pod, solver):
    """getModes with one volume field per snapshot (moving meshes, ITHACAPOD.C:1392-1541): correlation matrix as one
    contraction of the row-scaled snapshots, modes = (S Q) lambda^-1/2."""
    Nc, m, r = 4001, 18, 5
    S = synth.snapshots(Nc, 1, m, 2 * r, 88)
    rng = np.random.default_rng(88)
    V0 = synth.volumes(Nc, 88)
    Vs = np.asfortranarray(V0[:, None] * (1.0 + 0.2 * rng.uniform(-1, 1, (Nc, m))))
    Vs[::97, 3] *= -1.0   # the reference takes |V_i V_j|
    Sbc = np.asfortranarray(rng.standard_normal((17, m)))
    out = pkg.getModesVolumes(pod, S, Vs, r, solver, Sbc)
    refs = [po.getmodes_volumes(S, Vs, r, solver)] + ([po.ref_getmodes_volumes(S, Vs, r, solver)] if po.ref_available() else [])
    for ref in refs:
        assert rel_max(out["G"], ref["G"]) <= 1e-13
        assert np.all(np.abs(out["evals_norm"] - ref["evals_norm"]) <= 1e-10 * ref["evals_norm"])
        assert modes_rel(out["modes"], ref["modes"]) <= 1e-8
    ref = refs[0]
    s = np.sign(np.einsum("ij,ij->j", out["modes"], ref["modes"]))
    assert rel_max(out["modes_bc"] * s, (Sbc @ ref["evecs"]) / np.sqrt(ref["evals"])) <= 1e-8


def test_deim_modes_are_getmodes(pkg, pod):
    """DEIMmodes (ITHACAPOD.C:1739-1937) shares getModes' arithmetic line for line."""
    Nc, dim, m, r = 1501, 1, 15, 4
    S = synth.snapshots(Nc, dim, m, 2 * r, 91)
    V = synth.volumes(Nc, 91)
    out = pkg.DEIMmodes(pod, S, V, dim, r, "eigen", "Frobenius")
    ref = po.getmodes(S, V, dim, r, "eigen", "Frobenius")
    assert np.all(np.abs(out["evals_norm"][:r] - ref["evals_norm"][:r]) <= 1e-10 * ref["evals_norm"][:r])
    assert modes_rel(out["modes"], ref["modes"]) <= 1e-8


def test_blocked_cov_matrix_checkpoint_and_resume(pkg, pod, tmp_path):
    """PODTemplate::buildCovMatrix's checkpoint / resume protocol (PODTemplate.C:419-616): an interrupted run leaves
    covMatrix<field>_temp_<i>_<j>.npy files, the next run restarts from the oldest of them at the following block pair and
    ends with the same matrix; the finished covMatrix<field>.npy is read back by later runs without touching the snapshots."""
    Nc, dim, m, nBlocks = 2003, 3, 23, 4          # q = 5, remainder block of 3
    S = synth.snapshots(Nc, dim, m, 8, 17)
    V = synth.volumes(Nc, 17)
    folder, name = str(tmp_path / "CovMatrices") + "/", "covMatrixU"
    ref = po.gram(S, V, dim)
    reads = []

    def source(first, nb):
        reads.append((first, nb))
        return [np.ascontiguousarray(S[:, j]) for j in range(first, first + nb)]

    class Crash(Exception):
        pass

    def crash_at(i, j):
        if (i, j) == (3, 1):
            raise Crash()

    with pytest.raises(Crash):
        pkg.buildCovMatrix(pod, source, m, V, dim, nBlocks, folder, name, on_checkpoint=crash_at)
    left = sorted(os.listdir(folder))
    # pairs saved: (1,0) (2,0) (2,1) (3,0) (3,1); the last keep_temp + 1 = 4 files stay
    assert left == [name + "_temp_%d_%d.npy" % ij for ij in [(2, 0), (2, 1), (3, 0), (3, 1)]]
    T = np.load(folder + name + "_temp_2_0.npy")
    assert T.shape == (m, m) and T[0, 1] == pkg.COV_UNREAL and T[12, 14] == pkg.COV_UNREAL      # upper half / block (2,2)'s upper
    assert T[12, 3] != pkg.COV_UNREAL and T[12, 7] == pkg.COV_UNREAL and T[17, 0] == pkg.COV_UNREAL  # (2,0) done, (2,1), (3,0) not
    assert abs(T[12, 3] - ref[12, 3]) <= 1e-13 * np.abs(ref).max()
    reads.clear()
    G = pkg.buildCovMatrix(pod, source, m, V, dim, nBlocks, folder, name)
    # restart from the oldest file (2,0): the next pair is (2,1); blocks 0 and 1 are never read as the row block again
    assert reads[0] == (10, 5) and reads[1] == (5, 5)
    assert rel_max(G, ref) <= 1e-13 and np.array_equal(G, G.T)
    assert sorted(os.listdir(folder)) == [name + ".npy"]
    assert np.array_equal(np.load(folder + name + ".npy"), G)
    # the matrix is installed as the Gram of the eigensolve
    ev, Q = pod.eig(5)
    ref_ev = np.linalg.eigvalsh(ref)[::-1]
    assert np.all(np.abs(ev[:5] - ref_ev[:5]) <= 1e-10 * ref_ev[0])
    reads.clear()
    G2 = pkg.buildCovMatrix(pod, source, m, V, dim, nBlocks, folder, name)
    assert reads == [] and np.array_equal(G2, G)


def test_nested_snapshot_matrix_cormatrix_and_streamed_mean(pkg, pod):
    """getNestedSnapshotMatrix (ITHACAPOD.C:41-82), corMatrix (:955-1008) and getMeanMemoryEfficient (:660-697): host loops of
    the reference over kernels that are already on the path."""
    Nc, dim, Npar, Nt, nout = 1501, 3, 3, 12, 4
    m = Npar * Nt
    S = synth.snapshots(Nc, dim, m, 10, 33)
    V = synth.volumes(Nc, 33)
    out = pkg.getNestedSnapshotMatrix(pod, S, V, dim, Npar, nout, "eigen")
    assert out.shape == (Nc * dim, Npar * nout)
    for i in range(Npar):
        ref = po.getweightedmodes(np.asfortranarray(S[:, Nt * i:Nt * (i + 1)]), V, dim, nout, "eigen", None)
        assert modes_rel(out[:, nout * i:nout * (i + 1)], ref["modes"]) <= 1e-8
    assert rel_max(pkg.corMatrix(pod, S, V, dim), po.gram(S, V, dim)) <= 1e-13
    mean = pkg.getMeanMemoryEfficient(pod, lambda first, nb: [np.ascontiguousarray(S[:, j]) for j in range(first, first + nb)],
                                      m, V, dim, block=7)
    assert np.abs(mean - S.mean(axis=1)).max() <= 1e-13 * np.abs(S).max()


@pytest.mark.parametrize("m", [5, 57, 128, 129])
def test_small_matrix_cholesky_matches_the_blocked_one(pkg, pod, m, monkeypatch):
    """m <= 128: init + Cholesky run as ONE single-CTA launch in shared memory (chol_small_kernel); ITHACA_B200_CHOL=blocked
    keeps the multi-launch blocked factorisation. Same eigenpairs either way (129 takes the blocked path regardless)."""
    rng = np.random.default_rng(100 + m)
    A = rng.standard_normal((m + 3, m))
    G = A.T @ A
    w = np.linalg.eigvalsh(G)[::-1]
    out = {}
    for mode in ("small", "blocked"):
        if mode == "blocked":
            monkeypatch.setenv("ITHACA_B200_CHOL", "blocked")
        else:
            monkeypatch.delenv("ITHACA_B200_CHOL", raising=False)
        pod.set_gram(G)
        ev, Q = pod.eig(m)
        assert np.abs(ev - w).max() <= 1e-12 * w[0]
        assert np.abs(G @ Q - Q * ev).max() <= 1e-12 * w[0]
        out[mode] = ev
    assert np.abs(out["small"] - out["blocked"]).max() <= 1e-13 * w[0]
