"""GPU parity tests added in round 2: more modes than one lift launch holds (nmodes = 0 -> m - 2 / m, ITHACAPOD.C:111-130),
boundary rows after the snapshot-blocked lift, the row-streamed path (pod_stream_*), the benchmark shapes
(m = 1000 vector = C3, m = 2000 scalar = C5) against the compiled reference arithmetic, and the device-side synthetic generator.
Tolerances are BASELINE.json's: Gram 1e-13, eigenvalues 1e-10 relative, sign-aligned modes 1e-8 max relative."""
import ctypes
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "golden"))
from make_golden import SHAPE_CASES, inputs  # noqa: E402
from oracle import pod_oracle as po, synth  # noqa: E402

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def rel_max(A, B):
    return np.abs(A - B).max() / np.abs(B).max()


def modes_rel(A, B):
    A = po.sign_align(A, B)
    return (np.abs(A - B).max(axis=0) / np.abs(B).max(axis=0)).max()


def well_separated(Nc, dim, m, seed, floor=1e-3):
    """Snapshots whose Gram has ALL m eigenvalues separated (relative gaps >= ~1/m): every mode is well defined, so
    the 1e-8 gate applies to all of them (nmodes = 0 keeps m - 2 or m modes)."""
    rng = np.random.default_rng(seed)
    Nd = Nc * dim
    U, _ = np.linalg.qr(rng.standard_normal((Nd, m)))
    W, _ = np.linalg.qr(rng.standard_normal((m, m)))
    sig = np.linspace(1.0, floor ** 0.5, m) * (1.0 + 0.05 * rng.random(m) / m)
    V = rng.random(Nc) + 0.5
    S = (U / np.sqrt(np.repeat(V, dim))[:, None] * sig) @ W.T      # S^T V S = W diag(sig^2) W^T
    return np.asfortranarray(S), V


@pytest.mark.parametrize("solver,norm", [("spectra", "L2"), ("eigen", "Frobenius"), ("eigen", "L2")])
def test_getmodes_default_nmodes_more_than_one_lift_launch(pkg, pod, solver, norm):
    """getModes(..., nmodes = 0) on 200 snapshots: 198 (spectra) or 200 (eigen) modes, the lift runs as column chunks
    of <= 104 (lift_ensure refused r > 104). Internal field and boundary rows against the oracle."""
    Nc, dim, m = 1501, 3, 200
    S, V = well_separated(Nc, dim, m, 5)
    rng = np.random.default_rng(1)
    Sbc = np.asfortranarray(rng.standard_normal((37, m)))
    out = pkg.getModes(pod, S, V, dim, 0, solver, norm, Sbc)
    ref = po.getmodes(S, V, dim, 0, solver, norm, Sbc)
    r = m - 2 if solver == "spectra" else m
    assert out["modes"].shape[1] == r == ref["modes"].shape[1]
    assert rel_max(out["G"], ref["G"]) <= 1e-13
    assert np.all(np.abs(out["evals_norm"] - ref["evals_norm"]) <= 1e-10 * ref["evals_norm"])
    assert np.all(np.abs(out["normfact"] - ref["normfact"]) <= 1e-9 * ref["normfact"])
    assert modes_rel(out["modes"], ref["modes"]) <= 1e-8
    s = np.sign(np.einsum("ij,ij->j", out["modes"], ref["modes"]))
    assert rel_max(out["modes_bc"] * s, ref["modes_bc"]) <= 1e-8


@pytest.mark.parametrize("r", [105, 111, 208])
def test_lift_column_chunks_against_lift_oracle(pkg, pod, r):
    """r = 105 (104 + 1), 111 (104 + 7), 208 (two full chunks): lift + norms with the GPU's own eigenvectors."""
    Nc, dim, m = 1203, 1, 208
    rng = np.random.default_rng(r)
    S = np.asfortranarray(rng.standard_normal((Nc, m)))
    V = rng.random(Nc) + 0.5
    for norm in ("L2", "Frobenius"):
        out = pkg.getModes(pod, S, V, dim, r, "eigen", norm)
        Phi, nf, _ = po.lift(S, V, dim, out["evecs"], norm)
        assert np.all(np.abs(out["normfact"] - nf) <= 1e-12 * nf)
        assert (np.abs(out["modes"] - Phi).max(axis=0) / np.abs(Phi).max(axis=0)).max() <= 1e-11
    f = 1.0 + np.arange(r)
    pod.set_volumes(V); pod.upload(S, dim); pod.gram(); _, Q = pod.eig(r)
    Phi = pod.lift_scaled(f)                                          # getWeightedModes' lift with more than 104 factors
    assert rel_max(Phi, (S @ Q) * f) <= 1e-12


def test_blocked_path_lifts_boundary_rows(pkg, pod):
    """pod_block_lift_finish left have_lift unset, so the snapshot-blocked path had no boundary values
    (the reference assigns SnapMatrixBC * eigvec / normFactor on every patch, ITHACAPOD.C:573-578)."""
    Nc, dim, m, r = 2003, 3, 70, 12
    S = synth.snapshots(Nc, dim, m, 2 * r, 31)
    V = synth.volumes(Nc, 31)
    rng = np.random.default_rng(2)
    Sbc = np.asfortranarray(rng.standard_normal((29, m)))
    src = lambda first, nb: [np.ascontiguousarray(S[:, j]) for j in range(first, first + nb)]
    out = pkg.getModesMemoryEfficient(pod, src, m, V, dim, r, "eigen", "L2", block=32, Sbc=Sbc)
    ref = po.getmodes(S, V, dim, r, "eigen", "L2", Sbc)
    assert modes_rel(out["modes"], ref["modes"]) <= 1e-8
    s = np.sign(np.einsum("ij,ij->j", out["modes"], ref["modes"]))
    assert rel_max(out["modes_bc"] * s, ref["modes_bc"]) <= 1e-8


@pytest.mark.parametrize("Nc,dim,m,r,chunk,norm", [(5003, 3, 48, 10, 1700, "L2"), (4001, 1, 130, 20, 1000, "Frobenius"),
                                                     (999, 3, 33, 5, 999, "L2"), (1200, 1, 210, 120, 500, "L2")])
def test_row_streamed_pod_matches_oracle(pkg, pod, Nc, dim, m, r, chunk, norm):
    """pod_stream_*: ragged last chunk, one chunk only, more chunks than slots, r > 104; whole getModes against the oracle and
    against the resident path (the chunk sums differ from the one-launch Gram only by the summation order)."""
    if r > 100:
        S, V = well_separated(Nc, dim, m, 9)
    else:
        S = synth.snapshots(Nc, dim, m, 2 * r, 41)
        V = synth.volumes(Nc, 41)
    rng = np.random.default_rng(3)
    Sbc = np.asfortranarray(rng.standard_normal((11, m)))
    out = pkg.getModesStreamed(pod, S, V, dim, r, "eigen", norm, chunk_cells=chunk, Sbc=Sbc)
    ref = po.getmodes(S, V, dim, r, "eigen", norm, Sbc)
    assert rel_max(out["G"], ref["G"]) <= 1e-13
    assert np.all(np.abs(out["evals_norm"][:r] - ref["evals_norm"][:r]) <= 1e-10 * ref["evals_norm"][:r])
    assert np.all(np.abs(out["normfact"] - ref["normfact"]) <= 1e-9 * ref["normfact"])
    assert modes_rel(out["modes"], ref["modes"]) <= 1e-8
    s = np.sign(np.einsum("ij,ij->j", out["modes"], ref["modes"]))
    assert rel_max(out["modes_bc"] * s, ref["modes_bc"]) <= 1e-8
    res = pkg.getModes(pod, S, V, dim, r, "eigen", norm)
    assert rel_max(out["G"], res["G"]) <= 1e-13


def test_row_streamed_call_order_errors(pkg, pod):
    L, h = pod.L, pod.h
    pod.stream_begin(8, 1, 100)
    assert L.ithaca_b200_pod_stream_gram_add(h, 0) == 1           # empty slot
    assert L.ithaca_b200_pod_stream_gram_finish(h, None) == 1     # nothing added
    assert L.ithaca_b200_pod_stream_lift_add(h, 0, 2, 0) == 1     # no eigensolve yet
    assert L.ithaca_b200_pod_gram(h, None) == 1                   # resident-path call in streamed mode
    assert L.ithaca_b200_pod_stream_slot(h, 2, 0, 10, None, None, None) == 1
    assert L.ithaca_b200_pod_stream_slot(h, 0, 95, 10, None, None, None) == 1   # beyond the shard


# ---- benchmark shapes against the compiled reference (fixtures generated by tests/golden/make_golden.py) ----------------
@pytest.mark.parametrize("case", SHAPE_CASES, ids=[c[0] for c in SHAPE_CASES])
def test_benchmark_shapes_against_reference_fixture(pkg, pod, case):
    """no oracle comparison at the C3 shape (m = 1000, dim 3: the 25-tile-row Gram schedule with 4x3
    super-tiles, eig_jacobi_reg_kernel<32>, lift_kernel<7>) nor at m = 2000 (50 tile rows, <64>, <13>). Few cells keep it
    cheap; N_d > m so the Gram has full rank. The fixture holds the reference's own Eigen/Spectra results."""
    name, Nc, dim, m, r, seed, solver, gs = case
    g = np.load(os.path.join(GOLD, "pod_%s.npz" % name))
    S, V, _ = inputs(Nc, dim, m, r, seed, 0)
    out = pkg.getModes(pod, S, V, dim, r, solver, "L2")
    assert rel_max(out["G"][::gs, ::gs], g["G_sub"]) <= 1e-13
    assert abs(np.trace(out["G"]) - g["G_trace"]) <= 1e-13 * g["G_trace"]
    assert rel_max(out["G"], po.gram(S, V, dim)) <= 1e-13          # every entry, against the numpy restatement
    assert np.all(np.abs(out["evals_norm"][:r] - g["evals_norm"][:r]) <= 1e-10 * g["evals_norm"][:r])
    assert np.all(np.abs(out["normfact"] - g["normfact"]) <= 1e-9 * g["normfact"])
    st = int(g["stride"])
    assert modes_rel(out["modes"][::st], g["modes_sub"]) <= 1e-8


# ---- the device-side synthetic generator of the benchmark (measurement helper) ---------------------------------------
def _philox4x32_10(c0, c1, c2, c3, k0, k1):
    M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
    c0, c1, c2, c3 = [np.asarray(x, dtype=np.uint64) for x in (c0, c1, c2, c3)]
    k0, k1 = np.uint64(k0), np.uint64(k1)
    mask = np.uint64(0xFFFFFFFF)
    for _ in range(10):
        p0, p1 = M0 * c0, M1 * c2
        hi0, lo0, hi1, lo1 = p0 >> np.uint64(32), p0 & mask, p1 >> np.uint64(32), p1 & mask
        c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
        k0, k1 = (k0 + np.uint64(0x9E3779B9)) & mask, (k1 + np.uint64(0xBB67AE85)) & mask
    return c0, c1, c2, c3


def _unit_variate(rows, col, seed):
    rows = np.asarray(rows, dtype=np.uint64)
    w = _philox4x32_10(rows & np.uint64(0xFFFFFFFF), rows >> np.uint64(32), np.full(rows.shape, col, dtype=np.uint64),
                       np.zeros(rows.shape, dtype=np.uint64), seed & 0xFFFFFFFF, seed >> 32)
    s = sum(x.astype(np.float64) for x in w) / 4294967296.0
    return 1.7320508075688772 * (s - 2.0)


def test_synthetic_generator_is_chunking_independent_and_as_documented(pkg, pod):
    torch = pytest.importorskip("torch")
    dev = torch.device("cuda", 0)
    Nd_global, row0, nrows, m, R, eps, seed = 100_003, 20_011, 3_001, 37, 12, 1e-3, (7 << 32) + 20251104
    rng = np.random.default_rng(0)
    coef, _ = np.linalg.qr(rng.standard_normal((m, R)))
    dcoef = torch.tensor(np.asfortranarray(coef).T.copy(), device=dev)     # (R, m) row-major = m x R col-major
    ld = (nrows + 15) // 16 * 16
    A = torch.zeros((m, ld), device=dev, dtype=torch.float64)
    pod.synth_snapshots(A.data_ptr(), ld, row0, nrows, Nd_global, m, dcoef.data_ptr(), R, eps, seed)
    # the same rows produced as two chunks with another leading dimension
    cut = 1_234
    ld1, ld2 = 1_248, 1_776
    B1 = torch.zeros((m, ld1), device=dev, dtype=torch.float64)
    B2 = torch.zeros((m, ld2), device=dev, dtype=torch.float64)
    pod.synth_snapshots(B1.data_ptr(), ld1, row0, cut, Nd_global, m, dcoef.data_ptr(), R, eps, seed)
    pod.synth_snapshots(B2.data_ptr(), ld2, row0 + cut, nrows - cut, Nd_global, m, dcoef.data_ptr(), R, eps, seed)
    torch.cuda.synchronize()
    pod.timings()                                                           # synchronises the library's stream
    a = A[:, :nrows].cpu().numpy()
    b = np.concatenate([B1[:, :cut].cpu().numpy(), B2[:, :nrows - cut].cpu().numpy()], axis=1)
    assert np.array_equal(a, b)
    # the documented formula
    gi = row0 + np.arange(nrows)
    x = (gi + 0.5) / Nd_global
    U = np.stack([np.sin((k + 1) * np.pi * x) / (k + 1) for k in range(R)], axis=1)
    want = (U @ coef.T).T
    for j in (0, 5, m - 1):
        w = _unit_variate(gi, j, seed)
        assert np.abs(a[j] - (want[j] + eps * w)).max() <= 1e-12
    noise = (a - want) / eps
    assert abs(noise.mean()) < 0.02 and abs(noise.var() - 1.0) < 0.02
    # volumes
    Vd = torch.zeros(5_000, device=dev, dtype=torch.float64)
    pod.synth_volumes(Vd.data_ptr(), 1_000, 5_000, 50_000, seed)
    pod.timings()
    v = Vd.cpu().numpy()
    assert np.abs(v - np.exp(0.5 * _unit_variate(1_000 + np.arange(5_000), 0xFFFFFFFF, seed)) * 0.8824969025845955 / 50_000).max() <= 1e-18
    assert abs(v.sum() * 10 - 1.0) < 0.05


def test_row_streamed_in_place_producer(pkg, pod):
    """pod_stream_slot: the generator writes each chunk straight into the slot (no staging copy); same modes as the resident
    path on the same generated data."""
    torch = pytest.importorskip("torch")
    dev = torch.device("cuda", 0)
    Nc, dim, m, r, R, seed = 9_001, 3, 64, 8, 16, 99
    Nd = Nc * dim
    rng = np.random.default_rng(4)
    coef, _ = np.linalg.qr(rng.standard_normal((m, R)))
    dcoef = torch.tensor(np.asfortranarray(coef).T.copy(), device=dev)
    chunks = [(0, 4000), (4000, 4000), (8000, 1001)]
    pod.stream_begin(m, dim, Nc)

    def fill(slot, c0, nc):
        Sp, ld, Vp = pod.stream_slot(slot, c0, nc)
        pod.synth_snapshots(Sp, ld, c0 * dim, nc * dim, Nd, m, dcoef.data_ptr(), R, 1e-6, seed)
        pod.synth_volumes(Vp, c0, nc, Nc, seed)

    for k, (c0, nc) in enumerate(chunks):
        fill(k & 1, c0, nc)
        pod.stream_gram_add(k & 1)
    G = pod.stream_gram_finish()
    ev, Q = pod.eig(r)
    for k, (c0, nc) in enumerate(chunks):
        fill(k & 1, c0, nc)
        pod.stream_lift_add(k & 1, r, "L2")
    Phi, nf = pod.stream_lift_finish(r)
    # resident path on the same data
    ld = (Nd + 15) // 16 * 16
    A = torch.zeros((m, ld), device=dev, dtype=torch.float64)
    Vd = torch.zeros(Nc, device=dev, dtype=torch.float64)
    pod.synth_snapshots(A.data_ptr(), ld, 0, Nd, Nd, m, dcoef.data_ptr(), R, 1e-6, seed)
    pod.synth_volumes(Vd.data_ptr(), 0, Nc, Nc, seed)
    pod.timings()
    S = np.asfortranarray(A[:, :Nd].cpu().numpy().T)
    V = Vd.cpu().numpy()
    ref = po.getmodes(S, V, dim, r, "eigen", "L2")
    assert rel_max(G, ref["G"]) <= 1e-13
    assert np.all(np.abs(nf - ref["normfact"]) <= 1e-9 * ref["normfact"])
    assert modes_rel(Phi, ref["modes"]) <= 1e-8


def test_getmodes_fields2_overload(pkg, pod):
    """getModes(snapshots, modes, fields2, ...) (ITHACAPOD.C:1940-2086): eigenvectors of the vector fields' correlation matrix
    applied to a second list of another field type (flux-like: its own length, no volumes), boundary rows included."""
    Nc, m, r = 3001, 40, 9
    S2 = synth.snapshots(Nc, 3, m, 2 * r, 51)            # fields2: volVectorField snapshots
    V = synth.volumes(Nc, 51)
    rng = np.random.default_rng(6)
    nF = 8123                                            # snapshots: e.g. surfaceScalarField (internal faces), another size
    S = np.asfortranarray(rng.standard_normal((nF, m)))
    Sbc = np.asfortranarray(rng.standard_normal((77, m)))
    for solver in ("spectra", "eigen"):
        out = pkg.getModesFields2(pod, S, S2, V, 3, r, solver, Sbc)
        ref = po.getmodes_fields2(S, S2, V, 3, r, solver, Sbc)
        assert rel_max(out["G"], ref["G"]) <= 1e-13
        assert len(out["evals_norm"]) == r
        assert np.all(np.abs(out["evals_norm"] - ref["evals_norm"]) <= 1e-10 * ref["evals_norm"])
        assert modes_rel(out["modes"], ref["modes"]) <= 1e-8
        s = np.sign(np.einsum("ij,ij->j", out["modes"], ref["modes"]))
        assert rel_max(out["modes_bc"] * s, ref["modes_bc"]) <= 1e-8
    assert pod.L.ithaca_b200_pod_gram(pod.h, None) == 1   # the resident matrix is the second list now: no Gram of it
    with pytest.raises(pkg.IthacaB200Error):
        pod.upload_second(np.asfortranarray(S[:, :m - 1]))   # another number of snapshots


def test_row_streamed_identities_at_scale(pkg, pod):
    """Size-independent properties of the row-streamed path at a size no oracle run covers (2.4 GB: 400k cells x 256 vector
    snapshots, five ragged chunks produced in place): sum(lambda) = trace(G), normfact_i^2 = lambda_i (|S q_i|_V^2 = q_i^T G q_i),
    V-orthonormal modes, and the same eigenvalues whatever the chunking (chunk sums differ only in the summation order)."""
    torch = pytest.importorskip("torch")
    dev = torch.device("cuda", 0)
    Nc, dim, m, r, R, seed = 400_003, 3, 256, 24, 48, 7
    Nd = Nc * dim
    rng = np.random.default_rng(12)
    coef, _ = np.linalg.qr(rng.standard_normal((m, R)))
    dcoef = torch.tensor(np.asfortranarray(coef).T.copy(), device=dev)

    def run(chunk):
        chunks = [(c0, min(chunk, Nc - c0)) for c0 in range(0, Nc, chunk)]
        pod.stream_begin(m, dim, Nc)

        def fill(slot, c0, nc):
            Sp, ld, Vp = pod.stream_slot(slot, c0, nc)
            pod.synth_snapshots(Sp, ld, c0 * dim, nc * dim, Nd, m, dcoef.data_ptr(), R, 1e-6, seed)
            pod.synth_volumes(Vp, c0, nc, Nc, seed)

        for k, (c0, nc) in enumerate(chunks):
            fill(k & 1, c0, nc)
            pod.stream_gram_add(k & 1)
        G = pod.stream_gram_finish()
        ev, Q = pod.eig(r)
        for k, (c0, nc) in enumerate(chunks):
            fill(k & 1, c0, nc)
            pod.stream_lift_add(k & 1, r, "L2")
        Phi, nf = pod.stream_lift_finish(r)
        return G, ev, nf, Phi, Q

    G, ev, nf, Phi, Q = run(90_001)
    assert abs(ev.sum() - np.trace(G)) <= 1e-12 * np.trace(G)
    qGq = np.einsum("ij,ij->j", Q[:, :r], G @ Q[:, :r])   # on failure: tells an eigenvector error from a lift / norm error
    assert np.all(np.abs(nf * nf / ev[:r] - 1.0) <= 1e-9), (np.abs(qGq / ev[:r] - 1).max(), np.abs(nf * nf / qGq - 1).max())
    Vd = torch.zeros(Nc, device=dev, dtype=torch.float64)
    pod.synth_volumes(Vd.data_ptr(), 0, Nc, Nc, seed)
    pod.timings()
    W = np.repeat(Vd.cpu().numpy(), dim)
    assert np.abs(Phi.T @ (W[:, None] * Phi) - np.eye(r)).max() <= 1e-9
    G1, ev1, nf1, _, _ = run(Nc)                             # one chunk = the resident arithmetic
    assert rel_max(G, G1) <= 1e-13
    assert np.all(np.abs(ev[:r] - ev1[:r]) <= 1e-10 * ev1[:r])
    # bit-reproducible run to run (fixed summation orders everywhere): the same five-chunk pass three more times. This is the
    # check that catches a race in the TMA rings -- the stage-release race fixed in round 2 (common.cuh)
    # put one wrong 8 x 24 fragment into one pass in five at exactly this size
    for _ in range(3):
        G2, ev2, nf2, Phi2, _ = run(90_001)
        assert np.array_equal(G2, G) and np.array_equal(ev2, ev)
        assert np.array_equal(nf2, nf)
        assert np.array_equal(Phi2, Phi)
