"""GPU parity tests of the Galerkin projection (K5-K9) through the C ABI against oracle/fvc_oracle.c.
Tolerance (BASELINE.json): projected operators 1e-10 relative (max |delta| / max |op|); face addressing bit-exact."""
import os

import numpy as np
import pytest

from oracle import fvc, polymesh as pm

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
TOL = 1e-10


def rel(A, B):
    return np.abs(A - B).max() / np.abs(B).max()


def meshes():
    return {
        "tutorial04": lambda: pm.load_compact(os.path.join(GOLD, "mesh_04unsteadyNS.npz")),
        "tutorial01": lambda: pm.load_compact(os.path.join(GOLD, "mesh_01POD.npz")),        # 10 000 hex cells, lid-driven cavity
        "tutorial03": lambda: pm.load_compact(os.path.join(GOLD, "mesh_03steadyNS.npz")),   # 9 000 cells, backward-facing step
        "box": lambda: pm.synthetic_hex(9, 7, 5, kind="box", seed=3, perturb=0.15),
        "cyl": lambda: pm.synthetic_hex(12, 40, 6, kind="cylinder", seed=4, perturb=0.2),
    }


@pytest.fixture(scope="module", params=["tutorial04", "tutorial01", "tutorial03", "box", "cyl"])
def case(request, pkg, pod):
    mesh = meshes()[request.param]()
    N, Np = (7, 3) if request.param != "cyl" else (11, 4)
    U, Ub, P, Pb = fvc.synthetic_basis(mesh, N, Np, seed=11)
    pr = pkg.Projector(pod)
    pr.set_mesh(mesh, corrected=True)
    pr.set_basis("U", U, Ub)
    pr.set_basis("P", P, Pb)
    return mesh, fvc.Oracle(mesh, corrected=True), pr, (U, Ub, P, Pb)


def test_addressing_bit_exact(case):
    mesh, orc, pr, _ = case
    off, ent, other, digest = pr.addressing()
    o2, e2, t2 = orc.cell_faces()
    assert np.array_equal(off, o2) and np.array_equal(ent, e2) and np.array_equal(other, t2)
    assert digest != 0
    # the digest is a pure function of the arrays
    assert pr.addressing()[3] == digest


def test_operator_fields_match_oracle(case):
    mesh, orc, pr, (U, Ub, P, Pb) = case
    phi = orc.flux(U[2], Ub[2])
    assert rel(pr.fvc("flux", 2), phi) <= 1e-13
    assert rel(pr.fvc("div_phi", 2, 4), orc.div_phi(phi, U[4], Ub[4])) <= 1e-12
    assert rel(pr.fvc("laplacian", 3), orc.laplacian(U[3], Ub[3])) <= 1e-12
    assert rel(pr.fvc("grad_p", 1), orc.grad(P[1], Pb[1])) <= 1e-12
    assert rel(pr.fvc("div_u", 5), orc.div(U[5], Ub[5])) <= 1e-12


def test_projected_operators_match_oracle(case):
    mesh, orc, pr, (U, Ub, P, Pb) = case
    assert rel(pr.project("M"), orc.mass_term(U)) <= TOL
    assert rel(pr.project("B"), orc.diffusive_term(U, Ub)) <= TOL
    assert rel(pr.project("K"), orc.pressure_gradient_term(U, P, Pb)) <= TOL
    assert rel(pr.project("P"), orc.divergence_term(U, Ub, P)) <= TOL
    Cr = orc.convective_term_tens(U, Ub, cache=True)
    Cm = pr.project("Cmat")  # materialised div(phi_j, U_k) fields (the reference's _cache structure)
    assert Cm.shape == Cr.shape and rel(Cm, Cr) <= TOL
    t = pr.timings()
    assert t["contraction"] > 0 and t["div_phi"] > 0
    C = pr.project("C")      # fused face-based contraction
    assert C.shape == Cr.shape and rel(C, Cr) <= TOL
    t = pr.timings()
    assert t["contraction"] > 0 and t["face_records"] > 0


def test_ppe_operators_match_oracle(case, monkeypatch):
    """steadyNS::projectPPE's additional operators: D, G (divMomentum), BC3, BC4."""
    mesh, orc, pr, (U, Ub, P, Pb) = case
    monkeypatch.delenv("ITHACA_B200_G", raising=False)
    D = pr.project("D")
    assert rel(D, orc.laplacian_pressure(P, Pb)) <= TOL and np.array_equal(D, D.T)
    assert rel(pr.project("BC4"), orc.pressure_bc4(Ub, Pb)) <= TOL
    assert rel(pr.project("BC3"), orc.pressure_bc3(U, Ub, P, Pb)) <= TOL
    assert rel(pr.project("BC1"), orc.pressure_bc1(U, Ub, Pb)) <= TOL
    BC2 = pr.project("BC2")
    BC2r = orc.pressure_bc2(U, Ub, Pb)
    assert BC2.shape == BC2r.shape and rel(BC2, BC2r) <= TOL
    # fvc-only operators of discretizeThenProject (flux method): adjoint of the outer div on the device
    assert rel(pr.project("BP"), orc.diffusive_term_flux_method(U, Ub, P)) <= TOL
    Cf = pr.project("Cf")
    Cfr = orc.convective_term_flux_tens(U, Ub, P)
    assert Cf.shape == Cfr.shape and rel(Cf, Cfr) <= TOL
    vls, vlsb = U[1] + 0.5 * U[2], Ub[1] + 0.5 * Ub[2]
    assert rel(pr.project_background(vls, vlsb), orc.convective_background(U, Ub, vls, vlsb)) <= TOL
    assert rel(pr.project_background(vls, vlsb, divergent=True), orc.convective_background(U, Ub, vls, vlsb, P, Pb)) <= TOL
    # penalty terms of one (patch, component) pair: plain patch sums (steadyNS.C:1515-1582)
    name, typ, sl = [p for p in mesh.patch_slices() if p[1] != "empty" and p[2].stop > p[2].start][0]
    X = np.stack([u.reshape(-1, 3)[sl, 1] for u in Ub], axis=1)
    vec, mat = pr.bc_velocity(sl.start, sl.stop - sl.start, 1)
    assert np.abs(vec - X.sum(axis=0)).max() <= 1e-12 * np.abs(X).sum(axis=0).max()
    assert rel(mat, X.T @ X) <= 1e-12
    G = pr.project("G")
    Gr = orc.div_momentum(U, Ub, P, Pb)
    assert G.shape == Gr.shape and rel(G, Gr) <= TOL          # fused face-based contraction
    assert pr.timings()["face_records"] > 0
    monkeypatch.setenv("ITHACA_B200_G", "mat")
    Gm = pr.project("G")                                        # materialised div(phi_j, U_k) fields
    assert rel(Gm, Gr) <= TOL and pr.timings()["div_phi"] > 0
    monkeypatch.delenv("ITHACA_B200_G", raising=False)


def test_uncorrected_laplacian(pkg, pod):
    mesh = meshes()["cyl"]()
    U, Ub, _, _ = fvc.synthetic_basis(mesh, 5, 0, seed=2)
    pr = pkg.Projector(pod)
    pr.set_mesh(mesh, corrected=False)
    pr.set_basis("U", U, Ub)
    orc = fvc.Oracle(mesh, corrected=False)
    assert rel(pr.project("B"), orc.diffusive_term(U, Ub)) <= TOL
    # and it differs from the corrected operator on this non-orthogonal mesh (the switch is live)
    assert rel(pr.project("B"), fvc.Oracle(mesh, corrected=True).diffusive_term(U, Ub)) > 1e-6


def test_reference_loop_nest_equals_hoisted(case):
    """C from the GPU equals the reference's own N^3 loop nest (convective_term_tens, steadyNS.C:997-1036)."""
    mesh, orc, pr, (U, Ub, P, Pb) = case
    n = 3
    pr2 = pr
    pr2.set_basis("U", U[:n], Ub[:n])
    C = pr2.project("C")
    assert rel(C, orc.convective_term_tens(U[:n], Ub[:n], cache=False)) <= TOL
    pr2.set_basis("U", U, Ub)


def test_properties_at_scale(pkg, pod):
    """200k-cell unstructured cylinder mesh, N = 50: properties that need no oracle pass at this size.
    M equals the V-weighted Gram (symmetric); P row sums obey Gauss' theorem; C is linear in j."""
    mesh = pm.synthetic_hex(50, 200, 20, kind="cylinder", seed=9, perturb=0.2)
    N, Np = 50, 10
    U, Ub, P, Pb = fvc.synthetic_basis(mesh, N, Np, seed=5)
    U[7] = U[3] + 2.0 * U[5]
    Ub[7] = Ub[3] + 2.0 * Ub[5]
    pr = pkg.Projector(pod)
    pr.set_mesh(mesh)
    pr.set_basis("U", U, Ub)
    pr.set_basis("P", [np.ones(mesh.nCells)] + P[1:], [np.ones(mesh.nBnd)] + Pb[1:])
    M = pr.project("M")
    assert np.array_equal(M, M.T)
    Vd = np.repeat(mesh.V, 3)
    assert abs(M[3, 5] - np.sum(Vd * U[3] * U[5])) <= 1e-12 * abs(M).max()
    Pm = pr.project("P")  # P(0, j) with P_0 = 1: integral of div(U_j) = boundary flux of U_j
    bflux = np.array([(Ub[j].reshape(-1, 3) * mesh.Sf[mesh.nInt:]).sum() for j in range(N)])
    assert np.abs(Pm[0] - bflux).max() <= 1e-10 * np.abs(bflux).max()
    C = pr.project("C")
    assert rel(pr.project("Cmat"), C) <= TOL  # the two formulations agree at scale
    others = [i for i in range(N) if i != 7]
    lin = C[others, 3, :] + 2.0 * C[others, 5, :]
    assert np.abs(C[others, 7, :] - lin).max() <= 1e-10 * np.abs(C).max()
    off, ent, other, digest = pr.addressing()
    o2, e2, t2 = fvc.Oracle(mesh).cell_faces()
    assert np.array_equal(off, o2) and np.array_equal(ent, e2) and np.array_equal(other, t2)


def test_projection_error_paths(pkg):
    p = pkg.Pod(0)
    pr = pkg.Projector(p)
    with pytest.raises(pkg.IthacaB200Error):
        pr.N = 2
        pr.project("M")  # no mesh
    mesh = meshes()["box"]()
    pr.set_mesh(mesh)
    U, Ub, P, Pb = fvc.synthetic_basis(mesh, 2, 1, seed=1)
    pr.set_basis("U", U, Ub)
    pr.project("M")
    with pytest.raises(pkg.IthacaB200Error):
        pr.project("K")  # no pressure basis
    p.close()


@pytest.mark.parametrize("nranks,name,method", [(2, "tutorial04", "label"), (3, "cyl", "locality"), (4, "tutorial04", "locality")])
def test_sharded_projection_sums_to_the_global_operators(pkg, pod, nranks, name, method):
    """every rank projects on its shard mesh (owned cells + 2 halo layers, halo cells weigh 0); the sum
    over ranks -- what the per-operator all-reduce forms -- equals the single-mesh operators to 1e-10."""
    mesh = meshes()[name]()
    N, Np = 6, 3
    U, Ub, P, Pb = fvc.synthetic_basis(mesh, N, Np, seed=13)
    orc = fvc.Oracle(mesh, corrected=True)
    ref = {"M": orc.mass_term(U), "B": orc.diffusive_term(U, Ub), "K": orc.pressure_gradient_term(U, P, Pb),
           "P": orc.divergence_term(U, Ub, P), "C": orc.convective_term_tens(U, Ub, cache=True),
           "D": orc.laplacian_pressure(P, Pb), "G": orc.div_momentum(U, Ub, P, Pb), "BC3": orc.pressure_bc3(U, Ub, P, Pb),
           "BC4": orc.pressure_bc4(Ub, Pb), "BC1": orc.pressure_bc1(U, Ub, Pb), "BC2": orc.pressure_bc2(U, Ub, Pb),
           "BP": orc.diffusive_term_flux_method(U, Ub, P), "Cf": orc.convective_term_flux_tens(U, Ub, P)}
    ref["Cmat"] = ref["C"]
    acc = {k: np.zeros_like(v) for k, v in ref.items()}
    pr = pkg.Projector(pod)
    for rank in range(nranks):
        part = pkg.Partition(mesh, rank, nranks, 2, method=method)
        pr.set_mesh_shard(mesh, part, corrected=True)
        pr.set_basis("U", [part.gather(u, "cell", 3) for u in U], [part.gather(u, "bnd", 3) for u in Ub])
        pr.set_basis("P", [part.gather(p, "cell", 1) for p in P], [part.gather(p, "bnd", 1) for p in Pb])
        for k in acc:
            acc[k] += pr.project(k)
        part.close()
    for k in acc:
        assert rel(acc[k], ref[k]) <= TOL, k
    # a plain mesh_upload afterwards restores the unsharded behaviour
    pr.set_mesh(mesh, corrected=True)
    pr.set_basis("U", U, Ub)
    assert rel(pr.project("C"), ref["C"]) <= TOL


def test_cell_major_and_field_major_laplacians_agree_bitwise(pkg, pod, monkeypatch):
    """The cell-major pass (all fields per cell) and the field-major kernels do the same arithmetic in the same order."""
    mesh = meshes()["cyl"]()
    U, Ub, _, _ = fvc.synthetic_basis(mesh, 9, 0, seed=21)
    pr = pkg.Projector(pod)
    for corrected in (True, False):
        pr.set_mesh(mesh, corrected=corrected)
        pr.set_basis("U", U, Ub)
        monkeypatch.delenv("ITHACA_B200_LAP", raising=False)
        a, Ba = pr.fvc("laplacian", 4), pr.project("B")
        monkeypatch.setenv("ITHACA_B200_LAP", "cm")
        b, Bb = pr.fvc("laplacian", 4), pr.project("B")
        assert np.array_equal(a, b) and np.array_equal(Ba, Bb)
    monkeypatch.delenv("ITHACA_B200_LAP", raising=False)


# ---- the scheme set of tutorials/CFD/04unsteadyNS/system/fvSchemes ----------------------------
@pytest.fixture(scope="module", params=[("tutorial04", 1.0, 1.0), ("tutorial04", 0.5, 1.0), ("cyl", 1.0, 0.5), ("box", 0.7, -1.0)])
def scheme_case(request, pkg, pod):
    name, k, psi = request.param
    mesh = meshes()[name]()
    U, Ub, P, Pb = fvc.synthetic_basis(mesh, 6, 3, seed=21)
    rng = np.random.default_rng(5)   # rough modes: the gradient limiter is active in most cells for k = 1
    U = [u + 0.2 * rng.standard_normal(u.shape) for u in U]
    P = [p + 0.2 * rng.standard_normal(p.shape) for p in P]
    pr = pkg.Projector(pod)
    pr.set_mesh(mesh, corrected=True)
    pr.set_schemes(grad_limit_k=k, sngrad_psi=psi, C=mesh.Cc, Cf=mesh.Cf)
    pr.set_basis("U", U, Ub)
    pr.set_basis("P", P, Pb)
    return mesh, fvc.Oracle(mesh, corrected=True, grad_limit_k=k, sngrad_psi=psi), fvc.Oracle(mesh, corrected=True), pr, (U, Ub, P, Pb), k


def test_limited_schemes_fields_match_oracle(scheme_case):
    mesh, orc, plain, pr, (U, Ub, P, Pb), k = scheme_case
    g = orc.grad(P[1], Pb[1])
    assert rel(pr.fvc("grad_p", 1), g) <= 1e-12
    if k == 1.0:
        assert rel(g, plain.grad(P[1], Pb[1])) > 1e-3      # the limiter really acts on these modes
    assert rel(pr.fvc("laplacian", 3), orc.laplacian(U[3], Ub[3])) <= 1e-11


def test_limited_schemes_operators_match_oracle(scheme_case):
    mesh, orc, plain, pr, (U, Ub, P, Pb), k = scheme_case
    assert rel(pr.project("K"), orc.pressure_gradient_term(U, P, Pb)) <= TOL
    assert rel(pr.project("B"), orc.diffusive_term(U, Ub)) <= TOL
    D = pr.project("D")
    assert rel(D, orc.laplacian_pressure(P, Pb)) <= TOL and np.array_equal(D, D.T)
    assert rel(pr.project("G"), orc.div_momentum(U, Ub, P, Pb)) <= TOL
    assert rel(pr.project("BC3"), orc.pressure_bc3(U, Ub, P, Pb)) <= TOL
    assert rel(pr.project("BC1"), orc.pressure_bc1(U, Ub, Pb)) <= TOL
    assert rel(pr.project("BP"), orc.diffusive_term_flux_method(U, Ub, P)) <= TOL
    vls, vlsb = U[1] + 0.5 * U[2], Ub[1] + 0.5 * Ub[2]
    assert rel(pr.project_background(vls, vlsb, divergent=True), orc.convective_background(U, Ub, vls, vlsb, P, Pb)) <= TOL
    # operators without a gradient or snGrad in them are untouched by the scheme selection
    assert rel(pr.project("P"), plain.divergence_term(U, Ub, P)) <= TOL
    if k == 1.0:
        assert rel(pr.project("K"), plain.pressure_gradient_term(U, P, Pb)) > 1e-6


def test_set_schemes_argument_checks(pkg, pod):
    mesh = meshes()["box"]()
    pr = pkg.Projector(pod)
    pr.set_mesh(mesh, corrected=False)
    with pytest.raises(pkg.IthacaB200Error, match="should be >= 0 and <= 1"):
        pr.set_schemes(grad_limit_k=1.5, C=mesh.Cc, Cf=mesh.Cf)
    with pytest.raises(pkg.IthacaB200Error, match="nonOrthCorrectionVectors"):
        pr.set_schemes(sngrad_psi=1.0)
    with pytest.raises(pkg.IthacaB200Error, match="needs mesh.C"):
        pr.set_schemes(grad_limit_k=0.5)
    # k below SMALL is the unlimited gradient (cellMDLimitedGrads.C), and a new mesh resets the selection
    U, Ub, P, Pb = fvc.synthetic_basis(mesh, 2, 2, seed=1)
    pr.set_schemes(grad_limit_k=1e-16, C=mesh.Cc, Cf=mesh.Cf)
    pr.set_basis("U", U, Ub)
    pr.set_basis("P", P, Pb)
    assert rel(pr.fvc("grad_p", 0), fvc.Oracle(mesh, corrected=False).grad(P[0], Pb[0])) <= 1e-13


def test_sharded_projection_with_limited_schemes(pkg, pod):
    """Shards + mesh_set_schemes: the limiter walks each cell's faces in the shard's own face order (ascending global index
    inside the internal and the boundary group), which is the global order for every owned and first-layer cell, so the
    shard sums equal the single-mesh operators with the same schemes."""
    mesh = meshes()["tutorial04"]()
    U, Ub, P, Pb = fvc.synthetic_basis(mesh, 5, 3, seed=17)
    rng = np.random.default_rng(9)
    U = [u + 0.2 * rng.standard_normal(u.shape) for u in U]
    P = [p + 0.2 * rng.standard_normal(p.shape) for p in P]
    orc = fvc.Oracle(mesh, corrected=True, grad_limit_k=1.0, sngrad_psi=0.5)
    ref = {"B": orc.diffusive_term(U, Ub), "K": orc.pressure_gradient_term(U, P, Pb), "D": orc.laplacian_pressure(P, Pb),
           "BC3": orc.pressure_bc3(U, Ub, P, Pb), "G": orc.div_momentum(U, Ub, P, Pb)}
    acc = {k: np.zeros_like(v) for k, v in ref.items()}
    pr = pkg.Projector(pod)
    nranks = 3
    for rank in range(nranks):
        part = pkg.Partition(mesh, rank, nranks, 2, method="locality")
        pr.set_mesh_shard(mesh, part, corrected=True)
        cpa, fpa = part.addressing()
        pr.set_schemes(1.0, 0.5, C=mesh.Cc[cpa], Cf=mesh.Cf[np.abs(fpa.astype(np.int64)) - 1])
        pr.set_basis("U", [part.gather(u, "cell", 3) for u in U], [part.gather(u, "bnd", 3) for u in Ub])
        pr.set_basis("P", [part.gather(p, "cell", 1) for p in P], [part.gather(p, "bnd", 1) for p in Pb])
        for k in acc:
            acc[k] += pr.project(k)
        part.close()
    for k in acc:
        assert rel(acc[k], ref[k]) <= TOL, k


@pytest.mark.parametrize("name,scheme,k", [("tutorial04", "upwind", 0.0), ("cyl", "linearUpwind", 0.0), ("box", "linearUpwind", 1.0)])
def test_upwind_and_linear_upwind_convection_match_oracle(pkg, pod, name, scheme, k):
    """divSchemes 'Gauss upwind' / 'Gauss linearUpwind grad(U)' (SURVEY 8(f) rank 4): the div field, C and G (upwind: the fused
    contraction over doubled face records; linearUpwind: through materialised div fields), BC2 and the background terms follow the scheme;
    linearUpwind takes its gradient from the selected gradient scheme (k = 1: cellMDLimited)."""
    mesh = meshes()[name]()
    U, Ub, P, Pb = fvc.synthetic_basis(mesh, 6, 3, seed=23)
    rng = np.random.default_rng(6)
    U = [u + 0.1 * rng.standard_normal(u.shape) for u in U]
    pr = pkg.Projector(pod)
    pr.set_mesh(mesh, corrected=True)
    pr.set_schemes(grad_limit_k=k, C=mesh.Cc, Cf=mesh.Cf)
    pr.set_convection(scheme)
    pr.set_basis("U", U, Ub)
    pr.set_basis("P", P, Pb)
    orc = fvc.Oracle(mesh, corrected=True, grad_limit_k=k, convection=scheme)
    lin = fvc.Oracle(mesh, corrected=True, grad_limit_k=k)
    phi = orc.flux(U[2], Ub[2])
    d = orc.div_phi(phi, U[4], Ub[4])
    assert rel(pr.fvc("div_phi", 2, 4), d) <= 1e-12
    assert rel(d, lin.div_phi(phi, U[4], Ub[4])) > 1e-3
    Cr = orc.convective_term_tens(U, Ub, cache=True)
    assert rel(pr.project("C"), Cr) <= TOL
    assert (pr.timings()["face_records"] > 0) == (scheme == "upwind")     # upwind: fused contraction over doubled face records
    assert rel(pr.project("Cmat"), Cr) <= TOL
    assert rel(Cr, lin.convective_term_tens(U, Ub, cache=True)) > 1e-4
    assert rel(pr.project("G"), orc.div_momentum(U, Ub, P, Pb)) <= TOL
    assert rel(pr.project("BC2"), orc.pressure_bc2(U, Ub, Pb)) <= TOL
    vls, vlsb = U[1] + 0.5 * U[2], Ub[1] + 0.5 * Ub[2]
    assert rel(pr.project_background(vls, vlsb), orc.convective_background(U, Ub, vls, vlsb)) <= TOL
    if scheme == "upwind":
        assert rel(pr.project("Cf"), orc.convective_term_flux_tens(U, Ub, P)) <= TOL
    else:
        with pytest.raises(pkg.IthacaB200Error, match="linearUpwind"):
            pr.project("Cf")
    # operators without a convection term are untouched; back to linear restores the fused path
    assert rel(pr.project("B"), lin.diffusive_term(U, Ub)) <= TOL
    pr.set_convection("linear")
    assert rel(pr.project("C"), lin.convective_term_tens(U, Ub, cache=True)) <= TOL


def test_linear_upwind_needs_centres(pkg, pod):
    mesh = meshes()["box"]()
    pr = pkg.Projector(pod)
    pr.set_mesh(mesh, corrected=True)
    with pytest.raises(pkg.IthacaB200Error, match="needs mesh.C"):
        pr.set_convection("linearUpwind")
    pr.set_convection("upwind")


@pytest.mark.parametrize("name,scheme", [("tutorial04", "linear"), ("cyl", "linear"), ("box", "upwind")])
def test_consistent_flux_method_uses_the_uploaded_fluxes(pkg, pod, name, scheme):
    """fluxMethod == "consistent" (steadyNS.C:1010-1015): C(i,j,k) = int U_i . div(L_PHImodes[j], U_k) with caller-supplied flux
    fields (there was no ABI entry). Fused and materialised forms against the reference's loop nest over the
    oracle's div(phi, U); G, the divergence term (which must keep the interpolated flux) and the reset are checked too."""
    mesh = meshes()[name]()
    N, Np = 5, 3
    U, Ub, P, Pb = fvc.synthetic_basis(mesh, N, Np, seed=17)
    orc = fvc.Oracle(mesh, corrected=True, convection=scheme)
    rng = np.random.default_rng(8)
    nF = mesh.nInt + mesh.nBnd
    mask = np.concatenate([np.ones(mesh.nInt), np.asarray(mesh.bmask, dtype=np.float64)])
    # "POD modes of phi": the interpolated fluxes perturbed face by face (sign changes included), 0 on `empty` faces
    Phi = [mask * (orc.flux(U[j], Ub[j]) * (1.0 + 0.3 * rng.standard_normal(nF)) + 1e-3 * rng.standard_normal(nF)) for j in range(N)]
    pr = pkg.Projector(pod)
    pr.set_mesh(mesh, corrected=True)
    pr.set_convection(scheme)
    pr.set_basis("U", U, Ub)
    pr.set_basis("P", P, Pb)
    C_interp = pr.project("C")
    P_interp = pr.project("P")
    pr.set_flux(Phi)
    Cr = orc.convective_term_tens_consistent(U, Ub, Phi)
    assert rel(pr.project("C"), Cr) <= TOL
    assert rel(pr.project("Cmat"), Cr) <= TOL
    assert rel(pr.fvc("div_phi", 1, 3), orc.div_phi(Phi[1], U[3], Ub[3])) <= 1e-12
    assert rel(pr.fvc("flux", 2), orc.flux(U[2], Ub[2])) <= 1e-13          # fvc_eval(FLUX) stays the interpolated flux
    assert np.array_equal(pr.project("P"), P_interp)                       # fvc::div(U_j) does not take L_PHImodes
    assert rel(C_interp, Cr) > 1e-3                                        # the uploaded fluxes really differ
    pr.set_flux(None)
    assert np.array_equal(pr.project("C"), C_interp)
    with pytest.raises(pkg.IthacaB200Error):
        pr.set_flux(Phi[:-1])
    pr.set_convection("linear")
