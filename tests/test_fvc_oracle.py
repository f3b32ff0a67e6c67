"""CPU tests of the projection oracle (oracle/fvc_oracle.c + oracle/polymesh.py) against known answers.
The reference ships no golden values for the projected operators and OpenFOAM is not
installable here, so the oracle is checked through closed-form properties of the discrete operators."""
import os

import numpy as np
import pytest

from oracle import fvc, polymesh as pm

REF = "/root/reference/tutorials/CFD"
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def lin_vec(X, A, b):
    return np.ascontiguousarray((X @ A.T + b).ravel())


@pytest.fixture(scope="module")
def box():
    return pm.synthetic_hex(7, 6, 5, kind="box", seed=3, perturb=0.0)


@pytest.fixture(scope="module")
def cyl():
    return pm.synthetic_hex(8, 20, 3, kind="cylinder", seed=4, perturb=0.2)


def test_cavity_mesh_known_volume():
    if not os.path.isdir(REF):
        pytest.skip("reference tree absent")
    m = pm.load_reference_mesh(os.path.join(REF, "01POD"))
    assert (m.nCells, m.nInt, m.nBnd) == (10000, 19800, 20400)
    assert np.abs(m.V - 1e-8).max() < 1e-20  # 100x100x1 block of a 0.1x0.1x0.01 box (system/blockMeshDict)
    assert np.abs(m.w - 0.5).max() < 1e-12 and np.abs(m.corr).max() < 1e-12
    assert m.bmask.sum() == 400


def test_fixture_mesh_matches_reference_files():
    m = pm.load_compact(os.path.join(GOLD, "mesh_04unsteadyNS.npz"))
    assert (m.nCells, m.nInt, m.nBnd) == (11644, 23266, 23886)
    if os.path.isdir(REF):
        r = pm.load_reference_mesh(os.path.join(REF, "04unsteadyNS"))
        assert np.array_equal(m.owner, r.owner) and np.array_equal(m.neighbour, r.neighbour)
        assert np.array_equal(m.Sf, r.Sf) and np.array_equal(m.V, r.V)


@pytest.mark.parametrize("which", ["box", "cyl"])
def test_closed_cells_and_constant_fields(which, box, cyl):
    m = box if which == "box" else cyl
    o = fvc.Oracle(m)
    s = np.zeros((m.nCells, 3))
    np.add.at(s, m.owner, m.Sf)
    np.add.at(s, m.neighbour, -m.Sf[:m.nInt])
    assert np.abs(s).max() < 1e-13 * m.magSf.max()
    U = np.tile([1.0, -2.0, 0.5], m.nCells)
    Ub = np.tile([1.0, -2.0, 0.5], m.nBnd)
    scale = m.magSf.max() / m.V.min()
    assert np.abs(o.div(U, Ub)).max() < 1e-12 * scale
    assert np.abs(o.laplacian(U, Ub)).max() < 1e-10 * scale
    assert np.abs(o.grad(np.ones(m.nCells), np.ones(m.nBnd))).max() < 1e-12 * scale
    assert np.abs(o.grad_vec(U, Ub)).max() < 1e-12 * scale


def test_linear_fields_exact_on_orthogonal_mesh(box):
    m = box
    o = fvc.Oracle(m)
    rng = np.random.default_rng(0)
    A, b = rng.standard_normal((3, 3)), rng.standard_normal(3)
    U, Ub = lin_vec(m.Cc, A, b), lin_vec(m.Cf[m.nInt:], A, b)
    T = o.grad_vec(U, Ub).reshape(-1, 3, 3)  # T[c][i][j] = d_i u_j = A[j][i]
    assert np.abs(T - A.T).max() < 1e-11
    assert np.abs(o.div(U, Ub) - np.trace(A)).max() < 1e-11
    g = rng.standard_normal(3)
    p, pb = m.Cc @ g + 0.3, m.Cf[m.nInt:] @ g + 0.3
    assert np.abs(o.grad(p, pb).reshape(-1, 3) - g).max() < 1e-11
    assert np.abs(o.laplacian(U, Ub)).max() < 1e-9  # laplacian of a linear field vanishes


def test_gauss_theorem_and_flux_conservation(cyl):
    m = cyl
    o = fvc.Oracle(m)
    U, Ub, _, _ = fvc.synthetic_basis(m, 2, 0, seed=1)
    phi = o.flux(U[0], Ub[0])
    d = o.div(U[0], Ub[0])
    assert abs((m.V * d).sum() - phi[m.nInt:].sum()) < 1e-11 * np.abs(phi).sum()
    dv = o.div_phi(phi, U[1], Ub[1]).reshape(-1, 3)
    bsum = (phi[m.nInt:, None] * Ub[1].reshape(-1, 3)).sum(0)
    assert np.abs((m.V[:, None] * dv).sum(0) - bsum).max() < 1e-11 * np.abs(phi).sum()


def test_builders_consistent(cyl):
    m = cyl
    o = fvc.Oracle(m)
    U, Ub, P, Pb = fvc.synthetic_basis(m, 4, 2, seed=2)
    C1 = o.convective_term_tens(U, Ub, cache=False)
    C2 = o.convective_term_tens(U, Ub, cache=True)
    assert np.array_equal(C1, C2)  # convective_term_tens == convective_term_tens_cache (steadyNS.C:997 vs :1038)
    # C(i,j,k) is linear in the convecting field j
    U2 = [U[0], U[1] + 2.0 * U[2], U[2], U[3]]
    Ub2 = [Ub[0], Ub[1] + 2.0 * Ub[2], Ub[2], Ub[3]]
    C3 = o.convective_term_tens(U2, Ub2)
    ii = [0, 2, 3]  # test functions that did not change
    assert np.abs(C3[ii, 1, 0] - (C2[ii, 1, 0] + 2.0 * C2[ii, 2, 0])).max() < 1e-12 * np.abs(C2).max()
    M = o.mass_term(U)
    Vd = np.repeat(m.V, 3)
    assert np.allclose(M, np.array([[np.sum(Vd * a * b) for b in U] for a in U]), rtol=1e-13)
    B = o.diffusive_term(U, Ub)
    assert np.allclose(B[:, 1], [np.sum(Vd * a * o.laplacian(U[1], Ub[1])) for a in U], rtol=1e-13)
    K = o.pressure_gradient_term(U, P, Pb)
    assert K.shape == (4, 2)
    Pm = o.divergence_term(U, Ub, P)
    assert Pm.shape == (2, 4)
    assert np.allclose(Pm[1, 2], np.sum(m.V * P[1] * o.div(U[2], Ub[2])), rtol=1e-13)


def test_uncorrected_equals_corrected_on_orthogonal_mesh(box):
    U, Ub, _, _ = fvc.synthetic_basis(box, 1, 0, seed=5)
    a = fvc.Oracle(box, corrected=True).laplacian(U[0], Ub[0])
    b = fvc.Oracle(box, corrected=False).laplacian(U[0], Ub[0])
    assert np.abs(a - b).max() < 1e-9 * np.abs(a).max()


def test_cell_faces_csr(cyl):
    m = cyl
    off, ent, other = fvc.Oracle(m).cell_faces()
    assert off[-1] == 2 * m.nInt + m.nBnd
    for c in [0, 1, m.nCells // 2, m.nCells - 1]:
        e = ent[off[c]:off[c + 1]]
        f = np.abs(e) - 1
        assert np.all(np.diff(f) > 0)  # ascending face order
        for fe, ee, oc in zip(f, e, other[off[c]:off[c + 1]]):
            if ee > 0:
                assert m.owner[fe] == c and (oc == (m.neighbour[fe] if fe < m.nInt else -(fe - m.nInt + 1)))
            else:
                assert m.neighbour[fe] == c and oc == m.owner[fe]


def test_ppe_operators_known_answers(box, cyl):
    """D, G, BC3, BC4 of steadyNS::projectPPE (steadyNS.C:1176-1513) against closed forms:
    D is the V-weighted Gram of the pressure gradients; on the orthogonal box mesh linear fields have exact gradients, so
    D(i,j) = a_i . a_j * |Omega|; BC4 with P = 1 is the boundary flux = integral of div(U) (Gauss);
    BC3 = closed-surface integral of n . (grad p x curl U) = 0 for a rigid rotation (constant curl) and a linear p;
    G(i,j,k) = integral of grad(P_i) . div(phi_j, U_k) from its definition."""
    m = box
    o = fvc.Oracle(m)
    X, Xb = m.Cc, m.Cf[m.nInt:]
    a = np.array([[1.0, -2.0, 0.5], [0.3, 0.7, -1.1]])
    P = [np.ascontiguousarray(X @ a[i] + 0.2 * i) for i in range(2)]
    Pb = [np.ascontiguousarray(Xb @ a[i] + 0.2 * i) for i in range(2)]
    D = o.laplacian_pressure(P, Pb)
    assert np.abs(D - (a @ a.T) * m.V.sum()).max() <= 1e-12 * np.abs(D).max()
    assert np.array_equal(D, D.T)
    # rigid rotation + translation: U = w x x + t, curl U = 2 w
    w, t = np.array([0.4, -0.3, 0.9]), np.array([0.1, 0.2, -0.3])
    A = np.array([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]])
    U = [lin_vec(X, A, t), lin_vec(X, 2 * A, -t)]
    Ub = [lin_vec(Xb, A, t), lin_vec(Xb, 2 * A, -t)]
    T = o.grad_vec(U[0], Ub[0]).reshape(-1, 3, 3)
    assert np.abs(T - A.T).max() <= 1e-12  # T[i][d] = d_i U_d
    BC3 = o.pressure_bc3(U, Ub, P, Pb)
    scale = 2 * np.linalg.norm(w) * np.linalg.norm(a[0]) * m.magSf[m.nInt:].sum()
    assert np.abs(BC3).max() <= 1e-12 * scale
    ones, onesb = np.ones(m.nCells), np.ones(m.nBnd)
    BC4 = o.pressure_bc4(Ub, [onesb] + Pb)
    Pm = o.divergence_term(U, Ub, [ones])
    assert np.abs(BC4[0] - Pm[0]).max() <= 1e-12 * m.magSf.sum()
    # non-trivial BC3 / G on the non-orthogonal cylinder mesh: definition-level cross-checks
    m2 = cyl
    o2 = fvc.Oracle(m2)
    U2, Ub2, P2, Pb2 = fvc.synthetic_basis(m2, 3, 2, seed=8)
    G = o2.div_momentum(U2, Ub2, P2, Pb2)
    assert G.shape == (2, 3, 3)
    phi = o2.flux(U2[1], Ub2[1])
    d = o2.div_phi(phi, U2[2], Ub2[2])
    g = o2.grad(P2[1], Pb2[1])
    assert abs(G[1, 1, 2] - np.sum(np.repeat(m2.V, 3) * g * d)) <= 1e-13 * np.abs(G).max()
    BC3 = o2.pressure_bc3(U2, Ub2, P2, Pb2)
    BC3s = o2.pressure_bc3([U2[1] * 2.0], [Ub2[1] * 2.0], [P2[0] - 3.0 * P2[1]], [Pb2[0] - 3.0 * Pb2[1]])
    assert abs(BC3s[0, 0] - 2.0 * (BC3[0, 1] - 3.0 * BC3[1, 1])) <= 1e-12 * np.abs(BC3).max()  # bilinear


def test_background_convection_and_bc1_definitions(cyl):
    """convective_background / divergent_convective_background are slices of -C / -G with vls as the advecting field
    (steadyNS.C:1256-1293 against :1038-1083 and :1207-1254); BC1 with P = 1 is the boundary sum of lap(U)[owner] . Sf."""
    m = cyl
    o = fvc.Oracle(m)
    U, Ub, P, Pb = fvc.synthetic_basis(m, 3, 2, seed=8)
    C = o.convective_term_tens(U, Ub)
    assert np.array_equal(o.convective_background(U, Ub, U[1], Ub[1]), -C[:, 1, :])
    G = o.div_momentum(U, Ub, P, Pb)
    assert np.array_equal(o.convective_background(U, Ub, U[2], Ub[2], P, Pb), -G[:, 2, :])
    BC1 = o.pressure_bc1(U, Ub, [np.ones(m.nBnd)] + Pb)
    lap = o.laplacian(U[0], Ub[0]).reshape(-1, 3)
    act = m.bmask.astype(bool)
    own = m.owner[m.nInt:]
    want = np.sum((lap[own] * m.Sf[m.nInt:]).sum(axis=1)[act])
    assert abs(BC1[0, 0] - want) <= 1e-12 * np.abs(BC1).max()


def test_bc2_definition(cyl):
    """pressureBC2 (steadyNS.C:1396-1437): slice (i, j, k) = boundary sum of div(phi_j, U_k)[owner] . Sf P_i -- restated here
    from the oracle's own flux / div_phi fields; linear in P_i and U_k."""
    m = cyl
    o = fvc.Oracle(m)
    U, Ub, P, Pb = fvc.synthetic_basis(m, 3, 2, seed=9)
    BC2 = o.pressure_bc2(U, Ub, Pb)
    assert BC2.shape == (2, 3, 3)
    act = m.bmask.astype(bool)
    own = m.owner[m.nInt:]
    Sb = m.Sf[m.nInt:]
    for (i, j, k) in [(0, 0, 0), (1, 2, 1), (0, 1, 2)]:
        d = o.div_phi(o.flux(U[j], Ub[j]), U[k], Ub[k]).reshape(-1, 3)
        want = np.sum(((d[own] * Sb).sum(axis=1) * Pb[i])[act])
        assert abs(BC2[i, j, k] - want) <= 1e-12 * np.abs(BC2).max()
    lin = o.pressure_bc2([U[0], U[1] - 2.0 * U[2]], [Ub[0], Ub[1] - 2.0 * Ub[2]], [Pb[0] + 0.5 * Pb[1]])
    want = BC2[0, 0, 1] - 2.0 * BC2[0, 0, 2] + 0.5 * (BC2[1, 0, 1] - 2.0 * BC2[1, 0, 2])
    assert abs(lin[0, 0, 1] - want) <= 1e-12 * np.abs(BC2).max()


def test_flux_method_operators_definitions(cyl):
    """diffusive_term_flux_method / convective_term_flux_tens (steadyNS.C:1584-1610, :1704-1746): int P_i div(A) with A an
    fvc:: result whose boundary value is its owner cell's. Known answer (Gauss theorem): with P = 1 the volume integral of
    div(A) is the boundary sum of A[owner] . Sf, i.e. BP[P=1] == BC1[P=1] and Cf[P=1] == BC2[P=1]."""
    m = cyl
    o = fvc.Oracle(m)
    U, Ub, P, Pb = fvc.synthetic_basis(m, 3, 2, seed=10)
    one_c, one_b = np.ones(m.nCells), np.ones(m.nBnd)
    BP = o.diffusive_term_flux_method(U, Ub, [one_c] + P)
    Cf = o.convective_term_flux_tens(U, Ub, [one_c] + P)
    BC1 = o.pressure_bc1(U, Ub, [one_b])
    BC2 = o.pressure_bc2(U, Ub, [one_b])
    assert BP.shape == (3, 3) and Cf.shape == (3, 3, 3)
    assert np.abs(BP[0] - BC1[0]).max() <= 1e-11 * max(np.abs(BC1).max(), np.abs(BP).max())
    assert np.abs(Cf[0] - BC2[0]).max() <= 1e-11 * max(np.abs(BC2).max(), np.abs(Cf).max())
    # definition, restated from the oracle's own fields
    own = m.owner[m.nInt:]
    L = o.laplacian(U[1], Ub[1])
    d = o.div(L, np.ascontiguousarray(L.reshape(-1, 3)[own].ravel()))
    assert abs(BP[2, 1] - np.sum(m.V * P[1] * d)) <= 1e-12 * np.abs(BP).max()
    A = o.div_phi(o.flux(U[2], Ub[2]), U[0], Ub[0])
    d = o.div(A, np.ascontiguousarray(A.reshape(-1, 3)[own].ravel()))
    assert abs(Cf[1, 2, 0] - np.sum(m.V * P[0] * d)) <= 1e-12 * np.abs(Cf).max()


def test_consistent_flux_nest_reduces_to_the_interpolated_one():
    """steadyNS.C:1010-1015: with L_PHImodes[j] = linearInterpolate(U_j) & Sf the "consistent" loop nest is the other branch;
    C is linear in the flux field."""
    from oracle import fvc, polymesh as pm
    mesh = pm.synthetic_hex(6, 5, 4, kind="box", seed=3, perturb=0.1)
    U, Ub, P, Pb = fvc.synthetic_basis(mesh, 3, 2, seed=1)
    orc = fvc.Oracle(mesh)
    Phi = [orc.flux(U[j], Ub[j]) for j in range(3)]
    C = orc.convective_term_tens_consistent(U, Ub, Phi)
    Cr = orc.convective_term_tens(U, Ub)
    assert np.abs(C - Cr).max() <= 1e-13 * np.abs(Cr).max()
    C2 = orc.convective_term_tens_consistent(U, Ub, [2.0 * p for p in Phi])
    assert np.abs(C2 - 2.0 * C).max() <= 1e-13 * np.abs(C).max()
    # slabs of the outer loop over host threads give the same entries (CPU-baseline harness)
    assert np.array_equal(orc.convective_term_tens(U, Ub, threads=3), Cr)
    assert np.array_equal(orc.diffusive_term(U, Ub, threads=2), orc.diffusive_term(U, Ub))
