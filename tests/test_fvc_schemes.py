"""CPU tests of the oracle's restatement of the non-"Gauss linear" schemes that tutorials/CFD/04unsteadyNS/system/fvSchemes
ships: gradSchemes "cellMDLimited Gauss linear k" and snGrad "limited psi". OpenFOAM is not in this
container (parity UNPINNED at the OpenFOAM boundary, see oracle/fvc_oracle.c); these are definition-level checks: an
independent cell-by-cell restatement of the limiter, its defining properties, and the limiting cases of both schemes."""
import os

import numpy as np
import pytest

from oracle import fvc, polymesh as pm

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def cellwise_limited_grad(mesh, orc_plain, p, pb, k):
    """cellMDLimited as a loop over CELLS (each walking its own faces in ascending face order): what the device kernel
    does, against the oracle's face loop (OpenFOAM's structure)."""
    g = orc_plain.grad(p, pb).reshape(-1, 3).copy()
    off, ent, other = orc_plain.cell_faces()
    for c in range(mesh.nCells):
        vals = [p[c]]
        faces = []
        for e in range(off[c], off[c + 1]):
            f = abs(ent[e]) - 1
            if other[e] >= 0:
                vals.append(p[other[e]])
            else:
                b = -other[e] - 1
                if not mesh.bmask[b]:
                    continue
                vals.append(pb[b])
            faces.append(f)
        mx, mn = max(vals) - p[c], min(vals) - p[c]
        if k < 1.0:
            d = (1.0 / k - 1.0) * (mx - mn)
            mx, mn = mx + d, mn - d
        for f in faces:
            dcf = mesh.Cf[f] - mesh.Cc[c]
            ex = dcf @ g[c]
            if ex > mx:
                g[c] = g[c] + dcf * (mx - ex) / (dcf @ dcf)
            elif ex < mn:
                g[c] = g[c] + dcf * (mn - ex) / (dcf @ dcf)
    return g.reshape(-1)


@pytest.fixture(scope="module")
def box():
    mesh = pm.synthetic_hex(7, 6, 5, kind="box", seed=5, perturb=0.2)
    U, Ub, P, Pb = fvc.synthetic_basis(mesh, 4, 3, seed=2)
    return mesh, U, Ub, P, Pb


@pytest.mark.parametrize("k", [1.0, 0.5])
def test_limiter_matches_cellwise_restatement(box, k):
    mesh, U, Ub, P, Pb = box
    plain, lim = fvc.Oracle(mesh), fvc.Oracle(mesh, grad_limit_k=k)
    rng = np.random.default_rng(1)
    p = P[1] + 0.3 * rng.standard_normal(mesh.nCells)   # rough field: the limiter is active in most cells
    g = lim.grad(p, Pb[1])
    assert np.abs(g - cellwise_limited_grad(mesh, plain, p, Pb[1], k)).max() <= 1e-13 * np.abs(g).max()
    changed = np.abs(g - plain.grad(p, Pb[1])).reshape(-1, 3).max(axis=1) > 1e-9 * np.abs(g).max()
    assert changed.mean() > (0.3 if k == 1.0 else -1)   # k = 0.5 triples the admissible interval: rarely active on hexahedra
    # the vector gradient limits every component with that component's own bounds: column d of the tensor is the limited
    # gradient of the scalar field U_d
    u = U[2].reshape(-1, 3) + 0.2 * rng.standard_normal((mesh.nCells, 3))
    ub = Ub[2].reshape(-1, 3)
    T = lim.grad_vec(u.reshape(-1), ub.reshape(-1)).reshape(-1, 3, 3)
    for d in range(3):
        gd = lim.grad(np.ascontiguousarray(u[:, d]), np.ascontiguousarray(ub[:, d])).reshape(-1, 3)
        assert np.abs(T[:, :, d] - gd).max() <= 1e-13 * np.abs(gd).max()


def test_limiter_bounds_the_last_face_and_never_touches_admissible_gradients(box):
    mesh, U, Ub, P, Pb = box
    lim = fvc.Oracle(mesh, grad_limit_k=1.0)
    rng = np.random.default_rng(2)
    p = rng.standard_normal(mesh.nCells)
    pb = rng.standard_normal(mesh.nBnd)
    g = lim.grad(p, pb).reshape(-1, 3)
    off, ent, other = lim.cell_faces()
    for c in range(mesh.nCells):
        nb = [p[other[e]] if other[e] >= 0 else pb[-other[e] - 1] for e in range(off[c], off[c + 1])
              if other[e] >= 0 or mesh.bmask[-other[e] - 1]]
        mx, mn = max(nb + [p[c]]) - p[c], min(nb + [p[c]]) - p[c]
        last = [abs(ent[e]) - 1 for e in range(off[c], off[c + 1]) if other[e] >= 0 or mesh.bmask[-other[e] - 1]][-1]
        ex = (mesh.Cf[last] - mesh.Cc[c]) @ g[c]
        assert mn - 1e-12 <= ex <= mx + 1e-12
    # a linear field on an orthogonal uniform mesh: face-centre extrapolations are midpoints of neighbour values, the
    # limiter must leave the (exact) gradient alone
    um = pm.synthetic_hex(6, 5, 4, kind="box", seed=1, perturb=0.0)
    a = np.array([0.7, -1.3, 2.1])
    q, qb = um.Cc @ a + 0.4, um.Cf[um.nInt:] @ a + 0.4
    gl = fvc.Oracle(um, grad_limit_k=1.0).grad(q, qb).reshape(-1, 3)
    # (boundary faces extrapolate exactly onto their bound, so rounding may trigger clips of a few ulp)
    assert np.abs(gl - fvc.Oracle(um).grad(q, qb).reshape(-1, 3)).max() <= 1e-13
    assert np.abs(gl - a).max() <= 1e-12


def test_limiter_off_and_widening(box):
    mesh, U, Ub, P, Pb = box
    rng = np.random.default_rng(3)
    p = rng.standard_normal(mesh.nCells)
    g0 = fvc.Oracle(mesh).grad(p, Pb[0])
    assert np.array_equal(fvc.Oracle(mesh, grad_limit_k=0.0).grad(p, Pb[0]), g0)
    # smaller k widens the admissible interval: the correction shrinks monotonically (k -> 0: unlimited)
    d = [np.abs(fvc.Oracle(mesh, grad_limit_k=k).grad(p, Pb[0]) - g0).sum() for k in (1.0, 0.9, 0.75, 0.5, 1e-3)]
    assert d[0] > d[1] > d[2] >= d[3] >= d[4] == 0.0


def test_limited_sngrad_limits():
    mesh = pm.load_compact(os.path.join(GOLD, "mesh_04unsteadyNS.npz"))
    U, Ub, P, Pb = fvc.synthetic_basis(mesh, 3, 1, seed=4)
    corrected = fvc.Oracle(mesh, corrected=True).laplacian(U[1], Ub[1])
    uncorrected = fvc.Oracle(mesh, corrected=False).laplacian(U[1], Ub[1])
    # "limited 1" (the tutorial's setting) is "corrected" unless a face's uncorrected snGrad vanishes; "limited 0" is "uncorrected"
    assert np.array_equal(fvc.Oracle(mesh, sngrad_psi=1.0).laplacian(U[1], Ub[1]), corrected)
    assert np.array_equal(fvc.Oracle(mesh, sngrad_psi=0.0).laplacian(U[1], Ub[1]), uncorrected)
    half = fvc.Oracle(mesh, sngrad_psi=0.5).laplacian(U[1], Ub[1])
    assert np.abs(half - corrected).max() <= np.abs(uncorrected - corrected).max() * (1 + 1e-12)
    # a constant field has zero snGrad everywhere: the limiter is 0 and the result is exactly 0 for any psi
    one = np.tile([1.0, -2.0, 0.5], mesh.nCells)
    oneb = np.tile([1.0, -2.0, 0.5], mesh.nBnd)
    assert np.abs(fvc.Oracle(mesh, sngrad_psi=1.0, grad_limit_k=0.5).laplacian(one, oneb)).max() <= 1e-9


def test_limited_schemes_in_the_term_builders():
    """The builders pick the schemes up: with k = 1 and rough modes K, D and B differ from the plain 'Gauss linear' ones and
    stay self-consistent (D symmetric, K linear in the test function). With the tutorial's own k = 0.5 the admissible
    interval is three times the neighbourhood range and smooth modes are never clipped: the operators are unchanged."""
    mesh = pm.load_compact(os.path.join(GOLD, "mesh_04unsteadyNS.npz"))
    U, Ub, P, Pb = fvc.synthetic_basis(mesh, 4, 3, seed=6)
    rng = np.random.default_rng(8)
    P = [p + 0.2 * rng.standard_normal(p.shape) for p in P]
    U = [u + 0.2 * rng.standard_normal(u.shape) for u in U]
    plain, lim = fvc.Oracle(mesh), fvc.Oracle(mesh, grad_limit_k=1.0, sngrad_psi=1.0)
    K0, K1 = plain.pressure_gradient_term(U, P, Pb), lim.pressure_gradient_term(U, P, Pb)
    assert np.abs(K1 - K0).max() > 1e-6 * np.abs(K0).max()
    D1 = lim.laplacian_pressure(P, Pb)
    assert np.allclose(D1, D1.T, rtol=0, atol=1e-15 * np.abs(D1).max())
    B0, B1 = plain.diffusive_term(U, Ub), lim.diffusive_term(U, Ub)
    assert 0 < np.abs(B1 - B0).max() < 0.5 * np.abs(B0).max()
    K2 = lim.pressure_gradient_term([U[0] + 2.0 * U[1]], P, Pb)
    assert np.abs(K2[0] - (K1[0] + 2.0 * K1[1])).max() <= 1e-12 * np.abs(K1).max()
    Us, Ubs, Ps, Pbs = fvc.synthetic_basis(mesh, 4, 3, seed=6)
    t04 = fvc.Oracle(mesh, grad_limit_k=0.5, sngrad_psi=1.0)
    assert np.array_equal(t04.pressure_gradient_term(Us, Ps, Pbs), plain.pressure_gradient_term(Us, Ps, Pbs))


def test_upwind_and_linear_upwind_convection_definitions():
    """divSchemes 'Gauss upwind' / 'Gauss linearUpwind grad(U)' in fvc::div(phi, U): a face-by-face numpy restatement,
    the Gauss theorem for a constant field (any scheme reproduces div(phi) U0), and linearUpwind's exactness for linear
    fields on an orthogonal uniform mesh (where it must coincide with 'linear')."""
    mesh = pm.synthetic_hex(7, 6, 5, kind="box", seed=5, perturb=0.2)
    U, Ub, P, Pb = fvc.synthetic_basis(mesh, 3, 1, seed=2)
    lin, upw, lup = fvc.Oracle(mesh), fvc.Oracle(mesh, convection="upwind"), fvc.Oracle(mesh, convection="linearUpwind")
    phi = lin.flux(U[0], Ub[0])
    u, ub = U[1].reshape(-1, 3), Ub[1].reshape(-1, 3)
    o, n, nI = mesh.owner[:mesh.nInt], mesh.neighbour, mesh.nInt
    up = np.where(phi[:nI] >= 0, o, n)
    T = lin.grad_vec(U[1], Ub[1]).reshape(-1, 3, 3)
    for orc, face_val in ((upw, u[up]), (lup, u[up] + np.einsum("fi,fid->fd", mesh.Cf[:nI] - mesh.Cc[up], T[up]))):
        acc = np.zeros((mesh.nCells, 3))
        np.add.at(acc, o, phi[:nI, None] * face_val)
        np.add.at(acc, n, -phi[:nI, None] * face_val)
        on = mesh.bmask.astype(bool)
        np.add.at(acc, mesh.owner[nI:][on], (phi[nI:, None] * ub)[on])
        ref = (acc / mesh.V[:, None]).reshape(-1)
        got = orc.div_phi(phi, U[1], Ub[1])
        assert np.abs(got - ref).max() <= 1e-12 * np.abs(ref).max()
    assert np.abs(upw.div_phi(phi, U[1], Ub[1]) - lin.div_phi(phi, U[1], Ub[1])).max() > 1e-3 * np.abs(lin.div_phi(phi, U[1], Ub[1])).max()
    # constant field: div(phi, U0) = U0 div(phi) whatever the face interpolation (linearUpwind: grad of a constant is 0)
    c0 = np.array([0.3, -1.1, 2.0])
    Uc, Ucb = np.tile(c0, mesh.nCells), np.tile(c0, mesh.nBnd)
    divphi = np.zeros(mesh.nCells)
    np.add.at(divphi, o, phi[:nI]); np.add.at(divphi, n, -phi[:nI]); np.add.at(divphi, mesh.owner[nI:], phi[nI:] * mesh.bmask)
    divphi /= mesh.V
    for orc in (lin, upw, lup):
        d = orc.div_phi(phi, Uc, Ucb).reshape(-1, 3)
        assert np.abs(d - divphi[:, None] * c0).max() <= 1e-11 * max(np.abs(divphi).max(), 1.0) * np.abs(c0).max()
    # linear field, orthogonal uniform mesh: the upwind-side extrapolation to the face centre is exact, as is 'linear'
    um = pm.synthetic_hex(6, 5, 4, kind="box", seed=1, perturb=0.0)
    A = np.array([[0.7, -1.3, 2.1], [0.2, 0.5, -0.4], [1.0, 0.0, 0.3]])
    Ul, Ulb = (um.Cc @ A).reshape(-1), (um.Cf[um.nInt:] @ A).reshape(-1)
    Uq, Uqb, _, _ = fvc.synthetic_basis(um, 1, 0, seed=3)
    ph = fvc.Oracle(um).flux(Uq[0], Uqb[0])
    a, b = fvc.Oracle(um).div_phi(ph, Ul, Ulb), fvc.Oracle(um, convection="linearUpwind").div_phi(ph, Ul, Ulb)
    assert np.abs(a - b).max() <= 1e-11 * np.abs(a).max()
