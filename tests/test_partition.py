"""CPU tests of the shard partitioner (host code of libithaca_b200.so) against oracle/partition.py: the addressing
arrays must agree integer for integer (BASELINE.json: 'face addressing and partitioning are bit-exact'), and the
halo must be deep enough that every operator field on OWNED cells equals the global-mesh value."""
import importlib
import os

import numpy as np
import pytest

from oracle import fvc, partition as op, polymesh as pm

GOLD = os.path.join(os.path.dirname(__file__), "golden")
pkg = importlib.import_module("ithaca-fv_b200")

MESHES = {
    "tutorial04": lambda: pm.load_compact(os.path.join(GOLD, "mesh_04unsteadyNS.npz")),
    "box": lambda: pm.synthetic_hex(9, 7, 5, kind="box", seed=3, perturb=0.15),
    "cyl": lambda: pm.synthetic_hex(12, 40, 6, kind="cylinder", seed=4, perturb=0.2),
}


@pytest.fixture(scope="module", params=sorted(MESHES))
def mesh(request):
    return MESHES[request.param]()


@pytest.mark.parametrize("nranks,halo", [(1, 2), (2, 1), (2, 2), (3, 2), (8, 2), (5, 0)])
def test_addressing_matches_oracle_bit_exact(mesh, nranks, halo):
    owned = np.zeros(mesh.nCells, dtype=np.int64)
    for rank in range(nranks):
        ref = op.partition(mesh.nCells, mesh.owner, mesh.neighbour, mesh.nInt, rank, nranks, halo)
        p = pkg.Partition(mesh, rank, nranks, halo)
        cpa, fpa = p.addressing()
        assert (p.nLocalCells, p.nOwned, p.nLocalInt, p.nLocalBnd, p.nCut) == (
            ref["cpa"].size, ref["nOwned"], ref["nLocalInt"], ref["nLocalBnd"], ref["nCut"])
        assert cpa.dtype == np.int32 and np.array_equal(cpa, ref["cpa"])
        assert fpa.dtype == np.int32 and np.array_equal(fpa, ref["fpa"])
        b, e = pkg.cell_range(mesh.nCells, rank, nranks)
        assert np.array_equal(cpa[:p.nOwned], np.arange(b, e))
        owned[cpa[:p.nOwned]] += 1
        if nranks == 1:
            assert p.nCut == 0 and p.nLocalCells == mesh.nCells
            assert np.array_equal(fpa, np.arange(1, mesh.nInt + mesh.nBnd + 1))
        p.close()
    assert np.all(owned == 1)  # every cell owned by exactly one rank


def test_locality_ordering_matches_restatement(mesh):
    """The library's Cuthill-McKee ordering equals the oracle's restatement label for label."""
    o2n = pkg.mesh_renumbering(mesh)
    ref = op.cuthill_mckee(mesh.nCells, mesh.owner, mesh.neighbour, mesh.nInt)
    assert o2n.dtype == np.int32 and np.array_equal(o2n, ref)
    assert np.array_equal(np.sort(o2n), np.arange(mesh.nCells))


@pytest.mark.parametrize("nranks", [2, 3, 8])
def test_locality_partition_matches_oracle_bit_exact(mesh, nranks):
    o2n = op.cuthill_mckee(mesh.nCells, mesh.owner, mesh.neighbour, mesh.nInt)
    owned = np.zeros(mesh.nCells, dtype=np.int64)
    shard_cells = 0
    for rank in range(nranks):
        ref = op.partition(mesh.nCells, mesh.owner, mesh.neighbour, mesh.nInt, rank, nranks, 2, old2new=o2n)
        p = pkg.Partition(mesh, rank, nranks, 2, method="locality")
        cpa, fpa = p.addressing()
        assert np.array_equal(cpa, ref["cpa"]) and np.array_equal(fpa, ref["fpa"])
        assert (p.nOwned, p.nLocalInt, p.nLocalBnd, p.nCut) == (ref["nOwned"], ref["nLocalInt"], ref["nLocalBnd"], ref["nCut"])
        b, e = pkg.cell_range(mesh.nCells, rank, nranks)
        assert np.array_equal(np.sort(o2n[cpa[:p.nOwned]]), np.arange(b, e))
        owned[cpa[:p.nOwned]] += 1
        shard_cells += p.nLocalCells
        p.close()
    assert np.all(owned == 1)
    # compact shards: the halo stays a fraction of the mesh (a label range of a randomly relabelled mesh drags in all of it)
    if nranks == 2:
        assert shard_cells < 1.7 * mesh.nCells


def test_owned_cells_see_all_their_faces(mesh):
    nranks, halo = 3, 2
    for rank in range(nranks):
        p = pkg.Partition(mesh, rank, nranks, halo)
        cpa, fpa = p.addressing()
        g = np.abs(fpa) - 1
        # cut faces only touch cells of the outermost halo layer: no owned cell on either side
        cut = g[p.nLocalInt + p.nLocalBnd:]
        b, e = pkg.cell_range(mesh.nCells, rank, nranks)
        for arr in (mesh.owner[cut], mesh.neighbour[cut[cut < mesh.nInt]]):
            assert not np.any((arr >= b) & (arr < e))
        # flipped sign exactly when the shard-side cell is the global neighbour
        inside = np.zeros(mesh.nCells, dtype=bool)
        inside[cpa] = True
        s = fpa[p.nLocalInt + p.nLocalBnd:]
        assert np.array_equal(s > 0, inside[mesh.owner[cut]])
        assert np.array_equal(s < 0, inside[mesh.neighbour[cut]])
        # number of faces per owned cell is the global one
        cnt_g = np.bincount(mesh.owner, minlength=mesh.nCells) + np.bincount(mesh.neighbour, minlength=mesh.nCells)
        loc = g[:p.nLocalInt + p.nLocalBnd]
        own_l = mesh.owner[loc]
        nei_l = mesh.neighbour[loc[:p.nLocalInt]]
        cnt_l = np.bincount(own_l, minlength=mesh.nCells) + np.bincount(nei_l, minlength=mesh.nCells)
        assert np.array_equal(cnt_l[b:e], cnt_g[b:e])
        p.close()


def test_gather_matches_oracle(mesh):
    rng = np.random.default_rng(5)
    p = pkg.Partition(mesh, 1, 3, 2)
    sm = op.ShardMesh(mesh, op.partition(mesh.nCells, mesh.owner, mesh.neighbour, mesh.nInt, 1, 3, 2))
    for dim in (1, 3):
        f = rng.standard_normal(mesh.nCells * dim)
        fb = rng.standard_normal(mesh.nBnd * dim)
        assert np.array_equal(p.gather(f, "cell", dim), sm.cell_field(f, dim))
        assert np.array_equal(p.gather(fb, "bnd", dim), sm.bnd_field(fb, dim))
    p.close()


def test_halo_depth_suffices_for_every_operator(mesh):
    """fvc operators evaluated by the oracle ON THE SHARD MESH equal the global ones on owned cells (2 halo layers:
    the corrected laplacian needs grad in the first layer), and the integrals add up over the ranks."""
    N, Np, nranks = 4, 2, 3
    U, Ub, P, Pb = fvc.synthetic_basis(mesh, N, Np, seed=7)
    org = fvc.Oracle(mesh, corrected=True)
    lap_g = org.laplacian(U[1], Ub[1]).reshape(-1, 3)
    phi_g = org.flux(U[2], Ub[2])
    div_g = org.div_phi(phi_g, U[3], Ub[3]).reshape(-1, 3)
    grad_g = org.grad(P[1], Pb[1]).reshape(-1, 3)
    B_g = org.diffusive_term(U, Ub)
    B_sum = np.zeros_like(B_g)
    for rank in range(nranks):
        part = op.partition(mesh.nCells, mesh.owner, mesh.neighbour, mesh.nInt, rank, nranks, 2)
        sm = op.ShardMesh(mesh, part)
        ors = fvc.Oracle(sm, corrected=True)
        nO, cpa = sm.nOwned, part["cpa"]
        Us = [sm.cell_field(u, 3) for u in U]
        Ubs = [sm.bnd_field(u, 3) for u in Ub]
        lap = ors.laplacian(Us[1], Ubs[1]).reshape(-1, 3)
        scale = np.abs(lap_g).max()
        assert np.abs(lap[:nO] - lap_g[cpa[:nO]]).max() <= 1e-13 * scale
        phi = ors.flux(Us[2], Ubs[2])
        div = ors.div_phi(phi, Us[3], Ubs[3]).reshape(-1, 3)
        assert np.abs(div[:nO] - div_g[cpa[:nO]]).max() <= 1e-13 * np.abs(div_g).max()
        grad = ors.grad(sm.cell_field(P[1], 1), sm.bnd_field(Pb[1], 1)).reshape(-1, 3)
        assert np.abs(grad[:nO] - grad_g[cpa[:nO]]).max() <= 1e-13 * np.abs(grad_g).max()
        # owned-cell part of B: sum_c V_c U_i . lap(U_j) over owned cells only
        Vw = np.repeat(np.where(np.arange(sm.nCells) < nO, sm.V, 0.0), 3)
        L = np.stack([ors.laplacian(Us[j], Ubs[j]) for j in range(N)], axis=1)
        B_sum += np.stack(Us, axis=1).T @ (Vw[:, None] * L)
    assert np.abs(B_sum - B_g).max() <= 1e-12 * np.abs(B_g).max()


def test_partition_rejects_bad_arguments(mesh):
    with pytest.raises(pkg.IthacaB200Error):
        pkg.Partition(mesh, 3, 3, 2)
    bad = type("M", (), dict(nCells=mesh.nCells, nInt=mesh.nInt, nBnd=mesh.nBnd, owner=mesh.owner.copy(), neighbour=mesh.neighbour.copy()))
    bad.neighbour[0] = mesh.nCells
    with pytest.raises(pkg.IthacaB200Error):
        pkg.Partition(bad, 0, 2, 2)
