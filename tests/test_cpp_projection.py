"""The C++ host shim of the projection (host/steadyNS_b200.H = steadyNS::projectSUP / projectPPE and their term builders
above the C ABI) built with g++ against plain views of the mesh and the basis. CPU: it compiles, links libithaca_b200.so
and aborts loudly (M_Assert behaviour, no fallback) without a GPU. GPU: the operator files it leaves under
ITHACAoutput/Matrices/ (the reference's names and binary layout) hold the oracle's operators to 1e-10, and a second run
reads them back instead of recomputing (the reference's cache-or-compute branch, steadyNS.C:384-467)."""
import os
import struct
import subprocess

import numpy as np
import pytest

from oracle import fvc, polymesh as pm

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "tests", "cpp", "test_projectsup")
LIBDIR = os.path.join(ROOT, "ithaca-fv_b200", "lib")
GOLD = os.path.join(os.path.dirname(__file__), "golden")
TOL = 1e-10


def build():
    src = os.path.join(ROOT, "tests", "cpp", "test_projectsup.cpp")
    hdr = os.path.join(ROOT, "host", "steadyNS_b200.H")
    if (not os.path.exists(BIN)) or os.path.getmtime(BIN) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.check_call(["g++", "-O2", "-std=c++17", src, "-o", BIN, "-L" + LIBDIR, "-lithaca_b200",
                               "-Wl,-rpath," + LIBDIR, "-Wl,-rpath,/usr/local/cuda/lib64", "-L/usr/local/cuda/lib64"])
    return BIN


def write_input(path, mesh, lift, U, sup, P, NU, NP, NSUP, ppe=False, penalty=False, inlet=(), exports=0, k=0.0, psi=-1.0):
    """lift / U / sup: lists of (internal, boundary) vector fields; P: scalar fields."""
    c = lambda a, dt=np.float64: np.ascontiguousarray(a, dtype=dt).tobytes()
    patches = mesh.patch_slices()
    with open(path, "wb") as f:
        f.write(struct.pack("16q", mesh.nCells, mesh.nInt, mesh.nBnd, len(lift), len(U), len(sup), len(P), NU, NP, NSUP, 1,
                            int(ppe), int(penalty), len(patches), len(inlet), exports))
        f.write(struct.pack("2d", k, psi))
        f.write(c(mesh.owner, np.int32) + c(mesh.neighbour, np.int32) + c(mesh.Sf) + c(mesh.magSf) + c(mesh.w) + c(mesh.dc)
                + c(mesh.dcB) + c(mesh.corr) + c(mesh.V) + c(mesh.bmask, np.uint8) + c(mesh.Cc) + c(mesh.Cf))
        f.write(c([s.start for _, _, s in patches], np.int64) + c([s.stop - s.start for _, _, s in patches], np.int64))
        f.write(c(np.array(inlet, dtype=np.int64).reshape(-1), np.int64))
        for fi, fb in list(lift) + list(U) + list(sup) + list(P):
            f.write(c(fi) + c(fb))


def case():
    mesh = pm.load_compact(os.path.join(GOLD, "mesh_04unsteadyNS.npz"))
    Uf, Ub, Pf, Pb = fvc.synthetic_basis(mesh, 8, 3, seed=31)
    fields = list(zip(Uf, Ub))
    return mesh, fields[:1], fields[1:6], fields[6:8], list(zip(Pf, Pb))     # 1 lift, 5 U modes, 2 supremizers, 3 P modes


def test_projection_shim_builds_and_fails_loudly_without_gpu(tmp_path):
    import torch
    exe = build()
    if torch.cuda.is_available():
        pytest.skip("GPU present: the loud-failure path is for GPU-less hosts")
    mesh, lift, U, sup, P = case()
    write_input(tmp_path / "in.bin", mesh, lift, U, sup, P, 2, 2, 1)
    r = subprocess.run([exe, str(tmp_path / "in.bin"), str(tmp_path / "Matrices") + "/"], capture_output=True, text=True)
    assert r.returncode != 0
    assert "no CPU fallback" in r.stderr


@pytest.mark.gpu
def test_projectSUP_files_match_oracle_and_are_read_back(pkg, tmp_path):
    exe = build()
    mesh, lift, U, sup, P = case()
    NU, NP, NSUP = 4, 3, 2
    name, typ, sl = [p for p in mesh.patch_slices() if p[1] != "empty" and p[2].stop > p[2].start][0]
    pidx = [p[0] for p in mesh.patch_slices()].index(name)
    write_input(tmp_path / "in.bin", mesh, lift, U, sup, P, NU, NP, NSUP, penalty=True, inlet=[(pidx, 0), (pidx, 1)], exports=1 | 4 | 8)
    mdir = str(tmp_path / "ITHACAoutput" / "Matrices") + "/"
    r = subprocess.run([exe, str(tmp_path / "in.bin"), mdir], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "computed 5 read 0" in r.stdout
    L = lift + U[:NU] + sup[:NSUP]                      # L_U_SUPmodes (steadyNS.C:561-585)
    Uf, Ub = [a for a, _ in L], [b for _, b in L]
    Pf, Pb = [a for a, _ in P[:NP]], [b for _, b in P[:NP]]
    orc = fvc.Oracle(mesh, corrected=True)
    rel = lambda A, B: np.abs(A - B).max() / np.abs(B).max()
    tag = "1_%d_%d" % (NU, NSUP)
    assert rel(pkg.ReadDenseMatrix(mdir, "B_" + tag), orc.diffusive_term(Uf, Ub)) <= TOL
    assert rel(pkg.ReadDenseMatrix(mdir, "M_" + tag), orc.mass_term(Uf)) <= TOL
    assert rel(pkg.ReadDenseMatrix(mdir, "K_%s_%d" % (tag, NP)), orc.pressure_gradient_term(Uf, Pf, Pb)) <= TOL
    assert rel(pkg.ReadDenseMatrix(mdir, "P_%s_%d" % (tag, NP)), orc.divergence_term(Uf, Ub, Pf)) <= TOL
    C = pkg.ReadDenseTensor(mdir, "C_%s_t" % tag)
    Cr = orc.convective_term_tens(Uf, Ub, cache=True)
    assert C.shape == Cr.shape and rel(C, Cr) <= TOL
    # exports: python text and cnpy .npy of the same operators
    assert np.array_equal(np.load(mdir + "C.npy"), C) and os.path.exists(mdir + "B_mat.py")
    # projectSUP's export block as coded (steadyNS.C:680, :698-699): the C tensor goes out in the "python" format under every
    # switch, into Matrices/ for exportPython and into Matrices/C for exportTxt; the matrices follow the switch
    assert os.path.exists(mdir + "C_mat.py") and os.path.exists(mdir + "C/C_mat.py") and os.path.exists(mdir + "B_mat.txt")
    assert not os.path.exists(mdir + "C0_mat.txt") and not os.path.exists(mdir + "C/C0_mat.txt")
    # penalty terms: the leading NU + NSUP fields of L_U_SUPmodes, one file per inletIndex row
    X = np.stack([b.reshape(-1, 3)[sl, 1] for b in Ub[:NU + NSUP]], axis=1)
    v1 = np.loadtxt(mdir + "bcVelVec/bcVelVec1_mat.txt")
    m1 = np.loadtxt(mdir + "bcVelMat/bcVelMat1_mat.txt")
    assert np.abs(v1 - X.sum(axis=0)).max() <= 1e-12 * np.abs(X).sum(axis=0).max() and rel(m1, X.T @ X) <= 1e-12
    # second run: every operator file exists -> read, nothing recomputed
    r2 = subprocess.run([exe, str(tmp_path / "in.bin"), mdir], capture_output=True, text=True)
    assert r2.returncode == 0 and "computed 0 read 5" in r2.stdout
    # a missing file is recomputed alone
    os.remove(mdir + "K_%s_%d" % (tag, NP))
    r3 = subprocess.run([exe, str(tmp_path / "in.bin"), mdir], capture_output=True, text=True)
    assert "computed 1 read 4" in r3.stdout


@pytest.mark.gpu
def test_projectPPE_with_tutorial04_schemes(pkg, tmp_path):
    """projectPPE (the method tutorial 04 selects) with that tutorial's fvSchemes (cellMDLimited gradient, limited snGrad)."""
    exe = build()
    mesh, lift, U, sup, P = case()
    rng = np.random.default_rng(3)
    P = [(a + 0.2 * rng.standard_normal(a.shape), b) for a, b in P]      # rough pressure modes: the limiter acts
    NU, NP = 5, 3
    write_input(tmp_path / "in.bin", mesh, lift, U, sup, P, NU, NP, 0, ppe=True, k=1.0, psi=1.0)
    mdir = str(tmp_path / "ITHACAoutput" / "Matrices") + "/"
    r = subprocess.run([exe, str(tmp_path / "in.bin"), mdir], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "computed 8 read 0" in r.stdout
    L = lift + U[:NU]
    Uf, Ub = [a for a, _ in L], [b for _, b in L]
    Pf, Pb = [a for a, _ in P[:NP]], [b for _, b in P[:NP]]
    orc = fvc.Oracle(mesh, corrected=True, grad_limit_k=1.0, sngrad_psi=1.0)
    rel = lambda A, B: np.abs(A - B).max() / np.abs(B).max()
    tag = "1_%d_0" % NU
    assert rel(pkg.ReadDenseMatrix(mdir, "B_" + tag), orc.diffusive_term(Uf, Ub)) <= TOL
    assert rel(pkg.ReadDenseMatrix(mdir, "K_%s_%d" % (tag, NP)), orc.pressure_gradient_term(Uf, Pf, Pb)) <= TOL
    assert rel(pkg.ReadDenseMatrix(mdir, "D_%d" % NP), orc.laplacian_pressure(Pf, Pb)) <= TOL
    assert rel(pkg.ReadDenseMatrix(mdir, "BC3_1_%d_%d" % (NU, NP)), orc.pressure_bc3(Uf, Ub, Pf, Pb)) <= TOL
    assert rel(pkg.ReadDenseMatrix(mdir, "BC4_1_%d_%d" % (NU, NP)), orc.pressure_bc4(Ub, Pb)) <= TOL
    G = pkg.ReadDenseTensor(mdir, "G_%s_%d_t" % (tag, NP))
    assert rel(G, orc.div_momentum(Uf, Ub, Pf, Pb)) <= TOL
    plain = fvc.Oracle(mesh, corrected=True)
    assert rel(pkg.ReadDenseMatrix(mdir, "K_%s_%d" % (tag, NP)), plain.pressure_gradient_term(Uf, Pf, Pb)) > 1e-6
