"""The C++ host shim (host/ITHACAPOD_b200.H = ITHACAPOD::getModes above the C ABI) built with g++ against a minimal
field type. CPU: it compiles, links libithaca_b200.so and aborts loudly (M_Assert behaviour, no fallback) without a GPU.
GPU: modes, eigenvalue files and patch values match the oracle."""
import os
import struct
import subprocess

import numpy as np
import pytest

from oracle import pod_oracle as po, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "tests", "cpp", "test_getmodes")
LIBDIR = os.path.join(ROOT, "ithaca-fv_b200", "lib")


def build():
    src = os.path.join(ROOT, "tests", "cpp", "test_getmodes.cpp")
    hdr = os.path.join(ROOT, "host", "ITHACAPOD_b200.H")
    if (not os.path.exists(BIN)) or os.path.getmtime(BIN) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.check_call(["g++", "-O2", "-std=c++17", src, "-o", BIN, "-L" + LIBDIR, "-lithaca_b200",
                               "-Wl,-rpath," + LIBDIR, "-Wl,-rpath,/usr/local/cuda/lib64", "-L/usr/local/cuda/lib64"])
    return BIN


def write_input(path, S, V, dim, nmodes, patches, solver, norm, podex=0):
    """patches: list of (rows x m array, lifted flag); rows = faces*dim."""
    Nd, m = S.shape
    with open(path, "wb") as f:
        f.write(struct.pack("8q", Nd // dim, dim, m, nmodes, len(patches), 0 if solver == "spectra" else 1,
                            0 if norm == "L2" else 1, podex))
        f.write(struct.pack("%dq" % len(patches), *[p[0].shape[0] // dim for p in patches]))
        f.write(struct.pack("%dq" % len(patches), *[int(p[1]) for p in patches]))
        f.write(np.ascontiguousarray(V, dtype=np.float64).tobytes())
        for j in range(m):
            f.write(np.ascontiguousarray(S[:, j]).tobytes())
            for p, _ in patches:
                f.write(np.ascontiguousarray(p[:, j]).tobytes())


def read_output(path, Nd, patch_rows):
    raw = open(path, "rb").read()
    nm, ne, exported, reads = struct.unpack("4q", raw[:32])
    a = np.frombuffer(raw[32:], dtype=np.float64)
    ev, cum, nf = a[:ne], a[ne:2 * ne], a[2 * ne:2 * ne + nm]
    rest = a[2 * ne + nm:]
    per = Nd + sum(patch_rows)
    modes = rest.reshape(nm, per) if nm else np.zeros((0, per))
    return dict(nm=nm, ev=ev, cum=cum, nf=nf, exported=exported, reads=reads, modes=modes[:, :Nd].T, patches=modes[:, Nd:].T)


def test_shim_builds_and_fails_loudly_without_gpu(tmp_path):
    import torch
    exe = build()
    if torch.cuda.is_available():
        pytest.skip("GPU present: the loud-failure path is for GPU-less hosts")
    S = synth.snapshots(50, 1, 6, 3, 1)
    V = synth.volumes(50, 1)
    write_input(tmp_path / "in.bin", S, V, 1, 2, [], "eigen", "L2")
    r = subprocess.run([exe, str(tmp_path / "in.bin"), str(tmp_path / "out.bin"), str(tmp_path / "out")], capture_output=True, text=True)
    assert r.returncode != 0
    assert "no CPU fallback" in r.stderr or "no CUDA device" in r.stderr


def test_shim_reads_existing_modes_branch(tmp_path):
    """podex != 0 -> 'Reading the existing modes' (ITHACAPOD.C:309-322): no device work at all."""
    exe = build()
    S = synth.snapshots(50, 1, 6, 3, 1)
    V = synth.volumes(50, 1)
    write_input(tmp_path / "in.bin", S, V, 1, 2, [], "eigen", "L2", podex=1)
    r = subprocess.run([exe, str(tmp_path / "in.bin"), str(tmp_path / "out.bin"), str(tmp_path / "out")], capture_output=True, text=True)
    assert r.returncode == 0 and "Reading the existing modes" in r.stdout
    out = read_output(tmp_path / "out.bin", 50, [])
    assert out["reads"] == 1 and out["nm"] == 0


@pytest.mark.gpu
@pytest.mark.parametrize("solver,norm,dim", [("spectra", "L2", 3), ("eigen", "Frobenius", 1)])
def test_shim_matches_oracle(tmp_path, solver, norm, dim):
    exe = build()
    Nc, m, r = 3001, 30, 7
    S = synth.snapshots(Nc, dim, m, 2 * r, 5)
    V = synth.volumes(Nc, 5)
    rng = np.random.default_rng(3)
    patches = [(rng.standard_normal((11 * dim, m)), 1), (rng.standard_normal((5 * dim, m)), 0), (rng.standard_normal((8 * dim, m)), 1)]
    write_input(tmp_path / "in.bin", S, V, dim, r, patches, solver, norm)
    res = subprocess.run([exe, str(tmp_path / "in.bin"), str(tmp_path / "out.bin"), str(tmp_path / "out")], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    out = read_output(tmp_path / "out.bin", Nc * dim, [p[0].shape[0] for p in patches])
    Sbc = np.vstack([p[0] for p in patches])
    ref = po.getmodes(S, V, dim, r, solver, norm, Sbc)
    assert out["nm"] == r and out["exported"] == r
    assert len(out["ev"]) == len(ref["evals_norm"])
    assert np.all(np.abs(out["ev"][:r] - ref["evals_norm"][:r]) <= 1e-10 * ref["evals_norm"][:r])
    assert np.abs(out["cum"] - ref["cum"]).max() <= 1e-12
    A = po.sign_align(out["modes"], ref["modes"])
    assert (np.abs(A - ref["modes"]).max(axis=0) / np.abs(ref["modes"]).max(axis=0)).max() <= 1e-8
    s = np.sign(np.einsum("ij,ij->j", out["modes"], ref["modes"]))
    pb = out["patches"] * s
    r0 = 0
    for p, lifted in patches:
        rows = p.shape[0]
        if lifted:  # boundary rows of the modes = Sbc Q / normfact (ITHACAPOD.C:245-264)
            assert np.abs(pb[r0:r0 + rows] - ref["modes_bc"][r0:r0 + rows]).max() <= 1e-8 * np.abs(ref["modes_bc"]).max()
        else:       # untouched patch keeps the values of snapshots[0] (clone, ITHACAPOD.C:270)
            assert np.array_equal(out["patches"][r0:r0 + rows], np.repeat(p[:, :1], r, axis=1))
        r0 += rows
    # Eigenvalues_<name> file: MatrixMarket array (EigenFunctions.H:466-497)
    lines = open(tmp_path / "out" / "POD" / "Eigenvalues_U").read().split("\n")
    assert lines[0] == "%%MatrixMarket matrix array real general" and lines[1] == "%d 1" % len(ref["evals_norm"])
    vals = np.array([float(x) for x in lines[2:2 + len(ref["evals_norm"])]])
    assert np.all(np.abs(vals[:r] - ref["evals_norm"][:r]) <= 1e-10 * ref["evals_norm"][:r])
    assert os.path.exists(tmp_path / "out" / "POD" / "CumEigenvalues_U")


@pytest.mark.gpu
def test_shim_parallel_run_two_ranks(tmp_path):
    """Pstream::parRun() in the shim: two processes, one GPU and one cell shard each, the NCCL id carried by the
    shim's transport hook; every rank gets the global eigenvalues and its rows of the global modes."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import importlib
    pkg = importlib.import_module("ithaca-fv_b200")
    exe = build()
    Nc, dim, m, r = 4001, 3, 24, 6
    S = synth.snapshots(Nc, dim, m, 2 * r, 15)
    V = synth.volumes(Nc, 15)
    ref = po.getmodes(S, V, dim, r, "spectra", "L2")
    procs = []
    for rank in range(2):
        b, e = pkg.cell_range(Nc, rank, 2)
        write_input(tmp_path / ("in%d.bin" % rank), np.asfortranarray(S[b * dim:e * dim]), V[b:e], dim, r, [], "spectra", "L2")
        procs.append(subprocess.Popen([exe, str(tmp_path / ("in%d.bin" % rank)), str(tmp_path / ("out%d.bin" % rank)),
                                       str(tmp_path / ("out%d" % rank)), str(rank), "2", str(tmp_path / "nccl.id")],
                                      stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    outs = []
    for rank, p in enumerate(procs):
        so, se = p.communicate(timeout=300)
        assert p.returncode == 0, se[-1500:]
        b, e = pkg.cell_range(Nc, rank, 2)
        outs.append(read_output(tmp_path / ("out%d.bin" % rank), (e - b) * dim, []))
    modes = np.concatenate([o["modes"] for o in outs], axis=0)
    for o in outs:
        assert np.all(np.abs(o["ev"][:r] - ref["evals_norm"][:r]) <= 1e-10 * ref["evals_norm"][:r])
        assert np.all(np.abs(o["nf"] - ref["normfact"]) <= 1e-9 * ref["normfact"])
    A = po.sign_align(modes, ref["modes"])
    assert (np.abs(A - ref["modes"]).max(axis=0) / np.abs(ref["modes"]).max(axis=0)).max() <= 1e-8
