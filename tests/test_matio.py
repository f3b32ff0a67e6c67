"""On-disk formats of the reduced operators, host code behind the C ABI, no GPU needed.
The reference's writers need OpenFOAM headers (ITHACAstream.H includes fvCFD.H; cnpy.H includes IFstream.H) and cannot be
compiled here, so the checks are (i) byte layouts stated by the reference source (ITHACAstream.H:722-809: Eigen::Index
header + column-major doubles) rebuilt independently with numpy/struct, (ii) numpy itself reading the .npy files, and
(iii) the text exports evaluated / parsed back."""
import os
import struct

import numpy as np
import pytest


@pytest.fixture()
def folder(tmp_path):
    return str(tmp_path) + "/ITHACAoutput/Matrices/"


def test_dense_matrix_bytes_and_roundtrip(pkg, folder):
    rng = np.random.default_rng(3)
    for shape in [(5, 3), (1, 7), (4, 1), (0, 0)]:
        A = rng.standard_normal(shape)
        pkg.SaveDenseMatrix(A, folder, "B_3_2")
        raw = open(folder + "B_3_2", "rb").read()
        # rows, cols as 8-byte Eigen::Index, then Matrix.data() (column-major)
        assert raw == struct.pack("<qq", *shape) + np.asfortranarray(A).tobytes(order="F")
        B = pkg.ReadDenseMatrix(folder, "B_3_2")
        assert B.shape == shape and np.array_equal(A, B)


def test_dense_tensor_bytes_and_roundtrip(pkg, folder):
    os.makedirs(folder)
    T = np.random.default_rng(4).standard_normal((3, 4, 5))
    pkg.SaveDenseTensor(T, folder, "C_3_2_t")
    raw = open(folder + "C_3_2_t", "rb").read()
    # DSizes<Index,3>, then Tensor.data(): C(i,j,k) at i + 3*(j + 4*k)
    assert raw[:24] == struct.pack("<qqq", 3, 4, 5)
    flat = np.frombuffer(raw[24:], dtype="<f8")
    assert flat[1 + 3 * (2 + 4 * 3)] == T[1, 2, 3]
    assert np.array_equal(pkg.ReadDenseTensor(folder, "C_3_2_t"), T)


def test_read_missing_file_fails_like_reference(pkg, folder):
    with pytest.raises(pkg.IthacaB200Error, match="rerun the Offline Stage"):
        pkg.ReadDenseMatrix(folder, "nope")
    with pytest.raises(pkg.IthacaB200Error):
        pkg.SaveDenseTensor(np.zeros((1, 1, 1)), folder + "missing_dir/", "C")  # SaveDenseTensor does not mkDir


def test_npy_is_read_by_numpy(pkg, tmp_path):
    rng = np.random.default_rng(5)
    for shape in [(6,), (4, 3), (2, 3, 4), (50, 50, 50), (1, 10)]:
        A = rng.standard_normal(shape)
        f = str(tmp_path / "A.npy")
        pkg.save_npy(A, f)
        B = np.load(f)
        assert B.shape == shape and not np.isfortran(B) and np.array_equal(A, B)
        raw = open(f, "rb").read()
        hl = struct.unpack("<H", raw[8:10])[0]
        # cnpy's header: version 1.0, preamble + dictionary a multiple of 16, newline-terminated
        assert raw[:8] == b"\x93NUMPY\x01\x00" and (10 + hl) % 16 == 0 and raw[10 + hl - 1:10 + hl] == b"\n"
        assert raw[10:10 + hl].startswith(b"{'descr': '<f8', 'fortran_order': False, 'shape': (")


def test_export_matrix_text_formats(pkg, folder):
    A = np.array([[1.0, -2.5, 1.0 / 3.0], [1e-12, 12345678.9012345, 0.0]])
    pkg.exportMatrix(A, "B", "python", folder)
    txt = open(folder + "/B_mat.py").read()
    assert txt == "B=np.array([[1,-2.5,0.3333333333],\n[1e-12,12345678.9,0]])\n"
    ns = {"np": np}
    exec(txt, ns)
    assert np.allclose(ns["B"], A, rtol=1e-9)
    pkg.exportMatrix(A, "B", "matlab", folder)
    assert open(folder + "/B_mat.m").read() == "B=[ 1 -2.5 0.3333333333;\n 1e-12 12345678.9 0];\n"
    pkg.exportMatrix(A, "B", "eigen", folder)
    txt = open(folder + "/B_mat.txt").read()
    assert not txt.endswith("\n")
    assert np.array_equal(np.loadtxt(folder + "/B_mat.txt"), A)  # precision 20 round-trips doubles
    with pytest.raises(pkg.IthacaB200Error, match="was not implemented"):
        pkg.exportMatrix(A, "B", "csv", folder)


def test_export_tensor_text_formats(pkg, folder):
    T = np.random.default_rng(6).standard_normal((2, 3, 4))
    pkg.exportTensor(T, "C", "python", folder)
    ns = {"np": np}
    exec(open(folder + "/C_mat.py").read(), ns)
    assert ns["C"].shape == (2, 3, 4) and np.allclose(ns["C"], T, rtol=1e-9)
    pkg.exportTensor(T, "C", "matlab", folder)
    txt = open(folder + "/C_mat.m").read()
    assert txt.startswith("C(1,:,:)=[ ") and txt.count("];\n") == 2
    pkg.exportTensor(T, "C", "eigen", folder)
    for i in range(2):
        assert np.array_equal(np.loadtxt(folder + "/C%d_mat.txt" % i), T[i])
