"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol the header declares,
its host-only entry points work, and compute calls fail loudly (no fallback) without a GPU."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "ithaca_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(ithaca_b200_\w+)\s*\(", txt)))


def test_library_exports_every_declared_symbol(pkg):
    L = pkg.lib()
    syms = header_symbols()
    assert len(syms) >= 18
    for s in syms:
        assert hasattr(L, s), "missing export " + s
    assert sorted(pkg.SYMBOLS) == syms


def test_no_torch_types_in_header():
    txt = open(os.path.join(ROOT, "include", "ithaca_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)  # signatures only, comments stripped
    assert "torch" not in txt and "at::" not in txt and "std::" not in txt


def test_cell_range_partition_is_exact(pkg):
    for n, R in [(10, 3), (20_000_000, 8), (7, 8), (50_000_001, 8)]:
        ranges = [pkg.cell_range(n, r, R) for r in range(R)]
        assert ranges[0][0] == 0 and ranges[-1][1] == n
        for a, b in zip(ranges[:-1], ranges[1:]):
            assert a[1] == b[0]
        for r, (b, e) in enumerate(ranges):
            assert b == (r * n) // R and e == ((r + 1) * n) // R


def test_postprocess_eigenvalues_matches_oracle(pkg):
    from oracle import pod_oracle as po
    ev = np.array([5.0, 2.0, 1.0, 0.1, -0.05])
    for solver, nm in (("eigen", 3), ("spectra", 3), ("eigen", 5)):
        G = np.diag(ev)
        ref_ev, _ = po.eig(G, nm, solver)
        ref_n, ref_c = po.normalise_eigenvalues(ref_ev)
        out_n, out_c = pkg.postprocess_eigenvalues(ev, nm, solver)
        assert np.allclose(out_n, ref_n, rtol=1e-15, atol=0) and np.allclose(out_c, ref_c, rtol=1e-15)


def test_nmodes_asserts_mirror_reference(pkg):
    with pytest.raises(pkg.IthacaB200Error, match="Snapshots - 2"):
        pkg.check_nmodes(10, 9, "spectra")
    with pytest.raises(pkg.IthacaB200Error, match="number of Snapshots"):
        pkg.check_nmodes(10, 11, "eigen")
    assert pkg.check_nmodes(10, 0, "spectra") == 8


def test_compute_fails_loudly_without_gpu(pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(pkg.IthacaB200Error):
        pkg.Pod(0)
    with pytest.raises(pkg.IthacaB200Error):
        pkg.dmma_peak(0)


def test_product_does_not_import_oracle():
    pat = re.compile(r"(import\s+oracle|from\s+oracle|oracle/_ref|libpod_ref|libfvc_oracle|#include\s+\"[^\"]*oracle)")
    for dirpath, _, files in os.walk(os.path.join(ROOT, "ithaca-fv_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".H", ".C", ".cpp")):
                assert not pat.search(open(os.path.join(dirpath, f)).read()), f
