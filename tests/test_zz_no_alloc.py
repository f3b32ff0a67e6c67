"""The allocation-free entry point ithaca_b200_col_norms_noalloc (no counterpart in the reference):
called through the C ABI, two calls in a row, checked against the norm factors of the existing entry point
ithaca_b200_pod_lift at the tolerance the lift tests use (1e-12 relative). Collected last on purpose (
a failure here must not hide other tests). Passes alone in a fresh process and after the other tests."""
import importlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _case(pkg, pod, torch, dev, Nc, dim, m, r, seed, norm):
    """Unnormalised modes S Q on the device + the existing entry point's norm factors of exactly those columns."""
    rng = np.random.default_rng(seed)
    S = np.asfortranarray(rng.standard_normal((Nc * dim, m)))
    V = rng.random(Nc) + 0.5
    pod.set_volumes(V)
    pod.upload(S, dim)
    pod.gram(want=False)
    pod.eig(r)
    Phi = pod.lift_scaled(np.ones(r))                 # S Q, unnormalised (existing entry point pod_lift_scaled)
    _, nf = pod.lift(r, norm, want=False)             # existing entry point: sqrt(phi^T diag(V) phi) / sqrt(phi^T phi)
    Nd = Nc * dim
    ld = (Nd + 15) // 16 * 16
    dPhi = torch.zeros((r, ld), device=dev, dtype=torch.float64)
    dPhi[:, :Nd] = torch.from_numpy(np.ascontiguousarray(Phi.T)).to(dev)
    dV = torch.from_numpy(V).to(dev)
    return dPhi, dV, Nd, ld, nf


def test_col_norms_noalloc_matches_pod_lift_on_two_calls_in_a_row():
    torch = pytest.importorskip("torch")
    pkg = importlib.import_module("ithaca-fv_b200")
    dev = torch.device("cuda", 0)
    pod = pkg.Pod(0)
    try:
        # inputs and outputs of both calls are prepared first: between the two calls nothing else touches the GPU
        a = _case(pkg, pod, torch, dev, 6011, 3, 40, 37, 1, "L2")
        b = _case(pkg, pod, torch, dev, 3001, 1, 24, 9, 2, "Frobenius")    # fewer columns, fewer row slabs: stale scratch around it
        out_a = torch.full((37,), -1.0, device=dev, dtype=torch.float64)
        out_b = torch.full((9,), -1.0, device=dev, dtype=torch.float64)
        torch.cuda.synchronize()
        dPhi, dV, Nd, ld, nf_a = a
        pod.col_norms_noalloc(dPhi.data_ptr(), Nd, ld, 37, dV.data_ptr(), 3, "L2", out_a.data_ptr())
        free0 = torch.cuda.mem_get_info(0)[0]
        dPhi2, dV2, Nd2, ld2, nf_b = b
        pod.col_norms_noalloc(dPhi2.data_ptr(), Nd2, ld2, 9, dV2.data_ptr(), 1, "Frobenius", out_b.data_ptr())
        free1 = torch.cuda.mem_get_info(0)[0]
        ra, rb = out_a.cpu().numpy(), out_b.cpu().numpy()
        assert np.all(np.abs(ra - nf_a) <= 1e-12 * nf_a), np.abs(ra / nf_a - 1).max()
        assert np.all(np.abs(rb - nf_b) <= 1e-12 * nf_b), np.abs(rb / nf_b - 1).max()
        assert free1 == free0                                              # the second call allocated no device memory
        # the first case again after the smaller one: the shared scratch keeps stale values, the result must not
        out_a.fill_(-1.0)
        torch.cuda.synchronize()
        pod.col_norms_noalloc(dPhi.data_ptr(), Nd, ld, 37, dV.data_ptr(), 3, "L2", out_a.data_ptr())
        assert np.array_equal(out_a.cpu().numpy(), ra)                     # fixed summation order: bit-reproducible
        # caller-supplied stream: enqueue only, the caller synchronises
        st = torch.cuda.Stream(device=dev)
        out_b.fill_(-1.0)
        torch.cuda.synchronize()
        pod.col_norms_noalloc(dPhi2.data_ptr(), Nd2, ld2, 9, dV2.data_ptr(), 1, "Frobenius", out_b.data_ptr(), stream=st.cuda_stream)
        st.synchronize()
        assert np.array_equal(out_b.cpu().numpy(), rb)
    finally:
        pod.close()


def test_col_norms_noalloc_rejects_sizes_beyond_the_fixed_scratch():
    torch = pytest.importorskip("torch")
    pkg = importlib.import_module("ithaca-fv_b200")
    dev = torch.device("cuda", 0)
    pod = pkg.Pod(0)
    try:
        n = 257
        dPhi = torch.zeros((n, 64), device=dev, dtype=torch.float64)
        dV = torch.ones(64, device=dev, dtype=torch.float64)
        out = torch.zeros(n, device=dev, dtype=torch.float64)
        with pytest.raises(pkg.IthacaB200Error, match="exceed the fixed scratch of IB200_NOALLOC_MAX_COLS = 256"):
            pod.col_norms_noalloc(dPhi.data_ptr(), 64, 64, n, dV.data_ptr(), 1, "L2", out.data_ptr())
        pod.col_norms_noalloc(dPhi.data_ptr(), 64, 64, 256, dV.data_ptr(), 1, "L2", out.data_ptr())   # the maximum itself is accepted
        with pytest.raises(pkg.IthacaB200Error):
            pod.col_norms_noalloc(dPhi.data_ptr(), 64, 32, 4, dV.data_ptr(), 1, "L2", out.data_ptr())  # ld < nRows
        with pytest.raises(pkg.IthacaB200Error):
            pod.col_norms_noalloc(dPhi.data_ptr(), 64, 64, 4, None, 1, "L2", out.data_ptr())           # L2 without volumes
    finally:
        pod.close()
