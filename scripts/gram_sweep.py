#!/usr/bin/env python
"""GPU-box sweep of the Gram kernel: correctness against torch FP64 matmul and TFLOP/s at several shapes.
usage: python scripts/gram_sweep.py [shape ...]   shape = cells:dim:m"""
import importlib, os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
pkg = importlib.import_module("ithaca-fv_b200")
if os.environ.get("GRAM_SWEEP_LIB"):
    pkg.LIB_PATH = os.environ["GRAM_SWEEP_LIB"]   # A/B of two builds on one box
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
shapes = sys.argv[1:] or ["10000:3:100", "4099:3:37", "2000000:1:200", "250000:3:1000", "100000:1:2000", "400000:3:50"]
peak = pkg.dmma_peak(0)
print("dmma peak %.2f TFLOP/s" % peak)
pod = pkg.Pod(0)
for sh in shapes:
    cells, dim, m = (int(x) for x in sh.split(":"))
    Nd = cells * dim
    g = torch.Generator(device=dev); g.manual_seed(1)
    S = torch.randn(m, Nd, generator=g, device=dev, dtype=torch.float64)
    V = torch.rand(cells, generator=g, device=dev, dtype=torch.float64) + 0.5
    torch.cuda.synchronize()  # the library's stream does not wait for torch's
    pod.set_volumes(V)
    pod.upload([S[j] for j in range(m)], dim, cells)
    G = pod.gram(want=True)
    ref = ((S * V.repeat_interleave(dim)[None, :]) @ S.T).cpu().numpy()
    err = np.abs(G - ref).max() / np.abs(ref).max()
    ts = []
    for _ in range(5):
        pod.gram(want=False)
        ts.append(pod.timings()["gram_kernel"])
    t = sorted(ts)[len(ts) // 2]
    tf = Nd * m * (m + 1) / (t * 1e-3) / 1e12
    print("cells=%d dim=%d m=%d  rel err %.2e  gram_kernel %.3f ms (min %.3f)  %.2f TFLOP/s = %.1f%% of DMMA peak" %
          (cells, dim, m, err, t, min(ts), tf, 100 * tf / peak), flush=True)
    del S, V
    torch.cuda.empty_cache()
