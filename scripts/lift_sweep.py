#!/usr/bin/env python
"""GPU-box timing of the lift kernel (Phi = S Q_r + weighted column norms): python scripts/lift_sweep.py [cells:dim:m:r ...]
Knobs for A/B runs: ITHACA_B200_LIFT_PERSM=1 (one CTA per SM, 8 stages), ITHACA_B200_LIFT_STAGES=n."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
pkg = importlib.import_module("ithaca-fv_b200")
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
shapes = sys.argv[1:] or ["1000000:3:1000:50", "2000000:1:200:20", "1500000:1:2000:100"]
peak = pkg.dmma_peak(0)
pod = pkg.Pod(0)
for sh in shapes:
    cells, dim, m, r = (int(x) for x in sh.split(":"))
    Nd = cells * dim
    g = torch.Generator(device=dev); g.manual_seed(1)
    S = torch.randn(m, Nd, generator=g, device=dev, dtype=torch.float64)
    V = torch.rand(cells, generator=g, device=dev, dtype=torch.float64) + 0.5
    torch.cuda.synchronize()
    pod.set_volumes(V)
    pod.upload([S[j] for j in range(m)], dim, cells)
    pod.gram(want=False)
    pod.eig(r)
    ts = []
    for _ in range(5):
        pod.lift(r, "L2", want=False)
        ts.append(pod.timings()["lift"])
    t = sorted(ts)[len(ts) // 2]
    rp = (r + 7) // 8 * 8
    print("cells=%d dim=%d m=%d r=%d  lift %.3f ms (min %.3f)  %.2f TB/s of S, %.2f TFLOP/s on %d padded columns = %.1f%% of DMMA peak" %
          (cells, dim, m, r, t, min(ts), Nd * m * 8 / (t * 1e-3) / 1e12, 2.0 * Nd * m * rp / (t * 1e-3) / 1e12, rp,
           100 * 2.0 * Nd * m * rp / (t * 1e-3) / 1e12 / peak), flush=True)
    del S, V
    torch.cuda.empty_cache()
