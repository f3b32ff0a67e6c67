#!/usr/bin/env python
"""Which 40x40 blocks of the Gram are wrong? usage: gram_debug.py cells:dim:m ..."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
pkg = importlib.import_module("ithaca-fv_b200")
dev = torch.device("cuda", 0)
for sh in sys.argv[1:]:
    cells, dim, m = (int(x) for x in sh.split(":"))
    pod = pkg.Pod(0)
    Nd = cells * dim
    g = torch.Generator(device=dev); g.manual_seed(1)
    S = torch.randn(m, Nd, generator=g, device=dev, dtype=torch.float64)
    V = torch.rand(cells, generator=g, device=dev, dtype=torch.float64) + 0.5
    torch.cuda.synchronize()  # the library's stream does not wait for torch's
    pod.set_volumes(V); pod.upload([S[j] for j in range(m)], dim, cells)
    G = pod.gram(want=True)
    ref = ((S * V.repeat_interleave(dim)[None, :]) @ S.T).cpu().numpy()
    sc = np.abs(ref).max()
    nb = (m + 39) // 40
    print(sh, "rel err", np.abs(G - ref).max() / sc)
    bad = []
    for i in range(nb):
        for j in range(i + 1):
            e = np.abs(G[40*i:40*i+40, 40*j:40*j+40] - ref[40*i:40*i+40, 40*j:40*j+40]).max() / sc
            if e > 1e-12:
                z = np.abs(G[40*i:40*i+40, 40*j:40*j+40]).max()
                bad.append((i, j, float("%.2g" % e), float("%.2g" % z)))
    print(" bad blocks (i,j,err,|G|max):", bad[:40], "count", len(bad), "of", nb * (nb + 1) // 2)
    pod.close()
