#!/usr/bin/env python
"""C4 projection stages under different cluster sizes of the two-level cell ordering (ITHACA_B200_RENUMBER_BLOCK)."""
import importlib, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench_proj
from oracle import polymesh as pm
pkg = importlib.import_module("ithaca-fv_b200")
dev = torch.device("cuda", 0)
pod = pkg.Pod(0)
nr, nth, nz = bench_proj.DIMS[sys.argv[1] if len(sys.argv) > 1 else "c4"]
mesh = pm.synthetic_hex(nr, nth, nz, kind="cylinder", seed=20251104, perturb=0.2)
U, Ub, P, Pb = bench_proj.device_basis(torch, dev, mesh, 50, 10, 4)
pr = pkg.Projector(pod)
ref = None
for blk in [int(x) for x in os.environ.get("BLOCKS", "0,256,64,1024,4096").split(",")]:
    os.environ["ITHACA_B200_RENUMBER_BLOCK"] = str(blk)
    pr.set_mesh(mesh, corrected=True)
    pr.set_basis("U", U, Ub)
    pr.set_basis("P", P, Pb)
    out = {"block": blk}
    for op, key in (("B", "laplacian"), ("K", "grad_p"), ("P", "div_u"), ("C", "face_records"), ("Cmat", "div_phi")):
        ts = []
        for _ in range(2):
            if op == "Cmat":
                pr.set_basis("U", U, Ub)
            R = pr.project(op)
            ts.append(pr.timings()[key])
        out[key] = min(ts)
        if op == "P":
            out["flux"] = pr.timings()["flux"]
        if op == "B":
            if ref is None:
                ref = R
            out["B_rel_vs_first"] = float(np.abs(R - ref).max() / np.abs(ref).max())
    print(json.dumps(out), flush=True)
