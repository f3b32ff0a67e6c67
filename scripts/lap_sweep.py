#!/usr/bin/env python
"""Laplacian stage (operator B) of the C4 projection under different ITHACA_B200_LAP_BATCH values (run under gpurun)."""
import importlib, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
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
pr.set_mesh(mesh, corrected=True)
pr.set_basis("U", U, Ub)
ref = None
for batch in [int(x) for x in os.environ.get("LAP_SWEEP", "25,8,50,0").split(",")]:   # 0 = cell-major pass, else field-major with that batch
    os.environ["ITHACA_B200_LAP"] = "fm" if batch else "cm"
    if batch:
        os.environ["ITHACA_B200_LAP_BATCH"] = str(batch)
    ts = []
    for _ in range(3):
        B = pr.project("B")
        ts.append(pr.timings()["laplacian"])
    if ref is None:
        ref = B
    print(json.dumps({"batch": batch, "laplacian_ms": ts, "same_as_first": bool((B == ref).all())}), flush=True)
