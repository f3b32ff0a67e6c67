#!/usr/bin/env python
"""A/B of the eigensolve's barrier (cluster vs cooperative grid) on benchmark-like Gram matrices: python scripts/eig_ab.py [m ...]"""
import importlib, json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "exp"))
from jacobi_sim import bench_gram
pkg = importlib.import_module("ithaca-fv_b200")
pod = pkg.Pod(0)
for m in [int(a) for a in sys.argv[1:]] or [100, 200, 256]:
    G = bench_gram(m, Nd=20000)
    w = np.linalg.eigvalsh(G)[::-1]
    for name, env in (("cluster", {}), ("grid", {"ITHACA_B200_EIG_SYNC": "grid"})):
        os.environ.pop("ITHACA_B200_EIG_SYNC", None)
        os.environ.update(env)
        ts = []
        for _ in range(5):
            pod.set_gram(G)
            ev, Q = pod.eig(m)
            ts.append(pod.timings()["eig"])
        print(json.dumps({"m": m, "sync": name, "eig_ms_min": min(ts[1:]), "eig_ms": ts, "sweeps": pod.sweeps,
                          "evals_abs_rel": float(np.abs(ev - w).max() / w[0]), "orth": float(np.abs(Q.T @ Q - np.eye(m)).max()),
                          "resid": float(np.abs(G @ Q - Q * ev).max() / w[0])}), flush=True)
