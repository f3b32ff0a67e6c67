#!/usr/bin/env python
"""A/B timing + accuracy of the eigensolve kernels on snapshot-like Gram matrices (run under gpurun).
python scripts/eig_bench.py [m ...]   -> one JSON line per (m, variant)."""
import importlib, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("ithaca-fv_b200")


def gram_like(m, seed=3):
    """G = A^T A of a low-rank-plus-noise snapshot set (spectrum 2^-k/2 down to a 1e-12 noise floor)."""
    rng = np.random.default_rng(seed)
    n = 4 * m
    R = min(m, 40)
    U = np.linalg.qr(rng.standard_normal((n, R)))[0]
    Vt = np.linalg.qr(rng.standard_normal((m, R)))[0]
    S = (U * 2.0 ** (-0.5 * np.arange(R))) @ Vt.T + 1e-6 * rng.standard_normal((n, m))
    return S.T @ S


def main():
    ms = [int(a) for a in sys.argv[1:]] or [200, 1000]
    pod = pkg.Pod(0)
    for m in ms:
        G = gram_like(m)
        w_ref = np.linalg.eigvalsh(G)[::-1]
        variants = [("reg", {}), ("reg_chol_blocked", {"ITHACA_B200_CHOL": "blocked"}), ("block", {"ITHACA_B200_EIG": "block"})]
        if m <= 1024:
            for b in (4, 8, 12):
                variants.append(("reg_b%d" % b, {"ITHACA_B200_EIG_B": str(b)}))
        for name, env in variants:
            for k in ("ITHACA_B200_EIG", "ITHACA_B200_EIG_B", "ITHACA_B200_CHOL"):
                os.environ.pop(k, None)
            os.environ.update(env)
            pod.set_gram(G)
            ts = []
            for _ in range(4):
                pod.set_gram(G)
                ev, Q = pod.eig(min(m, 50))
                ts.append(pod.timings()["eig"])
            r = Q.shape[1]
            res = np.abs(G @ Q - Q * ev[:r]).max() / w_ref[0]
            orth = np.abs(Q.T @ Q - np.eye(r)).max()
            rel = np.abs(ev - w_ref).max() / w_ref[0]
            print(json.dumps({"m": m, "variant": name, "eig_ms_min": min(ts[1:]), "eig_ms": ts, "sweeps": pod.sweeps,
                              "eval_err_rel_max": rel, "residual": res, "orth": orth}), flush=True)
    pod.close()


if __name__ == "__main__":
    main()
