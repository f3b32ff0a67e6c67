"""Convergence experiment (CPU, numpy): sweeps of one-sided Jacobi on W = G under different stopping rules."""
import sys, numpy as np
sys.path.insert(0, "/root/repo")
from oracle import synth, pod_oracle as po

def jacobi(G, mode="rel", maxsweeps=40, sort=False, kappa=1.0):
    m = G.shape[0]
    W = G.copy(); Q = np.eye(m)
    if sort:
        p = np.argsort(-np.diag(G)); W = W[:, p]; Q = Q[:, p]
    tol = np.sqrt(m) * 2.0**-52
    lam1 = np.sqrt((W * W).sum(0).max())
    n = m + (m & 1)
    rots_per_sweep = []
    for sweep in range(maxsweeps):
        rots = 0
        for rd in range(n - 1):
            for p in range(n // 2):
                if p == 0: i, j = n - 1, rd
                else: i, j = (rd + p) % (n - 1), (rd - p + n - 1) % (n - 1)
                if i >= m or j >= m: continue
                if i > j: i, j = j, i
                a = W[:, i] @ W[:, i]; b = W[:, j] @ W[:, j]; c = W[:, i] @ W[:, j]
                if not a * b > 0: continue
                if abs(c) <= tol * np.sqrt(a * b): continue
                if mode == "abs" and abs(c) <= kappa * 2.0**-52 * lam1 * (np.sqrt(a) + np.sqrt(b)): continue
                z = (b - a) / (2 * c); t = np.sign(z) / (abs(z) + np.sqrt(1 + z * z)) if z != 0 else 1.0
                cs = 1 / np.sqrt(1 + t * t); sn = cs * t
                wi, wj = W[:, i].copy(), W[:, j].copy(); W[:, i] = cs * wi - sn * wj; W[:, j] = sn * wi + cs * wj
                qi, qj = Q[:, i].copy(), Q[:, j].copy(); Q[:, i] = cs * qi - sn * qj; Q[:, j] = sn * qi + cs * qj
                rots += 1
        rots_per_sweep.append(rots)
        if rots == 0: break
    lam = (Q * W).sum(0)
    o = np.argsort(-lam)
    return lam[o], Q[:, o], rots_per_sweep

if __name__ == "__main__":
    m = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    S = synth.snapshots(3000, 1, m, 20, 3); V = synth.volumes(3000, 3)
    G = po.gram(S, V, 1)
    w = np.linalg.eigvalsh(G)[::-1]
    for mode, sort, kappa in [("abs", False, 0.25), ("abs", False, 1.0), ("abs", True, 1.0), ("abs", False, 8.0)]:
        lam, Q, rots = jacobi(G, mode, sort=sort, kappa=kappa)
        print(mode, "sort" if sort else "", kappa, "sweeps", len(rots), rots[:12], "...", "top10 rel err %.1e" % np.abs(lam[:10] / w[:10] - 1).max(),
              "abs err/l1 %.1e" % (np.abs(lam - w).max() / w[0]), "orth %.1e" % np.abs(Q.T @ Q - np.eye(m)).max(),
              "resid %.1e" % (np.abs(G @ Q - Q * lam).max() / w[0]))
