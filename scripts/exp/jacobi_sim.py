"""Sweep-count simulation of the one-sided Jacobi eigensolve (vectorised round-robin rounds) on benchmark-like Gram matrices:
which stopping / rotation criteria and which preconditioning cut sweeps. python scripts/exp/jacobi_sim.py [m]"""
import sys
import numpy as np


def bench_gram(m, Nd=30000, R=64, eps=1e-6, seed=0):
    rng = np.random.default_rng(seed)
    coef, _ = np.linalg.qr(rng.standard_normal((m, R)))
    x = (np.arange(Nd) + 0.5) / Nd
    U = np.stack([np.sin((k + 1) * np.pi * x) / (k + 1) for k in range(R)], axis=1)
    S = U @ coef.T + eps * rng.standard_normal((Nd, m))
    V = np.exp(0.5 * rng.standard_normal(Nd)); V /= V.sum()
    return S.T @ (V[:, None] * S)


def rounds(n):
    nn = n + (n & 1)
    out = []
    for rd in range(nn - 1):
        I, J = [], []
        for p in range(nn // 2):
            if p == 0: i, j = nn - 1, rd
            else: i, j = (rd + p) % (nn - 1), (rd - p + nn - 1) % (nn - 1)
            if i >= n or j >= n: continue
            if i > j: i, j = j, i
            I.append(i); J.append(j)
        out.append((np.array(I), np.array(J)))
    return out


def jacobi(X, r=None, mode="rel", early=False, maxsweeps=40, absfac=1.0):
    """mode 'rel': rotate iff |c| > tol sqrt(ab). 'hybrid': pairs whose LARGER norm^2 is below thr_r (the r-th largest current
    norm^2 times 1e-2) use the absolute test |c| > absfac * m eps max_norm^2 instead."""
    m, n = X.shape; X = X.copy()
    eps = 2.0 ** -52
    tol = np.sqrt(m) * eps
    R = rounds(n); hist = []; maxcos_hist = []
    for sweep in range(maxsweeps):
        rots = 0; maxcos = 0.0
        nrm = (X * X).sum(0)
        thr = np.sort(nrm)[::-1][min(r, n) - 1] * 1e-2 if (mode == "hybrid" and r) else 0.0
        amax = nrm.max()
        for I, J in R:
            xi, xj = X[:, I], X[:, J]
            a = (xi * xi).sum(0); b = (xj * xj).sum(0); c = (xi * xj).sum(0)
            ab = a * b
            cosv = np.where(ab > 0, np.abs(c) / np.sqrt(np.maximum(ab, 1e-300)), 0.0)
            need = (ab > 0) & (np.abs(c) > tol * np.sqrt(ab))
            if mode == "hybrid":
                small = np.maximum(a, b) < thr
                need &= ~small | (np.abs(c) > absfac * m * eps * amax)
            if need.any():
                maxcos = max(maxcos, cosv[need].max())
            d = b - a
            with np.errstate(all="ignore"):
                t = 2 * c / (d + np.copysign(np.sqrt(d * d + 4 * c * c), d))
            t = np.where(need, t, 0.0)
            cs = 1 / np.sqrt(1 + t * t); sn = cs * t
            X[:, I] = cs * xi - sn * xj; X[:, J] = sn * xi + cs * xj
            rots += int(need.sum())
        hist.append(rots); maxcos_hist.append(maxcos)
        if rots == 0: break
        if early and maxcos < 1e-8: break
    return X, hist, maxcos_hist


def report(name, G, X, hist, r, perm=None):
    w, Qr = np.linalg.eigh(G); w = w[::-1]; Qr = Qr[:, ::-1]
    lam = (X * X).sum(0); o = np.argsort(-lam); lam = lam[o]
    U = X[:, o] / np.sqrt(lam)
    if perm is not None:
        Q = np.zeros_like(U); Q[perm] = U; U = Q
    s = np.sign((U[:, :r] * Qr[:, :r]).sum(0))
    print("%-34s sweeps %2d  evals top-r rel %.1e  all abs/l1 %.1e  modes top-r %.1e  orth top-r %.1e  orth all %.1e  rots %s" % (
        name, len(hist), np.abs(lam[:r] / w[:r] - 1).max(), np.abs(lam - w).max() / w[0], np.abs(U[:, :r] * s - Qr[:, :r]).max(),
        np.abs(U[:, :r].T @ U[:, :r] - np.eye(r)).max(), np.abs(U.T @ U - np.eye(len(lam))).max(), hist[-4:]))


if __name__ == "__main__":
    m = int(sys.argv[1]) if len(sys.argv) > 1 else 200
    r = max(10, m // 10)
    G = bench_gram(m)
    sc = 1.0 / np.diag(G).max()
    L = np.linalg.cholesky(G * sc) / np.sqrt(sc)
    for name, kw in [("baseline rel", {}), ("rel + early exit", dict(early=True)), ("hybrid r", dict(mode="hybrid", r=r)),
                     ("hybrid r + early", dict(mode="hybrid", r=r, early=True))]:
        X, hist, mc = jacobi(L, **kw)
        report(name, G, X, hist, r)
    # de Rijk: columns sorted by decreasing diagonal before the factorisation
    p = np.argsort(-np.diag(G)); Lp = np.linalg.cholesky(G[np.ix_(p, p)])
    X, hist, mc = jacobi(Lp)
    report("presorted diag + rel", G, X, hist, r, perm=p)
    # second preconditioning step: Cholesky of L^T L (one LR step), Jacobi on L2, back-transform U = L V / sigma
    L2 = np.linalg.cholesky(L.T @ L)
    X, hist, mc = jacobi(L2)
    print("LR step then rel: sweeps", len(hist), hist[-4:])
