"""Convergence experiment: pivoted Cholesky G = P L L^T P^T, then one-sided Jacobi on the columns of L."""
import sys, numpy as np
sys.path.insert(0, "/root/repo")
from oracle import synth, pod_oracle as po

def pchol(G):
    m = G.shape[0]; A = G.copy(); L = np.zeros((m, m)); piv = np.arange(m); rank = m
    d = np.diag(A).copy()
    for k in range(m):
        p = k + np.argmax(d[k:])
        if d[p] <= 0: rank = k; break
        if p != k:
            A[[k, p]] = A[[p, k]]; A[:, [k, p]] = A[:, [p, k]]; L[[k, p]] = L[[p, k]]; piv[[k, p]] = piv[[p, k]]; d[[k, p]] = d[[p, k]]
        L[k, k] = np.sqrt(d[k])
        L[k+1:, k] = (A[k+1:, k] - L[k+1:, :k] @ L[k, :k]) / L[k, k]
        d[k+1:] -= L[k+1:, k] ** 2
    return L[:, :rank], piv, rank

def jacobi_cols(X, maxsweeps=40):
    m, n = X.shape; X = X.copy()
    tol = np.sqrt(m) * 2.0**-52
    nn = n + (n & 1); hist = []
    for sweep in range(maxsweeps):
        rots = 0
        for rd in range(nn - 1):
            for p in range(nn // 2):
                if p == 0: i, j = nn - 1, rd
                else: i, j = (rd + p) % (nn - 1), (rd - p + nn - 1) % (nn - 1)
                if i >= n or j >= n: continue
                if i > j: i, j = j, i
                a = X[:, i] @ X[:, i]; b = X[:, j] @ X[:, j]; c = X[:, i] @ X[:, j]
                if not a * b > 0 or abs(c) <= tol * np.sqrt(a * b): continue
                z = (b - a) / (2 * c); t = np.sign(z) / (abs(z) + np.sqrt(1 + z * z)) if z != 0 else 1.0
                cs = 1 / np.sqrt(1 + t * t); sn = cs * t
                xi, xj = X[:, i].copy(), X[:, j].copy(); X[:, i] = cs * xi - sn * xj; X[:, j] = sn * xi + cs * xj
                rots += 1
        hist.append(rots)
        if rots == 0: break
    return X, hist

if __name__ == "__main__":
    m = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    S = synth.snapshots(3000, 1, m, 20, 3); V = synth.volumes(3000, 3)
    G = po.gram(S, V, 1)
    w, Qr = np.linalg.eigh(G); w = w[::-1]; Qr = Qr[:, ::-1]
    L, piv, rank = pchol(G)
    print("rank", rank, "chol resid %.1e" % (np.abs(L @ L.T - G[np.ix_(piv, piv)]).max() / w[0]))
    X, hist = jacobi_cols(L)
    lam = (X * X).sum(0); o = np.argsort(-lam); lam = lam[o]
    U = X[:, o] / np.sqrt(lam); Q = np.zeros_like(U); Q[piv] = U
    print("sweeps", len(hist), hist, "top10 rel err %.1e" % np.abs(lam[:10] / w[:10] - 1).max(), "abs err/l1 %.1e" % (np.abs(lam - w[:rank]).max() / w[0]),
          "orth %.1e" % np.abs(Q.T @ Q - np.eye(rank)).max(), "resid %.1e" % (np.abs(G @ Q - Q * lam).max() / w[0]))
    s = np.sign((Q[:, :10] * Qr[:, :10]).sum(0))
    print("top10 evec err %.1e" % np.abs(Q[:, :10] * s - Qr[:, :10]).max())
    # unpivoted
    try:
        L0 = np.linalg.cholesky(G)
        X, hist = jacobi_cols(L0)
        lam = np.sort((X * X).sum(0))[::-1]
        print("unpivoted: sweeps", len(hist), hist, "top10 rel err %.1e" % np.abs(lam[:10] / w[:10] - 1).max())
    except np.linalg.LinAlgError as e:
        print("unpivoted cholesky failed", e)
    # sorted-diagonal (one-off permutation by initial diagonal) then unpivoted
    p = np.argsort(-np.diag(G)); L1 = np.linalg.cholesky(G[np.ix_(p, p)])
    X, hist = jacobi_cols(L1)
    lam = np.sort((X * X).sum(0))[::-1]
    print("presorted: sweeps", len(hist), hist, "top10 rel err %.1e" % np.abs(lam[:10] / w[:10] - 1).max())
