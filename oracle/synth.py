"""TEST INFRASTRUCTURE ONLY -- seeded synthetic snapshot sets.

S = sum_{k<R} sigma_k u_k v_k^T + eps*W : smooth spatial functions u_k of a hashed cell coordinate,
orthonormal temporal coefficients v_k, i.i.d. noise W. sigma_k = 1/(k+1): eigenvalues of the Gram decay
like (k+1)^-2, so the kept modes stay far above the FP64 noise floor of the reference's own
eigensolver (the 1e-10 / 1e-8 parity gates are only meaningful there) and neighbouring eigenvalues are
separated by >> 1e-8 * lambda_1.  Volumes V_c = exp(N(0, 0.5^2)) scaled to sum 1.
Vector fields: 3 interleaved components per cell (quirk Q6), each with its own u_k.
"""
import numpy as np


def volumes(Nc, seed):
    rng = np.random.default_rng(seed + 7)
    V = np.exp(0.5 * rng.standard_normal(Nc))
    return V / V.sum()


def snapshots(Nc, dim, m, R, seed, eps=1e-6):
    """Returns S (Nd, m) Fortran-ordered (each snapshot contiguous, like an OpenFOAM field)."""
    rng = np.random.default_rng(seed)
    Nd = Nc * dim
    x = rng.random(Nc)
    y = rng.random(Nc)
    R = min(R, m)
    Vt, _ = np.linalg.qr(rng.standard_normal((m, R)))
    S = np.zeros((Nd, m), order="F")
    for k in range(R):
        sig = 1.0 / (k + 1)
        u = np.empty(Nd)
        for d in range(dim):
            u[d::dim] = np.sin((k + 1) * np.pi * x + 0.3 * d) * np.cos((k + 2) * y + 0.7 * d)
        S += sig * np.outer(u, Vt[:, k])
    S += eps * rng.standard_normal((Nd, m))
    return np.asfortranarray(S)
