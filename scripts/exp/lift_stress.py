"""Repeat pass 2 (fill + lift) of the row-streamed POD on one Gram / eigensolve and compare every repetition bit for bit with the
first one: run-to-run differences, where (rows, columns) and how large. Debugging aid for an intermittent norm-factor mismatch."""
import importlib
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
pkg = importlib.import_module("ithaca-fv_b200")
if os.environ.get("LIFT_STRESS_LIB"):
    pkg.LIB_PATH = os.environ["LIFT_STRESS_LIB"]   # A/B of two builds on one box

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
chunk = int(sys.argv[2]) if len(sys.argv) > 2 else 90_001
dev = torch.device("cuda", 0)
Nc, dim, m, r, R, seed = 400_003, 3, 256, 24, 48, 7
Nd = Nc * dim
rng = np.random.default_rng(12)
coef, _ = np.linalg.qr(rng.standard_normal((m, R)))
dcoef = torch.tensor(np.asfortranarray(coef).T.copy(), device=dev)
torch.cuda.synchronize()
pod = pkg.Pod(0)
chunks = [(c0, min(chunk, Nc - c0)) for c0 in range(0, Nc, chunk)]


def fill(slot, c0, nc):
    Sp, ld, Vp = pod.stream_slot(slot, c0, nc)
    pod.synth_snapshots(Sp, ld, c0 * dim, nc * dim, Nd, m, dcoef.data_ptr(), R, 1e-6, seed)
    pod.synth_volumes(Vp, c0, nc, Nc, seed)


pod.stream_begin(m, dim, Nc)
for k, (c0, nc) in enumerate(chunks):
    fill(k & 1, c0, nc)
    pod.stream_gram_add(k & 1)
G = pod.stream_gram_finish()
ev, Q = pod.eig(r)
ref = None
nbad = 0
gbad = 0
for rep in range(reps):
    if os.environ.get("LIFT_STRESS_GRAM"):
        pod.stream_begin(m, dim, Nc)
        for k, (c0, nc) in enumerate(chunks):
            fill(k & 1, c0, nc)
            pod.stream_gram_add(k & 1)
        G2 = pod.stream_gram_finish()
        if not np.array_equal(G2, G):
            gbad += 1
            print("rep", rep, "GRAM DIFF", np.abs(G2 - G).max() / np.abs(G).max(), np.argwhere(G2 != G)[:6].tolist(), flush=True)
        pod.eig(r)
    for k, (c0, nc) in enumerate(chunks):
        fill(k & 1, c0, nc)
        pod.stream_lift_add(k & 1, r, "L2")
    Phi, nf = pod.stream_lift_finish(r)
    un = Phi * nf                      # unnormalised modes (up to one rounding)
    if ref is None:
        ref = (un.copy(), nf.copy())
        print("rep 0 nf2/ev-1", np.abs(nf * nf / ev[:r] - 1).max(), flush=True)
        continue
    d = np.abs(un - ref[0])
    tol = 1e-12 * np.abs(ref[0]).max()
    if d.max() > tol or not np.array_equal(nf, ref[1]):
        nbad += 1
        rows = np.nonzero(d.max(axis=1) > tol)[0]
        cols = np.nonzero(d.max(axis=0) > tol)[0]
        print("rep", rep, "DIFF max", d.max(), "nf rel", np.abs(nf / ref[1] - 1).max(), "rows", len(rows),
              (rows[:6], rows[-6:]) if len(rows) else None, "cols", cols, "nf2/ev-1", np.abs(nf * nf / ev[:r] - 1).max(), flush=True)
        if len(rows):
            runs = np.split(rows, np.nonzero(np.diff(rows) > 1)[0] + 1)
            print("   runs", [(int(a[0]), int(a[-1]), int(a[0]) % 128, int(a[0]) // 128) for a in runs[:12]], "nruns", len(runs), flush=True)
print("reps", reps, "bad", nbad, "gram bad", gbad)
pod.close()
