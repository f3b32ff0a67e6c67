"""world_size-2 test of the N>1 path's host logic on CPU (gloo): the contiguous cell-range rule of the C ABI
(ithaca_b200_cell_range), ONE sum all-reduce of the m x m partial Grams and one of the r partial squared norms
reproduce the serial result (the reference's Pstream reduces: ITHACAcoeffsMass.C:112-115, ITHACAPOD.C:146-149,
:235-243). The per-shard arithmetic here is the oracle (no GPU in this container); the GPU path with NCCL is
exercised by bench.py --gpus N on the GPU box."""
import importlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, tmp):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pkg = importlib.import_module("ithaca-fv_b200")
    from oracle import pod_oracle as po, synth
    Nc, dim, m, r = 1001, 3, 12, 4
    S = synth.snapshots(Nc, dim, m, 2 * r, 42)
    V = synth.volumes(Nc, 42)
    b, e = pkg.cell_range(Nc, rank, world)
    Sl, Vl = S[b * dim:e * dim], V[b:e]
    G = torch.from_numpy(po.gram(Sl, Vl, dim))
    dist.all_reduce(G)  # the ONE Gram all-reduce
    ev, Q = po.eig(G.numpy(), r, "eigen")  # replicated eigensolve: identical input on every rank
    Phi = Sl @ Q
    n2 = torch.from_numpy(np.einsum("ij,i,ij->j", Phi, po.mass_fv(Vl, dim), Phi))
    dist.all_reduce(n2)  # partial squared norms, then sqrt (ITHACAPOD.C:235-243)
    Phi = Phi / np.sqrt(n2.numpy())
    np.savez(os.path.join(tmp, "rank%d.npz" % rank), b=b, e=e, G=G.numpy(), ev=ev, Phi=Phi)
    dist.destroy_process_group()


def test_two_rank_sharded_pod_equals_serial(tmp_path):
    sys.path.insert(0, ROOT)
    from oracle import pod_oracle as po, synth
    world, port = 2, 29500 + os.getpid() % 400
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    Nc, dim, m, r = 1001, 3, 12, 4
    S = synth.snapshots(Nc, dim, m, 2 * r, 42)
    V = synth.volumes(Nc, 42)
    ref = po.getmodes(S, V, dim, r, "eigen", "L2")
    parts = [np.load(tmp_path / ("rank%d.npz" % k)) for k in range(world)]
    assert parts[0]["b"] == 0 and parts[0]["e"] == parts[1]["b"] and parts[1]["e"] == Nc  # bit-exact partition
    assert np.array_equal(parts[0]["G"], parts[1]["G"])  # every rank holds the same reduced Gram
    assert np.abs(parts[0]["G"] - ref["G"]).max() <= 1e-13 * np.abs(ref["G"]).max()
    Phi = np.vstack([p["Phi"] for p in parts])
    A = po.sign_align(Phi, ref["modes"])
    assert (np.abs(A - ref["modes"]).max(axis=0) / np.abs(ref["modes"]).max(axis=0)).max() <= 1e-8


def _proj_worker(rank, world, port, tmp):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pkg = importlib.import_module("ithaca-fv_b200")
    from oracle import fvc, partition as op, polymesh as pm
    mesh = pm.synthetic_hex(8, 12, 4, kind="cylinder", seed=4, perturb=0.2)
    N = 4
    U, Ub, _, _ = fvc.synthetic_basis(mesh, N, 0, seed=3)
    part = pkg.Partition(mesh, rank, world, 2)          # the product's host partitioner
    cpa, fpa = part.addressing()
    sm = op.ShardMesh(mesh, dict(cpa=cpa, fpa=fpa, nOwned=part.nOwned, nLocalInt=part.nLocalInt, nLocalBnd=part.nLocalBnd,
                                 nCut=part.nCut))
    ors = fvc.Oracle(sm, corrected=True)                 # per-shard arithmetic: the oracle (no GPU here)
    Us = [part.gather(u, "cell", 3) for u in U]
    Ubs = [part.gather(u, "bnd", 3) for u in Ub]
    Vw = np.repeat(np.where(np.arange(sm.nCells) < part.nOwned, sm.V, 0.0), 3)
    L = np.stack([ors.laplacian(Us[j], Ubs[j]) for j in range(N)], axis=1)
    B = torch.from_numpy(np.stack(Us, axis=1).T @ (Vw[:, None] * L))
    dist.all_reduce(B)                                   # ONE all-reduce of the whole operator
    owned = torch.zeros(mesh.nCells, dtype=torch.int64)
    owned[torch.from_numpy(cpa[:part.nOwned].astype(np.int64))] = 1
    dist.all_reduce(owned)
    np.savez(os.path.join(tmp, "proj%d.npz" % rank), B=B.numpy(), owned=owned.numpy())
    dist.destroy_process_group()


def test_two_rank_sharded_projection_equals_serial(tmp_path):
    sys.path.insert(0, ROOT)
    from oracle import fvc, polymesh as pm
    world, port = 2, 30100 + os.getpid() % 400
    mp.spawn(_proj_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    mesh = pm.synthetic_hex(8, 12, 4, kind="cylinder", seed=4, perturb=0.2)
    U, Ub, _, _ = fvc.synthetic_basis(mesh, 4, 0, seed=3)
    ref = fvc.Oracle(mesh, corrected=True).diffusive_term(U, Ub)
    parts = [np.load(tmp_path / ("proj%d.npz" % k)) for k in range(world)]
    assert np.array_equal(parts[0]["B"], parts[1]["B"])
    assert np.all(parts[0]["owned"] == 1)                # every cell owned exactly once across the ranks
    assert np.abs(parts[0]["B"] - ref).max() <= 1e-12 * np.abs(ref).max()
