#!/usr/bin/env python
"""bench.py -- POD time-to-modes (ITHACAPOD::getModes hot path) on N B200s, one rank per GPU.

Headline workload (BASELINE.json's metric): C3 = 20M cells x 1000 FP64 vector snapshots, 50 modes, the cells sharded over
the N ranks (STRONG scaling: the same 480 GB problem at every N). A "step" is one pass of the hot path over the rank's shard:
Gram (DMMA) -> one NCCL all-reduce of the m x m matrix -> Jacobi eigensolve -> lift + normalise.
  * shard <= ~125 GB (N = 4, 8): the shard is RESIDENT in HBM before the timed region (one chunk);
  * shard  > HBM     (N = 1, 2): ROW-STREAMED through the ithaca_b200_pod_stream_* calls: chunks of consecutive cells pass
    through two device slots, pass 1 sums the chunk Grams, pass 2 lifts; every chunk is produced in place by the device-side
    counter-based generator (ithaca_b200_synth_*, outside the timed segments) because the set cannot exist anywhere at once.
`value`  = device time of the library's calls per step (CUDA events on the library's stream), max over ranks.
`e2e`    = the same through the C ABI with HOST buffers: every chunk is copied host -> device from pinned memory inside the
           timed region (twice per step when streamed, minus the chunks still resident for pass 2), the modes are copied back.
Sub-objects: `c2` (2M cells x 200 scalar snapshots, 20 modes PER RANK: the largest single-GPU configuration, with its own
device-resident value, e2e, roofline) and `projection` (C4, bench_proj.py).
Other workloads: --workload c5 (50M cells x 2000 scalar snapshots over 8 ranks = 6.25M cells per rank, weak), c2, c1.
`--impl reference` times the reference's own CPU expressions (oracle/_ref: its vendored Eigen/Spectra) with one cell shard
per host thread (= one MPI rank per core, the reference's only parallelism) on a bounded sample of the same workload.
"""
import argparse
import importlib
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# stdout carries ONE JSON line: NCCL's debug output (the version banner at NCCL_DEBUG >= VERSION) goes to stderr
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")

import numpy as np  # noqa: E402

WORKLOADS = {
    "c1": dict(cells=10_000, dim=3, m=100, r=10, scaling="weak", per_rank=True,
               name="C1 10k cells x 100 vector snapshots, 10 modes (per rank)"),
    "c2": dict(cells=2_000_000, dim=1, m=200, r=20, scaling="weak", per_rank=True,
               name="C2 2M cells x 200 scalar FP64 snapshots, 20 modes (per rank)"),
    "c3": dict(cells=20_000_000, dim=3, m=1000, r=50, scaling="strong", per_rank=False,
               name="C3 20M cells x 1000 vector FP64 snapshots, 50 modes (cell-sharded over the ranks)"),
    "c5": dict(cells=6_250_000, dim=1, m=2000, r=100, scaling="weak", per_rank=True,
               name="C5 50M cells x 2000 scalar FP64 snapshots over 8 ranks = 6.25M cells (100 GB) per rank, 100 modes"),
}
METRIC = "pod_time_to_modes"
UNIT = "s"
RESIDENT_MAX_BYTES = 125e9     # a shard up to this size stays in HBM as one chunk
CHUNK_BYTES = 60e9             # chunk size of the row-streamed path (two slots + the resident modes must fit 180 GB)
E2E_CHUNK_BYTES = 30e9         # chunk size of the end-to-end leg (host pool, upload / compute overlap)
SYNTH_R, SYNTH_EPS = 64, 1e-6  # rank of the smooth part and noise level of the synthetic snapshots
SEED = 20251104


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-steps", type=int, default=3, help="end-to-end steps of the headline workload (each moves the whole shard over PCIe)")
    ap.add_argument("--no-c2", action="store_true", help="skip the C2 sub-object")
    ap.add_argument("--no-projection", action="store_true", help="skip the C4 Galerkin-projection leg")
    ap.add_argument("--no-checks", action="store_true", help="skip the small parity check against the oracle")
    ap.add_argument("--projection", default="c4", choices=["c4", "c4small"])
    ap.add_argument("--cells", type=int, default=0, help="override the workload's cell count (debugging)")
    return ap.parse_args()


def workload_config(wl, world):
    """The `config` object of the JSON line: identical in both arms."""
    cells_total = wl["cells"] * world if wl["per_rank"] else wl["cells"]
    per_rank_gb = cells_total / world * wl["dim"] * wl["m"] * 8 / 1e9
    return {"workload": wl["name"], "cells_total": cells_total, "dim": wl["dim"], "snapshots": wl["m"], "modes": wl["r"],
            "ranks": world, "l2": "inputs larger than L2 (%.1f GB per rank), no flush needed" % per_rank_gb}


# ------------------------------------------------------------------------------------------------
# clocks sampler (pynvml; falls back to nothing if NVML is unavailable)
# ------------------------------------------------------------------------------------------------
class Clocks:
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
               0x80: "hw_power_brake", 0x2: "applications_clocks_setting"}

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.dev = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.dev, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        while not self._stop.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.dev, self.nv.NVML_CLOCK_SM))
                bits = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.dev)
                for b, n in self.REASONS.items():
                    if bits & b:
                        self.reasons.add(n)
            except Exception:
                pass
            self._stop.wait(0.05)

    def start(self):
        if self.nv:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()

    def stop(self):
        if self._thr:
            self._stop.set()
            self._thr.join()
        return {"sm_mhz": (statistics.median(self.samples) if self.samples else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons)}


# ------------------------------------------------------------------------------------------------
# CPU arm: the reference's expressions on its vendored Eigen/Spectra (oracle/_ref), else the numpy port
# ------------------------------------------------------------------------------------------------
def host_threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


def cpu_time_to_modes(wl, cells_full, threads, rows_per_thread, seed=5):
    """Times Gram + eigensolve + lift of the reference on `threads` cell shards of `rows_per_thread` cells
    each (one shard per host thread = one MPI rank per core), then scales the two stages that are linear in
    the cell count to the full workload (the eigensolve is run at full size m x m).
    Returns (seconds_full, detail)."""
    from oracle import pod_oracle as po
    dim, m, r = wl["dim"], wl["m"], wl["r"]
    kind = "reference" if po.ref_available() else "port"
    rng = np.random.default_rng(seed)
    shards = []
    for t in range(threads):
        S = np.asfortranarray(rng.standard_normal((rows_per_thread * dim, m)))
        V = rng.random(rows_per_thread) + 0.5
        shards.append((S, V))
    # the reference's Gram branch is chosen by the FULL per-rank size (ITHACAcoeffsMass.C:95): V.size() > 1e6
    per_rank_Nd = cells_full * dim / threads
    branch = 1 if per_rank_Nd > 1e6 else 0
    Gs = [None] * threads

    def run_gram(t):
        S, V = shards[t]
        Gs[t] = po.ref_gram(S, V, dim, branch) if kind == "reference" else po.gram(S, V, dim)

    t0 = time.perf_counter()
    th = [threading.Thread(target=run_gram, args=(t,)) for t in range(threads)]
    [x.start() for x in th]
    [x.join() for x in th]
    G = sum(Gs)  # = Pstream reduce
    t_gram = time.perf_counter() - t0
    t0 = time.perf_counter()
    ev, Q = (po.ref_eig(G, r, "spectra") if kind == "reference" else po.eig(G, r, "spectra"))
    t_eig = time.perf_counter() - t0

    def run_lift(t):
        S, V = shards[t]
        if kind == "reference":
            po.ref_lift(S, V, dim, Q, "L2")
        else:
            po.lift(S, V, dim, Q, "L2")

    t0 = time.perf_counter()
    th = [threading.Thread(target=run_lift, args=(t,)) for t in range(threads)]
    [x.start() for x in th]
    [x.join() for x in th]
    t_lift = time.perf_counter() - t0
    scale = cells_full / float(threads * rows_per_thread)
    total = (t_gram + t_lift) * scale + t_eig
    detail = dict(kind=kind, cores=threads, gram_s=t_gram * scale, eig_s=t_eig, lift_s=t_lift * scale,
                  gram_branch=("rank1-loop" if branch else "gemm"), scale=scale,
                  sample="%d threads x %d cells x %d snapshots (dim %d), Gram+lift scaled x%.1f to %d cells, "
                         "eigensolve %dx%d in full" % (threads, rows_per_thread, m, dim, scale, cells_full, m, m))
    return total, detail


def run_reference_arm(args, wl, rank, world):
    if rank != 0:
        return
    threads = host_threads()
    cfg = workload_config(wl, world)
    cells_total = cfg["cells_total"]
    rows = max(64, min(cells_total // threads, int(2.5e9 / (2.0 * wl["m"] * wl["m"] * wl["dim"]))))
    times, detail = [], None
    for it in range(args.warmup + args.steps):
        t, detail = cpu_time_to_modes(wl, cells_total, threads, rows, seed=it)
        if it >= args.warmup:
            times.append(t)
    v = statistics.mean(times)
    # for information: the GEMM branch of the same expression (taken by the reference only when a rank holds <= 1e6 values)
    gemm_s = None
    try:
        from oracle import pod_oracle as po
        if detail["gram_branch"] != "gemm" and po.ref_available():
            rng = np.random.default_rng(99)
            S = np.asfortranarray(rng.standard_normal((rows * wl["dim"], wl["m"])))
            V = rng.random(rows) + 0.5
            t0 = time.perf_counter()
            po.ref_gram(S, V, wl["dim"], 0)
            gemm_s = (time.perf_counter() - t0) * cells_total / float(threads * rows)
    except Exception:
        pass
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": v * 1e3, "higher_is_better": False, "scaling": wl["scaling"],
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": cfg,
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": detail["cores"], "kind": detail["kind"],
                             "sample": detail["sample"], "stages_s": {k: detail[k] for k in ("gram_s", "eig_s", "lift_s")},
                             "gram_branch": detail["gram_branch"], "extrapolated": True,
                             "gram_s_if_gemm_branch": gemm_s},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------------------
class _RawDevice:
    """A raw device allocation as a torch-importable array (plumbing: lets torch copy a slot to pinned host memory)."""

    def __init__(self, ptr, shape):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<f8", "data": (int(ptr), False), "version": 3,
                                         "strides": None}


def mem_available_bytes():
    try:
        for ln in open("/proc/meminfo"):
            if ln.startswith("MemAvailable:"):
                return int(ln.split()[1]) * 1024
    except Exception:
        pass
    return 64 << 30


HOST_MEM_AT_START = mem_available_bytes()   # before any rank pins memory: the same figure on every rank of the box


class Env:
    """One rank's plumbing: torch.distributed (NCCL) for the harness, library handles with their own communicator."""

    def __init__(self, rank, local_rank, world):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.rank, self.local_rank, self.world = rank, local_rank, world
        self.multi = world > 1
        self.pkg = importlib.import_module("ithaca-fv_b200")
        torch.cuda.set_device(local_rank)
        self.dev = torch.device("cuda", local_rank)
        if self.multi:
            dist.init_process_group("nccl", device_id=self.dev)

    def barrier(self):
        if self.multi:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, vals):
        t = self.torch.tensor(list(vals), device=self.dev, dtype=self.torch.float64)
        if self.multi:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return t.tolist()

    def new_pod(self):
        """A library handle with its own NCCL communicator (the unique id travels over torch.distributed)."""
        torch = self.torch
        pod = self.pkg.Pod(self.local_rank)
        if self.multi:
            import ctypes
            idt = torch.zeros(128, dtype=torch.uint8, device=self.dev)
            if self.rank == 0:
                buf = (ctypes.c_char * 128)()
                rc = self.pkg.lib().ithaca_b200_nccl_unique_id(buf)
                assert rc == 0, "ncclGetUniqueId failed"
                idt.copy_(torch.frombuffer(bytearray(bytes(buf)), dtype=torch.uint8))
            self.dist.broadcast(idt, 0)
            pod.comm_init(bytes(idt.cpu().numpy().tobytes()), self.rank, self.world)
        return pod

    def close(self):
        if self.multi:
            self.dist.destroy_process_group()


# ---- small parity check against the oracle, through the real communicator (the NCCL path was never compared) ----
def parity_check(env, pod):
    """A small cell-sharded POD (resident and row-streamed) and a cell-sharded projection through the library's own
    all-reduces, compared on rank 0 with the oracle (oracle/_ref when built): G 1e-13, eigenvalues 1e-10, sign-aligned
    modes 1e-8, projected operators 1e-10. Replaces reduce(M, sumOp) (ITHACAcoeffsMass.C:112-115, ITHACAPOD.C:146-149), the
    norm reduce (ITHACAPOD.C:235-243) and the per-entry gSum of fvc::domainIntegrate (steadyNS.C:896-897 et al.)."""
    from oracle import fvc, polymesh as opm, pod_oracle as po, synth
    torch, dist, pkg, rank, world, dev = env.torch, env.dist, env.pkg, env.rank, env.world, env.dev
    Nc, dim, m, r = 40_009, 3, 64, 10
    S = synth.snapshots(Nc, dim, m, 2 * r, 4242)
    V = synth.volumes(Nc, 4242)
    b, e = pkg.cell_range(Nc, rank, world)
    Sl = np.asfortranarray(S[b * dim:e * dim])
    Vl = np.ascontiguousarray(V[b:e])
    rows_max = max((pkg.cell_range(Nc, q, world)[1] - pkg.cell_range(Nc, q, world)[0]) * dim for q in range(world))

    def gather_modes(Phi):
        buf = torch.zeros((r, rows_max), device=dev, dtype=torch.float64)
        buf[:, :Phi.shape[0]] = torch.from_numpy(np.ascontiguousarray(Phi.T)).to(dev)
        if not env.multi:
            return Phi
        parts = [torch.zeros_like(buf) for _ in range(world)]
        dist.all_gather(parts, buf)
        cols = []
        for q in range(world):
            qb, qe = pkg.cell_range(Nc, q, world)
            cols.append(parts[q][:, :(qe - qb) * dim].cpu().numpy().T)
        return np.concatenate(cols, axis=0)

    out = pkg.getModes(pod, Sl, Vl, dim, r, "spectra", "L2")
    modes = gather_modes(out["modes"])
    outs = pkg.getModesStreamed(pod, Sl, Vl, dim, r, "spectra", "L2", chunk_cells=max(1, (e - b) // 3 + 1))
    modes_s = gather_modes(outs["modes"])

    mesh = opm.load_compact(os.path.join(ROOT, "tests", "golden", "mesh_04unsteadyNS.npz"))
    U, Ub, P, Pb = fvc.synthetic_basis(mesh, 6, 3, seed=13)
    pr = pkg.Projector(pod)
    part = None
    if env.multi:
        part = pkg.Partition(mesh, rank, world, 2, method="locality")
        pr.set_mesh_shard(mesh, part, corrected=True)
        pr.set_basis("U", [part.gather(u, "cell", 3) for u in U], [part.gather(u, "bnd", 3) for u in Ub])
        pr.set_basis("P", [part.gather(p, "cell", 1) for p in P], [part.gather(p, "bnd", 1) for p in Pb])
    else:
        pr.set_mesh(mesh, corrected=True)
        pr.set_basis("U", U, Ub)
        pr.set_basis("P", P, Pb)
    ops = {k: pr.project(k) for k in ("M", "B", "K", "P", "C")}   # all-reduced inside ithaca_b200_project
    if part is not None:
        part.close()
    res = None
    if rank == 0:
        ref = po.ref_getmodes(S, V, dim, r, "spectra", "L2") if po.ref_available() else po.getmodes(S, V, dim, r, "spectra", "L2")
        rel = lambda A, B: float(np.abs(A - B).max() / np.abs(B).max())

        def mrel(A, B):
            A = po.sign_align(A, B)
            return float((np.abs(A - B).max(axis=0) / np.abs(B).max(axis=0)).max())

        orc = fvc.Oracle(mesh, corrected=True)
        oref = {"M": orc.mass_term(U), "B": orc.diffusive_term(U, Ub), "K": orc.pressure_gradient_term(U, P, Pb),
                "P": orc.divergence_term(U, Ub, P), "C": orc.convective_term_tens(U, Ub, cache=True)}
        res = {"oracle": "reference (oracle/_ref: the reference's Eigen/Spectra expressions)" if po.ref_available() else "port (numpy restatement)",
               "case": "POD %d cells x %d vector snapshots, %d modes, cells sharded by ithaca_b200_cell_range over %d ranks; "
                       "projection M, B, K, P, C on the tutorial-04 mesh, %d U / %d p fields, %s"
                       % (Nc, m, r, world, len(U), len(P), "halo-sharded, one ncclAllReduce per operator" if env.multi else "one rank"),
               "gram_rel": rel(out["G"], ref["G"]),
               "evals_rel": float((np.abs(out["evals_norm"][:r] - ref["evals_norm"][:r]) / ref["evals_norm"][:r]).max()),
               "normfact_rel": float((np.abs(out["normfact"] - ref["normfact"]) / ref["normfact"]).max()),
               "modes_rel": mrel(modes, ref["modes"]),
               "streamed_gram_rel": rel(outs["G"], ref["G"]), "streamed_modes_rel": mrel(modes_s, ref["modes"])}
        for k in oref:
            res[k + "_rel"] = rel(ops[k], oref[k])
        res["pass"] = bool(res["gram_rel"] <= 1e-13 and res["evals_rel"] <= 1e-10 and res["modes_rel"] <= 1e-8 and
                           res["normfact_rel"] <= 1e-9 and res["streamed_gram_rel"] <= 1e-13 and res["streamed_modes_rel"] <= 1e-8 and
                           all(res[k + "_rel"] <= 1e-10 for k in oref))
    return res


# ---- headline: cell-sharded POD through the row-streamed calls (one resident chunk when the shard fits) ----
class ShardedPod:
    def __init__(self, env, pod, wl, cells_override=0):
        torch = env.torch
        self.env, self.pod, self.wl = env, pod, wl
        self.dim, self.m, self.r = wl["dim"], wl["m"], wl["r"]
        cells = cells_override or wl["cells"]
        if wl["per_rank"]:
            self.cells_global = cells * env.world
            self.cell0, self.cells = cells * env.rank, cells
        else:
            self.cells_global = cells
            b, e = env.pkg.cell_range(cells, env.rank, env.world)
            self.cell0, self.cells = b, e - b
        self.Nd = self.cells * self.dim
        self.Nd_global = self.cells_global * self.dim
        self.shard_bytes = float(self.Nd) * self.m * 8
        n = 1 if self.shard_bytes <= RESIDENT_MAX_BYTES else int(-(-self.shard_bytes // CHUNK_BYTES))
        cc = -(-self.cells // n)
        self.chunks = [(c0, min(cc, self.cells - c0)) for c0 in range(0, self.cells, cc)]
        self.resident = len(self.chunks) == 1
        g = torch.Generator()
        g.manual_seed(SEED)                              # the same coefficient matrix on every rank
        coef, _ = torch.linalg.qr(torch.randn(self.m, SYNTH_R, generator=g, dtype=torch.float64))
        self.dcoef = coef.t().contiguous().to(env.dev)   # (R, m) row-major = m x R column-major
        self.stream = torch.cuda.ExternalStream(pod.stream, device=env.dev)
        pod.stream_begin(self.m, self.dim, self.cells)
        self.in_slot = {}                                # chunk index -> slot currently holding it (device-generated)
        self.ptrs = {}                                   # slot -> (S address, ld, V address) of its current occupant
        if self.resident:
            self.fill(0, 0)
            pod.timings()                                # synchronises the library's stream

    def fill(self, slot, k):
        """Chunk k produced in place by the device-side generator (on the library's stream)."""
        c0, nc = self.chunks[k]
        Sp, ld, Vp = self.pod.stream_slot(slot, c0, nc)
        self.pod.synth_snapshots(Sp, ld, (self.cell0 + c0) * self.dim, nc * self.dim, self.Nd_global, self.m,
                                 self.dcoef.data_ptr(), SYNTH_R, SYNTH_EPS, SEED)
        self.pod.synth_volumes(Vp, self.cell0 + c0, nc, self.cells_global, SEED)
        self.in_slot = {q: s for q, s in self.in_slot.items() if s != slot}
        self.in_slot[k] = slot
        self.ptrs[slot] = (Sp, ld, Vp)

    def ensure(self, k):
        """Chunk k in slot k & 1 (generated unless it is still there)."""
        if self.in_slot.get(k) != (k & 1):
            self.fill(k & 1, k)
        return k & 1

    def _ev(self):
        return self.env.torch.cuda.Event(enable_timing=True)

    def lift_order(self):
        """Pass 2 visits the two chunks still resident from pass 1 first (no refill / re-upload), then the rest."""
        n = len(self.chunks)
        return [k for k in (n - 1, n - 2) if k >= 0] + list(range(n - 3, -1, -1))

    def step_device(self):
        """One step with the chunks produced on the device between the timed segments.
        Returns (library ms, {stage: ms}, gram kernel ms)."""
        pod, r = self.pod, self.r
        if self.resident:
            e0, e1 = self._ev(), self._ev()
            e0.record(self.stream)
            pod.stream_gram_add(0)
            pod.stream_gram_finish(want=False)
            tg = pod.timings()
            pod.eig(r)
            te = pod.timings()
            pod.stream_lift_add(0, r, "L2")
            pod.stream_lift_finish(r, want=False)
            e1.record(self.stream)
            e1.synchronize()
            tl = pod.timings()
            total = e0.elapsed_time(e1)
            st = {"gram": tg["gram"], "allreduce": tg["allreduce"], "eig": te["eig"], "scale": tl["scale"]}
            st["lift"] = total - sum(st.values())
            return total, st, tg["gram_kernel"]
        seg = []      # (stage, e0, e1)
        gk = 0.0

        def timed(stage, fn):
            e0, e1 = self._ev(), self._ev()
            e0.record(self.stream)
            fn()
            e1.record(self.stream)
            seg.append((stage, e0, e1))

        for k in range(len(self.chunks)):
            slot = self.ensure(k)
            timed("gram", lambda: pod.stream_gram_add(slot))
            gk += pod.timings()["gram_kernel"]
        timed("allreduce", lambda: pod.stream_gram_finish(want=False))
        timed("eig", lambda: pod.eig(r))
        for k in self.lift_order():
            slot = self.ensure(k)
            timed("lift", lambda: pod.stream_lift_add(slot, r, "L2"))
        timed("scale", lambda: pod.stream_lift_finish(r, want=False))
        self.env.torch.cuda.synchronize()
        st = {}
        for stage, a, b in seg:
            st[stage] = st.get(stage, 0.0) + a.elapsed_time(b)
        return sum(st.values()), st, gk

    def step_checked(self):
        """One UNTIMED step that brings the small results back: G (m x m), all eigenvalues, the r norm factors. Size-independent
        identities at the workload's full size: sum(lambda) = trace(G) (eigensolve), normfact_i^2 = lambda_i (because
        |S q_i|_V^2 = q_i^T G q_i: ties the Gram, its all-reduce, the eigensolve, the lift and the norm all-reduce together),
        G symmetric."""
        pod, r = self.pod, self.r
        for k in range(len(self.chunks)):
            pod.stream_gram_add(self.ensure(k))
        G = pod.stream_gram_finish(want=True)
        ev, _ = pod.eig(r)
        for k in self.lift_order():
            pod.stream_lift_add(self.ensure(k), r, "L2")
        _, nf = pod.stream_lift_finish(r, want=False)
        tr = float(np.trace(G))
        res = {"trace_vs_eigenvalue_sum_rel": abs(float(ev.sum()) - tr) / tr,
               "normfact_sq_vs_eigenvalue_rel": float(np.max(np.abs(nf * nf / ev[:r] - 1.0))),
               "gram_asymmetry_rel": float(np.abs(G - G.T).max() / np.abs(G).max()),
               "eigenvalues_descending": bool(np.all(np.diff(ev) <= 0)),
               "lambda_1": float(ev[0]), "lambda_r": float(ev[r - 1]), "lambda_min": float(ev[-1])}
        res["pass"] = bool(res["trace_vs_eigenvalue_sum_rel"] <= 1e-10 and res["normfact_sq_vs_eigenvalue_rel"] <= 1e-8 and
                           res["gram_asymmetry_rel"] <= 1e-13 and res["eigenvalues_descending"])
        return res

    # ---- end to end: host buffers in, host buffers out ----
    def host_setup(self):
        """Pinned host copies of the shard (or, when host RAM is smaller than the shard, of its first chunks: the pool is then
        read periodically) and of the modes. Filled from the device generator, outside the timed region.
        The end-to-end leg has its own chunking (<= E2E_CHUNK_BYTES per chunk, also when the device-resident leg holds the shard
        as ONE chunk): the upload of chunk k+1 then overlaps the contraction of chunk k, and the host pool stays small."""
        torch, env = self.env.torch, self.env
        local_world = int(os.environ.get("LOCAL_WORLD_SIZE", env.world))
        modes_bytes = self.r * self.Nd * 8
        budget = HOST_MEM_AT_START * 0.6 / local_world - modes_bytes
        # streamed shards keep the device leg's chunking (the two chunks that stay resident for pass 2 are as large as possible);
        # a resident shard is cut into <= E2E_CHUNK_BYTES pieces; either is halved until one chunk fits the host pool
        n = len(self.chunks) if not self.resident else int(-(-self.shard_bytes // E2E_CHUNK_BYTES))
        while self.shard_bytes / n > budget and n < 4096:
            n *= 2
        cc = -(-self.cells // n)
        self.e2e_chunks = [(c0, min(cc, self.cells - c0)) for c0 in range(0, self.cells, cc)]
        chunk_bytes = self.e2e_chunks[0][1] * self.dim * self.m * 8
        if budget < chunk_bytes:
            raise RuntimeError("host RAM: %.0f GB available per rank, one chunk needs %.0f GB" % (budget / 1e9, chunk_bytes / 1e9))
        npool = int(min(len(self.e2e_chunks), budget // chunk_bytes))
        self.pool_chunks = npool
        pool_cells = sum(nc for _, nc in self.e2e_chunks[:npool])
        self.S_host = torch.empty((self.m, pool_cells * self.dim), dtype=torch.float64, pin_memory=True)
        self.V_host = torch.empty(pool_cells, dtype=torch.float64, pin_memory=True)
        self.modes_host = torch.empty((self.r, self.Nd), dtype=torch.float64, pin_memory=True)
        self.pool_off = []
        off = 0
        self.in_slot = {}                                # slot 0 is reused as the generator's target
        for k in range(npool):
            c0, nc = self.e2e_chunks[k]
            Sp, ld, Vp = self.pod.stream_slot(0, c0, nc)
            self.pod.synth_snapshots(Sp, ld, (self.cell0 + c0) * self.dim, nc * self.dim, self.Nd_global, self.m,
                                     self.dcoef.data_ptr(), SYNTH_R, SYNTH_EPS, SEED)
            self.pod.synth_volumes(Vp, self.cell0 + c0, nc, self.cells_global, SEED)
            self.pod.timings()                           # the generator has finished
            St = torch.as_tensor(_RawDevice(Sp, (self.m, ld)), device=env.dev)
            Vt = torch.as_tensor(_RawDevice(Vp, (nc,)), device=env.dev)
            self.S_host[:, off * self.dim:(off + nc) * self.dim].copy_(St[:, :nc * self.dim])
            self.V_host[off:off + nc].copy_(Vt)
            torch.cuda.synchronize()
            self.pool_off.append(off)
            off += nc
        return {"pinned_snapshot_pool_GB": self.S_host.numel() * 8 / 1e9, "pool_chunks": npool, "chunks": len(self.e2e_chunks),
                "chunk_GB": chunk_bytes / 1e9, "periodic_pool": npool < len(self.e2e_chunks), "pinned_modes_GB": modes_bytes / 1e9,
                "host_mem_available_at_start_GB": HOST_MEM_AT_START / 1e9}

    def e2e_lift_order(self):
        n = len(self.e2e_chunks)
        return [k for k in (n - 1, n - 2) if k >= 0] + list(range(n - 3, -1, -1))

    def step_e2e(self):
        """Every chunk host -> device from the pinned pool (pipelined with the contraction), modes device -> host."""
        pod, r = self.pod, self.r
        base = [self.S_host[j].data_ptr() for j in range(self.m)]
        vbase = self.V_host.data_ptr()
        uploads = 0
        chunks = self.e2e_chunks

        def up(slot, k):
            nonlocal uploads
            c0, nc = chunks[k]
            off = self.pool_off[k % self.pool_chunks]    # a pool smaller than the shard is read periodically (same sizes: only the
            pod.stream_upload(slot, c0, nc, vbase + 8 * off, [b + 8 * off * self.dim for b in base])   # last chunk may be shorter)
            uploads += 1

        self.in_slot = {}                                # the slots are overwritten with host data
        n = len(chunks)
        up(0, 0)
        for k in range(n):
            if k + 1 < n:
                up((k + 1) & 1, k + 1)
            pod.stream_gram_add(k & 1)
        pod.stream_gram_finish(want=False)
        ev, Q = pod.eig(r)
        order = self.e2e_lift_order()
        slot_of = {k: k & 1 for k in order[:2]}
        pending = order[len(slot_of):]
        for k in order:
            pod.stream_lift_add(slot_of[k], r, "L2")
            if pending:
                kn = pending.pop(0)
                up(slot_of[k], kn)
                slot_of[kn] = slot_of[k]
        pod.stream_lift_finish(r, out=[self.modes_host[i] for i in range(r)])
        return uploads, ev


def run_headline(args, env, wl):
    torch = env.torch
    pod = env.new_pod()
    checks = None
    if not args.no_checks:
        checks = parity_check(env, pod)
        pod.close()
        pod = env.new_pod()                              # a fresh handle: nothing of the check stays allocated
    peak_dmma = env.pkg.dmma_peak(env.local_rank)
    sp = ShardedPod(env, pod, wl, args.cells)
    m, r, dim = sp.m, sp.r, sp.dim

    for _ in range(args.warmup):
        sp.step_device()
    clocks = Clocks(env.local_rank)
    env.barrier()
    pod.launch_count(reset=True)
    clocks.start()
    t_wall = time.perf_counter()
    tot, gk, stages = 0.0, 0.0, {}
    for _ in range(args.steps):
        t, st, g = sp.step_device()
        tot += t
        gk += g
        for k, v in st.items():
            stages[k] = stages.get(k, 0.0) + v
    env.barrier()
    wall = (time.perf_counter() - t_wall) / args.steps
    clk = clocks.stop()
    launches = pod.launch_count()
    ms_step, gram_kernel_ms, wall = env.max_over_ranks([tot / args.steps, gk / args.steps, wall])
    stages = {k: v / args.steps for k, v in stages.items()}
    full_size = None if args.no_checks else sp.step_checked()       # untimed; every rank takes part in the all-reduces

    e2e = None
    if not args.no_e2e:
        try:
            info = sp.host_setup()
            esteps = max(1, min(args.steps, args.e2e_steps))
            uploads, _ = sp.step_e2e()                   # warm-up: first touch of the pinned pages, NVLink/PCIe paths
            env.barrier()
            e0, e1 = sp._ev(), sp._ev()
            t0 = time.perf_counter()
            e0.record(sp.stream)
            for _ in range(esteps):
                uploads, _ = sp.step_e2e()
            e1.record(sp.stream)
            env.barrier()
            wall_e = (time.perf_counter() - t0) / esteps
            dev_e = e0.elapsed_time(e1) / 1e3 / esteps
            dev_e, wall_e = env.max_over_ranks([dev_e, wall_e])
            chunk_b = [nc * dim * m * 8 + nc * 8 for _, nc in sp.e2e_chunks]
            n = len(sp.e2e_chunks)
            h2d = sum(chunk_b) + sum(chunk_b[k] for k in sp.e2e_lift_order()[min(2, n):])
            e2e = {"value": wall_e, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int((r * sp.Nd + m + m * r + r) * 8),
                   "device_s": dev_e, "steps": esteps, "chunk_uploads_per_step": uploads,
                   "h2d_GBps_per_rank": h2d / wall_e / 1e9, "host": info,
                   "note": "host wall clock around the pipelined calls (upload of chunk k+1 overlaps the contraction of chunk k); "
                           "pass 2 lifts the two chunks still resident first and re-uploads only the others"}
        except RuntimeError as ex:
            e2e = {"value": None, "unit": UNIT, "error": str(ex)[:200]}
        sp.S_host = sp.V_host = sp.modes_host = None
    cfg = workload_config(wl, env.world)
    if args.cells:
        cfg["cells_override"] = args.cells
    Nd = sp.Nd
    gram_flops = float(Nd) * m * (m + 1)                 # lower triangle incl. diagonal, mul + add
    achieved = gram_flops / (gram_kernel_ms * 1e-3) / 1e12
    lift_bytes, lift_flops = 8.0 * Nd * (m + r), 2.0 * Nd * m * r
    hbm_peak = load_peaks().get("hbm_gbs", 6650.0)
    line = {
        "metric": METRIC, "value": ms_step / 1e3, "unit": UNIT, "n_gpus": env.world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_step, "higher_is_better": False, "scaling": wl["scaling"], "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": cfg,
        "path": {"kind": "resident (the rank's shard is one chunk in HBM before the timed region)" if sp.resident else
                         "row-streamed (ithaca_b200_pod_stream_*: chunks of consecutive cells through two device slots, pass 1 Gram, pass 2 lift)",
                 "cells_per_rank": sp.cells, "shard_GB": sp.shard_bytes / 1e9, "chunks": len(sp.chunks),
                 "chunk_cells": sp.chunks[0][1],
                 "value_definition": "sum of the device times of the library's calls in a step (CUDA events on the library's stream), max over ranks" +
                                     ("" if sp.resident else "; every chunk is regenerated in place by the device-side counter-based generator between the timed segments "
                                      "(the 480 GB set cannot exist at once)"),
                 "wall_s_per_step_including_generation": wall,
                 "eigensolver": "one-sided Jacobi (%d sweeps)" % pod.sweeps},
        "roofline": {"bound": "tensor", "kernel": "gram_kernel", "achieved": achieved, "peak": peak_dmma, "unit": "TFLOP/s",
                     "frac": achieved / peak_dmma, "frac_of_nominal_40": achieved / 40.0, "traffic": gram_traffic(args.workload),
                     "peak_source": "measured live: register-resident DMMA.8x8x4 microbenchmark (ithaca_b200_dmma_peak); "
                                    "MEASURED_PEAKS.json holds no FP64 figure; nominal B200 FP64 tensor = 40 TFLOP/s",
                     "algorithmic_flops_per_step": gram_flops, "launches_per_step": len(sp.chunks), "kernel_ms": gram_kernel_ms,
                     "algorithmic_flops_per_launch": gram_flops / len(sp.chunks), "kernel_ms_per_launch": gram_kernel_ms / len(sp.chunks),
                     "algorithmic_bytes_per_launch": (8.0 * Nd * m + 8.0 * sp.cells) / len(sp.chunks),
                     "traffic_note": "dram bytes (read + write) of ONE launch over a 60 GB chunk (2.5M cells x 1000 vector snapshots), ncu --set "
                                     "full, profiles/r2w_pod_full.md: 6.7x the algorithmic bytes -- at m = 1000 the 31 jobs of the lower triangle "
                                     "re-read their column panels and a 60 GB chunk does not fit the 126 MB L2; the kernel stays tensor-bound "
                                     "(DRAM 21 % of peak, tensor pipe 87.9 % active)"},
        "roofline_lift": {"bound": "tensor" if m >= 1000 else "hbm", "kernel": "lift_kernel", "ms": stages.get("lift"),
                          "GBps": lift_bytes / (stages["lift"] * 1e-3) / 1e9, "frac_hbm": lift_bytes / (stages["lift"] * 1e-3) / 1e9 / hbm_peak,
                          "tflops": lift_flops / (stages["lift"] * 1e-3) / 1e12,
                          "frac_dmma": lift_flops / (stages["lift"] * 1e-3) / 1e12 / peak_dmma,
                          "algorithmic_bytes_per_step": lift_bytes, "algorithmic_flops_per_step": lift_flops},
        "gram_tflops_total": achieved * env.world,
        "stages_ms": stages, "gpu_launches": int(launches), "clocks": clk, "e2e": e2e,
    }
    if not args.no_checks:
        line["checks"] = {"multi_gpu_parity" if env.multi else "parity_small": checks, "full_size_identities": full_size}
    pod.close()
    return line, peak_dmma, hbm_peak


def load_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


def gram_traffic(workload):
    tf = os.path.join(ROOT, "profiles", "gram_traffic.json")
    try:
        return json.load(open(tf)).get(workload)
    except Exception:
        return None


# ---- C2 (per rank): device-resident value, e2e with pinned host buffers, roofline of the Gram and lift kernels ----
def run_c2(args, env, pod, wl, peak_dmma, hbm_peak, steps, warmup):
    torch = env.torch
    dim, m, r, cells = wl["dim"], wl["m"], wl["r"], wl["cells"]
    Nd = cells * dim
    ld = (Nd + 15) // 16 * 16
    g = torch.Generator()
    g.manual_seed(SEED + 1)
    coef, _ = torch.linalg.qr(torch.randn(m, SYNTH_R, generator=g, dtype=torch.float64))
    dcoef = coef.t().contiguous().to(env.dev)
    S = torch.zeros((m, ld), device=env.dev, dtype=torch.float64)
    V = torch.zeros(cells, device=env.dev, dtype=torch.float64)
    cells_global = cells * env.world
    pod.synth_snapshots(S.data_ptr(), ld, env.rank * Nd, Nd, cells_global * dim, m, dcoef.data_ptr(), SYNTH_R, SYNTH_EPS, SEED + 1)
    pod.synth_volumes(V.data_ptr(), env.rank * cells, cells, cells_global, SEED + 1)
    pod.timings()
    stream = torch.cuda.ExternalStream(pod.stream, device=env.dev)
    snaps_dev = [S[j, :Nd] for j in range(m)]
    pod.set_volumes(V)
    pod.upload(snaps_dev, dim, cells)

    def step_resident():
        pod.gram(want=False)
        pod.eig(r)
        pod.lift(r, "L2", want=False)

    for _ in range(warmup):
        step_resident()
    env.barrier()
    pod.launch_count(reset=True)
    stage_acc = {}
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        step_resident()
        for k, v in pod.timings().items():
            stage_acc[k] = stage_acc.get(k, 0.0) + v
    e1.record(stream)
    env.barrier()
    launches = pod.launch_count()
    stages = {k: v / steps for k, v in stage_acc.items()}
    ms_step, gram_kernel_ms = env.max_over_ranks([e0.elapsed_time(e1) / steps, stages["gram_kernel"]])
    e2e = None
    if not args.no_e2e:
        try:
            S_host = torch.empty((m, Nd), dtype=torch.float64, pin_memory=True)
            S_host.copy_(S[:, :Nd])
            V_host = torch.empty(cells, dtype=torch.float64, pin_memory=True)
            V_host.copy_(V)
            modes_host = torch.empty((r, Nd), dtype=torch.float64, pin_memory=True)
            torch.cuda.synchronize()
            snaps_host = [S_host[j] for j in range(m)]
            modes_out = [modes_host[i] for i in range(r)]

            def step_e2e():
                pod.set_volumes(V_host)
                pod.upload(snaps_host, dim, cells)
                pod.gram(want=False)
                ev, Q = pod.eig(r)
                pod.lift(r, "L2", out=modes_out)
                return ev

            for _ in range(max(1, min(warmup, 3))):
                step_e2e()
            env.barrier()
            t0 = time.perf_counter()
            e0.record(stream)
            for _ in range(steps):
                step_e2e()
            e1.record(stream)
            env.barrier()
            wall = (time.perf_counter() - t0) / steps
            dev_s, wall = env.max_over_ranks([e0.elapsed_time(e1) / 1e3 / steps, wall])
            st = pod.timings()
            h2d = (m * Nd + cells) * 8
            e2e = {"value": dev_s, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": (r * Nd + m + m * r + r) * 8,
                   "wall_s": wall, "stages_ms": st, "h2d_GBps_rank0": h2d / (st["upload"] * 1e-3) / 1e9 if st["upload"] > 0 else None,
                   "d2h_GBps_rank0": r * Nd * 8 / (st["download"] * 1e-3) / 1e9 if st["download"] > 0 else None}
            up_min, = [(-x) for x in env.max_over_ranks([-(e2e["h2d_GBps_rank0"] or 0.0)])]
            e2e["h2d_GBps_slowest_rank"] = up_min
        except RuntimeError as ex:
            e2e = {"value": None, "unit": UNIT, "error": str(ex)[:120]}
    gram_flops = float(Nd) * m * (m + 1)
    achieved = gram_flops / (gram_kernel_ms * 1e-3) / 1e12
    lift_bytes = 8.0 * Nd * (m + r)
    out = {"config": workload_config(wl, env.world), "scaling": "weak", "value": ms_step / 1e3, "unit": UNIT, "ms_per_step": ms_step,
           "steps": steps, "warmup": warmup, "eigensolver": "one-sided Jacobi (%d sweeps)" % pod.sweeps,
           "roofline": {"bound": "tensor", "kernel": "gram_kernel", "achieved": achieved, "peak": peak_dmma, "unit": "TFLOP/s",
                        "frac": achieved / peak_dmma, "traffic": gram_traffic("c2"), "algorithmic_flops_per_launch": gram_flops,
                        "kernel_ms": gram_kernel_ms},
           "roofline_lift": {"bound": "hbm", "kernel": "lift_kernel", "achieved": lift_bytes / (stages["lift"] * 1e-3) / 1e9,
                             "peak": hbm_peak, "unit": "GB/s", "frac": lift_bytes / (stages["lift"] * 1e-3) / 1e9 / hbm_peak,
                             "algorithmic_bytes_per_launch": lift_bytes},
           "stages_ms": stages, "gpu_launches": int(launches), "e2e": e2e}
    del S, V, snaps_dev
    torch.cuda.empty_cache()
    return out


def run_b200(args, wl, rank, local_rank, world):
    env = Env(rank, local_rank, world)
    torch = env.torch
    line, peak_dmma, hbm_peak = run_headline(args, env, wl)
    torch.cuda.empty_cache()
    pod = env.new_pod()
    if not args.no_c2 and args.workload != "c2":
        try:
            line["c2"] = run_c2(args, env, pod, WORKLOADS["c2"], peak_dmma, hbm_peak, steps=max(args.steps, 10), warmup=max(args.warmup, 3))
        except Exception as ex:  # never lose the headline
            line["c2"] = {"error": str(ex)[:200]}
    if not args.no_projection:
        import bench_proj
        try:
            line["projection"], _, _, _ = bench_proj.run(pod, torch, env.dev, args.projection, hbm_peak=hbm_peak, dmma_peak=peak_dmma,
                                                         rank=rank, world=world, dist=env.dist if env.multi else None,
                                                         cpu_baseline=(not args.no_cpu_baseline and world == 1))
        except Exception as ex:
            line["projection"] = {"error": str(ex)[:200]}
        env.barrier()
    pod.close()
    ok = True
    if rank == 0:
        if not args.no_cpu_baseline and world == 1:
            cfg = line["config"]
            rows = max(64, min(cfg["cells_total"], int(2.0e10 / (2.0 * wl["m"] * wl["m"] * wl["dim"]))))  # ~10-20 s of single-core Gram
            v, d = cpu_time_to_modes(wl, cfg["cells_total"], 1, rows)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": d["cores"], "kind": d["kind"], "sample": d["sample"],
                                    "stages_s": {k: d[k] for k in ("gram_s", "eig_s", "lift_s")},
                                    "gram_branch": d["gram_branch"], "extrapolated": True, "host_threads_available": host_threads()}
        print(json.dumps(line), flush=True)
        chk = (line.get("checks") or {})
        res = chk.get("multi_gpu_parity") or chk.get("parity_small")
        ok = ((res is None) or bool(res.get("pass"))) and bool((chk.get("full_size_identities") or {"pass": True}).get("pass"))
    flag = env.max_over_ranks([0.0 if ok else 1.0])[0]
    env.close()
    if flag > 0:
        sys.stderr.write("bench.py: parity check against the oracle FAILED (see checks in the JSON line)\n")
        sys.exit(3)


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference_arm(args, wl, rank, world)
        return
    run_b200(args, wl, rank, local_rank, world)


if __name__ == "__main__":
    main()
