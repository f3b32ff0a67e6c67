#!/usr/bin/env python
"""Summarise ncu output brought back in gpurun_out/ into profiles/ (tracked).
  python scripts/ncu_summary.py launches gpurun_out/launches_r1.csv profiles/r1_launches.md
  python scripts/ncu_summary.py full gpurun_out/prof_r1_pod.ncu-rep profiles/r1_pod_full.md
"""
import collections
import csv
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor_subpipe_dmma.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__inst_executed.sum", "sm__cycles_elapsed.max"]


def launches(src, dst):
    lines = [l for l in open(src) if not l.startswith("==")]
    agg = collections.OrderedDict()
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        v = v / 1e3 if u == "ns" else (v * 1e3 if u == "ms" else v)
        a = agg.setdefault(row["Kernel Name"][:90], [0, 0.0])
        a[0] += 1
        a[1] += v
    foreign = ("at::", "cutlass", "geqr2", "larft", "lacpy", "orgqr", "batch_eye", "copy_info", "gemmcn", "elementwise_kernel", "nccl")
    ours = {k: v for k, v in agg.items()
            if "dmma_peak" not in k and ("ib200::" in k or not any(f in k for f in foreign))}
    tot = sum(v[1] for v in ours.values())
    with open(dst, "w") as f:
        f.write("# ncu launch list (gpu__time_duration.sum, --clock-control none; cold-cache, serialised)\n\n")
        f.write("source: `%s`; shares are over this library's kernels of the step (torch data-generation and the\n"
                "DMMA-peak microbenchmark listed below for completeness).\n\n" % src)
        f.write("| kernel | launches | total us | us/launch | share of step |\n|---|---|---|---|---|\n")
        for k, (n, t) in sorted(ours.items(), key=lambda x: -x[1][1]):
            f.write("| `%s` | %d | %.1f | %.1f | %.1f%% |\n" % (k, n, t, t / n, 100 * t / tot))
        f.write("\nother kernels in the capture:\n\n| kernel | launches | total us |\n|---|---|---|\n")
        for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
            if k not in ours:
                f.write("| `%s` | %d | %.1f |\n" % (k, n, t))
    print(open(dst).read())


def full(src, dst):
    raw = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    with open(dst, "w") as f:
        f.write("# ncu --set full summary\n\nsource: `%s` (per launch)\n" % src)
        for r in rows[2:]:
            name = r[hdr.index("Kernel Name")]
            f.write("\n## `%s`\n\n| metric | value | unit |\n|---|---|---|\n" % name[:100])
            for k in KEYS:
                if k in hdr:
                    i = hdr.index(k)
                    f.write("| %s | %s | %s |\n" % (k, r[i], units[i]))
    print(open(dst).read())


def full_many(dst, traffic_json, *srcs):
    """Several per-kernel reports -> one markdown; traffic_json gets {kernel: dram bytes (read + write) per launch}."""
    import json
    import re
    traffic = {}
    with open(dst, "w") as f:
        f.write("# ncu --set full summaries (one capture per kernel, --clock-control none, per launch)\n")
        for src in srcs:
            raw = open(src).read() if src.endswith(".csv") else subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
            rows = list(csv.reader([l for l in raw.splitlines() if not l.startswith("==")]))
            if len(rows) < 3:
                f.write("\n(no launches captured in `%s`)\n" % src)
                continue
            hdr, units = rows[0], rows[1]
            for r in rows[2:]:
                name = r[hdr.index("Kernel Name")]
                f.write("\n## `%s`  (%s)\n\n| metric | value | unit |\n|---|---|---|\n" % (name[:100], src.split("/")[-1]))
                vals = {}
                for k in KEYS:
                    if k in hdr:
                        i = hdr.index(k)
                        f.write("| %s | %s | %s |\n" % (k, r[i], units[i]))
                        vals[k] = (r[i], units[i])

                def nbytes(k):
                    if k not in vals:
                        return 0.0
                    v, u = vals[k]
                    v = float(v.replace(",", ""))
                    return v * {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(u, 1.0)

                short = re.sub(r"<.*", "", name.split("(")[0]).split("::")[-1]
                t = nbytes("dram__bytes_read.sum") + nbytes("dram__bytes_write.sum")
                traffic.setdefault(short, []).append(t)
    if traffic_json != "-":
        json.dump({k: {"dram_bytes_per_launch": v, "source": dst} for k, v in traffic.items()}, open(traffic_json, "w"), indent=1)
    print(open(dst).read())


if __name__ == "__main__":
    if sys.argv[1] == "full_many":
        full_many(sys.argv[2], sys.argv[3], *sys.argv[4:])
    else:
        {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3])
