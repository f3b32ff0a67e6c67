#!/usr/bin/env python
"""One compact block per kernel launch of an .ncu-rep: python scripts/ncu_brief.py file.ncu-rep [name-substring]"""
import csv, subprocess, sys
rep = sys.argv[1]
sub = sys.argv[2] if len(sys.argv) > 2 else ""
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[0]
col = {h: i for i, h in enumerate(hdr)}
def g(r, k):
    try:
        return float(r[col[k]].replace(",", ""))
    except Exception:
        return float("nan")
for r in rows[2:]:
    name = r[col["Kernel Name"]]
    if sub not in name:
        continue
    st = {k.replace("smsp__pcsamp_warps_issue_stalled_", ""): g(r, k) for k in hdr if k.startswith("smsp__pcsamp_warps_issue_stalled_") and "not_issued" not in k}
    tot = sum(v for v in st.values() if v == v) or 1
    top = sorted(st.items(), key=lambda kv: -kv[1])[:5]
    print("%s\n  %.3f ms grid %d regs %d | tensor %.1f%% fp64 %.1f%% issue %.1f%% warps %.1f%% | dram rd %.3f GB wr %.3f GB (%.1f%% peak) L2hit %.0f%% L1hit %.0f%% | smem wavefronts %.3g conflicts %.3g | active/elapsed cyc %.3g/%.3g\n  stalls: %s" % (
        name[:100], g(r, "gpu__time_duration.sum") * (1e-6 if rows[1][col["gpu__time_duration.sum"]] == "ns" else 1e-3 if rows[1][col["gpu__time_duration.sum"]] in ("us", "usecond") else 1),
        g(r, "launch__grid_size"), g(r, "launch__registers_per_thread"),
        g(r, "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"), g(r, "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"),
        g(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"), g(r, "sm__warps_active.avg.pct_of_peak_sustained_active"),
        g(r, "dram__bytes_read.sum") * {"Gbyte": 1, "Mbyte": 1e-3, "Kbyte": 1e-6, "byte": 1e-9}.get(rows[1][col["dram__bytes_read.sum"]], 1),
        g(r, "dram__bytes_write.sum") * {"Gbyte": 1, "Mbyte": 1e-3, "Kbyte": 1e-6, "byte": 1e-9}.get(rows[1][col["dram__bytes_write.sum"]], 1),
        g(r, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"), g(r, "lts__t_sector_hit_rate.pct"), g(r, "l1tex__t_sector_hit_rate.pct"),
        g(r, "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"), g(r, "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"),
        g(r, "sm__cycles_active.avg"), g(r, "sm__cycles_elapsed.max"),
        ", ".join("%s %.0f%%" % (k, 100 * v / tot) for k, v in top)))
