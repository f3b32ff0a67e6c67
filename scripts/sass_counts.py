#!/usr/bin/env python
"""Per-kernel SASS instruction counts of libithaca_b200.so (cuobjdump -sass, sm_100a): the mnemonics that prove the FP64 tensor
path (DMMA), the TMA path (UTMALDG = cp.async.bulk.tensor, UBLKCP = cp.async.bulk), mbarriers (SYNCS) and warp shuffles.
  python scripts/sass_counts.py > profiles/r2_sass_counts.md"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "ithaca-fv_b200", "lib", "libithaca_b200.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
MN = ["DMMA", "UTMALDG", "UBLKCP", "SYNCS", "DFMA", "SHFL", "LDS", "STS", "LDG", "STG", "BAR"]
cur, counts, total = None, collections.OrderedDict(), {}
for ln in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", ln)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(.*", "", cur).replace("ib200::", "").replace("void ", "")
        counts[cur] = collections.Counter()
        total[cur] = 0
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", ln)
    if m and cur:
        total[cur] += 1
        op = m.group(1)
        for k in MN:
            if op == k or op.startswith(k + "."):
                counts[cur][k] += 1
print("# SASS instruction counts per kernel (cuobjdump -sass ithaca-fv_b200/lib/libithaca_b200.so, sm_100a)\n")
print("DMMA = FP64 tensor-core MMA (mma.sync.m8n8k4.f64); UTMALDG = TMA tensor load (cp.async.bulk.tensor); UBLKCP = bulk copy "
      "(cp.async.bulk); SYNCS = mbarrier operations. Static counts (unrolled loop bodies), not executed counts.\n")
print("| kernel | instr | " + " | ".join(MN) + " |\n|---|---|" + "---|" * len(MN))
for k, c in counts.items():
    if any(c[x] for x in ("DMMA", "UTMALDG", "UBLKCP")) or total[k] > 400:
        print("| `%s` | %d | " % (k, total[k]) + " | ".join(str(c[x]) for x in MN) + " |")
