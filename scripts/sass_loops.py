#!/usr/bin/env python
"""Summarise the loops of one kernel's SASS (cuobjdump -sass): instruction mix per backward branch.
usage: sass_loops.py file.o kernel_substring"""
import bisect, re, subprocess, sys
obj, name = sys.argv[1], sys.argv[2]
txt = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
blocks = re.split(r"\n\s*Function : ", txt)
for b in blocks:
    if name not in b.split("\n", 1)[0]:
        continue
    ins = [l for l in b.splitlines() if re.search(r"/\*[0-9a-f]{4,5}\*/\s+\S", l)]
    A = [int(re.search(r"/\*([0-9a-f]{4,5})\*/", l).group(1), 16) for l in ins]
    print(b.split("\n", 1)[0][:90], "instructions", len(ins), "DMMA", sum("DMMA" in l for l in ins),
          "LDL", sum("LDL" in l for l in ins), "STL", sum("STL" in l for l in ins))
    for i, l in enumerate(ins):
        m = re.search(r"BRA\S*\s+(?:!?U?P\d,\s*)?0x([0-9a-f]+)", l)
        if not m:
            continue
        tgt = int(m.group(1), 16)
        if tgt < A[i]:
            body = ins[bisect.bisect_left(A, tgt):i + 1]
            keys = ["DMMA", "LDS", "LDL", "STL", "DMUL", "WARPSYNC", "SYNCS", "LDG", "UTMALDG"]
            c = {k: sum(k in x for x in body) for k in keys}
            if c["DMMA"] or len(body) > 40:
                print("  loop %#x..%#x n=%d " % (tgt, A[i], len(body)) + " ".join("%s=%d" % kv for kv in c.items() if kv[1]))
