"""Order of the consumer release in the SASS of the TMA ring kernels (the delayed release of common.cuh): in every consumer loop
body (the instructions from one SYNCS...TRYWAIT to the next one, or to the loop's backward branch) that holds DMMAs, the
consumer's SYNCS.ARRIVE.TRANS64.A1T0 -- the release of the PREVIOUS stage -- must come before the body's first LDS and first
DMMA, i.e. directly behind the wait, and no second release may follow the DMMAs. Usage: sass_ring_check.py file.o kernel_substring
Prints one line per such loop body; exit code 1 on any violation (a release scheduled among or behind the body's own loads is the
write-after-read race described in common.cuh)."""
import re
import shutil
import subprocess
import sys


def kernels(obj):
    cu = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    out = subprocess.run([cu, "-sass", obj], capture_output=True, text=True, check=True).stdout
    ks, name = {}, None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = m.group(1)
            ks[name] = []
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if name and m:
            ks[name].append((int(m.group(1), 16), m.group(2).strip()))
    return ks


def bodies(ops):
    """Split at every full-barrier TRYWAIT (a loop head of a consumer or producer loop)."""
    idx = [i for i, (_, o) in enumerate(ops) if "TRYWAIT" in o]
    for a, b in zip(idx, idx[1:] + [len(ops)]):
        yield ops[a:b]


def check(obj, sub):
    bad, seen = [], 0
    for name, ops in kernels(obj).items():
        if sub not in name:
            continue
        for body in bodies(ops):
            # cut the body at its first backward branch: what follows belongs to other code
            head = body[0][0]
            end = len(body)
            for i, (addr, o) in enumerate(body):
                m = re.search(r"BRA(?:\.\w+)* (?:!?U?P\d+, )?0x([0-9a-f]+)", o)
                if m and int(m.group(1), 16) <= head and i > 0:
                    end = i + 1
                    break
            body = body[:end]
            dm = [i for i, (_, o) in enumerate(body) if o.split()[0].startswith("DMMA") or " DMMA" in o]
            ar = [i for i, (_, o) in enumerate(body) if "SYNCS.ARRIVE.TRANS64.A1T0" in o]
            ld = [i for i, (_, o) in enumerate(body) if re.match(r"(@!?U?P\d+ )?LDS", o)]
            if not dm:
                continue
            seen += 1
            ok = len(ar) == 1 and ld and ar[0] < ld[0] and ar[0] < dm[0]
            print(f"{name[:60]} body@{head:#x}: {len(dm)} DMMA (first +{dm[0]}), first LDS +{ld[0] if ld else None}, "
                  f"release at {['+%d' % a for a in ar]} {'ok' if ok else 'NOT DIRECTLY BEHIND THE WAIT'}")
            if not ok:
                bad.append((name, head))
    return seen, bad


if __name__ == "__main__":
    seen, bad = check(sys.argv[1], sys.argv[2])
    print("consumer loop bodies with DMMAs:", seen, "violations:", len(bad))
    sys.exit(1 if bad or not seen else 0)
