#!/usr/bin/env python
"""SASS evidence of libetl_b200.so without a GPU: per kernel, registers / shared memory / stack (cuobjdump
-res-usage) and counts of the mnemonics that matter here (cuobjdump -sass): UBLKCP = cp.async.bulk (the TMA engine's
1-D bulk copy, loads and stores), SYNCS = mbarrier operations, ATOMS.CAST / ATOMS.CAS = compare-and-swap shared atomics
(the fp64 add loop), ATOMS.POPC / ATOMS.ADD = native 32-bit shared atomics, REDG / ATOMG = global reductions / atomics,
VOTE / MATCH / REDUX / SHFL = warp collectives, LDL / STL = local memory (spills).

    python tools/sass_summary.py > profiles/r02_sass_mnemonics.txt      (round 1: profiles/r01_sass_mnemonics.txt, counts only)
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "pyspark-tf-gke_b200", "libetl_b200.so")
KEEP = ("UBLKCP", "SYNCS", "ATOMS", "ATOMG", "REDG", "RED", "VOTE", "MATCH", "REDUX", "SHFL", "BAR", "LDG", "STG", "LDS", "STS",
        "LDL", "STL", "DADD", "DMUL", "DFMA", "UTMALDG", "UTCMMA", "HMMA")


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def main():
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout
    usage = {}
    cur = None
    for line in res.splitlines():
        m = re.match(r"\s*Function (\S+):", line)
        if m:
            cur = m.group(1)
        elif cur and "REG:" in line:
            usage[cur] = dict(re.findall(r"(\w+(?:\[\d+\])?):(\d+)", line))
            cur = None
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    counts = collections.defaultdict(collections.Counter)
    cur = None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]+)", line)
        if cur and m:
            op = m.group(1)
            base = op.split(".")[0]
            if base in KEEP:
                key = base
                if base == "ATOMS":
                    key = ".".join(op.split(".")[:2])
                counts[cur][key] += 1
    names = demangle(sorted(set(usage) | set(counts)))
    print(__doc__.split("\n\n")[0])
    print()
    for mangled in sorted(names, key=lambda k: names[k]):
        u = usage.get(mangled, {})
        name = re.sub(r"\(.*", "", names[mangled])
        print(name)
        print("    " + "  ".join(f"{k}={u[k]}" for k in ("REG", "SHARED", "STACK", "LOCAL") if k in u))
        c = counts.get(mangled)
        if c:
            print("    " + "  ".join(f"{k}={v}" for k, v in sorted(c.items())))


if __name__ == "__main__":
    sys.exit(main())
