#!/usr/bin/env python
"""SASS mnemonic counts per kernel of the built library -> profiles/r01_sass_mnemonics.txt
(cuobjdump -sass; runs without a GPU): python tools/sass_mnemonics.py > profiles/r01_sass_mnemonics.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "pyspark-tf-gke_b200", "libetl_b200.so")
KEEP = ("UBLKCP", "SYNCS", "ATOMS", "ATOMG", "REDG", "RED", "LDG", "STG", "LDS", "STS", "BAR", "VOTE", "DADD", "DMUL", "DFMA", "CCTL",
        "REDUX", "MATCH", "SHFL")


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
    kernels, cur = collections.OrderedDict(), None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = kernels.setdefault(re.sub(r"\(.*", "", demangle(m.group(1))), collections.Counter())
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and cur is not None:
            op = m.group(1)
            base = op.split(".")[0]
            if base in KEEP:
                key = ".".join(op.split(".")[:2]) if base == "ATOMS" else base
                cur[key] += 1
    print("SASS mnemonic counts per kernel of libetl_b200.so (cuobjdump -sass, sm_100a).")
    print("UBLKCP = cp.async.bulk (TMA engine, 1-D bulk copy); SYNCS = mbarrier ops; ATOMS.CAST = CAS-loop shared atomic "
          "(fp64/64-bit),\nATOMS.ADD/POPC = native 32-bit shared atomic; REDG = fire-and-forget global reduction; "
          "VOTE = warp ballot.\n")
    for name in sorted(kernels):
        print(name)
        print("    " + "  ".join(f"{k}={v}" for k, v in sorted(kernels[name].items())))


if __name__ == "__main__":
    main()
