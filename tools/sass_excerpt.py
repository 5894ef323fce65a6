#!/usr/bin/env python
"""Per-kernel SASS evidence for the shipped library (the .so is git-ignored): counts of the instructions that prove which
hardware path a kernel takes — tcgen05 (UTCHMMA / UTCBAR / LDTM / STTM), TMA (UTMALDG / UTMASTG / UBLKCP), vector
reductions (REDG), shared-memory atomics (ATOMS), 128/256-bit global accesses, packed fp32 (FFMA2 ...), MUFU.

    python tools/sass_excerpt.py > profiles/r2_sass_excerpt.txt      (needs cuobjdump; no GPU)"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "learnable-triangulation-pytorch_b200", "liblt_b200.so")
PAT = [("UTCHMMA", r"\bUTC[A-Z]*MMA"), ("UTCBAR", r"\bUTCBAR"), ("LDTM", r"\bLDTM"), ("STTM", r"\bSTTM"),
       ("UTMALDG", r"\bUTMALDG"), ("UTMASTG", r"\bUTMASTG"), ("UBLKCP", r"\bUBLKCP"), ("REDG", r"\bREDG?\.E"),
       ("ATOMS", r"\bATOMS"), ("ATOMG", r"\bATOMG"), ("LDG.128", r"\bLDG\.E\.128"), ("LDG.256", r"\bLDG\.E\.(ENL2\.)?256"),
       ("STG.128", r"\bSTG\.E\.128"), ("STG.256", r"\bSTG\.E\.(ENL2\.)?256"), ("LDGSTS", r"\bLDGSTS"),
       ("FFMA2", r"\bF(FMA|MUL|ADD)2"), ("MUFU", r"\bMUFU"), ("BAR", r"\bBAR\.SYNC"), ("SHFL", r"\bSHFL"),
       ("REDUX", r"\bREDUX"), ("LDL/STL", r"\b(LDL|STL)\b")]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    name, counts, order = None, collections.OrderedDict(), []
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip() or m.group(1)
            name = name.replace("(anonymous namespace)::", "").replace("void ", "").replace("lt::", "")
            name = re.sub(r"\((const |float|UnprojArgs|lt|KArgs|int|unsigned|long|double|void).*", "", name)
            counts[name] = collections.Counter()
            continue
        if name is None or "/*" not in line:
            continue
        ins = line.split("*/", 1)[-1]
        counts[name]["total"] += 1 if re.search(r"^\s+[@A-Z]", ins) else 0
        for key, pat in PAT:
            if re.search(pat, ins):
                counts[name][key] += 1
    print("# SASS instruction counts per kernel of liblt_b200.so (sm_100a), `cuobjdump -sass`; zero columns omitted per row")
    for n, c in counts.items():
        cols = " ".join("%s=%d" % (k, c[k]) for k, _ in PAT if c[k])
        print("%-90s instr=%-5d %s" % (n[:90], c["total"], cols))


if __name__ == "__main__":
    sys.exit(main())
