#!/usr/bin/env python
"""Per-kernel SASS evidence of libnrb200.so (runs without a GPU): instruction counts of the mnemonics that prove the
Blackwell paths - UTCHMMA (tcgen05.mma), LDTM / STTM (tcgen05.ld / st), UTCBAR (tcgen05.commit), UBLKCP (cp.async.bulk),
UTMALDG (TMA tensor loads), HMMA (mma.sync), FADD2 / FMUL2 (packed fp32x2), FMNMX3 (3-input min/max) - plus registers.
  python tools/sass_summary.py > profiles/r2_sass_summary.txt
"""
import os
import re
import subprocess
import sys
from collections import OrderedDict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "neural-renderer_b200", "libnrb200.so")
MNEMONICS = ["UTCHMMA", "LDTM", "STTM", "UTCBAR", "UBLKCP", "UTMALDG", "HMMA", "LDSM", "FADD2", "FMUL2", "FMNMX3", "LDG.E.128", "ATOM", "RED"]


def clean(n):
    n = n.replace("nrb::", "").replace("(anonymous namespace)::", "")
    n = re.sub(r"^void ", "", n)
    n = re.sub(r"\((?!.*<).*$", "", n) if "<" not in n else re.sub(r">\(.*$", ">", n)
    return n


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], stdout=subprocess.PIPE, text=True).stdout
    kernels = OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = subprocess.run(["c++filt", m.group(1)], stdout=subprocess.PIPE, text=True).stdout.strip()
            cur = clean(cur)
            kernels[cur] = {k: 0 for k in MNEMONICS}
            kernels[cur]["instructions"] = 0
            continue
        if cur and re.search(r"/\*[0-9a-f]{4,}\*/", line):
            kernels[cur]["instructions"] += 1
            for k in MNEMONICS:
                if re.search(r"\b" + re.escape(k), line):
                    kernels[cur][k] += 1
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True).stdout
    regs = {}
    name = None
    for line in res.splitlines():
        m = re.search(r"Function (\S+):", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], stdout=subprocess.PIPE, text=True).stdout.strip()
            name = clean(name)
        m = re.search(r"REG:(\d+).*?SHARED:(\d+)", line)
        if m and name:
            regs[name] = (int(m.group(1)), int(m.group(2)))
    print(f"# SASS summary of {os.path.relpath(LIB, ROOT)} (sm_100a), cuobjdump -sass / -res-usage")
    cols = ["instructions", "REG"] + MNEMONICS
    print("kernel".ljust(34) + " ".join(c.rjust(9) for c in cols))
    for k, v in kernels.items():
        row = [v["instructions"], regs.get(k, ("-", 0))[0]] + [v[m] for m in MNEMONICS]
        print(k[:33].ljust(34) + " ".join(str(x).rjust(9) for x in row))


if __name__ == "__main__":
    main()
