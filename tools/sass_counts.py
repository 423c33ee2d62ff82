#!/usr/bin/env python
"""SASS evidence for the built library: per kernel, the counts of the mnemonics that prove the Blackwell paths are in the binary
(tcgen05 MMA = UTCIMMA / UTCHMMA, TMA = UTMALDG, TMEM loads = LDTM, tcgen05.commit = UTCBAR, mbarrier = SYNCS, cp.async = LDGSTS,
POPC, DFMA, MUFU.LG2, REDUX/CREDUX, ATOM/RED), plus registers / shared memory / stack from cuobjdump -res-usage.

    python tools/sass_counts.py > profiles/r02_sass_counts.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "spydrpick_b200", "libspydrpick_b200.so")
KEYS = ["UTCIMMA", "UTCHMMA", "UTMALDG", "LDTM", "UTCBAR", "SYNCS", "LDGSTS", "POPC", "DFMA", "MUFU.LG2", "I2FP", "REDUX", "ATOM", "RED.", "SHFL", "LDS", "STS", "LDG", "STG"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return [re.sub(r"\(.*$", "", x).replace("void ", "").replace("spk::", "") for x in out]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout
    usage = {}
    cur = None
    for ln in res.splitlines():
        m = re.match(r"\s*Function (\S+):", ln)
        if m:
            cur = m.group(1)
        elif cur and "REG:" in ln:
            usage[cur] = " ".join(x for x in ln.split() if x.split(":")[0] in ("REG", "STACK", "SHARED"))
    counts = collections.OrderedDict()
    total = {}
    cur = None
    for ln in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", ln)
        if m:
            cur = m.group(1); counts[cur] = collections.Counter(); total[cur] = 0
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,6}\*/\s+(@!?U?P\d\s+)?([A-Z0-9_.]+)", ln)
        if cur and m:
            op = m.group(2); total[cur] += 1
            for k in KEYS:
                if op.startswith(k) or (k == "MUFU.LG2" and op.startswith("MUFU.LG2")):
                    counts[cur][k] += 1
    names = list(counts)
    pretty = demangle(names)
    print("# %s: cuobjdump -sass / -res-usage, sm_100a; static instruction counts per kernel" % os.path.relpath(LIB, ROOT))
    print("# arch lines: " + ", ".join(sorted(set(re.findall(r"arch = (sm_\w+)", sass)))))
    for nm, pn in sorted(zip(names, pretty), key=lambda t: -total[t[0]]):
        c = counts[nm]
        hits = ", ".join("%s %d" % (k, c[k]) for k in KEYS if c[k])
        print("%-60s %6d instr | %-28s | %s" % (pn[:60], total[nm], usage.get(nm, ""), hits))


if __name__ == "__main__":
    main()
