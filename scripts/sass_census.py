#!/usr/bin/env python
"""SASS census of the shipped kernels (evidence that the default kernels are real sm_100a TMA code and that the
bit-exact tier is uncontracted): per kernel, the number of TMA loads (UTMALDG), bulk copies (UBLKCP), mbarrier ops
(SYNCS), fp64 adds / multiplies / FMAs, shared loads and stores, local-memory spills.

    python scripts/sass_census.py [--write profiles/r02_sass_census.txt]
"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "neptune_pde_solver_b200", "lib", "libneptune_rt_b200.so")

# the heat / Jacobi / 9-point / 27-coefficient bodies of BASELINE.json's configs (catalogue keys)
BODIES = {"c3 heat_star 3d f64": "16c0ec9675fa5599"}
MNEMONICS = ["UTMALDG", "UBLKCP", "SYNCS", "DADD", "DMUL", "DFMA", "FADD", "FMUL", "FFMA", "LDS", "STS", "LDG", "STG", "LDL", "STL", "BAR"]


def kernels_of(lib=LIB):
    out = subprocess.run(["cuobjdump", "-elf", lib], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    return sorted(set(re.findall(r"\.text\.(_ZN7neptune\w+)", out)))


def census(symbol, lib=LIB):
    out = subprocess.run(["cuobjdump", "-sass", "-fun", symbol, lib], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    ops = re.findall(r"^\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_]+)", out, flags=re.M)
    c = {m: 0 for m in MNEMONICS}
    for o in ops:
        if o in c:
            c[o] += 1
    c["total"] = len(ops)
    return c


def demangle(sym):
    try:
        return subprocess.run(["c++filt", sym], stdout=subprocess.PIPE, text=True).stdout.strip()
    except OSError:
        return sym


def main():
    key = BODIES["c3 heat_star 3d f64"]
    rows = []
    for s in kernels_of():
        if key not in s:
            continue
        rows.append((demangle(s).split("(")[0].replace("void neptune::", ""), census(s)))
    lines = ["# SASS census, body 0x%s (3-D 7-point fp64 heat step), %s" % (key, os.path.relpath(LIB, ROOT)),
             "# %-86s %s" % ("kernel", " ".join("%7s" % m for m in MNEMONICS + ["total"]))]
    for name, c in rows:
        lines.append("%-88s %s" % (name[:88], " ".join("%7d" % c[m] for m in MNEMONICS + ["total"])))
    text = "\n".join(lines) + "\n"
    if "--write" in sys.argv:
        open(sys.argv[sys.argv.index("--write") + 1], "w").write(text)
    print(text)


if __name__ == "__main__":
    main()
