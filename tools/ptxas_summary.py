#!/usr/bin/env python
"""Dev helper: summarise ptxas -v output (registers / stack / spills per kernel) from build.log."""
import re, sys
lines = open(sys.argv[1] if len(sys.argv) > 1 else "build.log").read().split("\n")
name = None
stack = spill = "?"
for l in lines:
    m = re.search(r"Compiling entry function '([^']+)'", l)
    if m:
        name = m.group(1)
    m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores", l)
    if m:
        stack, spill = m.group(1), m.group(2)
    m = re.search(r"Used (\d+) registers", l)
    if m and name:
        k = re.search(r"(k_\w+?)I(Lb[01]E)?(Li\d+E)?(Lb[01]E)?", name)
        short = re.sub(r"^_ZN2rm\d+_GLOBAL__N__\w+?_cu_\w{8}\d+", "", name)[:40]
        print(f"{short:42s} regs={m.group(1):>3s} stack={stack:>4s} spill={spill}")
        name = None
