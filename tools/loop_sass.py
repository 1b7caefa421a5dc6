#!/usr/bin/env python
"""Dev tool: opcode histogram of the interpreter loop(s) of one kernel in a cubin (loops found by their LDS.128 fetches).
usage: loop_sass.py <cubin> <kernel-name-substring>"""
import re, subprocess, sys, collections
txt = subprocess.run(["cuobjdump", "-sass", sys.argv[1]], capture_output=True, text=True).stdout
for f in re.split(r'\n\s*Function : ', txt)[1:]:
    name = f.split('\n', 1)[0]
    if sys.argv[2] not in name:
        continue
    ins = [re.sub(r'/\*[0-9a-fx]+\*/', '', l).strip().rstrip(';').strip() for l in f.split('\n') if re.match(r'\s+/\*[0-9a-f]{4,5}\*/', l)]
    lds = [i for i, s in enumerate(ins) if s.startswith('LDS.128')]
    # group LDS.128 into loops: consecutive fetches closer than 40 instructions
    groups, cur = [], []
    for i in lds:
        if cur and i - cur[-1] > 40:
            groups.append(cur); cur = []
        cur.append(i)
    if cur: groups.append(cur)
    print(f"{name[-60:]}: {len(ins)} instructions, {len(groups)} interpreter loop(s)")
    for g in groups:
        a = g[0] - 4
        # loop end = first backward branch after the fetches whose target precedes them
        b = next((i for i in range(g[-1], len(ins)) if re.match(r'(@!?U?P\d+\s+)?BRA', ins[i]) and i - g[-1] > 100), len(ins) - 1)
        c = collections.Counter(re.split(r'[ .]', re.sub(r'^@!?U?P\d+\s+', '', s))[0] for s in ins[a:b + 1])
        movs = c.get('MOV', 0) + sum(1 for s in ins[a:b + 1] if 'IMAD.MOV' in s)
        print(f"   loop at {g[0]}..{b}: {b + 1 - a} instr, MOV-like {movs}, {dict(c.most_common(10))}")
