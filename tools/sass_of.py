#!/usr/bin/env python
"""Dev helper: dump the SASS of one kernel of librm_b200.so (substring match on the mangled name)."""
import re, subprocess, sys
so, pat = sys.argv[1], sys.argv[2]
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
for f in re.split(r'\n\s*Function : ', txt)[1:]:
    name = f.split('\n', 1)[0]
    if pat in name:
        ins = [re.sub(r'/\*[0-9a-fx]+\*/', '', l).strip().rstrip(';').strip()
               for l in f.split('\n') if re.match(r'\s+/\*[0-9a-f]{4,5}\*/', l)]
        print(f"# {name} : {len(ins)} instructions")
        for i, x in enumerate(ins):
            print(f"{i:5d} {x}")
