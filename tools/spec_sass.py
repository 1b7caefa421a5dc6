#!/usr/bin/env python
"""Dev helper (no GPU): the generated specialised kernel of a scene compiled with nvcc for sm_100a, and its SASS
summarised — registers, spill, instruction mix, local-memory instructions, constant-bank operands, size of the march loop.

  python tools/spec_sass.py [--scene S2] [--mode fast|exact] [--rays 0] [--literals] [--dump out.sass] [--minb N]
"""
import argparse
import collections
import os
import re
import subprocess
import sys
import tempfile
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default="S2")
    ap.add_argument("--mode", default="fast")
    ap.add_argument("--rays", type=int, default=0)
    ap.add_argument("--literals", action="store_true")
    ap.add_argument("--minb", type=int, default=0)
    ap.add_argument("--dump", default=None)
    ap.add_argument("--src", default=None, help="write the generated source here")
    ap.add_argument("--lib", default=None, help="another build of librm_b200.so (its generator)")
    a = ap.parse_args()
    if a.minb:
        os.environ["RM_SPEC_MINB"] = str(a.minb)
    import helpers as h
    import rm_import
    rm = rm_import.load()
    if a.lib:
        rm.LIB_PATH = Path(a.lib).resolve()
    cmd, words = h.SCENES[a.scene]()
    src = rm.spec_source(cmd, words, (rm.MODE_FAST if a.mode == "fast" else rm.MODE_EXACT) | (rm.SPEC_LITERAL_TIER if a.literals else 0), a.rays)
    if a.src:
        Path(a.src).write_text(src)
    with tempfile.TemporaryDirectory() as td:
        cu = Path(td) / "rm_spec.cu"
        cu.write_text(src)
        cubin = Path(td) / "rm_spec.cubin"
        r = subprocess.run(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-cubin", "-O3", "-std=c++17", "-lineinfo", "--fmad=true",
                            "-Xptxas", "-v", f"-I{ROOT / 'ray-marching_b200' / 'csrc'}", "-o", str(cubin), str(cu)], capture_output=True, text=True)
        if r.returncode != 0:
            print(r.stderr)
            return 1
        print("\n".join(l for l in r.stderr.splitlines() if "registers" in l or "spill" in l))
        sass = subprocess.run(["cuobjdump", "-sass", str(cubin)], capture_output=True, text=True).stdout
    if a.dump:
        Path(a.dump).write_text(sass)
    ins = []
    for l in sass.splitlines():
        m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", l)
        if m:
            ins.append((int(m.group(1), 16), m.group(2).strip()))
    print(f"{len(ins)} SASS instructions, {ins[-1][0] + 16 if ins else 0} bytes")
    # loops = backward branches; the march loop is the first big one (the normal taps' copies of map_scene follow it)
    loops = []
    for addr, t in ins:
        mm = re.search(r"BRA(?:\.U)?(?:\s+!?U?P\d,)?\s+0x([0-9a-f]+)", t)
        if mm and int(mm.group(1), 16) < addr:
            tgt = int(mm.group(1), 16)
            body = [x for ad, x in ins if tgt <= ad <= addr]
            loops.append((tgt, addr, body))
            print(f"loop {tgt:#x}..{addr:#x}: {len(body)} instructions, STL {sum('STL' in x for x in body)} LDL {sum('LDL' in x for x in body)} "
                  f"LDCU {sum('LDCU' in x for x in body)} LDC {sum(x.split()[0].startswith('LDC') and 'LDCU' not in x for x in body)} MUFU {sum('MUFU' in x for x in body)}")
    big = [l for l in loops if len(l[2]) > 0.5 * max(len(x[2]) for x in loops)] if loops else []
    loop = big[0][2] if big else [t for _, t in ins]
    print(f"march loop taken as the first large one: {len(loop)} instructions, {len(loop) * 16} bytes")
    def mix(lst):
        c = collections.Counter()
        for t in lst:
            t = re.sub(r"^@!?U?P\d+\s+", "", t)
            c[t.split()[0].split(".")[0]] += 1
        return c
    cm = mix(loop)
    print("loop mix:", ", ".join(f"{k} {v}" for k, v in cm.most_common(24)))
    allm = mix([t for _, t in ins])
    print("local memory: STL", allm.get("STL", 0), "LDL", allm.get("LDL", 0), "| in loop: STL", cm.get("STL", 0), "LDL", cm.get("LDL", 0))
    cb = sum(1 for t in loop if re.search(r"c\[0x0\]\[0x", t))
    ur = sum(1 for t in loop if re.search(r"\bUR\d+", t))
    print(f"loop instructions with a c[0x0][..] operand: {cb}; with a uniform-register operand: {ur}; ULDC/LDCU: {cm.get('ULDC', 0) + cm.get('LDCU', 0)}; LDC: {cm.get('LDC', 0)}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
