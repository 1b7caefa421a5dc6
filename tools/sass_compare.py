#!/usr/bin/env python
"""Compares the SASS of every kernel in two builds of librm_b200.so, ignoring kernel names (template signatures may
change) and addresses: prints which kernels of the FIRST library have no instruction-for-instruction equal in the SECOND,
and which kernels of the second are new.  Used to show that a change left the measured kernels untouched
(e.g. the N4 EXT instantiations: all 50 earlier kernels byte-identical, 16 new ones).

  python tools/sass_compare.py before/librm_b200.so ray-marching_b200/librm_b200.so
"""
import hashlib
import re
import subprocess
import sys


def kernels(so: str) -> dict:
    txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
    out, name, body = {}, None, []
    for line in txt.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            if name:
                out[name] = body
            name, body = m.group(1), []
        elif name:
            mm = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(.*?);", line)
            if mm:
                body.append(mm.group(1))
    if name:
        out[name] = body
    return out


def demangle(n: str) -> str:
    return subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()


def main():
    a, b = kernels(sys.argv[1]), kernels(sys.argv[2])
    digest = lambda body: hashlib.sha1("\n".join(body).encode()).hexdigest()  # noqa: E731
    hb = {digest(v) for v in b.values()}
    ha = {digest(v) for v in a.values()}
    changed = [k for k, v in a.items() if digest(v) not in hb]
    new = [k for k, v in b.items() if digest(v) not in ha]
    print(f"{len(a)} kernels in {sys.argv[1]}, {len(b)} in {sys.argv[2]}")
    print(f"kernels of the first with no identical SASS in the second: {len(changed)}")
    for k in changed:
        print("   ", demangle(k)[:160], f"({len(a[k])} instructions)")
    print(f"kernels only in the second: {len(new)}")
    for k in new:
        print("   ", demangle(k)[:160], f"({len(b[k])} instructions)")
    return 1 if changed else 0


if __name__ == "__main__":
    sys.exit(main())
