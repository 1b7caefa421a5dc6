#!/usr/bin/env python
"""Dev tool: time the default kernel on truncated S2 chains (n leaves = n fused instructions) at 1920x1080, fast mode.
Run once with RM_DIRECT_R=4 and once with RM_DIRECT_R=8 to find where 8 rays per thread starts to win."""
import os, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import numpy as np
import rm_import, helpers as h
rm = rm_import.load()
W, H = 1920, 1080
u = h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)
out = []
with rm.Context(0) as ctx:
    ctx.set_mode(rm.MODE_FAST)
    for n in (1, 2, 3, 4, 5, 6, 8, 10, 12, 16, 24, 32):
        cmd, words = h.scene_s2(n_leaves=n)
        ctx.prepare(u, cmd, words)
        for _ in range(3): ctx.paint(W, H)
        ts = []
        for _ in range(7):
            ctx.mark(0); ctx.paint(W, H); ctx.mark(1); ctx.sync(); ts.append(ctx.elapsed_ms(0, 1))
        out.append((n, float(np.median(ts))))
print("R=%s " % os.environ.get("RM_DIRECT_R", "auto") + " ".join(f"{n}:{t:.3f}" for n, t in out))
