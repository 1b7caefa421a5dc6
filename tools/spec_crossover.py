#!/usr/bin/env python
"""Dev tool: where does program specialisation stop paying?  Times S2-style chains of n leaves (1920x1080, fast mode)
through the interpreter and through the specialised kernel.  RM_SPEC_R=2 (or 4) in the environment forces the rays per
thread and bypasses the generator's code-size refusal, so chains beyond the budget can be measured too."""
import os, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import numpy as np
import rm_import, helpers as h
rm = rm_import.load()
W, H = 1920, 1080
u = h.orbit_uniforms(W, H, yaw=0.6, pitch=0.4)


def timed(ctx):
    for _ in range(2): ctx.paint(W, H)
    ts = []
    for _ in range(5):
        ctx.mark(0); ctx.paint(W, H); ctx.mark(1); ctx.sync(); ts.append(ctx.elapsed_ms(0, 1))
    return float(np.median(ts))


print("RM_SPEC_R=%s" % os.environ.get("RM_SPEC_R", "auto"))
with rm.Context(0, max_words=3072) as ctx:
    ctx.set_mode(rm.MODE_FAST)
    for n in [int(a) for a in sys.argv[1:]] or (16, 32, 48, 64, 96, 128):
        cmd, words = h.scene_s2(n_leaves=n)
        ctx.prepare(u, cmd, words)
        ctx.set_specialize(False)
        t_i = timed(ctx)
        try:
            ctx.specialize()
            info = ctx.spec_info()
            t_s = timed(ctx)
            print(f"leaves {n:4d} words {len(words):5d}: interpreter {t_i:8.3f} ms  specialised {t_s:8.3f} ms  x{t_i / t_s:.2f}  "
                  f"(rays/thread {info['rays_per_thread']}, cubin {info['cubin_bytes']} B, compile {info['compile_ms']:.0f} ms)")
        except rm.RmError as e:
            print(f"leaves {n:4d} words {len(words):5d}: interpreter {t_i:8.3f} ms  specialised: {str(e)[:90]}")
